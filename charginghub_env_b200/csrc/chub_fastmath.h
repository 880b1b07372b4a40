// chub_fastmath.h -- lean fp64 log / exp / divide for the hub step kernel (host + device).
//
// The reference evaluates its charging curves and polar normals with glibc's double-precision
// log/exp and IEEE division, then narrows to float.  CUDA's generic log()/exp()/'/' cost 100+/50/25
// instructions each because they handle denormals, infinities and NaNs; on this path every argument
// is a positive normal number in a known range.  These versions keep < 1 ulp (log, exp) and correctly
// rounded (divide, barring astronomically rare double-rounding cases) accuracy -- far inside what the
// fp32 narrowing needs -- in a fraction of the instructions.  The same source compiles for the host,
// where tests/test_fastmath.py checks it against libm over the argument ranges the kernel uses.
#pragma once
#include <stdint.h>
#include <string.h>
#if defined(__CUDACC__)
#define CHUB_FM __host__ __device__ __forceinline__
#else
#define CHUB_FM static inline
#endif
#include <math.h>

struct ChubLogEntry { double invc, logc; };  // 1/c rounded, and -log(invc)
#if !defined(__cplusplus)
typedef struct ChubLogEntry ChubLogEntry;
#endif
#define CHUB_LOG_N 128
#define CHUB_LOG_OFF 0x3fe6000000000000ull  // 0.6875: z = x * 2^-k lands in [0.6875, 1.375)

// fp64 constants of the hot device code, read from the constant bank.  A double literal whose low word is not zero
// costs the kernel TWO instructions (UMOV pairs into uniform registers) every time an expression uses it -- 8 % of all
// instructions the step kernel issued in a busy hour, more than its fp64 arithmetic proper.  From a __constant__ table
// two neighbouring doubles arrive in one LDCU.128, so the entries are listed in the order the code uses them.  Values are
// written exactly as the expressions they replace were (the compiler folds them the same way), so results do not change;
// the host build of these functions (tests/fastmath_check.cpp) sees the literals.
#define CHUB_DK_LIST(X)                                                                                                \
    /* chub_log */                                                                                                    \
    X(LN2_HI, 0x1.62e42fefa3800p-1) X(LN2_LO, 0x1.ef35793c76730p-45)                                                  \
    X(LG6, -1.0 / 6) X(LG5, 1.0 / 5) X(LG4, -1.0 / 4) X(LG3, 1.0 / 3)                                                 \
    /* chub_exp */                                                                                                    \
    X(INV_LN2, 0x1.71547652b82fep+0) X(EX13, 1.0 / 6227020800.0) X(EX12, 1.0 / 479001600.0) X(EX11, 1.0 / 39916800.0) \
    X(EX10, 1.0 / 3628800.0) X(EX9, 1.0 / 362880.0) X(EX8, 1.0 / 40320.0) X(EX7, 1.0 / 5040.0) X(EX6, 1.0 / 720.0)    \
    X(EX5, 1.0 / 120.0) X(EX4, 1.0 / 24.0) X(EX3, 1.0 / 6.0)                                                          \
    CHUB_DK_CURVES(X)
#ifndef CHUB_DK_CURVES
#define CHUB_DK_CURVES(X)
#endif
struct ChubDK {
#define CHUB_DK_FIELD(n, v) double n;
    CHUB_DK_LIST(CHUB_DK_FIELD)
#undef CHUB_DK_FIELD
};
#if defined(__CUDACC__)
static __constant__ ChubDK kDK = {
#define CHUB_DK_INIT(n, v) (v),
    CHUB_DK_LIST(CHUB_DK_INIT)
#undef CHUB_DK_INIT
};
#endif
#if defined(__CUDA_ARCH__)
#define DK(n) (kDK.n)
#else
#define CHUB_DK_LIT(n, v) static const double CHUB_DKV_##n = (v);
CHUB_DK_LIST(CHUB_DK_LIT)
#undef CHUB_DK_LIT
#define DK(n) (CHUB_DKV_##n)
#endif

CHUB_FM uint64_t chub_d2u(double x) {
#if defined(__CUDA_ARCH__)
    return (uint64_t) __double_as_longlong(x);
#else
    uint64_t u; memcpy(&u, &x, 8); return u;
#endif
}
CHUB_FM double chub_u2d(uint64_t u) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long) u);
#else
    double x; memcpy(&x, &u, 8); return x;
#endif
}

// @region chub_log
// log(x) for positive normal x.  Table-driven reduction (128 subintervals of [0.6875, 1.375)),
// r = z*invc - 1 with |r| < 0.004, then log1p(r) by a degree-6 Taylor polynomial.
CHUB_FM double chub_log(double x, const ChubLogEntry *tab) {
    const uint64_t ix = chub_d2u(x);
    const uint64_t tmp = ix - CHUB_LOG_OFF;
    const int i = (int) ((tmp >> 45) & (CHUB_LOG_N - 1));
    const int64_t k = (int64_t) tmp >> 52;
    const double z = chub_u2d(ix - (tmp & (0xfffull << 52)));
    const ChubLogEntry e = tab[i];
    const double r = fma(z, e.invc, -1.0);
    const double kd = (double) k;
    const double w = fma(kd, DK(LN2_HI), e.logc);  // k*ln2_hi is exact for |k| < 2^11
    const double hi = w + r;
    const double lo = (w - hi) + r + kd * DK(LN2_LO);
    const double r2 = r * r;
    // log1p(r) - r = r^2 (-1/2 + r/3 - r^2/4 + r^3/5 - r^4/6)
    const double p = fma(fma(fma(fma(DK(LG6), r, DK(LG5)), r, DK(LG4)), r, DK(LG3)), r, -0.5);
    return hi + fma(r2, p, lo);
}

// log of the polar method's r2 in (0, 1].  glibc's log / logf return exactly +0 for 1, and the normal variate then is
// y * sqrt(-0) = 0; the table path above returns 1.8e-18 there, whose square root argument is NEGATIVE -> NaN.  r2 == 1
// happens once in about 2^24 float attempts (x^2 + y^2 rounds to 1.0f), i.e. every few steps somewhere in a large batch.
CHUB_FM double chub_log_r2(double r2, const ChubLogEntry *tab) { return r2 == 1.0 ? 0.0 : chub_log(r2, tab); }

// @region chub_exp
// exp(a) for |a| < 700 (no overflow / underflow handling).
CHUB_FM double chub_exp(double a) {
    const double kd = rint(a * DK(INV_LN2));  // a / ln2
    double r = fma(-kd, DK(LN2_HI), a);
    r = fma(-kd, DK(LN2_LO), r);
    double p = DK(EX13);  // 1/13!
    p = fma(p, r, DK(EX12));
    p = fma(p, r, DK(EX11));
    p = fma(p, r, DK(EX10));
    p = fma(p, r, DK(EX9));
    p = fma(p, r, DK(EX8));
    p = fma(p, r, DK(EX7));
    p = fma(p, r, DK(EX6));
    p = fma(p, r, DK(EX5));
    p = fma(p, r, DK(EX4));
    p = fma(p, r, DK(EX3));
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    const int64_t k = (int64_t) kd;
    return chub_u2d(chub_d2u(p) + ((uint64_t) k << 52));  // p * 2^k
}

// @region chub_div
// n / d for normal operands whose quotient is normal: reciprocal by Newton from a hardware seed,
// then one fma residual correction (Markstein).
CHUB_FM double chub_div(double n, double d) {
#if defined(__CUDA_ARCH__)
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    double e = fma(-d, r, 1.0);  // seed has ~20 good bits; two Newton steps reach full precision
    r = fma(r, e, r);
    e = fma(-d, r, 1.0);
    r = fma(r, e, r);
#else
    double r = 1.0 / d;
#endif
    const double q = n * r;
    return fma(fma(-q, d, n), r, q);
}

// n / D for a compile-time constant D: rD must be the correctly rounded 1/D.
CHUB_FM double chub_div_const(double n, double D, double rD) {
    const double q = n * rD;
    return fma(fma(-q, D, n), rD, q);
}

// host: fill the log table (long double libm so logc carries no visible error)
static inline void chub_build_log_table(ChubLogEntry *tab) {
    for (int i = 0; i < CHUB_LOG_N; ++i) {
        const double lo = chub_u2d(CHUB_LOG_OFF + ((uint64_t) i << 45));
        const double hi = chub_u2d(CHUB_LOG_OFF + ((uint64_t) (i + 1) << 45));
        const double c = 0.5 * (lo + hi);
        tab[i].invc = 1.0 / c;
        tab[i].logc = (double) (-logl((long double) tab[i].invc));
    }
}
