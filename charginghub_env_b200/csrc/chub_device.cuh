// chub_device.cuh -- device-side building blocks of the hub step kernel (sm_100a).
//
// Everything here is arithmetic the reference performs per pile / per hub; the citations give
// the reference lines whose RESULTS each function reproduces.  The code is written for the GPU,
// not transcribed: polynomials are fp64 Horner chains with explicit fma(), pow() never appears,
// logf/expf with integer arguments come from host-built tables, and the libstdc++ polar normal
// is re-derived on top of a 32-bit minstd state.  The translation unit is compiled with
// -fmad=false so that every operation NOT written as fma() rounds exactly like the reference's
// separate multiply/add (x86-64 SSE2, no contraction); fp64 results then narrow to fp32 at the
// same points the reference's `float` returns do.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/chub.h"
#include "../../include/chub_rng.h"
// fp64 constants of the curves and the canonical variates, appended to the constant-bank table of chub_fastmath.h (see there)
#define CHUB_DK_CURVES(X)                                                                                              \
    /* canon_d_of */                                                                                                  \
    X(CAN_M, 2147483646.0) X(CAN_D, 4611686009837453312.0) X(CAN_RD, 1.0 / 4611686009837453312.0)                     \
    /* slow_step SEG 0, slow_t2p */                                                                                   \
    X(SC4, -0.0004112) X(SP4, -0.002056) X(SC3, 0.0023025) X(SP3, 0.00921) X(SC2, 0.03562 / 3) X(SP2, 0.03562)        \
    X(SC1, 0.011895) X(SP1, 0.02379) X(SC0, 6.007)                                                                    \
    /* slow_step SEG 1 */                                                                                             \
    X(S_END, 3.67 * 4) X(SL_A, -4.041) X(SL_X0, 0.485) X(SL_B, 19.140115) X(SL_C, 11.943306312699628) X(SL_P, 21.1)   \
    X(S_SCALE, 19.285746346634653) X(S_RSCALE, 1.0 / 19.285746346634653) X(S_SEG, 2.33 * 4)                           \
    /* slow_s2t */                                                                                                    \
    X(ST_BRK, 73.89239629561729) X(ST_B, 0.1295) X(ST_Q3, -0.001529) X(ST_Q2, 0.1871) X(ST_Q1, -10.15) X(ST_Q0, 213.1) \
    X(ST_ADD, 0.1787983924863248) X(ST_SLOPE, 0.2012016075138625) X(ST_W, 26.10760370438271)                          \
    X(ST_RW, 1.0 / 26.10760370438271)                                                                                 \
    /* fast_step, fast_t2p */                                                                                         \
    X(F_K, 0.053) X(F_A1, 0.7194 / 0.053) X(F_B1, 50.15) X(F_A, 0.7194) X(F_P0, 47.78) X(F_END, 51.2 / 15)            \
    X(F_SEG, 28.7 / 15)                                                                                               \
    X(FA4, 0.0002253) X(FI4, 0.0002253 / 5) X(FA3, 0.03572) X(FI3, -(0.03572 / 4)) X(FA2, 2.016) X(FI2, 2.016 / 3)    \
    X(FA1, 48.76) X(FI1, -(48.76 / 2)) X(FA0, 457.7)                                                                  \
    /* fast_s2t */                                                                                                    \
    X(FS_N2, (double) -18.18f) X(FS_D1, (double) -23.9f) X(FS_N1, (double) 9.559f) X(FS_D0, (double) 56.48f)          \
    X(FS_N0, (double) 48.99f) X(FS_DC, (double) 18.96f) X(FS_NC, (double) 29.09f)                                     \
    /* hub section (chub_kernels.cu): hydrogen system, income, observation */                                         \
    X(H_R5, 1.0 / 5.0) X(H_R210, 1.0 / 210.0) X(H_R92, 1.0 / 92.0) X(H_001, 0.01) X(H_R185, 1.0 / 18.5)               \
    X(H_R70, 1.0 / 70.0) X(H_01, 0.1) X(H_R900, 1.0 / 900.0) X(H_MOL, 2.02) X(H_RMOL, 1.0 / 2.02) X(H_EA, 0.326)      \
    X(H_EB, 1.476) X(H_R1000, 1.0 / 1000.0) X(H_08, 0.8) X(H_FC, 119.6) X(H_RFC, 1.0 / 119.6) X(H_M6, -6.0 / 1000.0)  \
    X(H_I42, 0.42 / 4.0) X(H_I21, 0.21 / 4.0) X(H_P6, 6.0 / 1000.0) X(H_M10, -10.0 / 1000.0) X(H_R50, 1.0 / 50.0)
#include "chub_fastmath.h"

// ---------------------------------------------------------------------------------------------
// constant-memory tables, filled by the host (chub_engine.cu) with this machine's glibc
struct ChubConst {
    float renege[16];      // (float)(0.1 * logf(i + 1))            CHS.hpp:1287
    float balk[64];        // expf((float)-(0.01 * k))               CHS.hpp:1299
    // the same two tests on the integer draw r = rand() % 1000 (u = r / 999.f is increasing in r):
    //   u >  renege[w]  <=>  r >  renege_k[w]        u <= balk[j]  <=>  r <= balk_k[j]
    short renege_k[16], balk_k[64];
    double price[96];      // price_after_MAD_96                     Aggregator_Simple.py:9-15
    double time_sin[96];   // sin(2 pi t / 96)                       evcssp_manager.py:319-320
    double price_mean, price_std, r_price_std;
    // curve constants that the reference evaluates with libm at run time
    float fast_aa1_c;      // (float)((0.7194/0.053) * exp(0))       CHS.hpp:693
    float fast_aa2_c;      // aa_part2_c                             CHS.hpp:698-700
    float fast_scale;      // 100 / aa_part2(51.2f)  (fp32 division) CHS.hpp:652
    float slow_c2_end;     // c_part2((float)3.67)                   CHS.hpp:576
    double slow_stt1_a;    // -4.276 * pow(10,-5)                    CHS.hpp:581
    double slow_stt2_a;    //  4.742 * pow(10,-6)                    CHS.hpp:585
    // J2601 lookup (hydro_sys.py:232-233)
    double j_x[10], j_y[10];
    double j_slope[8];     // (y[2+a]-y[1+a])/(x[2+a]-x[1+a])
    double topoff_time;    // (87.4 - 69) / 7.2                      hydro_sys.py:311
    double fcev_cap;       // 6.3 * 1000                             hydro_sys.py:221
};

// per hub-class constants (derived on the host from chub_hub_class; hydro_sys.py:9-158)
struct ChubClass {
    int n[2];
    int type[2];
    int constant;
    float perm_f;          // (float) fcev_permeate, as Boost.Python would narrow it
    float possible_in_f;   // (float) 0.3
    float transformer[2];  // CHS.hpp:1153-1154 / 1463-1464
    double half[2];        // transformer / 2 (evcssp_manager.py:377)
    double r_half[2], r_el_cells, r_el_den, r_cap_mass;  // correctly rounded reciprocals for chub_div_const
    double v_h_max, v_M, cap_mass, cap_floor /* cap_mass*0.1 */, floor_01 /* 0.1*cap_mass */;
    double init_soc, init_mass, loss, fc_cells, cpr_w12;
    double el_cells;       // integer valued
    double el_den;         // v_M * 1000 * 60
};

#ifdef CHUB_KERNEL_TU  // device code below is compiled only into chub_kernels.cu
__constant__ ChubConst kC;
__device__ ChubLogEntry kLogTab[CHUB_LOG_N];  // reduction table of chub_log (lanes index it divergently)
// The step kernel keeps a copy at the start of its dynamic shared memory (a 16-byte gather from there is two
// instructions and a short-scoreboard wait); the reset kernel reads the global copy.  LT selects which.
extern __shared__ __align__(16) unsigned char chub_smem[];
template <int LT>
__device__ __forceinline__ const ChubLogEntry *log_tab() {
    if (LT) return reinterpret_cast<const ChubLogEntry *>(chub_smem);
    return kLogTab;
}

// ---------------------------------------------------------------------------------------------
// per-hub random streams (include/chub_rng.h)
// @region rng/normals
struct HubRng {
    uint32_t k0, k1, uctr, lcg;
};

__device__ __forceinline__ int rng_u1000(HubRng &r) {  // rand() % 1000, CHS.hpp:41
    return (int) chub_urand(r.k0, r.k1, r.uctr++);   // already in [0, 1000): the % 1000 of CHS.hpp:41 is the identity on it
}
__device__ __forceinline__ float rng_u01(HubRng &r) {  // uniform_rand(0,1): tr*(1-0)+0
    float tr = (float) rng_u1000(r) / 999.0f;
    return tr * 1.0f + 0.0f;
}
// The minstd-driven normals are deliberately NOT inlined: they are called from several places
// (both station types, reset, FCEV arrivals) and one shared copy keeps the step kernel inside the
// instruction cache.  State goes in and out by value so nothing is forced into local memory.

// libstdc++ generate_canonical<double,53> over minstd_rand0: two draws (bits/random.tcc:3349-3381)
__device__ __forceinline__ double canon_d(uint32_t &lcg) {
    lcg = chub_minstd_next(lcg);
    double a = (double) (lcg - 1u);
    lcg = chub_minstd_next(lcg);
    double b = (double) (lcg - 1u);
    double sum = a + b * DK(CAN_M);
    // sum / D with D = (2^31-2)^2 rounded to double = 2^62 - 2^33: reciprocal estimate plus one fma residual
    // correction (Markstein) gives the correctly rounded quotient without the generic division routine
    const double D = DK(CAN_D), rD = DK(CAN_RD);
    double q0 = sum * rD;
    double ret = fma(fma(-q0, D, sum), rD, q0);
    if (ret >= 1.0) ret = 0.99999999999999988898;
    return ret;
}
// the same two, from states already drawn (the step kernel's add_car chain computes the states in parallel)
__device__ __forceinline__ double canon_d_of(uint32_t s1, uint32_t s2) {
    const double sum = (double) (s1 - 1u) + (double) (s2 - 1u) * DK(CAN_M);
    const double D = DK(CAN_D), rD = DK(CAN_RD);
    const double q0 = sum * rD;
    double ret = fma(fma(-q0, D, sum), rD, q0);
    if (ret >= 1.0) ret = 0.99999999999999988898;
    return ret;
}
__device__ __forceinline__ float canon_f_of(uint32_t s1) {
    float ret = (float) (s1 - 1u) * 4.656612873077392578125e-10f;
    if (ret >= 1.0f) ret = 0.99999994f;
    return ret;
}
// generate_canonical<float,24>: one draw, divisor float(2147483646) = 2^31
__device__ __forceinline__ float canon_f(uint32_t &lcg) {
    lcg = chub_minstd_next(lcg);
    float ret = (float) (lcg - 1u) * 4.656612873077392578125e-10f;  // exact: division by 2^31
    if (ret >= 1.0f) ret = 0.99999994f;
    return ret;
}

// mk_soc (CHS.hpp:804-814): std::normal_distribution<double>(7, 3)(e) on a FRESH distribution object --
// Marsaglia polar, second variate discarded (bits/random.tcc:1811-1844) -- clamped and mapped to an SOC.
// Returns {soc bits, new minstd state}.
template <int LT>
__device__ __noinline__ uint2 draw_mk_soc(uint32_t lcg) {
    // The accept/reject walk only needs to know whether x^2 + y^2 <= 1.  With c = (lo + hi (2^31 - 2)) / (2^31 - 2)^2
    // = hi / (2^31 - 2) to within 5e-10 (hi = the second state of the pair, minus 1), an fp32 estimate of r2 is good
    // to ~1e-6; the attempt is evaluated exactly, like the reference, only inside a +-4e-6 band around the boundary.
    // Rejected attempts (21 %) then cost a few fp32 operations instead of two fp64 canonical variates.
    uint32_t s0;
    for (;;) {
        s0 = lcg;
        const uint32_t s2 = chub_minstd_jump(s0, CHUB_MINSTD_A2), s4 = chub_minstd_jump(s0, CHUB_MINSTD_A4);
        lcg = s4;
        const float xa = fmaf((float) (s2 - 1u), 9.31322575e-10f, -1.0f), ya = fmaf((float) (s4 - 1u), 9.31322575e-10f, -1.0f);
        const float r2a = fmaf(xa, xa, ya * ya);
        bool ok = r2a < 1.0f;
        if (fabsf(r2a - 1.0f) < 4e-6f || r2a < 4e-6f) {
            const double xe = 2.0 * canon_d_of(chub_minstd_next(s0), s2) - 1.0;
            const double ye = 2.0 * canon_d_of(chub_minstd_jump(s0, CHUB_MINSTD_A3), s4) - 1.0;
            const double r2e = xe * xe + ye * ye;
            ok = !(r2e > 1.0 || r2e == 0.0);
        }
        if (ok) break;
    }
    const double x = 2.0 * canon_d_of(chub_minstd_next(s0), chub_minstd_jump(s0, CHUB_MINSTD_A2)) - 1.0;
    const double y = 2.0 * canon_d_of(chub_minstd_jump(s0, CHUB_MINSTD_A3), lcg) - 1.0;
    const double r2 = x * x + y * y;
    double mult = sqrt(chub_div(-2.0 * chub_log_r2(r2, log_tab<LT>()), r2));
    float de = (float) ((y * mult) * 3.0 + 7.0);
    de = fminf(fmaxf(de, 1.0f), 10.0f);
    return make_uint2(__float_as_uint((float) (75.0 - 5.0 * (double) de)), lcg);
}
// The serial part of the two polar normals of add_car -- walking the minstd stream to the attempt that is accepted -- lives
// in the step kernel (chub_kernels.cu, add_car: one loop trip per attempt of either kind); the variate itself is a pure
// function of the state its accepted attempt started from and is evaluated later, by any lane (mk_soc_from / late_time_from).
// mk_soc (CHS.hpp:804-814) from the start state of its accepted attempt
template <int LT>
__device__ __forceinline__ float mk_soc_from(uint32_t s0) {
    const double x = 2.0 * canon_d_of(chub_minstd_next(s0), chub_minstd_jump(s0, CHUB_MINSTD_A2)) - 1.0;
    const double y = 2.0 * canon_d_of(chub_minstd_jump(s0, CHUB_MINSTD_A3), chub_minstd_jump(s0, CHUB_MINSTD_A4)) - 1.0;
    const double r2 = x * x + y * y;
    const double mult = sqrt(chub_div(-2.0 * chub_log_r2(r2, log_tab<LT>()), r2));
    float de = (float) ((y * mult) * 3.0 + 7.0);
    de = fminf(fmaxf(de, 1.0f), 10.0f);
    return (float) (75.0 - 5.0 * (double) de);
}
// mk_late_time (CHS.hpp:816-830) from the start state of its accepted attempt
template <int LT>
__device__ __forceinline__ int late_time_from(uint32_t s0) {
    const float x = 2.0f * canon_f_of(chub_minstd_next(s0)) - 1.0f, y = 2.0f * canon_f_of(chub_minstd_jump(s0, CHUB_MINSTD_A2)) - 1.0f;
    const float r2 = x * x + y * y;
    const float lg = (float) chub_log_r2((double) r2, log_tab<LT>());
    const float v = (y * sqrtf(-2.0f * lg / r2)) * 2.0f + 2.0f;
    const int d = (int) roundf(v);
    return d > 0 ? d : 0;
}

// float normal with unit-free (mean, sd) (CHS.hpp:819-820, 833-834).  logf is taken as the correctly
// rounded fp32 log (fp64 log narrowed), which is what glibc's logf returns except on rare hard cases.
// Returns {value bits, new minstd state}.
template <int LT>
__device__ __noinline__ uint2 draw_normal_f(uint32_t lcg, float mean, float sd) {
    float x, y, r2;
    do {
        x = 2.0f * canon_f(lcg) - 1.0f;
        y = 2.0f * canon_f(lcg) - 1.0f;
        r2 = x * x + y * y;
    } while (r2 > 1.0f || r2 == 0.0f);
    float lg = (float) chub_log_r2((double) r2, log_tab<LT>());
    float mult = sqrtf(-2.0f * lg / r2);
    float ret = y * mult;
    return make_uint2(__float_as_uint(ret * sd + mean), lcg);
}

template <int LT = 0>
__device__ __forceinline__ float rng_mk_soc(HubRng &r) {
    uint2 v = draw_mk_soc<LT>(r.lcg);
    r.lcg = v.y;
    return __uint_as_float(v.x);
}
template <int LT = 0>
__device__ __forceinline__ int rng_late_time(HubRng &r) {  // CHS.hpp:816-830 ("slow" for both pile types)
    uint2 v = draw_normal_f<LT>(r.lcg, 2.0f, 2.0f);
    r.lcg = v.y;
    int d = (int) roundf(__uint_as_float(v.x));
    return d > 0 ? d : 0;
}
__device__ __forceinline__ int rng_init_cars(HubRng &r, int mu) {  // CHS.hpp:832-842, theta 3
    uint2 v = draw_normal_f<0>(r.lcg, (float) mu, 1.0f);
    r.lcg = v.y;
    int t = (int) roundf(__uint_as_float(v.x));
    return min(max(t, mu - 3), mu + 3);
}

// ---------------------------------------------------------------------------------------------
// charging curves.  Time unit = one 15-minute step.  fp32 in/out, fp64 inside.

// @region slow curves
// --- slow charger (CHS.hpp:495-588) ---
__device__ __forceinline__ float slow_s2t(float s) {
    if (s < 0.0f) return 0.0f;
    double d = (double) s;
    // a s^2 + 0.1295 s with the square and the linear term rounded on their own, as the reference's pow(soc, 2) form
    // does: the shorter d * fma(a, d, 0.1295) differs from glibc on ONE reachable argument (61.03515625f), this form on
    // none (tests/test_gpu_curves.py walks all of them)
    if (d <= DK(ST_BRK)) return (float) fma(kC.slow_stt1_a, d * d, DK(ST_B) * d);
    if (d <= 100.0) {
        double p = fma(fma(fma(fma(kC.slow_stt2_a, d, DK(ST_Q3)), d, DK(ST_Q2)), d, DK(ST_Q1)), d, DK(ST_Q0));
        float q = (float) (p + DK(ST_ADD));
        return (float) ((double) q + chub_div_const(DK(ST_SLOPE) * (d - DK(ST_BRK)), DK(ST_W), DK(ST_RW)));
    }
    return 14.68f;
}
// time -> (soc, power) in one go: car_step needs both at the same time (CHS.hpp:901-905).
// The argument is always s2t(soc) + 1 >= 1, and the curve's breakpoints (2.33*4, 3.67*4 as doubles) are
// not fp32 values, so "<" and "<=" select the same side and one segment index decides both outputs:
//   SEG 0: t < 2.33*4 (quartics)      SEG 1: otherwise (log / rational, or the flat tail past 3.67*4)
// The caller sorts piles by segment so that a warp runs ONE of the two bodies, not both.
// (double) t < 2.33 * 4 for a float t is t < the smallest float above that double (tests/test_host_logic.py)
#define CHUB_SLOW_SEG_THR 0x1.2a3d72p+3f
__device__ __forceinline__ int slow_seg(float t) { return t < CHUB_SLOW_SEG_THR ? 0 : 1; }
template <int SEG>
__device__ __forceinline__ void slow_step(float t, float &soc, float &power) {
    const double y = (double) (t * 0.25f);  // charge_time / 4 is exact in fp32
    float ch;
    if (SEG == 0) {
        ch = (float) (y * fma(fma(fma(fma(DK(SC4), y, DK(SC3)), y, DK(SC2)), y, DK(SC1)), y, DK(SC0)));
        power = (float) fma(fma(fma(fma(DK(SP4), y, DK(SP3)), y, DK(SP2)), y, DK(SP1)), y, DK(SC0));
    } else if ((double) t <= DK(S_END)) {
        ch = (float) (DK(SL_A) * y + DK(SL_B) * chub_log(fabs(y - DK(SL_X0)), log_tab<1>()) + DK(SL_C));
        power = (float) chub_div(DK(SL_A) * y + DK(SL_P), y - DK(SL_X0));
    } else {
        ch = kC.slow_c2_end;
        power = 0.0f;
    }
    soc = (float) chub_div_const((double) (100.0f * ch), DK(S_SCALE), DK(S_RSCALE));
}
__device__ __forceinline__ float slow_t2p(float t) {
    double y = (double) (t * 0.25f), td = (double) t;
    if (td < DK(S_SEG)) return (float) fma(fma(fma(fma(DK(SP4), y, DK(SP3)), y, DK(SP2)), y, DK(SP1)), y, DK(SC0));
    if (td < DK(S_END)) return (float) chub_div(DK(SL_A) * y + DK(SL_P), y - DK(SL_X0));
    return 0.0f;
}

// @region fast curves
// --- fast charger (CHS.hpp:621-725) ---
__device__ __forceinline__ float fast_s2t(float s) {
    if (s <= 0.0f) return 0.0f;
    if (s >= 100.0f) return (float) (51.2 / 15);
    // norm_soc: fp32 arithmetic.  a / 31.48f as two fmas with the rounded reciprocal: equal to the IEEE quotient for
    // every float |a| in [2^-60, 128] (tools/verify_divf.c, exhaustive); a = s - 61.43f is 0 or >= 2^-19 in magnitude
    const float za = s - 61.43f, zq = za * (1.0f / 31.48f);
    const float zf = fmaf(fmaf(-zq, 31.48f, za), 1.0f / 31.48f, zq);
    double z = (double) zf, z2 = z * z;
    // p4*x and q3*x are fp32 products in the reference (float locals times a float argument)
    double p4z = (double) (-62.97f * zf), q3z = (double) (-50.12f * zf);
    double num = fma(fma(DK(FS_N2), z, DK(FS_N1)), z, DK(FS_N0)) * z2 + p4z + DK(FS_NC);
    double den = fma(fma(z, 1.0, DK(FS_D1)), z, DK(FS_D0)) * z2 + q3z + DK(FS_DC);
    return (float) chub_div(num, den);
}
// a_part2 (CHS.hpp:688-690), the quartic power fit of the fast charger's second segment.  It falls to zero at 51.2
// minutes, where its terms (~1500 each) cancel down to ~0.02: there the rounding of every partial result shows in the
// float32 value, and the fma-Horner chain differs from the reference on six reachable arguments.  Evaluated in the
// reference's own order -- powers by repeated multiplication, terms added left to right -- it equals glibc's pow form
// on every float32 argument of the segment (exhaustive: tests/test_gpu_curves.py).
__device__ __forceinline__ float fast_a2(double x) {
    const double x2 = x * x, x3 = x2 * x, x4 = x2 * x2;
    return (float) ((((DK(FA4) * x4 - DK(FA3) * x3) + DK(FA2) * x2) - DK(FA1) * x) + DK(FA0));
}
__device__ __forceinline__ float fast_aa2(double x) {
    double p = x * fma(fma(fma(fma(DK(FI4), x, DK(FI3)), x, DK(FI2)), x, DK(FI1)), x, DK(FA0));
    return (float) (p + (double) kC.fast_aa2_c);
}
// SEG 0: t < 28.7/15 (exponential part)   SEG 1: otherwise (polynomial part, or full past 51.2/15); t >= 1
#define CHUB_FAST_SEG_THR 0x1.e9d038p+0f
__device__ __forceinline__ int fast_seg(float t) { return t < CHUB_FAST_SEG_THR ? 0 : 1; }
template <int SEG>
__device__ __forceinline__ void fast_step(float t, float &soc, float &power) {
    const double x = (double) (t * 15.0f);
    if (SEG == 0) {
        const double ex = chub_exp(DK(F_K) * x);
        const float a1 = (float) (DK(F_A1) * ex + DK(F_B1) * x - (double) kC.fast_aa1_c);
        soc = a1 * kC.fast_scale;
        power = (float) (DK(F_A) * ex + DK(F_P0));
    } else if ((double) t < DK(F_END)) {
        soc = fast_aa2(x) * kC.fast_scale;
        power = fast_a2(x);
    } else {
        soc = 100.0f;
        power = 0.0f;
    }
}
__device__ __forceinline__ float fast_t2p(float t) {
    double x = (double) (t * 15.0f), td = (double) t;
    if (td >= 0.0 && td < DK(F_SEG)) return (float) (DK(F_A) * chub_exp(DK(F_K) * x) + DK(F_P0));
    if (td >= 0.0 && td < DK(F_END)) return fast_a2(x);
    return 0.0f;
}

// @region const/dispatch
// --- constant-power variants (CHS.hpp:496-501, 514-521, 528-535, 622-627, 640-647, 662-669) ---
__device__ __forceinline__ float const_s2t(float s, double tmax) {
    if (s <= 0.0f) return 0.0f;
    if (s >= 100.0f) return (float) tmax;
    return (float) (tmax * (double) s / 100.0);
}
__device__ __forceinline__ void const_step(float t, double tmax, float pconst, float &soc, float &power) {
    if (t <= 0.0f) soc = 0.0f;
    else if ((double) t >= tmax) soc = 100.0f;
    else soc = (float) ((double) (100.0f * t) / tmax);
    power = (tmax >= (double) t && t >= 0.0f) ? pconst : 0.0f;
}

#define CHUB_SLOW_TMAX 14.68
#define CHUB_FAST_TMAX 3.4133333333333336
#define CHUB_SLOW_PCONST ((float) 5.254973139368931)
#define CHUB_FAST_PCONST ((float) 36.44764034125146)

// dispatch on (station type, constant_charging), both compile-time in the step kernel
template <int TYPE, int CP>
__device__ __forceinline__ float curve_s2t(float s) {
    if (CP) return const_s2t(s, TYPE == CHUB_SLOW ? CHUB_SLOW_TMAX : CHUB_FAST_TMAX);
    return TYPE == CHUB_SLOW ? slow_s2t(s) : fast_s2t(s);
}
template <int TYPE, int CP>
__device__ __forceinline__ int curve_seg(float t) {
    if (CP) return 0;
    return TYPE == CHUB_SLOW ? slow_seg(t) : fast_seg(t);
}
template <int TYPE, int CP, int SEG>
__device__ __forceinline__ void curve_step(float t, float &soc, float &power) {
    if (CP) {
        const_step(t, TYPE == CHUB_SLOW ? CHUB_SLOW_TMAX : CHUB_FAST_TMAX,
                   TYPE == CHUB_SLOW ? CHUB_SLOW_PCONST : CHUB_FAST_PCONST, soc, power);
    } else if (TYPE == CHUB_SLOW) {
        slow_step<SEG>(t, soc, power);
    } else {
        fast_step<SEG>(t, soc, power);
    }
}
template <int TYPE, int CP>
__device__ __forceinline__ float curve_t2p(float t) {
    if (CP) {
        const double tmax = TYPE == CHUB_SLOW ? CHUB_SLOW_TMAX : CHUB_FAST_TMAX;
        return (tmax >= (double) t && t >= 0.0f) ? (TYPE == CHUB_SLOW ? CHUB_SLOW_PCONST : CHUB_FAST_PCONST) : 0.0f;
    }
    return TYPE == CHUB_SLOW ? slow_t2p(t) : fast_t2p(t);
}
// run-time flavour for the table-building kernel
__device__ __forceinline__ float curve_s2t_rt(int type, int cp, float s) {
    if (type == CHUB_SLOW) return cp ? curve_s2t<CHUB_SLOW, 1>(s) : curve_s2t<CHUB_SLOW, 0>(s);
    return cp ? curve_s2t<CHUB_FAST, 1>(s) : curve_s2t<CHUB_FAST, 0>(s);
}
#endif  // CHUB_KERNEL_TU
