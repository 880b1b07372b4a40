// Host check of csrc/chub_fastmath.h against libm over the argument ranges the step kernel uses.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <random>
#include "../charginghub_env_b200/csrc/chub_fastmath.h"

static double ulp_err(double got, double want) {
    if (got == want) return 0.0;
    double u = std::nextafter(std::fabs(want), INFINITY) - std::fabs(want);
    return std::fabs(got - want) / u;
}

int main(int argc, char **argv) {
    long n = argc > 1 ? atol(argv[1]) : 2000000;
    ChubLogEntry tab[CHUB_LOG_N];
    chub_build_log_table(tab);
    std::mt19937_64 g(12345);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    double max_log_a = 0, max_log_b = 0, max_exp = 0;
    long f_log = 0, f_exp = 0, f_norm = 0, bad_div = 0;
    for (long i = 0; i < n; ++i) {
        // slow charger: log|y - 0.485| with y in (2.33, 3.67]  (CHS.hpp:564-566)
        float yf = (float) (2.33 + (3.67 - 2.33) * U(g));
        double a = std::fabs((double) yf - 0.485);
        double l1 = chub_log(a, tab), l0 = std::log(a);
        max_log_a = std::fmax(max_log_a, ulp_err(l1, l0));
        float c1 = (float) (-4.041 * yf + 19.140115 * l1 + 11.943306312699628);
        float c0 = (float) (-4.041 * yf + 19.140115 * l0 + 11.943306312699628);
        f_log += c1 != c0;
        // polar normal: log(r2), r2 in (0, 1], both tiny and near-1 values
        double r2 = (i & 1) ? U(g) : std::pow(U(g), 12.0);
        if (r2 <= 0) r2 = 1e-300;
        double m1 = chub_log(r2, tab), m0 = std::log(r2);
        max_log_b = std::fmax(max_log_b, std::fabs(m1 - m0) / std::fmax(std::fabs(m0), 1e-3) / 2.220446049250313e-16);
        float y = (float) U(g);
        float v1 = (float) (y * std::sqrt(-2 * m1 / r2) * 3.0 + 7.0), v0 = (float) (y * std::sqrt(-2 * m0 / r2) * 3.0 + 7.0);
        f_norm += v1 != v0;
        // fast charger: exp(0.053 x), x in [0, 28.7]  (CHS.hpp:684-695)
        float xf = (float) (28.7 * U(g));
        double e1 = chub_exp(0.053 * xf), e0 = std::exp(0.053 * xf);
        max_exp = std::fmax(max_exp, ulp_err(e1, e0));
        f_exp += (float) (0.7194 * e1 + 47.78) != (float) (0.7194 * e0 + 47.78);
        // division by constants
        double num = 100.0 * U(g);
        bad_div += chub_div_const(num, 19.285746346634653, 1.0 / 19.285746346634653) != num / 19.285746346634653;
        bad_div += chub_div_const(num, 26.10760370438271, 1.0 / 26.10760370438271) != num / 26.10760370438271;
        // every literal / class divisor the hub section uses (csrc/chub_kernels.cu DIVC / chub_div_const)
        static const double DS[] = {900.0, 2.02, 1000.0, 0.8, 119.6, 50.0, 5.0, 210.0, 92.0, 18.5, 70.0,
                                    179440.0, 771589.0, 24.436 * 1000 * 60, 445000.0, 178000.0, 89000000.0,
                                    364.47637939453125, 65.687164306640625, 0.05334155497008051};
        {
            double big = std::ldexp(U(g), (int) (40 * U(g)) - 10);
            for (double Dv : DS) bad_div += chub_div_const(big, Dv, 1.0 / Dv) != big / Dv;
        }
        double den = 0.5 + 40 * U(g);
        bad_div += chub_div(num, den) != num / den;
    }
    // the float polar normal exactly as the kernel forms it (csrc/chub_device.cuh late_time_from), on EVERY float r2 of
    // the last 2^16 below and including 1.0f and on the smallest ones: x^2 + y^2 rounds to exactly 1.0f once in about
    // 2^24 attempts, where log must be +0 (glibc) and not the table path's 1.8e-18 (sqrt of a negative: NaN)
    long bad_edge = 0;
    for (int k = 0; k < 65536 + 4096; ++k) {
        float r2f = k < 65536 ? std::nextafterf(1.0f, 0.0f) : 1e-38f;
        if (k < 65536) { r2f = 1.0f; for (int j = 0; j < k; ++j) r2f = std::nextafterf(r2f, 0.0f); if (k > 64) r2f = 1.0f - k * 5.9604645e-08f; }
        else r2f = std::ldexp(1.0f + (k - 65536) / 4096.0f, -120 + (k % 97));
        for (float y : {0.94084024f, -0.3f, 1e-3f}) {
            float lg1 = (float) chub_log_r2((double) r2f, tab), lg0 = logf(r2f);
            float v1 = (y * sqrtf(-2.0f * lg1 / r2f)) * 2.0f + 2.0f, v0 = (y * sqrtf(-2.0f * lg0 / r2f)) * 2.0f + 2.0f;
            // the variate is only ever used rounded to an integer (CHS.hpp:828, 835); glibc's logf is one ulp off the
            // correctly rounded value on a few arguments, which moves v by an ulp but not its rounding
            bad_edge += std::isnan(v1) || std::isnan(v0) || std::fabs(v1) > 1e5f ? std::isnan(v1) : (int) roundf(v1) != (int) roundf(v0);
            if (getenv("CHUB_VERBOSE") && std::isnan(v1)) fprintf(stderr, "k %d r2 %.9g y %g lg1 %.9g lg0 %.9g v1 %.9g v0 %.9g\n", k, r2f, y, lg1, lg0, v1, v0);
        }
    }
    bad_edge += chub_log_r2(1.0, tab) != 0.0 || std::isnan(std::sqrt(chub_div(-2.0 * chub_log_r2(1.0, tab), 1.0)));
    printf("{\"normal_edge_mismatch\": %ld, ", bad_edge);
    printf("\"n\": %ld, \"log_curve_max_ulp\": %.3f, \"log_normal_max_relerr_eps\": %.3f, \"exp_max_ulp\": %.3f, "
           "\"float_mismatch_log\": %ld, \"float_mismatch_normal\": %ld, \"float_mismatch_exp\": %ld, \"div_mismatch\": %ld}\n",
           n, max_log_a, max_log_b, max_exp, f_log, f_norm, f_exp, bad_div);
    return 0;
}
