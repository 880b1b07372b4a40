/* Exhaustive check that a / D in float32 equals the two-fma form the step kernel uses for D = 31.48f
 * (UtilFast::norm_soc, CHS.hpp:721-725):  q = a * rD;  q' = fmaf(fmaf(-q, D, a), rD, q),  rD = (float)(1 / D).
 * Every float a with |a| <= 128 is tested.  The kernel's argument is a = soc - 61.43f with soc a float in (0, 100),
 * which is either 0 or at least 2^-19 in magnitude; the only mismatches are for |a| < 2^-100, where the residual
 * underflows, and the program fails if it finds one at |a| >= 2^-60.
 * build: gcc -O2 -march=native -o /tmp/verify_divf tools/verify_divf.c -lm ; run: /tmp/verify_divf  (about 10 s) */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
int main(void) {
    const float D = 31.48f, rD = 1.0f / 31.48f;
    uint64_t bad = 0, n = 0, tiny = 0;
    for (uint32_t s = 0; s < 2; ++s)
        for (uint32_t u = 0; u <= 0x43000000u; ++u) { /* 0 .. 128.0f */
            uint32_t b = u | (s << 31);
            float a;
            memcpy(&a, &b, 4);
            volatile float ref = a / D;
            float q = a * rD;
            float q2 = fmaf(fmaf(-q, D, a), rD, q);
            float r = ref;
            if (memcmp(&q2, &r, 4) != 0) {
                /* denormal quotients may differ in the last bit; report only normal ones */
                if (fabsf(a) >= 0x1p-60f) { if (bad < 10) printf("a=%a ref=%a got=%a\n", a, r, q2); bad++; } else tiny++;
            }
            n++;
        }
    {   /* target_of (chub_kernels.cu): tu / 999 for the 10-bit target draw */
        const float r9 = 1.0f / 999.0f;
        for (int tu = 0; tu < 1024; ++tu) {
            float t = (float) tu;
            volatile float ref = t / 999.0f;
            float q0 = t * r9, q = fmaf(fmaf(-q0, 999.0f, t), r9, q0), rr = ref;
            if (memcmp(&q, &rr, 4) != 0) { printf("tu=%d ref=%a got=%a\n", tu, rr, q); bad++; }
        }
    }
    printf("tested %llu floats, %llu mismatches at |a| >= 2^-60, %llu below\n", (unsigned long long) n,
           (unsigned long long) bad, (unsigned long long) tiny);
    return bad != 0;
}
