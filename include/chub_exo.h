/* chub_exo.h -- the per-hub random stream of the on-device exogenous generator (PV / wind / price noise).
 *
 * The reference draws its renewable day choice from Python's `random` and its three Ornstein-Uhlenbeck noises
 * from numpy's global generator (renewable.py:21-76, evcssp_manager.py:344-361).  Neither can be reproduced on
 * a GPU, and both are process-global, so a batched engine needs its own per-hub stream.  This header fixes it:
 *   X stream: counter based, the same double-PCG hash as the U stream on keys derived from the hub's keys;
 *   integers:  chub_xrand31(...) % n;
 *   normals:   Marsaglia polar method over pairs of X draws, one attempt = two counters, evaluated in fp64 with
 *              the lean log / divide of chub_fastmath.h, so host and device produce bit-identical variates.
 * Plain C: the CUDA engine and the C oracle both include it.
 */
#ifndef CHUB_EXO_H
#define CHUB_EXO_H

#include "chub_rng.h"
#include "../charginghub_env_b200/csrc/chub_fastmath.h"

CHUB_HD uint32_t chub_xrand31(uint32_t key0, uint32_t key1, uint32_t counter) {
    return chub_rand31(key0 ^ 0x9E3779B9u, key1 + 0x85EBCA6Bu, counter);
}

/* uniform in (-1, 1): (r + 0.5) / 2^30 - 1, exact in double */
CHUB_HD double chub_xsym(uint32_t r31) { return ((double) r31 + 0.5) * (1.0 / 1073741824.0) - 1.0; }

/* standard normal; advances *counter by two per attempt */
CHUB_HD double chub_xnormal(uint32_t key0, uint32_t key1, uint32_t *counter, const ChubLogEntry *log_table) {
    double x, y, r2;
    do {
        x = chub_xsym(chub_xrand31(key0, key1, *counter));
        y = chub_xsym(chub_xrand31(key0, key1, *counter + 1u));
        *counter += 2u;
        const double xx = x * x, yy = y * y;
        r2 = xx + yy;
    } while (r2 >= 1.0 || r2 == 0.0);
    return x * sqrt(chub_div(-2.0 * chub_log_r2(r2, log_table), r2));
}

#endif /* CHUB_EXO_H */
