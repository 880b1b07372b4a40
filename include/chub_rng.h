/* chub_rng.h -- the two per-hub random streams of the charging-hub step path.
 *
 * The reference draws from two PROCESS-GLOBAL streams:
 *   U stream: libc rand(), used only as rand() % 1000 / 999.f   (CHS.hpp:35-44)
 *   E stream: std::default_random_engine e  == minstd_rand0     (CHS.hpp:25)
 * A batched engine needs one independent pair per hub.  This header fixes them:
 *   U: counter based, rand() := chub_urand(key0, key1, counter), a PCG output hash of the counter reduced to [0, 1000)
 *      (the reference only ever forms rand() % 1000);
 *   E: the same minstd_rand0 recurrence x <- 16807 x mod (2^31-1), one 31-bit state per hub.
 * It is plain C so the CUDA engine (device code), the C oracle and the reference shim (which
 * interposes rand() with chub_rand31) all consume bit-identical streams.
 */
#ifndef CHUB_RNG_H
#define CHUB_RNG_H

#include <stdint.h>

#if defined(__CUDACC__)
#define CHUB_HD __host__ __device__ __forceinline__
#else
#define CHUB_HD static inline
#endif

/* PCG-RXS-M-XS style 32-bit output hash (bijective on uint32). */
CHUB_HD uint32_t chub_pcg32(uint32_t v) {
    uint32_t s = v * 747796405u + 2891336453u;
    uint32_t w = ((s >> ((s >> 28) + 4u)) ^ s) * 277803737u;
    return (w >> 22) ^ w;
}

/* The value the reference would have obtained from rand(): uniform in [0, 2^31). */
CHUB_HD uint32_t chub_rand31(uint32_t key0, uint32_t key1, uint32_t counter) {
    return chub_pcg32(chub_pcg32(counter ^ key0) + key1) >> 1;
}

/* The value the reference obtains from rand() IN THE U STREAM.  CHS.hpp:41 uses it only as rand() % 1000, so the stream is
 * free to deliver that residue directly: one hash round of the counter, the second key folded in, and a multiply-shift
 * range reduction to [0, 1000) -- half the instructions of chub_rand31(...) % 1000 on the device, where a busy hub draws
 * forty of these per step.  The reference shim's rand() returns exactly this value, the oracle and the engine reduce it
 * with the reference's own % 1000 (a no-op).  (The X stream of chub_exo.h keeps the full-range chub_rand31.) */
CHUB_HD uint32_t chub_urand(uint32_t key0, uint32_t key1, uint32_t counter) {
    uint32_t h = chub_pcg32(counter ^ key0) ^ key1;
    return (uint32_t) (((uint64_t) h * 1000u) >> 32);
}

/* minstd_rand0 step; state stays in [1, 2^31-2]. */
CHUB_HD uint32_t chub_minstd_next(uint32_t x) {
    /* 2^31 == 1 (mod 2^31-1): fold the high bits instead of dividing; identical to the % form */
    uint64_t p = (uint64_t) x * 16807ull;
    uint32_t r = (uint32_t) (p & 0x7fffffffull) + (uint32_t) (p >> 31);
    return r >= 2147483647u ? r - 2147483647u : r;
}

/* k steps at once: x * (16807^k mod (2^31-1)) mod (2^31-1) for a precomputed multiplier A < 2^31.
 * The product has up to 62 bits, so the fold is applied twice.  Lets a caller compute the next
 * few states independently of each other instead of as a dependent chain. */
#define CHUB_MINSTD_A2 282475249u
#define CHUB_MINSTD_A3 1622650073u
#define CHUB_MINSTD_A4 984943658u
CHUB_HD uint32_t chub_minstd_jump(uint32_t x, uint32_t A) {
    uint64_t p = (uint64_t) x * (uint64_t) A;
    uint32_t r = (uint32_t) (p & 0x7fffffffull) + (uint32_t) (p >> 31);   /* < 2^32 */
    r = (r & 0x7fffffffu) + (r >> 31);                                    /* <= 2^31 */
    return r >= 2147483647u ? r - 2147483647u : r;
}

CHUB_HD uint64_t chub_splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

/* Stream parameters of a hub: a pure function of (base seed, GLOBAL hub id), so results do not
 * depend on how hubs are sharded over GPUs. */
CHUB_HD void chub_hub_streams(uint64_t base_seed, uint64_t hub_id,
                              uint32_t *key0, uint32_t *key1, uint32_t *e_state) {
    uint64_t s = chub_splitmix64(base_seed ^ (hub_id * 0xD1342543DE82EF95ull));
    uint64_t r = chub_splitmix64(s);
    *key0 = (uint32_t) s;
    *key1 = (uint32_t) (s >> 32);
    *e_state = 1u + (uint32_t) (r % 2147483646ull);
}

#endif /* CHUB_RNG_H */
