/* abi_driver.c -- a plain C99 caller of include/chub.h (no Python, no torch, no CUDA headers): what a C or C++
 * host of the reference's step/reset path would write.  Reads the static tables, traces and action rows from a
 * binary file written by tests/test_c_driver.py, runs reset + T host-buffer steps, and writes obs / reward / done of
 * every step to the output file.  Usage: abi_driver in.bin out.bin
 * in.bin: int32 {H, T, L, n0, n1}, double cdf[96*301], price[96], price_mean, price_std, sin[96], pv[L], wd[L], pr[L],
 *         uint64 seed, float actions[T][H][n0+n1+2]. */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "chub.h"

#define CHECK(call)                                                                  \
    do {                                                                             \
        int rc_ = (call);                                                            \
        if (rc_ != CHUB_OK) {                                                        \
            fprintf(stderr, "%s -> %d: %s\n", #call, rc_, chub_last_error());        \
            return 2;                                                                \
        }                                                                            \
    } while (0)

static int rd(FILE *f, void *p, size_t n) { return fread(p, 1, n, f) == n ? 0 : -1; }

int main(int argc, char **argv) {
    if (argc != 3) return 1;
    FILE *f = fopen(argv[1], "rb");
    if (!f) return 1;
    int32_t hd[5];
    if (rd(f, hd, sizeof hd)) return 1;
    const int H = hd[0], T = hd[1], L = hd[2], n0 = hd[3], n1 = hd[4], A = n0 + n1 + 2;
    double *cdf = malloc(sizeof(double) * 96 * 301), price[96], pm, ps, sn[96];
    double *pv = malloc(sizeof(double) * L), *wd = malloc(sizeof(double) * L), *pr = malloc(sizeof(double) * L);
    uint64_t seed;
    if (rd(f, cdf, sizeof(double) * 96 * 301) || rd(f, price, sizeof price) || rd(f, &pm, 8) || rd(f, &ps, 8) ||
        rd(f, sn, sizeof sn) || rd(f, pv, 8 * (size_t) L) || rd(f, wd, 8 * (size_t) L) || rd(f, pr, 8 * (size_t) L) ||
        rd(f, &seed, 8))
        return 1;

    chub_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.n_hubs = H; cfg.device = 0; cfg.hub_id_base = 0; cfg.n_classes = 1; cfg.action_stride = A; cfg.auto_reset = 1;
    chub_engine *e = NULL;
    CHECK(chub_create(&cfg, &e));
    chub_hub_class c;   /* the kwargs of the reference's test/env_test.py:10-30 */
    memset(&c, 0, sizeof c);
    c.n_piles[0] = n0; c.n_piles[1] = n1; c.station_type[0] = CHUB_FAST; c.station_type[1] = CHUB_SLOW;
    c.hydro_prod_rate = 100; c.hydro_store_vlt = 25.0; c.init_soc = 0.2; c.fc_max_power = 100; c.fcev_permeate = 0.01;
    CHECK(chub_set_hub_classes(e, &c, 1, NULL));
    chub_tables tb;
    tb.arrival_cdf = cdf; tb.price = price; tb.price_mean = pm; tb.price_std = ps; tb.time_sin = sn;
    CHECK(chub_set_tables(e, &tb));
    CHECK(chub_set_traces(e, pv, wd, pr, L, 0));
    CHECK(chub_seed(e, seed));

    float *act, *obs, *rew;
    uint8_t *done;
    CHECK(chub_alloc_host((int64_t) sizeof(float) * H * A, (void **) &act));
    CHECK(chub_alloc_host((int64_t) sizeof(float) * H * CHUB_OBS_DIM, (void **) &obs));
    CHECK(chub_alloc_host((int64_t) sizeof(float) * H, (void **) &rew));
    CHECK(chub_alloc_host((int64_t) H, (void **) &done));
    FILE *o = fopen(argv[2], "wb");
    if (!o) return 1;
    CHECK(chub_reset_host(e, obs, NULL));
    fwrite(obs, sizeof(float), (size_t) H * CHUB_OBS_DIM, o);
    for (int t = 0; t < T; ++t) {
        if (rd(f, act, sizeof(float) * (size_t) H * A)) return 1;
        CHECK(chub_step_host(e, act, obs, rew, done, NULL));
        fwrite(obs, sizeof(float), (size_t) H * CHUB_OBS_DIM, o);
        fwrite(rew, sizeof(float), (size_t) H, o);
        fwrite(done, 1, (size_t) H, o);
    }
    int64_t step_index = 0, launches = 0;
    uint32_t err = 0;
    CHECK(chub_counters(e, &step_index, &launches, &err));
    double rep[CHUB_REPORT_LEN];
    CHECK(chub_hub_report(e, H - 1, rep, NULL));
    fwrite(rep, sizeof(double), CHUB_REPORT_LEN, o);
    fclose(o);
    fclose(f);
    printf("steps %lld launches %lld error_bits %u\n", (long long) step_index, (long long) launches, err);
    chub_free_host(act); chub_free_host(obs); chub_free_host(rew); chub_free_host(done);
    CHECK(chub_destroy(e));
    free(cdf); free(pv); free(wd); free(pr);
    return 0;
}
