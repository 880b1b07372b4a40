/* exo_oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle/Makefile).
 *
 * Scalar CPU restatement of the reference's exogenous inputs for ONE hub at a time: ReNew day choice and
 * PV / wind noise (renewable.py:21-53), the OU_Noise process (renewable.py:56-76) and the held price noise of
 * make_state (evcssp_manager.py:344-361), in the reference's call order (reset: renew_reset, make_state(0),
 * price_count = 0, evcssp_manager.py:304-316).  The random numbers come from the per-hub X stream of
 * include/chub_exo.h instead of Python's `random` / numpy's global generator; everything else is the
 * reference's arithmetic, statement by statement.  tests/test_exo_oracle.py pins this file against
 * charginghub_env_b200/exogenous.py (itself pinned on the reference run live, tests/test_exo_live_reference.py) on
 * identical variates.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/chub_exo.h"

typedef struct {
    uint32_t k0, k1, ctr;
    int pv_day, wd_day, price_count;
    double ou_pv, ou_wd, ou_price, price_hold; /* OU states persist over resets (renewable.py:34-35) */
} exo_hub;

typedef struct {
    int H, n_pv, n_wd;
    double *pv_days, *wd_days; /* [n][96] */
    double price[96];
    double renew_fluctuate, price_fluctuate;
    ChubLogEntry log_table[CHUB_LOG_N];
    exo_hub *hub;
} exo_engine;

exo_engine *exo_create(int H, const double *pv_days, int n_pv, const double *wd_days, int n_wd, const double *price96,
                       double renew_fluctuate, double price_fluctuate) {
    exo_engine *X = (exo_engine *) calloc(1, sizeof(exo_engine));
    X->H = H; X->n_pv = n_pv; X->n_wd = n_wd;
    X->pv_days = (double *) malloc(sizeof(double) * 96 * n_pv);
    X->wd_days = (double *) malloc(sizeof(double) * 96 * n_wd);
    memcpy(X->pv_days, pv_days, sizeof(double) * 96 * n_pv);
    memcpy(X->wd_days, wd_days, sizeof(double) * 96 * n_wd);
    memcpy(X->price, price96, sizeof(double) * 96);
    X->renew_fluctuate = renew_fluctuate; X->price_fluctuate = price_fluctuate;
    chub_build_log_table(X->log_table);
    X->hub = (exo_hub *) calloc((size_t) H, sizeof(exo_hub));
    return X;
}
void exo_destroy(exo_engine *X) {
    if (!X) return;
    free(X->pv_days); free(X->wd_days); free(X->hub); free(X);
}
void exo_seed_base(exo_engine *X, uint64_t base_seed, uint64_t hub_id_base) {
    for (int h = 0; h < X->H; ++h) {
        uint32_t e;
        exo_hub *g = &X->hub[h];
        memset(g, 0, sizeof *g);
        chub_hub_streams(base_seed, hub_id_base + (uint64_t) h, &g->k0, &g->k1, &e);
    }
}

static double normal(exo_engine *X, exo_hub *g) { return chub_xnormal(g->k0, g->k1, &g->ctr, X->log_table); }

/* OU_Noise.sample, renewable.py:71-76: dx = theta * (mu - state) + sigma * N(0,1); state += dx */
static double ou_sample(exo_engine *X, exo_hub *g, double *state, double theta, double sigma) {
    double dx = theta * (0.0 - *state) + sigma * normal(X, g);
    *state = *state + dx;
    return *state;
}

/* make_state's exogenous part, evcssp_manager.py:344-361, with ReNew.get_pv_power / get_wd_power inlined */
static void values(exo_engine *X, exo_hub *g, int time, double last_table_price, double *pv, double *wd, double *price) {
    double temp = X->pv_days[g->pv_day * 96 + time];
    if (temp > 0 && g->pv_day % 2 == 0) temp += ou_sample(X, g, &g->ou_pv, 0.01, 1.0) * (1 + X->renew_fluctuate);
    *pv = (temp > 0 ? temp : 0.0) * 5; /* max(0, temp) * scale_pv */
    temp = X->wd_days[g->wd_day * 96 + time];
    temp += ou_sample(X, g, &g->ou_wd, 0.01, 1.5) * (1 + X->renew_fluctuate);
    *wd = (temp > 0 ? temp : 0.0) * 1; /* max(0, temp) * scale_wd */
    if (g->price_count % 4 == 0) g->price_hold = ou_sample(X, g, &g->ou_price, 0.1, 0.005) * (1 + X->price_fluctuate);
    *price = last_table_price + g->price_hold;
    g->price_count += 1;
}

/* reset(): renew_reset (random.randint(0, n - 1) twice), make_state(0), price_count = 0 */
void exo_reset(exo_engine *X, const uint8_t *mask_or_null, double *pv, double *wd, double *price) {
    for (int h = 0; h < X->H; ++h) {
        if (mask_or_null && !mask_or_null[h]) continue;
        exo_hub *g = &X->hub[h];
        g->pv_day = (int) (chub_xrand31(g->k0, g->k1, g->ctr) % (uint32_t) X->n_pv);
        g->wd_day = (int) (chub_xrand31(g->k0, g->k1, g->ctr + 1u) % (uint32_t) X->n_wd);
        g->ctr += 2u;
        values(X, g, 0, X->price[95], &pv[h], &wd[h], &price[h]);
        g->price_count = 0;
    }
}
/* the make_state(time_next) at the end of step(): the price list's last entry is table[time_prev]
 * (Aggregator_Simple.py:147) */
void exo_step(exo_engine *X, const int32_t *time_prev, double *pv, double *wd, double *price) {
    for (int h = 0; h < X->H; ++h) {
        const int tp = time_prev[h], tn = (tp + 1) % 96;
        values(X, &X->hub[h], tn, X->price[tp], &pv[h], &wd[h], &price[h]);
    }
}
void exo_get_days(exo_engine *X, int32_t *pv_day, int32_t *wd_day, uint32_t *ctr) {
    for (int h = 0; h < X->H; ++h) {
        pv_day[h] = X->hub[h].pv_day; wd_day[h] = X->hub[h].wd_day; ctr[h] = X->hub[h].ctr;
    }
}
/* n standard normals of one stream, for pinning against the Python generator */
void exo_normals(uint32_t k0, uint32_t k1, uint32_t ctr0, int n, double *out, uint32_t *ctr_after) {
    ChubLogEntry tab[CHUB_LOG_N];
    chub_build_log_table(tab);
    uint32_t c = ctr0;
    for (int i = 0; i < n; ++i) out[i] = chub_xnormal(k0, k1, &c, tab);
    *ctr_after = c;
}
uint32_t exo_rand31(uint32_t k0, uint32_t k1, uint32_t ctr) { return chub_xrand31(k0, k1, ctr); }
