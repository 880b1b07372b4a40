// chub_engine.cu -- host side of libchub.so: the C ABI of include/chub.h over the sm_100a kernels.
//
// Owns all device memory of one engine (one shard of hubs on one GPU), derives the per-class
// constants and lookup tables on the host with this machine's libm (so table entries equal what
// the reference computes at run time), launches the step / reset kernels stream-ordered, and
// exposes outputs as raw device pointers and DLPack tensors.  No CPU stepping path exists here:
// if CUDA is unavailable every entry point fails with CHUB_ERR_CUDA.
#include <dlfcn.h>
#include <nvtx3/nvToolsExt.h>

#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "chub_hostpack.h"
#include "chub_kernels.cuh"

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) {
    g_err = msg;
    return code;
}
#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(CHUB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));        \
    } while (0)

#define CHUB_HOST_INLINE_PACK 4096   // up to this many hubs the calling thread packs the rows itself: waking a pool costs more
#define CHUB_HOST_KSTREAMS 4   // ranges are independent: their kernels may share the SMs
#define CHUB_TIMER_RING 128
struct chub_engine {
    chub_config cfg;
    int H = 0, Hpad = 0, S0 = 0, S1 = 0;
    int refcount = 1;  // engine handle + live DLPack views
    bool destroyed = false;
    std::vector<ChubClass> classes;
    std::vector<double> el_tables;
    std::vector<int32_t> hub_class;
    ChubParams P;
    // device
    uint4 *pile = nullptr; float *arrive = nullptr; uint4 *hubA = nullptr; double *store = nullptr;
    uint2 *keys = nullptr, *occ = nullptr; int *cls = nullptr; double2 *stats = nullptr; float4 *last_stats = nullptr;
    uint32_t *err = nullptr; float *qsoc = nullptr; ChubClass *d_classes = nullptr; double *d_el = nullptr;
    uint16_t *lut = nullptr; float *s2t_lut = nullptr; double *pv = nullptr, *wd = nullptr, *price_tr = nullptr;
    float *obs = nullptr, *reward = nullptr; uint8_t *done = nullptr; double *diag = nullptr; float *st_power = nullptr;
    long long *step_index = nullptr;
    unsigned long long *queue = nullptr;   // [0] ticket counter, [1] its value at the start of the epoch (chub_step_kernel)
    unsigned char *lpt = nullptr;          // the step kernel's cost-binned unit lists
    int lpt_cap = 0;
    int num_sms = 0;
    float *actions_stage = nullptr;  // device copy target of chub_step_host when the rows travel unpacked (host_threads == 0)
    uint4 *packed_stage = nullptr;   // device copy target of chub_step_host: one packed record per hub
    uint4 *packed_host = nullptr;    // pinned: where the pool packs the caller's rows
    ChubPackPool *pool = nullptr;
    int host_threads = -1;           // -1: not chosen yet (CHUB_HOST_THREADS or the machine's core count); 0: unpacked rows
    int host_ranges = 0;             // hub ranges of the chub_step_host pipeline
    std::vector<int32_t> na_row;     // per-hub number of pile actions (only with more than one class)
    double *report_dev = nullptr, *report_host = nullptr;   // chub_hub_report: device gather target, pinned landing area
    int exo_mode = 0, n_pv_days = 0, n_wd_days = 0;
    double exo_renew_gain = 1.0, exo_price_gain = 1.0;
    double *exo_d = nullptr, *pv_days = nullptr, *wd_days = nullptr;
    uint2 *exo_i = nullptr;
    float *final_soc = nullptr;
    unsigned long long *act_bits = nullptr;  // pile on-bits of the scalar-load step (chub_step_loads)
    unsigned long long *static_bits = nullptr;   // the same in the library's file-scope array (chub_step_loads_static); not owned
    cudaError_t static_bits_err = cudaSuccess;
    cudaStream_t host_streams[2 + CHUB_HOST_KSTREAMS] = {};  // chub_step_host pipeline: copy in, copy out, step kernels (round robin)
    cudaEvent_t host_events[2 * CHUB_PACK_MAX_RANGES + 2] = {};
    int trace_len = 0, trace_per_hub = 0;
    bool have_tables = false, have_classes = false, have_traces = false, have_seed = false;
    long long launches = 0;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    int graph_steps = 0;
    // the one collective of the path (chub_gather_stats / chub_reduce_stats): snapshot + communication stream
    float4 *stats_stage = nullptr;
    double *stats_partial = nullptr, *stats_sums = nullptr;
    cudaStream_t comm_stream = nullptr;
    cudaEvent_t comm_ready = nullptr, comm_done = nullptr;
    bool comm_pending = false;
    // per-launch timers (chub_set_timing / chub_step_times): CUDA events around the most recent step launches
    bool timing = false;
    cudaEvent_t time_ev[2 * CHUB_TIMER_RING] = {};
    long long timed = 0;
};
struct NvtxRange {   // shows up in nsys / ncu --nvtx; costs a pointer test when no tool is attached
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};
static void drop_graph(chub_engine *e) {   // a captured rollout bakes device pointers and launch parameters
    if (e->graph_exec) { cudaGraphExecDestroy(e->graph_exec); e->graph_exec = nullptr; }
    if (e->graph) { cudaGraphDestroy(e->graph); e->graph = nullptr; }
    e->graph_steps = 0;
}

extern "C" const char *chub_last_error(void) { return g_err.c_str(); }
extern "C" int chub_version(void) { return 100; }

static void engine_free(chub_engine *e) {
    cudaSetDevice(e->cfg.device);
    if (e->pool) chub_pack_pool_destroy(e->pool);
    if (e->packed_host) cudaFreeHost(e->packed_host);
    if (e->report_host) cudaFreeHost(e->report_host);
    if (e->graph_exec) cudaGraphExecDestroy(e->graph_exec);
    if (e->graph) cudaGraphDestroy(e->graph);
    for (cudaStream_t st : e->host_streams)
        if (st) cudaStreamDestroy(st);
    for (cudaEvent_t ev : e->host_events)
        if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : e->time_ev)
        if (ev) cudaEventDestroy(ev);
    if (e->comm_stream) cudaStreamDestroy(e->comm_stream);
    if (e->comm_ready) cudaEventDestroy(e->comm_ready);
    if (e->comm_done) cudaEventDestroy(e->comm_done);
    void *ptrs[] = {e->pile, e->arrive, e->hubA, e->store, e->keys, e->occ, e->cls, e->stats, e->last_stats, e->err, e->qsoc,
                    e->d_classes, e->d_el, e->lut, e->s2t_lut, e->pv, e->wd, e->price_tr, e->obs, e->reward, e->done, e->diag,
                    e->st_power, e->step_index, e->queue, e->lpt, e->actions_stage, e->exo_d, e->exo_i, e->pv_days, e->wd_days, e->act_bits, e->final_soc, e->packed_stage, e->report_dev,
                    e->stats_stage, e->stats_partial, e->stats_sums};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    delete e;
}
static void engine_unref(chub_engine *e) {
    if (--e->refcount == 0) engine_free(e);
}

template <class T>
static cudaError_t dalloc(T **p, size_t n) {
    cudaError_t r = cudaMalloc((void **) p, n * sizeof(T));
    if (r == cudaSuccess) r = cudaMemset(*p, 0, n * sizeof(T));
    return r;
}

extern "C" int chub_create(const chub_config *cfg, chub_engine **out) {
    if (!cfg || !out) return fail(CHUB_ERR_ARG, "chub_create: null argument");
    if (cfg->n_hubs <= 0) return fail(CHUB_ERR_ARG, "chub_create: n_hubs must be positive");
    if (cfg->action_stride < 3) return fail(CHUB_ERR_ARG, "chub_create: action_stride too small");
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return fail(CHUB_ERR_CUDA, "chub_create: no CUDA device (this engine has no CPU path)");
    if (cfg->device < 0 || cfg->device >= ndev) return fail(CHUB_ERR_ARG, "chub_create: bad device ordinal");
    CK(cudaSetDevice(cfg->device));
    chub_engine *e = new chub_engine();
    e->cfg = *cfg;
    e->H = cfg->n_hubs;
    e->Hpad = (cfg->n_hubs + 31) / 32 * 32;
    cudaDeviceGetAttribute(&e->num_sms, cudaDevAttrMultiProcessorCount, cfg->device);
    if (e->num_sms <= 0) e->num_sms = 148;
    // set-up, this device current: the device address of chub_step_loads_static's scratch (asked once per device; a
    // failure only disables that entry point)
    e->static_bits_err = chub_static_act_bits(cfg->device, &e->static_bits);
    const size_t Hp = (size_t) e->Hpad;
    cudaError_t r = cudaSuccess;
    auto A = [&](cudaError_t x) { if (r == cudaSuccess) r = x; };
    A(dalloc(&e->hubA, Hp)); A(dalloc(&e->store, Hp)); A(dalloc(&e->keys, Hp)); A(dalloc(&e->occ, Hp));
    A(dalloc(&e->stats, 2 * Hp)); A(dalloc(&e->last_stats, Hp)); A(dalloc(&e->err, Hp)); A(dalloc(&e->final_soc, Hp));
    A(dalloc(&e->qsoc, Hp * CHUB_FCEV_QCAP)); A(dalloc(&e->lut, (size_t) 96 * 1000)); A(dalloc(&e->s2t_lut, (size_t) 4000));
    A(dalloc(&e->obs, Hp * CHUB_OBS_DIM)); A(dalloc(&e->reward, Hp)); A(dalloc(&e->done, Hp));
    A(dalloc(&e->step_index, 1)); A(dalloc(&e->queue, 2));
    e->lpt_cap = (e->Hpad / 32 + 3) / 4;   // units of >= 4 tiles (the smallest launch shape)
    A(dalloc(&e->lpt, CHUB_LPT_BYTES(e->lpt_cap)));
    if (cfg->diagnostics) { A(dalloc(&e->diag, Hp * CHUB_D_COUNT)); A(dalloc(&e->st_power, Hp * 6)); }
    if (r != cudaSuccess) {
        engine_free(e);
        return fail(CHUB_ERR_CUDA, std::string("chub_create: allocation failed: ") + cudaGetErrorString(r));
    }
    *out = e;
    return CHUB_OK;
}

extern "C" int chub_destroy(chub_engine *e) {
    if (!e) return CHUB_OK;
    if (e->destroyed) return fail(CHUB_ERR_STATE, "chub_destroy: already destroyed");
    e->destroyed = true;
    cudaSetDevice(e->cfg.device);
    cudaDeviceSynchronize();
    engine_unref(e);
    return CHUB_OK;
}

// ---- per-class constants (hydro_sys.py:9-36, 57-72, 91-102, 133-158; CHS.hpp:1153-1154, 1463-1464) ----
static double host_el_power(const ChubClass &c, double m) {
    if (c.el_cells == 0.0) return 0.0;
    double per_cell = m / c.el_cells;
    double mol = per_cell / 2.02;
    double litres = mol * c.v_M;
    double ml_min = litres * 1000 * 60;
    double cur = ml_min * 2 * 96487 / c.el_den;
    double w = std::pow(cur, 2) * 0.326 + cur * 1.476;
    return c.el_cells * w / 1000;
}
static void derive_class(const chub_hub_class &in, ChubClass &c, double *table) {
    memset(&c, 0, sizeof c);
    for (int k = 0; k < 2; ++k) {
        c.n[k] = in.n_piles[k];
        c.type[k] = in.station_type[k];
        float unit = (in.station_type[k] == CHUB_SLOW) ? (float) 5.254973139368931 : (float) 36.44764034125146;
        c.transformer[k] = unit * (float) in.n_piles[k];
        c.half[k] = (double) c.transformer[k] / 2;
    }
    c.constant = in.constant_charging ? 1 : 0;
    c.perm_f = (float) in.fcev_permeate;
    c.possible_in_f = (float) 0.3;
    const double prod = in.hydro_prod_rate > 0 ? in.hydro_prod_rate : 430.0;
    c.v_h_max = 0.089 * prod * 1000 / 3600;
    c.v_M = 0.082 * (273 + 25) / 1;
    c.el_den = c.v_M * 1000 * 60;
    {   // one reference cell produces 10 mL/min (hydro_sys.py:22-24)
        double litres_s = (10.0 / 1000) / 60;
        double grams_s = (litres_s / c.v_M) * 2.02;
        c.el_cells = std::ceil(c.v_h_max / grams_s);
    }
    {   // two-stage adiabatic compression 1 -> sqrt(200) -> 200 bar (hydro_sys.py:57-77)
        const double alpha = 1.4, R = 0.082, T = 273 + 25;
        const double p_mid = std::sqrt(1.0 * 200.0);
        const double k1 = alpha / (alpha - 1), k2 = k1 * R * T, k3 = (alpha - 1) / alpha;
        c.cpr_w12 = k2 * (-1 + std::pow(p_mid / 1.0, k3)) + k2 * (-1 + std::pow(200.0 / p_mid, k3));
    }
    const double vol = in.hydro_store_vlt > 0 ? in.hydro_store_vlt : 5000.0;
    c.cap_mass = (0.089 * (200 / 1)) * (vol * 1000);
    c.cap_floor = c.cap_mass * 0.1;
    c.floor_01 = 0.1 * c.cap_mass;
    c.init_soc = in.init_soc;
    c.init_mass = in.init_soc * c.cap_mass;
    c.loss = in.hydro_loss;
    c.fc_cells = in.fc_max_power < 0 ? 100.0 : in.fc_max_power;
    for (int k = 0; k < 2; ++k) c.r_half[k] = c.half[k] != 0.0 ? 1.0 / c.half[k] : 0.0;
    c.r_el_cells = c.el_cells != 0.0 ? 1.0 / c.el_cells : 0.0;
    c.r_el_den = 1.0 / c.el_den;
    c.r_cap_mass = 1.0 / c.cap_mass;
    if (in.el_table) {
        memcpy(table, in.el_table, sizeof(double) * CHUB_EL_TABLE);
    } else {  // hy_step(0.01*i) on an unclipped store (hydro_sys.py:153-157, 172-186)
        for (int i = 0; i < 101; ++i) {
            double flow = (0.01 * i * c.v_h_max * (15 * 60)) / (15 * 60);
            if (c.v_h_max < flow) flow = c.v_h_max;
            double n = flow / 2.02;
            table[i] = host_el_power(c, flow) + (n * c.cpr_w12 / 0.8) / 1000;
        }
        table[101] = table[100];
    }
}

extern "C" int chub_set_hub_classes(chub_engine *e, const chub_hub_class *classes, int32_t n_classes,
                                    const int32_t *hub_class) {
    if (!e || !classes || n_classes < 1) return fail(CHUB_ERR_ARG, "chub_set_hub_classes: bad argument");
    if (n_classes > 1 && !hub_class) return fail(CHUB_ERR_ARG, "chub_set_hub_classes: hub_class required");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    drop_graph(e);   // the pile arrays are reallocated below
    int S0 = 0, S1 = 0;
    e->classes.assign(n_classes, ChubClass());
    e->el_tables.assign((size_t) n_classes * CHUB_EL_TABLE, 0.0);
    for (int k = 0; k < n_classes; ++k) {
        const chub_hub_class &in = classes[k];
        for (int s = 0; s < 2; ++s) {
            if (in.n_piles[s] < 0 || in.n_piles[s] > CHUB_MAX_PILES)
                return fail(CHUB_ERR_ARG, "chub_set_hub_classes: piles per station must be in [0, 32]");
            if (in.station_type[s] != CHUB_FAST && in.station_type[s] != CHUB_SLOW)
                return fail(CHUB_ERR_ARG, "chub_set_hub_classes: station type must be CHUB_FAST or CHUB_SLOW");
        }
        if (in.n_piles[0] + in.n_piles[1] < 1)
            return fail(CHUB_ERR_ARG, "chub_set_hub_classes: a hub needs at least one pile");  // evcssp_manager.py:336
        if (in.n_piles[0] + in.n_piles[1] + 2 > e->cfg.action_stride)
            return fail(CHUB_ERR_ARG, "chub_set_hub_classes: action_stride smaller than n0+n1+2");
        if (!(in.init_soc >= 0.1 && in.init_soc <= 1.0))
            return fail(CHUB_ERR_ARG, "chub_set_hub_classes: init_soc must be in [0.1, 1]");  // hydro_sys.py:137
        derive_class(in, e->classes[k], &e->el_tables[(size_t) k * CHUB_EL_TABLE]);
        S0 = std::max(S0, in.n_piles[0]);
        S1 = std::max(S1, in.n_piles[1]);
    }
    if (hub_class)
        for (int h = 0; h < e->H; ++h)
            if (hub_class[h] < 0 || hub_class[h] >= n_classes)
                return fail(CHUB_ERR_ARG, "chub_set_hub_classes: class index out of range");
    if (e->pile) { cudaFree(e->pile); e->pile = nullptr; }
    if (e->arrive) { cudaFree(e->arrive); e->arrive = nullptr; }
    if (e->d_classes) { cudaFree(e->d_classes); e->d_classes = nullptr; }
    if (e->d_el) { cudaFree(e->d_el); e->d_el = nullptr; }
    if (e->cls) { cudaFree(e->cls); e->cls = nullptr; }
    e->S0 = S0; e->S1 = S1;
    const size_t slots = (size_t) (S0 + S1) * e->Hpad;
    CK(dalloc(&e->pile, slots));
    CK(dalloc(&e->arrive, slots));
    CK(dalloc(&e->d_classes, (size_t) n_classes));
    CK(dalloc(&e->d_el, (size_t) n_classes * CHUB_EL_TABLE));
    CK(cudaMemcpy(e->d_classes, e->classes.data(), sizeof(ChubClass) * n_classes, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(e->d_el, e->el_tables.data(), sizeof(double) * e->el_tables.size(), cudaMemcpyHostToDevice));
    e->hub_class.assign(e->H, 0);
    if (n_classes > 1) {
        e->hub_class.assign(hub_class, hub_class + e->H);
        CK(dalloc(&e->cls, (size_t) e->Hpad));
        CK(cudaMemcpy(e->cls, hub_class, sizeof(int32_t) * e->H, cudaMemcpyHostToDevice));
    }
    e->na_row.clear();
    if (n_classes > 1) {
        e->na_row.resize(e->H);
        for (int h = 0; h < e->H; ++h) e->na_row[h] = classes[hub_class[h]].n_piles[0] + classes[hub_class[h]].n_piles[1];
    }
    e->cfg.n_classes = n_classes;
    e->have_classes = true;
    return CHUB_OK;
}

extern "C" int chub_get_class_table(chub_engine *e, int32_t k, double *out) {
    if (!e || !out || !e->have_classes || k < 0 || k >= (int) e->classes.size())
        return fail(CHUB_ERR_ARG, "chub_get_class_table: bad argument");
    memcpy(out, &e->el_tables[(size_t) k * CHUB_EL_TABLE], sizeof(double) * CHUB_EL_TABLE);
    return CHUB_OK;
}

extern "C" int64_t chub_algorithmic_bytes(chub_engine *e, int32_t k) {
    if (!e || !e->have_classes || k < 0 || k >= (int) e->classes.size()) return -1;
    const int64_t n = e->classes[k].n[0] + e->classes[k].n[1];
    return 2 * (16 * n + 80) + 4 * (n + 2) + 57;
}

extern "C" int32_t chub_unit_hubs(chub_engine *e) {
    if (!e || !e->have_classes) return -1;
    ChubParams P;
    memset(&P, 0, sizeof P);   // the shape depends on the batch size, the pile counts and the action stride only
    P.H = e->H; P.Hpad = e->Hpad; P.S0 = e->S0; P.S1 = e->S1; P.h_begin = 0; P.h_end = e->H;
    P.action_stride = e->cfg.action_stride; P.num_sms = e->num_sms;
    const char *b = getenv("CHUB_BLOCK");
    P.block = b ? atoi(b) : 0;
    return (chub_step_shape(P) / 100) * 32;
}

// ---- static tables ----
static float h_fast_aa1(float x, float c1) { return (float) ((0.7194 / 0.053) * std::exp(0.053 * x) + 50.15 * x - c1); }
static double h_fast_anti(double x) {  // antiderivative of the quartic power fit, libm pow like the reference
    return (0.0002253 / 5) * std::pow(x, 5) - (0.03572 / 4) * std::pow(x, 4) + (2.016 / 3) * std::pow(x, 3) -
           (48.76 / 2) * std::pow(x, 2) + 457.7 * x;
}

extern "C" int chub_set_tables(chub_engine *e, const chub_tables *t) {
    if (!e || !t || !t->arrival_cdf || !t->price || !t->time_sin) return fail(CHUB_ERR_ARG, "chub_set_tables: null argument");
    CK(cudaSetDevice(e->cfg.device));
    ChubConst c;
    memset(&c, 0, sizeof c);
    for (int i = 0; i < 16; ++i) c.renege[i] = (float) (0.1 * logf((float) (i + 1)));
    for (int k = 0; k < 64; ++k) c.balk[k] = expf((float) -(0.01 * k));
    {   // integer thresholds: the largest r in 0..999 with uniform_rand(0,1) = r/999.f*(1-0)+0 <= table value (-1 if none)
        auto thr = [](float v) {
            int best = -1;
            for (int r = 0; r < 1000; ++r) {
                volatile float tr = (float) r / 999.0f;
                volatile float u = tr * 1.0f + 0.0f;
                if (u <= v) best = r;
            }
            return (short) best;
        };
        for (int i = 0; i < 16; ++i) c.renege_k[i] = thr(c.renege[i]);
        for (int k = 0; k < 64; ++k) c.balk_k[k] = thr(c.balk[k]);
    }
    memcpy(c.price, t->price, sizeof c.price);
    memcpy(c.time_sin, t->time_sin, sizeof c.time_sin);
    c.price_mean = t->price_mean;
    c.price_std = t->price_std;
    c.r_price_std = 1.0 / t->price_std;
    c.fast_aa1_c = (float) ((0.7194 / 0.053) * std::exp(0));
    c.fast_aa2_c = (float) (h_fast_aa1((float) 28.7, c.fast_aa1_c) - h_fast_anti(28.7));
    {
        float x = (float) 51.2;
        float full = (float) ((0.0002253 / 5) * std::pow(x, 5) - (0.03572 / 4) * std::pow(x, 4) +
                              (2.016 / 3) * std::pow(x, 3) - (48.76 / 2) * std::pow(x, 2) + 457.7 * x + c.fast_aa2_c);
        c.fast_scale = 100 / full;
    }
    {
        float x = (float) 3.67;
        c.slow_c2_end = (float) (-4.041 * x + 19.140115 * std::log(std::fabs(x - 0.485)) + 11.943306312699628);
    }
    c.slow_stt1_a = -4.276 * std::pow(10, -5);
    c.slow_stt2_a = 4.742 * std::pow(10, -6);
    const double jx[10] = {0.50, 5.00, 10.0, 15.0, 20.0, 30.0, 40.0, 50.0, 60.0, 70.0};
    const double jy[10] = {87.4, 81.0, 86.8, 86.1, 85.4, 83.8, 82.2, 80.4, 78.5, 76.1};
    for (int i = 0; i < 10; ++i) { c.j_x[i] = jx[i]; c.j_y[i] = jy[i]; }
    for (int a = 0; a < 8; ++a) c.j_slope[a] = (jy[2 + a] - jy[1 + a]) / (jx[2 + a] - jx[1 + a]);
    c.topoff_time = (87.4 - 69) / 7.2;
    c.fcev_cap = 6.3 * 1000;
    CK(chub_upload_const(c));
    // traffic-count LUT: for every possible U value r/999.f the first column whose CDF entry is >= it
    std::vector<uint16_t> lut((size_t) 96 * 1000);
    for (int tt = 0; tt < 96; ++tt) {
        const double *row = t->arrival_cdf + (size_t) tt * 301;
        for (int r = 0; r < 1000; ++r) {
            float u = (float) r / 999.0f;
            u = u * (1.0f - 0.0f) + 0.0f;
            int j = 0;
            while (j < 301 && !(row[j] >= (double) u)) ++j;
            lut[(size_t) tt * 1000 + r] = (uint16_t) (j < 301 ? j : 300);
        }
    }
    CK(cudaMemcpy(e->lut, lut.data(), lut.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
    chub_launch_s2t_lut(e->s2t_lut, nullptr);  // device-built so the step kernel and the table agree bit for bit
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    e->have_tables = true;
    return CHUB_OK;
}

extern "C" int chub_set_traces(chub_engine *e, const double *pv, const double *wd, const double *price, int32_t length,
                               int32_t per_hub) {
    if (!e || !pv || !wd || !price || length < 1) return fail(CHUB_ERR_ARG, "chub_set_traces: bad argument");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    drop_graph(e);   // a captured rollout holds the old trace pointers
    for (double **p : {&e->pv, &e->wd, &e->price_tr})
        if (*p) { cudaFree(*p); *p = nullptr; }
    const size_t row = per_hub ? (size_t) e->Hpad : 1;
    const size_t n = (size_t) length * row;
    const double *src[3] = {pv, wd, price};
    double **dst[3] = {&e->pv, &e->wd, &e->price_tr};
    for (int a = 0; a < 3; ++a) {
        CK(dalloc(dst[a], n));
        if (per_hub)
            CK(cudaMemcpy2D(*dst[a], row * sizeof(double), src[a], (size_t) e->H * sizeof(double),
                            (size_t) e->H * sizeof(double), length, cudaMemcpyHostToDevice));
        else
            CK(cudaMemcpy(*dst[a], src[a], n * sizeof(double), cudaMemcpyHostToDevice));
    }
    e->trace_len = length;
    e->trace_per_hub = per_hub ? 1 : 0;
    e->have_traces = true;
    e->exo_mode = 0;
    return CHUB_OK;
}

extern "C" int chub_set_exogenous_generator(chub_engine *e, const double *pv_days, int32_t n_pv, const double *wd_days,
                                            int32_t n_wd, double renew_fluctuate, double price_fluctuate) {
    if (!e || !pv_days || !wd_days || n_pv < 1 || n_wd < 1 || n_pv > 1023 || n_wd > 1023)
        return fail(CHUB_ERR_ARG, "chub_set_exogenous_generator: bad argument (1..1023 days per table)");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    drop_graph(e);
    for (double **p : {&e->pv_days, &e->wd_days, &e->exo_d})
        if (*p) { cudaFree(*p); *p = nullptr; }
    if (e->exo_i) { cudaFree(e->exo_i); e->exo_i = nullptr; }
    CK(dalloc(&e->pv_days, (size_t) n_pv * 96));
    CK(dalloc(&e->wd_days, (size_t) n_wd * 96));
    CK(cudaMemcpy(e->pv_days, pv_days, sizeof(double) * 96 * n_pv, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(e->wd_days, wd_days, sizeof(double) * 96 * n_wd, cudaMemcpyHostToDevice));
    CK(dalloc(&e->exo_d, (size_t) 7 * e->Hpad));   // zeroed: OU states start at their mean, like OU_Noise.reset
    CK(dalloc(&e->exo_i, (size_t) e->Hpad));
    e->n_pv_days = n_pv; e->n_wd_days = n_wd;
    e->exo_renew_gain = 1 + renew_fluctuate; e->exo_price_gain = 1 + price_fluctuate;
    e->exo_mode = 1;
    if (!e->have_traces) {   // the kernels take the trace pointers unconditionally: give them one slot
        const double z = 0.0;
        int rc = chub_set_traces(e, &z, &z, &z, 1, 0);
        if (rc) return rc;
        e->exo_mode = 1;
    }
    return CHUB_OK;
}

extern "C" int chub_poke_trace(chub_engine *e, int32_t slot, double pv, double wd, double price, void *stream) {
    if (!e || !e->have_traces || e->trace_per_hub || slot < 0 || slot >= e->trace_len)
        return fail(CHUB_ERR_ARG, "chub_poke_trace: needs shared traces and a slot inside them");
    CK(cudaSetDevice(e->cfg.device));
    cudaStream_t s = (cudaStream_t) stream;
    CK(cudaMemcpyAsync(e->pv + slot, &pv, sizeof(double), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(e->wd + slot, &wd, sizeof(double), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(e->price_tr + slot, &price, sizeof(double), cudaMemcpyHostToDevice, s));
    CK(cudaStreamSynchronize(s));  // the three doubles live on this stack frame
    return CHUB_OK;
}

static int upload_streams(chub_engine *e, const std::vector<uint2> &keys, const std::vector<uint4> &hubA) {
    CK(cudaMemcpy(e->keys, keys.data(), sizeof(uint2) * e->H, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(e->hubA, hubA.data(), sizeof(uint4) * e->H, cudaMemcpyHostToDevice));
    if (e->exo_d) {   // a new seed restarts the generator's streams and noise processes
        CK(cudaMemset(e->exo_d, 0, sizeof(double) * 7 * e->Hpad));
        CK(cudaMemset(e->exo_i, 0, sizeof(uint2) * e->Hpad));
    }
    e->have_seed = true;
    return CHUB_OK;
}

extern "C" int chub_seed(chub_engine *e, uint64_t base_seed) {
    if (!e) return fail(CHUB_ERR_ARG, "chub_seed: null engine");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    std::vector<uint2> keys(e->H);
    std::vector<uint4> hubA(e->H);
    for (int h = 0; h < e->H; ++h) {
        uint32_t k0, k1, e0;
        chub_hub_streams(base_seed, (uint64_t) e->cfg.hub_id_base + (uint64_t) h, &k0, &k1, &e0);
        keys[h] = make_uint2(k0, k1);
        hubA[h] = make_uint4(e0, 0u, 0u, 128u | (128u << 8));
    }
    return upload_streams(e, keys, hubA);
}

extern "C" int chub_seed_explicit(chub_engine *e, const uint32_t *key0, const uint32_t *key1, const uint32_t *uctr,
                                  const uint32_t *lcg) {
    if (!e || !key0 || !key1 || !lcg) return fail(CHUB_ERR_ARG, "chub_seed_explicit: null argument");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    std::vector<uint2> keys(e->H);
    std::vector<uint4> hubA(e->H);
    for (int h = 0; h < e->H; ++h) {
        if (lcg[h] < 1u || lcg[h] > 2147483646u) return fail(CHUB_ERR_ARG, "chub_seed_explicit: minstd state out of range");
        keys[h] = make_uint2(key0[h], key1[h]);
        hubA[h] = make_uint4(lcg[h], uctr ? uctr[h] : 0u, 0u, 128u | (128u << 8));
    }
    return upload_streams(e, keys, hubA);
}

// ---- launches ----
static int fill_params(chub_engine *e, ChubParams &P) {
    if (!e->have_classes) return fail(CHUB_ERR_STATE, "call chub_set_hub_classes first");
    if (!e->have_tables) return fail(CHUB_ERR_STATE, "call chub_set_tables first");
    if (!e->have_traces) return fail(CHUB_ERR_STATE, "call chub_set_traces first");
    if (!e->have_seed) return fail(CHUB_ERR_STATE, "call chub_seed first");
    memset(&P, 0, sizeof P);
    P.H = e->H; P.Hpad = e->Hpad; P.S0 = e->S0; P.S1 = e->S1;
    P.h_begin = 0; P.h_end = e->H;
    P.n_classes = (int) e->classes.size(); P.action_stride = e->cfg.action_stride;
    P.auto_reset = e->cfg.auto_reset; P.diagnostics = e->cfg.diagnostics;
    P.trace_len = e->trace_len; P.trace_per_hub = e->trace_per_hub;
    P.step_offset = 0; P.step_index = e->step_index; P.num_sms = e->num_sms;
    P.queue = getenv("CHUB_STATIC_GRID") ? nullptr : e->queue; P.queue_base = e->queue + 1; P.queue_stride = 0;
    P.lpt = e->lpt; P.lpt_cap = e->lpt_cap;
    {   // experiment knobs (defaults are the measured best)
        const char *b = getenv("CHUB_BLOCK"), *y = getenv("CHUB_PHASE_SYNC");
        P.block = b ? atoi(b) : 0;
        P.phase_sync = y ? atoi(y) : 0;   // experiment knob: bit 0 = hand units out in index order (ignore the cost lists)
    }
    P.pile = e->pile; P.arrive = e->arrive; P.hubA = e->hubA; P.store = e->store; P.keys = e->keys; P.occ = e->occ; P.cls = e->cls;
    P.stats = e->stats; P.last_stats = e->last_stats; P.err = e->err; P.qsoc = e->qsoc;
    P.classes = e->d_classes; P.el_table = e->d_el; P.lut = e->lut; P.s2t_lut = e->s2t_lut;
    P.pv = e->pv; P.wd = e->wd; P.price_tr = e->price_tr;
    P.exo_mode = e->exo_mode; P.n_pv_days = e->n_pv_days; P.n_wd_days = e->n_wd_days;
    P.exo_renew_gain = e->exo_renew_gain; P.exo_price_gain = e->exo_price_gain;
    P.exo_d = e->exo_d; P.exo_i = e->exo_i; P.pv_days = e->pv_days; P.wd_days = e->wd_days;
    P.actions = nullptr; P.obs = e->obs; P.reward = e->reward; P.done = e->done;
    P.diag = e->diag; P.st_power = e->st_power; P.reset_mask = nullptr; P.final_soc = e->final_soc;
    return CHUB_OK;
}

// One step launch (+ the tick that advances the trace index), with the optional per-launch timer around the step kernel.
static int run_step(chub_engine *e, const ChubParams &P, cudaStream_t s) {
    NvtxRange range("chub_step");
    int slot = -1;
    if (e->timing) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(s, &cs);
        if (cs == cudaStreamCaptureStatusNone) {   // events inside a capture carry no time
            slot = (int) (e->timed % CHUB_TIMER_RING);
            for (int k = 0; k < 2; ++k)
                if (!e->time_ev[2 * slot + k]) CK(cudaEventCreate(&e->time_ev[2 * slot + k]));
            CK(cudaEventRecord(e->time_ev[2 * slot], s));
        }
    }
    unsigned long long tickets = 0;
    bool ticked = false;
    ChubParams Q = P;
    Q.self_tick = 1;   // the kernel's last block advances the step index: one launch per step, no tick kernel behind it
    CK(chub_launch_step(Q, s, &tickets, &ticked));
    if (slot >= 0) {
        CK(cudaEventRecord(e->time_ev[2 * slot + 1], s));
        e->timed += 1;
    }
    if (!ticked) chub_launch_tick(e->step_index, 1, e->queue + 1, tickets, s);
    CK(cudaGetLastError());
    e->launches += 1;
    return CHUB_OK;
}

extern "C" int chub_reset(chub_engine *e, const uint8_t *mask_dev, int64_t step_index, void *stream) {
    if (!e) return fail(CHUB_ERR_ARG, "chub_reset: null engine");
    CK(cudaSetDevice(e->cfg.device));
    ChubParams P;
    int rc = fill_params(e, P);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t) stream;
    if (step_index >= 0) {
        long long v = step_index;
        CK(cudaMemcpyAsync(e->step_index, &v, sizeof v, cudaMemcpyHostToDevice, s));
        CK(cudaStreamSynchronize(s));  // v lives on this stack frame
    }
    P.reset_mask = mask_dev;
    NvtxRange range("chub_reset");
    chub_launch_reset(P, s);
    CK(cudaGetLastError());
    return CHUB_OK;
}

extern "C" int chub_step(chub_engine *e, const float *actions_dev, void *stream) {
    if (!e || !actions_dev) return fail(CHUB_ERR_ARG, "chub_step: null argument");
    CK(cudaSetDevice(e->cfg.device));
    ChubParams P;
    int rc = fill_params(e, P);
    if (rc) return rc;
    P.actions = actions_dev;
    cudaStream_t s = (cudaStream_t) stream;
    return run_step(e, P, s);
}

extern "C" int chub_step_into(chub_engine *e, const float *actions_dev, float *obs_out_dev, float *reward_out_dev,
                              uint8_t *done_out_dev, void *stream) {
    if (!e || !actions_dev) return fail(CHUB_ERR_ARG, "chub_step_into: null argument");
    CK(cudaSetDevice(e->cfg.device));
    ChubParams P;
    int rc = fill_params(e, P);
    if (rc) return rc;
    P.actions = actions_dev;
    if (obs_out_dev) P.obs = obs_out_dev;
    if (reward_out_dev) P.reward = reward_out_dev;
    if (done_out_dev) P.done = done_out_dev;
    cudaStream_t s = (cudaStream_t) stream;
    return run_step(e, P, s);
}

// Scalar-load step: rows [load0 kW, load1 kW, a[-2], a[-1]]; the dispatch kernel turns the two station power targets
// into pile on-bits (evs_step(float), CHS.hpp:1169-1186 / 1478-1495), then the ordinary step kernel runs on them.
extern "C" int chub_step_loads(chub_engine *e, const float *loads_dev, void *stream) {
    if (!e || !loads_dev) return fail(CHUB_ERR_ARG, "chub_step_loads: null argument");
    CK(cudaSetDevice(e->cfg.device));
    ChubParams P;
    int rc = fill_params(e, P);
    if (rc) return rc;
    if (!e->act_bits) {
        CK(cudaMalloc((void **) &e->act_bits, sizeof(unsigned long long) * (size_t) e->Hpad));
        CK(cudaMemset(e->act_bits, 0, sizeof(unsigned long long) * (size_t) e->Hpad));
    }
    P.actions = loads_dev;
    cudaStream_t s = (cudaStream_t) stream;
    chub_launch_dispatch(P, e->act_bits, s);
    P.act_bits = e->act_bits;
    return run_step(e, P, s);
}

// chub_step_loads without any allocation: the dispatch kernel's output goes to the library's file-scope device array
// (chub_noalloc.cu) instead of a buffer malloc'ed on first use.  Launches only -- no cudaMalloc / cudaFree, no event or
// stream creation (the per-launch timers of chub_set_timing are not applied here), nothing from an allocating library.
extern "C" int chub_step_loads_static(chub_engine *e, const float *loads_dev, float *obs_out_dev, float *reward_out_dev,
                                      uint8_t *done_out_dev, void *stream) {
    if (!e || !loads_dev) return fail(CHUB_ERR_ARG, "chub_step_loads_static: null argument");
    if (e->H > CHUB_STATIC_MAX_HUBS)
        return fail(CHUB_ERR_ARG, "chub_step_loads_static: the engine holds " + std::to_string(e->H) + " hubs, the static scratch "
                                  "CHUB_STATIC_MAX_HUBS = " + std::to_string(CHUB_STATIC_MAX_HUBS) + "; use chub_step_loads");
    if (e->static_bits_err != cudaSuccess || !e->static_bits)
        return fail(CHUB_ERR_CUDA, std::string("chub_step_loads_static: the static scratch has no device address on this device: ") +
                                   cudaGetErrorString(e->static_bits_err));
    CK(cudaSetDevice(e->cfg.device));
    ChubParams P;
    int rc = fill_params(e, P);
    if (rc) return rc;
    P.actions = loads_dev;
    if (obs_out_dev) P.obs = obs_out_dev;
    if (reward_out_dev) P.reward = reward_out_dev;
    if (done_out_dev) P.done = done_out_dev;
    cudaStream_t s = (cudaStream_t) stream;
    chub_launch_dispatch(P, e->static_bits, s);     // writes the word of every hub < n_hubs: nothing must start at zero
    P.act_bits = e->static_bits;
    unsigned long long tickets = 0;
    bool ticked = false;
    P.self_tick = 1;
    CK(chub_launch_step(P, s, &tickets, &ticked));
    if (!ticked) chub_launch_tick(e->step_index, 1, e->queue + 1, tickets, s);
    CK(cudaGetLastError());
    e->launches += 1;
    return CHUB_OK;
}

extern "C" int chub_situation_width(chub_engine *e, int32_t *s0, int32_t *s1) {
    if (!e || !e->have_classes) return fail(CHUB_ERR_STATE, "chub_situation_width: call chub_set_hub_classes first");
    if (s0) *s0 = e->S0;
    if (s1) *s1 = e->S1;
    return CHUB_OK;
}

extern "C" int chub_situation(chub_engine *e, float *out_dev, void *stream) {
    if (!e || !out_dev) return fail(CHUB_ERR_ARG, "chub_situation: null argument");
    CK(cudaSetDevice(e->cfg.device));
    ChubParams P;
    int rc = fill_params(e, P);
    if (rc) return rc;
    chub_launch_situation(P, out_dev, (cudaStream_t) stream);
    CK(cudaGetLastError());
    return CHUB_OK;
}

// Packed-action step: one 16-byte record per hub instead of a float row -- what a policy that already emits
// switch decisions (or the host pipeline below) hands over.
extern "C" int chub_step_packed(chub_engine *e, const void *packed_dev, void *stream) {
    if (!e || !packed_dev) return fail(CHUB_ERR_ARG, "chub_step_packed: null argument");
    if (reinterpret_cast<uintptr_t>(packed_dev) & 15u) return fail(CHUB_ERR_ARG, "chub_step_packed: records must be 16-byte aligned");
    CK(cudaSetDevice(e->cfg.device));
    ChubParams P;
    int rc = fill_params(e, P);
    if (rc) return rc;
    P.act_packed = (const uint4 *) packed_dev;
    cudaStream_t s = (cudaStream_t) stream;
    return run_step(e, P, s);
}

// Threads that pack the caller's rows inside chub_step_host; 0 sends the float rows over the bus as they are.
extern "C" int chub_set_host_threads(chub_engine *e, int32_t n_threads, int32_t n_ranges) {
    if (!e || n_threads < 0 || n_threads > 64) return fail(CHUB_ERR_ARG, "chub_set_host_threads: threads must be in [0, 64]");
    if (n_ranges < 0 || n_ranges > CHUB_PACK_MAX_RANGES) return fail(CHUB_ERR_ARG, "chub_set_host_threads: too many ranges");
    if (e->pool && chub_pack_pool_threads(e->pool) != n_threads) { chub_pack_pool_destroy(e->pool); e->pool = nullptr; }
    e->host_threads = n_threads;
    if (n_ranges) e->host_ranges = n_ranges;
    return CHUB_OK;
}

// cpulist ("0-15,32-47") of the NUMA node of a CUDA device, from sysfs; false when unknown or the host has one node
static bool gpu_local_cpulist(int device, char *out, size_t cap) {
    char bus[32] = {0};
    if (cudaDeviceGetPCIBusId(bus, sizeof bus, device) != cudaSuccess) return false;
    for (char *c = bus; *c; ++c) *c = (char) tolower(*c);
    char path[128];
    snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/numa_node", bus);
    FILE *f = fopen(path, "r");
    if (!f) return false;
    int node = -1;
    const int got = fscanf(f, "%d", &node);
    fclose(f);
    if (got != 1 || node < 0) return false;
    f = fopen("/sys/devices/system/node/node1/cpulist", "r");   // a second node exists?
    if (!f) return false;
    fclose(f);
    snprintf(path, sizeof path, "/sys/devices/system/node/node%d/cpulist", node);
    f = fopen(path, "r");
    if (!f) return false;
    const bool ok = fgets(out, (int) cap, f) != nullptr;
    fclose(f);
    return ok;
}

static int host_pipeline_setup(chub_engine *e) {
    if (e->host_threads < 0) {
        const char *t = getenv("CHUB_HOST_THREADS"), *r = getenv("CHUB_HOST_RANGES");
        int hw = (int) std::thread::hardware_concurrency();
        const char *lw = getenv("LOCAL_WORLD_SIZE");   // one process per GPU (torchrun): the ranks share the host's cores
        if (lw && atoi(lw) > 1) hw = std::max(2, hw / atoi(lw));
        e->host_threads = t ? atoi(t) : std::max(1, std::min(std::min(hw - 1, 16), e->H / 8192));  // small batches: one helper
        if (e->host_threads < 0 || e->host_threads > 64) e->host_threads = 1;
        if (r && atoi(r) >= 1 && atoi(r) <= CHUB_PACK_MAX_RANGES) e->host_ranges = atoi(r);
    }
    if (!e->host_streams[0]) {
        for (int i = 0; i < 2 + CHUB_HOST_KSTREAMS; ++i) CK(cudaStreamCreateWithFlags(&e->host_streams[i], cudaStreamNonBlocking));
        for (int i = 0; i < 2 * CHUB_PACK_MAX_RANGES + 2; ++i) CK(cudaEventCreateWithFlags(&e->host_events[i], cudaEventDisableTiming));
    }
    if (e->host_threads == 0) {
        if (!e->actions_stage) CK(dalloc(&e->actions_stage, (size_t) e->Hpad * (size_t) e->cfg.action_stride));
        return CHUB_OK;
    }
    if (!e->pool && e->H > CHUB_HOST_INLINE_PACK) {
        // the packing threads read the caller's rows and write the pinned record buffer: keep them on the cores of the
        // NUMA node the GPU hangs off (no-op on single-node hosts and where sysfs does not say)
        char cpus[512];
        e->pool = chub_pack_pool_create_on(e->host_threads, gpu_local_cpulist(e->cfg.device, cpus, sizeof cpus) ? cpus : nullptr);
    }
    if (!e->packed_stage) CK(dalloc(&e->packed_stage, (size_t) e->Hpad));
    if (!e->packed_host) CK(cudaHostAlloc((void **) &e->packed_host, sizeof(uint4) * (size_t) e->Hpad, cudaHostAllocDefault));
    return CHUB_OK;
}

// The packing alone (host rows -> host records), for callers that want to overlap it themselves and for the tests.
extern "C" int chub_pack_actions(chub_engine *e, const float *actions_host, void *packed_host) {
    if (!e || !actions_host || !packed_host || !e->have_classes) return fail(CHUB_ERR_ARG, "chub_pack_actions: bad argument");
    const ChubClass &c0 = e->classes[0];
    chub_pack_rows(actions_host, e->cfg.action_stride, e->na_row.empty() ? nullptr : e->na_row.data(), c0.n[0] + c0.n[1], e->H,
                   packed_host);
    return CHUB_OK;
}

// Host-buffer step.  The batch is cut into hub ranges that flow through a pipeline -- rows thresholded into packed
// records by the pool's threads (chub_hostpack.cpp), records host->device, step kernel, obs/reward/done device->host --
// so the host's memory reads, both PCIe directions and the kernel overlap; hubs are independent, so stepping them
// range by range is the same computation as one launch.
static int step_host_impl(chub_engine *e, const float *actions_host, const uint4 *records_host, float *obs_host,
                          float *reward_host, uint8_t *done_host, void *stream) {
    CK(cudaSetDevice(e->cfg.device));
    cudaStream_t s = (cudaStream_t) stream;
    ChubParams P;
    int rc = fill_params(e, P);
    if (rc) return rc;
    rc = host_pipeline_setup(e);
    if (rc) return rc;
    NvtxRange range("chub_step_host");
    const bool packed = records_host != nullptr || e->host_threads > 0;
    if (packed) P.act_packed = e->packed_stage;
    else P.actions = e->actions_stage;
    if (records_host && !e->packed_stage) CK(dalloc(&e->packed_stage, (size_t) e->Hpad));
    const int R = e->host_ranges ? e->host_ranges : (packed ? 16 : 8);   // whole rows over PCIe: fewer, larger copies
    cudaStream_t s_in = e->host_streams[0], s_out = e->host_streams[1], *s_k = e->host_streams + 2;
    cudaEvent_t *ev_in = e->host_events, *ev_k = e->host_events + CHUB_PACK_MAX_RANGES;
    cudaEvent_t ev_fork = e->host_events[2 * CHUB_PACK_MAX_RANGES], ev_join = e->host_events[2 * CHUB_PACK_MAX_RANGES + 1];
    const int A = e->cfg.action_stride;
    int per = (e->H + R - 1) / R;
    per = (per + 1023) / 1024 * 1024;           // whole warps (and whole blocks of any launch shape)
    const bool pooled = packed && !records_host && e->pool != nullptr;
    const uint4 *records = records_host ? records_host : e->packed_host;
    if (packed && !pooled && !records_host) {
        rc = chub_pack_actions(e, actions_host, e->packed_host);
        if (rc) return rc;
    }
    struct Drain {   // whatever path leaves this function, nobody reads the caller's rows afterwards
        ChubPackPool *p;
        ~Drain() { if (p) chub_pack_pool_drain(p); }
    } drain{nullptr};
    if (pooled) {                               // the workers start on range 0 while the streams are being forked
        const ChubClass &c0 = e->classes[0];
        if (chub_pack_pool_start(e->pool, actions_host, A, e->na_row.empty() ? nullptr : e->na_row.data(), c0.n[0] + c0.n[1],
                                 e->H, per, e->packed_host) < 0)
            return fail(CHUB_ERR_STATE, "chub_step_host: range count out of bounds");
        drain.p = e->pool;
    }
    CK(cudaEventRecord(ev_fork, s));            // everything queued on the caller's stream comes first
    CK(cudaStreamWaitEvent(s_in, ev_fork, 0));
    CK(cudaStreamWaitEvent(s_out, ev_fork, 0));
    for (int k = 0; k < CHUB_HOST_KSTREAMS; ++k) CK(cudaStreamWaitEvent(s_k[k], ev_fork, 0));
    int n_ranges = 0, out_begin = 0;
    const int out_every = std::max(1, ((e->H + per - 1) / per + 7) / 8);
    for (int c = 0, b = 0; b < e->H; ++c, b += per) {
        cudaStream_t sk = s_k[c % CHUB_HOST_KSTREAMS];
        n_ranges = c + 1;
        const int n = (b + per <= e->H ? per : e->H - b);
        if (packed) {
            if (pooled) chub_pack_pool_wait(e->pool, c);
            CK(cudaMemcpyAsync(e->packed_stage + b, records + b, sizeof(uint4) * (size_t) n, cudaMemcpyHostToDevice, s_in));
        } else {
            CK(cudaMemcpyAsync(e->actions_stage + (size_t) b * A, actions_host + (size_t) b * A, sizeof(float) * (size_t) n * A,
                               cudaMemcpyHostToDevice, s_in));
        }
        CK(cudaEventRecord(ev_in[c], s_in));
        CK(cudaStreamWaitEvent(sk, ev_in[c], 0));
        P.h_begin = b; P.h_end = b + n;
        P.queue = nullptr;                      // ranges run side by side on several streams: static block -> tile mapping
        CK(chub_launch_step(P, sk));
        CK(cudaEventRecord(ev_k[c], sk));
        CK(cudaStreamWaitEvent(s_out, ev_k[c], 0));
        // results go back in at most 8 pieces: below a few MB a copy costs its set-up, not its bytes
        // (the last two ranges go back on their own: whatever follows the last kernel is the tail of the step)
        if ((c + 1) % out_every != 0 && b + 2 * (size_t) per < (size_t) e->H) continue;
        const int ob = out_begin, on = b + n - out_begin;
        out_begin = b + n;
        if (obs_host) CK(cudaMemcpyAsync(obs_host + (size_t) ob * CHUB_OBS_DIM, e->obs + (size_t) ob * CHUB_OBS_DIM,
                                         sizeof(float) * (size_t) on * CHUB_OBS_DIM, cudaMemcpyDeviceToHost, s_out));
        if (reward_host) CK(cudaMemcpyAsync(reward_host + ob, e->reward + ob, sizeof(float) * (size_t) on, cudaMemcpyDeviceToHost, s_out));
        if (done_host) CK(cudaMemcpyAsync(done_host + ob, e->done + ob, (size_t) on, cudaMemcpyDeviceToHost, s_out));
    }
    for (int c = std::max(0, n_ranges - CHUB_HOST_KSTREAMS); c < n_ranges; ++c) CK(cudaStreamWaitEvent(s, ev_k[c], 0));
    CK(cudaEventRecord(ev_join, s_out));
    CK(cudaStreamWaitEvent(s, ev_join, 0));
    chub_launch_tick(e->step_index, 1, nullptr, 0, s);      // after the last range: every launch above read the same step index
    CK(cudaGetLastError());
    e->launches += 1;
    CK(cudaStreamSynchronize(s));
    return CHUB_OK;
}

extern "C" int chub_step_host(chub_engine *e, const float *actions_host, float *obs_host, float *reward_host,
                              uint8_t *done_host, void *stream) {
    if (!e || !actions_host) return fail(CHUB_ERR_ARG, "chub_step_host: null argument");
    return step_host_impl(e, actions_host, nullptr, obs_host, reward_host, done_host, stream);
}

// The same pipeline for a host producer that already holds switch decisions: 16-byte records in, nothing to threshold.
extern "C" int chub_step_host_packed(chub_engine *e, const void *records_host, float *obs_host, float *reward_host,
                                     uint8_t *done_host, void *stream) {
    if (!e || !records_host) return fail(CHUB_ERR_ARG, "chub_step_host_packed: null argument");
    if (reinterpret_cast<uintptr_t>(records_host) & 15u) return fail(CHUB_ERR_ARG, "chub_step_host_packed: records must be 16-byte aligned");
    return step_host_impl(e, nullptr, (const uint4 *) records_host, obs_host, reward_host, done_host, stream);
}

extern "C" int chub_reset_host(chub_engine *e, float *obs_host, void *stream) {
    int rc = chub_reset(e, nullptr, 0, stream);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t) stream;
    if (obs_host) CK(cudaMemcpyAsync(obs_host, e->obs, sizeof(float) * (size_t) e->H * CHUB_OBS_DIM, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return CHUB_OK;
}

extern "C" int chub_rollout_capture(chub_engine *e, const float *actions_dev, int32_t n_steps, void *stream) {
    if (!e || !actions_dev || n_steps < 1) return fail(CHUB_ERR_ARG, "chub_rollout_capture: bad argument");
    CK(cudaSetDevice(e->cfg.device));
    ChubParams P;
    int rc = fill_params(e, P);
    if (rc) return rc;
    P.actions = actions_dev;
    drop_graph(e);
    cudaStream_t s = (cudaStream_t) stream;
    cudaStream_t cap = s;
    bool own = false;
    if (cap == nullptr) {  // the legacy default stream cannot be captured
        CK(cudaStreamCreateWithFlags(&cap, cudaStreamNonBlocking));
        own = true;
    }
    CK(cudaStreamBeginCapture(cap, cudaStreamCaptureModeThreadLocal));
    unsigned long long tickets = 0;
    bool ticked = false;
    P.self_tick = 1;
    for (int j = 0; j < n_steps; ++j) {
        // persistent launches advance the step index themselves; for static grids the step offset is baked per node
        // and one tick closes the replay
        P.step_offset = ticked ? 0 : j;
        chub_launch_step(P, cap, &tickets, &ticked);
    }
    if (!ticked) chub_launch_tick(e->step_index, n_steps, e->queue + 1, tickets * (unsigned long long) n_steps, cap);
    cudaError_t ce = cudaStreamEndCapture(cap, &e->graph);
    if (own) cudaStreamDestroy(cap);
    if (ce != cudaSuccess) return fail(CHUB_ERR_CUDA, std::string("capture: ") + cudaGetErrorString(ce));
    CK(cudaGraphInstantiate(&e->graph_exec, e->graph, 0));
    e->graph_steps = n_steps;
    return CHUB_OK;
}

extern "C" int chub_rollout(chub_engine *e, void *stream) {
    if (!e || !e->graph_exec) return fail(CHUB_ERR_STATE, "chub_rollout: capture a rollout first");
    CK(cudaSetDevice(e->cfg.device));
    NvtxRange range("chub_rollout");
    CK(cudaGraphLaunch(e->graph_exec, (cudaStream_t) stream));
    e->launches += e->graph_steps;
    return CHUB_OK;
}

extern "C" int chub_counters(chub_engine *e, int64_t *step_index, int64_t *launches, uint32_t *error_or) {
    if (!e) return fail(CHUB_ERR_ARG, "chub_counters: null engine");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    if (step_index) {
        long long v = 0;
        CK(cudaMemcpy(&v, e->step_index, sizeof v, cudaMemcpyDeviceToHost));
        *step_index = v;
    }
    if (launches) *launches = e->launches;
    if (error_or) {
        std::vector<uint32_t> er(e->H);
        CK(cudaMemcpy(er.data(), e->err, sizeof(uint32_t) * e->H, cudaMemcpyDeviceToHost));
        uint32_t o = 0;
        for (uint32_t v : er) o |= v;
        *error_or = o;
    }
    return CHUB_OK;
}

// ---- output buffers, DLPack ----
struct BufInfo { void *ptr; int64_t bytes; int ndim; int64_t shape[2]; uint8_t code, bits; };
static bool buf_info(chub_engine *e, int which, BufInfo &b) {
    const int64_t H = e->H;
    switch (which) {
        case CHUB_BUF_OBS: b = {e->obs, H * CHUB_OBS_DIM * 4, 2, {H, CHUB_OBS_DIM}, 2, 32}; return true;
        case CHUB_BUF_REWARD: b = {e->reward, H * 4, 1, {H, 1}, 2, 32}; return true;
        case CHUB_BUF_DONE: b = {e->done, H, 1, {H, 1}, 1, 8}; return true;
        case CHUB_BUF_STATS: b = {e->stats, H * 32, 2, {H, 4}, 2, 64}; return true;
        case CHUB_BUF_LAST_STATS: b = {e->last_stats, H * 16, 2, {H, 4}, 2, 32}; return true;
        case CHUB_BUF_ERROR: b = {e->err, H * 4, 1, {H, 1}, 1, 32}; return true;
        case CHUB_BUF_FINAL_SOC: b = {e->final_soc, H * 4, 1, {H, 1}, 2, 32}; return true;
    }
    return false;
}

extern "C" int chub_buffer(chub_engine *e, int32_t which, void **dev_ptr, int64_t *bytes) {
    BufInfo b;
    if (!e || !buf_info(e, which, b)) return fail(CHUB_ERR_ARG, "chub_buffer: bad argument");
    if (dev_ptr) *dev_ptr = b.ptr;
    if (bytes) *bytes = b.bytes;
    return CHUB_OK;
}

// DLPack v0.8 structures (public ABI of dlpack.h, declared here to avoid a dependency)
extern "C" {
typedef struct { int32_t device_type; int32_t device_id; } ChubDLDevice;
typedef struct { uint8_t code; uint8_t bits; uint16_t lanes; } ChubDLDataType;
typedef struct {
    void *data; ChubDLDevice device; int32_t ndim; ChubDLDataType dtype;
    int64_t *shape; int64_t *strides; uint64_t byte_offset;
} ChubDLTensor;
typedef struct ChubDLManagedTensor {
    ChubDLTensor dl_tensor; void *manager_ctx; void (*deleter)(struct ChubDLManagedTensor *);
} ChubDLManagedTensor;
}
struct DLHolder { ChubDLManagedTensor t; int64_t shape[2]; chub_engine *e; };
static void dl_deleter(ChubDLManagedTensor *t) {
    DLHolder *hld = (DLHolder *) t->manager_ctx;
    engine_unref(hld->e);
    delete hld;
}

extern "C" int chub_tensor(chub_engine *e, int32_t which, void **dl_managed_tensor) {
    BufInfo b;
    if (!e || !dl_managed_tensor || !buf_info(e, which, b)) return fail(CHUB_ERR_ARG, "chub_tensor: bad argument");
    DLHolder *hld = new DLHolder();
    hld->e = e;
    hld->shape[0] = b.shape[0]; hld->shape[1] = b.shape[1];
    hld->t.dl_tensor.data = b.ptr;
    hld->t.dl_tensor.device = {2 /* kDLCUDA */, e->cfg.device};
    hld->t.dl_tensor.ndim = b.ndim;
    hld->t.dl_tensor.dtype = {b.code, b.bits, 1};
    hld->t.dl_tensor.shape = hld->shape;
    hld->t.dl_tensor.strides = nullptr;
    hld->t.dl_tensor.byte_offset = 0;
    hld->t.manager_ctx = hld;
    hld->t.deleter = dl_deleter;
    e->refcount += 1;
    *dl_managed_tensor = &hld->t;
    return CHUB_OK;
}

// ---- read-back ----
extern "C" int chub_hub_report(chub_engine *e, int32_t hub, double *host_out, void *stream) {
    if (!e || !host_out || hub < 0 || hub >= e->H) return fail(CHUB_ERR_ARG, "chub_hub_report: bad argument");
    CK(cudaSetDevice(e->cfg.device));
    ChubParams P;
    int rc = fill_params(e, P);
    if (rc) return rc;
    if (!e->report_dev) CK(dalloc(&e->report_dev, (size_t) 32));
    if (!e->report_host) CK(cudaHostAlloc((void **) &e->report_host, sizeof(double) * 32, cudaHostAllocDefault));
    cudaStream_t s = (cudaStream_t) stream;
    chub_launch_report(P, hub, e->report_dev, s);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(e->report_host, e->report_dev, sizeof(double) * CHUB_REPORT_LEN, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    memcpy(host_out, e->report_host, sizeof(double) * CHUB_REPORT_LEN);
    return CHUB_OK;
}

extern "C" int chub_get_field(chub_engine *e, int32_t field, void *host_out) {
    if (!e || !host_out || !e->have_classes) return fail(CHUB_ERR_ARG, "chub_get_field: bad argument");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    const int H = e->H, Hp = e->Hpad;
    int32_t *oi = (int32_t *) host_out;
    float *of = (float *) host_out;
    double *od = (double *) host_out;
    if (field < 32) {
        const size_t slots = (size_t) (e->S0 + e->S1) * Hp;
        std::vector<uint4> pile(slots);
        std::vector<float> arrive(slots);
        CK(cudaMemcpy(pile.data(), e->pile, slots * sizeof(uint4), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(arrive.data(), e->arrive, slots * sizeof(float), cudaMemcpyDeviceToHost));
        for (int h = 0; h < H; ++h) {
            const ChubClass &c = e->classes[e->hub_class[h]];
            for (int s = 0; s < 2; ++s)
                for (int i = 0; i < CHUB_MAX_PILES; ++i) {
                    const size_t o = ((size_t) h * 2 + s) * CHUB_MAX_PILES + i;
                    uint4 r = make_uint4(0, 0, 0, 0);
                    float arr = 0.f;
                    if (i < c.n[s]) {
                        const size_t idx = chub_pile_index(h, (s ? e->S0 : 0) + i, e->S0 + e->S1);
                        r = pile[idx];
                        arr = arrive[idx];
                    }
                    const bool car = r.w & PW_CAR;
                    float fx, fy, fz;
                    memcpy(&fx, &r.x, 4); memcpy(&fy, &r.y, 4); memcpy(&fz, &r.z, 4);
                    switch (field) {
                        case CHUB_F_PILE_CAR: oi[o] = car; break;
                        case CHUB_F_PILE_CHARGE: oi[o] = (r.w & PW_CHARGE) ? 1 : 0; break;
                        case CHUB_F_PILE_URGENT: oi[o] = (r.w & PW_URGENT) ? 1 : 0; break;
                        case CHUB_F_PILE_STAY: oi[o] = car ? (int) PW_STAY(r.w) : 0; break;
                        case CHUB_F_PILE_ALREADY: oi[o] = car ? (int) PW_ALREADY(r.w) : 0; break;
                        case CHUB_F_PILE_SOC: of[o] = car ? fx : 0.f; break;
                        case CHUB_F_PILE_TARGET: of[o] = car ? (float) PW_TU(r.w) / 999.0f * 20.0f + 80.0f : 0.f; break;
                        case CHUB_F_PILE_POWER: of[o] = car ? fz : 0.f; break;
                        case CHUB_F_PILE_ARRIVE: of[o] = car ? arr : 0.f; break;
                        default: return fail(CHUB_ERR_ARG, "chub_get_field: unknown pile field");
                    }
                }
        }
        return CHUB_OK;
    }
    if (field >= CHUB_F_ST_MIN_POWER && field <= CHUB_F_ST_MAX_POWER) {
        if (!e->st_power) return fail(CHUB_ERR_STATE, "chub_get_field: station powers need diagnostics=1");
        std::vector<float> sp((size_t) 6 * Hp);
        CK(cudaMemcpy(sp.data(), e->st_power, sp.size() * sizeof(float), cudaMemcpyDeviceToHost));
        const int col = field - CHUB_F_ST_MIN_POWER;
        for (int h = 0; h < H; ++h)
            for (int s = 0; s < 2; ++s) of[(size_t) h * 2 + s] = sp[(size_t) (3 * s + col) * Hp + h];
        return CHUB_OK;
    }
    if (field >= CHUB_F_DIAG) {
        const int k = field - CHUB_F_DIAG;
        if (!e->diag || k >= CHUB_D_COUNT) return fail(CHUB_ERR_STATE, "chub_get_field: diagnostics column unavailable");
        CK(cudaMemcpy(od, e->diag + (size_t) k * Hp, sizeof(double) * H, cudaMemcpyDeviceToHost));
        return CHUB_OK;
    }
    if (field == CHUB_F_HUB_STORE) {
        CK(cudaMemcpy(od, e->store, sizeof(double) * H, cudaMemcpyDeviceToHost));
        return CHUB_OK;
    }
    std::vector<uint4> hubA(H);
    CK(cudaMemcpy(hubA.data(), e->hubA, sizeof(uint4) * H, cudaMemcpyDeviceToHost));
    std::vector<uint4> pile;
    if (field == CHUB_F_ST_CAR_NUMBER) {
        pile.resize((size_t) (e->S0 + e->S1) * Hp);
        CK(cudaMemcpy(pile.data(), e->pile, pile.size() * sizeof(uint4), cudaMemcpyDeviceToHost));
    }
    for (int h = 0; h < H; ++h) {
        const uint4 a = hubA[h];
        switch (field) {
            case CHUB_F_ST_LINE: oi[2 * h] = (a.z >> 8) & 0xff; oi[2 * h + 1] = (a.z >> 16) & 0xff; break;
            case CHUB_F_ST_FLOW_IN: oi[2 * h] = (int) (a.w & 0xff) - 128; oi[2 * h + 1] = (int) ((a.w >> 8) & 0xff) - 128; break;
            case CHUB_F_ST_CAR_NUMBER: {
                const ChubClass &c = e->classes[e->hub_class[h]];
                for (int s = 0; s < 2; ++s) {
                    int cnt = 0;
                    for (int i = 0; i < c.n[s]; ++i)
                        cnt += (pile[chub_pile_index(h, (s ? e->S0 : 0) + i, e->S0 + e->S1)].w & PW_CAR) ? 1 : 0;
                    oi[2 * h + s] = cnt;
                }
                break;
            }
            case CHUB_F_HUB_T: oi[h] = a.z & 0xff; break;
            case CHUB_F_HUB_UCTR: oi[h] = (int32_t) a.y; break;
            case CHUB_F_HUB_LCG: oi[h] = (int32_t) a.x; break;
            case CHUB_F_FCEV_ARRIVE: oi[h] = a.w >> 24; break;
            case CHUB_F_FCEV_QN: oi[h] = a.z >> 24; break;
            default: return fail(CHUB_ERR_ARG, "chub_get_field: unknown field");
        }
    }
    return CHUB_OK;
}

// ---- checkpoint blob: a header, then every mutable state array and the outputs of the last step, in a fixed order ----
struct StateHeader {
    uint32_t magic, version;
    int32_t H, Hpad, S0, S1, n_classes, exo, diagnostics, reserved;
    int64_t bytes;       // of the whole blob, header included
    int64_t pad[2];
};
#define CHUB_STATE_MAGIC 0x42554843u   // "CHUB"
#define CHUB_STATE_VERSION 2u
struct Seg { void *p; size_t bytes; };
static std::vector<Seg> state_segments(chub_engine *e) {
    const size_t Hp = e->Hpad, slots = (size_t) (e->S0 + e->S1) * Hp;
    std::vector<Seg> v = {{e->pile, slots * sizeof(uint4)}, {e->arrive, slots * sizeof(float)}, {e->hubA, Hp * sizeof(uint4)},
            {e->store, Hp * sizeof(double)}, {e->keys, Hp * sizeof(uint2)}, {e->occ, Hp * sizeof(uint2)}, {e->stats, 2 * Hp * sizeof(double2)},
            {e->last_stats, Hp * sizeof(float4)}, {e->err, Hp * sizeof(uint32_t)},
            {e->qsoc, Hp * CHUB_FCEV_QCAP * sizeof(float)}, {e->step_index, sizeof(long long)},
            // what the last step returned: a trainer that resumes mid-episode feeds obs to its policy before it steps again
            {e->obs, Hp * CHUB_OBS_DIM * sizeof(float)}, {e->reward, Hp * sizeof(float)}, {e->done, Hp}, {e->final_soc, Hp * sizeof(float)}};
    if (e->exo_d) {   // generator state travels with the checkpoint
        v.push_back({e->exo_d, 7 * Hp * sizeof(double)});
        v.push_back({e->exo_i, Hp * sizeof(uint2)});
    }
    return v;
}
static StateHeader state_header(chub_engine *e) {
    StateHeader h;
    memset(&h, 0, sizeof h);
    h.magic = CHUB_STATE_MAGIC; h.version = CHUB_STATE_VERSION;
    h.H = e->H; h.Hpad = e->Hpad; h.S0 = e->S0; h.S1 = e->S1; h.n_classes = (int32_t) e->classes.size();
    h.exo = e->exo_d ? 1 : 0; h.diagnostics = e->cfg.diagnostics;
    size_t n = sizeof(StateHeader);
    for (const Seg &s : state_segments(e)) n += s.bytes;
    h.bytes = (int64_t) n;
    return h;
}
extern "C" int64_t chub_state_bytes(chub_engine *e) {
    if (!e || !e->have_classes) return -1;
    return state_header(e).bytes;
}
extern "C" int chub_get_state(chub_engine *e, void *host_blob) {
    if (!e || !host_blob || !e->have_classes) return fail(CHUB_ERR_ARG, "chub_get_state: bad argument");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    char *o = (char *) host_blob;
    const StateHeader h = state_header(e);
    memcpy(o, &h, sizeof h);
    o += sizeof h;
    for (const Seg &s : state_segments(e)) {
        CK(cudaMemcpy(o, s.p, s.bytes, cudaMemcpyDeviceToHost));
        o += s.bytes;
    }
    return CHUB_OK;
}
extern "C" int chub_set_state(chub_engine *e, const void *host_blob) {
    if (!e || !host_blob || !e->have_classes) return fail(CHUB_ERR_ARG, "chub_set_state: bad argument");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    const char *o = (const char *) host_blob;
    StateHeader got;
    memcpy(&got, o, sizeof got);
    const StateHeader want = state_header(e);
    if (got.magic != CHUB_STATE_MAGIC) return fail(CHUB_ERR_ARG, "chub_set_state: not a chub state blob");
    if (got.version != CHUB_STATE_VERSION) return fail(CHUB_ERR_ARG, "chub_set_state: blob written by another layout version");
    if (got.H != want.H || got.Hpad != want.Hpad || got.S0 != want.S0 || got.S1 != want.S1 || got.n_classes != want.n_classes ||
        got.exo != want.exo || got.bytes != want.bytes)
        return fail(CHUB_ERR_ARG, "chub_set_state: blob belongs to an engine of another shape (hubs, piles, classes or generator)");
    o += sizeof got;
    for (const Seg &s : state_segments(e)) {
        CK(cudaMemcpy(s.p, o, s.bytes, cudaMemcpyHostToDevice));
        o += s.bytes;
    }
    e->have_seed = true;
    return CHUB_OK;
}

// ---- per-launch timers ----
extern "C" int chub_set_timing(chub_engine *e, int32_t on) {
    if (!e) return fail(CHUB_ERR_ARG, "chub_set_timing: null engine");
    e->timing = on != 0;
    e->timed = 0;
    return CHUB_OK;
}
extern "C" int chub_step_times(chub_engine *e, float *ms_out, int32_t max_n, int32_t *n_out) {
    if (!e || !ms_out || !n_out || max_n < 0) return fail(CHUB_ERR_ARG, "chub_step_times: bad argument");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    const long long have = e->timed < CHUB_TIMER_RING ? e->timed : CHUB_TIMER_RING;
    const int n = (int) (have < max_n ? have : max_n);
    for (int j = 0; j < n; ++j) {   // oldest first
        const int slot = (int) ((e->timed - n + j) % CHUB_TIMER_RING);
        CK(cudaEventElapsedTime(&ms_out[j], e->time_ev[2 * slot], e->time_ev[2 * slot + 1]));
    }
    *n_out = n;
    return CHUB_OK;
}

// ---- self-checking build ----
extern "C" int chub_checks(chub_engine *e, int32_t *compiled_in, int64_t *failures) {
    if (!e) return fail(CHUB_ERR_ARG, "chub_checks: null engine");
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaDeviceSynchronize());
    if (compiled_in) *compiled_in = chub_checks_compiled_in();
    if (failures) {
        unsigned long long n = 0;
        CK(chub_read_check_failures(&n));
        *failures = (int64_t) n;
    }
    return CHUB_OK;
}

// ---- curve sweep (parity instrument) ----
extern "C" int chub_debug_curve(chub_engine *e, int32_t which, uint32_t first_bits, int64_t n, float *out_dev, void *stream) {
    if (!e || !out_dev || n < 1 || which < 0 || which > 15) return fail(CHUB_ERR_ARG, "chub_debug_curve: bad argument");
    if (!e->have_tables) return fail(CHUB_ERR_STATE, "chub_debug_curve: call chub_set_tables first (curve constants)");
    CK(cudaSetDevice(e->cfg.device));
    chub_launch_curve_sweep(which, first_bits, (long long) n, out_dev, (cudaStream_t) stream);
    CK(cudaGetLastError());
    return CHUB_OK;
}

// ---- the one collective of the path: episode statistics, once per rollout (the quantities mirror
// evcssp_manager.py:259-297).  NCCL is resolved at run time from the library the caller's communicator came from
// (libnccl.so.2 already in the process, else the system one), so libchub.so itself links no NCCL. ----
typedef struct ncclComm *chub_nccl_comm;
struct NcclApi {
    int (*GroupStart)();
    int (*GroupEnd)();
    int (*Send)(const void *, size_t, int, int, chub_nccl_comm, cudaStream_t);
    int (*Recv)(void *, size_t, int, int, chub_nccl_comm, cudaStream_t);
    int (*AllReduce)(const void *, void *, size_t, int, int, chub_nccl_comm, cudaStream_t);
    int (*CommCount)(chub_nccl_comm, int *);
    int (*CommUserRank)(chub_nccl_comm, int *);
    const char *(*GetErrorString)(int);
    bool ok = false;
};
#define CHUB_NCCL_FLOAT32 7   // ncclFloat32 (nccl.h ncclDataType_t)
#define CHUB_NCCL_FLOAT64 8   // ncclFloat64
#define CHUB_NCCL_SUM 0       // ncclSum
static NcclApi g_nccl;
static int nccl_load() {
    if (g_nccl.ok) return CHUB_OK;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);   // the copy the communicator's owner (e.g. torch) loaded
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW);
    if (!h) return fail(CHUB_ERR_UNSUPPORTED, std::string("NCCL not available: ") + dlerror());
    bool all = true;
    auto sym = [&](const char *name) { void *p = dlsym(h, name); if (!p) all = false; return p; };
    g_nccl.GroupStart = (int (*)()) sym("ncclGroupStart");
    g_nccl.GroupEnd = (int (*)()) sym("ncclGroupEnd");
    g_nccl.Send = (int (*)(const void *, size_t, int, int, chub_nccl_comm, cudaStream_t)) sym("ncclSend");
    g_nccl.Recv = (int (*)(void *, size_t, int, int, chub_nccl_comm, cudaStream_t)) sym("ncclRecv");
    g_nccl.AllReduce = (int (*)(const void *, void *, size_t, int, int, chub_nccl_comm, cudaStream_t)) sym("ncclAllReduce");
    g_nccl.CommCount = (int (*)(chub_nccl_comm, int *)) sym("ncclCommCount");
    g_nccl.CommUserRank = (int (*)(chub_nccl_comm, int *)) sym("ncclCommUserRank");
    g_nccl.GetErrorString = (const char *(*)(int)) sym("ncclGetErrorString");
    if (!all) return fail(CHUB_ERR_UNSUPPORTED, "NCCL library lacks a needed symbol");
    g_nccl.ok = true;
    return CHUB_OK;
}
#define NK(call)                                                                                   \
    do {                                                                                           \
        int r_ = (call);                                                                           \
        if (r_ != 0) return fail(CHUB_ERR_CUDA, std::string(#call) + ": " + g_nccl.GetErrorString(r_)); \
    } while (0)

static int comm_setup(chub_engine *e) {
    if (!e->comm_stream) {
        CK(cudaStreamCreateWithFlags(&e->comm_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&e->comm_ready, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&e->comm_done, cudaEventDisableTiming));
    }
    if (!e->stats_stage) CK(dalloc(&e->stats_stage, (size_t) e->Hpad));
    return CHUB_OK;
}

extern "C" int chub_gather_stats(chub_engine *e, void *nccl_comm, int32_t root, const int32_t *hubs_per_rank, float *recv_dev,
                                 void *stream) {
    if (!e || !nccl_comm) return fail(CHUB_ERR_ARG, "chub_gather_stats: null argument");
    CK(cudaSetDevice(e->cfg.device));
    int rc = nccl_load();
    if (rc) return rc;
    rc = comm_setup(e);
    if (rc) return rc;
    NvtxRange range("chub_gather_stats");
    chub_nccl_comm comm = (chub_nccl_comm) nccl_comm;
    int nranks = 0, rank = 0;
    NK(g_nccl.CommCount(comm, &nranks));
    NK(g_nccl.CommUserRank(comm, &rank));
    if (root < 0 || root >= nranks) return fail(CHUB_ERR_ARG, "chub_gather_stats: root outside the communicator");
    if (rank == root && !recv_dev) return fail(CHUB_ERR_ARG, "chub_gather_stats: the root needs a receive buffer");
    if (hubs_per_rank && hubs_per_rank[rank] != e->H) return fail(CHUB_ERR_ARG, "chub_gather_stats: hubs_per_rank disagrees with this engine");
    cudaStream_t s = (cudaStream_t) stream;
    if (e->comm_pending) CK(cudaStreamWaitEvent(s, e->comm_done, 0));   // the previous gather still reads the snapshot
    // snapshot on the caller's stream (a few microseconds), transfer on the engine's communication stream: the steps
    // the caller queues next overlap the transfer and may finish episodes without disturbing what is being sent
    CK(cudaMemcpyAsync(e->stats_stage, e->last_stats, sizeof(float4) * (size_t) e->H, cudaMemcpyDeviceToDevice, s));
    CK(cudaEventRecord(e->comm_ready, s));
    CK(cudaStreamWaitEvent(e->comm_stream, e->comm_ready, 0));
    NK(g_nccl.GroupStart());
    if (rank == root) {
        size_t at = 0;
        for (int r = 0; r < nranks; ++r) {
            const size_t n = (size_t) (hubs_per_rank ? hubs_per_rank[r] : e->H) * 4;
            if (r == rank) CK(cudaMemcpyAsync(recv_dev + at, e->stats_stage, n * sizeof(float), cudaMemcpyDeviceToDevice, e->comm_stream));
            else NK(g_nccl.Recv(recv_dev + at, n, CHUB_NCCL_FLOAT32, r, comm, e->comm_stream));
            at += n;
        }
    } else {
        NK(g_nccl.Send(e->stats_stage, (size_t) e->H * 4, CHUB_NCCL_FLOAT32, root, comm, e->comm_stream));
    }
    NK(g_nccl.GroupEnd());
    CK(cudaEventRecord(e->comm_done, e->comm_stream));
    e->comm_pending = true;
    return CHUB_OK;
}

extern "C" int chub_gather_wait(chub_engine *e, void *stream) {
    if (!e) return fail(CHUB_ERR_ARG, "chub_gather_wait: null engine");
    if (!e->comm_pending) return CHUB_OK;
    CK(cudaSetDevice(e->cfg.device));
    CK(cudaStreamWaitEvent((cudaStream_t) stream, e->comm_done, 0));
    e->comm_pending = false;
    return CHUB_OK;
}

extern "C" int chub_reduce_stats(chub_engine *e, void *nccl_comm, double *sums_dev, void *stream) {
    if (!e || !sums_dev) return fail(CHUB_ERR_ARG, "chub_reduce_stats: null argument");
    CK(cudaSetDevice(e->cfg.device));
    if (!e->stats_partial) CK(dalloc(&e->stats_partial, (size_t) 4 * CHUB_STATS_BLOCKS));
    NvtxRange range("chub_reduce_stats");
    cudaStream_t s = (cudaStream_t) stream;
    chub_launch_stats_sums(e->last_stats, e->H, e->stats_partial, sums_dev, s);
    CK(cudaGetLastError());
    if (nccl_comm) {
        int rc = nccl_load();
        if (rc) return rc;
        NK(g_nccl.AllReduce(sums_dev, sums_dev, 4, CHUB_NCCL_FLOAT64, CHUB_NCCL_SUM, (chub_nccl_comm) nccl_comm, s));
    }
    return CHUB_OK;
}

// pinned host memory for the callers of chub_step_host
extern "C" int chub_alloc_host(int64_t bytes, void **out) {
    if (!out || bytes <= 0) return fail(CHUB_ERR_ARG, "chub_alloc_host: bad argument");
    CK(cudaMallocHost(out, (size_t) bytes));
    return CHUB_OK;
}
extern "C" int chub_free_host(void *p) {
    if (p) CK(cudaFreeHost(p));
    return CHUB_OK;
}
