/* chub.h -- C ABI of the B200 batched charging-hub stepping engine (libchub.so).
 *
 * This is the drop-in boundary for the reference's step/reset path.  In the reference that path
 * crosses a Boost.Python module, `pyevstation` (evcssp_env_cpp/envs/lion_cpp20/main.cpp:19-290),
 * roughly ten times per env step and per hub: SlowChargeStation/FastChargeStation.evs_step /
 * evs_reset (main.cpp:186-233, called from Aggregator_Simple.py:141-142,168-169), Vector_float
 * appends (Aggregator_Simple.py:132-139), PoissonNumber.hv_car_number_wrt_poisson and
 * CarArriveRandom.mk_soc (hydro_sys.py:251,259), and ~20 attribute reads
 * (evcssp_manager.py:366-368, Aggregator_Simple.py:208-218).  Here the WHOLE of
 * EvcsspManagerEnv_v6.step()/reset() (evcssp_manager.py:136-316) for a batch of hubs is one call.
 *
 * Conventions: plain C, no exceptions; every function returns CHUB_OK (0) or a negative error and
 * records a message readable with chub_last_error(); all device work is stream ordered on the
 * cudaStream_t passed in (as void*; NULL = legacy default stream); the engine owns all device
 * memory; pointers named *_dev are device pointers owned by the caller.  One engine per CUDA
 * device; calls on one engine must be serialised by the caller.
 */
#ifndef CHUB_H
#define CHUB_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CHUB_OK 0
#define CHUB_ERR_ARG (-1)
#define CHUB_ERR_CUDA (-2)
#define CHUB_ERR_STATE (-3)
#define CHUB_ERR_UNSUPPORTED (-4)

#define CHUB_MAX_PILES 32   /* piles per station (fast or slow) */
#define CHUB_OBS_DIM 13     /* evcssp_manager.py:74-95 with both stations populated */
#define CHUB_EL_TABLE 102   /* hy_power_speed_list, hydro_sys.py:154-157 */
#define CHUB_FCEV_QCAP 16   /* FCEV backlog entries kept per hub (reference list is unbounded); oracle/hub_oracle.c uses the same */
#define CHUB_DAY 96         /* steps per episode, evcssp_manager.py:53 */

#define CHUB_FAST 0
#define CHUB_SLOW 1

/* sticky per-hub error bits (the reference would print or raise) */
#define CHUB_HUBERR_FCEV_OVERFLOW 1u   /* FCEV list longer than CHUB_FCEV_QCAP (hydro_sys.py:266-285) */
#define CHUB_HUBERR_STORE_ASSERT 2u    /* hydro_sys.py:111-112 assert would have fired */
#define CHUB_HUBERR_COUNTER 4u         /* a stay/queue counter left its packed range */
#define CHUB_HUBERR_BAD_LOAD 8u        /* chub_step_loads got a NaN power target (treated as "nothing beyond min_power") */

typedef struct chub_engine chub_engine;

/* Engine-wide configuration. */
typedef struct chub_config {
    int32_t n_hubs;          /* hubs stepped by THIS engine (the local shard) */
    int32_t device;          /* CUDA device ordinal */
    int64_t hub_id_base;     /* global id of local hub 0; seeds are f(base_seed, global id) */
    int32_t n_classes;       /* number of hub parameter classes (>=1) */
    int32_t action_stride;   /* floats per hub row of the action matrix, >= max(n0+n1)+2 */
    int32_t auto_reset;      /* 1: a hub that reports done is reset inside the same launch */
    int32_t diagnostics;     /* 1: also write the fp64 diagnostic columns (parity tests) */
    int32_t reserved[4];
} chub_config;

/* One hub parameter class == one set of the reference's constructor kwargs
 * (evcssp_manager.py:25-27, test/env_test.py:10-30).  Negative optional values mean "None". */
typedef struct chub_hub_class {
    int32_t n_piles[2];          /* station_list */
    int32_t station_type[2];     /* station_type_list: CHUB_FAST / CHUB_SLOW */
    int32_t constant_charging;
    int32_t reserved;
    double hydro_prod_rate;      /* m^3/h; <=0 -> 430 (hydro_sys.py:140-143) */
    double hydro_store_vlt;      /* m^3;   <=0 -> 5000 (hydro_sys.py:96) */
    double init_soc;             /* store SOC at reset, in [0.1, 1] (hydro_sys.py:137) */
    double fc_max_power;         /* kW; <0 -> 100 (hydro_sys.py:401-404) */
    double fcev_permeate;        /* hydro_sys.py:248 */
    double hydro_loss;           /* hydro_sys.py:121 */
    /* optional explicit electrolyser power table (hy_power_speed_list); NULL -> the engine builds
     * the table the reference obtains when the store is not clipped during construction */
    const double *el_table;      /* [CHUB_EL_TABLE] or NULL */
} chub_hub_class;

/* Static tables (host pointers, copied). */
typedef struct chub_tables {
    const double *arrival_cdf;   /* [96][301] car_flow_possibility_list as parsed by CHS.hpp:93-175 */
    const double *price;         /* [96] price_after_MAD_96 (Aggregator_Simple.py:9-15) */
    double price_mean, price_std;/* evcssp_manager.py:45-46 */
    const double *time_sin;      /* [96] sin(2*pi*t/96) (evcssp_manager.py:319-320) */
} chub_tables;

/* Which buffer chub_buffer()/chub_tensor() exposes. */
enum chub_buffer_id {
    CHUB_BUF_OBS = 0,        /* float32 [n_hubs][13] */
    CHUB_BUF_REWARD = 1,     /* float32 [n_hubs] */
    CHUB_BUF_DONE = 2,       /* uint8   [n_hubs] */
    CHUB_BUF_STATS = 3,      /* float64 [n_hubs][4]: running episode return, income, grid energy drawn (kW-steps), incomplete
                              * departures -- fp64 like the reference's accumulators (evcssp_manager.py:259-269) */
    CHUB_BUF_LAST_STATS = 4, /* float32 [n_hubs][4]: the same, frozen at the last completed episode (the gather's payload) */
    CHUB_BUF_ERROR = 5,      /* uint32  [n_hubs] sticky error bits */
    CHUB_BUF_FINAL_SOC = 6   /* float32 [n_hubs] hydrogen store SOC when the last finished episode ended: what the
                              * reference's end-of-episode SOC deviation / test_penalty is formed from
                              * (evcssp_manager.py:275-290); an output like `done`, not part of the checkpoint */
};

/* Field ids for chub_get_field (read-back for parity tests, introspection, checkpoints).
 * Output is hub-major, dense, host memory. */
enum chub_field_id {
    CHUB_F_PILE_CAR = 0, CHUB_F_PILE_CHARGE, CHUB_F_PILE_URGENT, CHUB_F_PILE_STAY, CHUB_F_PILE_ALREADY, /* int32 [H][2][32] */
    CHUB_F_PILE_SOC = 16, CHUB_F_PILE_TARGET, CHUB_F_PILE_POWER, CHUB_F_PILE_ARRIVE,                   /* float [H][2][32] */
    CHUB_F_ST_LINE = 32, CHUB_F_ST_FLOW_IN, CHUB_F_ST_CAR_NUMBER,                                      /* int32 [H][2] */
    CHUB_F_ST_MIN_POWER = 48, CHUB_F_ST_CHARGE_POWER, CHUB_F_ST_MAX_POWER,                             /* float [H][2], diagnostics only */
    CHUB_F_HUB_T = 64, CHUB_F_HUB_UCTR, CHUB_F_HUB_LCG, CHUB_F_FCEV_ARRIVE, CHUB_F_FCEV_QN,            /* int32 [H] */
    CHUB_F_HUB_STORE = 80,                                                                             /* double [H] */
    CHUB_F_DIAG = 96   /* CHUB_F_DIAG + k: double [H], diagnostics column k (enum chub_diag_id) */
};

enum chub_diag_id {
    CHUB_D_REWARD = 0, CHUB_D_INCOME, CHUB_D_HY_FLOW, CHUB_D_H2_POWER, CHUB_D_FC_POWER, CHUB_D_HY_USE,
    CHUB_D_NOT_MEET, CHUB_D_HY_TO_USE, CHUB_D_DEMAND, CHUB_D_EV_GRID0, CHUB_D_EV_GRID1, CHUB_D_H2_GRID,
    CHUB_D_HY_ACT, CHUB_D_STORE_SOC, CHUB_D_COUNT
};

const char *chub_last_error(void);
int chub_version(void);

int chub_create(const chub_config *cfg, chub_engine **out);
int chub_destroy(chub_engine *e);

/* classes[n_classes]; hub_class = per-hub class index [n_hubs] (host) or NULL when n_classes == 1 */
int chub_set_hub_classes(chub_engine *e, const chub_hub_class *classes, int32_t n_classes, const int32_t *hub_class);
int chub_set_tables(chub_engine *e, const chub_tables *t);
/* read back the electrolyser table the engine uses for class k (host, [CHUB_EL_TABLE]) */
int chub_get_class_table(chub_engine *e, int32_t k, double *out);

/* Exogenous traces (the reference's ReNew + OU noise + price list, renewable.py:21-76,
 * evcssp_manager.py:344-361).  pv is the scaled PV power (x5), wd the wind power, price = table[t-1] + noise.
 * Addressing: a hub whose time slot is t when absolute step k starts sees trace[(96 * floor((k - t) / 96) + t) % length]
 * in the state it leaves and slot + 1 in the state it enters -- for hubs in lock step with the step index (reset at a
 * multiple of 96: every full reset, every auto-reset) that is simply trace[k % length] and trace[(k + 1) % length];
 * hubs reset in the middle of a day (masked reset, truncation) keep reading the slots of THEIR time of day.
 * A rollout longer than the trace wraps around: make `length` a multiple of 96 then.
 * per_hub = 0: arrays [length];  per_hub = 1: arrays [length][n_hubs].  Host pointers, copied.
 * Invalidates a captured rollout (chub_rollout_capture), like every setter that reallocates device memory. */
int chub_set_traces(chub_engine *e, const double *pv, const double *wd, const double *price,
                    int32_t length, int32_t per_hub);

/* Overwrite one slot of SHARED traces (an online caller that generates the exogenous values step by
 * step, like the single-hub gym wrapper, keeps a short ring and pokes the next slot before each step). */
int chub_poke_trace(chub_engine *e, int32_t slot, double pv, double wd, double price, void *stream);

/* On-device exogenous generator, instead of traces: every hub draws its own PV day and wind day at each reset
 * (ReNew.renew_reset, renewable.py:51-53), adds Ornstein-Uhlenbeck noise to them (ReNew.get_pv_power /
 * get_wd_power + OU_Noise, renewable.py:37-76) and to the price, whose noise is resampled every fourth state and
 * held in between (make_state, evcssp_manager.py:344-361), inside the step / reset kernels.  The variates come
 * from the per-hub X stream of include/chub_exo.h (keyed like the other streams, so chub_seed fixes them); the OU
 * states survive resets like the reference's.  pv_days [n_pv_days][96] (unscaled: the x5 is applied inside),
 * wd_days [n_wd_days][96]: host pointers, copied.  The price table is the one of chub_set_tables.
 * With auto-reset the finished episode's last make_state is generated (and discarded) before the new episode's
 * first one, which is what a reference loop "step ... done -> reset" consumes. */
int chub_set_exogenous_generator(chub_engine *e, const double *pv_days, int32_t n_pv_days, const double *wd_days,
                                 int32_t n_wd_days, double renew_fluctuate, double price_fluctuate);

/* Streams: either derive (key0,key1,minstd state) from base_seed and the global hub id
 * (include/chub_rng.h), or set them explicitly (host arrays [n_hubs]; uctr may be NULL = 0). */
int chub_seed(chub_engine *e, uint64_t base_seed);
int chub_seed_explicit(chub_engine *e, const uint32_t *key0, const uint32_t *key1, const uint32_t *uctr,
                       const uint32_t *lcg);

/* reset(): EvcsspManagerEnv_v6.reset (evcssp_manager.py:304-316) for every hub, or for hubs whose
 * mask byte is non-zero (mask_dev: device uint8 [n_hubs] or NULL).  Also rewinds the absolute
 * step index to `step_index` (use 0; negative = leave it).  Fills the obs buffer.
 * With a mask (cutting episodes short, TimeLimit-style) the reward / done of the last step stay as they are and the
 * masked hubs' running statistics become their CHUB_BUF_LAST_STATS (and CHUB_BUF_FINAL_SOC the store's present SOC),
 * as at `done`; a full reset clears reward and done. */
int chub_reset(chub_engine *e, const uint8_t *mask_dev, int64_t step_index, void *stream);

/* step(action): EvcsspManagerEnv_v6.step (evcssp_manager.py:136-302) for every hub.
 * actions_dev: float32 [n_hubs][action_stride], row = [pile actions of station 0, station 1, a[-2], a[-1]]
 * in [-1,1] exactly as the reference's action vector.  Fills obs / reward / done buffers. */
int chub_step(chub_engine *e, const float *actions_dev, void *stream);

/* chub_step with the results written straight into buffers of the caller -- the slices of a trainer's rollout storage
 * (obs_out_dev float32 [n_hubs][13], reward_out_dev float32 [n_hubs], done_out_dev uint8 [n_hubs]; NULL = the engine's
 * own buffer, which is then the only one updated for that quantity).  Saves the copy kernels of a rollout loop; with
 * auto-reset the obs written is the first obs of the next episode, as in chub_step. */
int chub_step_into(chub_engine *e, const float *actions_dev, float *obs_out_dev, float *reward_out_dev,
                   uint8_t *done_out_dev, void *stream);

/* The stations' scalar-load step, evs_step(float) (CHS.hpp:1169-1186 slow, 1478-1495 fast; bound in main.cpp:196-197):
 * each station gets ONE power target in kW instead of per-pile switches; it is clamped to [min_power, max_power]
 * (catch_load, CHS.hpp:358-366) and handed out to the cars in order of urgency (assign_on_off / rank_power_add,
 * CHS.hpp:1318-1362, 1375-1402).  The rest of the hub step is chub_step's.
 * loads_dev: float32 [n_hubs][4], row = [target of station 0, target of station 1, a[-2], a[-1]]; the last two are the
 * electrolyser / fuel-cell actions in [-1,1] as in chub_step.  Targets must be finite. */
int chub_step_loads(chub_engine *e, const float *loads_dev, void *stream);

/* chub_step_loads for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * chub_step_loads keeps the dispatch kernel's output (one 64-bit word of pile on-bits per hub) in a buffer
 * it cudaMalloc's on its first call; this variant keeps it in a file-scope device array of the library
 * (csrc/chub_noalloc.cu) of CHUB_STATIC_MAX_HUBS words = 8 MiB, which every GPU that loads libchub.so holds until the
 * process exits.  It calls no cudaMalloc / cudaFree, creates no stream or event, and calls into no allocating library;
 * it launches the dispatch kernel and the step kernel (which advances the step index itself; behind a static grid a
 * one-thread tick kernel follows) on `stream` and returns.
 * loads_dev: as chub_step_loads.  obs_out_dev / reward_out_dev / done_out_dev: as chub_step_into (NULL = the engine's
 * own buffer).  Engines with more than CHUB_STATIC_MAX_HUBS hubs are refused with CHUB_ERR_ARG (use chub_step_loads).
 * The array is shared by every engine on a device and by every call, and keeps the last call's contents; nothing in it
 * has to start at zero (every hub's word is written before it is read, stream ordered), but calls that OVERLAP on the
 * GPU -- two engines of one device on different streams, one engine driven from two streams -- race on it.  Its device
 * address is resolved at chub_create (set-up), once per device.  chub_set_timing's per-launch timers do not cover it. */
#define CHUB_STATIC_MAX_HUBS 1048576
int chub_step_loads_static(chub_engine *e, const float *loads_dev, float *obs_out_dev, float *reward_out_dev,
                           uint8_t *done_out_dev, void *stream);

/* step(action) on packed actions: packed_dev = [n_hubs] 16-byte records {uint32 on-bits of piles 0..31, uint32 on-bits of
 * piles 32..63, float a[-2], float a[-1]}, pile j = entry j of the chub_step row (station 0's piles first); bit j is
 * what action_to_real (evcssp_manager.py:384-404) makes of row entry j: 1 iff a[j] >= -2^-25.  For producers that
 * already hold switch decisions, and the device side of chub_step_host.  Must be 16-byte aligned. */
int chub_step_packed(chub_engine *e, const void *packed_dev, void *stream);

/* The same through HOST buffers (what a CPU caller of the reference has): takes the float action rows of chub_step from
 * host memory, steps, copies obs/reward/done out, and synchronises the stream.  Any out pointer may be NULL.
 * PCIe bounds this call, so the rows are thresholded into chub_step_packed records by a pool of host threads and only
 * those cross the bus (16 instead of 4*action_stride bytes per hub); ranges of hubs flow through pack -> copy in ->
 * kernel -> copy out concurrently.  chub_set_host_threads chooses the pool size (0 = send the float rows unpacked;
 * default: CHUB_HOST_THREADS or min(cores - 1, 16)) and the number of ranges (0 = keep; default 16 packed, 8 unpacked; at most 32).
 * One process per GPU on a shared host should divide the cores between the ranks (the default does, by LOCAL_WORLD_SIZE).
 * chub_pack_actions is the packing alone: [n_hubs][action_stride] host rows -> [n_hubs] host records. */
int chub_step_host(chub_engine *e, const float *actions_host, float *obs_host, float *reward_host,
                   uint8_t *done_host, void *stream);
/* chub_step_host for a HOST producer that already holds switch decisions: records_host = [n_hubs] 16-byte records of
 * chub_step_packed in host memory (pinned for full overlap; 16-byte aligned).  Same pipeline -- records host->device,
 * step kernel, results device->host, range by range -- minus the thresholding: 16 bytes per hub cross the host's memory
 * system instead of 4 * action_stride, which is what bounds chub_step_host when several GPUs share one host. */
int chub_step_host_packed(chub_engine *e, const void *records_host, float *obs_host, float *reward_host,
                          uint8_t *done_host, void *stream);
int chub_reset_host(chub_engine *e, float *obs_host, void *stream);
int chub_set_host_threads(chub_engine *e, int32_t n_threads, int32_t n_ranges);
int chub_pack_actions(chub_engine *e, const float *actions_host, void *packed_host);
/* The thresholding itself, engine-free (csrc/chub_hostpack.cpp): rows float32 [n_rows][stride]; na_row = per-row number
 * of pile actions or NULL for na_all everywhere; out = [n_rows] records. */
void chub_pack_rows(const float *rows, int32_t stride, const int32_t *na_row, int32_t na_all, int64_t n_rows, void *out);

/* n_steps step launches captured once into a CUDA graph and replayed; the action buffer is re-read
 * at every step, so a policy writing into it on the same stream may be captured by the caller
 * instead.  chub_rollout replays the captured graph. */
int chub_rollout_capture(chub_engine *e, const float *actions_dev, int32_t n_steps, void *stream);
int chub_rollout(chub_engine *e, void *stream);

/* Raw device pointer + byte size of an engine-owned output buffer. */
int chub_buffer(chub_engine *e, int32_t which, void **dev_ptr, int64_t *bytes);
/* DLPack export (DLManagedTensor*, version 0.8 layout) of the same buffers; the tensor is a view and
 * keeps the engine alive until its deleter runs. */
int chub_tensor(chub_engine *e, int32_t which, void **dl_managed_tensor);

/* The stations' `situation` table (CHS.hpp:204-231; what main.cpp:54-74, 142-184 exposes as nested maps) as dense
 * device matrices for richer observations and debugging: out_dev = float32 [CHUB_SIT_COLUMNS][n_hubs][S], S = the
 * widest station 0 + the widest station 1 over the hub classes (chub_situation_width); station 0's piles come first.
 * Piles without a car and slots beyond a station's pile count read 0, as after reset_position.  Asynchronous. */
enum chub_situation_column {
    CHUB_SIT_CAR = 0, CHUB_SIT_CHARGE, CHUB_SIT_EMERGENCY, CHUB_SIT_ASSIGN /* always 0 on this path */, CHUB_SIT_POWER,
    CHUB_SIT_SOC, CHUB_SIT_INIT_SOC, CHUB_SIT_TARGET_SOC, CHUB_SIT_COLUMNS
};
int chub_situation_width(chub_engine *e, int32_t *s0, int32_t *s1);
int chub_situation(chub_engine *e, float *out_dev, void *stream);

/* One hub's scalars in one stream-ordered call -- what the single-hub gym wrapper reads after every step (the reference's
 * real_state, evcssp_manager.py:364-372, and the attributes step() sets, evcssp_manager.py:229-297) -- instead of one
 * chub_get_field round trip per quantity.  host_out[CHUB_REPORT_LEN] doubles: [0..13] diagnostics columns (enum
 * chub_diag_id; 0 unless diagnostics=1), [14..16] min / charge / max power of station 0, [17..19] of station 1 (0 unless
 * diagnostics=1), [20..21] line, [22..23] flow_in_number, [24] t, [25] store mass in g, [26] FCEV arrivals of the step,
 * [27] FCEV backlog, [28] sticky error bits.  Synchronises the stream. */
#define CHUB_REPORT_LEN 29
int chub_hub_report(chub_engine *e, int32_t hub, double *host_out, void *stream);

/* Parity / checkpoint read-back (synchronises).  The state blob opens with a header (magic, layout version, hubs, pile
 * slots, classes, generator flag, total size) that chub_set_state validates, and carries every mutable state array plus
 * what the last step returned (obs, reward, done, final SOC), so a resumed trainer sees the observation it left. */
int chub_get_field(chub_engine *e, int32_t field, void *host_out);
int64_t chub_state_bytes(chub_engine *e);
int chub_get_state(chub_engine *e, void *host_blob);
int chub_set_state(chub_engine *e, const void *host_blob);

/* The one collective of the path: the per-hub statistics of the last finished episode -- return, income,
 * grid energy, incomplete departures; what the reference prints at `done`, evcssp_manager.py:259-297, CHS.hpp:916-921 --
 * gathered on `root` once per rollout.  nccl_comm is an ncclComm_t (as void*) spanning the ranks' engines; NCCL is
 * resolved at run time from the libnccl.so.2 already loaded in the process (torch's, when the communicator is torch's:
 * ProcessGroupNCCL._comm_ptr()), else the system one -- libchub.so links none.
 * hubs_per_rank: host int32 [nranks] (NULL = every rank holds as many hubs as this engine); recv_dev: device float32
 * [sum of hubs][4] in rank order, needed on the root only.
 * The statistics are snapshotted on `stream` and travel on the engine's own communication stream, so steps queued on
 * `stream` afterwards overlap the transfer; chub_gather_wait(e, stream) makes `stream` wait for its completion (call it
 * before recv_dev is read, and before the timed region of a measurement ends).
 * chub_reduce_stats: the 32-byte variant -- column sums over all hubs of all ranks, fp64 [4] in sums_dev on every rank
 * (fixed summation order; nccl_comm NULL = this engine's hubs only), on `stream`. */
int chub_gather_stats(chub_engine *e, void *nccl_comm, int32_t root, const int32_t *hubs_per_rank, float *recv_dev, void *stream);
int chub_gather_wait(chub_engine *e, void *stream);
int chub_reduce_stats(chub_engine *e, void *nccl_comm, double *sums_dev, void *stream);

/* Per-launch timers: with timing on, every step call records CUDA events around its step kernel on
 * the stream it launches on; chub_step_times synchronises the device and returns the durations (ms, oldest first) of the
 * most recent launches, at most 128 and at most max_n.  Calls made while the stream is being captured are not timed. */
int chub_set_timing(chub_engine *e, int32_t on);
int chub_step_times(chub_engine *e, float *ms_out, int32_t max_n, int32_t *n_out);

/* Parity instrument: the charging curves exactly as the step kernel compiles them (CHS.hpp:495-588, 621-725), evaluated
 * for the n consecutive float32 bit patterns starting at first_bits -> out_dev float32 [n].  which: 0 slow time->power (as
 * add_car uses it), 1 slow car_step SOC, 2 slow SOC->time, 3 fast time->power, 4 fast car_step SOC, 5 fast SOC->time,
 * 6 slow car_step power, 7 fast car_step power; +8 = constant_charging variants.  tests/test_gpu_curves.py walks every
 * reachable argument with it. */
int chub_debug_curve(chub_engine *e, int32_t which, uint32_t first_bits, int64_t n, float *out_dev, void *stream);

/* Self-checking build: libchub_checked.so is this library compiled with -DCHUB_CHECKS, which makes the kernels assert that
 * every index into their shared-memory rings, posted-power rows, pile tiles and the unit queue stays inside its array
 * (what compute-sanitizer's memcheck would watch; that tool is unavailable on the development pool).  compiled_in = 1
 * in that build; failures = violations counted on this device since the library was loaded (0 is the only good value). */
int chub_checks(chub_engine *e, int32_t *compiled_in, int64_t *failures);

/* Absolute step index, number of step kernel launches so far, and sum of sticky error bits. */
int chub_counters(chub_engine *e, int64_t *step_index, int64_t *launches, uint32_t *error_or);

/* Pinned host memory for chub_step_host / chub_reset_host callers. */
int chub_alloc_host(int64_t bytes, void **out);
int chub_free_host(void *p);

/* Algorithmic HBM bytes of one hub-step for class k: B(nf,ns) of SURVEY.md section 8d. */
int64_t chub_algorithmic_bytes(chub_engine *e, int32_t k);

/* Hubs per scheduling unit of a whole-batch step launch: the step kernel hands its persistent blocks units of this many
 * consecutive hubs (one 32-hub tile per warp).  Hubs that share a unit are stepped side by side on one SM, so a caller
 * that can choose the order of its hubs (which are independent: evcssp_manager.py:25-27 takes no neighbour) does best
 * to keep hubs of similar size and phase together in runs of this length.  Informational; no reference counterpart. */
int32_t chub_unit_hubs(chub_engine *e);

#ifdef __cplusplus
}
#endif
#endif /* CHUB_H */
