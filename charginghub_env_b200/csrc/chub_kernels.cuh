// chub_kernels.cuh -- launch parameters and HBM layout of the hub step / reset kernels.
#pragma once
#include "chub_device.cuh"

// HBM layout.  Pile state is tiled by warp: the S = S0 + S1 pile rows of 32 consecutive hubs are one contiguous
// block of S * 512 bytes (a few DRAM pages, a handful of TLB entries per warp instead of one 8 MB-strided row each):
//   pile[chub_pile_index(h, slot, S)]   uint4 {soc f32, s2t(soc) f32, power f32, packed}  slot = station0 piles, then station1 piles
//   arrive[chub_pile_index(h, slot, S)] float  arrival SOC ("init_soc" column; written once per car, never read by the step)
// Per-hub arrays are plain [Hpad] (a warp reads 32 consecutive elements):
//   hubA[h]                   uint4 {minstd state, U counter, packed1, packed2}
//   store[h]                  double hydrogen store mass in grams
//   keys[h]                   uint2  U-stream keys (read only)
//   occ[h]                    uint2  occupancy masks of the two stations: the step only touches pile rows that hold a car
//   stats[2h], stats[2h+1]    double2 x 2 {episode return, income | grid energy, incomplete departures}, running (fp64 like
//                             the reference's Python accumulators); last_stats[h] float4: the same at the last `done`
//   qsoc[(j) * Hpad + h]      float  FCEV backlog arrival SOCs (cold; touched only when a backlog exists)
// packed pile word : bits 0-5 stay_time, 6-11 already_stay_time, 12 car, 13 charge, 14 urgent,
//                    15-19 ceil(needed_time) cache (0 = needed_time <= 0), 20-29 target draw u (target = u/999*20+80)
// packed1          : bits 0-7 t (time hole), 8-15 line[0], 16-23 line[1], 24-31 FCEV backlog length
// packed2          : bits 0-7 flow_in[0]+128, 8-15 flow_in[1]+128, 16-23 FCEV line, 24-31 FCEV arrivals
#define PW_CAR (1u << 12)
#define PW_CHARGE (1u << 13)
#define PW_URGENT (1u << 14)
#define PW_STAY(w) ((w) & 0x3fu)
#define PW_ALREADY(w) (((w) >> 6) & 0x3fu)
#define PW_CN(w) (((w) >> 15) & 0x1fu)
#define PW_TU(w) (((w) >> 20) & 0x3ffu)


// element index of pile `slot` of hub `h`: [h / 32][slot][h % 32]
__host__ __device__ inline size_t chub_pile_index(int h, int slot, int S) {
    return ((size_t) (h >> 5) * (size_t) S + (size_t) slot) * 32 + (size_t) (h & 31);
}

struct ChubParams {
    int H, Hpad, S0, S1;
    int h_begin, h_end;               // hub range of this launch (h_begin a multiple of 32); the whole batch by default
    int n_classes, action_stride, auto_reset, diagnostics;
    int trace_len, trace_per_hub;
    int block, phase_sync;            // launch shape override of the step kernel (see chub_launch_step); phase_sync: unused
    int num_sms;                      // SMs of the device (launch-shape choice)
    int ws_bytes;                     // bytes of shared-memory scratch per warp (set by chub_launch_step)
    int step_offset;                  // added to *step_index (graph-captured rollouts bake 0..n-1)
    int self_tick;                    // the step kernel's last block advances *step_index (and the queue's epoch) itself: no tick launch
    const long long *step_index;      // device: absolute step index of the state being left
    unsigned long long *queue;        // device: ticket counter of the persistent step kernel; NULL = static block -> tile mapping
    const unsigned long long *queue_base;   // device: the counter's value when the current epoch (since the last tick) began
    unsigned long long queue_stride;  // tickets one launch draws (set by chub_launch_step)
    unsigned char *lpt;               // device: two unit lists binned by measured cost (see chub_step_kernel)
    int lpt_cap;                      // units a bin's list can hold
    uint4 *pile;
    float *arrive;
    uint4 *hubA;
    double *store;
    const uint2 *keys;
    uint2 *occ;                       // per hub: occupancy bit mask of station 0 / station 1 piles (bit i = a car on pile i)
    const int *cls;                   // NULL when n_classes == 1
    double2 *stats;
    float4 *last_stats;
    uint32_t *err;
    float *qsoc;
    const ChubClass *classes;
    const double *el_table;           // [n_classes][102]
    const uint16_t *lut;              // [96][1000] traffic count for u = r/999.f (CHS.hpp:731-743)
    const float *s2t_lut;             // [2 constant][2 type][1000] soc_to_time(target(u))
    const double *pv, *wd, *price_tr;
    // on-device exogenous generator (exo_mode != 0): per-hub state and the day tables
    int exo_mode, n_pv_days, n_wd_days;
    double exo_renew_gain, exo_price_gain;   // 1 + renew_fluctuate, 1 + price_fluctuate
    double *exo_d;                    // [7][Hpad]: OU states pv, wd, price; held price noise; current pv, wd, price
    uint2 *exo_i;                     // [Hpad]: {pv_day | wd_day << 10 | price_count mod 4 << 20, X-stream counter}
    const double *pv_days, *wd_days;  // [n][96]
    const float *actions;
    float *final_soc;                    // store SOC at the end of the last finished episode (evcssp_manager.py:275-290)
    const unsigned long long *act_bits;  // scalar-load mode: pile on-bits from chub_dispatch_kernel; actions = [H][4] rows
    const uint4 *act_packed;             // packed-action mode: [H] records {on-bits lo, hi, a[-2], a[-1]} (chub_step_packed)
    float *obs;
    float *reward;
    uint8_t *done;
    double *diag;                     // [CHUB_D_COUNT][Hpad] or NULL
    float *st_power;                  // [6][Hpad] (min,charge,max) x 2 stations, or NULL
    const uint8_t *reset_mask;        // reset kernel only
};

// tickets: what a later tick has to add to the queue's epoch base; self_ticked: the launch advances the step index itself
// (asked for with p.self_tick, granted when the launch is a persistent one: its last block out is known)
cudaError_t chub_launch_step(const ChubParams &p, cudaStream_t s, unsigned long long *tickets = nullptr, bool *self_ticked = nullptr);
int chub_step_shape(const ChubParams &p);   // warps per block * 100 + blocks per SM * 10 of a launch with these parameters
void chub_launch_reset(const ChubParams &p, cudaStream_t s);
void chub_launch_situation(const ChubParams &p, float *out, cudaStream_t s);
void chub_launch_report(const ChubParams &p, int h, double *out, cudaStream_t s);
void chub_launch_dispatch(const ChubParams &p, unsigned long long *act_bits, cudaStream_t s);
void chub_launch_s2t_lut(float *out, cudaStream_t s);
void chub_launch_tick(long long *step_index, int n, unsigned long long *queue_base, unsigned long long tickets, cudaStream_t s);
#define CHUB_STATS_BLOCKS 296
#define CHUB_LPT_BYTES(cap) (2 * (272 + (size_t) 64 * (cap) * 4))   // two LptList + their 64 bins of `cap` ints
void chub_launch_stats_sums(const float4 *stats, int H, double *partial, double *sums, cudaStream_t s);
void chub_launch_curve_sweep(int which, uint32_t first_bits, long long n, float *out, cudaStream_t s);
cudaError_t chub_upload_const(const ChubConst &c);
// chub_noalloc.cu: device address of the file-scope scratch of chub_step_loads_static on the CURRENT device `device`
cudaError_t chub_static_act_bits(int device, unsigned long long **ptr);
cudaError_t chub_read_check_failures(unsigned long long *out);
int chub_checks_compiled_in();
