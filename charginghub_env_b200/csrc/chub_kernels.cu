// chub_kernels.cu -- the fused hub step kernel and the reset kernel (sm_100a).
//
// One thread advances one hub through the whole of EvcsspManagerEnv_v6.step()
// (evcssp_manager.py:136-302): both EV stations (CHS.hpp evs_step), the hydrogen system
// (hydro_sys.py hy_step / use_cell), renewable netting, reward, done, fused auto-reset and the
// observation.  A warp therefore holds 32 independent hubs and every state array is laid out
// with the hub index fastest, so each per-pile access of a warp is one contiguous 512-byte row
// read or written with 128-bit vector loads/stores.  Action rows (row-major [H][A], what a policy
// emits) are read cooperatively: the warp walks its 32 rows with coalesced 128-byte loads and
// turns each row into a 64-bit on/off mask with __ballot_sync, so the per-pile loop never
// touches the action matrix.  Charger assignment ("first k empty piles", CHS.hpp:417-451) is a
// running count inside the single ascending pass; no list is built.
#define CHUB_KERNEL_TU
#include "chub_kernels.cuh"

cudaError_t chub_upload_const(const ChubConst &c) {
    ChubLogEntry tab[CHUB_LOG_N];
    chub_build_log_table(tab);
    cudaError_t r = cudaMemcpyToSymbol(kLogTab, tab, sizeof tab);
    if (r != cudaSuccess) return r;
    return cudaMemcpyToSymbol(kC, &c, sizeof(ChubConst));
}

struct StationOut {
    float mn, mx, now;  // min_power, max_power, charge_power (CHS.hpp:1257-1259)
    int number;         // car_number
    int line;           // queue length after assign_car
    int flow;           // flow_in_number.back()
    int incomplete;     // departures below target (the reference prints, CHS.hpp:916-921)
};

__device__ __forceinline__ float target_of(uint32_t tu) {  // uniform_rand(80, 100), CHS.hpp:41-42, 867
    return (float) tu / 999.0f * 20.0f + 80.0f;
}

// One EV station of one hub.  RESET = false: evs_step(vector) (CHS.hpp:1188-1207 / 1499-1518);
// RESET = true: evs_reset (CHS.hpp:1209-1231 / 1520-1542).
//
// Work is split so that the lanes of a warp (32 different hubs) execute the expensive code together:
//   scan   : ascending read of the packed words only -> bit masks (occupied, charging, departing)
//   loop A : each lane walks ITS OWN charging piles (car_step + urgency)            -- convergent fp64
//   loop B : each lane walks ITS OWN newly assigned piles in ascending order (add_car, serial RNG)
//   pass C : ascending coalesced sweep: non-charging bookkeeping (remove_car) and the station sums in
//            the reference's fp32 summation order (calculate_output)
// "First k empty piles" (assign_car / find_empty, CHS.hpp:417-451) is bit arithmetic on the occupancy mask.
template <bool RESET, int TYPE, int CP>
__device__ __forceinline__ StationOut station_pass(const ChubParams &P, int h, int slot0, int n, int t, int line,
                                                   uint32_t act_bits, HubRng &rng, uint32_t &err) {
    StationOut o;
    o.incomplete = 0;
    // ---- station-level draws of receive_car (CHS.hpp:1272-1306), before any per-car draw ----
    int in_car;
    if (RESET) {
        in_car = rng_init_cars(rng, n / 2);
        line = 0;
    } else {
        int N = (int) __ldg(&P.lut[t * 1000 + rng_u1000(rng)]);
        in_car = (int) roundf((TYPE == CHUB_FAST ? 0.15f : 0.1f) * 0.2f * (float) N);
    }
    int kept = 0;
    for (int w = 0; w < line; ++w) {  // reneging: keep if u > 0.1*logf(w+1)
        float a = rng_u01(rng);
        if (a > kC.renege[w]) kept += 1;
    }
    line = kept;
    int true_in = 0;
    for (int k = 0; k < in_car; ++k) {  // balking: stay if u <= expf(-0.01*(line+k)) and k <= n
        float a = rng_u01(rng);
        if (a <= kC.balk[min(line + k, 63)] && k <= n) true_in += 1;
    }
    o.flow = (TYPE == CHUB_SLOW) ? true_in : in_car;  // the slow/fast asymmetry, CHS.hpp:1306 vs 1617
    const int want = line + o.flow;

    uint4 *row = P.pile + (size_t) slot0 * P.Hpad + h;
    const size_t ld = (size_t) P.Hpad;
    const float *tlut = P.s2t_lut + (CP * 2 + TYPE) * 1000;  // s2t(target) for the 1000 possible targets
    const uint32_t nmask = n >= 32 ? 0xffffffffu : ((1u << n) - 1u);

    // ---- scan ----
    uint32_t car_mask = 0u, chg_mask = 0u, dep_mask = 0u;
    if (!RESET) {
#pragma unroll 4
        for (int i = 0; i < n; ++i) {
            const uint32_t w = row[i * ld].w;
            const uint32_t bit = 1u << i;
            if (w & PW_CAR) {
                car_mask |= bit;
                // judge_feasibility + assign_on_off_piece (CHS.hpp:1404-1413, 1364-1373)
                if ((w & PW_URGENT) || (act_bits & bit)) chg_mask |= bit;
                if ((int) PW_STAY(w) - ((int) PW_ALREADY(w) + 1) <= 0) dep_mask |= bit;  // remove_car
            }
        }
    }
    const uint32_t occ_after = car_mask & ~dep_mask;
    const int empties = n - __popc(occ_after);
    const int assigned = min(want, empties);  // assign_car, wait == true (CHS.hpp:417-430); may be negative
    o.line = min(want - assigned, 10);
    uint32_t add_mask = 0u;
    {
        uint32_t freeb = ~occ_after & nmask;
        for (int k = 0; k < assigned; ++k) {  // the first `assigned` empty piles, ascending
            uint32_t b = freeb & (0u - freeb);
            add_mask |= b;
            freeb ^= b;
        }
    }

    // ---- loop A: car_step for this hub's charging piles (CHS.hpp:900-905 / 1065-1070) ----
    if (!RESET) {
        uint32_t m = chg_mask;
        uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
        if (m) nxt = row[(__ffs(m) - 1) * ld];
        while (m) {
            const int i = __ffs(m) - 1;
            m &= m - 1u;
            uint4 r = nxt;
            if (m) nxt = row[(__ffs(m) - 1) * ld];  // prefetch this lane's next charging pile
            const uint32_t w = r.w;
            float soc, power;
            curve_step<TYPE, CP>(__uint_as_float(r.y) + 1.0f, soc, power);
            const float tsoc = curve_s2t<TYPE, CP>(soc);
            const uint32_t tu = PW_TU(w);
            const float needed = __ldg(&tlut[tu]) - tsoc;
            const int cn = needed > 0.0f ? (int) ceilf(needed) : 0;
            const int stay = PW_STAY(w), already = PW_ALREADY(w) + 1;
            if (stay - already <= 0) {
                if (ceilf(soc) < floorf(target_of(tu))) o.incomplete += 1;
                r = make_uint4(0u, 0u, 0u, 0u);
            } else {
                const bool urgent = cn > 0 && (stay - already) <= cn;
                r = make_uint4(__float_as_uint(soc), __float_as_uint(tsoc), __float_as_uint(power),
                               (uint32_t) stay | ((uint32_t) already << 6) | PW_CAR | PW_CHARGE |
                                   (urgent ? PW_URGENT : 0u) | ((uint32_t) cn << 15) | (tu << 20));
            }
            row[i * ld] = r;
        }
    }
    // ---- loop B: add_car for the newly assigned piles, ascending (CHS.hpp:864-877 / 1029-1042) ----
    {
        uint32_t m = add_mask;
        while (m) {
            const int i = __ffs(m) - 1;
            m &= m - 1u;
            const float arrive = rng_mk_soc(rng);                 // E
            const uint32_t tu = (uint32_t) rng_u1000(rng);        // U  (target = u/999*20+80)
            const float tsoc = curve_s2t<TYPE, CP>(arrive);
            const float needed = __ldg(&tlut[tu]) - tsoc;
            const int must_needed = (int) ceilf(needed);          // calculate_min_charging_time
            int stay = must_needed + rng_late_time(rng);          // E
            const float power = curve_t2p<TYPE, CP>(tsoc);
            int cn = needed > 0.0f ? must_needed : 0;
            if (stay > 63 || stay < 1 || cn > 31) {
                err |= CHUB_HUBERR_COUNTER;
                stay = min(max(stay, 1), 63);
                cn = min(cn, 31);
            }
            const bool urgent = cn > 0 && stay <= cn;
            row[i * ld] = make_uint4(__float_as_uint(arrive), __float_as_uint(tsoc), __float_as_uint(power),
                                     (uint32_t) stay | PW_CAR | (urgent ? PW_URGENT : 0u) | ((uint32_t) cn << 15) |
                                         (tu << 20));
            P.arrive[(size_t) (slot0 + i) * ld + h] = arrive;
        }
    }
    // ---- pass C: remaining bookkeeping + calculate_output in ascending fp32 order (CHS.hpp:1233-1261) ----
    float mn = 0.0f, mx = 0.0f, now = 0.0f;
    int number = 0;
    const uint32_t done_mask = chg_mask | add_mask;
    for (int i = 0; i < n; ++i) {
        const uint32_t bit = 1u << i;
        if (RESET && !(add_mask & bit)) {
            row[i * ld] = make_uint4(0u, 0u, 0u, 0u);
            continue;
        }
        if (!RESET && !((car_mask | add_mask) & bit)) continue;  // stays empty: nothing to read
        uint4 r = row[i * ld];
        uint32_t w = r.w;
        if (!(done_mask & bit)) {  // occupied, not charging this step: remove_car only
            const int stay = PW_STAY(w), already = PW_ALREADY(w) + 1;
            if (dep_mask & bit) {
                if (ceilf(__uint_as_float(r.x)) < floorf(target_of(PW_TU(w)))) o.incomplete += 1;
                w = 0u;
                r = make_uint4(0u, 0u, 0u, 0u);
            } else {
                const int cn = PW_CN(w);
                const bool urgent = cn > 0 && (stay - already) <= cn;
                w = (w & ~(PW_CHARGE | PW_URGENT | (0x3fu << 6))) | ((uint32_t) already << 6) | (urgent ? PW_URGENT : 0u);
                r.w = w;
            }
            row[i * ld] = r;
        }
        if (w & PW_CAR) {
            const float power = __uint_as_float(r.z);
            number += 1;
            mx += power;
            if (w & PW_URGENT) mn += power;
            if (w & PW_CHARGE) now += power;
        }
    }
    o.mn = mn; o.mx = mx; o.now = now; o.number = number;
    return o;
}

template <bool RESET>
__device__ __forceinline__ StationOut station_any(const ChubParams &P, int h, int slot0, int n, int type, int cp, int t,
                                                  int line, uint32_t act_bits, HubRng &rng, uint32_t &err) {
    if (cp) {  // constant-power charging (constant_charging=True): cold instantiations
        if (type == CHUB_SLOW) return station_pass<RESET, CHUB_SLOW, 1>(P, h, slot0, n, t, line, act_bits, rng, err);
        return station_pass<RESET, CHUB_FAST, 1>(P, h, slot0, n, t, line, act_bits, rng, err);
    }
    if (type == CHUB_SLOW) return station_pass<RESET, CHUB_SLOW, 0>(P, h, slot0, n, t, line, act_bits, rng, err);
    return station_pass<RESET, CHUB_FAST, 0>(P, h, slot0, n, t, line, act_bits, rng, err);
}

// ---------------------------------------------------------------------------------------------
// hydrogen side (hydro_sys.py), fp64 like the reference's Python floats; every expression keeps
// Python's evaluation order and -fmad=false keeps each multiply/add separately rounded.

// x / D for a literal D: exactly the IEEE quotient (Markstein correction), 3 instructions instead of ~25
#define DIVC(x, D) chub_div_const((x), (D), 1.0 / (D))

__device__ __forceinline__ double py_max(double a, double b) { return b > a ? b : a; }  // Python max(a, b)
__device__ __forceinline__ double py_min(double a, double b) { return b < a ? b : a; }  // Python min(a, b)

__device__ __forceinline__ double el_power(const ChubClass &c, double m) {  // Electrolyser.get_power, hydro_sys.py:38-48
    if (c.el_cells == 0.0) return 0.0;
    double v_H_mass = chub_div_const(m, c.el_cells, c.r_el_cells);
    double v_H_mol = DIVC(v_H_mass, 2.02);
    double v_H_L = v_H_mol * c.v_M;
    double v_H = v_H_L * 1000.0 * 60.0;
    double temp = chub_div_const(v_H * 2.0 * 96487.0, c.el_den, c.r_el_den);
    double power = (temp * temp) * 0.326 + temp * 1.476;
    return DIVC(c.el_cells * power, 1000.0);
}
__device__ __forceinline__ double cpr_power(const ChubClass &c, double m) {  // Compressor.generate_W, hydro_sys.py:74-82
    double n = DIVC(m, 2.02);
    return DIVC(DIVC(n * c.cpr_w12, 0.8), 1000.0);
}

// SAE J2601 table: initial pressure -> target pressure (hydro_sys.py:338-388)
__device__ __forceinline__ double j2601_target(double p) {
    if (p < kC.j_x[1]) return kC.j_y[0];
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        bool hit = (a < 7) ? (kC.j_x[1 + a] <= p && p < kC.j_x[2 + a]) : (kC.j_x[1 + a] <= p && p <= kC.j_x[2 + a]);
        if (hit) return kC.j_slope[a] * (p - kC.j_x[1 + a]) + kC.j_y[1 + a];
    }
    return p;
}
// one arriving FCEV: fueling minutes and grams (hydro_sys.py:259-265, 308-321)
__device__ __forceinline__ void fcev_need(float soc_f, double &minutes, double &grams) {
    double soc = (double) soc_f;
    if (soc < 0.5) soc = 0.5;
    double p0 = (soc * 0.01) * 70.0;
    double tp = j2601_target(p0);
    if (p0 < 5.0) minutes = DIVC(69.0 - p0, 18.5) + kC.topoff_time;
    else minutes = DIVC(tp - p0, 18.5);
    grams = kC.fcev_cap * DIVC(tp, 70.0) - kC.fcev_cap * DIVC(p0, 70.0);
}

// HyFCEVStation.hvs_step with a backlog or several arrivals (hydro_sys.py:266-285); rare, kept
// out of line so its local arrays do not cost the common path registers.
__device__ __noinline__ double fcev_general(const ChubParams &P, int h, int arrivals, int &qn, int &fline,
                                            HubRng &rng, uint32_t &err) {
    float qs[CHUB_FCEV_QCAP];
    double tm[CHUB_FCEV_QCAP], ms[CHUB_FCEV_QCAP];
    int n = qn;
    for (int j = 0; j < n; ++j) {
        qs[j] = P.qsoc[(size_t) j * P.Hpad + h];
        fcev_need(qs[j], tm[j], ms[j]);
    }
    for (int k = 0; k < arrivals; ++k) {
        float s = rng_mk_soc(rng);
        if (n < CHUB_FCEV_QCAP) {
            qs[n] = s;
            fcev_need(s, tm[n], ms[n]);
            n += 1;
        } else {
            err |= CHUB_HUBERR_FCEV_OVERFLOW;
        }
    }
    double total_time = 0.0, total_mass = 0.0;
    for (int j = 0; j < n; ++j) total_time += tm[j];
    for (int j = 0; j < n; ++j) total_mass += ms[j];
    if (total_time > 15.0) {
        int num = 0;
        for (int i = 1; i <= arrivals - 1; ++i) {
            int keep = max(n - i, 0);
            double s = 0.0;
            for (int j = 0; j < keep; ++j) s += tm[j];
            if (s <= 15.0) {
                fline = i;
                num = keep;
                break;
            }
        }
        for (int j = num; j < n; ++j) P.qsoc[(size_t) (j - num) * P.Hpad + h] = qs[j];
        n -= num;
    } else {
        fline = 0;
        n = 0;
    }
    qn = n;
    return total_mass;
}

struct HubVars {  // per-hub registers that live across the stages
    HubRng rng;
    int t, line0, line1, qn, fline, farr, flow0, flow1;
    double store;
    uint32_t err;
};

__device__ __forceinline__ void hub_load(const ChubParams &P, int h, HubVars &v) {
    uint4 a = P.hubA[h];
    uint2 k = __ldg(&P.keys[h]);
    v.rng.k0 = k.x; v.rng.k1 = k.y; v.rng.lcg = a.x; v.rng.uctr = a.y;
    v.t = a.z & 0xff; v.line0 = (a.z >> 8) & 0xff; v.line1 = (a.z >> 16) & 0xff; v.qn = a.z >> 24;
    v.flow0 = (int) (a.w & 0xff) - 128; v.flow1 = (int) ((a.w >> 8) & 0xff) - 128;
    v.fline = (a.w >> 16) & 0xff; v.farr = a.w >> 24;
    v.store = P.store[h];
    v.err = 0u;
}
__device__ __forceinline__ void hub_store(const ChubParams &P, int h, const HubVars &v) {
    uint32_t z = (uint32_t) v.t | ((uint32_t) v.line0 << 8) | ((uint32_t) v.line1 << 16) | ((uint32_t) v.qn << 24);
    uint32_t w = (uint32_t) ((v.flow0 + 128) & 0xff) | ((uint32_t) ((v.flow1 + 128) & 0xff) << 8) |
                 ((uint32_t) (v.fline & 0xff) << 16) | ((uint32_t) (v.farr & 0xff) << 24);
    P.hubA[h] = make_uint4(v.rng.lcg, v.rng.uctr, z, w);
    P.store[h] = v.store;
    if (v.err) P.err[h] |= v.err;
}

__device__ __forceinline__ void trace_at(const ChubParams &P, int h, long long k, double &pv, double &wd, double &pr) {
    size_t i = (size_t) (k % (long long) P.trace_len);
    if (P.trace_per_hub) i = i * (size_t) P.Hpad + h;
    pv = __ldg(&P.pv[i]); wd = __ldg(&P.wd[i]); pr = __ldg(&P.price_tr[i]);
}

// observation, evcssp_manager.py:318-342 (state_norm) over make_state's real_state
__device__ __forceinline__ void write_obs(const ChubParams &P, int h, const ChubClass &c, int t, const StationOut &s0,
                                          const StationOut &s1, double store_soc, double pv, double wd, double pr) {
    float o[CHUB_OBS_DIM];
#pragma unroll
    for (int j = 0; j < CHUB_OBS_DIM; ++j) o[j] = 0.0f;
    int k = 0;
    o[k++] = (float) kC.time_sin[t];
    o[k++] = (float) chub_div_const(pr - kC.price_mean, kC.price_std, kC.r_price_std);
    if (c.n[0] > 0) {
        o[k++] = (float) chub_div_const((double) s0.mn - c.half[0], c.half[0], c.r_half[0]);
        o[k++] = (float) chub_div_const((double) s0.now - c.half[0], c.half[0], c.r_half[0]);
        o[k++] = (float) chub_div_const((double) s0.mx - c.half[0], c.half[0], c.r_half[0]);
        o[k++] = (float) DIVC((double) s0.line, 5.0);
    }
    if (c.n[1] > 0) {
        o[k++] = (float) chub_div_const((double) s1.mn - c.half[1], c.half[1], c.r_half[1]);
        o[k++] = (float) chub_div_const((double) s1.now - c.half[1], c.half[1], c.r_half[1]);
        o[k++] = (float) chub_div_const((double) s1.mx - c.half[1], c.half[1], c.r_half[1]);
        o[k++] = (float) DIVC((double) s1.line, 5.0);
    }
    o[k++] = (float) store_soc;
    o[k++] = (float) DIVC(pv, 210.0);
    o[k++] = (float) DIVC(wd, 92.0);
    float *dst = P.obs + (size_t) h * CHUB_OBS_DIM;
#pragma unroll
    for (int j = 0; j < CHUB_OBS_DIM; ++j) dst[j] = o[j];
}

// EvcsspManagerEnv_v6.reset (evcssp_manager.py:304-316) of one hub, from and to HBM: both stations
// (evs_reset), hydrogen store, FCEV queue, episode statistics, first observation of the new episode.
// Out of line on purpose: the step kernel calls it once per hub per 96 steps (fused auto-reset).
__device__ __noinline__ void hub_reset_full(const ChubParams &P, int h, long long k_obs) {
    const int ci = P.cls ? __ldg(&P.cls[h]) : 0;
    const ChubClass &c = P.classes[ci];
    HubVars v;
    hub_load(P, h, v);
    StationOut s0 = station_any<true>(P, h, 0, c.n[0], c.type[0], c.constant, 0, 0, 0u, v.rng, v.err);
    StationOut s1 = station_any<true>(P, h, P.S0, c.n[1], c.type[1], c.constant, 0, 0, 0u, v.rng, v.err);
    v.t = 0;
    v.line0 = s0.line; v.line1 = s1.line;
    v.flow0 = s0.flow; v.flow1 = s1.flow;
    v.qn = 0; v.fline = 0; v.farr = 0;
    v.store = c.init_mass;
    P.stats[h] = make_float4(0.f, 0.f, 0.f, 0.f);
    double pv, wd, pr;
    trace_at(P, h, k_obs, pv, wd, pr);
    write_obs(P, h, c, 0, s0, s1, c.init_soc, pv, wd, pr);
    if (P.diag) {
        const size_t s = (size_t) P.Hpad;
        float *q = P.st_power + h;
        q[0 * s] = s0.mn; q[1 * s] = s0.now; q[2 * s] = s0.mx; q[3 * s] = s1.mn; q[4 * s] = s1.now; q[5 * s] = s1.mx;
    }
    hub_store(P, h, v);
}

// ---------------------------------------------------------------------------------------------
// Cooperative read of the warp's 32 action rows.  Each row is loaded with coalesced 128-byte
// requests (lane = column) and binarised exactly like action_to_real on a float32 vector
// (evcssp_manager.py:384-404): on iff (a + 1)/2 >= 0.5 in fp32.  __ballot_sync packs a row into
// masks that lane r keeps; the two analog entries come back by shuffle.
__device__ __forceinline__ void load_actions(const ChubParams &P, int h, bool active, int na,
                                             unsigned long long &bits, float &a_m2, float &a_m1) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int h0 = h - lane;
    bits = 0ull; a_m2 = 0.0f; a_m1 = 0.0f;
    const int my_na = active ? na : -2;  // -2: no hub in this lane -> every column predicate is false
    // eight rows per batch: all their loads are issued before the first value is used, so the warp
    // pays four memory round trips for its 32 rows instead of one per row and chunk
    for (int r0 = 0; r0 < 32; r0 += 8) {
        float v[8][3];
        int rna[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            rna[j] = __shfl_sync(full, my_na, r0 + j);
            const float *rowp = P.actions + (size_t) (h0 + r0 + j) * P.action_stride;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const int idx = c * 32 + lane;
                v[j][c] = (idx < rna[j] + 2) ? __ldg(rowp + idx) : -1.0f;
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int n = rna[j];
            const unsigned m0 = __ballot_sync(full, lane < n && (v[j][0] + 1.0f) / 2.0f >= 0.5f);
            const unsigned m1 = __ballot_sync(full, 32 + lane < n && (v[j][1] + 1.0f) / 2.0f >= 0.5f);
            const int c2 = n >> 5, c1 = (n + 1) >> 5;  // chunks holding action[-2] and action[-1]
            const float t2 = c2 == 0 ? v[j][0] : (c2 == 1 ? v[j][1] : v[j][2]);
            const float t1 = c1 == 0 ? v[j][0] : (c1 == 1 ? v[j][1] : v[j][2]);
            const float e2 = __shfl_sync(full, t2, n & 31);
            const float e1 = __shfl_sync(full, t1, (n + 1) & 31);
            if (lane == r0 + j) {
                bits = (unsigned long long) m0 | ((unsigned long long) m1 << 32);
                a_m2 = e2;  // action[-2] -> electrolyser set-point (the reference swaps the last two)
                a_m1 = e1;  // action[-1] -> fuel-cell set-point
            }
        }
    }
}

template <int BLOCK, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) chub_step_kernel(const __grid_constant__ ChubParams P) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = h < P.H;
    const int ci = (active && P.cls) ? __ldg(&P.cls[h]) : 0;
    const ChubClass &c = P.classes[ci];
    const int n0 = c.n[0], n1 = c.n[1];
    unsigned long long bits;
    float a_m2, a_m1;
    load_actions(P, h, active, n0 + n1, bits, a_m2, a_m1);

    HubVars v;
    v.err = 0u; v.t = 0; v.line0 = v.line1 = 0;
    if (active) hub_load(P, h, v);
    const long long k = *P.step_index + P.step_offset;
    double pv_t = 0.0, wd_t = 0.0, pr_t = 0.0;
    if (active) trace_at(P, h, k, pv_t, wd_t, pr_t);

    // ---- the two EV stations, station 0 first (Aggregator_Simple.py:141-142) ----
    // phase_sync keeps the warps of a block inside the same code region (instruction-cache locality)
    StationOut s0 = {}, s1 = {};
    if (P.phase_sync & 1) __syncthreads();
    if (active && !(P.phase_sync & 2)) s0 = station_any<false>(P, h, 0, n0, c.type[0], c.constant, v.t, v.line0, (uint32_t) bits, v.rng, v.err);
    if (P.phase_sync & 1) __syncthreads();
    if (active && !(P.phase_sync & 4)) s1 = station_any<false>(P, h, P.S0, n1, c.type[1], c.constant, v.t, v.line1,
                                        (uint32_t) (bits >> n0), v.rng, v.err);
    if (P.phase_sync & 1) __syncthreads();
    if (!active) return;
    if (P.phase_sync & 8) {  // timing ablation only: stations without the hub section
        hub_store(P, h, v);
        return;
    }
    v.line0 = s0.line; v.line1 = s1.line;
    v.flow0 = s0.flow; v.flow1 = s1.flow;

    // ---- electrolyser set-point under the grid limit (evcssp_manager.py:157-180) ----
    const double cp0 = (double) s0.now, cp1 = (double) s1.now;
    const double charging = 0.0 + cp0 + cp1;
    double renew = wd_t + pv_t;
    const double limit = py_max(0.0, 2000.0 + renew - charging);
    const double a_el = (double) ((a_m2 + 1.0f) / 2.0f);
    const double a_fc = (double) ((a_m1 + 1.0f) / 2.0f);
    int idx = (int) ceil(a_el * 100.0);
    idx = min(max(idx, 0), CHUB_EL_TABLE - 1);
    const double *tab = P.el_table + (size_t) ci * CHUB_EL_TABLE;
    double act = a_el;
    if (__ldg(&tab[idx]) > limit) {
        act = 0.01 * 100.0;  // hy_power_speed_list_input[-1]
        for (int ind = 0; ind < CHUB_EL_TABLE; ++ind) {
            if (__ldg(&tab[ind]) >= limit) {
                if (ind > 0) act = 0.01 * (double) (ind - 1);  // ind == 0 picks input[-1] (Python index -1)
                break;
            }
        }
    }

    // ---- HySystem.hy_step (hydro_sys.py:160-195) ----
    // FCEV arrivals first: 1 U draw, then one E normal per arriving car (hydro_sys.py:253-265)
    double demand = 0.0;
    {
        int N = (int) __ldg(&P.lut[v.t * 1000 + rng_u1000(v.rng)]);
        float pin = c.possible_in_f > 1.0f ? 0.3f : c.possible_in_f;
        float per = c.perm_f > 1.0f ? 0.01f : c.perm_f;
        int arrivals = (int) roundf(pin * per * (float) N);
        v.farr = arrivals;
        if (v.qn == 0 && arrivals <= 1) {
            if (arrivals == 1) {
                double minutes, grams;
                fcev_need(rng_mk_soc(v.rng), minutes, grams);  // one car always fits the 15-minute window
                demand = 0.0 + grams;
            }
            v.fline = 0;
        } else {
            demand = fcev_general(P, h, arrivals, v.qn, v.fline, v.rng, v.err);
        }
    }
    const double must = py_max(0.0, c.cap_mass * 0.1 - v.store);
    const double upper = py_max(c.cap_mass - v.store, 0.0);
    const double charge_temp = py_max(py_min(act * c.v_h_max * 900.0, upper), must);
    double flow = DIVC(charge_temp, 900.0);
    flow = py_min(flow, c.v_h_max);
    const double ele_kw = el_power(c, flow);
    const double cpr_kw = cpr_power(c, flow);
    // HyStore.sty_step (hydro_sys.py:104-126)
    v.store += flow * 15.0 * 60.0;
    if (c.cap_mass - v.store < 0.0) v.err |= CHUB_HUBERR_STORE_ASSERT;
    double lower = py_max(0.0, v.store - 0.1 * c.cap_mass);
    const double hy_use = py_min(lower, demand);
    const double not_meet = demand - hy_use;
    v.store -= hy_use;
    v.store -= v.store * c.loss;
    double store_soc = chub_div_const(v.store, c.cap_mass, c.r_cap_mass);
    const double h2_kw = ele_kw + cpr_kw;
    const bool gen_hy = flow > 0.5;

    // ---- renewables cover the electrolyser first, then the EV load (evcssp_manager.py:183-213) ----
    double h2_grid = h2_kw, ev_sum = charging, l0 = cp0, l1 = cp1;
    if (renew >= h2_grid) {
        renew -= h2_grid;
        h2_grid = 0.0;
        ev_sum = py_max(0.0, ev_sum - renew);
        const double s = 0.0 + l0 + l1;
        if (s > 0.0) {
            const double rate = chub_div(ev_sum, s);
            l0 = rate * l0;
            l1 = rate * l1;
        }
    } else {
        h2_grid -= renew;
    }
    // ---- fuel cell (evcssp_manager.py:216-227, hydro_sys.py:409-430) ----
    double fc = gen_hy ? 0.0 : c.fc_cells * a_fc;
    if (fc > c.fc_cells) fc = c.fc_cells;
    else if (fc < 0.0) fc = 0.0;
    else if (fc > ev_sum) fc = ev_sum;
    double hy_fc = DIVC(fc * 1500.0, 119.6);
    hy_fc = py_min(v.store, hy_fc);
    v.store -= hy_fc;
    const double fc_kw = py_max(0.0, fc);
    ev_sum -= fc_kw;
    {
        const double s = 0.0 + l0 + l1;
        if (s > 0.0) {
            const double rate = chub_div(ev_sum, s);
            l0 = rate * l0;
            l1 = rate * l1;
        }
    }
    // ---- income and reward (evcssp_manager.py:234-269) ----
    const double hy_loss = -6.0 / 1000.0 * hy_fc;
    const double p4 = pr_t * 0.25;
    const double inc_fast = 0.42 / 4.0 * cp0 - p4 * l0;
    const double inc_slow = 0.21 / 4.0 * cp1 - p4 * l1;
    const double inc_evs = inc_fast + inc_slow;
    const double serve = 0.8 * (double) (s0.flow + s1.flow);
    const double inc_hys = 6.0 / 1000.0 * hy_use;
    const double nm_loss = -10.0 / 1000.0 * not_meet;
    const double hy_cost = -p4 * h2_grid;
    const double income = inc_hys + inc_evs + serve + hy_cost;
    const double reward = DIVC(inc_hys + inc_evs + serve + hy_cost + 1.0 * hy_loss + nm_loss, 50.0);
    const bool done = v.t + 1 >= CHUB_DAY;
    v.t = (v.t + 1) % CHUB_DAY;

    if (P.diag) {  // fp64 columns for the parity tests
        double *d = P.diag + h;
        const size_t s = (size_t) P.Hpad;
        d[CHUB_D_REWARD * s] = reward;   d[CHUB_D_INCOME * s] = income;   d[CHUB_D_HY_FLOW * s] = flow;
        d[CHUB_D_H2_POWER * s] = h2_kw;  d[CHUB_D_FC_POWER * s] = fc_kw;  d[CHUB_D_HY_USE * s] = hy_use;
        d[CHUB_D_NOT_MEET * s] = not_meet; d[CHUB_D_HY_TO_USE * s] = hy_fc; d[CHUB_D_DEMAND * s] = demand;
        d[CHUB_D_EV_GRID0 * s] = l0;     d[CHUB_D_EV_GRID1 * s] = l1;     d[CHUB_D_H2_GRID * s] = h2_grid;
        d[CHUB_D_HY_ACT * s] = act;      d[CHUB_D_STORE_SOC * s] = store_soc;
        float *q = P.st_power + h;
        q[0 * s] = s0.mn; q[1 * s] = s0.now; q[2 * s] = s0.mx; q[3 * s] = s1.mn; q[4 * s] = s1.now; q[5 * s] = s1.mx;
    }

    // ---- episode statistics (evcssp_manager.py:259-269, CHS.hpp:916-921) ----
    float4 st = P.stats[h];
    st.x += (float) reward;
    st.y += (float) income;
    st.z += (float) (l0 + l1 + h2_grid);
    st.w += (float) (s0.incomplete + s1.incomplete);
    if (done) {
        P.last_stats[h] = st;
        st = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    P.stats[h] = st;
    P.reward[h] = (float) reward;
    P.done[h] = done ? 1 : 0;
    hub_store(P, h, v);
    if (done && P.auto_reset) {  // fused auto-reset: the returned obs is the first one of the next episode
        hub_reset_full(P, h, k + 1);
    } else {
        double pv_n, wd_n, pr_n;
        trace_at(P, h, k + 1, pv_n, wd_n, pr_n);
        write_obs(P, h, c, v.t, s0, s1, store_soc, pv_n, wd_n, pr_n);
    }
}

__global__ void __launch_bounds__(128) chub_reset_kernel(const __grid_constant__ ChubParams P) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= P.H) return;
    if (P.reset_mask && !P.reset_mask[h]) return;
    P.reward[h] = 0.0f;
    P.done[h] = 0;
    if (P.diag) {
        const int ci = P.cls ? __ldg(&P.cls[h]) : 0;
        const size_t s = (size_t) P.Hpad;
        for (int j = 0; j < CHUB_D_COUNT; ++j) P.diag[j * s + h] = 0.0;
        P.diag[CHUB_D_STORE_SOC * s + h] = P.classes[ci].init_soc;
    }
    hub_reset_full(P, h, *P.step_index);
}

// s2t(target) for the 1000 possible target SOCs x {fast, slow} x {curve, constant power}
__global__ void chub_s2t_lut_kernel(float *out) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 4000) return;
    const int cp = idx / 2000, type = (idx / 1000) % 2, u = idx % 1000;
    out[idx] = curve_s2t_rt(type, cp, target_of((uint32_t) u));
}
void chub_launch_s2t_lut(float *out, cudaStream_t s) { chub_s2t_lut_kernel<<<(4000 + 127) / 128, 128, 0, s>>>(out); }

__global__ void chub_tick_kernel(long long *step_index, int n) { *step_index += n; }

#define CHUB_LAUNCH(B, M) chub_step_kernel<B, M><<<(p.H + (B) - 1) / (B), B, 0, s>>>(p)
void chub_launch_step(const ChubParams &p, cudaStream_t s) {
    switch (p.block) {  // block size * 100 + min blocks per SM (register budget = 65536 / (block * minb))
        case 12808: CHUB_LAUNCH(128, 8); break;
        case 12810: CHUB_LAUNCH(128, 10); break;
        case 12812: CHUB_LAUNCH(128, 12); break;
        case 25604: CHUB_LAUNCH(256, 4); break;
        case 25605: CHUB_LAUNCH(256, 5); break;
        case 25606: CHUB_LAUNCH(256, 6); break;
        case 51203: CHUB_LAUNCH(512, 3); break;
        case 102401: CHUB_LAUNCH(1024, 1); break;
        case 76801: CHUB_LAUNCH(768, 1); break;
        default: CHUB_LAUNCH(512, 2); break;
    }
}
void chub_launch_reset(const ChubParams &p, cudaStream_t s) {
    const int block = 128;
    chub_reset_kernel<<<(p.H + block - 1) / block, block, 0, s>>>(p);
}
void chub_launch_tick(long long *step_index, int n, cudaStream_t s) { chub_tick_kernel<<<1, 1, 0, s>>>(step_index, n); }
