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
#include <algorithm>
#include "chub_kernels.cuh"
#include "../../include/chub_exo.h"

cudaError_t chub_upload_const(const ChubConst &c) {
    ChubLogEntry tab[CHUB_LOG_N];
    chub_build_log_table(tab);
    cudaError_t r = cudaMemcpyToSymbol(kLogTab, tab, sizeof tab);
    if (r != cudaSuccess) return r;
    return cudaMemcpyToSymbol(kC, &c, sizeof(ChubConst));
}

// Self-checking build (-DCHUB_CHECKS -> libchub_checked.so): compute-sanitizer is not available on the GPU pool this engine
// is developed on, so the invariants it would watch -- every index into the per-warp shared-memory rings, the posted-power
// rows, the warp's pile tile and the unit queue stays inside its array -- are asserted by the kernel itself and counted;
// tests/test_gpu_checked_build.py runs the parity scenarios through that build and wants zero.  Compiled out otherwise.
__device__ unsigned long long kCheckFails;
#ifdef CHUB_CHECKS
#define CHUB_CHECK(cond) do { if (!(cond)) atomicAdd(&kCheckFails, 1ull); } while (0)
#else
#define CHUB_CHECK(cond) do { } while (0)
#endif
cudaError_t chub_read_check_failures(unsigned long long *out) { return cudaMemcpyFromSymbol(out, kCheckFails, sizeof *out); }
int chub_checks_compiled_in() {
#ifdef CHUB_CHECKS
    return 1;
#else
    return 0;
#endif
}

// @region misc
struct StationOut {
    float mn, mx, now;  // min_power, max_power, charge_power (CHS.hpp:1257-1259)
    int number;         // car_number
    int line;           // queue length after assign_car
    int flow;           // flow_in_number.back()
    int incomplete;     // departures below target (the reference prints, CHS.hpp:916-921)
    uint32_t occ;       // occupancy mask after the call
};

__device__ __forceinline__ float target_of(uint32_t tu) {  // uniform_rand(80, 100), CHS.hpp:41-42, 867
    // tu / 999 correctly rounded without the generic division: reciprocal multiply + one residual correction, equal to
    // the IEEE quotient for every tu in 0..1023 (tools/verify_divf.c checks all of them)
    const float t = (float) tu, q0 = t * (1.0f / 999.0f);
    const float q = fmaf(fmaf(-q0, 999.0f, t), 1.0f / 999.0f, q0);
    return q * 20.0f + 80.0f;
}

// evs_reset of one EV station of one hub (CHS.hpp:1209-1231 / 1520-1542), one thread per hub: draw the
// initial car count, fill the first piles in ascending order (add_car, serial E/U/E draws per car), clear the
// rest, and form the station sums.  Runs once per hub per episode (reset kernel, fused auto-reset).
// @region station_reset
template <int TYPE, int CP>
__device__ __forceinline__ StationOut station_reset(const ChubParams &P, int h, int slot0, int n, HubRng &rng,
                                                    uint32_t &err) {
    StationOut o;
    o.incomplete = 0;
    const int in_car = rng_init_cars(rng, n / 2);
    int true_in = 0;
    for (int k = 0; k < in_car; ++k) {  // balking against an empty line (CHS.hpp:1294-1304)
        if (rng_u1000(rng) <= (int) kC.balk_k[min(k, 63)] && k <= n) true_in += 1;
    }
    o.flow = (TYPE == CHUB_SLOW) ? true_in : in_car;  // the slow/fast asymmetry, CHS.hpp:1306 vs 1617
    const int assigned = min(o.flow, n);
    o.line = min(o.flow - assigned, 10);
    const size_t at0 = chub_pile_index(h, slot0, P.S0 + P.S1);
    uint4 *row = P.pile + at0;
    const size_t ld = 32;
    const float *tlut = P.s2t_lut + (CP * 2 + TYPE) * 1000;
    float mx = 0.0f, mn = 0.0f;
    for (int i = 0; i < n; ++i) {
        if (i >= assigned) {
            row[i * ld] = make_uint4(0u, 0u, 0u, 0u);
            continue;
        }
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
                                 (uint32_t) stay | PW_CAR | (urgent ? PW_URGENT : 0u) | ((uint32_t) cn << 15) | (tu << 20));
        P.arrive[at0 + (size_t) i * ld] = arrive;
        mx += power;
        if (urgent) mn += power;
    }
    o.mn = mn; o.mx = mx; o.now = 0.0f; o.number = max(assigned, 0);
    o.occ = assigned >= 32 ? 0xffffffffu : ((1u << max(assigned, 0)) - 1u);
    return o;
}

__device__ __forceinline__ StationOut station_reset_any(const ChubParams &P, int h, int slot0, int n, int type, int cp,
                                                        HubRng &rng, uint32_t &err) {
    if (cp) {  // constant-power charging (constant_charging=True)
        if (type == CHUB_SLOW) return station_reset<CHUB_SLOW, 1>(P, h, slot0, n, rng, err);
        return station_reset<CHUB_FAST, 1>(P, h, slot0, n, rng, err);
    }
    if (type == CHUB_SLOW) return station_reset<CHUB_SLOW, 0>(P, h, slot0, n, rng, err);
    return station_reset<CHUB_FAST, 0>(P, h, slot0, n, rng, err);
}

// ---------------------------------------------------------------------------------------------
// evs_step(vector) of one EV station (CHS.hpp:1188-1207 / 1499-1518) for the 32 hubs of a warp.
//
// A thread owns a hub, but the expensive per-car arithmetic is NOT done hub by hub: the number of
// charging piles and of arriving cars differs from hub to hub, and a per-lane loop leaves most lanes
// idle while the busiest hub finishes (measured: 18 of 32 lanes active).  Instead the warp pools the
// work of its 32 hubs in shared memory and runs it 32 items at a time:
//   scan     : lanes = hubs, ascending coalesced read of the pile rows.  Non-charging cars are finished on
//              the spot (remove_car bookkeeping); charging piles are appended to one of two work lists,
//              by the curve segment their new charge time falls in, so a batch runs ONE curve body.
//   flat A   : whenever a list holds 32 items the warp executes car_step + urgency for them
//              (lanes = piles of any hub), writes the pile records and posts power / urgency to the owner.
//   chain    : the serial part of add_car -- only the minstd accept/reject walk of the two polar normals
//              per car (the variates' values never influence how many draws are consumed) -- as a
//              per-lane state machine, one ATTEMPT per iteration, recording the minstd state each accepted
//              attempt started from.  The double-precision attempt is decided from an fp32 estimate of r2
//              (exact evaluation only inside a +-4e-6 band around the accept boundary).
//   flat B   : the variates themselves (log/sqrt), curves and record packing of the new cars, lanes = new cars
//              of any hub.
//   sums     : each owner adds its piles' power in the reference's ascending fp32 order (calculate_output).
// Per-warp scratch lives in shared memory (WarpScratch); warps never synchronise with each other.
// @region scratch
#define LA_RING 64    // >= 31 waiting + 32 appended per scanned row; a batch is up to 32 items
#define CHUB_SMEM_TABLES (CHUB_LOG_N * 16 + 8000)   // log table + s2t(target) tables ahead of the warps' scratch
template <int Q>
struct WarpScratch {
    float am[2][32];                  // the two analog action entries of each hub, parked here during the stations
    // ---- everything below doubles as the staging area of the warp's action rows ----
    uint32_t park[2][5][32];          // the stations' outputs, parked until the hub section (registers are scarce)
    uint32_t inc_s[32], err_s[32];    // posted to the owner lane: incomplete departures found by a car_step batch, error bits
    uint32_t la_t[2][LA_RING], la_w[2][LA_RING], la_id[2][LA_RING];  // car_step rings per curve segment: new charge time
                                      // (f32 bits), packed pile word, lane | pile << 5
    uint32_t pw[1][32];               // [pile][lane], max(S0, S1) rows in fact: power (f32 bits, >= 0) of the car now on that
                                      // pile; sign bit = urgent
};
// The warp's scratch is addressed through ONE 32-bit offset into chub_smem that stays in a register: a C++ reference
// to the struct made the compiler re-derive the shared-window address (S2R tid, S2UR cta id, LDC ws_bytes, IMAD) at
// almost every use inside the scan loop.
template <int Q>
struct Scr {
    uint32_t base;
    int lane;   // threadIdx.x & 31, read from the special register once
    __device__ __forceinline__ WarpScratch<Q> *operator->() const {
        return reinterpret_cast<WarpScratch<Q> *>(__cvta_shared_to_generic((size_t) base));
    }
};
template <int Q>
static size_t warp_scratch_bytes(int rows) { return offsetof(WarpScratch<Q>, pw) + (size_t) rows * 128; }
#define PWB_URGENT 0x80000000u

// @region flat A shell (car_step batch)
template <int TYPE, int CP, int SEG, int Q>
__device__ __forceinline__ void flat_car_step(const ChubParams &P, Scr<Q> S, const float *tlut, uint4 *pile0,
                                              int head, int count) {
    const int lane = S.lane;
    __syncwarp();
    if (lane < count) {
        const int sl = (head + lane) & (LA_RING - 1);
        const float tt = __uint_as_float(S->la_t[SEG][sl]);
        const uint32_t w = S->la_w[SEG][sl];
        const uint32_t id = S->la_id[SEG][sl], owner = id & 31u;
        CHUB_CHECK(count >= 1 && count <= 32 && (uint32_t) sl < LA_RING && id < 32u * (uint32_t) max(P.S0, P.S1));
        float soc, power;
        curve_step<TYPE, CP, SEG>(tt, soc, power);            // car_step, CHS.hpp:900-905 / 1065-1070
        const float tsoc = curve_s2t<TYPE, CP>(soc);
        const uint32_t tu = PW_TU(w);
        const float needed = tlut[tu] - tsoc;                 // calculate_needed, CHS.hpp:879-898
        const int cn = needed > 0.0f ? (int) ceilf(needed) : 0;
        const int left = (int) PW_STAY(w) - (int) PW_ALREADY(w) - 1;
        uint4 r = make_uint4(0u, 0u, 0u, 0u);
        if (left <= 0) {                                      // remove_car, CHS.hpp:912-930
            if (ceilf(soc) < floorf(target_of(tu))) atomicAdd(&S->inc_s[owner], 1u);
        } else {
            const uint32_t urgent = (cn > 0 && left <= cn) ? 1u : 0u;
            r = make_uint4(__float_as_uint(soc), __float_as_uint(tsoc), __float_as_uint(power),
                           ((w + (1u << 6)) & 0xfffu) | PW_CAR | PW_CHARGE | (urgent << 14) | ((uint32_t) cn << 15) |
                               (w & (0x3ffu << 20)));
            (&S->pw[0][0])[id] = __float_as_uint(power) | (urgent << 31);   // id = pile * 32 + owner lane
        }
        pile0[id] = r;
    }
    __syncwarp();
}

// @region flat B shell (add_car batch)
// add_car (CHS.hpp:864-877 / 1029-1042) for up to 128 new cars of the warp's hubs, lanes = cars of any hub.  The
// owners have walked their streams (which attempt of each polar normal is accepted, which target draw); the
// variates, the curves and the pile record are computed here.  Items sit in the (idle) car_step rings:
// la_t = start state of the mk_soc attempt, la_w = of the mk_late_time attempt, la_id = lane | pile << 5 | u << 10.
template <int TYPE, int CP, int Q>
__device__ __forceinline__ void flat_add_car(Scr<Q> S, const float *tlut, uint4 *pile0, float *arrive0, int total) {
    const int lane = S.lane;
    const uint32_t *const sd = &S->la_t[0][0], *const sf = &S->la_w[0][0], *const sx = &S->la_id[0][0];
    for (int j = lane; j < total; j += 32) {
        const uint32_t x = sx[j], owner = x & 31u, i = (x >> 5) & 31u, tu = x >> 10;
        CHUB_CHECK(total <= 2 * LA_RING && tu < 1000u);
        const float arrive = mk_soc_from<1>(sd[j]);
        const int late = late_time_from<1>(sf[j]);
        const float tsoc = curve_s2t<TYPE, CP>(arrive);
        const float needed = tlut[tu] - tsoc;
        const int must_needed = (int) ceilf(needed);          // calculate_min_charging_time
        int stay = must_needed + late;
        const float power = curve_t2p<TYPE, CP>(tsoc);
        int cn = needed > 0.0f ? must_needed : 0;
        if (stay > 63 || stay < 1 || cn > 31) {
            atomicOr(&S->err_s[owner], (uint32_t) CHUB_HUBERR_COUNTER);
            stay = min(max(stay, 1), 63);
            cn = min(cn, 31);
        }
        const uint32_t urgent = (cn > 0 && stay <= cn) ? 1u : 0u;
        const uint32_t at = i * 32u + owner;
        pile0[at] = make_uint4(__float_as_uint(arrive), __float_as_uint(tsoc), __float_as_uint(power),
                               (uint32_t) stay | PW_CAR | (urgent << 14) | ((uint32_t) cn << 15) | (tu << 20));
        arrive0[at] = arrive;
        (&S->pw[0][0])[at] = __float_as_uint(power) | (urgent << 31);
    }
    __syncwarp();
}

// All 32 lanes call this together.  `mine` = this lane's hub has a station of this (TYPE, CP) at this
// position; the other lanes own no piles here but still work on the pooled items.
// @region station prologue + receive_car draws
template <int TYPE, int CP, int Q>
__device__ __forceinline__ void station_step(const ChubParams &P, Scr<Q> S, const float *tl_s, bool mine, int h,
                                          int slot0, int n_in, int t, int line_in, uint32_t act_bits, uint32_t occ_in,
                                          HubRng &rng, uint32_t &err, StationOut &o) {
    const unsigned full = 0xffffffffu;
    const int lane = S.lane;
    const int h0 = h - lane;
    const int n = mine ? n_in : 0;
    // s2t(target) of the 1000 possible targets: shared-memory copy for curve charging, global (L1) for constant power
    const float *tlut = CP ? P.s2t_lut + (2 + TYPE) * 1000 : tl_s + TYPE * 1000;
    S->inc_s[lane] = 0u; S->err_s[lane] = 0u;

    // ---- station-level draws of receive_car (CHS.hpp:1272-1306), before any per-car draw ----
    int want = 0, flow = 0;
    if (mine) {
        const int N = (int) __ldg(&P.lut[t * 1000 + rng_u1000(rng)]);
        const int in_car = (int) roundf((TYPE == CHUB_FAST ? 0.15f : 0.1f) * 0.2f * (float) N);
        int kept = 0;
        for (int w = 0; w < line_in; ++w) {  // reneging: keep if u > 0.1*logf(w+1)
            if (rng_u1000(rng) > (int) kC.renege_k[w]) kept += 1;
        }
        int true_in = 0;
        for (int k = 0; k < in_car; ++k) {   // balking: stay if u <= expf(-0.01*(line+k)) and k <= n
            if (rng_u1000(rng) <= (int) kC.balk_k[min(kept + k, 63)] && k <= n) true_in += 1;
        }
        flow = (TYPE == CHUB_SLOW) ? true_in : in_car;  // the slow/fast asymmetry, CHS.hpp:1306 vs 1617
        want = kept + flow;
    }
    __syncwarp();

// @region scan
    // ---- scan + flat A ----
    // Charging power is never negative (curves and constants, tests/test_host_logic.py), so the sign bit of the
    // posted power carries the pile's urgency to the owner's summation.
    const size_t tile0 = chub_pile_index(h0, slot0, P.S0 + P.S1);
    uint4 *const pile0 = P.pile + tile0;             // the warp's tile of this station: [pile][32 hubs], contiguous
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t car_mask = 0u, chg_mask = 0u, dep_mask = 0u;
    int incomplete = 0;
    int cnt0 = 0, cnt1 = 0, head0 = 0, head1 = 0;    // warp-uniform ring state: first waiting slot (mod LA_RING), items waiting
    // Only rows on which some hub of the warp has a car are visited, and a lane only loads the piles its own
    // occupancy mask names: the traffic of a quiet hour is proportional to the cars, not to the pile count.
    const uint32_t occ = mine ? occ_in : 0u;
    asm volatile("" : "+r"(act_bits));               // this station's action bits stay in one register through the scan
    uint32_t rows = __reduce_or_sync(full, occ);
    uint4 *rp = pile0 + lane;
    asm volatile("" : "+l"(rp));                     // keep the tile pointer in a register pair instead of re-deriving it
    uint32_t *const pwp = &S->pw[0][lane];
    int inext = __ffs(rows) - 1;
    uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
    if (rows && ((occ >> inext) & 1u)) nxt = rp[inext * 32];
    while (rows) {
        const uint4 r = nxt;
        const int i = inext;
        rows &= rows - 1u;
        inext = __ffs(rows) - 1;
        if (rows && ((occ >> inext) & 1u)) nxt = rp[inext * 32];
        int seg = -1;
        float tt = 0.0f;
        const uint32_t w = r.w, bit = 1u << i;
        if (occ & bit) {                             // a car is on this pile of my hub (else r is stale: not loaded)
            car_mask |= bit;
            const int left = (int) PW_STAY(w) - (int) PW_ALREADY(w) - 1;
            if (left <= 0) dep_mask |= bit;
            // judge_feasibility + assign_on_off_piece (CHS.hpp:1404-1413, 1364-1373)
            if ((w & PW_URGENT) || (act_bits & bit)) {
                chg_mask |= bit;
                tt = __uint_as_float(r.y) + 1.0f;
                seg = curve_seg<TYPE, CP>(tt);
            } else if (left <= 0) {                            // remove_car of a car that is not charging
                if (ceilf(__uint_as_float(r.x)) < floorf(target_of(PW_TU(w)))) incomplete += 1;
                rp[i * 32] = make_uint4(0u, 0u, 0u, 0u);
            } else {
                const int cn = PW_CN(w);
                const uint32_t urgent = (cn > 0 && left <= cn) ? 1u : 0u;
                rp[i * 32].w = ((w & ~(PW_CHARGE | PW_URGENT)) + (1u << 6)) | (urgent << 14);
                pwp[i * 32] = r.z | (urgent << 31);
            }
        }
        const uint32_t m0 = __ballot_sync(full, seg == 0), m1 = CP ? 0u : __ballot_sync(full, seg == 1);
        const int need = rows ? 32 : 1;              // the last occupied row also flushes the rings
        if ((m0 | m1) != 0u || need == 1) {          // otherwise nobody charges on this row: nothing to queue
            CHUB_CHECK(cnt0 + __popc(m0) <= LA_RING && cnt1 + __popc(m1) <= LA_RING);
            if (seg >= 0) {
                CHUB_CHECK(mine && i < n);
                const int sl = seg * LA_RING + ((seg == 0 ? head0 + cnt0 + __popc(m0 & lt) : head1 + cnt1 + __popc(m1 & lt)) & (LA_RING - 1));
                (&S->la_t[0][0])[sl] = __float_as_uint(tt);
                (&S->la_w[0][0])[sl] = w;
                (&S->la_id[0][0])[sl] = (uint32_t) lane | ((uint32_t) i << 5);
            }
            cnt0 += __popc(m0);
            cnt1 += __popc(m1);
            while (cnt0 >= need) {
                const int c = min(cnt0, 32);
                flat_car_step<TYPE, CP, 0, Q>(P, S, tlut, pile0, head0, c);
                head0 += c; cnt0 -= c;
            }
            while (!CP && cnt1 >= need) {
                const int c = min(cnt1, 32);
                flat_car_step<TYPE, CP, 1, Q>(P, S, tlut, pile0, head1, c);
                head1 += c; cnt1 -= c;
            }
        }
    }

// @region assign_car
    // ---- assign_car, wait == true (CHS.hpp:417-451): the first `assigned` empty piles, ascending ----
    const uint32_t nmask = n >= 32 ? 0xffffffffu : ((1u << n) - 1u);
    const uint32_t occ_after = car_mask & ~dep_mask;
    const int empties = n - __popc(occ_after);
    const int assigned = min(want, empties);  // may be negative
    uint32_t add_mask = 0u;
    {
        uint32_t freeb = ~occ_after & nmask;
        for (int k = 0; k < assigned; ++k) {
            const uint32_t b = freeb & (0u - freeb);
            add_mask |= b;
            freeb ^= b;
        }
    }

    // The arrival SOC of a new car goes to a row of its own (`arrive`), 4 bytes into a sector nobody has read: a partial
    // write to a sector that is not in L2 ends as a read-modify-write at the memory.  The rows that will take new cars are
    // known now; lane r asks for row r's 128-byte line, and by the time the walk is done the stores find it in L2.
    {
        const uint32_t arows = __reduce_or_sync(full, add_mask);
        if ((arows >> lane) & 1u) asm volatile("prefetch.global.L2 [%0];" ::"l"(P.arrive + tile0 + (size_t) lane * 32));
    }

// @region add_car (per-lane stream walk)
    // ---- add_car for the newly assigned piles, ascending (CHS.hpp:864-877 / 1029-1042).  Per car the E stream serves
    // mk_soc then mk_late_time and the U stream the target draw in between.  Which attempt of each polar normal is
    // accepted is a serial walk through the hub's stream: one lane = one hub, four cars per round.  Everything else is
    // pooled over the warp (flat_add_car). ----
    {
        const int a = max(assigned, 0);
        uint32_t rem = add_mask;
        for (int kb = 0; __any_sync(full, kb < a); kb += 4) {
            const int q = min(max(a - kb, 0), 4);
            int incl = q;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int v = __shfl_up_sync(full, incl, d);
                if (lane >= d) incl += v;
            }
            const int total = __shfl_sync(full, incl, 31);
            int slot = incl - q;
            // One loop trip = one polar ATTEMPT of whichever kind the lane is at -- the double attempt of mk_soc (four
            // draws: x from the pair's second state, y from the fourth) or the float attempt of mk_late_time (two draws) --
            // through one body: the two differ in their jump multipliers and in the band test only.  A lane's trip count
            // is the SUM of its attempts (10.2 +- 1.7 for four cars) instead of the warp waiting, car by car and normal by
            // normal, for the lane with the most rejections (2.5 trips per normal with a dozen lanes at work).
            {
                int left = q;
                uint32_t st = rng.lcg, sd = 0u, tu = 0u;
                bool mode_f = false;
                while (left > 0) {
                    const uint32_t sx = chub_minstd_jump(st, mode_f ? 16807u : CHUB_MINSTD_A2);
                    const uint32_t sy = chub_minstd_jump(st, mode_f ? CHUB_MINSTD_A2 : CHUB_MINSTD_A4);
                    // 2 * canonical - 1 of generate_canonical<float,24> (bits/random.tcc: (s - 1) / 2^31, clamped below 1):
                    // one rounding, like the reference's 2.0f * ret - 1.0f (the scalings by powers of two are exact).  For
                    // the double attempt the same expression on the pair's high states estimates x^2 + y^2 to 1e-6 (draw_mk_soc, chub_device.cuh).
                    const float x = fmaf(fminf((float) (sx - 1u), 2147483520.0f), 0x1p-30f, -1.0f);
                    const float y = fmaf(fminf((float) (sy - 1u), 2147483520.0f), 0x1p-30f, -1.0f);
                    const float r2 = x * x + y * y;
                    bool ok = !(r2 > 1.0f || r2 == 0.0f);
                    if (!mode_f && (fabsf(r2 - 1.0f) < 4e-6f || r2 < 4e-6f)) {   // too close to call in fp32: like the reference
                        const double xe = 2.0 * canon_d_of(chub_minstd_next(st), sx) - 1.0;
                        const double ye = 2.0 * canon_d_of(chub_minstd_jump(st, CHUB_MINSTD_A3), sy) - 1.0;
                        const double r2e = xe * xe + ye * ye;
                        ok = !(r2e > 1.0 || r2e == 0.0);
                    }
                    const uint32_t s0 = st;
                    st = sy;
                    if (ok) {
                        if (!mode_f) {
                            sd = s0;                                  // E: mk_soc's accepted attempt starts here
                            tu = (uint32_t) rng_u1000(rng);           // U  (target = u/999*20+80)
                            mode_f = true;
                        } else {                                      // E: mk_late_time's accepted attempt starts at s0
                            const uint32_t i = (uint32_t) __ffs(rem) - 1u;
                            rem &= rem - 1u;
                            CHUB_CHECK(slot >= 0 && slot < 2 * LA_RING && i < (uint32_t) n);
                            (&S->la_t[0][0])[slot] = sd;
                            (&S->la_w[0][0])[slot] = s0;
                            (&S->la_id[0][0])[slot] = (uint32_t) lane | (i << 5) | (tu << 10);
                            slot += 1;
                            left -= 1;
                            mode_f = false;
                        }
                    }
                }
                rng.lcg = st;
            }
            __syncwarp();
            flat_add_car<TYPE, CP, Q>(S, tlut, pile0, P.arrive + tile0, total);
        }
    }

// @region station sums
    // ---- calculate_output in ascending fp32 order (CHS.hpp:1233-1261); adding +0.0f for a skipped pile is exact ----
    const uint32_t present = occ_after | add_mask;
    const uint32_t chg_after = chg_mask & ~dep_mask & ~add_mask;
    float mn = 0.0f, mx = 0.0f, now = 0.0f;
    for (uint32_t prow = __reduce_or_sync(full, present); prow; prow &= prow - 1u) {
        const int i = __ffs(prow) - 1;
        const uint32_t vb = pwp[i * 32];
        const bool here = (present >> i) & 1u;
        CHUB_CHECK(i < max(P.S0, P.S1) && (!here || (mine && i < n)));
        const float power = here ? __uint_as_float(vb & 0x7fffffffu) : 0.0f;
        mx += power;
        mn += (vb & PWB_URGENT) ? power : 0.0f;
        now += ((chg_after >> i) & 1u) ? power : 0.0f;
    }
    if (mine) {
        o.mn = mn; o.mx = mx; o.now = now;
        o.number = __popc(present);
        o.line = min(want - assigned, 10);
        o.flow = flow;
        o.incomplete = incomplete + (int) S->inc_s[lane];
        err |= S->err_s[lane];
        o.occ = present;
    }
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// hydrogen side (hydro_sys.py), fp64 like the reference's Python floats; every expression keeps
// Python's evaluation order and -fmad=false keeps each multiply/add separately rounded.

// x / D for a literal D: exactly the IEEE quotient (Markstein correction), 3 instructions instead of ~25
#define DIVC(x, D) chub_div_const((x), (D), 1.0 / (D))
// the same with the reciprocal (and, where its low word is not zero, the divisor) from the constant-bank table
#define DIVK(x, D, RD) chub_div_const((x), (D), DK(RD))

// @region hydrogen helpers
__device__ __forceinline__ double py_max(double a, double b) { return b > a ? b : a; }  // Python max(a, b)
__device__ __forceinline__ double py_min(double a, double b) { return b < a ? b : a; }  // Python min(a, b)

__device__ __forceinline__ double el_power(const ChubClass &c, double m) {  // Electrolyser.get_power, hydro_sys.py:38-48
    if (c.el_cells == 0.0) return 0.0;
    double v_H_mass = chub_div_const(m, c.el_cells, c.r_el_cells);
    double v_H_mol = DIVK(v_H_mass, DK(H_MOL), H_RMOL);
    double v_H_L = v_H_mol * c.v_M;
    double v_H = v_H_L * 1000.0 * 60.0;
    double temp = chub_div_const(v_H * 2.0 * 96487.0, c.el_den, c.r_el_den);
    double power = (temp * temp) * DK(H_EA) + temp * DK(H_EB);
    return DIVK(c.el_cells * power, 1000.0, H_R1000);
}
__device__ __forceinline__ double cpr_power(const ChubClass &c, double m) {  // Compressor.generate_W, hydro_sys.py:74-82
    double n = DIVK(m, DK(H_MOL), H_RMOL);
    return DIVK(chub_div_const(n * c.cpr_w12, DK(H_08), 1.0 / 0.8), 1000.0, H_R1000);
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
    double p0 = (soc * DK(H_001)) * 70.0;
    double tp = j2601_target(p0);
    if (p0 < 5.0) minutes = DIVK(69.0 - p0, 18.5, H_R185) + kC.topoff_time;
    else minutes = DIVK(tp - p0, 18.5, H_R185);
    grams = kC.fcev_cap * DIVK(tp, 70.0, H_R70) - kC.fcev_cap * DIVK(p0, 70.0, H_R70);
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

// @region hub load/store/trace/obs
struct HubVars {  // per-hub registers that live across the stages
    HubRng rng;
    int t, line0, line1, qn, fline, farr, flow0, flow1;
    uint32_t occ0, occ1;
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
    const uint2 oc = P.occ[h];
    v.occ0 = oc.x; v.occ1 = oc.y;
    v.err = 0u;   // v.store is loaded after the stations (it is not needed before, and registers are scarce there)
}
__device__ __forceinline__ void hub_store(const ChubParams &P, int h, const HubVars &v) {
    uint32_t z = (uint32_t) v.t | ((uint32_t) v.line0 << 8) | ((uint32_t) v.line1 << 16) | ((uint32_t) v.qn << 24);
    uint32_t w = (uint32_t) ((v.flow0 + 128) & 0xff) | ((uint32_t) ((v.flow1 + 128) & 0xff) << 8) |
                 ((uint32_t) (v.fline & 0xff) << 16) | ((uint32_t) (v.farr & 0xff) << 24);
    P.hubA[h] = make_uint4(v.rng.lcg, v.rng.uctr, z, w);
    P.occ[h] = make_uint2(v.occ0, v.occ1);
    P.store[h] = v.store;
    if (v.err) P.err[h] |= v.err;
}

// Exogenous traces are addressed by (the hub's day, the hub's time of day): slot = 96 * floor((k - t) / 96) + t', where k
// is the absolute step index, t the hub's time slot when step k starts and t' = t (state being left) or t + 1 (state
// being entered; 96 = slot 0 of the next day).  For hubs that run in lock step with the step index (reset at a multiple
// of 96, every test against the reference) this is simply k or k + 1.  A hub that was reset in the middle of a day
// (masked reset, TimeLimit truncation, a de-phased batch) keeps reading the slot that belongs to ITS time of day, so PV
// follows its sun and the price its tariff table (evcssp_manager.py:344-361 shows table[t - 1] at time t).
__device__ __forceinline__ void trace_at(const ChubParams &P, int h, long long k, int t, int t_read, double &pv, double &wd,
                                         double &pr) {
    const long long d = k - (long long) t;
    uint32_t m;
    if ((unsigned long long) d < 0x7fff0000ull) {   // every practical rollout: 32-bit arithmetic, division by a constant
        m = ((uint32_t) d / (uint32_t) CHUB_DAY) * (uint32_t) CHUB_DAY + (uint32_t) t_read;
        if (m >= (uint32_t) P.trace_len) m %= (uint32_t) P.trace_len;
    } else {
        const long long day = d >= 0 ? d / CHUB_DAY : -((CHUB_DAY - 1 - d) / CHUB_DAY);
        long long mm = (day * CHUB_DAY + (long long) t_read) % (long long) P.trace_len;
        if (mm < 0) mm += P.trace_len;
        m = (uint32_t) mm;
    }
    size_t i = (size_t) m;
    if (P.trace_per_hub) i = i * (size_t) P.Hpad + h;
    pv = __ldg(&P.pv[i]); wd = __ldg(&P.wd[i]); pr = __ldg(&P.price_tr[i]);
}

// ---------------------------------------------------------------------------------------------
// on-device exogenous generator (chub_set_exogenous_generator): the reference's ReNew / OU_Noise / make_state
// arithmetic on the hub's X stream (include/chub_exo.h)
struct ExoState {
    double ou_pv, ou_wd, ou_pr, hold;
    uint32_t days, ctr;   // pv_day | wd_day << 10 | (price_count mod 4) << 20
};
__device__ __forceinline__ void exo_load(const ChubParams &P, int h, ExoState &g) {
    const size_t s = (size_t) P.Hpad;
    g.ou_pv = P.exo_d[h]; g.ou_wd = P.exo_d[s + h]; g.ou_pr = P.exo_d[2 * s + h]; g.hold = P.exo_d[3 * s + h];
    const uint2 w = P.exo_i[h];
    g.days = w.x; g.ctr = w.y;
}
__device__ __forceinline__ void exo_store(const ChubParams &P, int h, const ExoState &g, double pv, double wd, double pr) {
    const size_t s = (size_t) P.Hpad;
    P.exo_d[h] = g.ou_pv; P.exo_d[s + h] = g.ou_wd; P.exo_d[2 * s + h] = g.ou_pr; P.exo_d[3 * s + h] = g.hold;
    P.exo_d[4 * s + h] = pv; P.exo_d[5 * s + h] = wd; P.exo_d[6 * s + h] = pr;
    P.exo_i[h] = make_uint2(g.days, g.ctr);
}
__device__ __forceinline__ void exo_current(const ChubParams &P, int h, double &pv, double &wd, double &pr) {
    const size_t s = (size_t) P.Hpad;
    pv = P.exo_d[4 * s + h]; wd = P.exo_d[5 * s + h]; pr = P.exo_d[6 * s + h];
}
// OU_Noise.sample (renewable.py:71-76): dx = theta (mu - x) + sigma N(0,1), mu = 0
__device__ __forceinline__ double exo_ou(double &x, double theta, double sigma, uint32_t k0, uint32_t k1, uint32_t &ctr) {
    const double dx = theta * (0.0 - x) + sigma * chub_xnormal(k0, k1, &ctr, kLogTab);
    x = x + dx;
    return x;
}
// the exogenous part of make_state(time) (evcssp_manager.py:344-361, renewable.py:37-49): out of line, once per hub-step
__device__ __noinline__ void exo_values(const ChubParams &P, ExoState &g, uint32_t k0, uint32_t k1, int time,
                                        double last_table_price, double &pv, double &wd, double &pr) {
    const int pv_day = g.days & 1023u, wd_day = (g.days >> 10) & 1023u;
    uint32_t pcount = (g.days >> 20) & 3u;
    double temp = __ldg(&P.pv_days[pv_day * 96 + time]);
    if (temp > 0.0 && (pv_day & 1) == 0) temp += exo_ou(g.ou_pv, 0.01, 1.0, k0, k1, g.ctr) * P.exo_renew_gain;
    pv = (temp > 0.0 ? temp : 0.0) * 5.0;
    temp = __ldg(&P.wd_days[wd_day * 96 + time]);
    temp += exo_ou(g.ou_wd, 0.01, 1.5, k0, k1, g.ctr) * P.exo_renew_gain;
    wd = (temp > 0.0 ? temp : 0.0) * 1.0;
    if (pcount == 0u) g.hold = exo_ou(g.ou_pr, 0.1, 0.005, k0, k1, g.ctr) * P.exo_price_gain;
    pr = last_table_price + g.hold;
    pcount = (pcount + 1u) & 3u;
    g.days = (g.days & 0xfffffu) | (pcount << 20);
}
// reset(): renew_reset, make_state(0), price_count = 0 (evcssp_manager.py:304-316)
__device__ __forceinline__ void exo_reset(const ChubParams &P, ExoState &g, uint32_t k0, uint32_t k1, double &pv, double &wd,
                                          double &pr) {
    const uint32_t pd = chub_xrand31(k0, k1, g.ctr) % (uint32_t) P.n_pv_days;
    const uint32_t wdd = chub_xrand31(k0, k1, g.ctr + 1u) % (uint32_t) P.n_wd_days;
    g.ctr += 2u;
    g.days = pd | (wdd << 10) | (g.days & (3u << 20));
    exo_values(P, g, k0, k1, 0, kC.price[95], pv, wd, pr);
    g.days &= 0xfffffu;
}

// observation, evcssp_manager.py:318-342 (state_norm) over make_state's real_state
// obs_row: where the 13 values go instead of the hub's row of P.obs (the step kernel parks them in shared memory and the warp
// then writes its 32 rows -- one contiguous block of the row-major matrix -- with coalesced stores); NULL = straight to P.obs
__device__ __forceinline__ void write_obs(const ChubParams &P, int h, const ChubClass &c, int t, const StationOut &s0,
                                          const StationOut &s1, double store_soc, double pv, double wd, double pr,
                                          float *obs_row = nullptr) {
    // written where it goes as it is computed (a local array indexed by a run-time count lives in local memory: 26 more
    // trips through the load/store unit per hub-step)
    float *const o = obs_row ? obs_row : P.obs + (size_t) h * CHUB_OBS_DIM;
    int k = 0;
    o[k++] = (float) kC.time_sin[t];
    o[k++] = (float) chub_div_const(pr - kC.price_mean, kC.price_std, kC.r_price_std);
    if (c.n[0] > 0) {
        o[k++] = (float) chub_div_const((double) s0.mn - c.half[0], c.half[0], c.r_half[0]);
        o[k++] = (float) chub_div_const((double) s0.now - c.half[0], c.half[0], c.r_half[0]);
        o[k++] = (float) chub_div_const((double) s0.mx - c.half[0], c.half[0], c.r_half[0]);
        o[k++] = (float) DIVK((double) s0.line, 5.0, H_R5);
    }
    if (c.n[1] > 0) {
        o[k++] = (float) chub_div_const((double) s1.mn - c.half[1], c.half[1], c.r_half[1]);
        o[k++] = (float) chub_div_const((double) s1.now - c.half[1], c.half[1], c.r_half[1]);
        o[k++] = (float) chub_div_const((double) s1.mx - c.half[1], c.half[1], c.r_half[1]);
        o[k++] = (float) DIVK((double) s1.line, 5.0, H_R5);
    }
    o[k++] = (float) store_soc;
    o[k++] = (float) DIVK(pv, 210.0, H_R210);
    o[k++] = (float) DIVK(wd, 92.0, H_R92);
    for (; k < CHUB_OBS_DIM; ++k) o[k] = 0.0f;
}

// EvcsspManagerEnv_v6.reset (evcssp_manager.py:304-316) of one hub, from and to HBM: both stations
// (evs_reset), hydrogen store, FCEV queue, episode statistics, first observation of the new episode.
// Out of line on purpose: the step kernel calls it once per hub per 96 steps (fused auto-reset).
// @region hub_reset_full
__device__ __noinline__ void hub_reset_full(const ChubParams &P, int h, long long k_obs, float *obs_row = nullptr) {
    const int ci = P.cls ? __ldg(&P.cls[h]) : 0;
    const ChubClass &c = P.classes[ci];
    HubVars v;
    hub_load(P, h, v);
    v.store = 0.0;
    StationOut s0 = station_reset_any(P, h, 0, c.n[0], c.type[0], c.constant, v.rng, v.err);
    StationOut s1 = station_reset_any(P, h, P.S0, c.n[1], c.type[1], c.constant, v.rng, v.err);
    v.t = 0;
    v.line0 = s0.line; v.line1 = s1.line;
    v.flow0 = s0.flow; v.flow1 = s1.flow;
    v.occ0 = s0.occ; v.occ1 = s1.occ;
    v.qn = 0; v.fline = 0; v.farr = 0;
    v.store = c.init_mass;
    P.stats[2 * h] = make_double2(0.0, 0.0);
    P.stats[2 * h + 1] = make_double2(0.0, 0.0);
    double pv, wd, pr;
    if (P.exo_mode) {
        ExoState g;
        exo_load(P, h, g);
        exo_reset(P, g, v.rng.k0, v.rng.k1, pv, wd, pr);
        exo_store(P, h, g, pv, wd, pr);
    } else {
        trace_at(P, h, k_obs, 0, 0, pv, wd, pr);
    }
    write_obs(P, h, c, 0, s0, s1, c.init_soc, pv, wd, pr, obs_row);
    if (P.diag) {
        const size_t s = (size_t) P.Hpad;
        float *q = P.st_power + h;
        q[0 * s] = s0.mn; q[1 * s] = s0.now; q[2 * s] = s0.mx; q[3 * s] = s1.mn; q[4 * s] = s1.now; q[5 * s] = s1.mx;
    }
    hub_store(P, h, v);
}

// ---------------------------------------------------------------------------------------------
// The warp's 32 action rows are one contiguous block of the row-major [H][stride] matrix a policy
// emits.  It is copied into shared memory with coalesced 128-bit loads and each lane then reads ITS
// row from there (stride odd -> conflict free), binarising exactly like action_to_real on a float32
// vector (evcssp_manager.py:384-404): on iff (a + 1)/2 >= 0.5 in fp32.  fl(a + 1) >= 1 holds exactly
// for a >= -2^-25 (the tie at 1 - 2^-25 rounds to even = 1.0), so the test is a single compare
// (tests/test_host_logic.py checks the equivalence around the boundary).  Rows that cannot be staged
// (partial last warp, very wide or misaligned matrices) are read straight from global memory.
// @region load_actions
__device__ __forceinline__ void load_actions(const ChubParams &P, float *stage, int stage_floats, int h, bool active,
                                             int na, unsigned long long &bits, float &a_m2, float &a_m1) {
    // a_m2 / a_m1 are returned in registers; the caller parks them in shared memory after the final __syncwarp
    const int lane = threadIdx.x & 31;
    const int h0 = h - lane;
    const int stride = P.action_stride;
    bits = 0ull; a_m2 = 0.0f; a_m1 = 0.0f;
    if (P.act_packed) {   // packed-action mode: the rows were thresholded by the producer, one 16-byte record per hub
        if (active) {
            const uint4 r = __ldcs(P.act_packed + h);
            bits = (unsigned long long) r.x | ((unsigned long long) r.y << 32);
            if (na < 64) bits &= (1ull << na) - 1ull;
            a_m2 = __uint_as_float(r.z);
            a_m1 = __uint_as_float(r.w);
        }
        return;
    }
    if (P.act_bits) {   // scalar-load mode: chub_dispatch_kernel already turned the two power targets into pile bits
        if (active) {
            bits = P.act_bits[h];
            a_m2 = __ldg(P.actions + (size_t) h * 4 + 2);
            a_m1 = __ldg(P.actions + (size_t) h * 4 + 3);
        }
        return;
    }
    const bool staged = (h0 + 32 <= P.h_end) && (32 * stride <= stage_floats) &&
                        ((reinterpret_cast<uintptr_t>(P.actions) & 15u) == 0);
    const float *src = P.actions + (size_t) h * stride;
    if (staged) {
        const uint4 *g = reinterpret_cast<const uint4 *>(P.actions + (size_t) h0 * stride);
        uint4 *d = reinterpret_cast<uint4 *>(stage);
        CHUB_CHECK(32 * stride <= stage_floats && h0 + 32 <= P.h_end);
        for (int c = lane; c < 8 * stride; c += 32) d[c] = __ldcs(g + c);
        __syncwarp();
        src = stage + lane * stride;
    }
    if (active) {
        // two 32-bit halves built from the top down (shift in one bit per entry): no 64-bit shifts, no per-entry bit index
        uint32_t lo = 0u, hi = 0u;
        const int n_lo = na < 32 ? na : 32;
#pragma unroll 4
        for (int j = na - 1; j >= 32; --j) hi = (hi << 1) | (src[j] >= -0x1p-25f ? 1u : 0u);
#pragma unroll 4
        for (int j = n_lo - 1; j >= 0; --j) lo = (lo << 1) | (src[j] >= -0x1p-25f ? 1u : 0u);
        bits = (unsigned long long) lo | ((unsigned long long) hi << 32);
        a_m2 = src[na];      // action[-2] -> electrolyser set-point (the reference swaps the last two)
        a_m1 = src[na + 1];  // action[-1] -> fuel-cell set-point
    }
    __syncwarp();  // the staging area is reused as the stations' scratch
}

// constant-power stations (constant_charging=True) are rare: one shared out-of-line copy per type.  State
// travels by value so that the caller's copies stay in registers.
// @region kernel head
struct StationIO {
    HubRng rng;
    uint32_t err;
    StationOut o;
};
template <int TYPE, int Q>
__device__ __noinline__ StationIO station_step_cp(const ChubParams &P, Scr<Q> S, const float *tl_s, bool mine, int h,
                                                  int slot0, int n, int t, int line, uint32_t act_bits, uint32_t occ_in, StationIO io) {
    station_step<TYPE, 1, Q>(P, S, tl_s, mine, h, slot0, n, t, line, act_bits, occ_in, io.rng, io.err, io.o);
    return io;
}

// One hub tile (32 consecutive hubs, one per lane) through the whole step.
template <int Q>
__device__ __forceinline__ void step_tile(const ChubParams &P, Scr<Q> S, const float *tl_s, const int h);
template <int Q>
__device__ __forceinline__ void flush_obs(const ChubParams &P, Scr<Q> S, int h0);

// The kernel is persistent: one block per SM pulls UNITS of WARPS consecutive tiles (what a block of the static grid would
// own) from a ticket counter until the batch is used up, heaviest units first.
// Why: a hub's cost follows its time of day (quiet night: 10 k warp instructions per tile, every pile occupied: 25 k, the
// resetting step: 32 k).  In a batch whose hubs are not all at the same time of day (staggered resets, truncation) units
// differ 3.5x in duration, and with only four units per SM the order they are handed out in decides how long the last SMs
// run alone (measured with the static grid: SMs 81 % busy).  Longest-first is the classic remedy; the cost of a unit is
// what it took in the PREVIOUS launch (a day moves one step at a time), measured with clock64 and binned into
// CHUB_LPT_BINS bins of 8192 cycles: a finished unit appends itself to its bin's list, the next launch walks the bins from
// the top.  No sort, one atomic per unit.  A unit stays on ONE SM with all its warps: warps at different times of day run
// different parts of a 64 KB loop body, and mixing them on one SM costs a quarter of the issue slots in instruction-cache
// misses (measured: no_instruction stalls 2.7 % -> 24 %; handing quarter-units to the four sub-cores separately: the same).
// Tickets: unit k of launch j of an epoch is atomicAdd(queue) - (*queue_base + j * stride); every block draws exactly one
// ticket past the end, so a launch advances the counter by stride = units + blocks and the tick kernel moves the base.
// The lists are double buffered by the parity of the absolute step; a list that does not hold exactly this launch's
// units (first launch, other batch range or shape) is ignored and units go out in index order.
// the absolute step index of this launch (*P.step_index + P.step_offset), read from global memory ONCE per block: every tile
// needs it for its trace slots and every unit draw for the parity of the cost lists, and a dependent trip to L2 each time
// is what the car-independent part of a step is made of (tools/empty_tile_cost.py)
__shared__ long long g_step_k;
#define CHUB_LPT_BINS 64
struct LptList {            // one per parity, in global memory: head[0] = units the list was built for, then the bins
    int n_units, done_blocks, pad[2];
    int count[CHUB_LPT_BINS];
    // int unit[CHUB_LPT_BINS][cap] follows
};
// Warp 0 of a block draws a ticket and turns it into a unit (-1: the batch is used up).  All 32 lanes call it.
template <int WARPS>
__device__ __noinline__ int draw_unit(const ChubParams &P, int lane) {
    const int n_units = (int) (((long long) (P.h_end - P.h_begin) + WARPS * 32 - 1) / (WARPS * 32));
    const int par = (int) (g_step_k & 1);
    const size_t list_bytes = sizeof(LptList) + (size_t) CHUB_LPT_BINS * P.lpt_cap * sizeof(int);
    const LptList *const rd = reinterpret_cast<const LptList *>(P.lpt + par * list_bytes);
    long long k = 0;
    if (lane == 0) {
        const unsigned long long first = *P.queue_base + (unsigned long long) P.step_offset * P.queue_stride;
        k = (long long) (atomicAdd(P.queue, 1ull) - first);
    }
    k = __shfl_sync(0xffffffffu, k, 0);
    if (k >= n_units) return -1;
    int u = (int) k;
    // bins from the top: lane l looks after bins 63 - 2l and 62 - 2l
    const int c_hi = rd->count[CHUB_LPT_BINS - 1 - 2 * lane], c_lo = rd->count[CHUB_LPT_BINS - 2 - 2 * lane];
    int incl = c_hi + c_lo;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    if (total == n_units && rd->n_units == n_units && !(P.phase_sync & 1)) {
        const int before = incl - c_hi - c_lo;
        const unsigned hit = __ballot_sync(0xffffffffu, (int) k < incl);
        const int owner = __ffs(hit) - 1;
        if (lane == owner) {
            const int r = (int) k - before;
            const int bin = r < c_hi ? CHUB_LPT_BINS - 1 - 2 * lane : CHUB_LPT_BINS - 2 - 2 * lane;
            u = reinterpret_cast<const int *>(rd + 1)[(size_t) bin * P.lpt_cap + (r < c_hi ? r : r - c_hi)];
        }
        u = __shfl_sync(0xffffffffu, u, owner);
    }
    return u;
}

template <int WARPS, int MINB, int Q>
__global__ void __launch_bounds__(WARPS * 32, MINB) chub_step_kernel(const __grid_constant__ ChubParams P) {
    unsigned char *const smem = chub_smem;
    const float *tl_s = reinterpret_cast<const float *>(chub_smem + CHUB_LOG_N * 16);
    // the first 2 KB of shared memory hold the log table (chub_device.cuh log_tab<1>), then one scratch area per warp
    for (int i = threadIdx.x; i < CHUB_LOG_N * 4; i += WARPS * 32)
        reinterpret_cast<uint32_t *>(smem)[i] = reinterpret_cast<const uint32_t *>(kLogTab)[i];
    float *const tl_w = reinterpret_cast<float *>(smem + CHUB_LOG_N * 16);   // then the two curve-charging s2t(target) tables
    for (int i = threadIdx.x; i < 2000; i += WARPS * 32) tl_w[i] = __ldg(&P.s2t_lut[i]);
    __shared__ int s_unit[2];
    __shared__ long long s_t0;
    if (threadIdx.x == 0) g_step_k = *P.step_index + P.step_offset;
    __syncthreads();
    Scr<Q> S;
    S.base = (uint32_t) __cvta_generic_to_shared(chub_smem + CHUB_SMEM_TABLES + (threadIdx.x >> 5) * P.ws_bytes);
    S.lane = threadIdx.x & 31;
    asm volatile("" : "+r"(S.base), "+r"(S.lane));

    // Nothing of the queue's bookkeeping stays in registers across step_tile (they are the scarce resource there) but the
    // loop count: warp 0 re-derives the rest from the launch parameters at every draw, and it draws the NEXT unit as soon as
    // its own tile is done, so a unit boundary costs one barrier and (mostly) no round trip.
    // Without a queue (ranged launches of the host pipeline, which run side by side on several streams) the mapping is
    // static: block b does unit b and leaves.  One loop serves both, so that the tile code exists once.
    if (threadIdx.x < 32) {
        const int u = P.queue ? draw_unit<WARPS>(P, threadIdx.x) : (int) blockIdx.x;
        if (threadIdx.x == 0) s_unit[0] = u;
    }
    __syncthreads();
    for (int it = 0;; ++it) {
        const int unit = s_unit[it & 1];
        if (unit < 0) break;
        CHUB_CHECK((long long) unit * WARPS * 32 < (long long) (P.h_end - P.h_begin));
        if (threadIdx.x == 0) s_t0 = clock64();
        const long long h0 = (long long) P.h_begin + ((long long) unit * WARPS + (threadIdx.x >> 5)) * 32;
        if (h0 < (long long) P.h_end) {   // the last unit may be short of tiles
            step_tile<Q>(P, S, tl_s, (int) h0 + S.lane);
            flush_obs<Q>(P, S, (int) h0);
        }
        if (threadIdx.x < 32) {
            // warp 0 draws the next unit when ITS tile is done, i.e. while the unit's stragglers are still at work: late
            // enough that units still go to whichever SM is about to be free (drawing at the start of the unit would
            // reserve two units per SM up front and cost the de-phased batch 14 %: measured), early enough to hide the
            // round trip to the ticket counter and the bin walk
            const int nxt = P.queue ? draw_unit<WARPS>(P, threadIdx.x) : -1;
            if (threadIdx.x == 0) s_unit[(it + 1) & 1] = nxt;
        }
        __syncthreads();          // the unit's cost is that of its slowest warp
        if (threadIdx.x == 0 && P.queue) {   // book it: into the bin of its cost, for the next launch to read
            const int par = (int) (g_step_k & 1);
            const size_t list_bytes = sizeof(LptList) + (size_t) CHUB_LPT_BINS * P.lpt_cap * sizeof(int);
            LptList *const wr = reinterpret_cast<LptList *>(P.lpt + (par ^ 1) * list_bytes);
            int bin = (int) ((clock64() - s_t0) >> 13);
            bin = bin < 0 ? 0 : (bin >= CHUB_LPT_BINS ? CHUB_LPT_BINS - 1 : bin);
            const int at = atomicAdd(&wr->count[bin], 1);
            if (at < P.lpt_cap) reinterpret_cast<int *>(wr + 1)[(size_t) bin * P.lpt_cap + at] = s_unit[it & 1];
        }
    }
    // the last block out tags the list it helped write and clears the one everybody has finished reading
    if (threadIdx.x == 0 && P.queue != nullptr) {
        const int n_units = (int) (((long long) (P.h_end - P.h_begin) + WARPS * 32 - 1) / (WARPS * 32));
        const int par = (int) (g_step_k & 1);
        const size_t list_bytes = sizeof(LptList) + (size_t) CHUB_LPT_BINS * P.lpt_cap * sizeof(int);
        LptList *const rd = reinterpret_cast<LptList *>(P.lpt + par * list_bytes);
        LptList *const wr = reinterpret_cast<LptList *>(P.lpt + (par ^ 1) * list_bytes);
        __threadfence();
        if (atomicAdd(&wr->done_blocks, 1) == (int) gridDim.x - 1) {
            wr->done_blocks = 0;
            wr->n_units = n_units;
            rd->n_units = -1;
            for (int b = 0; b < CHUB_LPT_BINS; ++b) rd->count[b] = 0;
            if (P.self_tick) {   // every block has read the step index for the last time: advance it, and the queue's epoch
                *const_cast<long long *>(P.step_index) = *P.step_index + 1;
                *const_cast<unsigned long long *>(P.queue_base) = *P.queue_base + P.queue_stride;
            }
        }
    }
}

template <int Q>
__device__ __forceinline__ void step_tile(const ChubParams &P, Scr<Q> S, const float *tl_s, const int h) {
    const unsigned full = 0xffffffffu;
    const bool active = h < P.h_end;
    const int ci = (active && P.cls) ? __ldg(&P.cls[h]) : 0;
    const ChubClass &c = P.classes[ci];
    const int n0 = c.n[0], n1 = c.n[1];
    unsigned long long bits;
    {
        float a_m2, a_m1;
        load_actions(P, reinterpret_cast<float *>(&S->park[0][0][0]), (P.ws_bytes - (int) offsetof(WarpScratch<Q>, park)) / 4, h,
                     active, n0 + n1, bits, a_m2, a_m1);
        S->am[0][S.lane] = a_m2;
        S->am[1][S.lane] = a_m1;
    }

    // During the stations only the random streams, the time slot and the two queue lengths are needed; the rest of
    // the hub record stays packed (registers are the scarce resource there) and is unpacked afterwards.
    HubRng rng;
    rng.k0 = rng.k1 = rng.uctr = 0u; rng.lcg = 1u;
    uint32_t az = 0u, aw = 0u, err = 0u;
    uint2 occ = make_uint2(0u, 0u);
    if (active) {
        const uint4 a = P.hubA[h];
        const uint2 key = __ldg(&P.keys[h]);
        occ = P.occ[h];
        rng.k0 = key.x; rng.k1 = key.y; rng.lcg = a.x; rng.uctr = a.y;
        az = a.z; aw = a.w;
        // what the hub section will ask for after the stations: on its way to L2 now (no registers held)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(P.store + h));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(P.stats + 2 * h));
    }

    // ---- the two EV stations, station 0 first (Aggregator_Simple.py:141-142); the whole warp takes part ----
    const int lane_ = S.lane;
#pragma unroll 1
    for (int s = 0; s < 2; ++s) {
        const int type = c.type[s], n = s ? n1 : n0, slot0 = s ? P.S0 : 0, line = (az >> (8 + 8 * s)) & 0xff, t = az & 0xff;
        const uint32_t abits = s ? (uint32_t) (bits >> n0) : (uint32_t) bits, occ_s = s ? occ.y : occ.x;
        const bool curve = active && !c.constant;
        StationOut o = {};
        const bool m_slow = curve && type == CHUB_SLOW, m_fast = curve && type == CHUB_FAST;
        if (__any_sync(full, m_fast)) station_step<CHUB_FAST, 0, Q>(P, S, tl_s, m_fast, h, slot0, n, t, line, abits, occ_s, rng, err, o);
        if (__any_sync(full, m_slow)) station_step<CHUB_SLOW, 0, Q>(P, S, tl_s, m_slow, h, slot0, n, t, line, abits, occ_s, rng, err, o);
        const bool c_slow = active && c.constant && type == CHUB_SLOW, c_fast = active && c.constant && type == CHUB_FAST;
        if (__any_sync(full, c_fast || c_slow)) {
            StationIO io;
            io.rng = rng; io.err = err; io.o = o;
            if (__any_sync(full, c_fast)) io = station_step_cp<CHUB_FAST, Q>(P, S, tl_s, c_fast, h, slot0, n, t, line, abits, occ_s, io);
            if (__any_sync(full, c_slow)) io = station_step_cp<CHUB_SLOW, Q>(P, S, tl_s, c_slow, h, slot0, n, t, line, abits, occ_s, io);
            rng = io.rng; err = io.err; o = io.o;
        }
        // the station's outputs wait in shared memory for the hub section
        S->park[s][0][lane_] = __float_as_uint(o.mn); S->park[s][1][lane_] = __float_as_uint(o.mx);
        S->park[s][2][lane_] = __float_as_uint(o.now); S->park[s][3][lane_] = o.occ;
        S->park[s][4][lane_] = (uint32_t) (o.line & 0xff) | ((uint32_t) (o.flow & 0xff) << 8) | ((uint32_t) o.incomplete << 16);
    }
    // what only the hub section needs is fetched or unpacked now: store mass, exogenous traces, the two analog actions
    HubVars v;
    v.rng = rng; v.err = err;
    v.t = az & 0xff; v.qn = az >> 24; v.fline = (aw >> 16) & 0xff; v.farr = aw >> 24;
    StationOut s0, s1;
    {
        const uint32_t p0 = S->park[0][4][lane_], p1 = S->park[1][4][lane_];
        s0.mn = __uint_as_float(S->park[0][0][lane_]); s0.mx = __uint_as_float(S->park[0][1][lane_]);
        s0.now = __uint_as_float(S->park[0][2][lane_]); s0.occ = S->park[0][3][lane_];
        s0.line = p0 & 0xff; s0.flow = (int) (signed char) ((p0 >> 8) & 0xff); s0.incomplete = p0 >> 16; s0.number = __popc(s0.occ);
        s1.mn = __uint_as_float(S->park[1][0][lane_]); s1.mx = __uint_as_float(S->park[1][1][lane_]);
        s1.now = __uint_as_float(S->park[1][2][lane_]); s1.occ = S->park[1][3][lane_];
        s1.line = p1 & 0xff; s1.flow = (int) (signed char) ((p1 >> 8) & 0xff); s1.incomplete = p1 >> 16; s1.number = __popc(s1.occ);
    }
    const float a_m2 = S->am[0][lane_], a_m1 = S->am[1][lane_];
    __syncwarp();   // every lane has read what the stations parked: the scratch now takes the observation rows
    if (!active) return;
    v.store = P.store[h];
    const long long k = g_step_k;
    double pv_t, wd_t, pr_t;
    if (P.exo_mode) exo_current(P, h, pv_t, wd_t, pr_t);
    else trace_at(P, h, k, v.t, v.t, pv_t, wd_t, pr_t);
    v.line0 = s0.line; v.line1 = s1.line;
    v.flow0 = s0.flow; v.flow1 = s1.flow;
    v.occ0 = s0.occ; v.occ1 = s1.occ;

// @region hub section
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
                if (ind > 0) act = DK(H_001) * (double) (ind - 1);  // ind == 0 picks input[-1] (Python index -1)
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
    const double must = py_max(0.0, c.cap_mass * DK(H_01) - v.store);
    const double upper = py_max(c.cap_mass - v.store, 0.0);
    const double charge_temp = py_max(py_min(act * c.v_h_max * 900.0, upper), must);
    double flow = DIVK(charge_temp, 900.0, H_R900);
    flow = py_min(flow, c.v_h_max);
    const double ele_kw = el_power(c, flow);
    const double cpr_kw = cpr_power(c, flow);
    // HyStore.sty_step (hydro_sys.py:104-126)
    v.store += flow * 15.0 * 60.0;
    if (c.cap_mass - v.store < 0.0) v.err |= CHUB_HUBERR_STORE_ASSERT;
    double lower = py_max(0.0, v.store - DK(H_01) * c.cap_mass);
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
    double hy_fc = DIVK(fc * 1500.0, DK(H_FC), H_RFC);
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
    const double hy_loss = DK(H_M6) * hy_fc;
    const double p4 = pr_t * 0.25;
    const double inc_fast = DK(H_I42) * cp0 - p4 * l0;
    const double inc_slow = DK(H_I21) * cp1 - p4 * l1;
    const double inc_evs = inc_fast + inc_slow;
    const double serve = DK(H_08) * (double) (s0.flow + s1.flow);
    const double inc_hys = DK(H_P6) * hy_use;
    const double nm_loss = DK(H_M10) * not_meet;
    const double hy_cost = -p4 * h2_grid;
    const double income = inc_hys + inc_evs + serve + hy_cost;
    const double reward = DIVK(inc_hys + inc_evs + serve + hy_cost + 1.0 * hy_loss + nm_loss, 50.0, H_R50);
    const bool done = v.t + 1 >= CHUB_DAY;
    const int t_left = v.t;
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
    // accumulated in fp64 like the reference's Python floats (acumulate_reward, cumulated_income, cumulated_draw_ele)
    double2 sa = P.stats[2 * h], sb = P.stats[2 * h + 1];
    sa.x += reward;
    sa.y += income;
    sb.x += l0 + l1 + h2_grid;
    sb.y += (double) (s0.incomplete + s1.incomplete);
    if (done) {
        P.final_soc[h] = (float) store_soc;
        P.last_stats[h] = make_float4((float) sa.x, (float) sa.y, (float) sb.x, (float) sb.y);
        sa = make_double2(0.0, 0.0);
        sb = make_double2(0.0, 0.0);
    }
    P.stats[2 * h] = sa;
    P.stats[2 * h + 1] = sb;
    P.reward[h] = (float) reward;
    P.done[h] = done ? 1 : 0;
    hub_store(P, h, v);
    double pv_n = 0.0, wd_n = 0.0, pr_n = 0.0;
    if (P.exo_mode) {   // make_state(time_next) of this step: the price list's last entry is table[time just left]
        ExoState g;
        exo_load(P, h, g);
        exo_values(P, g, v.rng.k0, v.rng.k1, v.t, kC.price[(v.t + CHUB_DAY - 1) % CHUB_DAY], pv_n, wd_n, pr_n);
        exo_store(P, h, g, pv_n, wd_n, pr_n);
    } else if (!(done && P.auto_reset)) {
        trace_at(P, h, k, t_left, t_left + 1, pv_n, wd_n, pr_n);
    }
    // the observation row waits in the (now idle) station scratch: lane l's 13 values at float 13 l (odd pitch: no conflicts)
    float *const obs_row = reinterpret_cast<float *>(&S->park[0][0][0]) + lane_ * CHUB_OBS_DIM;
    if (done && P.auto_reset) {  // fused auto-reset: the returned obs is the first one of the next episode
        hub_reset_full(P, h, k + 1, obs_row);
    } else {
        write_obs(P, h, c, v.t, s0, s1, store_soc, pv_n, wd_n, pr_n, obs_row);
    }
}

// second half of step_tile, by the whole warp: the 32 observation rows parked in shared memory are one contiguous block of
// the row-major [H][13] matrix -- 13 coalesced 128-byte stores instead of 13 x 32 scattered ones
template <int Q>
__device__ __forceinline__ void flush_obs(const ChubParams &P, Scr<Q> S, int h0) {
    __syncwarp();
    const int rows = min(32, P.h_end - h0);
    const float *src = reinterpret_cast<const float *>(&S->park[0][0][0]);
    float *dst = P.obs + (size_t) h0 * CHUB_OBS_DIM;
#pragma unroll
    for (int k = 0; k < CHUB_OBS_DIM; ++k) {
        const int e = k * 32 + S.lane;
        if (e < rows * CHUB_OBS_DIM) dst[e] = src[e];
    }
    __syncwarp();
}

// @region other kernels
__global__ void __launch_bounds__(128) chub_reset_kernel(const __grid_constant__ ChubParams P) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= P.H) return;
    if (P.reset_mask && !P.reset_mask[h]) return;
    if (P.reset_mask) {
        // A masked reset cuts episodes short (TimeLimit truncation): what the last step returned -- reward, done -- stays
        // as it is, and the statistics of the episode that ends here become the "last finished episode" like at done.
        const double2 sa = P.stats[2 * h], sb = P.stats[2 * h + 1];
        P.last_stats[h] = make_float4((float) sa.x, (float) sa.y, (float) sb.x, (float) sb.y);
        const ChubClass &c0 = P.classes[P.cls ? __ldg(&P.cls[h]) : 0];
        P.final_soc[h] = (float) chub_div_const(P.store[h], c0.cap_mass, c0.r_cap_mass);
    } else {
        P.reward[h] = 0.0f;
        P.done[h] = 0;
    }
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

__global__ void chub_tick_kernel(long long *step_index, int n, unsigned long long *queue_base, unsigned long long tickets) {
    *step_index += n;
    if (queue_base) *queue_base += tickets;   // the launches since the last tick drew this many tickets (see chub_step_kernel)
}

template <int Q>
static int step_scratch_bytes(const ChubParams &p) {
    int ws = (int) warp_scratch_bytes<Q>(p.S0 > p.S1 ? p.S0 : p.S1);
    // room to stage the warp's 32 action rows behind the parked analog entries, when that stays small
    const int stage_bytes = (int) offsetof(WarpScratch<Q>, park) + 128 * p.action_stride;
    if (stage_bytes > ws && stage_bytes <= 10240) ws = stage_bytes;
    return ws;
}
template <int WARPS, int MINB, int Q>
static cudaError_t launch_step_variant(const ChubParams &p0, cudaStream_t s, unsigned long long *tickets, bool *self_ticked) {
    ChubParams p = p0;
    p.ws_bytes = step_scratch_bytes<Q>(p);
    const size_t smem = CHUB_SMEM_TABLES + (size_t) WARPS * p.ws_bytes;
    auto kern = chub_step_kernel<WARPS, MINB, Q>;
    // per device and cheap; set on every launch so that engines on several devices of one process all get it
    cudaError_t r = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (r != cudaSuccess) return r;
    const int block = WARPS * 32;
    const long long hubs = (long long) p.h_end - p.h_begin;
    long long grid = (hubs + block - 1) / block;
    if (p.queue) {   // persistent: one block per SM draws units of one block's worth of tiles
        const long long units = grid;
        const long long sms = (long long) (p.num_sms > 0 ? p.num_sms : 148) * MINB;
        grid = std::min(sms, units);
        p.queue_stride = (unsigned long long) (units + grid);
        if (units > p.lpt_cap) p.queue = nullptr, grid = units;   // more units than the lists hold: static grid
        if (tickets) *tickets = p.queue ? p.queue_stride : 0;
    } else if (tickets) {
        *tickets = 0;
    }
    if (!p.queue) p.self_tick = 0;   // a static grid does not know its last block
    if (self_ticked) *self_ticked = p.self_tick != 0;
    kern<<<(unsigned) grid, block, smem, s>>>(p);
    return cudaSuccess;
}
// Launch shape.  One block per SM.  Large blocks run a tile cheaper than small ones (measured: 28 warps 1.00, 16 warps
// 1.12, 8 warps 1.2, 4 warps 1.3 per tile), so the shape is the largest that still gives every SM a block's worth of
// tiles; below that small blocks spread a small batch over the SMs.  CHUB_BLOCK overrides (warps * 100 + 10).
// ---- scalar-load dispatch: the stations' evs_step(float) (CHS.hpp:1169-1186 slow, 1478-1495 fast) ----
// The reference clamps the station's power target to [min_power, max_power] (catch_load, CHS.hpp:358-366), orders the
// piles by urgency (a multimap keyed by -emergency: ties keep pile order, CHS.hpp:1375-1390) and switches cars on along
// that order while the running fp32 sum of their posted powers stays within target + 1e-4 (CHS.hpp:1355-1360); with
// constant-power charging it switches on the first round(target / constant_power) cars instead (CHS.hpp:1338-1349).
// Everything after that -- car_step, remove_car, receive_car -- is the per-pile path, so this kernel only turns the two
// targets of a hub into the same on-bits that a per-pile action row would give and the step kernel runs unchanged.
// One thread per hub; urgencies and powers of its occupied piles wait in shared memory (one row per thread, read four
// at a time) for the selection loop.
// emergency = (needed / left)^2 is pow(float, 2) in double narrowed to float (CHS.hpp:890-891): the exact product
// rounded once, which is what the fp32 multiply gives.
// @region load dispatch
#define DISPATCH_WARPS 4
#define DISPATCH_LD 36   // floats per lane row: a multiple of 4 (128-bit reads) that keeps the quarter-warps on distinct banks
__device__ __forceinline__ int dispatch_pick(float *es, int n4) {
    float best = -1.0f;   // piles without a car and cars already taken hold -1
    int bi = 0;
    const float4 *row = reinterpret_cast<const float4 *>(es);
    for (int j = 0; j < n4; ++j) {
        const float4 v = row[j];
        if (v.x > best) { best = v.x; bi = 4 * j; }   // strict: equal urgencies keep pile order
        if (v.y > best) { best = v.y; bi = 4 * j + 1; }
        if (v.z > best) { best = v.z; bi = 4 * j + 2; }
        if (v.w > best) { best = v.w; bi = 4 * j + 3; }
    }
    es[bi] = -1.0f;
    return bi;
}

// calculate_needed (CHS.hpp:879-898 / 1044-1063) from a pile record: s2t(soc) is cached, s2t(target) is a table
__device__ __forceinline__ float pile_emergency(const float *tlut, const uint4 r) {
    const float needed = __ldg(&tlut[PW_TU(r.w)]) - __uint_as_float(r.y);
    const int left = (int) PW_STAY(r.w) - (int) PW_ALREADY(r.w);
    if (!(needed > 0.0f)) return 0.0f;
    if ((float) left <= ceilf(needed)) return 10.0f;
    const float q = needed / (float) left;
    return q * q;
}

__device__ __forceinline__ uint32_t dispatch_station(const ChubParams &P, float *es, float *ps, int h,
                                                     int slot0, int n, int type, int cp, uint32_t occ, float load) {
    const uint4 *rp = P.pile + chub_pile_index(h, slot0, P.S0 + P.S1);
    const int n4 = (n + 3) >> 2;
    for (int j = 0; j < n4; ++j) reinterpret_cast<float4 *>(es)[j] = make_float4(-1.0f, -1.0f, -1.0f, -1.0f);
    const float *tlut = P.s2t_lut + (cp * 2 + type) * 1000;
    float mn = 0.0f, mx = 0.0f;
    for (uint32_t m = occ; m;) {   // calculate_needed + calculate_output, ascending pile order, 8 records in flight
        int idx[8];
        uint4 r[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            idx[u] = -1;
            if (m) {
                idx[u] = __ffs(m) - 1;
                m &= m - 1u;
                r[u] = rp[idx[u] * 32];
            }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (idx[u] < 0) break;
            const float e = pile_emergency(tlut, r[u]);
            const float p = __uint_as_float(r[u].z);
            mx += p;
            if (e > 8.0f) mn += p;
            es[idx[u]] = e;
            ps[idx[u]] = p;
        }
    }
    if (load > mx) load = mx;                  // catch_load
    else if (load < mn) load = mn;
    uint32_t on = 0u;
    int cars = __popc(occ);
    if (cp) {
        const float cpw = type == CHUB_SLOW ? (float) 5.254973139368931 : (float) 36.44764034125146;
        int k = (int) roundf(load / cpw);
        for (; cars > 0 && k > 0; --k, --cars) on |= 1u << dispatch_pick(es, n4);
    } else {
        const double lim = (double) load + 0.0001;
        float cum = 0.0f;
        for (; cars > 0; --cars) {
            const int bi = dispatch_pick(es, n4);
            cum += ps[bi];
            if (!(lim >= (double) cum)) break;   // powers are >= 0: the running sum never comes back under the target
            on |= 1u << bi;
        }
    }
    return on;
}

__global__ void __launch_bounds__(DISPATCH_WARPS * 32) chub_dispatch_kernel(const __grid_constant__ ChubParams P,
                                                                           unsigned long long *out) {
    __shared__ __align__(16) float es[DISPATCH_WARPS * 32][DISPATCH_LD], ps[DISPATCH_WARPS * 32][DISPATCH_LD];   // [thread][pile]
    const int h = blockIdx.x * (DISPATCH_WARPS * 32) + threadIdx.x;
    if (h >= P.H) return;
    const ChubClass &c = P.classes[P.cls ? __ldg(&P.cls[h]) : 0];
    const uint2 occ = P.occ[h];
    float load0 = __ldg(P.actions + (size_t) h * 4), load1 = __ldg(P.actions + (size_t) h * 4 + 1);
    if (load0 != load0 || load1 != load1) {   // a NaN target would slip through catch_load's comparisons
        P.err[h] |= CHUB_HUBERR_BAD_LOAD;
        if (load0 != load0) load0 = 0.0f;
        if (load1 != load1) load1 = 0.0f;
    }
    const uint32_t on0 = dispatch_station(P, es[threadIdx.x], ps[threadIdx.x], h, 0, c.n[0], c.type[0], c.constant, occ.x, load0);
    const uint32_t on1 = dispatch_station(P, es[threadIdx.x], ps[threadIdx.x], h, P.S0, c.n[1], c.type[1], c.constant, occ.y, load1);
    out[h] = (unsigned long long) on0 | ((unsigned long long) on1 << c.n[0]);
}

void chub_launch_dispatch(const ChubParams &p, unsigned long long *act_bits, cudaStream_t s) {
    const int threads = DISPATCH_WARPS * 32;
    chub_dispatch_kernel<<<(p.H + threads - 1) / threads, threads, 0, s>>>(p, act_bits);
}

// ---- per-pile introspection: the stations' `situation` columns (CHS.hpp:204-231, read through main.cpp:54-74,
// 142-184) as dense device matrices.  out: float32 [CHUB_SIT_COLUMNS][H][S0+S1]; slots of station 0 come first, slots
// beyond a station's pile count and piles without a car read 0 like a reset position.
// @region situation
__global__ void __launch_bounds__(256) chub_situation_kernel(const __grid_constant__ ChubParams P, float *out) {
    const int S = P.S0 + P.S1;
    const long long idx = (long long) blockIdx.x * 256 + threadIdx.x;
    if (idx >= (long long) P.H * S) return;
    const int h = (int) (idx / S), slot = (int) (idx - (long long) h * S);
    const ChubClass &c = P.classes[P.cls ? __ldg(&P.cls[h]) : 0];
    const int st = slot >= P.S0, i = slot - (st ? P.S0 : 0);
    float v[CHUB_SIT_COLUMNS] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    if (i < c.n[st]) {
        const size_t at = chub_pile_index(h, slot, S);
        const uint4 r = P.pile[at];
        if (r.w & PW_CAR) {
            v[CHUB_SIT_CAR] = 1.0f;
            v[CHUB_SIT_CHARGE] = (r.w & PW_CHARGE) ? 1.0f : 0.0f;
            v[CHUB_SIT_EMERGENCY] = pile_emergency(P.s2t_lut + (c.constant * 2 + c.type[st]) * 1000, r);
            v[CHUB_SIT_POWER] = __uint_as_float(r.z);
            v[CHUB_SIT_SOC] = __uint_as_float(r.x);
            v[CHUB_SIT_INIT_SOC] = P.arrive[at];
            v[CHUB_SIT_TARGET_SOC] = target_of(PW_TU(r.w));
        }
    }
    const size_t plane = (size_t) P.H * S;
#pragma unroll
    for (int k = 0; k < CHUB_SIT_COLUMNS; ++k) out[k * plane + idx] = v[k];
}

void chub_launch_situation(const ChubParams &p, float *out, cudaStream_t s) {
    const long long n = (long long) p.H * (p.S0 + p.S1);
    chub_situation_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(p, out);
}

// One hub's scalars in one place (chub_hub_report): what the single-hub gym wrapper reads back after every step
// (evcssp_manager.py:364-372 real_state, the attributes set in step()).  out[0..13] = diagnostics columns (0 without
// diagnostics), [14..19] = min / charge / max power of station 0 then 1, [20..21] = line, [22..23] = flow_in,
// [24] = t, [25] = store mass (g), [26] = FCEV arrivals, [27] = FCEV backlog, [28] = sticky error bits.
__global__ void chub_report_kernel(const __grid_constant__ ChubParams P, int h, double *out) {
    const int k = threadIdx.x;
    const uint4 a = P.hubA[h];
    double v = 0.0;
    if (k < CHUB_D_COUNT) v = P.diag ? P.diag[(size_t) k * P.Hpad + h] : 0.0;
    else if (k < CHUB_D_COUNT + 6) v = P.st_power ? (double) P.st_power[(size_t) (k - CHUB_D_COUNT) * P.Hpad + h] : 0.0;
    else if (k == 20) v = (double) ((a.z >> 8) & 0xff);
    else if (k == 21) v = (double) ((a.z >> 16) & 0xff);
    else if (k == 22) v = (double) ((int) (a.w & 0xff) - 128);
    else if (k == 23) v = (double) ((int) ((a.w >> 8) & 0xff) - 128);
    else if (k == 24) v = (double) (a.z & 0xff);
    else if (k == 25) v = P.store[h];
    else if (k == 26) v = (double) (a.w >> 24);
    else if (k == 27) v = (double) (a.z >> 24);
    else if (k == 28) v = (double) P.err[h];
    if (k < CHUB_REPORT_LEN) out[k] = v;
}
void chub_launch_report(const ChubParams &p, int h, double *out, cudaStream_t s) { chub_report_kernel<<<1, 32, 0, s>>>(p, h, out); }

// ---- curve sweep (parity instrument, tests/test_gpu_curves.py): the charging curves exactly as the step kernel compiles
// them, evaluated for `n` consecutive float32 bit patterns starting at `first_bits`, so that a test can walk EVERY
// reachable argument and compare with glibc (CHS.hpp:495-588, 621-725).  which: 0 slow t2p (add_car), 1 slow car_step
// soc, 2 slow s2t, 3 fast t2p, 4 fast car_step soc, 5 fast s2t, 6 slow car_step power, 7 fast car_step power;
// +8 = the constant-power variants.
// @region curve sweep
template <int TYPE, int CP>
__device__ __forceinline__ float sweep_eval(int f, float x) {
    if (f == 0) return curve_t2p<TYPE, CP>(x);
    if (f == 2) return curve_s2t<TYPE, CP>(x);
    float soc, power;
    if (curve_seg<TYPE, CP>(x) == 0) curve_step<TYPE, CP, 0>(x, soc, power);
    else curve_step<TYPE, CP, 1>(x, soc, power);
    return f == 1 ? soc : power;
}
__global__ void __launch_bounds__(256) chub_curve_kernel(int which, uint32_t first_bits, long long n, float *out) {
    for (int i = threadIdx.x; i < CHUB_LOG_N * 4; i += 256)   // slow_step reads the log table from shared memory
        reinterpret_cast<uint32_t *>(chub_smem)[i] = reinterpret_cast<const uint32_t *>(kLogTab)[i];
    __syncthreads();
    const int cp = which >> 3, w = which & 7;
    const int type = (w == 0 || w == 1 || w == 2 || w == 6) ? CHUB_SLOW : CHUB_FAST;
    const int f = (w == 0 || w == 3) ? 0 : (w == 2 || w == 5) ? 2 : (w == 1 || w == 4) ? 1 : 3;
    for (long long i = (long long) blockIdx.x * 256 + threadIdx.x; i < n; i += (long long) gridDim.x * 256) {
        const float x = __uint_as_float(first_bits + (uint32_t) i);
        float y;
        if (type == CHUB_SLOW) y = cp ? sweep_eval<CHUB_SLOW, 1>(f, x) : sweep_eval<CHUB_SLOW, 0>(f, x);
        else y = cp ? sweep_eval<CHUB_FAST, 1>(f, x) : sweep_eval<CHUB_FAST, 0>(f, x);
        out[i] = y;
    }
}
void chub_launch_curve_sweep(int which, uint32_t first_bits, long long n, float *out, cudaStream_t s) {
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    chub_curve_kernel<<<(unsigned) blocks, 256, CHUB_LOG_N * 16, s>>>(which, first_bits, n, out);
}

// Column sums of the episode statistics (the 32-byte variant of the gather): fixed grid and a fixed summation order,
// so the result does not depend on scheduling.  fp64 accumulators over float4 rows.
// @region stats sums
__global__ void __launch_bounds__(256) chub_stats_partial_kernel(const float4 *stats, int H, double *partial) {
    __shared__ double sh[4][256];
    double a = 0.0, b = 0.0, c = 0.0, d = 0.0;
    for (int h = blockIdx.x * 256 + threadIdx.x; h < H; h += CHUB_STATS_BLOCKS * 256) {
        const float4 v = stats[h];
        a += (double) v.x; b += (double) v.y; c += (double) v.z; d += (double) v.w;
    }
    sh[0][threadIdx.x] = a; sh[1][threadIdx.x] = b; sh[2][threadIdx.x] = c; sh[3][threadIdx.x] = d;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if ((int) threadIdx.x < w)
            for (int k = 0; k < 4; ++k) sh[k][threadIdx.x] += sh[k][threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x < 4) partial[threadIdx.x * CHUB_STATS_BLOCKS + blockIdx.x] = sh[threadIdx.x][0];
}
__global__ void chub_stats_final_kernel(const double *partial, double *sums) {
    const int k = threadIdx.x;
    if (k >= 4) return;
    double a = 0.0;
    for (int j = 0; j < CHUB_STATS_BLOCKS; ++j) a += partial[k * CHUB_STATS_BLOCKS + j];
    sums[k] = a;
}
void chub_launch_stats_sums(const float4 *stats, int H, double *partial, double *sums, cudaStream_t s) {
    chub_stats_partial_kernel<<<CHUB_STATS_BLOCKS, 256, 0, s>>>(stats, H, partial);
    chub_stats_final_kernel<<<1, 32, 0, s>>>(partial, sums);
}

int chub_step_shape(const ChubParams &p) {
    int shape = p.block;
    if (shape == 0) {
        // One block per SM; a block of W warps steps units of W tiles, one tile per warp, unit after unit.  Large blocks
        // run a tile cheaper than small ones (measured cost per tile: 28 warps 1.00, 16 warps 1.12, 8 warps 1.2, 4 warps
        // 1.3), but what decides for a batch that is not many times the machine is how its units divide among the SMs:
        // 4,096 tiles are ONE round of 28-warp units on 147 SMs, but two rounds -- the second almost empty -- of 24-warp
        // units.  So: the shape with the least estimated time = rounds x tiles in flight per SM x cost per tile.
        const int ws = step_scratch_bytes<0>(p), sms = p.num_sms > 0 ? p.num_sms : 148;
        const long long tiles = ((long long) (p.h_end - p.h_begin) + 31) / 32;
        const int NC = 6;
        const int cand[NC] = {28, 24, 20, 16, 8, 4};
        const double cost[NC] = {1.00, 1.04, 1.08, 1.12, 1.2, 1.3};
        double best = 0.0;
        shape = 410;
        for (int k = 0; k < NC; ++k) {
            if (CHUB_SMEM_TABLES + cand[k] * ws + 1024 > 227 * 1024) continue;   // the warps' scratch must fit
            const long long units = (tiles + cand[k] - 1) / cand[k], rounds = (units + sms - 1) / sms;
            const double est = (double) rounds * cand[k] * cost[k];
            if (best == 0.0 || est < best) { best = est; shape = cand[k] * 100 + 10; }
        }
    }
    return shape;
}
cudaError_t chub_launch_step(const ChubParams &p, cudaStream_t s, unsigned long long *tickets, bool *self_ticked) {
    const int shape = chub_step_shape(p);
    switch (shape) {  // warps per block * 100 + min blocks per SM * 10
        case 410: return launch_step_variant<4, 1, 0>(p, s, tickets, self_ticked);
        case 810: return launch_step_variant<8, 1, 0>(p, s, tickets, self_ticked);
        case 1610: return launch_step_variant<16, 1, 0>(p, s, tickets, self_ticked);
        case 2010: return launch_step_variant<20, 1, 0>(p, s, tickets, self_ticked);
        case 2410: return launch_step_variant<24, 1, 0>(p, s, tickets, self_ticked);
        case 2810: return launch_step_variant<28, 1, 0>(p, s, tickets, self_ticked);
        default: return launch_step_variant<16, 1, 0>(p, s, tickets, self_ticked);
    }
}
void chub_launch_reset(const ChubParams &p, cudaStream_t s) {
    const int block = 128;
    chub_reset_kernel<<<(p.H + block - 1) / block, block, 0, s>>>(p);
}
void chub_launch_tick(long long *step_index, int n, unsigned long long *queue_base, unsigned long long tickets, cudaStream_t s) {
    chub_tick_kernel<<<1, 1, 0, s>>>(step_index, n, queue_base, tickets);
}
