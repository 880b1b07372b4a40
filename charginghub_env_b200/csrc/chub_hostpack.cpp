// chub_hostpack.cpp -- host side of chub_step_host: action rows -> 16-byte packed records, on a small thread pool.
//
// A per-pile action carries one bit: action_to_real (evcssp_manager.py:384-404) switches pile j on iff
// (a[j] + 1)/2 >= 0.5 in float32, i.e. iff a[j] >= -2^-25 (csrc/chub_kernels.cu load_actions, tests/test_host_logic.py).
// PCIe is what bounds the host-buffer step (98.6 MB of float rows per step at config 5), so the rows are thresholded
// HERE, next to the memory they live in, and only {64 on-bits, a[-2], a[-1]} = 16 bytes per hub cross the bus.
// This is input marshalling, not stepping: nothing of the hub dynamics runs on the host.
#include <immintrin.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#include "chub_hostpack.h"

static inline void pack_row_scalar(const float *row, int na, uint32_t *out) {
    uint64_t bits = 0;
    for (int j = 0; j < na; ++j) bits |= (uint64_t) (row[j] >= -0x1p-25f) << j;
    out[0] = (uint32_t) bits;
    out[1] = (uint32_t) (bits >> 32);
    memcpy(out + 2, row + na, 8);   // a[-2], a[-1] travel as they are
}

static void pack_rows_scalar(const float *rows, int stride, const int32_t *na_row, int na_all, int64_t n, uint32_t *out) {
    for (int64_t r = 0; r < n; ++r) pack_row_scalar(rows + r * stride, na_row ? na_row[r] : na_all, out + 4 * r);
}

#if defined(__x86_64__)
__attribute__((target("avx2"))) static void pack_rows_avx2(const float *rows, int stride, const int32_t *na_row, int na_all,
                                                           int64_t n, uint32_t *out) {
    const __m256 thr = _mm256_set1_ps(-0x1p-25f);
    for (int64_t r = 0; r < n; ++r) {
        const float *row = rows + r * stride;
        const int na = na_row ? na_row[r] : na_all;
        if (na < 8) {
            pack_row_scalar(row, na, out + 4 * r);
            continue;
        }
        uint64_t bits = 0;
        int j = 0;
        for (; j + 8 <= na; j += 8)   // _CMP_GE_OQ: false for NaN, like the kernel's compare
            bits |= (uint64_t) (uint32_t) _mm256_movemask_ps(_mm256_cmp_ps(_mm256_loadu_ps(row + j), thr, _CMP_GE_OQ)) << j;
        if (j < na)                   // the last, overlapping block re-derives bits already set: OR is idempotent
            bits |= (uint64_t) (uint32_t) _mm256_movemask_ps(_mm256_cmp_ps(_mm256_loadu_ps(row + na - 8), thr, _CMP_GE_OQ)) << (na - 8);
        uint32_t *o = out + 4 * r;
        o[0] = (uint32_t) bits;
        o[1] = (uint32_t) (bits >> 32);
        memcpy(o + 2, row + na, 8);
    }
}
#endif

extern "C" void chub_pack_rows(const float *rows, int32_t stride, const int32_t *na_row, int32_t na_all, int64_t n_rows,
                               void *out) {
#if defined(__x86_64__)
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2) return pack_rows_avx2(rows, stride, na_row, na_all, n_rows, (uint32_t *) out);
#endif
    pack_rows_scalar(rows, stride, na_row, na_all, n_rows, (uint32_t *) out);
}

// ---- the pool: every worker packs its slice of range 0, 1, 2, ... and bumps that range's counter; the caller waits
// for one range at a time and sends it on its way while the workers are already in the next one ----
static inline void cpu_relax() {
#if defined(__x86_64__)
    _mm_pause();
#endif
}

struct ChubPackPool {
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv;
    std::atomic<uint64_t> epoch{0};
    bool stop = false;
    // the job (written by the caller before the epoch moves)
    const float *rows = nullptr;
    const int32_t *na_row = nullptr;
    int na_all = 0, stride = 0;
    int64_t n_rows = 0, per = 0;
    uint32_t *out = nullptr;
    int n_ranges = 0;
    std::atomic<int> done[CHUB_PACK_MAX_RANGES];
    bool busy = false;   // a job was started and not every range of it has been waited for

    void slice(int worker, int range) {
        const int T = (int) workers.size();
        const int64_t b = (int64_t) range * per, n = (b + per <= n_rows ? per : n_rows - b);
        const int64_t lo = b + n * worker / T, hi = b + n * (worker + 1) / T;
        if (hi > lo) chub_pack_rows(rows + lo * stride, stride, na_row ? na_row + lo : nullptr, na_all, hi - lo, out + 4 * lo);
    }
    void run(int worker) {
        uint64_t seen = 0;
        for (;;) {
            int spins = 0;
            while (epoch.load(std::memory_order_acquire) == seen) {
                if (++spins < 4000) { cpu_relax(); continue; }   // a step loop comes back within microseconds ...
                std::unique_lock<std::mutex> lk(mu);            // ... anything slower finds the workers asleep
                cv.wait(lk, [&] { return stop || epoch.load(std::memory_order_acquire) != seen; });
                break;
            }
            {
                std::lock_guard<std::mutex> lk(mu);
                if (stop) return;
            }
            seen = epoch.load(std::memory_order_acquire);
            for (int r = 0; r < n_ranges; ++r) {
                slice(worker, r);
                done[r].fetch_add(1, std::memory_order_release);
            }
        }
    }
};

extern "C" ChubPackPool *chub_pack_pool_create(int32_t n_threads) {
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 64) n_threads = 64;
    ChubPackPool *p = new ChubPackPool();
    for (auto &d : p->done) d.store(0);
    for (int i = 0; i < n_threads; ++i) p->workers.emplace_back([p, i] { p->run(i); });
    return p;
}

extern "C" void chub_pack_pool_destroy(ChubPackPool *p) {
    if (!p) return;
    {
        std::lock_guard<std::mutex> lk(p->mu);
        p->stop = true;
        p->epoch.fetch_add(1, std::memory_order_release);
    }
    p->cv.notify_all();
    for (auto &t : p->workers) t.join();
    delete p;
}

extern "C" void chub_pack_pool_wait(ChubPackPool *p, int32_t r);
extern "C" int32_t chub_pack_pool_threads(const ChubPackPool *p) { return p ? (int32_t) p->workers.size() : 0; }

// Start packing n_rows rows in ranges of `per` rows (at most CHUB_PACK_MAX_RANGES).  Returns the number of ranges.
extern "C" int32_t chub_pack_pool_start(ChubPackPool *p, const float *rows, int32_t stride, const int32_t *na_row,
                                        int32_t na_all, int64_t n_rows, int64_t per, void *out) {
    int n_ranges = (int) ((n_rows + per - 1) / per);
    if (n_ranges > CHUB_PACK_MAX_RANGES || n_ranges < 1) return -1;
    if (p->busy) chub_pack_pool_wait(p, p->n_ranges - 1);   // the workers must be out of the previous job
    p->busy = true;
    p->rows = rows; p->stride = stride; p->na_row = na_row; p->na_all = na_all;
    p->n_rows = n_rows; p->per = per; p->out = (uint32_t *) out; p->n_ranges = n_ranges;
    for (int r = 0; r < n_ranges; ++r) p->done[r].store(0, std::memory_order_relaxed);
    {
        std::lock_guard<std::mutex> lk(p->mu);
        p->epoch.fetch_add(1, std::memory_order_release);
    }
    p->cv.notify_all();
    return n_ranges;
}

// Block until every worker has finished range r.
extern "C" void chub_pack_pool_wait(ChubPackPool *p, int32_t r) {
    const int T = (int) p->workers.size();
    int spins = 0;
    while (p->done[r].load(std::memory_order_acquire) < T) {
        if (++spins < 20000) cpu_relax();
        else std::this_thread::yield();
    }
    if (r == p->n_ranges - 1) p->busy = false;   // ranges finish in order on every worker
}
