// chub_hostpack.cpp -- host side of chub_step_host: action rows -> 16-byte packed records, on a small thread pool.
//
// A per-pile action carries one bit: action_to_real (evcssp_manager.py:384-404) switches pile j on iff
// (a[j] + 1)/2 >= 0.5 in float32, i.e. iff a[j] >= -2^-25 (csrc/chub_kernels.cu load_actions, tests/test_host_logic.py).
// PCIe is what bounds the host-buffer step (98.6 MB of float rows per step at config 5), so the rows are thresholded
// HERE, next to the memory they live in, and only {64 on-bits, a[-2], a[-1]} = 16 bytes per hub cross the bus.
// This is input marshalling, not stepping: nothing of the hub dynamics runs on the host.
#include <immintrin.h>
#include <pthread.h>
#include <sched.h>
#include <stddef.h>
#include <stdlib.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
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
#define CHUB_AVX2 __attribute__((target("avx2")))
#include <stdlib.h>
static const int g_prefetch_rows = getenv("CHUB_PACK_PREFETCH") ? atoi(getenv("CHUB_PACK_PREFETCH")) : 16;
CHUB_AVX2 static inline uint32_t on_bits8(const float *p, __m256 thr) {   // _CMP_GE_OQ: false for NaN, like the kernel's compare
    return (uint32_t) _mm256_movemask_ps(_mm256_cmp_ps(_mm256_loadu_ps(p), thr, _CMP_GE_OQ));
}

// Every row has the same na: NB = ceil(na / 8) blocks of eight entries, the last one placed at na - 8 (it overlaps its
// predecessor unless na is a multiple of 8; OR-ing a bit twice changes nothing).  NB is a compile-time constant so the
// row body is straight-line code.
template <int NB>
CHUB_AVX2 static void pack_rows_avx2_nb(const float *rows, int stride, int na, int64_t n, uint32_t *out) {
    const __m256 thr = _mm256_set1_ps(-0x1p-25f);
    const int last = na - 8;
    const int64_t pf = (int64_t) g_prefetch_rows * stride;
    for (int64_t r = 0; r < n; ++r) {
        const float *row = rows + r * stride;
        if (pf) {   // a core keeps few demand misses in flight; asking early for the lines of a later row adds to them
            const char *q = (const char *) (row + pf);
#pragma GCC unroll 4
            for (int k = 0; k < NB; k += 2) _mm_prefetch(q + 32 * k, _MM_HINT_T0);
        }
        uint64_t bits = (uint64_t) on_bits8(row + last, thr) << last;
#pragma GCC unroll 8
        for (int b = 0; b < NB - 1; ++b) bits |= (uint64_t) on_bits8(row + 8 * b, thr) << (8 * b);
        uint64_t tail;
        memcpy(&tail, row + na, 8);
        _mm_storeu_si128((__m128i *) (out + 4 * r), _mm_set_epi64x((long long) tail, (long long) bits));
    }
}

CHUB_AVX2 static void pack_rows_avx2(const float *rows, int stride, const int32_t *na_row, int na_all, int64_t n, uint32_t *out) {
    if (!na_row && na_all >= 8) {
        switch ((na_all + 7) / 8) {
        case 1: return pack_rows_avx2_nb<1>(rows, stride, na_all, n, out);
        case 2: return pack_rows_avx2_nb<2>(rows, stride, na_all, n, out);
        case 3: return pack_rows_avx2_nb<3>(rows, stride, na_all, n, out);
        case 4: return pack_rows_avx2_nb<4>(rows, stride, na_all, n, out);
        case 5: return pack_rows_avx2_nb<5>(rows, stride, na_all, n, out);
        case 6: return pack_rows_avx2_nb<6>(rows, stride, na_all, n, out);
        case 7: return pack_rows_avx2_nb<7>(rows, stride, na_all, n, out);
        default: return pack_rows_avx2_nb<8>(rows, stride, na_all, n, out);
        }
    }
    const __m256 thr = _mm256_set1_ps(-0x1p-25f);
    for (int64_t r = 0; r < n; ++r) {
        const float *row = rows + r * stride;
        const int na = na_row ? na_row[r] : na_all;
        if (na < 8) {
            pack_row_scalar(row, na, out + 4 * r);
            continue;
        }
        uint64_t bits = (uint64_t) on_bits8(row + na - 8, thr) << (na - 8);
        for (int j = 0; j + 8 < na; j += 8) bits |= (uint64_t) on_bits8(row + j, thr) << j;
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

// Work is handed out in blocks of CHUB_PACK_BLOCK rows from a per-range ticket counter, so a thread that the host
// schedules late (or a busy core) costs one block, not its share of the range; the caller packs along while it waits.
#define CHUB_PACK_BLOCK 256
struct alignas(64) PackRange {
    std::atomic<int64_t> next{0};   // next block of this range to hand out
    char pad0[56];
    std::atomic<int64_t> done{0};   // blocks of this range finished
    int64_t n_blocks = 0;
    char pad1[48];
};

struct ChubPackPool {
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv;
    std::atomic<uint64_t> epoch{0};
    std::atomic<int> idle{0};       // workers that are out of the current job
    bool stop = false;
    // the job (written by the caller before the epoch moves)
    const float *rows = nullptr;
    const int32_t *na_row = nullptr;
    int na_all = 0, stride = 0;
    int64_t n_rows = 0, per = 0;
    uint32_t *out = nullptr;
    int n_ranges = 0;
    PackRange range[CHUB_PACK_MAX_RANGES];

    void work_on(int r) {
        PackRange &R = range[r];
        const int64_t b = (int64_t) r * per, n = (b + per <= n_rows ? per : n_rows - b);
        for (;;) {
            const int64_t k = R.next.fetch_add(1, std::memory_order_relaxed);
            if (k >= R.n_blocks) return;
            const int64_t lo = b + k * CHUB_PACK_BLOCK, hi = std::min(b + n, lo + CHUB_PACK_BLOCK);
            chub_pack_rows(rows + lo * stride, stride, na_row ? na_row + lo : nullptr, na_all, hi - lo, out + 4 * lo);
            R.done.fetch_add(1, std::memory_order_release);
        }
    }
    void run() {
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
            for (int r = 0; r < n_ranges; ++r) work_on(r);
            idle.fetch_add(1, std::memory_order_release);
        }
    }
};

extern "C" ChubPackPool *chub_pack_pool_create(int32_t n_threads) {
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 64) n_threads = 64;
    ChubPackPool *p = new ChubPackPool();
    p->idle.store(n_threads);
    for (int i = 0; i < n_threads; ++i) p->workers.emplace_back([p] { p->run(); });
    return p;
}

extern "C" ChubPackPool *chub_pack_pool_create_on(int32_t n_threads, const char *cpulist) {
    ChubPackPool *p = chub_pack_pool_create(n_threads);
#if defined(__linux__)
    if (p && cpulist) {
        cpu_set_t set;
        CPU_ZERO(&set);
        int n_set = 0;
        for (const char *c = cpulist; *c;) {   // "a-b,c,d-e"
            char *end;
            long a = strtol(c, &end, 10);
            if (end == c) break;
            long b = a;
            if (*end == '-') { c = end + 1; b = strtol(c, &end, 10); }
            for (long k = a; k <= b && k < CPU_SETSIZE; ++k) { CPU_SET((int) k, &set); ++n_set; }
            c = (*end == ',') ? end + 1 : end;
            if (*end != ',') break;
        }
        if (n_set > 0)
            for (auto &t : p->workers) pthread_setaffinity_np(t.native_handle(), sizeof set, &set);
    }
#endif
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

extern "C" int32_t chub_pack_pool_threads(const ChubPackPool *p) { return p ? (int32_t) p->workers.size() : 0; }

// Start packing n_rows rows in ranges of `per` rows (at most CHUB_PACK_MAX_RANGES).  Returns the number of ranges.
extern "C" int32_t chub_pack_pool_start(ChubPackPool *p, const float *rows, int32_t stride, const int32_t *na_row,
                                        int32_t na_all, int64_t n_rows, int64_t per, void *out) {
    if (per < 1 || n_rows < 1) return -1;
    const int64_t nr = (n_rows + per - 1) / per;
    if (nr > CHUB_PACK_MAX_RANGES) return -1;
    const int n_ranges = (int) nr, T = (int) p->workers.size();
    while (p->idle.load(std::memory_order_acquire) < T) cpu_relax();   // every worker must be out of the previous job
    p->idle.store(0, std::memory_order_relaxed);
    p->rows = rows; p->stride = stride; p->na_row = na_row; p->na_all = na_all;
    p->n_rows = n_rows; p->per = per; p->out = (uint32_t *) out; p->n_ranges = n_ranges;
    for (int r = 0; r < n_ranges; ++r) {
        const int64_t b = (int64_t) r * per, n = (b + per <= n_rows ? per : n_rows - b);
        p->range[r].n_blocks = (n + CHUB_PACK_BLOCK - 1) / CHUB_PACK_BLOCK;
        p->range[r].next.store(0, std::memory_order_relaxed);
        p->range[r].done.store(0, std::memory_order_relaxed);
    }
    {
        std::lock_guard<std::mutex> lk(p->mu);
        p->epoch.fetch_add(1, std::memory_order_release);
    }
    p->cv.notify_all();
    return n_ranges;
}

// Block until every worker is out of the current job (the caller's rows are no longer being read).
extern "C" void chub_pack_pool_drain(ChubPackPool *p) {
    const int T = (int) p->workers.size();
    for (int r = 0; r < p->n_ranges; ++r) p->work_on(r);
    int spins = 0;
    while (p->idle.load(std::memory_order_acquire) < T) {
        if (++spins < 20000) cpu_relax();
        else std::this_thread::yield();
    }
}

// Block until range r is packed; the caller takes blocks of it too instead of idling.
extern "C" void chub_pack_pool_wait(ChubPackPool *p, int32_t r) {
    p->work_on(r);
    PackRange &R = p->range[r];
    int spins = 0;
    while (R.done.load(std::memory_order_acquire) < R.n_blocks) {
        if (++spins < 20000) cpu_relax();
        else std::this_thread::yield();
    }
}
