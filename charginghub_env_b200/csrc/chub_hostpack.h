// chub_hostpack.h -- host-side packing of action rows for chub_step_host (see chub_hostpack.cpp).
#ifndef CHUB_HOSTPACK_H
#define CHUB_HOSTPACK_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#define CHUB_PACK_MAX_RANGES 32
typedef struct ChubPackPool ChubPackPool;
// rows: float32 [n_rows][stride]; the row's first na entries are per-pile actions, then a[-2], a[-1].
// na_row: per-row na, or NULL -> na_all.  out: [n_rows] records {bits lo, bits hi, a[-2], a[-1]} (16 bytes each).
void chub_pack_rows(const float *rows, int32_t stride, const int32_t *na_row, int32_t na_all, int64_t n_rows, void *out);
ChubPackPool *chub_pack_pool_create(int32_t n_threads);
// the same with the workers confined to a set of CPUs given as a sysfs cpulist ("0-15,32-47"); NULL = anywhere
ChubPackPool *chub_pack_pool_create_on(int32_t n_threads, const char *cpulist);
void chub_pack_pool_destroy(ChubPackPool *p);
int32_t chub_pack_pool_threads(const ChubPackPool *p);
int32_t chub_pack_pool_start(ChubPackPool *p, const float *rows, int32_t stride, const int32_t *na_row, int32_t na_all,
                             int64_t n_rows, int64_t per, void *out);
// returns when range r (and only range r) is packed; the caller takes blocks of that range while it waits
void chub_pack_pool_wait(ChubPackPool *p, int32_t r);
// returns when no worker reads the job's rows any more (error paths of the caller)
void chub_pack_pool_drain(ChubPackPool *p);
#ifdef __cplusplus
}
#endif
#endif
