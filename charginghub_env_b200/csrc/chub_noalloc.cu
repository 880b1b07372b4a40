// chub_noalloc.cu -- the scratch of chub_step_loads_static (include/chub.h): a file-scope device array, so that the
// scalar-load step can run for callers that must not allocate (code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
//
// chub_step_loads keeps the dispatch kernel's output -- one 64-bit word of pile on-bits per hub -- in a buffer it
// cudaMalloc's on its first call.  Here the same words live in g_static_act_bits: CHUB_STATIC_MAX_HUBS x 8 bytes = 8 MiB
// that EVERY GPU which loads libchub.so holds until the process exits.
//
// The array's name is a host-side stand-in in host code, not a device address: a kernel handed `g_static_act_bits`
// from the host would fault (or, where the GPU can reach host memory, quietly use the wrong memory).  The device
// address comes from cudaGetSymbolAddress, asked once per device with that device current (chub_static_act_bits, called
// from chub_create = set-up), in THIS file because the symbol has internal linkage.
//
// Sharing: one array per device, shared by every engine on it and by every call; it keeps the last call's contents.
// Nothing has to start at zero -- the dispatch kernel writes the word of every hub below n_hubs before the step kernel
// reads it, stream ordered -- but two calls that overlap on the GPU (two engines of one device on different streams,
// or one engine driven from two streams) RACE on it.  Serialise them, or use chub_step_loads.
#include <cuda_runtime.h>

#include <mutex>

#include "chub_kernels.cuh"

static __device__ __align__(256) unsigned long long g_static_act_bits[CHUB_STATIC_MAX_HUBS];

#define CHUB_STATIC_MAX_DEVICES 64
static std::mutex g_static_mu;
static unsigned long long *g_static_ptr[CHUB_STATIC_MAX_DEVICES];
static cudaError_t g_static_err[CHUB_STATIC_MAX_DEVICES];
static bool g_static_asked[CHUB_STATIC_MAX_DEVICES];

// Device address of the array on `device`, which must be the CURRENT device of the calling thread.  The first call for
// a device asks the runtime (this may load the module on that device: do it at set-up, not inside a capture); later
// calls return the cached answer and touch no CUDA API.
cudaError_t chub_static_act_bits(int device, unsigned long long **ptr) {
    if (!ptr || device < 0 || device >= CHUB_STATIC_MAX_DEVICES) return cudaErrorInvalidValue;
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_asked[device]) {
        void *p = nullptr;
        g_static_err[device] = cudaGetSymbolAddress(&p, g_static_act_bits);
        g_static_ptr[device] = g_static_err[device] == cudaSuccess ? (unsigned long long *) p : nullptr;
        g_static_asked[device] = true;
    }
    *ptr = g_static_ptr[device];
    return g_static_err[device];
}
