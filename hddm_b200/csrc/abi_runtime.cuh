// abi_runtime.cuh -- host-side runtime shared by the extern "C" translation units (abi*.cu): the last-error
// string, status macros, the library stream, the grow-only device arena and page-locked staging buffer,
// and the argument checks several entry points share.  State is defined once, in abi.cu.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/hddm_b200.h"
#include "kernels.cuh"

namespace hddm {
namespace rt {

extern thread_local char g_err[512];

inline int fail(int code, const char* msg)
{
    snprintf(g_err, sizeof(g_err), "%s", msg);
    return code;
}

#define CUDA_TRY(expr)                                                                    \
    do {                                                                                  \
        cudaError_t e__ = (expr);                                                         \
        if (e__ != cudaSuccess) {                                                         \
            snprintf(g_err, sizeof(g_err), "CUDA error %s at %s:%d (%s)",                 \
                     cudaGetErrorName(e__), __FILE__, __LINE__, cudaGetErrorString(e__)); \
            return HDDM_ERR_CUDA;                                                         \
        }                                                                                 \
    } while (0)

#define TRY(expr)                 \
    do {                          \
        int rc__ = (expr);        \
        if (rc__) return rc__;    \
    } while (0)

struct Runtime {
    std::mutex mu;
    bool ready = false;
    int device = -1;  // the ONE device this process uses (stream, arena, kernel attributes belong to it)
    cudaStream_t stream = nullptr;
    char* arena = nullptr;  // grow-only device scratch, carved per call
    size_t arena_bytes = 0;
    unsigned long long arena_epoch = 0;  // counts ensure_arena calls: whoever keeps data there across calls checks it
    char* pinned = nullptr;  // grow-only page-locked staging for the small per-call inputs / outputs
    size_t pinned_bytes = 0;
};
extern Runtime g_rt;

// One process per GPU: all library state lives on the device that was current at first use.  A call made
// with another device current fails loudly instead of enqueueing on a stream of the wrong device.
inline int ensure_ready()
{
    if (g_rt.ready) {
        int cur = -1;
        if (cudaGetDevice(&cur) != cudaSuccess || cur != g_rt.device) {
            snprintf(g_err, sizeof(g_err), "hddm_b200 is bound to CUDA device %d but device %d is current "
                     "(one process per GPU; see hddm_b200_init)", g_rt.device, cur);
            return HDDM_ERR_ARG;
        }
        return HDDM_OK;
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        snprintf(g_err, sizeof(g_err), "no usable CUDA device (%s); hddm_b200 has no CPU fallback",
                 e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return HDDM_ERR_CUDA;
    }
    CUDA_TRY(cudaGetDevice(&g_rt.device));
    CUDA_TRY(cudaStreamCreateWithFlags(&g_rt.stream, cudaStreamNonBlocking));
    g_rt.ready = true;
    return HDDM_OK;
}

inline int ensure_arena(size_t bytes)
{
    ++g_rt.arena_epoch;
    if (bytes <= g_rt.arena_bytes) return HDDM_OK;
    if (g_rt.arena) {
        CUDA_TRY(cudaStreamSynchronize(g_rt.stream));
        CUDA_TRY(cudaFree(g_rt.arena));
        g_rt.arena = nullptr;
        g_rt.arena_bytes = 0;
    }
    const size_t want = bytes + bytes / 4 + (1 << 20);
    CUDA_TRY(cudaMalloc(&g_rt.arena, want));
    g_rt.arena_bytes = want;
    return HDDM_OK;
}

inline int ensure_pinned(size_t bytes)
{
    if (bytes <= g_rt.pinned_bytes) return HDDM_OK;
    if (g_rt.pinned) {
        CUDA_TRY(cudaStreamSynchronize(g_rt.stream));
        CUDA_TRY(cudaFreeHost(g_rt.pinned));
        g_rt.pinned = nullptr;
        g_rt.pinned_bytes = 0;
    }
    const size_t want = bytes + bytes / 2 + (64 << 10);
    CUDA_TRY(cudaMallocHost(&g_rt.pinned, want));
    g_rt.pinned_bytes = want;
    return HDDM_OK;
}

constexpr size_t kPinnedMax = 8u << 20;  // larger inputs are copied straight from the caller's buffer

inline size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

// Two-phase arena layout: reserve() everything, ensure_arena(total), then at().
struct Layout {
    size_t total = 0;
    size_t reserve(size_t bytes)
    {
        const size_t off = total;
        total += align_up(bytes ? bytes : 1);
        return off;
    }
    template <class T>
    static T* at(size_t off) { return reinterpret_cast<T*>(g_rt.arena + off); }
};

inline int check_numerics(int n_st, int n_sz, int use_adaptive)
{
    if (use_adaptive && (n_st > HDDM_MAX_DEPTH || n_sz > HDDM_MAX_DEPTH))
        return fail(HDDM_ERR_ARG, "adaptive recursion depth n_st/n_sz > 24 is not supported");
    return HDDM_OK;
}

inline Numerics make_numerics(double err, int n_st, int n_sz, int use_adaptive, double simps_err)
{
    Numerics nm;
    nm.err = err;
    nm.simps_err = simps_err;
    nm.n_st = n_st;
    nm.n_sz = n_sz;
    nm.use_adaptive = use_adaptive;
    return nm;
}

inline int64_t chunks_for(int64_t max_len)
{
    const int64_t c = (max_len + kTrialsPerBlock - 1) / kTrialsPerBlock;
    return c < 1 ? 1 : c;
}

inline bool out_of_support(double sv, double a, double z, double sz, double t, double st, double p_outlier)
{  // src/wfpt.pyx:395-396 == src/cdfdif_wrapper.pyx:23-24
    return (sv < 0) || (a <= 0) || (z < 0) || (z > 1) || (sz < 0) || (sz > 1) || (z + sz / 2. > 1) ||
           (z - sz / 2. < 0) || (t - st / 2. < 0) || (t < 0) || (st < 0) ||
           !((p_outlier >= 0) & (p_outlier <= 1));
}

}  // namespace rt
}  // namespace hddm
