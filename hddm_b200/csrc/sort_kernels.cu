// sort_kernels.cu -- one-time layout step for the likelihood sweep: trials ordered by boundary, then |rt|.
//
// The series branch and term count of every decision-time node are monotone in |rt| for a given
// parameter set (src/pdf.pxi:36-60 through tt = (|rt| - t') / a^2), so when the 32 lanes of a warp
// hold neighbouring |rt| they take the same branch with the same K and the FP64 pipe is not spent on
// masked-off lanes (profiles/r1_v3: 19.7 of 32 lanes active per DFMA on unsorted trials).  The sum
// over trials is order-independent up to rounding, and the order is fixed (stable sort), so results
// stay bit-reproducible.  Sorting is library plumbing (CUB), not part of the measured hot path: it
// runs once per trial vector, which is then reused by every parameter set / MCMC step.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_segmented_sort.cuh>

#include "kernels.cuh"

namespace hddm {

namespace {

__global__ void k_abs_keys(const double* __restrict__ x, unsigned long long* __restrict__ keys, int64_t n)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        // boundary-major, then |rt|: lower-boundary trials (x <= 0) first, each group ascending in |x| (NaNs last
        // in their group), so that a warp's lanes share the boundary as well as the series branch
        const double v = x[i];
        keys[i] = ((unsigned long long)__double_as_longlong(v) & 0x7fffffffffffffffULL) |
                  (v > 0 ? 0x8000000000000000ULL : 0ULL);
    }
}

size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

// Scratch layout: keys_in[n] | keys_out[n] | cub temp.
size_t sort_trials_scratch_bytes(int64_t n, int64_t n_seg)
{
    if (n <= 0) return 256;
    size_t tmp = 0;
    unsigned long long* k = nullptr;
    double* v = nullptr;
    if (n_seg > 0) {
        const int64_t* off = nullptr;
        cub::DeviceSegmentedSort::StableSortPairs(nullptr, tmp, k, k, v, v, (int)n, (int)n_seg, off, off + 1);
    } else {
        cub::DeviceRadixSort::SortPairs(nullptr, tmp, k, k, v, v, n, 0, 64);
    }
    return 2 * align256((size_t)n * sizeof(unsigned long long)) + align256(tmp) + 256;
}

// d_sorted[i] = the trials of d_x ordered by |x| (within each segment [seg[s], seg[s+1]) when d_seg is
// given; trials outside every segment are copied through).  d_sorted must not alias d_x.
cudaError_t sort_trials(const double* d_x, double* d_sorted, int64_t n, const int64_t* d_seg, int64_t n_seg,
                        void* d_scratch, size_t scratch_bytes, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    if (n > 2147483647LL && n_seg > 0) return cudaErrorInvalidValue;
    const size_t kb = align256((size_t)n * sizeof(unsigned long long));
    if (scratch_bytes < 2 * kb + 256) return cudaErrorInvalidValue;
    unsigned long long* keys_in = reinterpret_cast<unsigned long long*>(d_scratch);
    unsigned long long* keys_out = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(d_scratch) + kb);
    void* tmp = reinterpret_cast<char*>(d_scratch) + 2 * kb;
    size_t tmp_bytes = scratch_bytes - 2 * kb;
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_abs_keys<<<(unsigned)blocks, 256, 0, stream>>>(d_x, keys_in, n);
    if (n_seg > 0) {
        // segments may not cover all of x: start from a copy
        cudaError_t e = cudaMemcpyAsync(d_sorted, d_x, (size_t)n * sizeof(double), cudaMemcpyDeviceToDevice, stream);
        if (e != cudaSuccess) return e;
        return cub::DeviceSegmentedSort::StableSortPairs(tmp, tmp_bytes, keys_in, keys_out, d_x, d_sorted, (int)n,
                                                         (int)n_seg, d_seg, d_seg + 1, stream);
    }
    return cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys_in, keys_out, d_x, d_sorted, n, 0, 64, stream);
}

}  // namespace hddm
