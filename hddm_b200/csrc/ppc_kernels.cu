// ppc_kernels.cu -- posterior-predictive batch: many posterior draws per launch.
//
// The reference simulates one data set per posterior draw by calling gen_rts_from_cdf once per draw
// (kabuki post_pred_gen -> hddm/likelihoods.py:76-81 -> hddm/generate.py:190-194 -> src/wfpt.pyx:323-354):
// a density table on an RT grid, its running sum, an inverse-CDF lookup per simulated trial, then the
// summary statistics of hddm/utils.py:241-267 on the simulated RTs.  Here the D draws of a batch share the
// grid and go through each stage together: k_pdf_array with blockIdx.y = draw (wfpt_kernels.cu) fills
// tables[D][m]; k_cumsum_batch (one CTA per draw) normalises them; k_sample_batch looks the RTs up; and
// k_ppc_stats (one CTA per draw) sorts the draw's RTs in shared memory and reduces them to 15 statistics,
// so that only D x 15 doubles need to leave the GPU.
#include <curand_kernel.h>

#include "kernels.cuh"

namespace hddm {

namespace {

constexpr int kScanThreads = 1024;
#ifndef HDDM_SCAN_COLS
#define HDDM_SCAN_COLS 8
#endif
constexpr int kScanCols = HDDM_SCAN_COLS;       // slice elements staged per round (8: two CTAs per SM)
constexpr int kScanPitch = kScanCols + 1;       // odd pitch: the row-wise pass is free of bank conflicts
constexpr size_t kScanTileBytes = (size_t)kScanThreads * kScanPitch * sizeof(double);

// Rectangle-rule running sum from index 1 with table[0] = 0 (src/wfpt.pyx:333-336), one CTA per draw.  The
// arithmetic is that of k_scan_normalise<1> (gen_kernels.cu), element for element: thread j owns the contiguous
// slice j and adds it in ascending order, slice totals are combined in ascending order -- so a batch and D single
// calls give the same tables.  What differs is how the numbers travel: slices are staged through a shared-memory
// tile, 8 elements of every slice per round, read and written as 64-byte runs (a thread walking its own slice
// in global memory costs a sector per element and lane), and the division by the last entry (:337-340) is left to
// the lookup, which divides the entries it probes by last[d] -- the same quotient, 17 of them per simulated trial
// instead of one pass over the table.
__global__ void __launch_bounds__(kScanThreads) k_cumsum_batch(double* tables, int64_t m, double* __restrict__ last)
{
    extern __shared__ double tile[];  // [kScanThreads][kScanPitch]
    __shared__ double slice_total[kScanThreads];
    __shared__ double slice_offset[kScanThreads];
    double* table = tables + (int64_t)blockIdx.x * m;
    const int j = threadIdx.x;
    const int64_t per = (m - 1 + kScanThreads - 1) / kScanThreads;
    const int64_t lo = 1 + (int64_t)j * per;
    const int64_t hi = lo + per < m ? lo + per : m;
    const int mine = hi > lo ? (int)(hi - lo) : 0;  // elements of this thread's slice
    // tile <-> the slices' elements [c0, c0 + kScanCols).  Thread j moves column j % kScanCols of the rows
    // j / kScanCols + k * (kScanThreads / kScanCols): one column test per round, one add and one bound test per
    // element (the index arithmetic of the obvious loop was half of this kernel's instructions)
    constexpr int kRowsPerStep = kScanThreads / kScanCols;
    const int col = j % kScanCols, row0 = j / kScanCols;
    const int64_t g0 = 1 + (int64_t)row0 * per + col, gstep = (int64_t)kRowsPerStep * per;
    double* const t0 = tile + row0 * kScanPitch + col;
    auto stage = [&](int64_t c0, bool store) {
        if (!(c0 + col < per)) return;
        const int64_t i0 = g0 + c0;
#pragma unroll
        for (int k = 0; k < kScanCols; ++k) {
            const int64_t i = i0 + k * gstep;
            if (i < m) {
                if (store) table[i] = t0[k * kRowsPerStep * kScanPitch];
                else t0[k * kRowsPerStep * kScanPitch] = table[i];
            }
        }
    };
    double run = 0.0;
    for (int64_t c0 = 0; c0 < per; c0 += kScanCols) {
        stage(c0, false);
        __syncthreads();
        const int n = mine - (int)c0 < kScanCols ? mine - (int)c0 : kScanCols;
        for (int c = 0; c < n; ++c) run += tile[j * kScanPitch + c];
        __syncthreads();
    }
    slice_total[j] = run;
    __syncthreads();
    if (j == 0) {
        double off = 0.0;
        for (int s = 0; s < kScanThreads; ++s) {
            slice_offset[s] = off;
            off += slice_total[s];
        }
    }
    __syncthreads();
    run = slice_offset[j];
    for (int64_t c0 = 0; c0 < per; c0 += kScanCols) {
        stage(c0, false);
        __syncthreads();
        const int n = mine - (int)c0 < kScanCols ? mine - (int)c0 : kScanCols;
        for (int c = 0; c < n; ++c) {
            run += tile[j * kScanPitch + c];
            tile[j * kScanPitch + c] = run;
        }
        __syncthreads();
        stage(c0, true);
        __syncthreads();
    }
    if (j == 0) table[0] = 0.0;
    if (hi == m && mine > 0) last[blockIdx.x] = run;  // the thread that owns table[m - 1]
    if (m == 1 && j == 0) last[blockIdx.x] = 0.0;     // (a one-point grid: 0 / 0, as the reference)
}

// rt = x[searchsorted(table_d, f, 'left')] + sign(rt) * (t_d | u * st_d + (t_d - st_d/2))
// (src/wfpt.pyx:344-353).  blockIdx.y = draw.
__global__ void __launch_bounds__(256) k_sample_batch(const double* __restrict__ xgrid,
                                                      const double* __restrict__ tables, int64_t m,
                                                      const double* __restrict__ f,
                                                      const double* __restrict__ u_delay, int64_t S,
                                                      const double* __restrict__ t_par,
                                                      const double* __restrict__ st_par,
                                                      unsigned long long seed, int64_t draw0,
                                                      const double* __restrict__ last, double* __restrict__ rts)
{
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t d = blockIdx.y;
    if (s >= S) return;
    const double* table = tables + d * m;
    const double t = t_par[d], st = st_par[d];
    const double total = last[d];  // table[m - 1] of the running sum: entries are normalised as they are probed
    double target, u = 0.0;
    if (f) {
        target = f[d * S + s];
        if (st != 0) u = u_delay[d * S + s];
    } else {  // device uniforms in [0, 1): Philox4x32-10, one counter sequence per simulated trial
        curandStatePhilox4_32_10_t rng;
        curand_init(seed, (unsigned long long)((draw0 + d) * S + s), 0, &rng);
        const double2 r = curand_uniform2_double(&rng);  // (0, 1]
        target = 1.0 - r.x;
        u = 1.0 - r.y;
    }
    int64_t lo = 0, hi = m;
    while (lo < hi) {
        const int64_t mid = lo + ((hi - lo) >> 1);
        if (table[mid] / total < target) lo = mid + 1; else hi = mid;
    }
    if (lo >= m) lo = m - 1;
    const double rt = xgrid[lo];
    const double sgn = (double)((rt > 0) - (rt < 0));
    const double shift = (st == 0) ? t : (u * st + (t - st / 2.));
    rts[d * S + s] = rt + sgn * shift;
}

constexpr int kStatThreads = 512;

// Fixed-tree block sum (deterministic): every thread gets the total.
__device__ double block_sum(double v, double* scratch)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    __syncthreads();
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    double s = 0.0;
    for (int w = 0; w < kStatThreads / 32; ++w) s += scratch[w];
    return s;
}

// scipy.stats.scoreatpercentile(sorted, per) with the default 'fraction' interpolation, as called by
// hddm/utils.py:258,265: idx = per/100.*(L-1); whole idx -> that element, else the weighted pair.
// `at(k)` returns the k-th smallest element.
template <class At>
__device__ double score_at_percentile(At at, int64_t L, double per)
{
    if (L <= 0) return NAN;
    const double idx = __dmul_rn(per / 100., (double)(L - 1));
    const int64_t i = (int64_t)idx;
    if ((double)i == idx) return at(i);
    const double wl = (double)(i + 1) - idx, wr = idx - (double)i;
    return __dadd_rn(__dmul_rn(at(i), wl), __dmul_rn(at(i + 1), wr)) / __dadd_rn(wl, wr);
}

// One CTA per draw: bitonic sort of the draw's signed RTs in shared memory (zeros and NaNs keyed 0 sit
// between the two boundaries), then the statistics of hddm/utils.py:241-267.
__global__ void __launch_bounds__(kStatThreads) k_ppc_stats(const double* __restrict__ rts, int64_t S,
                                                            int npad, double* __restrict__ stats)
{
    extern __shared__ double key[];
    __shared__ double scratch[kStatThreads / 32];
    __shared__ int count[2];
    const double* x = rts + (int64_t)blockIdx.x * S;
    if (threadIdx.x < 2) count[threadIdx.x] = 0;
    __syncthreads();
    int n_neg = 0, n_pos = 0;
    for (int i = threadIdx.x; i < npad; i += kStatThreads) {
        double v = INFINITY;  // padding sorts to the end
        if (i < S) {
            v = x[i];
            if (v < 0) ++n_neg;
            else if (v > 0) ++n_pos;
            else v = 0.0;  // zero or NaN: belongs to neither boundary
        }
        key[i] = v;
    }
    atomicAdd(&count[0], n_neg);
    atomicAdd(&count[1], n_pos);
    __syncthreads();
    for (int k = 2; k <= npad; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < npad; i += kStatThreads) {
                const int l = i ^ j;
                if (l > i) {
                    const double a = key[i], b = key[l];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) {
                        key[i] = b;
                        key[l] = a;
                    }
                }
            }
            __syncthreads();
        }
    const int64_t L_lb = count[0], L_ub = count[1];
    const int64_t ub0 = S - L_ub;  // +inf padding lies beyond S; real +inf RTs do not occur
    // upper boundary: key[ub0 + k] ascending; lower boundary |x| ascending: -key[L_lb - 1 - k]
    double sum_ub = 0.0, sum_lb = 0.0;
    for (int64_t i = threadIdx.x; i < L_ub; i += kStatThreads) sum_ub += key[ub0 + i];
    for (int64_t i = threadIdx.x; i < L_lb; i += kStatThreads) sum_lb += key[i];
    const double mean_ub = L_ub ? block_sum(sum_ub, scratch) / (double)L_ub : NAN;
    const double mean_lb = L_lb ? block_sum(sum_lb, scratch) / (double)L_lb : NAN;  // of x[x<0]: negative
    double sq_ub = 0.0, sq_lb = 0.0;
    for (int64_t i = threadIdx.x; i < L_ub; i += kStatThreads) {
        const double dlt = key[ub0 + i] - mean_ub;
        sq_ub += dlt * dlt;
    }
    for (int64_t i = threadIdx.x; i < L_lb; i += kStatThreads) {
        const double dlt = key[i] - mean_lb;
        sq_lb += dlt * dlt;
    }
    const double var_ub = L_ub ? block_sum(sq_ub, scratch) / (double)L_ub : NAN;
    const double var_lb = L_lb ? block_sum(sq_lb, scratch) / (double)L_lb : NAN;
    if (threadIdx.x == 0) {
        double* out = stats + (int64_t)blockIdx.x * kPpcStats;
        const double q[5] = {10, 30, 50, 70, 90};
        out[0] = (double)L_ub / (double)S;  // np.mean(x > 0)
        out[1] = mean_ub;
        out[2] = sqrt(var_ub);              // np.std: population (ddof = 0)
        auto at_ub = [&](int64_t k) { return key[ub0 + k]; };
        auto at_lb = [&](int64_t k) { return -key[L_lb - 1 - k]; };
        for (int k = 0; k < 5; ++k) out[3 + k] = score_at_percentile(at_ub, L_ub, q[k]);
        out[8] = mean_lb;
        out[9] = sqrt(var_lb);
        for (int k = 0; k < 5; ++k) out[10 + k] = score_at_percentile(at_lb, L_lb, q[k]);
    }
}

}  // namespace

cudaError_t launch_cumsum_batch(double* tables, int64_t m, int64_t D, double* last, cudaStream_t stream)
{
    if (m <= 0 || D <= 0) return cudaSuccess;
    static bool opted_in = false;  // (one device per process: hddm_b200_init)
    if (!opted_in) {
        cudaError_t e = cudaFuncSetAttribute(k_cumsum_batch, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)kScanTileBytes);
        if (e != cudaSuccess) return e;
        opted_in = true;
    }
    k_cumsum_batch<<<(unsigned)D, kScanThreads, kScanTileBytes, stream>>>(tables, m, last);
    return cudaGetLastError();
}

cudaError_t launch_sample_rts_batch(const double* xgrid, const double* tables, int64_t m, const double* f,
                                    const double* u_delay, int64_t S, int64_t D, const double* t,
                                    const double* st, unsigned long long seed, int64_t draw0,
                                    const double* last, double* rts, cudaStream_t stream)
{
    if (S <= 0 || D <= 0) return cudaSuccess;
    if (D > 65535) return cudaErrorInvalidValue;
    dim3 grid((unsigned)((S + 255) / 256), (unsigned)D);
    k_sample_batch<<<grid, 256, 0, stream>>>(xgrid, tables, m, f, u_delay, S, t, st, seed, draw0, last, rts);
    return cudaGetLastError();
}

cudaError_t launch_ppc_stats(const double* rts, int64_t S, int64_t D, double* stats, cudaStream_t stream)
{
    if (D <= 0) return cudaSuccess;
    if (S <= 0 || S > kPpcMaxSamples) return cudaErrorInvalidValue;
    int npad = 2;
    while (npad < S) npad <<= 1;
    const size_t smem = (size_t)npad * sizeof(double);
    static bool opted_in = false;
    if (!opted_in) {
        cudaError_t e = cudaFuncSetAttribute(k_ppc_stats, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)(kPpcMaxSamples * sizeof(double)));
        if (e != cudaSuccess) return e;
        opted_in = true;
    }
    k_ppc_stats<<<(unsigned)D, kStatThreads, smem, stream>>>(rts, S, npad, stats);
    return cudaGetLastError();
}

}  // namespace hddm
