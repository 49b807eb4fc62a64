// gen_kernels.cu -- table post-processing and inverse-CDF sampling for the posterior-predictive
// path: gen_cdf_using_pdf (src/wfpt.pyx:389-414) and gen_rts_from_cdf (src/wfpt.pyx:323-354).
// The density tables themselves are produced by k_pdf_array (wfpt_kernels.cu); here one CTA
// turns a table of m densities into a normalised cumulative table with a fixed-order blocked
// scan (thread j owns the contiguous slice j, slice totals are combined in ascending order), so
// the result is deterministic; rounding differs from numpy's sequential cumsum only in the last
// bits (well inside the 1e-10 parity tolerance).
#include "kernels.cuh"

namespace hddm {

namespace {

constexpr int kScanThreads = 1024;

// kind 0: cumtrapz with unit spacing, out[0] keeps y[0] (src/wfpt.pyx:405-409)
// kind 1: rectangle rule from index 1 with out[0] = 0 (src/wfpt.pyx:333-336)
template <int KIND>
__global__ void __launch_bounds__(kScanThreads) k_scan_normalise(double* table, int64_t m)
{
    __shared__ double slice_total[kScanThreads];
    __shared__ double slice_offset[kScanThreads];
    const int j = threadIdx.x;
    const int64_t per = (m - 1 + kScanThreads - 1) / kScanThreads;  // elements 1..m-1 are scanned
    const int64_t lo = 1 + (int64_t)j * per;
    const int64_t hi = lo + per < m ? lo + per : m;
    const double before = (lo < m) ? table[lo - 1] : 0.0;  // original y[lo-1], read before any write
    double run = 0.0, prev = before;
    for (int64_t i = lo; i < hi; ++i) {
        const double cur = table[i];
        run += KIND == 0 ? (1.0 * (cur + prev)) / 2.0 : cur;
        prev = cur;
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
    prev = before;
    for (int64_t i = lo; i < hi; ++i) {
        const double cur = table[i];
        run += KIND == 0 ? (1.0 * (cur + prev)) / 2.0 : cur;
        prev = cur;
        table[i] = run;
    }
    if (KIND == 1 && j == 0) table[0] = 0.0;
    __syncthreads();
    const double last = table[m - 1];
    __syncthreads();
    for (int64_t i = j; i < m; i += kScanThreads) table[i] = table[i] / last;
}

// rt = x[searchsorted(table, f, 'left')] + sign(rt) * (t | u*st + (t - st/2))   (src/wfpt.pyx:344-353)
__global__ void __launch_bounds__(256) k_sample_rts(const double* __restrict__ xgrid,
                                                    const double* __restrict__ table, int64_t m,
                                                    const double* __restrict__ f,
                                                    const double* __restrict__ u_delay, int64_t samples,
                                                    double t, double st, double* __restrict__ rts)
{
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= samples) return;
    const double target = f[s];
    int64_t lo = 0, hi = m;
    while (lo < hi) {
        const int64_t mid = lo + ((hi - lo) >> 1);
        if (table[mid] < target) lo = mid + 1; else hi = mid;
    }
    if (lo >= m) lo = m - 1;  // f > table[-1] cannot happen for f in [0,1); stay in bounds anyway
    const double rt = xgrid[lo];
    const double sgn = (double)((rt > 0) - (rt < 0));
    const double shift = (st == 0) ? t : (u_delay[s] * st + (t - st / 2.));
    rts[s] = rt + sgn * shift;
}

}  // namespace

cudaError_t launch_cumtrapz_normalise(double* table, int64_t m, cudaStream_t stream)
{
    if (m <= 0) return cudaSuccess;
    k_scan_normalise<0><<<1, kScanThreads, 0, stream>>>(table, m);
    return cudaGetLastError();
}

cudaError_t launch_cumsum_normalise(double* table, int64_t m, cudaStream_t stream)
{
    if (m <= 0) return cudaSuccess;
    k_scan_normalise<1><<<1, kScanThreads, 0, stream>>>(table, m);
    return cudaGetLastError();
}

cudaError_t launch_sample_rts(const double* xgrid, const double* table, int64_t m, const double* f,
                              const double* u_delay, int64_t samples, double t, double st, double* rts,
                              cudaStream_t stream)
{
    if (samples <= 0) return cudaSuccess;
    k_sample_rts<<<(unsigned)((samples + 255) / 256), 256, 0, stream>>>(xgrid, table, m, f, u_delay,
                                                                      samples, t, st, rts);
    return cudaGetLastError();
}

}  // namespace hddm
