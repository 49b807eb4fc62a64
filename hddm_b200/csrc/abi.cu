// abi.cu -- the extern "C" boundary declared in include/hddm_b200.h.
//
// Host side of the drop-in: argument checks that the reference performs in Python/Cython
// (ValueError / AssertionError conditions become status codes), staging of host buffers into a
// grow-only device arena, kernel launches on one library stream, and the blocking copy-back the
// reference's synchronous API implies.  No CPU evaluation path exists here: if CUDA is not
// usable every entry point returns HDDM_ERR_CUDA.
#include <new>
#include <vector>

#include <cstdlib>

#include "abi_runtime.cuh"

using namespace hddm;

namespace hddm {
namespace rt {
thread_local char g_err[512] = "";
Runtime g_rt;
}  // namespace rt
}  // namespace hddm
using namespace hddm::rt;

namespace {

// Host buffers to stage, and how each of the 9 parameters is addressed inside them.
struct Upload {
    const void* host;
    size_t bytes;
};
struct ParSpec {
    int upload;      // index into the uploads
    int64_t offset;  // in doubles
    int64_t set_stride, trial_stride;
};

constexpr int64_t kZeroCopyMaxSets = 2048;  // above this the inputs are worth a real H2D copy

// Core of every "sum of log densities" entry point.  x is either a host buffer (h_x) staged
// into the arena or already device-resident (d_x).
int like_impl(const double* h_x, const double* d_x, int64_t n, const int64_t* h_seg, int64_t n_seg,
              const int64_t* h_node, const int32_t* h_cont, const Upload* uploads, int n_uploads, const ParSpec spec[kNumPar],
              int64_t P, int mode, const Numerics& nm, double* logp, unsigned long long* evals)
{
    if (n < 0 || P < 0 || (n > 0 && !h_x && !d_x) || (P > 0 && !logp))
        return fail(HDDM_ERR_ARG, "likelihood: bad buffer or size");
    TRY(check_numerics(nm.n_st, nm.n_sz, nm.use_adaptive));
    TRY(ensure_ready());
    if (P == 0) return HDDM_OK;

    int64_t max_len = n;
    if (h_seg) {
        max_len = 0;
        for (int64_t p = 0; p < n_seg; ++p) {
            const int64_t len = h_seg[p + 1] - h_seg[p];
            if (len < 0 || h_seg[p] < 0 || h_seg[p + 1] > n)
                return fail(HDDM_ERR_ARG, "batch: segment offsets must be ascending within [0, n]");
            if (len > max_len) max_len = len;
        }
        if (h_node)
            for (int64_t p = 0; p < P; ++p)
                if (h_node[p] < 0 || h_node[p] >= n_seg)
                    return fail(HDDM_ERR_ARG, "batch: set_node entries must index a segment");
    } else if (h_node) {
        return fail(HDDM_ERR_ARG, "batch: set_node needs segment offsets");
    }
    const int64_t chunks = chunks_for(max_len);
    if (chunks > 65535) return fail(HDDM_ERR_ARG, "more than 65535*2048 trials per parameter set");
    if (P > 2147483647LL) return fail(HDDM_ERR_ARG, "more than 2^31-1 parameter sets per launch");

    // Shared-parameter sums are order-independent: lay the trials out by |rt| (sort_kernels.cu).
    bool sortable = g_sort_trials && h_x && mode == kModeWienerLike && !h_cont && n >= kSortMin;
    for (int k = 0; k < 7; ++k) sortable = sortable && spec[k].trial_stride == 0;
    const size_t sort_bytes = sortable ? sort_trials_scratch_bytes(n, h_seg ? n_seg : 0) : 0;

    Layout L;
    const size_t off_x = h_x ? L.reserve((size_t)n * sizeof(double)) : 0;
    const size_t off_xs = sortable ? L.reserve((size_t)n * sizeof(double)) : 0;
    const size_t off_sort = sortable ? L.reserve(sort_bytes) : 0;
    const size_t off_cont = h_cont ? L.reserve((size_t)n * sizeof(int32_t)) : 0;
    // small inputs (segment table, node map, parameters) are contiguous: staged through page-locked
    // memory and uploaded with ONE copy; likewise logp + the eval counter come back with one copy
    const size_t small_lo = L.total;
    const size_t off_seg = h_seg ? L.reserve((size_t)(n_seg + 1) * sizeof(int64_t)) : 0;
    const size_t off_node = h_node ? L.reserve((size_t)P * sizeof(int64_t)) : 0;
    size_t off_up[8];
    for (int u = 0; u < n_uploads; ++u) off_up[u] = L.reserve(uploads[u].bytes);
    const size_t small_hi = L.total;
    const size_t off_partial = L.reserve((size_t)P * chunks * sizeof(double));
    const size_t off_logp = L.reserve((size_t)P * sizeof(double));
    const size_t off_evals = L.reserve(sizeof(unsigned long long));
    const size_t out_bytes = L.total - off_logp;
    TRY(ensure_arena(L.total));
    const bool staged = (small_hi - small_lo) <= kPinnedMax && out_bytes <= kPinnedMax;
    // Latency path (an MCMC step on 200-trial nodes is launch-bound): with few sets the kernels read the small
    // inputs from, and write the log-likelihoods to, page-locked host memory directly (UVA: cudaMallocHost
    // memory is device-accessible under the same address) -- no H2D / D2H copies, only launch + sync.
    const size_t in_bytes = small_hi - small_lo;
    const bool zero_copy = staged && P <= kZeroCopyMaxSets;
    if (staged) TRY(ensure_pinned(zero_copy ? align_up(in_bytes) + out_bytes : (in_bytes > out_bytes ? in_bytes : out_bytes)));

    cudaStream_t s = g_rt.stream;
    if (h_x && n > 0)
        CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_x), h_x, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
    if (h_cont && n > 0)
        CUDA_TRY(cudaMemcpyAsync(L.at<int32_t>(off_cont), h_cont, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    if (staged) {
        char* ph = g_rt.pinned - small_lo;
        if (h_seg) memcpy(ph + off_seg, h_seg, (size_t)(n_seg + 1) * sizeof(int64_t));
        if (h_node) memcpy(ph + off_node, h_node, (size_t)P * sizeof(int64_t));
        for (int u = 0; u < n_uploads; ++u)
            if (uploads[u].bytes) memcpy(ph + off_up[u], uploads[u].host, uploads[u].bytes);
        if (!zero_copy)
            CUDA_TRY(cudaMemcpyAsync(L.at<char>(small_lo), g_rt.pinned, in_bytes, cudaMemcpyHostToDevice, s));
    } else {
        if (h_seg)
            CUDA_TRY(cudaMemcpyAsync(L.at<int64_t>(off_seg), h_seg, (size_t)(n_seg + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
        if (h_node)
            CUDA_TRY(cudaMemcpyAsync(L.at<int64_t>(off_node), h_node, (size_t)P * sizeof(int64_t), cudaMemcpyHostToDevice, s));
        for (int u = 0; u < n_uploads; ++u)
            if (uploads[u].bytes)
                CUDA_TRY(cudaMemcpyAsync(L.at<char>(off_up[u]), uploads[u].host, uploads[u].bytes, cudaMemcpyHostToDevice, s));
    }
    if (evals) CUDA_TRY(cudaMemsetAsync(L.at<char>(off_evals), 0, sizeof(unsigned long long), s));

    // where the kernels find the small inputs: the arena copy, or the page-locked staging itself
    char* small_base = zero_copy ? g_rt.pinned - small_lo : g_rt.arena;
    auto small_at = [&](size_t off) { return small_base + off; };
    LikeArgs a;
    memset(&a, 0, sizeof(a));
    a.x = h_x ? L.at<double>(off_x) : d_x;
    a.seg = h_seg ? reinterpret_cast<const int64_t*>(small_at(off_seg)) : nullptr;
    a.set_node = h_node ? reinterpret_cast<const int64_t*>(small_at(off_node)) : nullptr;
    if (sortable) {
        // (the sort reads the segment table from device memory: sortable launches carry host trials and are
        // not the latency path; give it the arena copy)
        if (zero_copy && h_seg)
            CUDA_TRY(cudaMemcpyAsync(L.at<int64_t>(off_seg), h_seg, (size_t)(n_seg + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
        CUDA_TRY(sort_trials(L.at<double>(off_x), L.at<double>(off_xs), n, h_seg ? L.at<int64_t>(off_seg) : nullptr,
                             h_seg ? n_seg : 0, L.at<char>(off_sort), sort_bytes, s));
        a.x = L.at<double>(off_xs);
    }
    a.cont = h_cont ? L.at<int32_t>(off_cont) : nullptr;
    for (int k = 0; k < kNumPar; ++k) {
        a.par[k] = reinterpret_cast<const double*>(small_at(off_up[spec[k].upload])) + spec[k].offset;
        a.set_stride[k] = spec[k].set_stride;
        a.trial_stride[k] = spec[k].trial_stride;
    }
    a.n = n;
    a.n_sets = P;
    a.max_seg_len = max_len;
    a.chunks = (int)chunks;
    a.mode = mode;
    a.nm = nm;
    {   // simple-DDM sweeps (no sz, no st in any set) have a kernel of their own: the host can see it
        bool simple = (mode == kModeWienerLike || mode == kModeContaminant) && spec[4].trial_stride == 0 && spec[6].trial_stride == 0;
        for (int64_t p = 0; p < P && simple; ++p)
            for (int k = 4; k <= 6; k += 2) {
                const double val = static_cast<const double*>(uploads[spec[k].upload].host)[spec[k].offset + p * spec[k].set_stride];
                simple = simple && (val < 1e-3);
            }
        a.all_simple = simple ? 1 : 0;
    }
    a.partial = L.at<double>(off_partial);
    double* host_out = reinterpret_cast<double*>(g_rt.pinned + align_up(in_bytes));  // zero-copy result block
    a.logp = zero_copy ? host_out : L.at<double>(off_logp);
    a.evals = evals ? L.at<unsigned long long>(off_evals) : nullptr;
    CUDA_TRY(launch_like(a, s));
    if (zero_copy) {
        if (evals) CUDA_TRY(cudaMemcpyAsync(evals, a.evals, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
        CUDA_TRY(cudaStreamSynchronize(s));
        memcpy(logp, host_out, (size_t)P * sizeof(double));
        return HDDM_OK;
    }
    if (staged) {  // the upload has been consumed by the kernels queued before this copy: reuse the staging
        CUDA_TRY(cudaMemcpyAsync(g_rt.pinned, L.at<char>(off_logp), out_bytes, cudaMemcpyDeviceToHost, s));
        CUDA_TRY(cudaStreamSynchronize(s));
        memcpy(logp, g_rt.pinned, (size_t)P * sizeof(double));
        if (evals) memcpy(evals, g_rt.pinned + (off_evals - off_logp), sizeof(unsigned long long));
        return HDDM_OK;
    }
    CUDA_TRY(cudaMemcpyAsync(logp, a.logp, (size_t)P * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (evals) CUDA_TRY(cudaMemcpyAsync(evals, a.evals, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

int scalar_like(const double* h_x, const double* d_x, int64_t n, const double par[kNumPar], int mode,
                const int32_t* h_cont, const Numerics& nm, double* logp)
{
    const Upload up[1] = {{par, kNumPar * sizeof(double)}};
    ParSpec spec[kNumPar];
    for (int k = 0; k < kNumPar; ++k) spec[k] = {0, k, 0, 0};
    return like_impl(h_x, d_x, n, nullptr, 0, nullptr, h_cont, up, 1, spec, 1, mode, nm, logp, nullptr);
}

int batch_like(const double* h_x, const double* d_x, int64_t n, const int64_t* seg, int64_t n_seg,
               const int64_t* set_node, const double* params, int64_t P, const Numerics& nm, double* logp,
               unsigned long long* evals)
{
    if (P > 0 && !params) return fail(HDDM_ERR_ARG, "batch: null params");
    const Upload up[1] = {{params, (size_t)P * kNumPar * sizeof(double)}};
    ParSpec spec[kNumPar];
    for (int k = 0; k < kNumPar; ++k) spec[k] = {0, k * P, 1, 0};
    return like_impl(h_x, d_x, n, seg, n_seg, set_node, nullptr, up, 1, spec, P, kModeWienerLike, nm, logp, evals);
}

int pdf_array_impl(const double* x, int64_t n, const double par[kNumPar], const Numerics& nm, int logp,
                   double* y)
{
    if (n < 0 || (n > 0 && (!x || !y))) return fail(HDDM_ERR_ARG, "pdf_array: bad buffer");
    TRY(check_numerics(nm.n_st, nm.n_sz, nm.use_adaptive));
    TRY(ensure_ready());
    if (n == 0) return HDDM_OK;
    Layout L;
    const size_t off_x = L.reserve((size_t)n * sizeof(double));
    const size_t off_y = L.reserve((size_t)n * sizeof(double));
    TRY(ensure_arena(L.total));
    cudaStream_t s = g_rt.stream;
    CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_x), x, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
    PdfArgs a;
    a.x = L.at<double>(off_x);
    a.y = L.at<double>(off_y);
    a.n = n;
    a.logp = logp;
    a.evals = nullptr;
    memcpy(a.par, par, sizeof(a.par));
    a.nm = nm;
    CUDA_TRY(launch_pdf_array(a, s));
    CUDA_TRY(cudaMemcpyAsync(y, a.y, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

}  // namespace

struct hddm_trials {
    double* d_x;
    int64_t n;
    double* d_sorted;              // trials ordered by |rt| within the segments of `sorted_seg`
    std::vector<int64_t> sorted_seg;  // empty: one segment covering everything
    bool has_sorted;
};

extern "C" {

int hddm_b200_version(void) { return 100; }

const char* hddm_b200_last_error(void) { return g_err; }

int hddm_b200_init(int device)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (g_rt.ready && device >= 0 && device != g_rt.device) {
        snprintf(g_err, sizeof(g_err), "hddm_b200 is already bound to CUDA device %d; one process per GPU "
                 "(asked for device %d)", g_rt.device, device);
        return HDDM_ERR_ARG;
    }
    if (device >= 0) CUDA_TRY(cudaSetDevice(device));
    return ensure_ready();
}

int hddm_b200_device_count(int* count)
{
    if (!count) return fail(HDDM_ERR_ARG, "null count");
    *count = 0;
    CUDA_TRY(cudaGetDeviceCount(count));
    return HDDM_OK;
}

int hddm_b200_synchronize(void)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    TRY(ensure_ready());
    CUDA_TRY(cudaStreamSynchronize(g_rt.stream));
    return HDDM_OK;
}

int hddm_b200_fp64_peak(int iters, double* flops, float* ms)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    TRY(ensure_ready());
    if (!flops) return fail(HDDM_ERR_ARG, "null flops");
    *flops = fp64_peak_probe(iters > 0 ? iters : 4096, ms, g_rt.stream);
    CUDA_TRY(cudaGetLastError());
    return *flops > 0 ? HDDM_OK : fail(HDDM_ERR_CUDA, "fp64 probe failed");
}

// The two knobs below change perf-critical behaviour for every caller in the process: they are for A/B
// parity tests and profiling only and are refused unless HDDM_B200_DEBUG_KNOBS=1 is in the environment.
static bool debug_knobs_enabled()
{
    const char* v = getenv("HDDM_B200_DEBUG_KNOBS");
    return v && v[0] == '1';
}

int hddm_b200_set_kernel(int which)
{
    if (!debug_knobs_enabled())
        return fail(HDDM_ERR_ARG, "set_kernel is a debug knob: set HDDM_B200_DEBUG_KNOBS=1 to enable it");
    if (which < 0 || which > 4)
        return fail(HDDM_ERR_ARG, "set_kernel: 0 (auto), 1 (generic), 2 (no grid), 3 (grid, one pass) or "
                                  "4 (grid, two passes in one launch)");
    std::lock_guard<std::mutex> lock(g_rt.mu);
    g_kernel_choice = which;
    return HDDM_OK;
}

int hddm_b200_set_sort(int on)
{
    if (!debug_knobs_enabled())
        return fail(HDDM_ERR_ARG, "set_sort is a debug knob: set HDDM_B200_DEBUG_KNOBS=1 to enable it");
    std::lock_guard<std::mutex> lock(g_rt.mu);
    g_sort_trials = on ? 1 : 0;
    return HDDM_OK;
}

int hddm_series_table(double err, double* thr, int* code, int* n_thr)
{
    if (!thr || !code || !n_thr) return fail(HDDM_ERR_ARG, "series_table: null output");
    const SeriesTable& t = series_table_for(err);
    *n_thr = t.n_thr;
    for (int i = 0; i < t.n_thr; ++i) thr[i] = t.thr[i];
    for (int i = 0; i <= t.n_thr; ++i) code[i] = t.code[i];
    return HDDM_OK;
}

/* ---- wfpt module --------------------------------------------------------------------------- */

int hddm_pdf_array(const double* x, int64_t n, double v, double sv, double a, double z, double sz,
                   double t, double st, double err, int logp, int n_st, int n_sz, int use_adaptive,
                   double simps_err, double p_outlier, double w_outlier, double* y)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    const double par[kNumPar] = {v, sv, a, z, sz, t, st, p_outlier, w_outlier};
    return pdf_array_impl(x, n, par, make_numerics(err, n_st, n_sz, use_adaptive, simps_err), logp, y);
}

int hddm_full_pdf(double x, double v, double sv, double a, double z, double sz, double t, double st,
                  double err, int n_st, int n_sz, int use_adaptive, double simps_err, double* out)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!out) return fail(HDDM_ERR_ARG, "full_pdf: null out");
    const double par[kNumPar] = {v, sv, a, z, sz, t, st, 0.0, 0.0};
    return pdf_array_impl(&x, 1, par, make_numerics(err, n_st, n_sz, use_adaptive, simps_err), 0, out);
}

int hddm_wiener_like(const double* x, int64_t n, double v, double sv, double a, double z, double sz,
                     double t, double st, double err, int n_st, int n_sz, int use_adaptive,
                     double simps_err, double p_outlier, double w_outlier, double* logp)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    const double par[kNumPar] = {v, sv, a, z, sz, t, st, p_outlier, w_outlier};
    return scalar_like(x, nullptr, n, par, kModeWienerLike, nullptr,
                       make_numerics(err, n_st, n_sz, use_adaptive, simps_err), logp);
}

int hddm_wiener_like_multi(const double* x, int64_t n, const double* const par[7], const int64_t stride[7],
                           double err, int n_st, int n_sz, int use_adaptive, double simps_err,
                           double p_outlier, double w_outlier, double* logp)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!par || !stride) return fail(HDDM_ERR_ARG, "wiener_like_multi: null parameter table");
    double scalars[kNumPar] = {0, 0, 0, 0, 0, 0, 0, p_outlier, w_outlier};
    Upload up[8];
    ParSpec spec[kNumPar];
    int n_up = 1;
    for (int k = 0; k < 7; ++k) {
        if (!par[k]) return fail(HDDM_ERR_ARG, "wiener_like_multi: null parameter pointer");
        if (stride[k] == 0) {
            scalars[k] = par[k][0];
            spec[k] = {0, k, 0, 0};
        } else if (stride[k] == 1) {
            up[n_up] = {par[k], (size_t)n * sizeof(double)};
            spec[k] = {n_up, 0, 0, 1};
            ++n_up;
        } else {
            return fail(HDDM_ERR_ARG, "wiener_like_multi: stride must be 0 (scalar) or 1 (per trial)");
        }
    }
    spec[7] = {0, 7, 0, 0};
    spec[8] = {0, 8, 0, 0};
    up[0] = {scalars, sizeof(scalars)};
    return like_impl(x, nullptr, n, nullptr, 0, nullptr, nullptr, up, n_up, spec, 1, kModeMulti,
                     make_numerics(err, n_st, n_sz, use_adaptive, simps_err), logp, nullptr);
}

int hddm_wiener_like_contaminant(const double* x, const int32_t* cont_x, int64_t n, double v, double sv,
                                 double a, double z, double sz, double t, double st, double t_min,
                                 double t_max, double err, int n_st, int n_sz, int use_adaptive,
                                 double simps_err, double* logp)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (n > 0 && !cont_x) return fail(HDDM_ERR_ARG, "wiener_like_contaminant: null cont_x");
    int n_cont = 0;  // `cdef int n_cont = np.sum(cont_x)`, src/wfpt.pyx:370
    for (int64_t i = 0; i < n; ++i) n_cont += cont_x[i];
    const double par[kNumPar] = {v, sv, a, z, sz, t, st, 0.0, 0.0};
    TRY(scalar_like(x, nullptr, n, par, kModeContaminant, n > 0 ? cont_x : nullptr,
                    make_numerics(err, n_st, n_sz, use_adaptive, simps_err), logp));
    // src/wfpt.pyx:384 (never reached when a zero density returned -inf earlier, :377-379)
    if (!(std::isinf(*logp) && *logp < 0)) *logp += n_cont * log(0.5 * 1. / (t_max - t_min));
    return HDDM_OK;
}

int hddm_gen_cdf_using_pdf(const double* xgrid, int64_t m, double v, double sv, double a, double z,
                           double sz, double t, double st, double err, int n_st, int n_sz, int use_adaptive,
                           double simps_err, double p_outlier, double w_outlier, double* cdf)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (out_of_support(sv, a, z, sz, t, st, p_outlier))
        return fail(HDDM_ERR_SUPPORT, "at least one of the parameters is out of the support");
    if (m <= 0 || !xgrid || !cdf) return fail(HDDM_ERR_ARG, "gen_cdf_using_pdf: bad grid");
    const Numerics nm = make_numerics(err, n_st, n_sz, use_adaptive, simps_err);
    TRY(check_numerics(n_st, n_sz, use_adaptive));
    TRY(ensure_ready());
    Layout L;
    const size_t off_x = L.reserve((size_t)m * sizeof(double));
    const size_t off_y = L.reserve((size_t)m * sizeof(double));
    TRY(ensure_arena(L.total));
    cudaStream_t s = g_rt.stream;
    CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_x), xgrid, (size_t)m * sizeof(double), cudaMemcpyHostToDevice, s));
    PdfArgs pa;
    pa.x = L.at<double>(off_x);
    pa.y = L.at<double>(off_y);
    pa.n = m;
    pa.logp = 0;
    pa.evals = nullptr;
    const double par[kNumPar] = {v, sv, a, z, sz, t, st, p_outlier, w_outlier};
    memcpy(pa.par, par, sizeof(par));
    pa.nm = nm;
    CUDA_TRY(launch_pdf_array(pa, s));
    CUDA_TRY(launch_cumtrapz_normalise(pa.y, m, s));
    CUDA_TRY(cudaMemcpyAsync(cdf, pa.y, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

int hddm_gen_rts_from_cdf(const double* xgrid, int64_t m, double v, double sv, double a, double z, double sz,
                          double t, double st, const double* f, const double* u_delay, int64_t samples,
                          double* rts, double* table)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (m <= 0 || samples < 0 || !xgrid || (samples > 0 && (!f || !rts)) || (st != 0 && samples > 0 && !u_delay))
        return fail(HDDM_ERR_ARG, "gen_rts_from_cdf: bad buffer");
    TRY(ensure_ready());
    Layout L;
    const size_t off_x = L.reserve((size_t)m * sizeof(double));
    const size_t off_y = L.reserve((size_t)m * sizeof(double));
    const size_t off_f = L.reserve((size_t)samples * sizeof(double));
    const size_t off_u = L.reserve((size_t)samples * sizeof(double));
    const size_t off_r = L.reserve((size_t)samples * sizeof(double));
    TRY(ensure_arena(L.total));
    cudaStream_t s = g_rt.stream;
    CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_x), xgrid, (size_t)m * sizeof(double), cudaMemcpyHostToDevice, s));
    if (samples > 0) {
        CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_f), f, (size_t)samples * sizeof(double), cudaMemcpyHostToDevice, s));
        if (st != 0)
            CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_u), u_delay, (size_t)samples * sizeof(double), cudaMemcpyHostToDevice, s));
    }
    // table of full_pdf(x[i], v, sv, a, z, sz, 0, 0, 1e-4) with the cpdef defaults n=2, adaptive,
    // simps_err=1e-3 (src/wfpt.pyx:335, src/pdf.pxi:106-107)
    PdfArgs pa;
    pa.x = L.at<double>(off_x);
    pa.y = L.at<double>(off_y);
    pa.n = m;
    pa.logp = 0;
    pa.evals = nullptr;
    const double par[kNumPar] = {v, sv, a, z, sz, 0.0, 0.0, 0.0, 0.0};
    memcpy(pa.par, par, sizeof(par));
    pa.nm = make_numerics(1e-4, 2, 2, 1, 1e-3);
    CUDA_TRY(launch_pdf_array(pa, s));
    CUDA_TRY(launch_cumsum_normalise(pa.y, m, s));
    CUDA_TRY(launch_sample_rts(pa.x, pa.y, m, L.at<double>(off_f), L.at<double>(off_u), samples, t, st,
                               L.at<double>(off_r), s));
    if (samples > 0)
        CUDA_TRY(cudaMemcpyAsync(rts, L.at<double>(off_r), (size_t)samples * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (table) CUDA_TRY(cudaMemcpyAsync(table, pa.y, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

/* ---- cdfdif_wrapper module ------------------------------------------------------------------ */

int hddm_dmat_cdf_array(const double* x, int64_t n, double v, double sv, double a, double z, double sz,
                        double t, double st, double p_outlier, double w_outlier, double* y)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (n < 0 || (n > 0 && (!x || !y))) return fail(HDDM_ERR_ARG, "dmat_cdf_array: bad buffer");
    if (p_outlier > 0 && n > 0) {  // src/cdfdif_wrapper.pyx:20-21
        double peak = 0;
        bool has_nan = false;
        for (int64_t i = 0; i < n; ++i) {
            const double m = fabs(x[i]);
            has_nan |= (m != m);
            if (m > peak) peak = m;
        }
        if (has_nan || !(peak < (1. / (2 * w_outlier))))
            return fail(HDDM_ERR_RT_RANGE, "1. / (2*w_outlier) must be smaller than RT");
    }
    if (out_of_support(sv, a, z, sz, t, st, p_outlier))
        return fail(HDDM_ERR_SUPPORT, "at least one of the parameters is out of the support");
    TRY(ensure_ready());
    if (n == 0) return HDDM_OK;
    Layout L;
    const size_t off_x = L.reserve((size_t)n * sizeof(double));
    const size_t off_y = L.reserve((size_t)n * sizeof(double));
    TRY(ensure_arena(L.total));
    cudaStream_t s = g_rt.stream;
    CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_x), x, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
    CdfArgs ca;
    ca.x = L.at<double>(off_x);
    ca.y = L.at<double>(off_y);
    ca.n = n;
    const double par[kNumPar] = {v, sv, a, z, sz, t, st, p_outlier, w_outlier};
    memcpy(ca.par, par, sizeof(par));
    CUDA_TRY(launch_cdf_array(ca, s));
    CUDA_TRY(cudaMemcpyAsync(y, ca.y, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

/* ---- batched likelihood ---------------------------------------------------------------------- */

int hddm_wiener_like_batch(const double* x, int64_t n, const int64_t* seg, const double* params, int64_t P,
                           double err, int n_st, int n_sz, int use_adaptive, double simps_err, double* logp,
                           unsigned long long* evals)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    return batch_like(x, nullptr, n, seg, P, nullptr, params, P, make_numerics(err, n_st, n_sz, use_adaptive, simps_err),
                      logp, evals);
}

int hddm_wiener_like_multi_batch(const double* x, int64_t n, const double* v_trial, const double* params,
                                 int64_t P, double err, int n_st, int n_sz, int use_adaptive, double simps_err,
                                 double* logp, unsigned long long* evals)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (P > 0 && (!params || (n > 0 && !v_trial))) return fail(HDDM_ERR_ARG, "multi_batch: null params");
    const Upload up[2] = {{params, (size_t)P * kNumPar * sizeof(double)},
                          {v_trial, (size_t)P * (size_t)n * sizeof(double)}};
    ParSpec spec[kNumPar];
    for (int k = 0; k < kNumPar; ++k) spec[k] = {0, k * P, 1, 0};
    spec[0] = {1, 0, n, 1};
    return like_impl(x, nullptr, n, nullptr, 0, nullptr, nullptr, up, 2, spec, P, kModeMulti,
                     make_numerics(err, n_st, n_sz, use_adaptive, simps_err), logp, evals);
}

/* ---- device-resident trials ------------------------------------------------------------------ */

int hddm_trials_upload(const double* x, int64_t n, hddm_trials** out)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!out || n < 0 || (n > 0 && !x)) return fail(HDDM_ERR_ARG, "trials_upload: bad buffer");
    TRY(ensure_ready());
    hddm_trials* h = new (std::nothrow) hddm_trials{nullptr, n, nullptr, {}, false};
    if (!h) return fail(HDDM_ERR_ARG, "trials_upload: out of host memory");
    cudaError_t e = cudaMalloc(&h->d_x, (size_t)(n > 0 ? n : 1) * sizeof(double));
    if (e == cudaSuccess && n > 0) {
        e = cudaMemcpyAsync(h->d_x, x, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, g_rt.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(g_rt.stream);
    }
    if (e != cudaSuccess) {
        if (h->d_x) cudaFree(h->d_x);
        delete h;
        CUDA_TRY(e);
    }
    *out = h;
    return HDDM_OK;
}

int hddm_trials_free(hddm_trials* h)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!h) return HDDM_OK;
    if (g_rt.ready) cudaStreamSynchronize(g_rt.stream);
    cudaFree(h->d_x);
    if (h->d_sorted) cudaFree(h->d_sorted);
    delete h;
    return HDDM_OK;
}

}  // extern "C"

namespace {
// The handle's trials ordered by |rt| within `seg` (P+1 offsets, or nullptr), cached until the
// segmentation changes.  Falls back to the caller's order for short vectors or when disabled.
int sorted_view(hddm_trials* h, const int64_t* seg, int64_t P, const double** out)
{
    *out = h->d_x;
    if (!g_sort_trials || h->n < kSortMin) return HDDM_OK;
    const size_t n_off = seg ? (size_t)P + 1 : 0;
    if (seg)
        for (int64_t p = 0; p < P; ++p)
            if (seg[p] < 0 || seg[p + 1] < seg[p] || seg[p + 1] > h->n)
                return fail(HDDM_ERR_ARG, "batch: segment offsets must be ascending within [0, n]");
    const bool same = h->has_sorted && h->sorted_seg.size() == n_off &&
                      (n_off == 0 || memcmp(h->sorted_seg.data(), seg, n_off * sizeof(int64_t)) == 0);
    if (!same) {
        if (!h->d_sorted) CUDA_TRY(cudaMalloc(&h->d_sorted, (size_t)h->n * sizeof(double)));
        const size_t bytes = sort_trials_scratch_bytes(h->n, seg ? P : 0);
        Layout L;
        const size_t off_seg = L.reserve(n_off * sizeof(int64_t));
        const size_t off_sort = L.reserve(bytes);
        TRY(ensure_arena(L.total));
        cudaStream_t s = g_rt.stream;
        if (seg) CUDA_TRY(cudaMemcpyAsync(L.at<int64_t>(off_seg), seg, n_off * sizeof(int64_t), cudaMemcpyHostToDevice, s));
        CUDA_TRY(sort_trials(h->d_x, h->d_sorted, h->n, seg ? L.at<int64_t>(off_seg) : nullptr, seg ? P : 0,
                             L.at<char>(off_sort), bytes, s));
        CUDA_TRY(cudaStreamSynchronize(s));
        h->sorted_seg.assign(seg, seg + n_off);
        h->has_sorted = true;
    }
    *out = h->d_sorted;
    return HDDM_OK;
}
}  // namespace

extern "C" {

int hddm_trials_wiener_like(const hddm_trials* h, double v, double sv, double a, double z, double sz, double t,
                            double st, double err, int n_st, int n_sz, int use_adaptive, double simps_err,
                            double p_outlier, double w_outlier, double* logp)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!h) return fail(HDDM_ERR_ARG, "null trials handle");
    const double par[kNumPar] = {v, sv, a, z, sz, t, st, p_outlier, w_outlier};
    const double* dx = nullptr;
    TRY(ensure_ready());
    TRY(sorted_view(const_cast<hddm_trials*>(h), nullptr, 0, &dx));
    return scalar_like(nullptr, dx, h->n, par, kModeWienerLike, nullptr,
                       make_numerics(err, n_st, n_sz, use_adaptive, simps_err), logp);
}

int hddm_trials_wiener_like_batch(const hddm_trials* h, const int64_t* seg, const double* params, int64_t P,
                                  double err, int n_st, int n_sz, int use_adaptive, double simps_err,
                                  double* logp, unsigned long long* evals)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!h) return fail(HDDM_ERR_ARG, "null trials handle");
    if (P < 0) return fail(HDDM_ERR_ARG, "batch: negative P");
    const double* dx = nullptr;
    TRY(ensure_ready());
    TRY(sorted_view(const_cast<hddm_trials*>(h), seg, P, &dx));
    return batch_like(nullptr, dx, h->n, seg, P, nullptr, params, P,
                      make_numerics(err, n_st, n_sz, use_adaptive, simps_err), logp, evals);
}

int hddm_wiener_like_nodes(const double* x, int64_t n, const int64_t* node_off, int64_t n_nodes,
                           const int64_t* set_node, const double* params, int64_t P, double err, int n_st,
                           int n_sz, int use_adaptive, double simps_err, double* logp, unsigned long long* evals)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!node_off || n_nodes < 0 || (P > 0 && !set_node)) return fail(HDDM_ERR_ARG, "nodes: null node table");
    return batch_like(x, nullptr, n, node_off, n_nodes, set_node, params, P,
                      make_numerics(err, n_st, n_sz, use_adaptive, simps_err), logp, evals);
}

int hddm_trials_wiener_like_nodes(const hddm_trials* h, const int64_t* node_off, int64_t n_nodes,
                                  const int64_t* set_node, const double* params, int64_t P, double err,
                                  int n_st, int n_sz, int use_adaptive, double simps_err, double* logp,
                                  unsigned long long* evals)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!h) return fail(HDDM_ERR_ARG, "null trials handle");
    if (!node_off || n_nodes < 0 || P < 0 || (P > 0 && !set_node))
        return fail(HDDM_ERR_ARG, "nodes: null node table");
    const double* dx = nullptr;
    TRY(ensure_ready());
    TRY(sorted_view(const_cast<hddm_trials*>(h), node_off, n_nodes, &dx));
    return batch_like(nullptr, dx, h->n, node_off, n_nodes, set_node, params, P,
                      make_numerics(err, n_st, n_sz, use_adaptive, simps_err), logp, evals);
}

/* ---- device-pointer entry points ------------------------------------------------------------- */

int64_t hddm_sort_trials_scratch(int64_t n, int64_t n_seg)
{
    return (int64_t)sort_trials_scratch_bytes(n, n_seg);
}

int hddm_sort_trials_dev(const double* d_x, int64_t n, const int64_t* d_seg, int64_t n_seg, double* d_sorted,
                         void* d_scratch, int64_t scratch_bytes, void* stream)
{
    if (n < 0 || n_seg < 0 || (n > 0 && (!d_x || !d_sorted || !d_scratch)) || (n_seg > 0 && !d_seg) || d_x == d_sorted)
        return fail(HDDM_ERR_ARG, "sort_trials_dev: bad buffer");
    {  // binds / checks the process's one device (kernel attributes and pool settings are per device)
        std::lock_guard<std::mutex> lock(g_rt.mu);
        TRY(ensure_ready());
    }
    CUDA_TRY(sort_trials(d_x, d_sorted, n, d_seg, n_seg, d_scratch, (size_t)scratch_bytes, (cudaStream_t)stream));
    return HDDM_OK;
}

int64_t hddm_wiener_like_batch_scratch(int64_t max_seg_len, int64_t P)
{
    return (P > 0 ? P : 0) * chunks_for(max_seg_len);
}

int hddm_wiener_like_batch_dev(const double* d_x, int64_t n, const int64_t* d_seg, int64_t max_seg_len,
                               const double* d_params, int64_t P, double err, int n_st, int n_sz,
                               int use_adaptive, double simps_err, double* d_logp, double* d_scratch,
                               unsigned long long* d_evals, void* stream)
{
    if (n < 0 || P < 0 || (P > 0 && (!d_params || !d_logp || !d_scratch)) || (n > 0 && !d_x))
        return fail(HDDM_ERR_ARG, "batch_dev: bad buffer or size");
    {  // binds / checks the process's one device (kernel attributes and pool settings are per device)
        std::lock_guard<std::mutex> lock(g_rt.mu);
        TRY(ensure_ready());
    }
    TRY(check_numerics(n_st, n_sz, use_adaptive));
    if (P == 0) return HDDM_OK;
    const int64_t chunks = chunks_for(d_seg ? max_seg_len : n);
    if (chunks > 65535) return fail(HDDM_ERR_ARG, "more than 65535*2048 trials per parameter set");
    LikeArgs a;
    memset(&a, 0, sizeof(a));
    a.x = d_x;
    a.seg = d_seg;
    for (int k = 0; k < kNumPar; ++k) {
        a.par[k] = d_params + k * P;
        a.set_stride[k] = 1;
        a.trial_stride[k] = 0;
    }
    a.n = n;
    a.n_sets = P;
    a.max_seg_len = d_seg ? max_seg_len : n;
    a.chunks = (int)chunks;
    a.mode = kModeWienerLike;
    a.nm = make_numerics(err, n_st, n_sz, use_adaptive, simps_err);
    a.partial = d_scratch;
    a.logp = d_logp;
    a.evals = d_evals;
    CUDA_TRY(launch_like(a, (cudaStream_t)stream));
    return HDDM_OK;
}

// No counterpart in the reference: hddm_wiener_like_batch_dev for callers that may not allocate (include/hddm_b200.h).
// Nothing here or below it calls cudaMalloc*, cudaFree* or creates a stream: the library must already be bound to
// the device (hddm_b200_init at set-up), the partial sums live in the static array of wfpt_kernels.cu.
int hddm_wiener_like_batch_dev_static(const double* d_x, int64_t n, const int64_t* d_seg, int64_t max_seg_len,
                                      const double* d_params, int64_t P, double err, int n_st, int n_sz,
                                      int use_adaptive, double simps_err, double* d_logp,
                                      unsigned long long* d_evals, void* stream)
{
    if (n < 0 || P < 0 || (P > 0 && (!d_params || !d_logp)) || (n > 0 && !d_x) || (d_seg && max_seg_len < 0))
        return fail(HDDM_ERR_ARG, "batch_dev_static: bad buffer or size");
    {
        std::lock_guard<std::mutex> lock(g_rt.mu);
        if (!g_rt.ready)
            return fail(HDDM_ERR_ARG, "batch_dev_static: call hddm_b200_init(device) at set-up first "
                                      "(this entry point binds no device and creates nothing)");
        TRY(ensure_ready());  // bound already: only checks that the bound device is the current one
    }
    TRY(check_numerics(n_st, n_sz, use_adaptive));
    if (P == 0) return HDDM_OK;
    const int64_t chunks = chunks_for(d_seg ? max_seg_len : n);
    if (chunks > 65535) return fail(HDDM_ERR_ARG, "more than 65535*2048 trials per parameter set");
    if (P > kStaticPartialDoubles || P * chunks > kStaticPartialDoubles) {
        snprintf(g_err, sizeof(g_err), "batch_dev_static: %lld sets x %lld blocks of 2048 trials need %lld doubles of "
                 "scratch, the static array holds %lld (2 MiB); use hddm_wiener_like_batch_dev with caller scratch",
                 (long long)P, (long long)chunks, (long long)(P * chunks), (long long)kStaticPartialDoubles);
        return HDDM_ERR_ARG;
    }
    LikeArgs a;
    memset(&a, 0, sizeof(a));
    a.x = d_x;
    a.seg = d_seg;
    for (int k = 0; k < kNumPar; ++k) {
        a.par[k] = d_params + k * P;
        a.set_stride[k] = 1;
        a.trial_stride[k] = 0;
    }
    a.n = n;
    a.n_sets = P;
    a.max_seg_len = d_seg ? max_seg_len : n;
    a.chunks = (int)chunks;
    a.mode = kModeWienerLike;
    a.nm = make_numerics(err, n_st, n_sz, use_adaptive, simps_err);
    a.partial = nullptr;  // launch_like_static points it at its array
    a.logp = d_logp;
    a.evals = d_evals;
    CUDA_TRY(launch_like_static(a, (cudaStream_t)stream));
    return HDDM_OK;
}

int hddm_pdf_array_dev(const double* d_x, int64_t n, double v, double sv, double a, double z, double sz,
                       double t, double st, double err, int logp, int n_st, int n_sz, int use_adaptive,
                       double simps_err, double p_outlier, double w_outlier, double* d_y,
                       unsigned long long* d_evals, void* stream)
{
    if (n < 0 || (n > 0 && (!d_x || !d_y))) return fail(HDDM_ERR_ARG, "pdf_array_dev: bad buffer");
    {  // binds / checks the process's one device (kernel attributes and pool settings are per device)
        std::lock_guard<std::mutex> lock(g_rt.mu);
        TRY(ensure_ready());
    }
    TRY(check_numerics(n_st, n_sz, use_adaptive));
    PdfArgs pa;
    pa.x = d_x;
    pa.y = d_y;
    pa.n = n;
    pa.logp = logp;
    pa.evals = d_evals;
    const double par[kNumPar] = {v, sv, a, z, sz, t, st, p_outlier, w_outlier};
    memcpy(pa.par, par, sizeof(par));
    pa.nm = make_numerics(err, n_st, n_sz, use_adaptive, simps_err);
    CUDA_TRY(launch_pdf_array(pa, (cudaStream_t)stream));
    return HDDM_OK;
}

int hddm_dmat_cdf_array_dev(const double* d_x, int64_t n, double v, double sv, double a, double z, double sz,
                            double t, double st, double p_outlier, double w_outlier, double* d_y, void* stream)
{
    if (n < 0 || (n > 0 && (!d_x || !d_y))) return fail(HDDM_ERR_ARG, "dmat_cdf_array_dev: bad buffer");
    {  // binds / checks the process's one device (kernel attributes and pool settings are per device)
        std::lock_guard<std::mutex> lock(g_rt.mu);
        TRY(ensure_ready());
    }
    if (out_of_support(sv, a, z, sz, t, st, p_outlier))
        return fail(HDDM_ERR_SUPPORT, "at least one of the parameters is out of the support");
    CdfArgs ca;
    ca.x = d_x;
    ca.y = d_y;
    ca.n = n;
    const double par[kNumPar] = {v, sv, a, z, sz, t, st, p_outlier, w_outlier};
    memcpy(ca.par, par, sizeof(par));
    CUDA_TRY(launch_cdf_array(ca, (cudaStream_t)stream));
    return HDDM_OK;
}

}  // extern "C"
