// abi_rl.cu -- extern "C" entry points of the reinforcement-learning likelihoods
// (src/wfpt.pyx:79-155 wiener_like_rlddm, :158-241 wiener_like_rl, :277-320 wiener_like_multi_rlddm).
//
// Host work here is what the reference does in numpy before its loop: np.unique(split_by) and the
// boolean-mask selection of each condition's trials (src/wfpt.pyx:103,112-118) become one stable
// ordering of the trials by condition; everything numerical runs on the GPU (rl_kernels.cu for the
// Q recursion, the ordinary likelihood kernel for the densities).
#include <algorithm>
#include <new>
#include <mutex>
#include <numeric>
#include <vector>

#include "abi_runtime.cuh"

using namespace hddm;
using namespace hddm::rt;

namespace {

// Buffers queued for upload into the arena: reserve() all, commit() once, then dev<T>(slot).
struct Stager {
    struct Item {
        const void* host;
        size_t bytes, off;
    };
    std::vector<Item> items;
    Layout L;
    int add(const void* host, size_t bytes)
    {
        items.push_back({host, bytes, L.reserve(bytes)});
        return (int)items.size() - 1;
    }
    int scratch(size_t bytes) { return add(nullptr, bytes); }
    int commit(cudaStream_t s)
    {
        TRY(ensure_arena(L.total));
        for (const Item& it : items)
            if (it.host && it.bytes)
                CUDA_TRY(cudaMemcpyAsync(g_rt.arena + it.off, it.host, it.bytes, cudaMemcpyHostToDevice, s));
        return HDDM_OK;
    }
    template <class T>
    T* dev(int slot) const { return reinterpret_cast<T*>(g_rt.arena + items[slot].off); }
};

// The task of one subject in run order: conditions ascending (np.unique), trials of a condition in the
// caller's order; or, for the regression flavour, the caller's order cut where split_by changes.
struct RlTask {
    std::vector<int64_t> order;    // run order -> caller index (identity for the regression flavour)
    std::vector<int64_t> run_off;  // n_runs + 1
    std::vector<int32_t> resp, head;
    std::vector<double> fb, x;
};

int build_task(const double* x, const int64_t* response, const double* feedback, const int64_t* split_by,
               int64_t n, bool by_condition, RlTask& t)
{
    if (n < 0 || (n > 0 && (!response || !feedback || !split_by)))
        return fail(HDDM_ERR_ARG, "rl: null response / feedback / split_by");
    for (int64_t i = 0; i < n; ++i)
        if (response[i] != 0 && response[i] != 1)
            return fail(HDDM_ERR_ARG, "rl: response must be 0 (lower) or 1 (upper)");
    t.order.resize((size_t)n);
    std::iota(t.order.begin(), t.order.end(), (int64_t)0);
    if (by_condition)
        std::stable_sort(t.order.begin(), t.order.end(),
                         [&](int64_t l, int64_t r) { return split_by[l] < split_by[r]; });
    t.run_off.assign(1, 0);
    t.resp.resize((size_t)n);
    t.head.assign((size_t)n, 0);
    t.fb.resize((size_t)n);
    if (x) t.x.resize((size_t)n);
    for (int64_t j = 0; j < n; ++j) {
        const int64_t i = t.order[(size_t)j];
        if (j > 0 && split_by[i] != split_by[t.order[(size_t)j - 1]]) t.run_off.push_back(j);
        t.resp[(size_t)j] = (int32_t)response[i];
        t.fb[(size_t)j] = feedback[i];
        if (x) t.x[(size_t)j] = x[i];
    }
    if (n > 0) t.run_off.push_back(n);
    for (size_t r = 0; r + 1 < t.run_off.size(); ++r) t.head[(size_t)t.run_off[r]] = 1;
    return HDDM_OK;
}

constexpr int kRlRows = 12;  // v, sv, a, z, sz, t, st, p_outlier, w_outlier, q, alpha, pos_alpha

int rlddm_impl(const double* x, const int64_t* response, const double* feedback, const int64_t* split_by,
               int64_t n, const double* params, int64_t P, const Numerics& nm, double* logp,
               unsigned long long* evals)
{
    if (P < 0 || (P > 0 && (!params || !logp)) || (n > 0 && !x)) return fail(HDDM_ERR_ARG, "rlddm: bad buffer");
    TRY(check_numerics(nm.n_st, nm.n_sz, nm.use_adaptive));
    TRY(ensure_ready());
    if (P == 0) return HDDM_OK;
    RlTask t;
    TRY(build_task(x, response, feedback, split_by, n, true, t));
    const int64_t chunks = chunks_for(n);
    if (chunks > 65535) return fail(HDDM_ERR_ARG, "more than 65535*2048 trials per parameter set");
    if ((double)P * (double)n * 8.0 > 64e9) return fail(HDDM_ERR_ARG, "rlddm: P x n drift table above 64 GB");
    const int64_t n_runs = (int64_t)t.run_off.size() - 1;
    // The scan walks the task in run order; the densities do not care about order, and their kernel wants the
    // trials boundary-major and by |rt| (a warp then shares its series branch and evaluator class, sort_kernels.cu):
    // the scan scatters each trial's drift to the trial's place in that layout.
    std::vector<int32_t> pos((size_t)n), head_sorted((size_t)n);
    std::vector<double> x_sorted((size_t)n);
    {
        std::vector<int32_t> by_rt((size_t)n);
        std::iota(by_rt.begin(), by_rt.end(), 0);
        auto key = [&](int32_t j) {
            const double v = t.x[(size_t)j];
            uint64_t bits;
            const double mag = fabs(v);
            memcpy(&bits, &mag, sizeof(bits));
            return bits | (v > 0 ? (uint64_t)1 << 63 : (uint64_t)0);
        };
        std::stable_sort(by_rt.begin(), by_rt.end(), [&](int32_t l, int32_t r) { return key(l) < key(r); });
        for (int64_t k = 0; k < n; ++k) {
            const int32_t j = by_rt[(size_t)k];
            pos[(size_t)j] = (int32_t)k;
            x_sorted[(size_t)k] = t.x[(size_t)j];
            head_sorted[(size_t)k] = t.head[(size_t)j];
        }
    }

    Stager S;
    const int s_x = S.add(x_sorted.data(), (size_t)n * sizeof(double));
    const int s_resp = S.add(t.resp.data(), (size_t)n * sizeof(int32_t));
    const int s_head = S.add(head_sorted.data(), (size_t)n * sizeof(int32_t));
    const int s_pos = S.add(pos.data(), (size_t)n * sizeof(int32_t));
    const int s_fb = S.add(t.fb.data(), (size_t)n * sizeof(double));
    const int s_run = S.add(t.run_off.data(), t.run_off.size() * sizeof(int64_t));
    const int s_par = S.add(params, (size_t)P * kRlRows * sizeof(double));
    const int s_drift = S.scratch((size_t)P * (size_t)(n > 0 ? n : 1) * sizeof(double));
    const int s_partial = S.scratch((size_t)P * chunks * sizeof(double));
    const int s_logp = S.scratch((size_t)P * sizeof(double));
    const int s_evals = S.scratch(sizeof(unsigned long long));
    cudaStream_t s = g_rt.stream;
    TRY(S.commit(s));
    CUDA_TRY(cudaMemsetAsync(S.dev<char>(s_evals), 0, sizeof(unsigned long long), s));
    const double* d_par = S.dev<double>(s_par);

    RlScanArgs sc;
    memset(&sc, 0, sizeof(sc));
    sc.response = S.dev<int32_t>(s_resp);
    sc.feedback = S.dev<double>(s_fb);
    sc.run_off = S.dev<int64_t>(s_run);
    sc.n = n;
    sc.n_runs = n_runs;
    sc.n_sets = P;
    sc.v = d_par + 0 * P;
    sc.q = d_par + 9 * P;
    sc.alpha = d_par + 10 * P;
    sc.pos_alpha = d_par + 11 * P;
    sc.q_stride = sc.alpha_set_stride = sc.pos_stride = sc.v_set_stride = 1;
    sc.drift = S.dev<double>(s_drift);
    sc.pos = n > 0 ? S.dev<int32_t>(s_pos) : nullptr;
    CUDA_TRY(launch_rl_scan(sc, s));

    LikeArgs a;
    memset(&a, 0, sizeof(a));
    a.x = S.dev<double>(s_x);
    a.cont = n > 0 ? S.dev<int32_t>(s_head) : nullptr;  // src/wfpt.pyx:120-131: run heads only update Q
    for (int k = 0; k < kNumPar; ++k) {
        a.par[k] = d_par + k * P;
        a.set_stride[k] = 1;
    }
    a.par[0] = sc.drift;
    a.set_stride[0] = n;
    a.trial_stride[0] = 1;
    a.n = n;
    a.n_sets = P;
    a.chunks = (int)chunks;
    a.mode = kModeWienerLike;
    a.nm = nm;
    a.partial = S.dev<double>(s_partial);
    a.logp = S.dev<double>(s_logp);
    a.evals = S.dev<unsigned long long>(s_evals);
    CUDA_TRY(launch_like(a, s));
    CUDA_TRY(cudaMemcpyAsync(logp, a.logp, (size_t)P * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (evals) CUDA_TRY(cudaMemcpyAsync(evals, a.evals, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

constexpr int kRlChoiceRows = 7;  // q, alpha, pos_alpha, v, z, p_outlier, w_outlier

int rl_choice_impl(const int64_t* response, const double* feedback, const int64_t* split_by, int64_t n,
                   const double* params, int64_t P, double* logp)
{
    if (P < 0 || (P > 0 && (!params || !logp))) return fail(HDDM_ERR_ARG, "rl: bad buffer");
    TRY(ensure_ready());
    if (P == 0) return HDDM_OK;
    RlTask t;
    TRY(build_task(nullptr, response, feedback, split_by, n, true, t));
    const int64_t n_runs = (int64_t)t.run_off.size() - 1;
    Stager S;
    const int s_resp = S.add(t.resp.data(), (size_t)n * sizeof(int32_t));
    const int s_fb = S.add(t.fb.data(), (size_t)n * sizeof(double));
    const int s_run = S.add(t.run_off.data(), t.run_off.size() * sizeof(int64_t));
    const int s_par = S.add(params, (size_t)P * kRlChoiceRows * sizeof(double));
    const int s_partial = S.scratch((size_t)P * (size_t)(n_runs > 0 ? n_runs : 1) * sizeof(double));
    const int s_logp = S.scratch((size_t)P * sizeof(double));
    cudaStream_t s = g_rt.stream;
    TRY(S.commit(s));
    const double* d_par = S.dev<double>(s_par);
    RlChoiceArgs c;
    memset(&c, 0, sizeof(c));
    c.scan.response = S.dev<int32_t>(s_resp);
    c.scan.feedback = S.dev<double>(s_fb);
    c.scan.run_off = S.dev<int64_t>(s_run);
    c.scan.n = n;
    c.scan.n_runs = n_runs;
    c.scan.n_sets = P;
    c.scan.q = d_par + 0 * P;
    c.scan.alpha = d_par + 1 * P;
    c.scan.pos_alpha = d_par + 2 * P;
    c.scan.v = d_par + 3 * P;
    c.z = d_par + 4 * P;
    c.p_outlier = d_par + 5 * P;
    c.w_outlier = d_par + 6 * P;
    c.scan.q_stride = c.scan.alpha_set_stride = c.scan.pos_stride = c.scan.v_set_stride = 1;
    c.z_stride = c.pout_stride = c.wout_stride = 1;
    c.partial = S.dev<double>(s_partial);
    c.logp = S.dev<double>(s_logp);
    CUDA_TRY(launch_rl_choice(c, s));
    CUDA_TRY(cudaMemcpyAsync(logp, c.logp, (size_t)P * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

}  // namespace

extern "C" {

int hddm_wiener_like_rlddm(const double* x, const int64_t* response, const double* feedback,
                           const int64_t* split_by, int64_t n, double q, double alpha, double pos_alpha,
                           double v, double sv, double a, double z, double sz, double t, double st, double err,
                           int n_st, int n_sz, int use_adaptive, double simps_err, double p_outlier,
                           double w_outlier, double* logp)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    const double params[kRlRows] = {v, sv, a, z, sz, t, st, p_outlier, w_outlier, q, alpha, pos_alpha};
    return rlddm_impl(x, response, feedback, split_by, n, params, 1,
                      make_numerics(err, n_st, n_sz, use_adaptive, simps_err), logp, nullptr);
}

int hddm_wiener_like_rlddm_batch(const double* x, const int64_t* response, const double* feedback,
                                 const int64_t* split_by, int64_t n, const double* params, int64_t P,
                                 double err, int n_st, int n_sz, int use_adaptive, double simps_err,
                                 double* logp, unsigned long long* evals)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    return rlddm_impl(x, response, feedback, split_by, n, params, P,
                      make_numerics(err, n_st, n_sz, use_adaptive, simps_err), logp, evals);
}

int hddm_wiener_like_rl(const int64_t* response, const double* feedback, const int64_t* split_by, int64_t n,
                        double q, double alpha, double pos_alpha, double v, double z, double p_outlier,
                        double w_outlier, double* logp)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    const double params[kRlChoiceRows] = {q, alpha, pos_alpha, v, z, p_outlier, w_outlier};
    return rl_choice_impl(response, feedback, split_by, n, params, 1, logp);
}

int hddm_wiener_like_rl_batch(const int64_t* response, const double* feedback, const int64_t* split_by,
                              int64_t n, const double* params, int64_t P, double* logp)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    return rl_choice_impl(response, feedback, split_by, n, params, P, logp);
}

// ---- regression likelihood (HDDMRegressor): trials + design matrices resident on the device ---------------------

}  // extern "C"

struct hddm_regdata {
    int64_t n = 0;
    int n_terms = 0;
    int64_t K[kMaxRegTerms] = {0, 0, 0, 0};
    double* d_x = nullptr;                       // trials ordered boundary-major, then |rt| (one-time layout)
    double* d_design[kMaxRegTerms] = {nullptr, nullptr, nullptr, nullptr};  // [K][n] each, rows in the same order
    double* d_block = nullptr;                   // one allocation behind all of the above
    double* d_scratch = nullptr;                 // partial sums / logp / params / coefficients, grown on demand
    size_t scratch_bytes = 0;
};

extern "C" {

int hddm_regdata_upload(const double* x, int64_t n, int n_terms, const int64_t* K, const double* const* design,
                        hddm_regdata** out)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!out || n < 0 || n_terms < 0 || n_terms > kMaxRegTerms || (n > 0 && !x) || (n_terms > 0 && (!K || !design)))
        return fail(HDDM_ERR_ARG, "regdata_upload: bad argument (at most 4 regression terms)");
    for (int r = 0; r < n_terms; ++r)
        if (K[r] < 1 || K[r] > kMaxRegK || !design[r])
            return fail(HDDM_ERR_ARG, "regdata_upload: a design matrix needs 1..16 columns");
    TRY(ensure_ready());
    // One-time layout on the host: lower-boundary trials first, each group ascending in |rt| (what sort_trials does for
    // the uniform-parameter entry points), so that the lanes of a warp share boundary, series branch and term count;
    // the design rows follow the same permutation.
    std::vector<int64_t> order((size_t)n);
    std::iota(order.begin(), order.end(), (int64_t)0);
    auto key = [&](int64_t i) {
        uint64_t bits;
        const double v = x[i];
        memcpy(&bits, &v, sizeof(bits));
        return (bits & 0x7fffffffffffffffULL) | (v > 0 ? 0x8000000000000000ULL : 0ULL);
    };
    std::stable_sort(order.begin(), order.end(), [&](int64_t l, int64_t r) { return key(l) < key(r); });
    size_t doubles = (size_t)(n > 0 ? n : 1);
    for (int r = 0; r < n_terms; ++r) doubles += (size_t)K[r] * (size_t)(n > 0 ? n : 1);
    std::vector<double> host(doubles);
    for (int64_t j = 0; j < n; ++j) host[(size_t)j] = x[order[(size_t)j]];
    size_t off = (size_t)(n > 0 ? n : 1);
    hddm_regdata* h = new (std::nothrow) hddm_regdata();
    if (!h) return fail(HDDM_ERR_ARG, "regdata_upload: out of host memory");
    h->n = n;
    h->n_terms = n_terms;
    if (cudaMalloc(&h->d_block, doubles * sizeof(double)) != cudaSuccess) {
        delete h;
        (void)cudaGetLastError();
        return fail(HDDM_ERR_CUDA, "regdata_upload: out of device memory");
    }
    h->d_x = h->d_block;
    for (int r = 0; r < n_terms; ++r) {
        h->K[r] = K[r];
        h->d_design[r] = h->d_block + off;
        for (int64_t k = 0; k < K[r]; ++k)
            for (int64_t j = 0; j < n; ++j) host[off + (size_t)k * n + j] = design[r][(size_t)k * n + order[(size_t)j]];
        off += (size_t)K[r] * (size_t)(n > 0 ? n : 1);
    }
    const cudaError_t e = cudaMemcpy(h->d_block, host.data(), doubles * sizeof(double), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(h->d_block);
        delete h;
        CUDA_TRY(e);
    }
    *out = h;
    return HDDM_OK;
}

int hddm_regdata_free(hddm_regdata* h)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!h) return HDDM_OK;
    if (g_rt.ready) cudaStreamSynchronize(g_rt.stream);
    cudaFree(h->d_block);
    cudaFree(h->d_scratch);
    delete h;
    return HDDM_OK;
}

int hddm_regdata_like_batch(hddm_regdata* h, const int* outcome, const int* link, const double* const* coef,
                            const double* params, int64_t P, double err, int n_st, int n_sz, int use_adaptive,
                            double simps_err, double* logp, unsigned long long* evals)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!h || P < 0 || (P > 0 && (!params || !logp)) || (h->n_terms > 0 && (!outcome || !link || !coef)))
        return fail(HDDM_ERR_ARG, "regdata_like_batch: bad argument");
    for (int r = 0; r < h->n_terms; ++r) {
        if (outcome[r] < 0 || outcome[r] > 6 || link[r] < 0 || link[r] > kLinkExp || !coef[r])
            return fail(HDDM_ERR_ARG, "regdata_like_batch: outcome must be 0..6 (v sv a z sz t st), link 0..2");
        for (int q = 0; q < r; ++q)
            if (outcome[q] == outcome[r]) return fail(HDDM_ERR_ARG, "regdata_like_batch: an outcome is regressed twice");
    }
    const Numerics nm = make_numerics(err, n_st, n_sz, use_adaptive, simps_err);
    TRY(check_numerics(n_st, n_sz, use_adaptive));
    TRY(ensure_ready());
    if (P == 0) return HDDM_OK;
    const int64_t n = h->n, chunks = chunks_for(n);
    if (chunks > 65535) return fail(HDDM_ERR_ARG, "more than 65535*2048 trials per parameter set");
    // per-call device scratch: params [9][P] | coef per term [K][P] | partial [P][2 chunks] | logp [P] | evals
    size_t coef_doubles = 0;
    for (int r = 0; r < h->n_terms; ++r) coef_doubles += (size_t)h->K[r] * P;
    const size_t need = ((size_t)kNumPar * P + coef_doubles + (size_t)2 * P * chunks + (size_t)P + 2) * sizeof(double);
    cudaStream_t s = g_rt.stream;
    if (need > h->scratch_bytes) {
        CUDA_TRY(cudaStreamSynchronize(s));
        cudaFree(h->d_scratch);
        h->d_scratch = nullptr;
        h->scratch_bytes = 0;
        CUDA_TRY(cudaMalloc(&h->d_scratch, need + need / 4));
        h->scratch_bytes = need + need / 4;
    }
    double* d_par = h->d_scratch;
    double* d_coef = d_par + (size_t)kNumPar * P;
    double* d_partial = d_coef + coef_doubles;
    double* d_logp = d_partial + (size_t)2 * P * chunks;
    unsigned long long* d_evals = reinterpret_cast<unsigned long long*>(d_logp + P);
    CUDA_TRY(cudaMemcpyAsync(d_par, params, (size_t)kNumPar * P * sizeof(double), cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemsetAsync(d_evals, 0, sizeof(unsigned long long), s));
    LikeArgs a;
    memset(&a, 0, sizeof(a));
    a.x = h->d_x;
    for (int k = 0; k < kNumPar; ++k) {
        a.par[k] = d_par + (size_t)k * P;
        a.set_stride[k] = 1;
    }
    a.n_reg = h->n_terms;
    size_t off = 0;
    for (int r = 0; r < h->n_terms; ++r) {
        CUDA_TRY(cudaMemcpyAsync(d_coef + off, coef[r], (size_t)h->K[r] * P * sizeof(double), cudaMemcpyHostToDevice, s));
        a.reg[r].design = h->d_design[r];
        a.reg[r].coef = d_coef + off;
        a.reg[r].K = (int)h->K[r];
        a.reg[r].outcome = outcome[r];
        a.reg[r].link = link[r];
        off += (size_t)h->K[r] * P;
    }
    a.n = n;
    a.n_sets = P;
    a.chunks = (int)chunks;
    a.mode = kModeMulti;  // wiener_like_multi semantics per set (src/wfpt.pyx:244-274)
    a.nm = nm;
    a.partial = d_partial;
    a.logp = d_logp;
    a.evals = d_evals;
    CUDA_TRY(launch_like(a, s));
    CUDA_TRY(cudaMemcpyAsync(logp, d_logp, (size_t)P * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (evals) CUDA_TRY(cudaMemcpyAsync(evals, d_evals, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

// One-shot form of the above for a regression on the drift with the identity link (kept from round 1).
int hddm_wiener_like_regression_batch(const double* x, int64_t n, const double* design, int64_t K,
                                      const double* coef, const double* params, int64_t P, double err, int n_st,
                                      int n_sz, int use_adaptive, double simps_err, double* logp,
                                      unsigned long long* evals)
{
    if (n < 0 || K < 1 || P < 0 || (n > 0 && (!x || !design)) || (P > 0 && (!coef || !params || !logp))) {
        std::lock_guard<std::mutex> lock(g_rt.mu);
        return fail(HDDM_ERR_ARG, "regression_batch: bad buffer or size");
    }
    hddm_regdata* h = nullptr;
    TRY(hddm_regdata_upload(x, n, 1, &K, &design, &h));
    const int outcome = 0, link = kLinkIdentity;
    const int rc = hddm_regdata_like_batch(h, &outcome, &link, &coef, params, P, err, n_st, n_sz, use_adaptive, simps_err,
                                           logp, evals);
    hddm_regdata_free(h);
    return rc;
}

int hddm_wiener_like_multi_rlddm(const double* x, const int64_t* response, const double* feedback,
                                 const int64_t* split_by, int64_t n, double q, const double* const par[8],
                                 const int64_t stride[8], double err, int n_st, int n_sz, int use_adaptive,
                                 double simps_err, double p_outlier, double w_outlier, double* logp)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (!par || !stride || !logp || (n > 0 && !x)) return fail(HDDM_ERR_ARG, "multi_rlddm: bad buffer");
    const Numerics nm = make_numerics(err, n_st, n_sz, use_adaptive, simps_err);
    TRY(check_numerics(n_st, n_sz, use_adaptive));
    TRY(ensure_ready());
    RlTask t;
    TRY(build_task(x, response, feedback, split_by, n, false, t));
    const int64_t chunks = chunks_for(n);
    if (chunks > 65535) return fail(HDDM_ERR_ARG, "more than 65535*2048 trials per parameter set");
    const int64_t n_runs = (int64_t)t.run_off.size() - 1;

    // scalars: v, sv, a, z, sz, t, st, p_outlier, w_outlier, alpha, q
    double scalars[11] = {0, 0, 0, 0, 0, 0, 0, p_outlier, w_outlier, 0, q};
    Stager S;
    const int s_x = S.add(t.x.data(), (size_t)n * sizeof(double));
    const int s_resp = S.add(t.resp.data(), (size_t)n * sizeof(int32_t));
    const int s_fb = S.add(t.fb.data(), (size_t)n * sizeof(double));
    const int s_run = S.add(t.run_off.data(), t.run_off.size() * sizeof(int64_t));
    int slot[8];
    for (int k = 0; k < 8; ++k) {
        if (!par[k]) return fail(HDDM_ERR_ARG, "multi_rlddm: null parameter pointer");
        if (stride[k] == 0) {
            scalars[k < 7 ? k : 9] = par[k][0];
            slot[k] = -1;
        } else if (stride[k] == 1) {
            slot[k] = S.add(par[k], (size_t)n * sizeof(double));
        } else {
            return fail(HDDM_ERR_ARG, "multi_rlddm: stride must be 0 (scalar) or 1 (per trial)");
        }
    }
    const int s_scal = S.add(scalars, sizeof(scalars));
    const int s_drift = S.scratch((size_t)(n > 0 ? n : 1) * sizeof(double));
    const int s_partial = S.scratch((size_t)chunks * sizeof(double));
    const int s_logp = S.scratch(sizeof(double));
    cudaStream_t s = g_rt.stream;
    TRY(S.commit(s));
    const double* d_scal = S.dev<double>(s_scal);

    RlScanArgs sc;
    memset(&sc, 0, sizeof(sc));
    sc.response = S.dev<int32_t>(s_resp);
    sc.feedback = S.dev<double>(s_fb);
    sc.run_off = S.dev<int64_t>(s_run);
    sc.n = n;
    sc.n_runs = n_runs;
    sc.n_sets = 1;
    sc.q = d_scal + 10;
    sc.v = slot[0] >= 0 ? S.dev<double>(slot[0]) : d_scal + 0;
    sc.v_trial_stride = slot[0] >= 0 ? 1 : 0;
    sc.alpha = slot[7] >= 0 ? S.dev<double>(slot[7]) : d_scal + 9;
    sc.alpha_trial_stride = slot[7] >= 0 ? 1 : 0;
    sc.pos_alpha = nullptr;  // single learning rate, src/wfpt.pyx:316
    sc.drift = S.dev<double>(s_drift);
    CUDA_TRY(launch_rl_scan(sc, s));

    LikeArgs a;
    memset(&a, 0, sizeof(a));
    a.x = S.dev<double>(s_x);
    for (int k = 1; k < 7; ++k) {
        a.par[k] = slot[k] >= 0 ? S.dev<double>(slot[k]) : d_scal + k;
        a.trial_stride[k] = slot[k] >= 0 ? 1 : 0;
    }
    a.par[0] = sc.drift;
    a.trial_stride[0] = 1;
    a.par[7] = d_scal + 7;
    a.par[8] = d_scal + 8;
    a.n = n;
    a.n_sets = 1;
    a.chunks = (int)chunks;
    a.mode = kModeMultiRL;
    a.nm = nm;
    a.partial = S.dev<double>(s_partial);
    a.logp = S.dev<double>(s_logp);
    CUDA_TRY(launch_like(a, s));
    CUDA_TRY(cudaMemcpyAsync(logp, a.logp, sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

}  // extern "C"
