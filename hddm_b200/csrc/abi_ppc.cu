// abi_ppc.cu -- extern "C" entry points of the posterior-predictive batch (BASELINE
// config 5): simulated data sets and their summary statistics for many posterior draws, and the cdfdif
// quantile probabilities of many parameter sets, one launch sequence per slab of draws.
#include <cstring>
#include <vector>

#include "abi_runtime.cuh"

using namespace hddm;
using namespace hddm::rt;

namespace {

constexpr size_t kTableSlabBytes = (size_t)4 << 30;  // density tables of one slab of draws

// Draws [d0, d0 + Dn) of the batch: tables -> cumulative tables -> RTs (-> statistics).
int ppc_slab(const double* d_grid, int64_t m, const double* h_params, int64_t D, int64_t d0, int64_t Dn,
             const double* h_f, const double* h_u, unsigned long long seed, int64_t first_draw, int64_t S,
             double* h_rts, double* h_stats, cudaStream_t s)
{
    // per-slab parameter table for the density kernel: t = st = 0, no outlier mix (src/wfpt.pyx:335)
    std::vector<double> par((size_t)kNumPar * Dn, 0.0), tt((size_t)2 * Dn);
    for (int64_t d = 0; d < Dn; ++d) {
        for (int k = 0; k < 5; ++k) par[(size_t)k * Dn + d] = h_params[(size_t)k * D + d0 + d];
        tt[(size_t)d] = h_params[(size_t)5 * D + d0 + d];
        tt[(size_t)Dn + d] = h_params[(size_t)6 * D + d0 + d];
    }
    Layout L;
    const size_t off_tab = L.reserve((size_t)Dn * m * sizeof(double));
    const size_t off_par = L.reserve(par.size() * sizeof(double));
    const size_t off_tt = L.reserve(tt.size() * sizeof(double));
    const size_t off_f = h_f ? L.reserve((size_t)Dn * S * sizeof(double)) : 0;
    const size_t off_u = h_u ? L.reserve((size_t)Dn * S * sizeof(double)) : 0;
    const size_t off_r = L.reserve((size_t)Dn * S * sizeof(double));
    const size_t off_st = L.reserve((size_t)Dn * kPpcStats * sizeof(double));
    const size_t off_last = L.reserve((size_t)Dn * sizeof(double));
    // the grid lives at the front of the arena (reserved by the caller): keep it there
    const size_t grid_bytes = align_up((size_t)m * sizeof(double));
    char* base = g_rt.arena + grid_bytes;
    auto at = [&](size_t off) { return reinterpret_cast<double*>(base + off); };
    (void)d_grid;
    CUDA_TRY(cudaMemcpyAsync(at(off_par), par.data(), par.size() * sizeof(double), cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemcpyAsync(at(off_tt), tt.data(), tt.size() * sizeof(double), cudaMemcpyHostToDevice, s));
    if (h_f)
        CUDA_TRY(cudaMemcpyAsync(at(off_f), h_f + (size_t)d0 * S, (size_t)Dn * S * sizeof(double),
                                 cudaMemcpyHostToDevice, s));
    if (h_u)
        CUDA_TRY(cudaMemcpyAsync(at(off_u), h_u + (size_t)d0 * S, (size_t)Dn * S * sizeof(double),
                                 cudaMemcpyHostToDevice, s));
    PdfArgs pa;
    pa.x = reinterpret_cast<const double*>(g_rt.arena);
    pa.y = at(off_tab);
    pa.n = m;
    pa.logp = 0;
    pa.evals = nullptr;
    memset(pa.par, 0, sizeof(pa.par));
    pa.nm = make_numerics(1e-4, 2, 2, 1, 1e-3);  // cpdef defaults of full_pdf, src/pdf.pxi:106-107
    pa.par_sets = at(off_par);
    pa.n_sets = Dn;
    pa.x_set_stride = 0;
    CUDA_TRY(launch_pdf_array(pa, s));
    CUDA_TRY(launch_cumsum_batch(at(off_tab), m, Dn, at(off_last), s));
    CUDA_TRY(launch_sample_rts_batch(pa.x, at(off_tab), m, h_f ? at(off_f) : nullptr, h_u ? at(off_u) : nullptr,
                                     S, Dn, at(off_tt), at(off_tt) + Dn, seed, first_draw + d0, at(off_last), at(off_r), s));
    if (h_stats) {
        CUDA_TRY(launch_ppc_stats(at(off_r), S, Dn, at(off_st), s));
        CUDA_TRY(cudaMemcpyAsync(h_stats + (size_t)d0 * kPpcStats, at(off_st),
                                 (size_t)Dn * kPpcStats * sizeof(double), cudaMemcpyDeviceToHost, s));
    }
    if (h_rts)
        CUDA_TRY(cudaMemcpyAsync(h_rts + (size_t)d0 * S, at(off_r), (size_t)Dn * S * sizeof(double),
                                 cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));  // the host vectors above die with this frame
    return HDDM_OK;
}

int ppc_impl(const double* xgrid, int64_t m, const double* params, int64_t D, const double* f, const double* u,
             unsigned long long seed, int64_t first_draw, int64_t S, double* rts, double* stats)
{
    if (m <= 0 || D < 0 || S < 0 || !xgrid || (D > 0 && !params) || (D > 0 && S > 0 && !rts && !stats))
        return fail(HDDM_ERR_ARG, "gen_rts_from_cdf_batch: bad buffer or size");
    if (stats && (S < 1 || S > kPpcMaxSamples))
        return fail(HDDM_ERR_ARG, "gen_rts_from_cdf_batch: statistics need 1 <= samples <= 16384 per draw");
    if (f && !u)
        for (int64_t d = 0; d < D; ++d)
            if (params[(size_t)6 * D + d] != 0)
                return fail(HDDM_ERR_ARG, "gen_rts_from_cdf_batch: u_delay is required when any st != 0");
    TRY(ensure_ready());
    if (D == 0 || S == 0) return HDDM_OK;
    // slab size: bounded table bytes and the 65535 limit of gridDim.y
    int64_t slab = (int64_t)(kTableSlabBytes / ((size_t)m * sizeof(double)));
    if (slab < 1) slab = 1;
    if (slab > 32768) slab = 32768;
    if (slab > D) slab = D;
    const size_t per_draw = (size_t)m * 8 + (size_t)S * 8 * 3 + (kNumPar + 3 + kPpcStats) * 8;
    TRY(ensure_arena(align_up((size_t)m * sizeof(double)) + (size_t)slab * per_draw + (16 << 10)));
    cudaStream_t s = g_rt.stream;
    // A posterior-predictive loop passes the same grid call after call (np.arange(lb, ub, dt): 960 KB at dt = 1e-4):
    // it stays at the front of the arena as long as nobody else carved the arena in between and the bytes are equal.
    static std::vector<double> grid_seen;
    static unsigned long long grid_epoch = 0;
    static const char* grid_arena = nullptr;
    const bool resident = grid_arena == g_rt.arena && grid_epoch + 1 == g_rt.arena_epoch && (int64_t)grid_seen.size() == m &&
                          memcmp(grid_seen.data(), xgrid, (size_t)m * sizeof(double)) == 0;
    if (!resident) {
        CUDA_TRY(cudaMemcpyAsync(g_rt.arena, xgrid, (size_t)m * sizeof(double), cudaMemcpyHostToDevice, s));
        grid_seen.assign(xgrid, xgrid + m);
        grid_arena = g_rt.arena;
    }
    grid_epoch = g_rt.arena_epoch;
    for (int64_t d0 = 0; d0 < D; d0 += slab) {
        const int64_t Dn = D - d0 < slab ? D - d0 : slab;
        TRY(ppc_slab(reinterpret_cast<const double*>(g_rt.arena), m, params, D, d0, Dn, f, u, seed, first_draw, S, rts,
                     stats, s));
    }
    return HDDM_OK;
}

}  // namespace

extern "C" {

int hddm_gen_rts_from_cdf_batch(const double* xgrid, int64_t m, const double* params, int64_t D, const double* f,
                                const double* u_delay, int64_t samples, double* rts, double* stats)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (D > 0 && samples > 0 && !f) return fail(HDDM_ERR_ARG, "gen_rts_from_cdf_batch: null uniforms");
    return ppc_impl(xgrid, m, params, D, f, u_delay, 0ull, 0, samples, rts, stats);
}

int hddm_gen_rts_from_cdf_batch_seeded(const double* xgrid, int64_t m, const double* params, int64_t D,
                                       unsigned long long seed, int64_t first_draw, int64_t samples, double* rts,
                                       double* stats)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (first_draw < 0) return fail(HDDM_ERR_ARG, "gen_rts_from_cdf_batch_seeded: negative first_draw");
    return ppc_impl(xgrid, m, params, D, nullptr, nullptr, seed, first_draw, samples, rts, stats);
}

int hddm_ppc_stats(const double* rts, int64_t D, int64_t samples, double* stats)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (D < 0 || (D > 0 && (!rts || !stats))) return fail(HDDM_ERR_ARG, "ppc_stats: bad buffer");
    if (D > 0 && (samples < 1 || samples > kPpcMaxSamples))
        return fail(HDDM_ERR_ARG, "ppc_stats: 1 <= samples <= 16384 per data set");
    TRY(ensure_ready());
    if (D == 0) return HDDM_OK;
    Layout L;
    const size_t off_r = L.reserve((size_t)D * samples * sizeof(double));
    const size_t off_s = L.reserve((size_t)D * kPpcStats * sizeof(double));
    TRY(ensure_arena(L.total));
    cudaStream_t s = g_rt.stream;
    CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_r), rts, (size_t)D * samples * sizeof(double), cudaMemcpyHostToDevice, s));
    CUDA_TRY(launch_ppc_stats(L.at<double>(off_r), samples, D, L.at<double>(off_s), s));
    CUDA_TRY(cudaMemcpyAsync(stats, L.at<double>(off_s), (size_t)D * kPpcStats * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

int hddm_dmat_cdf_batch(const double* x, int64_t n, const double* params, int64_t D, double* y, int32_t* status)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (n < 0 || D < 0 || (n > 0 && D > 0 && (!x || !y || !params)) || (D > 0 && !status))
        return fail(HDDM_ERR_ARG, "dmat_cdf_batch: bad buffer");
    TRY(ensure_ready());
    if (D == 0) return HDDM_OK;
    // the per-call guards of src/cdfdif_wrapper.pyx:20-25, per parameter set
    std::vector<int32_t> ok((size_t)D);
    for (int64_t d = 0; d < D; ++d) {
        const double sv = params[1 * D + d], a = params[2 * D + d], z = params[3 * D + d], sz = params[4 * D + d],
                     t = params[5 * D + d], st = params[6 * D + d], po = params[7 * D + d], wo = params[8 * D + d];
        int code = HDDM_OK;
        if (po > 0) {
            double peak = 0;
            bool has_nan = false;
            for (int64_t i = 0; i < n; ++i) {
                const double mag = fabs(x[d * n + i]);
                has_nan |= (mag != mag);
                if (mag > peak) peak = mag;
            }
            if (n > 0 && (has_nan || !(peak < (1. / (2 * wo))))) code = HDDM_ERR_RT_RANGE;
        }
        if (code == HDDM_OK && out_of_support(sv, a, z, sz, t, st, po)) code = HDDM_ERR_SUPPORT;
        status[d] = code;
        ok[(size_t)d] = code == HDDM_OK;
    }
    if (n == 0) return HDDM_OK;
    Layout L;
    const size_t off_x = L.reserve((size_t)D * n * sizeof(double));
    const size_t off_y = L.reserve((size_t)D * n * sizeof(double));
    const size_t off_p = L.reserve((size_t)D * kNumPar * sizeof(double));
    const size_t off_ok = L.reserve((size_t)D * sizeof(int32_t));
    TRY(ensure_arena(L.total));
    cudaStream_t s = g_rt.stream;
    CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_x), x, (size_t)D * n * sizeof(double), cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_p), params, (size_t)D * kNumPar * sizeof(double), cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemcpyAsync(L.at<int32_t>(off_ok), ok.data(), (size_t)D * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemsetAsync(L.at<char>(off_y), 0xff, (size_t)D * n * sizeof(double), s));  // skipped sets read NaN
    CdfArgs ca;
    ca.x = L.at<double>(off_x);
    ca.y = L.at<double>(off_y);
    ca.n = n;
    memset(ca.par, 0, sizeof(ca.par));
    ca.par_sets = L.at<double>(off_p);
    ca.ok = L.at<int32_t>(off_ok);
    ca.n_sets = D;
    CUDA_TRY(launch_cdf_array(ca, s));
    CUDA_TRY(cudaMemcpyAsync(y, ca.y, (size_t)D * n * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

}  // extern "C"

// Shared body of the two simulator entry points.  switch_par == nullptr: no switch, u_init has 3 columns.
static int simulate_drift_impl(const double* params, const double* switch_par, int64_t D, int64_t samples, double dt,
                               double intra_sv, unsigned long long seed, const double* u_init,
                               const double* u_steps, int64_t max_steps, double* rts)
{
    std::lock_guard<std::mutex> lock(g_rt.mu);
    if (D < 0 || samples < 0 || max_steps < 1 || !(dt > 0) || !(intra_sv > 0) ||
        (D > 0 && samples > 0 && (!params || !rts)) || ((u_init == nullptr) != (u_steps == nullptr)))
        return fail(HDDM_ERR_ARG, "simulate_drift_batch: bad buffer, size, dt or intra_sv");
    if (switch_par)  // the reference draws blocks of round(t_switch / dt) steps: a block of 0 steps fails there
        for (int64_t d = 0; d < D; ++d)
            if (!(nearbyint(switch_par[2 * D + d] / dt) >= 1))
                return fail(HDDM_ERR_ARG, "simulate_drift_batch: round(t_switch / dt) must be >= 1");
    TRY(ensure_ready());
    const int64_t n = D * samples;
    if (n == 0) return HDDM_OK;
    const int cols = switch_par ? 4 : 3;
    Layout L;
    const size_t off_p = L.reserve((size_t)D * 7 * sizeof(double));
    const size_t off_w = switch_par ? L.reserve((size_t)D * 3 * sizeof(double)) : 0;
    const size_t off_r = L.reserve((size_t)n * sizeof(double));
    const size_t off_i = u_init ? L.reserve((size_t)n * cols * sizeof(double)) : 0;
    const size_t off_u = u_steps ? L.reserve((size_t)n * max_steps * sizeof(double)) : 0;
    TRY(ensure_arena(L.total));
    cudaStream_t s = g_rt.stream;
    CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_p), params, (size_t)D * 7 * sizeof(double), cudaMemcpyHostToDevice, s));
    if (switch_par)
        CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_w), switch_par, (size_t)D * 3 * sizeof(double), cudaMemcpyHostToDevice, s));
    if (u_init) {
        CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_i), u_init, (size_t)n * cols * sizeof(double), cudaMemcpyHostToDevice, s));
        CUDA_TRY(cudaMemcpyAsync(L.at<double>(off_u), u_steps, (size_t)n * max_steps * sizeof(double), cudaMemcpyHostToDevice, s));
    }
    SimArgs a;
    a.params = L.at<double>(off_p);
    a.switch_par = switch_par ? L.at<double>(off_w) : nullptr;
    a.init_cols = cols;
    a.n_draws = D;
    a.samples = samples;
    a.max_steps = max_steps;
    a.draw0 = 0;
    a.dt = dt;
    a.intra_sv = intra_sv;
    a.seed = seed;
    a.u_init = u_init ? L.at<double>(off_i) : nullptr;
    a.u_steps = u_steps ? L.at<double>(off_u) : nullptr;
    a.rts = L.at<double>(off_r);
    CUDA_TRY(launch_simulate_drift(a, s));
    CUDA_TRY(cudaMemcpyAsync(rts, a.rts, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return HDDM_OK;
}

extern "C" int hddm_simulate_drift_batch(const double* params, int64_t D, int64_t samples, double dt,
                                         double intra_sv, unsigned long long seed, const double* u_init,
                                         const double* u_steps, int64_t max_steps, double* rts)
{
    return simulate_drift_impl(params, nullptr, D, samples, dt, intra_sv, seed, u_init, u_steps, max_steps, rts);
}

extern "C" int hddm_simulate_drift_switch_batch(const double* params, const double* switch_par, int64_t D,
                                                int64_t samples, double dt, double intra_sv,
                                                unsigned long long seed, const double* u_init,
                                                const double* u_steps, int64_t max_steps, double* rts)
{
    if (D > 0 && !switch_par) {
        std::lock_guard<std::mutex> lock(g_rt.mu);
        return fail(HDDM_ERR_ARG, "simulate_drift_switch_batch: switch_par is NULL");
    }
    return simulate_drift_impl(params, switch_par, D, samples, dt, intra_sv, seed, u_init, u_steps, max_steps, rts);
}
