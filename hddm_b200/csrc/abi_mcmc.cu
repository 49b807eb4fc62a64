// abi_mcmc.cu -- native slice-within-Gibbs driver for the hierarchical full DDM of BASELINE.json configs[1]
// ("a minimal slice/Gibbs driver that evaluates all children of a group-level node and all
// chains per launch").  Host C++ around the CUDA likelihood: no Python between the launches of a sweep, which is
// what a single chain needs -- its launches are 4000 evaluations each and the time goes to what surrounds them.
//
// The model, the update order and every random number are those of hddm_b200/sampler.py (HierarchicalDDM, itself
// the restatement of hddm/models/hddm_info.py:103-175 and hddm/models/base.py:578-681 for a batched likelihood):
// priors, slice widths (hddm/models/hddm_info.py:103-105), stepping out + shrinkage with maxiter 5000 (:174),
// numpy's PCG64 stream (the caller passes the generator's 128-bit state and increment and gets them back), so a
// chain drawn here equals the chain sampler.py draws from the same generator state, draw for draw -- the test
// (tests/test_sampler.py) and the reason the Python sampler stays: it is the readable specification.
// Speculation (see slice_step_batched in sampler.py): the points a slice step probes do not depend on the values it
// finds, so each phase (stepping out, shrinkage) is ONE launch over all its candidate points.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <vector>

#include "abi_runtime.cuh"

using namespace hddm;
using namespace hddm::rt;

namespace {

typedef unsigned __int128 u128;

// numpy.random.PCG64: 128-bit LCG, XSL-RR output; Generator.random() = (next64 >> 11) * 2^-53.
struct Pcg64 {
    u128 state, inc;
    uint64_t next64()
    {
        const u128 mult = ((u128)0x2360ED051FC65DA4ULL << 64) | 0x4385DF649FCCF645ULL;
        state = state * mult + inc;
        const uint64_t hi = (uint64_t)(state >> 64), lo = (uint64_t)state;
        const uint64_t x = hi ^ lo;
        const unsigned rot = (unsigned)(hi >> 58);
        return (x >> rot) | (x << ((64 - rot) & 63));
    }
    double random() { return (double)(next64() >> 11) * (1.0 / 9007199254740992.0); }
};

// numpy's pairwise summation of a contiguous double vector (DOUBLE_pairwise_sum): what ndarray.sum does along a
// contiguous axis -- reproduced so that sums over subjects match sampler.py bit for bit.
double np_sum(const double* a, int64_t n)
{
    if (n < 8) {
        double res = 0.;
        for (int64_t i = 0; i < n; ++i) res += a[i];
        return res;
    }
    if (n <= 128) {
        double r[8];
        for (int k = 0; k < 8; ++k) r[k] = a[k];
        int64_t i;
        for (i = 8; i < n - (n % 8); i += 8)
            for (int k = 0; k < 8; ++k) r[k] += a[i + k];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    }
    int64_t n2 = n / 2;
    n2 -= n2 % 8;
    return np_sum(a, n2) + np_sum(a + n2, n - n2);
}

const double kInf = INFINITY;

double norm_logpdf(double x, double mu, double sd)
{
    const double q = (x - mu) / sd;
    return -0.5 * (q * q) - log(sd) - 0.9189385332046727;
}
double halfnorm_logpdf(double x, double sd)
{
    if (!(x > 0)) return -kInf;
    const double q = x / sd;
    return -0.5 * (q * q) - log(sd) - 0.2257913526447274;
}
double gamma_logpdf(double x, double shape, double rate)
{
    if (!((x > 0) && (shape > 0) && (rate > 0))) return -kInf;
    return shape * log(rate) - lgamma(shape) + (shape - 1) * log(x) - rate * x;
}
double gamma_mean_std(double x, double mean, double sd) { return gamma_logpdf(x, (mean * mean) / (sd * sd), mean / (sd * sd)); }
double beta_logpdf(double x, double a, double b)
{
    if (!((x > 0) && (x < 1))) return -kInf;
    return (a - 1) * log(x) + (b - 1) * log1p(-x) - (lgamma(a) + lgamma(b) - lgamma(a + b));
}

constexpr int kMaxIter = 5000;  // hddm/models/hddm_info.py:174

// target(K, X[K][n], out[K][n]); `batch` numbers the calls made inside one slice step from 0.
typedef std::function<int(int K, const double* X, double* out, int batch)> Target;

// One slice update of n independent units; mirrors sampler.slice_step_batched (hence sampler.slice_step).
// picks_batch / picks_row (may be null): where each unit's accepted point was evaluated (row -1: kept x0).
int slice_step(int64_t n, double* x, const double* lp0_in, const Target& target, double width, bool has_left,
               double left, Pcg64& rng, int m_out, int m_shrink, int* picks_batch, int* picks_row, int* n_batches)
{
    std::vector<double> x0(x, x + n), lp0(n), y(n), L(n), R(n);
    int batches = 0;
    std::vector<double> X, V;
    if (lp0_in) {
        memcpy(lp0.data(), lp0_in, sizeof(double) * n);
    } else {
        TRY(target(1, x0.data(), lp0.data(), batches));
        ++batches;
    }
    for (int64_t i = 0; i < n; ++i) y[i] = lp0[i] + log(rng.random());
    for (int64_t i = 0; i < n; ++i) L[i] = x0[i] - width * rng.random();
    std::vector<char> gl(n), gr(n, 1);
    for (int64_t i = 0; i < n; ++i) {
        R[i] = L[i] + width;
        if (has_left) L[i] = L[i] > left ? L[i] : left;  // np.maximum
        gl[i] = has_left ? (L[i] > left) : 1;
    }
    auto any = [&](const std::vector<char>& v) {
        for (char c : v)
            if (c) return true;
        return false;
    };
    int it = 0;
    while ((any(gl) || any(gr)) && it < kMaxIter) {
        // the next m_out rounds of both ends, as the sequential step would probe them if every unit kept growing
        X.assign((size_t)2 * m_out * n, 0.0);
        V.assign((size_t)2 * m_out * n, 0.0);
        std::vector<double> Lk(L), Rk(R);
        std::vector<char> g2(gl);
        for (int k = 0; k < m_out; ++k) {
            for (int64_t i = 0; i < n; ++i) {
                X[(size_t)(2 * k) * n + i] = g2[i] ? Lk[i] : x0[i];
                X[(size_t)(2 * k + 1) * n + i] = gr[i] ? Rk[i] : x0[i];
                if (g2[i]) Lk[i] = Lk[i] - width;
                if (has_left && Lk[i] <= left) {
                    Lk[i] = left;
                    g2[i] = 0;
                }
                if (gr[i]) Rk[i] = Rk[i] + width;
            }
        }
        TRY(target(2 * m_out, X.data(), V.data(), batches));
        ++batches;
        for (int k = 0; k < m_out; ++k) {
            if (!((any(gl) || any(gr)) && it < kMaxIter)) break;
            for (int64_t i = 0; i < n; ++i) {
                gl[i] = gl[i] && (V[(size_t)(2 * k) * n + i] > y[i]);
                gr[i] = gr[i] && (V[(size_t)(2 * k + 1) * n + i] > y[i]);
                if (gl[i]) L[i] = L[i] - width;
                if (has_left && L[i] <= left) {
                    L[i] = left;
                    gl[i] = 0;
                }
                if (gr[i]) R[i] = R[i] + width;
            }
            ++it;
        }
    }
    // shrinkage
    std::vector<double> x1(x0), lp1(lp0);
    std::vector<char> todo(n, 1);
    if (picks_batch)
        for (int64_t i = 0; i < n; ++i) picks_batch[i] = picks_row[i] = -1;
    it = 0;
    std::vector<double> U;
    while (any(todo) && it < kMaxIter) {
        const Pcg64 saved = rng;
        U.resize((size_t)m_shrink * n);
        for (size_t k = 0; k < U.size(); ++k) U[k] = rng.random();  // row r = the draws of round r
        X.assign((size_t)m_shrink * n, 0.0);
        V.assign((size_t)m_shrink * n, 0.0);
        std::vector<double> Lk(L), Rk(R);
        for (int r = 0; r < m_shrink; ++r)  // candidates of the next rounds if every pending unit keeps being rejected
            for (int64_t i = 0; i < n; ++i) {
                const double cand = todo[i] ? Lk[i] + (Rk[i] - Lk[i]) * U[(size_t)r * n + i] : x1[i];
                X[(size_t)r * n + i] = cand;
                if (todo[i]) {  // (a NaN candidate moves neither end, as numpy's two comparisons)
                    if (cand < x0[i]) Lk[i] = cand;
                    else if (cand >= x0[i]) Rk[i] = cand;
                }
            }
        TRY(target(m_shrink, X.data(), V.data(), batches));
        int used = 0;
        for (int r = 0; r < m_shrink; ++r) {
            if (!(any(todo) && it < kMaxIter)) break;
            for (int64_t i = 0; i < n; ++i) {
                if (!todo[i]) continue;
                const double cand = X[(size_t)r * n + i], lpc = V[(size_t)r * n + i];
                if (lpc > y[i]) {
                    x1[i] = cand;
                    lp1[i] = lpc;
                    todo[i] = 0;
                    if (picks_batch) {
                        picks_batch[i] = batches;
                        picks_row[i] = r;
                    }
                } else if (cand < x0[i]) {
                    L[i] = cand;
                } else if (cand >= x0[i]) {
                    R[i] = cand;
                }
            }
            ++it;
            ++used;
        }
        ++batches;
        if (used < m_shrink) {  // give back the draws the sequential step would not have made
            rng = saved;
            for (int64_t k = 0; k < (int64_t)used * n; ++k) rng.next64();
        }
    }
    memcpy(x, x1.data(), sizeof(double) * n);
    if (n_batches) *n_batches += batches;
    return HDDM_OK;
}

}  // namespace

// g: [9][C] rows v, v_std, a, a_std, t, t_std, sv, sz, st (group level); subj: [3][C][S] rows v, a, t (subject level).
extern "C" int hddm_hier_ddm_sample(const hddm_trials* trials, const int64_t* node_off, int64_t S, int64_t C, double* g,
                                    double* subj, double z, double p_outlier, double w_outlier, double err, int n_st,
                                    int n_sz, int use_adaptive, double simps_err, uint64_t* rng_state, int64_t n_iter,
                                    double* trace_g, double* trace_subj, int64_t* n_launches)
{
    if (!trials || !node_off || S < 1 || C < 1 || !g || !subj || !rng_state || n_iter < 0)
        return fail(HDDM_ERR_ARG, "hier_ddm_sample: bad argument");
    Pcg64 rng;
    rng.state = ((u128)rng_state[0] << 64) | rng_state[1];
    rng.inc = ((u128)rng_state[2] << 64) | rng_state[3];
    const int64_t CS = C * S;
    enum { GV = 0, GVS, GA, GAS, GT, GTS, GSV, GSZ, GST };
    double* G[9];
    for (int k = 0; k < 9; ++k) G[k] = g + k * C;
    double* SUB[3] = {subj, subj + CS, subj + 2 * CS};  // v, a, t
    const int sub_row[3] = {0, 2, 5};                   // rows of the [9][P] parameter layout: v, a, t
    const int sub_mean[3] = {GV, GA, GT}, sub_std[3] = {GVS, GAS, GTS};
    const double sub_width[3] = {1.5, 1.0, 0.01};       // hddm/models/hddm_info.py:103-105
    int64_t launches = 0;
    std::vector<double> params, ll_buf;
    std::vector<int64_t> set_node;
    std::vector<std::vector<double>> payloads;  // node log-likelihoods [K][C][S] of every batch of the current step

    // node log-likelihoods for K candidate values of parameter-layout row `row`: X is [K][C][S] (per_subject) or [K][C]
    auto node_ll_many = [&](int row, int K, const double* X, bool per_subject, double* out) -> int {
        const int64_t P = (int64_t)K * CS;
        params.resize((size_t)9 * P);
        set_node.resize((size_t)P);
        for (int k = 0; k < K; ++k)
            for (int64_t c = 0; c < C; ++c)
                for (int64_t s = 0; s < S; ++s) {
                    const int64_t p = ((int64_t)k * C + c) * S + s, cs = c * S + s;
                    double* col = params.data() + p;
                    col[0 * P] = SUB[0][cs]; col[1 * P] = G[GSV][c]; col[2 * P] = SUB[1][cs]; col[3 * P] = z;
                    col[4 * P] = G[GSZ][c]; col[5 * P] = SUB[2][cs]; col[6 * P] = G[GST][c];
                    col[7 * P] = p_outlier; col[8 * P] = w_outlier;
                    if (row >= 0) col[(int64_t)row * P] = per_subject ? X[((int64_t)k * C + c) * S + s] : X[(int64_t)k * C + c];
                    set_node[(size_t)p] = s;
                }
        ++launches;
        return hddm_trials_wiener_like_nodes(trials, node_off, S, set_node.data(), params.data(), P, err, n_st, n_sz,
                                             use_adaptive, simps_err, out, nullptr);
    };

    std::vector<double> node_ll((size_t)CS);
    TRY(node_ll_many(-1, 1, nullptr, true, node_ll.data()));
    std::vector<int> pb((size_t)CS), pr((size_t)CS);
    std::vector<double> lp0((size_t)CS), tmp;

    for (int64_t iter = 0; iter < n_iter; ++iter) {
        // subject-level v, a, t: C x S units in lock-step
        for (int f = 0; f < 3; ++f) {
            payloads.clear();
            const double* mean = G[sub_mean[f]];
            const double* sd = G[sub_std[f]];
            auto prior = [&](int64_t c, double x) { return f == 0 ? norm_logpdf(x, mean[c], sd[c]) : gamma_mean_std(x, mean[c], sd[c]); };
            Target target = [&](int K, const double* X, double* out, int) -> int {
                payloads.emplace_back((size_t)K * CS);
                std::vector<double>& ll = payloads.back();
                TRY(node_ll_many(sub_row[f], K, X, true, ll.data()));
                for (int k = 0; k < K; ++k)
                    for (int64_t c = 0; c < C; ++c)
                        for (int64_t s = 0; s < S; ++s) {
                            const int64_t i = ((int64_t)k * C + c) * S + s;
                            out[i] = ll[(size_t)i] + prior(c, X[i]);
                        }
                return HDDM_OK;
            };
            for (int64_t c = 0; c < C; ++c)
                for (int64_t s = 0; s < S; ++s) lp0[(size_t)(c * S + s)] = node_ll[(size_t)(c * S + s)] + prior(c, SUB[f][c * S + s]);
            TRY(slice_step(CS, SUB[f], lp0.data(), target, sub_width[f], false, 0.0, rng, 3, 6, pb.data(), pr.data(), nullptr));
            for (int64_t i = 0; i < CS; ++i)
                if (pr[(size_t)i] >= 0) node_ll[(size_t)i] = payloads[(size_t)pb[(size_t)i]][(size_t)pr[(size_t)i] * CS + i];
        }
        // group-only sv, sz, st: every observed node of the chain is a child
        const int grp[3] = {GSV, GSZ, GST};
        const int grp_row[3] = {1, 4, 6};
        const double grp_width[3] = {3.0, 1.1, 0.1};
        for (int f = 0; f < 3; ++f) {
            payloads.clear();
            auto prior = [&](double x) { return f == 0 ? halfnorm_logpdf(x, 2.0) : (f == 1 ? beta_logpdf(x, 1.0, 3.0) : halfnorm_logpdf(x, 0.3)); };
            Target target = [&](int K, const double* X, double* out, int) -> int {
                payloads.emplace_back((size_t)K * CS);
                std::vector<double>& ll = payloads.back();
                TRY(node_ll_many(grp_row[f], K, X, false, ll.data()));
                for (int k = 0; k < K; ++k)
                    for (int64_t c = 0; c < C; ++c) out[(int64_t)k * C + c] = np_sum(ll.data() + ((int64_t)k * C + c) * S, S) + prior(X[(int64_t)k * C + c]);
                return HDDM_OK;
            };
            for (int64_t c = 0; c < C; ++c) lp0[(size_t)c] = np_sum(node_ll.data() + c * S, S) + prior(G[grp[f]][c]);
            TRY(slice_step(C, G[grp[f]], lp0.data(), target, grp_width[f], true, 0.0, rng, 3, 6, pb.data(), pr.data(), nullptr));
            for (int64_t c = 0; c < C; ++c)
                if (pr[(size_t)c] >= 0)
                    memcpy(node_ll.data() + c * S, payloads[(size_t)pb[(size_t)c]].data() + ((int64_t)pr[(size_t)c] * C + c) * S, sizeof(double) * S);
        }
        // group means and spreads: children are the subject-level priors only (no likelihood)
        const double hyper_width[3] = {1.5, 1.0, 0.01}, std_width[3] = {1.0, 1.0, 0.15}, std_sd[3] = {2.0, 2.0, 1.0};
        for (int f = 0; f < 3; ++f) {
            const double* xs = SUB[f];
            auto subj_prior = [&](double x, double m, double sdv) { return f == 0 ? norm_logpdf(x, m, sdv) : gamma_mean_std(x, m, sdv); };
            auto hyper = [&](double x) { return f == 0 ? norm_logpdf(x, 2.0, 3.0) : (f == 1 ? gamma_mean_std(x, 1.5, 0.75) : gamma_mean_std(x, 0.4, 0.2)); };
            tmp.resize((size_t)S);
            Target tg = [&](int K, const double* X, double* out, int) -> int {
                for (int k = 0; k < K; ++k)
                    for (int64_t c = 0; c < C; ++c) {
                        for (int64_t s = 0; s < S; ++s) tmp[(size_t)s] = subj_prior(xs[c * S + s], X[(int64_t)k * C + c], G[sub_std[f]][c]);
                        out[(int64_t)k * C + c] = np_sum(tmp.data(), S) + hyper(X[(int64_t)k * C + c]);
                    }
                return HDDM_OK;
            };
            TRY(slice_step(C, G[sub_mean[f]], nullptr, tg, hyper_width[f], false, 0.0, rng, 3, 6, nullptr, nullptr, nullptr));
            Target ts = [&](int K, const double* X, double* out, int) -> int {
                for (int k = 0; k < K; ++k)
                    for (int64_t c = 0; c < C; ++c) {
                        for (int64_t s = 0; s < S; ++s) tmp[(size_t)s] = subj_prior(xs[c * S + s], G[sub_mean[f]][c], X[(int64_t)k * C + c]);
                        out[(int64_t)k * C + c] = np_sum(tmp.data(), S) + halfnorm_logpdf(X[(int64_t)k * C + c], std_sd[f]);
                    }
                return HDDM_OK;
            };
            TRY(slice_step(C, G[sub_std[f]], nullptr, ts, std_width[f], true, 0.0, rng, 3, 6, nullptr, nullptr, nullptr));
        }
        if (trace_g) memcpy(trace_g + iter * 9 * C, g, sizeof(double) * 9 * C);
        if (trace_subj) memcpy(trace_subj + iter * 3 * CS, subj, sizeof(double) * 3 * CS);
    }
    rng_state[0] = (uint64_t)(rng.state >> 64); rng_state[1] = (uint64_t)rng.state;
    rng_state[2] = (uint64_t)(rng.inc >> 64); rng_state[3] = (uint64_t)rng.inc;
    if (n_launches) *n_launches = launches;
    return HDDM_OK;
}
