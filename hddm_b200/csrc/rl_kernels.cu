// rl_kernels.cu -- the Q-learning recursions of the reinforcement-learning likelihoods
// (src/wfpt.pyx:79-155 wiener_like_rlddm, :158-241 wiener_like_rl, :277-320 wiener_like_multi_rlddm).
//
// The reference interleaves the Q-value update with one full_pdf per trial, which looks sequential.
// It is not: Q depends on (response, feedback, learning rates) only.  So the recursion runs first, as a
// scan with one thread per (parameter set, run) -- a run is one condition's trials (rlddm / rl) or a block
// of neighbouring trials with equal split_by (multi_rlddm) -- and writes the drift every trial sees;
// the expensive part, the densities, is then the ordinary one-thread-per-(trial, set) likelihood kernel
// with a per-trial drift column.  Threads are laid out set-fastest, so the lanes of a warp walk the same
// run and their loads of (response, feedback) are broadcasts.
//
// Arithmetic is kept un-contracted (__dmul_rn / __dadd_rn) where the reference's C has a separate
// multiply and add, so Q values follow the reference to the last bit given equal learning rates.
#include "kernels.cuh"

namespace hddm {

// src/wfpt.pyx:123-126: 2.718281828459**a / (1 + 2.718281828459**a) (the reference's truncated e).
__device__ __forceinline__ double rl_rate(double logit)
{
    const double e = pow(2.718281828459, logit);
    return e / (1.0 + e);
}

struct RlLane {
    double q0, q1;        // Q values: [0] lower bound, [1] upper bound
    double rate, rate_pos;
    bool dual;
};

__device__ __forceinline__ void rl_update(RlLane& s, int resp, double fb, double rate_neg)
{
    // src/wfpt.pyx:145-154: learning rate by the sign of the prediction error, then the delta rule
    const double qr = resp ? s.q1 : s.q0;
    const double alfa = (s.dual && fb > qr) ? s.rate_pos : rate_neg;
    const double nq = __dadd_rn(qr, __dmul_rn(alfa, __dadd_rn(fb, -qr)));
    if (resp) s.q1 = nq; else s.q0 = nq;
}

__global__ void __launch_bounds__(128) k_rl_scan(const RlScanArgs a)
{
    const int64_t id = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= a.n_sets * a.n_runs) return;
    const int64_t p = id % a.n_sets, run = id / a.n_sets;
    const int64_t lo = a.run_off[run], hi = a.run_off[run + 1];
    const double q = a.q[p * a.q_stride];
    RlLane s;
    s.q0 = q;
    s.q1 = q;
    const bool alpha_per_trial = a.alpha_trial_stride != 0;
    double rate_neg = alpha_per_trial ? 0.0 : rl_rate(a.alpha[p * a.alpha_set_stride]);
    s.dual = false;
    s.rate_pos = rate_neg;
    if (a.pos_alpha) {
        const double pa = a.pos_alpha[p * a.pos_stride];
        s.dual = true;
        s.rate_pos = (pa == 100.00) ? rate_neg : rl_rate(pa);  // src/wfpt.pyx:106-109
    }
    double* out = a.drift + p * a.n;
    for (int64_t i = lo; i < hi; ++i) {
        const double v = a.v[p * a.v_set_stride + i * a.v_trial_stride];
        out[a.pos ? a.pos[i] : i] = __dmul_rn(__dadd_rn(s.q1, -s.q0), v);  // (qs[1] - qs[0]) * v, src/wfpt.pyx:135
        if (alpha_per_trial) rate_neg = rl_rate(a.alpha[p * a.alpha_set_stride + i * a.alpha_trial_stride]);
        rl_update(s, a.response[i], a.feedback[i], rate_neg);
    }
}

cudaError_t launch_rl_scan(const RlScanArgs& a, cudaStream_t stream)
{
    const int64_t threads = a.n_sets * a.n_runs;
    if (threads <= 0) return cudaSuccess;
    k_rl_scan<<<(unsigned)((threads + 127) / 128), 128, 0, stream>>>(a);
    return cudaGetLastError();
}

// drift[p][i] = sum_k design[k][i] * coef[k][p]; grid (n / 256, P), coalesced in i.
__global__ void __launch_bounds__(256) k_design_drift(const double* __restrict__ design,
                                                      const double* __restrict__ coef, int64_t n, int64_t K,
                                                      int64_t P, double* __restrict__ drift)
{
    const int64_t p = blockIdx.y;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double acc = 0.0;
        for (int64_t k = 0; k < K; ++k) acc = fma(__ldg(design + k * n + i), __ldg(coef + k * P + p), acc);
        drift[p * n + i] = acc;
    }
}

cudaError_t launch_design_drift(const double* design, const double* coef, int64_t n, int64_t K, int64_t P,
                                double* drift, cudaStream_t stream)
{
    if (n <= 0 || P <= 0) return cudaSuccess;
    if (P > 65535) return cudaErrorInvalidValue;
    int64_t bx = (n + 255) / 256;
    if (bx > 1024) bx = 1024;
    k_design_drift<<<dim3((unsigned)bx, (unsigned)P), 256, 0, stream>>>(design, coef, n, K, P, drift);
    return cudaGetLastError();
}

// wiener_like_rl: per (set, run) sum of log P(response), -inf when a mixed probability is exactly 0.
__global__ void __launch_bounds__(128) k_rl_choice(const RlChoiceArgs c)
{
    const RlScanArgs& a = c.scan;
    const int64_t id = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= a.n_sets * a.n_runs) return;
    const int64_t p = id % a.n_sets, run = id / a.n_sets;
    const int64_t lo = a.run_off[run], hi = a.run_off[run + 1];
    const double q = a.q[p * a.q_stride], v = a.v[p * a.v_set_stride], z = c.z[p * c.z_stride];
    const double p_out = c.p_outlier[p * c.pout_stride], w_out = c.w_outlier[p * c.wout_stride];
    const double mix = __dmul_rn(w_out, p_out);
    RlLane s;
    s.q0 = q;
    s.q1 = q;
    const double rate_neg = rl_rate(a.alpha[p * a.alpha_set_stride]);
    const double pa = a.pos_alpha[p * a.pos_stride];
    s.dual = true;
    s.rate_pos = (pa == 100.00) ? rate_neg : rl_rate(pa);
    double total = 0.0;
    bool zero = false;
    for (int64_t i = lo; i < hi; ++i) {
        const int resp = a.response[i];
        if (i != lo) {  // src/wfpt.pyx:203-228: the first trial of a condition only updates Q
            const double drift = __dmul_rn(__dadd_rn(s.q1, -s.q0), v);
            double pr;
            if (drift == 0) {
                pr = 0.5;
            } else {
                const double up = (pow(2.718281828459, -2 * z * drift) - 1) / (pow(2.718281828459, -2 * drift) - 1);
                pr = resp == 1 ? up : 1 - up;
            }
            pr = __dadd_rn(__dmul_rn(pr, 1 - p_out), mix);
            zero |= (pr == 0);
            total += log(pr);
        }
        rl_update(s, resp, a.feedback[i], rate_neg);
    }
    c.partial[p * a.n_runs + run] = zero ? -INFINITY : total;
}

__global__ void __launch_bounds__(128) k_rl_choice_final(const RlChoiceArgs c)
{
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= c.scan.n_sets) return;
    const double p_out = c.p_outlier[p * c.pout_stride];
    double s = 0.0;
    bool neg_inf = !((p_out >= 0) & (p_out <= 1));  // src/wfpt.pyx:176-177
    for (int64_t r = 0; r < c.scan.n_runs; ++r) {    // conditions in np.unique order, as the reference sums
        const double v = c.partial[p * c.scan.n_runs + r];
        neg_inf |= (v == -INFINITY);
        s += v;
    }
    c.logp[p] = neg_inf ? -INFINITY : s;
}

cudaError_t launch_rl_choice(const RlChoiceArgs& c, cudaStream_t stream)
{
    if (c.scan.n_sets <= 0) return cudaSuccess;
    const int64_t threads = c.scan.n_sets * c.scan.n_runs;
    if (threads > 0) k_rl_choice<<<(unsigned)((threads + 127) / 128), 128, 0, stream>>>(c);
    k_rl_choice_final<<<(unsigned)((c.scan.n_sets + 127) / 128), 128, 0, stream>>>(c);
    return cudaGetLastError();
}

}  // namespace hddm
