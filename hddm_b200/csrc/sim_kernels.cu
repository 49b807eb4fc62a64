// sim_kernels.cu -- random-walk simulator of the full DDM, hddm/generate.py:208-319
// (_gen_rts_from_simulated_drift), one thread per simulated trial.
//
// Per trial the reference draws a non-decision time uniform in t +- st/2 (:244-248), a start point
// uniform in (z +- sz/2) a (:251-255), a drift normal(v, sv) (:266-269), then walks +-sqrt(dt)*intra_sv
// steps, up with probability (1 + sqrt(dt)/intra_sv * drift)/2 (:278,286-288), until the position leaves
// [0, a]; the crossing time is interpolated linearly inside the last step (:304-315) and signed by the
// boundary.  The position is accumulated step by step in FP64 exactly as np.cumsum does, so that with
// supplied uniforms (parity mode) the crossing step and the RT equal a numpy run of the reference loop.
// In production mode every trial owns a Philox4x32-10 counter sequence: 4 steps per generator call,
// each a 32-bit integer compared with the up-probability scaled to 2^32.
// v_switch variant (:234-241,272-276,283-285): the reference draws its steps in blocks of
// nn = round(t_switch / dt) and from the second block on walks with the switched drift
// normal(v_switch, V_switch), i.e. steps 0..nn-1 use the first up-probability, all later ones the second.
#include <curand_kernel.h>

#include "kernels.cuh"

namespace hddm {

__global__ void __launch_bounds__(128) k_simulate_drift(const SimArgs a)
{
    const int64_t id = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= a.n_draws * a.samples) return;
    const int64_t d = id / a.samples;
    const double v = a.params[0 * a.n_draws + d], sv = a.params[1 * a.n_draws + d];
    const double bound = a.params[2 * a.n_draws + d], z = a.params[3 * a.n_draws + d];
    const double sz = a.params[4 * a.n_draws + d], t = a.params[5 * a.n_draws + d];
    const double st = a.params[6 * a.n_draws + d];

    const bool sw = a.switch_par != nullptr;
    const double v_sw = sw ? a.switch_par[0 * a.n_draws + d] : 0.0, sv_sw = sw ? a.switch_par[1 * a.n_draws + d] : 0.0;
    // steps before the switch: int(round(t_switch / dt)), round-half-even like Python's round (:236-239)
    const int64_t n_first = sw ? (int64_t)rint(a.switch_par[2 * a.n_draws + d] / a.dt) : a.max_steps;

    curandStatePhilox4_32_10_t rng;
    double u_delay, u_start, n_drift, n_drift_sw = 0.0;
    if (a.u_init) {
        const double* u = a.u_init + id * a.init_cols;
        u_delay = u[0];
        u_start = u[1];
        n_drift = u[2];
        if (sw) n_drift_sw = u[3];
    } else {
        curand_init(a.seed, (unsigned long long)(a.draw0 * a.samples + id), 0, &rng);
        const double2 u = curand_uniform2_double(&rng);
        u_delay = 1.0 - u.x;  // [0, 1)
        u_start = 1.0 - u.y;
        const double2 g = curand_normal2_double(&rng);
        n_drift = g.x;
        n_drift_sw = g.y;
    }
    // scipy uniform.rvs(loc, scale) = loc + scale*u; norm.rvs(loc, scale) = loc + scale*n
    const double delay = __dadd_rn(__dadd_rn(t, __dmul_rn(st, u_delay)), -(st / 2.));
    double y = __dmul_rn(__dadd_rn(__dadd_rn(z, __dmul_rn(sz, u_start)), -(sz / 2.)), bound);
    const double drift = sv != 0 ? __dadd_rn(v, __dmul_rn(sv, n_drift)) : v;
    const double root_dt = sqrt(a.dt);
    const double step = __dmul_rn(root_dt, a.intra_sv);
    const double prob_up = __dmul_rn(0.5, __dadd_rn(1.0, __dmul_rn(root_dt / a.intra_sv, drift)));
    const double drift_sw = sv_sw != 0 ? __dadd_rn(v_sw, __dmul_rn(sv_sw, n_drift_sw)) : v_sw;  // :272-276
    const double prob_up_sw = sw ? __dmul_rn(0.5, __dadd_rn(1.0, __dmul_rn(root_dt / a.intra_sv, drift_sw))) : prob_up;

    int64_t k = 0;       // 0-based index of the step being taken
    double prev = y;
    bool crossed = false;
    if (a.u_steps) {
        const double* u = a.u_steps + id * a.max_steps;
        for (; k < a.max_steps; ++k) {
            prev = y;
            y = __dadd_rn(y, (u[k] < (k < n_first ? prob_up : prob_up_sw)) ? step : -step);
            if (y < 0 || y > bound) {
                crossed = true;
                break;
            }
        }
    } else {
        // P(r < thr) with r uniform on 0..2^32-1: thr = prob_up * 2^32 clamped to [0, 2^32]
        auto threshold = [](double p) {
            const double scaled = p * 4294967296.0;
            return scaled <= 0 ? 0ull : (scaled >= 4294967296.0 ? 4294967296ull : (unsigned long long)scaled);
        };
        const unsigned long long thr_first = threshold(prob_up), thr_sw = threshold(prob_up_sw);
        while (k < a.max_steps && !crossed) {
            const uint4 r = curand4(&rng);
            const unsigned int w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (!crossed && k < a.max_steps) {
                    prev = y;
                    y = __dadd_rn(y, ((unsigned long long)w[j] < (k < n_first ? thr_first : thr_sw)) ? step : -step);
                    if (y < 0 || y > bound) crossed = true; else ++k;
                }
            }
        }
    }
    double rt = NAN;  // no crossing within max_steps
    if (crossed) {
        // hddm/generate.py:304-315: slope through the last two positions, intercept at step index k
        const double m = __dadd_rn(y, -prev) / a.dt;
        const double b = __dadd_rn(y, -__dmul_rn(__dmul_rn(m, (double)k), a.dt));
        const double cross_t = (y < 0) ? __dadd_rn(0.0, -b) / m : __dadd_rn(bound, -b) / m;
        const double sgn = (double)((y > 0) - (y < 0));
        rt = __dmul_rn(__dadd_rn(cross_t, delay), sgn);
    }
    a.rts[id] = rt;
}

cudaError_t launch_simulate_drift(const SimArgs& a, cudaStream_t stream)
{
    const int64_t threads = a.n_draws * a.samples;
    if (threads <= 0) return cudaSuccess;
    k_simulate_drift<<<(unsigned)((threads + 127) / 128), 128, 0, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace hddm
