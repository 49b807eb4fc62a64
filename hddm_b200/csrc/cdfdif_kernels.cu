// cdfdif_kernels.cu -- full-DDM cumulative distribution (Tuerlinckx 2004) on the device.
//
// Mirrors the behaviour of src/cdfdif.c:59-221 behind src/cdfdif_wrapper.pyx:16-52.  Layout of
// the work: FIVE signed RTs per warp, six lanes each (the 6 Gauss-Hermite drift nodes), for the boundary
// probability and the branch "past the Ter window", whose Fourier series advance in lock step on the
// index v with the 3-term stopping rule evaluated per group; RTs "inside the window" (36 drift x
// start-point node pairs) are evaluated one at a time by the whole warp.  Node sums are reduced in the
// reference's node order (m ascending, i ascending), so the result does not depend on lane scheduling.
#include "kernels.cuh"

namespace hddm {

namespace {

constexpr double kPiLong = 3.1415926535897932384626433832795028841971693993751;
constexpr double kTiny = 1e-29;     // src/cdfdif.c:67
constexpr double kNearZero = 1e-7;  // :68
constexpr double kMinRt = 0.001;    // :69
constexpr int kMaxTerms = 5000;     // :71

__constant__ double c_gh_x[6] = {-2.3506049736744922818, -1.3358490740136970132, -.43607741192761650950,
                                 .43607741192761650950,  1.3358490740136970132,  2.3506049736744922818};
__constant__ double c_gh_w[6] = {.45300099055088421593e-2, .15706732032114842368, .72462959522439207571,
                                 .72462959522439207571,    .15706732032114842368, .45300099055088421593e-2};
__constant__ double c_gl_x[6] = {-.93246951420315193904, -.66120938646626381541, -.23861918608319693247,
                                 .23861918608319712676,  .66120938646626459256,  .93246951420315160597};
__constant__ double c_gl_w[6] = {.17132449237917049545, .36076157304813916138, .46791393457269092604,
                                 .46791393457269092604, .36076157304813843973, .17132449237917132812};

struct Hist {
    double h0, h1, h2;
    __device__ void push(double term)
    {
        h0 = h1;
        h1 = h2;
        h2 = h1 + term;
    }
    __device__ bool settled() const  // src/cdfdif.c:144-145, :177-178, :202-203
    {
        return (fabs(h0 - h1) < kTiny) && (fabs(h1 - h2) < kTiny) && (h2 > 0);
    }
};

// Sum `val` held by lanes [0, count) in ascending lane order (fixed order => deterministic and
// the reference's node order).
__device__ __forceinline__ double ordered_lane_sum(double val, int count)
{
    double s = 0.0;
    for (int l = 0; l < count; ++l) s += __shfl_sync(0xffffffffu, val, l);
    return s;
}

}  // namespace

// In-window branch of cdfdif (src/cdfdif.c:151-211: the RT falls inside the non-decision-time window) for ONE RT,
// evaluated by a whole warp: the 36 (drift, start point) node pairs are spread over the lanes (pair `lane`, and
// pair 32 + lane for lane < 4); every lane walks the series of its pair; node sums are assembled in the reference's
// order.  All arguments are warp-uniform.
__device__ __noinline__ double cdf_in_window(double a, double z, double sZ, double st, double nu, double eta, double t,
                                             int x, double lower_t, double Z_U, double Z_L, double base)
{
    const int lane = threadIdx.x & 31;
    const double a2 = a * a;
    const double sx = (double)(2 * x - 1);
    double pair_val[2] = {0.0, 0.0};
    bool any_zero_drift = false;
    for (int m = 0; m < 6; ++m) any_zero_drift |= !(fabs(1.41421356237309505 * c_gh_x[m] * eta + nu) > kNearZero);
    for (int round = 0; round < 2; ++round) {
        const int pair = round * 32 + lane;
        if (pair >= 36) break;
        const int m = pair / 6, i6 = pair % 6;
        const double g = 1.41421356237309505 * c_gh_x[m] * eta + nu;
        if (!(fabs(g) > kNearZero)) continue;  // handled by the zero-drift series below
        const double gz = (.5 * sZ * c_gl_x[i6]) + z;
        const double zzz = (a - gz) * x + gz * (1 - x);
        const double f = (double)(1 - 2 * x);
        const double sh = sinh(f * g * a / .01);
        const double ser = -((a * a2) / (f * g * kPiLong * .01)) * sinh(zzz * f * g / .01) / (sh * sh) +
                           (zzz * a2) / (f * g * kPiLong * .01) * cosh((a - zzz) * f * g / .01) / sh;
        Hist s = {0, 0, 0};
        for (int v = 0; v < kMaxTerms; ++v) {
            const double sifa = kPiLong * v / a;
            const double denom = (g * g * 100 + (kPiLong * v) * (kPiLong * v) / (a2 * 100));
            // exp(-.5 denom dt - 2 log(denom)) without the logarithm
            s.push(v * sin(sifa * zzz) * (exp(-.5 * denom * (t - lower_t)) / (denom * denom)));
            if (s.settled()) break;
        }
        pair_val[round] = .5 * c_gl_w[i6] * (ser - 4 * s.h2) * (kPiLong / 100) / (a2 * st) * exp(sx * zzz * g * 100);
    }
    double zero_branch = 0.0;
    if (any_zero_drift) {  // :185-207 (warp-uniform; every lane computes the same value)
        Hist s = {0, 0, 0};
        const double su = -(Z_U * Z_U) / (12 * a2) + (Z_U * Z_U * Z_U) / (12 * a * a2) -
                          (Z_U * Z_U * Z_U * Z_U) / (48 * a2 * a2);
        const double sl = -(Z_L * Z_L) / (12 * a2) + (Z_L * Z_L * Z_L) / (12 * a * a2) -
                          (Z_L * Z_L * Z_L * Z_L) / (48 * a2 * a2);
        for (int v = 1; v < kMaxTerms; ++v) {
            const double sifa = kPiLong * v / a;
            const double denom = (kPiLong * v) * (kPiLong * v) / (a2 * 100);
            const double vd = (double)v;
            s.push(1 / (kPiLong * kPiLong * kPiLong * kPiLong * vd * vd * vd * vd) *
                   (cos(sifa * Z_L) - cos(sifa * Z_U)) * exp(-.5 * denom * (t - lower_t)));
            if (s.settled()) break;
        }
        zero_branch = 400 * a2 * a * (sl - su - s.h2) / (st * sZ);
    }
    __syncwarp();
    // assemble in the reference's order: for m: sum_z = sum_i pair(m,i); sum_nu += sum_z*w_gh[m]
    double sum_nu = 0.0;
    for (int m = 0; m < 6; ++m) {
        const double g = 1.41421356237309505 * c_gh_x[m] * eta + nu;
        double sum_z = 0.0;
        if (fabs(g) > kNearZero) {
            for (int i6 = 0; i6 < 6; ++i6) {
                const int pair = m * 6 + i6;
                const double v0 = __shfl_sync(0xffffffffu, pair_val[0], pair & 31);
                const double v1 = __shfl_sync(0xffffffffu, pair_val[1], pair & 31);
                sum_z += pair >= 32 ? v1 : v0;
            }
        } else {
            sum_z = zero_branch;
        }
        sum_nu += sum_z * (c_gh_w[m] / 1.772453850905515882);
    }
    return base - sum_nu;
}

// cdfdif(|x|, x>0, par) and the wrapper's signed assembly + outlier mix.  A warp takes FIVE signed RTs: lanes
// 6 g .. 6 g + 5 own the six Gauss-Hermite drift nodes of RT g (lanes 30, 31 idle), so that the boundary probability
// (src/cdfdif.c:96-113) and the Fourier series past the non-decision-time window (:121-150) use 30 of 32 lanes
// instead of 6; the five series advance in lock step on the index v, each group stopping by its own 3-term rule
// (:144-146), the loop ending when every group has settled.  Node sums are taken in the reference's node order.
// RTs inside the window (:151-211, 36 node pairs each) are then evaluated one at a time by the whole warp.
// Per term the reference's exp(A - 3 log(denom) + log(w) - 2 log 100) is E_A w / (1e4 denom^3) with E_A = exp(A)
// hoisted, and exp(-.5 denom dt + log(1 - exp(-.5 denom W))) is exp(-.5 denom dt) (1 - exp(-.5 denom W)): two
// exponentials and no logarithm per term instead of four and two.
__global__ void __launch_bounds__(128) k_cdf_array(const CdfArgs q)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp_id = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t total = q.n * q.n_sets;
    const int64_t first = warp_id * 5;
    if (first >= total) return;
    const int grp = lane < 30 ? lane / 6 : 4;
    const int m6 = lane % 6;
    const int lead = grp * 6;
    const bool owner = lane < 30;
    int64_t i = first + grp;
    const bool have = owner && i < total;
    if (i >= total) i = total - 1;  // (idle lanes shadow a valid RT; their results are dropped)
    const int64_t set = q.par_sets ? i / q.n : 0;
    const bool ok = have && !(q.ok && q.ok[set] == 0);
    double par[kNumPar];
#pragma unroll
    for (int k = 0; k < kNumPar; ++k) par[k] = q.par_sets ? __ldg(q.par_sets + k * q.n_sets + set) : q.par[k];

    // wrapper rescale to DMAT units, src/cdfdif_wrapper.pyx:36-42 (par: v sv a z sz t st po wo)
    const double a = par[2] / 10.;
    const double Ter = par[5];
    const double eta = par[1] / 10. + 1e-10;
    const double z = par[3] * (par[2] / 10.);
    const double sZ = par[4] * (par[2] / 10.) + 1e-10;
    const double st = par[6] + 1e-10;
    const double nu = par[0] / 10.;

    const double xs = q.x[i];
    const double t = fabs(xs);
    const int x = xs > 0 ? 1 : 0;
    const double sx = (double)(2 * x - 1);  // +1 upper, -1 lower
    const double a2 = a * a;
    const double zc = (1 - x) * z + x * (a - z);
    const double Z_U = zc + sZ / 2, Z_L = zc - sZ / 2;
    const double lower_t = Ter - st / 2;

    const double gk = 1.41421356237309505 * c_gh_x[m6] * eta + nu;  // src/cdfdif.c:86
    const double wgh = c_gh_w[m6] / 1.772453850905515882;           // :88
    // sum over the six drift nodes of this lane's group, in node order
    auto group_sum = [&](double val) {
        double s = 0.0;
#pragma unroll
        for (int l = 0; l < 6; ++l) s += __shfl_sync(0xffffffffu, val, lead + l);
        return s;
    };

    // ---- probability of the boundary, src/cdfdif.c:96-113
    double prob = 0.0;
    for (int i6 = 0; i6 < 6; ++i6) {
        const double gz = (.5 * sZ * c_gl_x[i6]) + z;  // :92
        double term;
        if (fabs(gk) > kNearZero)
            term = (exp(-200 * gz * gk) - 1) / (exp(-200 * a * gk) - 1) * wgh;
        else
            term = gz / a * wgh;
        prob += group_sum(term) * c_gl_w[i6] / 2;
    }

    const bool in_range = ok && (t - Ter + st / 2 > kMinRt);  // :116
    const double upper_t = t < Ter + st / 2 ? t : Ter + st / 2;
    const double p1 = prob * (upper_t - lower_t) / st;
    const double p0 = (1 - prob) * (upper_t - lower_t) / st;
    const double base = p0 * (1 - x) + p1 * x;
    const bool past = in_range && (t > Ter + st / 2);
    double F = 0.0;
    if (__any_sync(0xffffffffu, past)) {  // :121-150
        Hist s = {0, 0, 0};
        bool running = past;
        const double EU = exp(sx * Z_U * gk * 100), EL = exp(sx * Z_L * gk * 100);
        const double cw = wgh / 1e4;
        const double dt = t - upper_t, W = upper_t - lower_t;
        for (int v = 0; v < kMaxTerms && __any_sync(0xffffffffu, running); ++v) {
            const double sifa = kPiLong * v / a;
            double su, cu, sl, cl;
            sincos(sifa * Z_U, &su, &cu);
            sincos(sifa * Z_L, &sl, &cl);
            const double denom = (100 * gk * gk + (kPiLong * kPiLong) * (v * v) / (100 * a2));
            const double w3 = cw / (denom * denom * denom);
            const double fact = (EU * w3) * (sx * gk * su * 100 - sifa * cu) - (EL * w3) * (sx * gk * sl * 100 - sifa * cl);
            const double exdif = exp(-.5 * denom * dt) * (1 - exp(-.5 * denom * W));
            const double sum_nu = group_sum(fact * exdif);
            if (running) {
                s.push(v * sum_nu);
                if (s.settled()) running = false;
            }
        }
        if (past) F = base - s.h2 * 4 * kPiLong / (a2 * sZ * st);
    }
    // ---- RTs inside the window: one at a time, the whole warp on its 36 node pairs
    const bool inwin = in_range && !past;
    unsigned int todo = __ballot_sync(0xffffffffu, inwin && lane == lead);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const double Fw = cdf_in_window(__shfl_sync(0xffffffffu, a, src), __shfl_sync(0xffffffffu, z, src),
                                        __shfl_sync(0xffffffffu, sZ, src), __shfl_sync(0xffffffffu, st, src),
                                        __shfl_sync(0xffffffffu, nu, src), __shfl_sync(0xffffffffu, eta, src),
                                        __shfl_sync(0xffffffffu, t, src), __shfl_sync(0xffffffffu, x, src),
                                        __shfl_sync(0xffffffffu, lower_t, src), __shfl_sync(0xffffffffu, Z_U, src),
                                        __shfl_sync(0xffffffffu, Z_L, src), __shfl_sync(0xffffffffu, base, src));
        if (lane == src) F = Fw;
    }
    F = F > kTiny ? F : 0;  // :218

    if (ok && lane == lead) {  // src/cdfdif_wrapper.pyx:46-51
        const double sgn = (double)((xs > 0) - (xs < 0));
        double y = (1 - prob) + sgn * F;
        const double po = par[7], wo = par[8];
        y = y * (1 - po) + (xs + (1. / (2 * wo))) * wo * po;
        q.y[i] = y;
    }
}

cudaError_t launch_cdf_array(const CdfArgs& a, cudaStream_t stream)
{
    const int64_t total = a.n * (a.par_sets ? a.n_sets : 1);
    if (total <= 0) return cudaSuccess;
    CdfArgs b = a;
    if (!b.par_sets) b.n_sets = 1;
    const int64_t warps = (total + 4) / 5;
    const int warps_per_block = 4;
    k_cdf_array<<<(unsigned)((warps + warps_per_block - 1) / warps_per_block), 128, 0, stream>>>(b);
    return cudaGetLastError();
}

}  // namespace hddm
