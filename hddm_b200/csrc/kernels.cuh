// kernels.cuh -- launch interfaces shared by the kernel translation units and the C-ABI (abi.cu).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>


namespace hddm {

constexpr int kMaxDepth = 24;  // max n_st / n_sz for the adaptive explicit stack (HDDM_MAX_DEPTH)

struct Numerics {
    double err;
    double simps_err;
    int n_st;
    int n_sz;
    int use_adaptive;
};

// Navarro-Fuss series selection as a function of normalised time tt for ONE value of err
// (src/pdf.pxi:36-54,60).  The reference recomputes two logs and four square roots per pdf_sv
// call only to arrive at a (branch, K) pair that is piecewise constant in tt; the host finds the
// break points once per err by evaluating the reference's own formulas (series_table.cu) and
// the kernels look the pair up: region r = #{i : tt >= thr[i]}, code[r] = +K (small-time
// series, K terms) or -K (large-time series, K terms).  n_thr < 0: no table (too many regions
// or K out of range) and the device evaluates the formulas directly.
constexpr int kMaxRegions = 32;
struct SeriesTable {
    double thr[kMaxRegions - 1];
    short code[kMaxRegions];
    int n_thr;
    int max_small_k;  // largest K of any small-time region
    int max_large_k;  // largest K of any large-time region
    double err;
};
// Host: cached per err; thread-safe.
const SeriesTable& series_table_for(double err);

// Parameter order used everywhere: v, sv, a, z, sz, t, st, p_outlier, w_outlier.
constexpr int kNumPar = 9;
constexpr int kTrialsPerBlock = 2048;  // one CTA reduces this many trials of one parameter set ...
// Build-time tuning knobs (python -m hddm_b200.build with HDDM_B200_NVCC_EXTRA="-DHDDM_BIG_BLOCK=..." and
// HDDM_B200_LIB=<other file> builds a variant to measure beside the default; profiles/r1_v7_summary.md).
#ifndef HDDM_BIG_BLOCK
#define HDDM_BIG_BLOCK 8192
#endif
#ifndef HDDM_UNROLL_J
#define HDDM_UNROLL_J 5
#endif
constexpr int kTrialsPerBigBlock = HDDM_BIG_BLOCK;  // ... or this many when the grid is large enough anyway (launch_like)
#ifndef HDDM_DEFER_SPAN
#define HDDM_DEFER_SPAN 1
#endif
constexpr int kBlocksPerDeferredCta = HDDM_DEFER_SPAN;  // second launch of a split sweep: blocks of trials per CTA (one table build)
constexpr int kUnrollJ = HDDM_UNROLL_J;  // first pass: the five decision-time nodes of a trial are unrolled (5) or looped (1)
#ifndef HDDM_PAIR_LOADS
#define HDDM_PAIR_LOADS 0
#endif
constexpr bool kPairLoads = HDDM_PAIR_LOADS != 0;  // first pass: adjacent trials per thread, one 16-byte RT load per two rounds
#ifndef HDDM_PDF_ILP
#define HDDM_PDF_ILP 2
#endif
constexpr int kPdfIlp = HDDM_PDF_ILP;  // k_pdf_array, one decision time per point: points per thread in flight
#ifndef HDDM_LIKE_THREADS
#define HDDM_LIKE_THREADS 256
#endif
constexpr int kLikeThreads = HDDM_LIKE_THREADS;  // >= 256 (build_gridset gives threads 0..255 a table entry each), multiple of 32

enum LikeMode : int {
    kModeWienerLike = 0,  // src/wfpt.pyx:54-76   outlier mix, p==0 -> -inf
    kModeMulti = 1,       // src/wfpt.pyx:244-274 +-999 sentinels, no zero guard
    kModeContaminant = 2, // src/wfpt.pyx:357-386 skip cont_x != 0, p==0 -> -inf, no mix
    kModeMultiRL = 3      // src/wfpt.pyx:277-320 as kModeMulti but every x is an RT (no +-999 sentinels)
};

// One launch evaluates `n_sets` parameter sets.  Parameter k of (set p, trial i) is read from
// par[k][p * set_stride[k] + i * trial_stride[k]] (i = absolute index into x):
//   scalar set                -> both strides 0
//   batch of sets (SoA [9][P]) -> set_stride 1, trial_stride 0
//   regression (per trial)    -> trial_stride 1 (set_stride = n for a batch of regressions)
// Set p sees trials [seg[s], seg[s+1]) of x with s = set_node ? set_node[p] : p, or all n trials when
// seg == nullptr.
// Regression term (HDDMRegressor: hddm/models/hddm_regression.py:26-36 regresses any of v, sv, a, z, sz, t, st): the
// per-trial value of parameter `outcome` is link(sum_k design[k][i] * coef[k][p]), k ascending, fused multiply-adds,
// evaluated inside the likelihood kernel (K numbers of the design row per trial instead of a [P][n] table).
constexpr int kMaxRegTerms = 4;
constexpr int kMaxRegK = 16;
enum RegLink : int { kLinkIdentity = 0, kLinkInvLogit = 1, kLinkExp = 2 };
struct RegTerm {
    const double* design;  // [K][n], rows in the order of the trials of x
    const double* coef;    // [K][n_sets]
    int K, outcome, link, pad;
};

struct LikeArgs {
    const double* x;
    const int64_t* seg;
    const int64_t* set_node;  // optional [n_sets]: set p sees segment set_node[p] (many sets per segment)
    const int32_t* cont;  // trials with cont[i] != 0 are skipped: contaminant flags (mode 2), the unscored
                          // first trial of each condition (wiener_like_rlddm); or nullptr
    const double* par[kNumPar];
    int64_t set_stride[kNumPar];
    int64_t trial_stride[kNumPar];
    int64_t n;
    int64_t n_sets;
    int64_t max_seg_len;  // longest segment (0: unknown, n is assumed): picks the one- or two-pass kernel
    int chunks;  // ceil(max segment length / kTrialsPerBlock): CTAs per set as the caller sized `partial`
    int trials_per_block;  // set by launch_like (callers leave 0)
    int mode;
    Numerics nm;
    SeriesTable tab;
    double* partial;             // [n_sets * chunks]
    double* logp;                // [n_sets]
    unsigned long long* evals;   // device counter of pdf_sv evaluations (may be nullptr)
    // Split two-pass sweep (set by launch_like): pass 1 leaves one bit per trial of its CTA's block
    // ("needs the complete evaluator"), a second launch scores those trials and writes partial2.
    unsigned int* defer_bits = nullptr;  // [n_sets * chunks][trials_per_block / 32]
    double* partial2 = nullptr;          // [n_sets * chunks2]
    int chunks2 = 0;                     // ceil(chunks / kBlocksPerDeferredCta)
    const void* set_tables = nullptr;    // [n_sets] GridSet tables built once per set by k_build_tables (else per CTA)
    int all_simple = 0;                  // the host saw every set's sz and st below the 1e-3 snap (src/pdf.pxi:122-125): k_like_simple
    int n_reg = 0;                       // regression terms (0: none); the outcomes' par[] entries are then ignored
    RegTerm reg[kMaxRegTerms] = {};
};

cudaError_t launch_like(const LikeArgs& a, cudaStream_t stream);
// Allocation-free variant (hddm_wiener_like_batch_static): same kernels, the per-CTA partial sums in a file-scope
// __device__ array of kStaticPartialDoubles doubles instead of caller / pool memory; a.partial is ignored.
// Returns cudaErrorInvalidValue when n_sets * chunks exceeds the array.
constexpr int64_t kStaticPartialDoubles = 1 << 18;  // 2 MiB per GPU that loads the library
cudaError_t launch_like_static(const LikeArgs& a, cudaStream_t stream);
extern int g_kernel_choice;  // 0 auto, 1 generic evaluator only, 2 no grid evaluator, 3 grid evaluator in one pass only,
                             // 4 grid evaluator with both passes in one launch

// par_sets == nullptr: one parameter set (par) over x[n] -> y[n].  Otherwise n_sets parameter sets, SoA
// [9][n_sets]; set s (blockIdx.y) reads x + s * x_set_stride (0: one grid shared by all sets) and writes
// y + s * n: the density tables of many posterior draws in one launch.
struct PdfArgs {
    const double* x;
    double* y;
    int64_t n;
    double par[kNumPar];
    int logp;
    Numerics nm;
    SeriesTable tab;
    unsigned long long* evals;
    const double* par_sets = nullptr;
    int64_t n_sets = 1;
    int64_t x_set_stride = 0;
};
cudaError_t launch_pdf_array(const PdfArgs& a, cudaStream_t stream);

// cdfdif (src/cdfdif.c:59-221 + src/cdfdif_wrapper.pyx:36-51): y[i] for signed RT x[i].
// par_sets != nullptr: n_sets parameter sets (SoA [9][n_sets]), x and y are [n_sets][n]; sets whose
// ok[s] == 0 (out of support, flagged by the host) are skipped and their y left untouched.
struct CdfArgs {
    const double* x;
    double* y;
    int64_t n;
    double par[kNumPar];
    const double* par_sets = nullptr;
    const int32_t* ok = nullptr;
    int64_t n_sets = 1;
};
cudaError_t launch_cdf_array(const CdfArgs& a, cudaStream_t stream);

// In-place table post-processing for gen_cdf_using_pdf / gen_rts_from_cdf (src/wfpt.pyx:333-353,
// :405-412): one CTA scans a table of m densities.
cudaError_t launch_cumtrapz_normalise(double* table, int64_t m, cudaStream_t stream);
cudaError_t launch_cumsum_normalise(double* table, int64_t m, cudaStream_t stream);
cudaError_t launch_sample_rts(const double* xgrid, const double* table, int64_t m, const double* f,
                              const double* u_delay, int64_t samples, double t, double st,
                              double* rts, cudaStream_t stream);

// Posterior-predictive batch (ppc_kernels.cu): D draws share the grid; tables is [D][m] densities in,
// rectangle-rule running sums out (src/wfpt.pyx:333-336 per draw) with last[d] = the final entry: the lookup
// below divides the entries it probes by it (:337-340).
cudaError_t launch_cumsum_batch(double* tables, int64_t m, int64_t D, double* last, cudaStream_t stream);
// rts[d][s] by inverse-CDF lookup (src/wfpt.pyx:344-353); t / st per draw; uniforms either supplied
// (f, u_delay: [D][S]) or, when f == nullptr, drawn on the device from Philox(seed, draw0 + d, s).
cudaError_t launch_sample_rts_batch(const double* xgrid, const double* tables, int64_t m, const double* f,
                                    const double* u_delay, int64_t S, int64_t D, const double* t,
                                    const double* st, unsigned long long seed, int64_t draw0,
                                    const double* last, double* rts, cudaStream_t stream);
// Summary statistics of hddm/utils.py:241-267 (gen_ppc_stats) per draw: stats[d][15] = accuracy,
// mean_ub, std_ub, 10/30/50/70/90 q_ub, mean_lb, std_lb, 10/30/50/70/90 q_lb.  S <= kPpcMaxSamples.
constexpr int kPpcStats = 15;
constexpr int64_t kPpcMaxSamples = 16384;
cudaError_t launch_ppc_stats(const double* rts, int64_t S, int64_t D, double* stats, cudaStream_t stream);

// Reinforcement-learning likelihoods (rl_kernels.cu; src/wfpt.pyx:79-241,277-320).  The Q-learning
// recursion depends only on (response, feedback, learning rates), never on the densities: it runs as a
// sequential scan per (parameter set, run of trials) that writes the per-trial drift, after which the
// densities of all trials are evaluated in parallel by the ordinary likelihood kernel.
struct RlScanArgs {
    const int32_t* response;   // [n] 0 = lower, 1 = upper (in run order)
    const double* feedback;    // [n]
    const int64_t* run_off;    // [n_runs + 1]: Q values restart at the head of every run
    int64_t n, n_runs, n_sets;
    const double* q;           // per set: set_stride 0 or 1
    const double* alpha;       // logit learning rate: per set, or per trial (alpha_trial_stride 1)
    const double* pos_alpha;   // nullptr: single learning rate; 100.0 means "same as alpha" (src/wfpt.pyx:106)
    const double* v;           // drift scale: per set or per trial
    int64_t q_stride, alpha_set_stride, alpha_trial_stride, pos_stride, v_set_stride, v_trial_stride;
    double* drift;             // out [n_sets][n]: v * (Q_upper - Q_lower) BEFORE the trial's update
    const int32_t* pos = nullptr;  // optional [n]: trial i's drift goes to column pos[i] (the likelihood's |rt| layout)
};
cudaError_t launch_rl_scan(const RlScanArgs& a, cudaStream_t stream);

// Regression drift built on the device: drift[p][i] = sum_k design[k][i] * coef[k][p] (k ascending, fused
// multiply-adds), the per-trial `v` column of HDDMRegressor (hddm/models/hddm_regression.py:26-36) for P
// coefficient proposals at once, so that only K x (n + P) numbers cross the host-device link instead of P x n.
cudaError_t launch_design_drift(const double* design, const double* coef, int64_t n, int64_t K, int64_t P,
                                double* drift, cudaStream_t stream);

// wiener_like_rl (src/wfpt.pyx:158-241): choice-only likelihood; the same scan, scoring each trial but
// the first of a run with the closed-form P(response), summed per set in run order.
struct RlChoiceArgs {
    RlScanArgs scan;           // drift unused
    const double* z;           // per set
    const double* p_outlier;
    const double* w_outlier;
    int64_t z_stride, pout_stride, wout_stride;
    double* partial;           // [n_sets][n_runs]
    double* logp;              // [n_sets]
};
cudaError_t launch_rl_choice(const RlChoiceArgs& a, cudaStream_t stream);

// Random-walk simulator (sim_kernels.cu; hddm/generate.py:208-319): rts[d][s] for n_draws parameter sets
// (SoA [7][n_draws]: v, sv, a, z, sz, t, st) x `samples` trials each.  u_init / u_steps == nullptr: Philox
// keyed by (seed, (draw0 + d) * samples + s).  Otherwise (parity mode) u_init is [n_draws*samples][3] =
// delay uniform, start-point uniform, standard-normal drift deviate, and u_steps is
// [n_draws*samples][max_steps] step uniforms.  A walk that has not crossed after max_steps gives NaN.
// switch_par != nullptr: the v_switch variant (hddm/generate.py:234-241,272-276,283-285), SoA [3][n_draws] =
// v_switch, V_switch, t_switch; u_init then has a fourth column, the standard-normal deviate of the switched drift.
struct SimArgs {
    const double* params;
    const double* switch_par = nullptr;
    int init_cols = 3;
    int64_t n_draws, samples, max_steps, draw0;
    double dt, intra_sv;
    unsigned long long seed;
    const double* u_init;
    const double* u_steps;
    double* rts;
};
cudaError_t launch_simulate_drift(const SimArgs& a, cudaStream_t stream);

double fp64_peak_probe(int iters, float* ms_out, cudaStream_t stream);

// Trials ordered by |rt| (sort_kernels.cu): a one-time layout step so that the lanes of a warp take the
// same series branch.  n_seg > 0: within each segment [d_seg[s], d_seg[s+1]).
size_t sort_trials_scratch_bytes(int64_t n, int64_t n_seg);
cudaError_t sort_trials(const double* d_x, double* d_sorted, int64_t n, const int64_t* d_seg, int64_t n_seg,
                        void* d_scratch, size_t scratch_bytes, cudaStream_t stream);
constexpr int64_t kSortMin = 64;  // shorter trial vectors are evaluated in the caller's order
extern int g_sort_trials;         // 1 (default): order trials by |rt| where the entry point allows it

}  // namespace hddm
