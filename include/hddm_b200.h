/*
 * hddm_b200.h -- C ABI of libhddm_b200.so: the B200 (sm_100a) implementation of HDDM's Wiener
 * first-passage-time hot path.  This is the drop-in boundary: each entry point below replaces
 * one Python-visible function of the reference's two native extension modules
 *     wfpt            (/root/reference/src/wfpt.pyx, built by setup.py:6)
 *     cdfdif_wrapper  (/root/reference/src/cdfdif_wrapper.pyx + src/cdfdif.c, setup.py:7)
 * with plain pointers and sizes (no Python, numpy or torch types).  INTEGRATION.md shows the
 * ctypes binding a maintainer adds on the reference side; hddm_b200/wfpt.py is that binding.
 *
 * Conventions
 *   - all arithmetic is FP64; RTs are signed (negative = lower boundary, hddm/utils.py:15-37);
 *   - parameter order is the reference's: v, sv, a, z, sz, t, st, then the numerics
 *     err, n_st, n_sz, use_adaptive, simps_err, then p_outlier, w_outlier;
 *   - every function returns an int status (HDDM_OK = 0); nothing throws across the ABI;
 *     hddm_b200_last_error() describes the last non-zero status of the calling thread;
 *   - "host" entry points take host pointers, copy to the current CUDA device, run the kernels
 *     on the library's stream and block until the result is back in the caller's buffer;
 *   - "_dev" entry points take device pointers and a cudaStream_t (as void*), enqueue only;
 *   - there is NO CPU fallback: without a usable CUDA device every call returns HDDM_ERR_CUDA.
 */
#ifndef HDDM_B200_H
#define HDDM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HDDM_OK 0
#define HDDM_ERR_CUDA 1      /* CUDA runtime / no device: Python side raises RuntimeError        */
#define HDDM_ERR_ARG 2       /* malformed argument (null pointer, negative size, depth > 24)      */
#define HDDM_ERR_SUPPORT 3   /* parameters out of support: reference raises ValueError            */
                             /*   (src/wfpt.pyx:395-398, src/cdfdif_wrapper.pyx:23-25)            */
#define HDDM_ERR_RT_RANGE 4  /* max|x| >= 1/(2 w_outlier) with p_outlier > 0: AssertionError      */
                             /*   (src/cdfdif_wrapper.pyx:20-21)                                  */

#define HDDM_MAX_DEPTH 24    /* largest n_st / n_sz accepted when use_adaptive != 0              */

/* ---- runtime ------------------------------------------------------------------------------ */
int hddm_b200_version(void);
const char *hddm_b200_last_error(void);
/* Bind the calling process to CUDA device `device` (-1: keep the current one) and create the
 * library stream.  Called lazily by every other entry point.  ONE PROCESS PER GPU: the stream, the
 * device arena and the kernels' attributes belong to the device bound here; a later init with another
 * device, or any call made while another device is current, returns HDDM_ERR_ARG. */
int hddm_b200_init(int device);
int hddm_b200_device_count(int *count);
int hddm_b200_synchronize(void);
/* FP64 roofline denominator measured on this device: achieved FLOP/s (2 per FMA) of a
 * register-resident dependent-FMA probe, and its duration in ms. */
int hddm_b200_fp64_peak(int iters, double *flops, float *ms);
/* Evaluator selection for the likelihood kernels: 0 = per launch (the grid evaluator when every
 * parameter but the drift is uniform per set and integration is adaptive; else the table-driven walk
 * when z and sz are uniform per set; else the generic one; the grid evaluator runs in two passes when
 * segments are longer than 512 trials, and as two launches when the sweep has thousands of 8192-trial
 * blocks), 1 = always the generic evaluator, 2 = never the grid evaluator, 3 = grid evaluator in one pass
 * only, 4 = its two passes always in one launch.  All run on the GPU; this exists for A/B parity tests
 * and profiling, acts on every caller in the process, and is therefore refused (HDDM_ERR_ARG) unless the
 * environment holds HDDM_B200_DEBUG_KNOBS=1. */
int hddm_b200_set_kernel(int which);
/* Trial layout: 1 (default) = entry points whose result is a sum over trials with set-uniform
 * parameters (wiener_like, wiener_like_batch, the trials handle) evaluate the trials ordered by |rt|
 * (within each segment), so the lanes of a warp share the series branch; 0 = caller's order.  The sum
 * is the same up to rounding and deterministic either way.  Debug knob: needs HDDM_B200_DEBUG_KNOBS=1. */
int hddm_b200_set_sort(int on);
/* The (branch, K) decision of ftt_01w (src/pdf.pxi:36-54,60) as the piecewise-constant function of
 * normalised time the kernels look up: region r = #{i < n_thr[0] : tt >= thr[i]}, code[r] = +K
 * (small-time series, K terms) or -K (large-time).  thr holds >= 31 doubles, code >= 32 ints.
 * n_thr[0] = -1 when no table exists for this err (the kernels then evaluate the formulas).  Pure
 * host computation: needs no CUDA device. */
int hddm_series_table(double err, double *thr, int *code, int *n_thr);

/* ---- wfpt module (src/wfpt.pyx) ----------------------------------------------------------- */

/* full_pdf, src/pdf.pxi:104-146 (cpdef, Python-visible).  out[0] = density. */
int hddm_full_pdf(double x, double v, double sv, double a, double z, double sz, double t, double st,
                  double err, int n_st, int n_sz, int use_adaptive, double simps_err, double *out);

/* pdf_array, src/wfpt.pyx:32-48.  y[i] = full_pdf(x[i]) * (1-p_outlier) + w_outlier*p_outlier,
 * log taken when logp == 1. */
int hddm_pdf_array(const double *x, int64_t n, double v, double sv, double a, double z, double sz,
                   double t, double st, double err, int logp, int n_st, int n_sz, int use_adaptive,
                   double simps_err, double p_outlier, double w_outlier, double *y);

/* wiener_like, src/wfpt.pyx:54-76.  logp[0] = sum of log mixed densities; -inf when any mixed
 * density is exactly 0 or p_outlier is outside [0,1]; 0.0 for n == 0. */
int hddm_wiener_like(const double *x, int64_t n, double v, double sv, double a, double z, double sz,
                     double t, double st, double err, int n_st, int n_sz, int use_adaptive,
                     double simps_err, double p_outlier, double w_outlier, double *logp);

/* wiener_like_multi, src/wfpt.pyx:244-274.  par[k] / stride[k] for k = v, sv, a, z, sz, t, st:
 * stride 0 broadcasts par[k][0], stride 1 reads one value per trial (the `multi` names).
 * x == +999 / -999 are missing-response sentinels (prob_ub / 1 - prob_ub, no outlier mix);
 * no zero-density guard (log(0) = -inf flows into the sum). */
int hddm_wiener_like_multi(const double *x, int64_t n, const double *const par[7],
                           const int64_t stride[7], double err, int n_st, int n_sz, int use_adaptive,
                           double simps_err, double p_outlier, double w_outlier, double *logp);

/* wiener_like_contaminant, src/wfpt.pyx:357-386.  cont_x is int32[n]. */
int hddm_wiener_like_contaminant(const double *x, const int32_t *cont_x, int64_t n, double v,
                                 double sv, double a, double z, double sz, double t, double st,
                                 double t_min, double t_max, double err, int n_st, int n_sz,
                                 int use_adaptive, double simps_err, double *logp);

/* gen_cdf_using_pdf, src/wfpt.pyx:389-414, on the caller's grid xgrid = linspace(-time, time,
 * 2N+1).  cdf[0] keeps the raw density, cdf[1:] = cumtrapz, all divided by cdf[m-1].
 * HDDM_ERR_SUPPORT mirrors the reference's ValueError. */
int hddm_gen_cdf_using_pdf(const double *xgrid, int64_t m, double v, double sv, double a, double z,
                           double sz, double t, double st, double err, int n_st, int n_sz,
                           int use_adaptive, double simps_err, double p_outlier, double w_outlier,
                           double *cdf);

/* gen_rts_from_cdf, src/wfpt.pyx:323-354, on the caller's grid xgrid = arange(lb, ub, dt) with the
 * caller's uniforms: f[samples], and u_delay[samples] when st != 0 (the reference draws them from
 * numpy's global RNG in that order).  table (optional, m doubles) receives the normalised table. */
int hddm_gen_rts_from_cdf(const double *xgrid, int64_t m, double v, double sv, double a, double z,
                          double sz, double t, double st, const double *f, const double *u_delay,
                          int64_t samples, double *rts, double *table);

/* ---- cdfdif_wrapper module (src/cdfdif_wrapper.pyx, src/cdfdif.c) -------------------------- */

/* dmat_cdf_array, src/cdfdif_wrapper.pyx:16-52 -> cdfdif, src/cdfdif.c:59-221. */
int hddm_dmat_cdf_array(const double *x, int64_t n, double v, double sv, double a, double z,
                        double sz, double t, double st, double p_outlier, double w_outlier, double *y);

/* ---- NEW: batched likelihood (no reference counterpart; north_star's batched entry point) --- */

/* P parameter sets in one launch, each with wiener_like semantics (src/wfpt.pyx:54-76).
 * params is SoA float64 [9][P]: rows v, sv, a, z, sz, t, st, p_outlier, w_outlier.
 * seg == NULL: every set is evaluated against all n trials; otherwise seg is int64[P+1] and set p
 * sees x[seg[p] .. seg[p+1]) (ragged: subjects x their own trials x many proposals).
 * evals (optional) receives the number of pdf_sv evaluations the launch made. */
int hddm_wiener_like_batch(const double *x, int64_t n, const int64_t *seg, const double *params,
                           int64_t P, double err, int n_st, int n_sz, int use_adaptive,
                           double simps_err, double *logp, unsigned long long *evals);

/* Same with per-trial drift (regression layout): v is float64[P][n] (row p = set p), the other
 * eight rows of params as above (row 0 of params is ignored).  wiener_like_multi semantics
 * (src/wfpt.pyx:244-274) per set, against all n trials. */
int hddm_wiener_like_multi_batch(const double *x, int64_t n, const double *v_trial,
                                 const double *params, int64_t P, double err, int n_st, int n_sz,
                                 int use_adaptive, double simps_err, double *logp,
                                 unsigned long long *evals);

/* ---- NEW: posterior-predictive batch (many posterior draws per call) ------------------------- */

/* gen_rts_from_cdf (src/wfpt.pyx:323-354) for D posterior draws on one shared grid xgrid[m]
 * (= arange(cdf_lb, cdf_ub, dt)): the per-draw loop of kabuki's post_pred_gen over Wfpt.random
 * (hddm/likelihoods.py:76-81 -> hddm/generate.py:190-194).  params is SoA float64 [7][D]: rows v, sv, a,
 * z, sz, t, st.  f and u_delay are the caller's uniforms, [D][samples] (u_delay may be NULL when every
 * st == 0).  rts (optional) receives [D][samples]; stats (optional) receives [D][15]: the statistics of
 * hddm/utils.py:241-267 (gen_ppc_stats) per simulated data set -- accuracy, mean_ub, std_ub, 10/30/50/
 * 70/90 q_ub, mean_lb, std_lb, 10/30/50/70/90 q_lb (scipy scoreatpercentile semantics; NaN where a
 * boundary has no responses; mean_lb is of the signed RTs, as in the reference).  With stats,
 * 1 <= samples <= 16384. */
int hddm_gen_rts_from_cdf_batch(const double *xgrid, int64_t m, const double *params, int64_t D,
                                const double *f, const double *u_delay, int64_t samples, double *rts,
                                double *stats);
/* Same with the uniforms drawn on the device: Philox4x32-10 keyed by seed, one counter sequence per
 * (global draw index, sample), where column d of params is global draw first_draw + d -- reproducible and
 * independent of slab size and of how the draws are sharded over GPUs (rank r passes the first index of
 * its block).  Not numpy's stream. */
int hddm_gen_rts_from_cdf_batch_seeded(const double *xgrid, int64_t m, const double *params, int64_t D,
                                       unsigned long long seed, int64_t first_draw, int64_t samples,
                                       double *rts, double *stats);
/* The statistics alone, for D data sets of `samples` signed RTs each ([D][samples] -> [D][15]); used for
 * the observed data of a posterior predictive check. */
int hddm_ppc_stats(const double *rts, int64_t D, int64_t samples, double *stats);
/* dmat_cdf_array (src/cdfdif_wrapper.pyx:16-52) for D parameter sets in one launch: x and y are
 * [D][n] (set d's own quantile edges, hddm/likelihoods.py:198-203), params is SoA float64 [9][D]
 * (v, sv, a, z, sz, t, st, p_outlier, w_outlier).  status[d] = HDDM_OK, HDDM_ERR_SUPPORT or
 * HDDM_ERR_RT_RANGE per set (where the reference raises ValueError / AssertionError); rows of y whose
 * status is not HDDM_OK are NaN. */
int hddm_dmat_cdf_batch(const double *x, int64_t n, const double *params, int64_t D, double *y,
                        int32_t *status);

/* Random-walk simulator, hddm/generate.py:208-319 (_gen_rts_from_simulated_drift; gen_rts(method='drift'),
 * used by HDDMRegressor's random(), hddm/models/hddm_regression.py:62): D parameter sets (SoA float64
 * [7][D]: v, sv, a, z, sz, t, st) x `samples` trials, one GPU thread per trial; rts is [D][samples] signed
 * RTs.  dt = step length in seconds (reference default 1e-4), intra_sv = within-trial noise (1.0).
 * u_init == NULL and u_steps == NULL: device Philox4x32-10 keyed by (seed, global trial index).
 * Otherwise both must be given (parity mode): u_init [D*samples][3] = delay uniform, start-point uniform,
 * standard-normal drift deviate; u_steps [D*samples][max_steps] = one uniform per step.  A walk that
 * has not left [0, a] after max_steps steps yields NaN. */
int hddm_simulate_drift_batch(const double *params, int64_t D, int64_t samples, double dt, double intra_sv,
                              unsigned long long seed, const double *u_init, const double *u_steps,
                              int64_t max_steps, double *rts);
/* The v_switch variant of the same function (hddm/generate.py:234-241,272-276,283-285): switch_par is SoA
 * float64 [3][D] = v_switch, V_switch, t_switch.  The reference walks in blocks of nn = round(t_switch / dt)
 * steps and uses the switched drift normal(v_switch, V_switch) from its second block on, i.e. from step nn.
 * Parity mode: u_init is [D*samples][4], the fourth column the standard-normal deviate of the switched
 * drift.  HDDM_ERR_ARG when round(t_switch / dt) < 1 (the reference fails on an empty block). */
int hddm_simulate_drift_switch_batch(const double *params, const double *switch_par, int64_t D, int64_t samples,
                                     double dt, double intra_sv, unsigned long long seed, const double *u_init,
                                     const double *u_steps, int64_t max_steps, double *rts);

/* ---- reinforcement-learning likelihoods (src/wfpt.pyx:79-241, 277-320) -------------------------- */

/* wiener_like_rlddm, src/wfpt.pyx:79-155.  response (0 = lower, 1 = upper) and split_by (condition
 * label) are int64[n] like the reference's np.ndarray[long]; feedback is float64[n].  Conditions are
 * visited in ascending label order, each with its trials in the caller's order; Q values restart at q
 * per condition; the first trial of a condition updates Q but is not scored; drift of a scored trial =
 * (Q_upper - Q_lower) * v; learning rate = logistic(alpha), or logistic(pos_alpha) after a positive
 * prediction error unless pos_alpha == 100.0.  -inf as hddm_wiener_like.  0.0 for n == 0. */
int hddm_wiener_like_rlddm(const double *x, const int64_t *response, const double *feedback,
                           const int64_t *split_by, int64_t n, double q, double alpha, double pos_alpha,
                           double v, double sv, double a, double z, double sz, double t, double st,
                           double err, int n_st, int n_sz, int use_adaptive, double simps_err,
                           double p_outlier, double w_outlier, double *logp);
/* NEW: P parameter sets against one subject's task in one launch pair.  params is SoA float64 [12][P]:
 * rows v, sv, a, z, sz, t, st, p_outlier, w_outlier, q, alpha, pos_alpha. */
int hddm_wiener_like_rlddm_batch(const double *x, const int64_t *response, const double *feedback,
                                 const int64_t *split_by, int64_t n, const double *params, int64_t P,
                                 double err, int n_st, int n_sz, int use_adaptive, double simps_err,
                                 double *logp, unsigned long long *evals);
/* wiener_like_rl, src/wfpt.pyx:158-241: choice-only likelihood (the reference accepts and ignores
 * err, n_st, n_sz, use_adaptive, simps_err; they are not part of this signature). */
int hddm_wiener_like_rl(const int64_t *response, const double *feedback, const int64_t *split_by,
                        int64_t n, double q, double alpha, double pos_alpha, double v, double z,
                        double p_outlier, double w_outlier, double *logp);
/* NEW: params SoA float64 [7][P]: rows q, alpha, pos_alpha, v, z, p_outlier, w_outlier. */
int hddm_wiener_like_rl_batch(const int64_t *response, const double *feedback, const int64_t *split_by,
                              int64_t n, const double *params, int64_t P, double *logp);
/* wiener_like_multi_rlddm, src/wfpt.pyx:277-320.  par[k] / stride[k] for k = v, sv, a, z, sz, t, st,
 * alpha (stride 0 scalar, 1 per trial).  Trials in the caller's order; Q values restart where split_by
 * differs from the previous trial; every trial is scored; single learning rate; no zero-density guard
 * and no +-999 sentinels. */
int hddm_wiener_like_multi_rlddm(const double *x, const int64_t *response, const double *feedback,
                                 const int64_t *split_by, int64_t n, double q, const double *const par[8],
                                 const int64_t stride[8], double err, int n_st, int n_sz, int use_adaptive,
                                 double simps_err, double p_outlier, double w_outlier, double *logp);

/* NEW: regression likelihood for many coefficient proposals (HDDMRegressor; hddm/models/hddm_regression.py:26-36 ->
 * wiener_like_multi, src/wfpt.pyx:244-274).  The reference regresses any of v, sv, a, z, sz, t, st and hands the
 * likelihood one per-trial column per outcome; here the linear predictors are evaluated INSIDE the likelihood
 * kernel: for regression term r, the value of parameter outcome[r] (0..6 = v sv a z sz t st) of trial i under
 * proposal p is link[r](sum_k design_r[k][i] * coef_r[k][p]), k ascending, with link 0 = identity, 1 = inverse
 * logit 1 / (1 + exp(-x)), 2 = exp.  No [P][n] table exists anywhere: no limit on P, K_r (n + P) numbers cross the
 * host-device link.  hddm_regdata_upload keeps the trials and the design matrices (design[r] is float64 [K[r]][n],
 * 1 <= K[r] <= 16, at most 4 terms) resident, laid out once by boundary and |rt|; hddm_regdata_like_batch then takes
 * the coefficients (coef[r] float64 [K[r]][P]) and the remaining parameters (params [9][P]; rows of regressed outcomes
 * are ignored) and returns wiener_like_multi's value per proposal (+-999 sentinels and all).  A regression on v alone
 * runs in the lean / grid evaluators; other outcomes in the table-driven or generic ones. */
typedef struct hddm_regdata hddm_regdata; /* opaque */
int hddm_regdata_upload(const double *x, int64_t n, int n_terms, const int64_t *K, const double *const *design,
                        hddm_regdata **out);
int hddm_regdata_free(hddm_regdata *h);
int hddm_regdata_like_batch(hddm_regdata *h, const int *outcome, const int *link, const double *const *coef,
                            const double *params, int64_t P, double err, int n_st, int n_sz, int use_adaptive,
                            double simps_err, double *logp, unsigned long long *evals);
/* One-shot form for a regression on v with the identity link: design float64 [K][n], coef float64 [K][P]. */
int hddm_wiener_like_regression_batch(const double *x, int64_t n, const double *design, int64_t K,
                                      const double *coef, const double *params, int64_t P, double err, int n_st,
                                      int n_sz, int use_adaptive, double simps_err, double *logp,
                                      unsigned long long *evals);

/* ---- device-resident trials (one upload per observed node, reused every MCMC step) --------- */
typedef struct hddm_trials hddm_trials; /* opaque */
int hddm_trials_upload(const double *x, int64_t n, hddm_trials **out);
int hddm_trials_free(hddm_trials *h);
int hddm_trials_wiener_like(const hddm_trials *h, double v, double sv, double a, double z, double sz,
                            double t, double st, double err, int n_st, int n_sz, int use_adaptive,
                            double simps_err, double p_outlier, double w_outlier, double *logp);
int hddm_trials_wiener_like_batch(const hddm_trials *h, const int64_t *seg, const double *params,
                                  int64_t P, double err, int n_st, int n_sz, int use_adaptive,
                                  double simps_err, double *logp, unsigned long long *evals);

/* Hierarchical layout: the observed nodes (subject x condition) are the segments
 * [node_off[s], node_off[s+1]) of x, s < n_nodes, and parameter set p is evaluated against node
 * set_node[p] -- many proposals / chains per node and all children of a group-level parameter in ONE
 * launch (the per-node logp of hddm/likelihoods.py:52-55, batched).  logp[p] as hddm_wiener_like. */
int hddm_wiener_like_nodes(const double *x, int64_t n, const int64_t *node_off, int64_t n_nodes,
                           const int64_t *set_node, const double *params, int64_t P, double err,
                           int n_st, int n_sz, int use_adaptive, double simps_err, double *logp,
                           unsigned long long *evals);
int hddm_trials_wiener_like_nodes(const hddm_trials *h, const int64_t *node_off, int64_t n_nodes,
                                  const int64_t *set_node, const double *params, int64_t P, double err,
                                  int n_st, int n_sz, int use_adaptive, double simps_err, double *logp,
                                  unsigned long long *evals);

/* NEW: native slice-within-Gibbs driver for the hierarchical full DDM of BASELINE config 2 (S subjects, one observed
 * node each, include = sv sz st; priors, slice widths and maxiter of hddm/models/hddm_info.py:103-175,
 * hddm/models/base.py:578-681), C chains in lock-step, no Python between the launches of a sweep.  `trials` holds all
 * subjects' signed RTs, subject s = [node_off[s], node_off[s+1]).  g is float64 [9][C] (rows v, v_std, a, a_std, t,
 * t_std, sv, sz, st) and subj float64 [3][C][S] (rows v, a, t): the chains' state, in and out.  rng_state = numpy
 * PCG64 {state hi, state lo, inc hi, inc lo}, in and out: the chain drawn equals the one hddm_b200/sampler.py
 * (HierarchicalDDM, the readable specification) draws from the same generator state.  trace_g [n_iter][9][C] and
 * trace_subj [n_iter][3][C][S] may be NULL.  n_launches (may be NULL) counts likelihood launches. */
int hddm_hier_ddm_sample(const hddm_trials *trials, const int64_t *node_off, int64_t S, int64_t C, double *g,
                         double *subj, double z, double p_outlier, double w_outlier, double err, int n_st, int n_sz,
                         int use_adaptive, double simps_err, uint64_t *rng_state, int64_t n_iter, double *trace_g,
                         double *trace_subj, int64_t *n_launches);


/* ---- device-pointer entry points (enqueue only; used under torch.distributed sharding) ----- */

/* One-time layout step for a device-resident trial vector that many launches will reuse: d_sorted =
 * the trials of d_x ordered by |rt| (within each segment [d_seg[s], d_seg[s+1]) when n_seg > 0).
 * hddm_sort_trials_scratch gives the BYTES of d_scratch needed.  d_sorted must not alias d_x. */
int64_t hddm_sort_trials_scratch(int64_t n, int64_t n_seg);
int hddm_sort_trials_dev(const double *d_x, int64_t n, const int64_t *d_seg, int64_t n_seg,
                         double *d_sorted, void *d_scratch, int64_t scratch_bytes, void *stream);

/* Elements (doubles) of scratch the batched launch needs for P sets whose longest segment has
 * max_seg_len trials. */
int64_t hddm_wiener_like_batch_scratch(int64_t max_seg_len, int64_t P);
/* All pointers are device pointers; d_seg may be NULL; d_evals may be NULL (a device counter that
 * is ADDED to).  stream is a cudaStream_t. */
int hddm_wiener_like_batch_dev(const double *d_x, int64_t n, const int64_t *d_seg, int64_t max_seg_len,
                               const double *d_params, int64_t P, double err, int n_st, int n_sz,
                               int use_adaptive, double simps_err, double *d_logp, double *d_scratch,
                               unsigned long long *d_evals, void *stream);
/* No counterpart in the reference.  hddm_wiener_like_batch_dev
 * for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc): same inputs and outputs, no d_scratch argument -- the per-block partial sums go to a static
 * __device__ array inside the library (2 MiB = 262144 doubles, held by every GPU that loads the library until
 * exit).  The call allocates nothing, frees nothing and creates nothing: the library must already be bound to the
 * current device (hddm_b200_init at set-up, else HDDM_ERR_ARG); the first call per device also looks up the
 * array's device address and uploads the 16 KB exponential table, so make one call before a capture.
 * LIMIT: P * ceil(longest segment / 2048) <= 262144, else HDDM_ERR_ARG with the sizes in hddm_b200_last_error().
 * All calls share the one array.  Nothing in it has to start at zero (each slot is written before it is read,
 * the previous call's contents are never read), but two calls that OVERLAP on the GPU (different streams)
 * race on it: keep calls to this entry point on one stream, or order them with events.
 * Sums are taken per 2048-trial block, then over blocks in ascending order: equal to hddm_wiener_like_batch_dev
 * within the summation tolerance (1e-9 relative), not bit for bit (that entry point picks other block sizes). */
int hddm_wiener_like_batch_dev_static(const double *d_x, int64_t n, const int64_t *d_seg, int64_t max_seg_len,
                                      const double *d_params, int64_t P, double err, int n_st, int n_sz,
                                      int use_adaptive, double simps_err, double *d_logp,
                                      unsigned long long *d_evals, void *stream);
int hddm_pdf_array_dev(const double *d_x, int64_t n, double v, double sv, double a, double z,
                       double sz, double t, double st, double err, int logp, int n_st, int n_sz,
                       int use_adaptive, double simps_err, double p_outlier, double w_outlier,
                       double *d_y, unsigned long long *d_evals, void *stream);
int hddm_dmat_cdf_array_dev(const double *d_x, int64_t n, double v, double sv, double a, double z,
                            double sz, double t, double st, double p_outlier, double w_outlier,
                            double *d_y, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* HDDM_B200_H */
