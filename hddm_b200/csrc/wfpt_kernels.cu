// wfpt_kernels.cu -- likelihood kernels: one thread per (trial, parameter set).
//
// k_like_partial   grid (n_sets, chunks): CTA (p, c) evaluates trials
//                  [lo_p + c*2048, lo_p + (c+1)*2048) of set p, thread j takes trials
//                  j, j+256, ... (coalesced 8-byte loads of x and of per-trial parameters),
//                  sums log-densities in that order, then a fixed xor-shuffle tree per warp and
//                  a fixed-order sum over the 8 warps -> partial[p*chunks + c].
// k_like_final     one warp per set adds the chunk partials in ascending order (lane-strided,
//                  then the same shuffle tree) and applies the reference's -inf rule.
// No atomics on floating point anywhere: results are bit-reproducible run to run and independent
// of how sets are sharded over GPUs.
#include "kernels.cuh"
#include "wfpt_device.cuh"
#include "wfpt_fast.cuh"
#include "wfpt_grid.cuh"
#include "wfpt_lean.cuh"

#include <cmath>
#include <mutex>
#include <vector>

namespace hddm {

int g_sort_trials = 1;
int g_kernel_choice = 0;  // 0: pick per launch; 1: always the generic evaluator; 2: never the grid evaluator

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

__device__ __forceinline__ double load_par(const LikeArgs& a, int k, int64_t p, int64_t i)
{
    return __ldg(a.par[k] + p * a.set_stride[k] + i * a.trial_stride[k]);
}

// KMAX > 0: the table-driven evaluator of wfpt_fast.cuh (z, sz uniform per set, adaptive
// integration, large-time series of at most KMAX terms).  KMAX == 0: the generic evaluator of
// wfpt_device.cuh (any per-trial parameter, fixed-grid Simpson, no table).
// GRID: additionally sv, a, t, st uniform per set -> the v3 grid evaluator of wfpt_grid.cuh.
// TWO_PASS (GRID only): see the trial loop.  Pays off when a thread owns several trials (long segments);
// short observed nodes (one round per thread) take the one-pass flavour, whose critical path is shorter.
// PASSES (GRID only): 0 one pass; 1 two passes in one launch; 2 / 3 the two passes as separate launches
// (kSplitFirst leaves a bit per deferred trial in a.defer_bits, kSplitSecond scores those trials).
enum : int { kOnePass = 0, kTwoPass = 1, kSplitFirst = 2, kSplitSecond = 3 };
#ifndef HDDM_SECOND_MIN_BLOCKS
#define HDDM_SECOND_MIN_BLOCKS 2
#endif
#ifndef HDDM_FIRST_MIN_BLOCKS
#define HDDM_FIRST_MIN_BLOCKS 2
#endif
template <int KMAX, bool GRID, int PASSES = kOnePass>
__global__ void __launch_bounds__(kLikeThreads, KMAX > 0 ? (PASSES == kSplitSecond ? HDDM_SECOND_MIN_BLOCKS
                                                             : (PASSES == kSplitFirst ? HDDM_FIRST_MIN_BLOCKS : 2)) : 1)
    k_like_partial(const __grid_constant__ LikeArgs a)
{
    constexpr bool TWO_PASS = PASSES != kOnePass;
    constexpr bool LEAN = GRID && KMAX == 4 && (PASSES == kTwoPass || PASSES == kSplitFirst);
    constexpr bool TREE = GRID && KMAX == 4 && (PASSES == kTwoPass || PASSES == kSplitSecond);  // deferred trials: lean_tree
    constexpr int KT = KMAX > 0 ? KMAX : 1;
    __shared__ GridSet<KT> gs;
    WTable<KT>* wt = gs.wt;
    const int64_t p = blockIdx.x;
    const int chunk = blockIdx.y;
    const int64_t node = a.set_node ? a.set_node[p] : p;
    const int64_t lo = a.seg ? a.seg[node] : 0;
    const int64_t hi = a.seg ? a.seg[node + 1] : a.n;
    // kSplitSecond: the CTA covers kBlocksPerDeferredCta consecutive blocks of the set's trials
    constexpr int kSpan = PASSES == kSplitSecond ? kBlocksPerDeferredCta : 1;
    const int64_t begin = lo + (int64_t)chunk * kSpan * a.trials_per_block;
    double* const slot = PASSES == kSplitSecond ? a.partial2 + p * a.chunks2 + chunk : a.partial + p * a.chunks + chunk;
    const bool single = PASSES < kSplitFirst && a.chunks == 1;  // this CTA's sum IS the set's log-likelihood
    const double p_out = load_par(a, 7, p, 0), w_out = load_par(a, 8, p, 0);
    // src/wfpt.pyx:50-51,63-64: p_outlier outside [0,1] -> -inf before any trial is looked at
    const bool multi = a.mode == kModeMulti || a.mode == kModeMultiRL;  // no range / zero guards
    const bool range_ok = multi || ((p_out >= 0) & (p_out <= 1));
    if (begin >= hi || !range_ok) {  // (ragged batch: shorter sets have empty trailing chunks)
        if (threadIdx.x == 0) {
            // (kSplitSecond adds nothing: the first launch's partial already carries the -inf)
            *slot = (range_ok || PASSES == kSplitSecond) ? 0.0 : -INFINITY;
            if (single) a.logp[p] = range_ok ? 0.0 : -INFINITY;  // single chunk: no k_like_final
        }
        return;
    }
    // Layout of a block's trials over (round, thread) in the two-pass flavours: thread t owns the ADJACENT trials
    // 2t and 2t + 1 of every 2 * kLikeThreads tile -- rounds 2k and 2k + 1 -- so that one 16-byte load brings both
    // RTs (coalesced: a warp reads 512 contiguous bytes).  The defer bitmap is indexed by (round, warp, lane).
    // (kPairLoads: measured 6 % slower on the headline sweep -- the pair carried across two rounds costs registers
    //  the first level does not have -- so the default stays one 8-byte load per trial and round, 256 contiguous
    //  bytes per warp; HDDM_PAIR_LOADS=1 builds the paired flavour.)
    auto trial_of = [](int r, int t) { return kPairLoads ? (r >> 1) * (2 * kLikeThreads) + 2 * t + (r & 1) : r * kLikeThreads + t; };
    auto rounds_for = [](int64_t cnt) {
        return kPairLoads ? 2 * (int)((cnt + 2 * kLikeThreads - 1) / (2 * kLikeThreads)) : (int)((cnt + kLikeThreads - 1) / kLikeThreads);
    };
    const int64_t span = (int64_t)kSpan * a.trials_per_block;
    const int64_t end = begin + span < hi ? begin + span : hi;
    const int words_per_block = a.trials_per_block / 32;
    if (PASSES == kSplitSecond) {  // nothing deferred in this CTA's blocks: leave before the table build
        unsigned int any = 0;
#pragma unroll
        for (int c = 0; c < kSpan; ++c) {  // (no early exit: the blocks' loads are independent and overlap)
            const int64_t b0 = begin + (int64_t)c * a.trials_per_block;
            const int64_t left = end - b0;
            const int64_t cnt = left < a.trials_per_block ? left : a.trials_per_block;
            const int words = cnt > 0 ? rounds_for(cnt) * (kLikeThreads / 32) : 0;
            const unsigned int* w = a.defer_bits + (p * a.chunks + (int64_t)chunk * kSpan + c) * words_per_block;
            for (int k = threadIdx.x; k < words; k += kLikeThreads) any |= w[k];
        }
        if (!__syncthreads_or(any != 0)) {
            if (threadIdx.x == 0) *slot = 0.0;
            return;
        }
    }

    // Set-uniform parameters are read once; per-trial ones (trial_stride != 0) inside the loop.
    double par[7];
    bool per_trial = false;
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        par[k] = load_par(a, k, p, 0);
        per_trial |= a.trial_stride[k] != 0;
    }
    // regression terms: this set's coefficients, once per CTA
    __shared__ double s_coef[kMaxRegTerms][kMaxRegK];
    const bool reg_v_only = a.n_reg == 1 && a.reg[0].outcome == 0 && !per_trial;  // the lean evaluators take it
    // ... and a drift COLUMN (one value per trial: the Q-learning scan of the RL-DDM, wiener_like_multi with 'v')
    bool col_v_only = a.n_reg == 0 && a.trial_stride[0] != 0;
#pragma unroll
    for (int k = 1; k < 7; ++k) col_v_only = col_v_only && a.trial_stride[k] == 0;
    const bool v_per_trial = reg_v_only || col_v_only;
    if (a.n_reg > 0) {
        for (int k = threadIdx.x; k < a.n_reg * kMaxRegK; k += kLikeThreads) {
            const RegTerm& t = a.reg[k / kMaxRegK];
            s_coef[k / kMaxRegK][k % kMaxRegK] = (k % kMaxRegK) < t.K ? __ldg(t.coef + (int64_t)(k % kMaxRegK) * a.n_sets + p) : 0.0;
        }
        per_trial = true;
        __syncthreads();
    }
    auto reg_value = [&](int r, int64_t i) {
        const RegTerm& t = a.reg[r];
        double acc = 0.0;
        for (int k = 0; k < t.K; ++k) acc = fma(__ldg(t.design + (int64_t)k * a.n + i), s_coef[r][k], acc);
        if (t.link == kLinkInvLogit) acc = 1.0 / (1.0 + exp(-acc));
        else if (t.link == kLinkExp) acc = exp(acc);
        return acc;
    };
    auto trial_drift = [&](int64_t i) { return reg_v_only ? reg_value(0, i) : load_par(a, 0, p, i); };

    if (GRID && a.set_tables) {  // split sweep: the set's tables were built once, by k_build_tables
        static_assert(sizeof(GridSet<KT>) % sizeof(uint4) == 0, "tables are copied 16 bytes at a time");
        const uint4* src = reinterpret_cast<const uint4*>(a.set_tables) + p * (sizeof(GridSet<KT>) / sizeof(uint4));
        uint4* dst = reinterpret_cast<uint4*>(&gs);
        for (int k = threadIdx.x; k < (int)(sizeof(GridSet<KT>) / sizeof(uint4)); k += kLikeThreads) dst[k] = __ldg(src + k);
        __syncthreads();
    } else if (GRID) {
        build_gridset<KT>(gs, par, a.tab,
                          (a.mode != kModeContaminant && p_out >= 0 && p_out <= 1) ? w_out * p_out : 0.0);
        if (LEAN) {  // (k_build_tables does the same for the shared tables)
            if (threadIdx.x == 0) build_leanset(gs, gs.lean, par[0], a.nm, !per_trial);
            __syncthreads();
        }
    } else if (KMAX > 0)
        build_wtables<KT>(wt, par[3], par[4]);
    // Lean first level (wfpt_lean.cuh): the set qualifies (CTA-uniform) -> stage the 2^(i/2048) table
    __shared__ double ebig[(LEAN || TREE) ? kExpBig : 1];
    // (uniform drift, or a drift regressed per trial; wiener_like / contaminant / wiener_like_multi semantics)
    const bool lean_ok = (LEAN || TREE) && gs.lean.ok != 0 && a.mode != kModeMultiRL &&
                         (v_per_trial ? gs.lean.uniform == 0 : (!per_trial && gs.lean.uniform != 0));
    const bool tree_ok = TREE && lean_ok && a.nm.n_st == 2;
    if (lean_ok) {
        for (int k = threadIdx.x; k < kExpBig; k += kLikeThreads) ebig[k] = g_exp2_big[k];
        __syncthreads();
    }

    // Stages of the two-pass flavours (GRID): 0 = first pass, the first 5 x 5 level only (*need: the trial goes
    // to stage 1); 1 = the complete grid evaluator (*need: out of its range, goes to stage 2); 2 = the v2
    // evaluator.  The one-pass flavours have stage 0 only, and it is complete.
    auto density = [&](double x, unsigned long long& ev, int stage, bool* need) {
        if (GRID && TWO_PASS) {
            if (stage == 0) return full_pdf_grid<KT, 1>(x, par[0], gs, a.nm, &a.tab, ev, need);
            if (stage == 1) return full_pdf_grid_slow<KT>(x, par[0], gs, a.nm, &a.tab, ev, need);
            return grid_fallback<KT>(x, par[0], gs, a.nm, &a.tab, ev);
        }
        if (GRID) return full_pdf_grid<KT, 0>(x, par[0], gs, a.nm, &a.tab, ev);
        if (KMAX > 0)
            return full_pdf_fast<(KMAX > 0 ? KMAX : 1)>(x, par[0], par[1], par[2], par[3], par[4], par[5],
                                                        par[6], a.nm, &a.tab, wt, ev);
        return full_pdf(x, par[0], par[1], par[2], par[3], par[4], par[5], par[6], a.nm, &a.tab, ev);
    };

    // sum of log(d) = log(prod of mantissas) + ln2 * (sum of exponents) for normal positive d: one log
    // per thread instead of one per trial; anything else (0, negative, NaN, inf, subnormal) takes log().
    double acc = 0.0, mant = 1.0;
    int expo = 0;
    auto add_log = [&](double d) {
        const int hi = __double2hiint(d);
        if ((unsigned)(hi - 0x00100000) < 0x7fe00000u) {
            expo += (hi >> 20) - 1023;
            mant *= __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(d));
        } else {
            acc += log(d);
        }
    };
    static_assert(kLikeThreads >= 256 && kLikeThreads % 32 == 0 && kTrialsPerBigBlock / kLikeThreads <= 256,
                  "table build needs 256 threads; mantissa product must stay finite (two passes)");
    unsigned long long evals = 0;
    bool hit_zero = false;
    // src/wfpt.pyx:70, spelled out so that every instantiation of this kernel rounds it the same way
    const double keep_mass = 1 - p_out, floor_mass = __dmul_rn(w_out, p_out);
    // One trial: load, evaluate, mix, accumulate.  Returns true when the trial was deferred.
    auto score = [&](int64_t i, int stage) {
        const double x = __ldg(a.x + i);
        if (per_trial) {
#pragma unroll
            for (int k = 0; k < (GRID ? 1 : 7); ++k)  // GRID: only the drift can vary per trial
                if (a.trial_stride[k] != 0) par[k] = load_par(a, k, p, i);
            for (int r = 0; r < a.n_reg; ++r) par[GRID ? 0 : a.reg[r].outcome] = reg_value(r, i);
        }
        bool need = false;
        double d;
        if (multi) {
            if (a.mode == kModeMultiRL || fabs(x) != 999.) {  // src/wfpt.pyx:261-266
                d = density(x, evals, stage, &need);
                d = fma(d, keep_mass, floor_mass);
            } else if (x == 999.) {  // :267-268
                d = prob_ub(par[0], GRID ? gs.raw[2] : par[2], GRID ? gs.raw[3] : par[3]);
            } else {  // :269-270
                d = 1 - prob_ub(par[0], GRID ? gs.raw[2] : par[2], GRID ? gs.raw[3] : par[3]);
            }
            if (!need) add_log(d);
        } else {
            if (a.cont && a.cont[i] != 0) return false;  // src/wfpt.pyx:374; :134 (first trial of a condition)
            d = density(x, evals, stage, &need);
            if (!need) {
                d = fma(d, keep_mass, floor_mass);  // src/wfpt.pyx:70 (contaminant: p_out = 0)
                hit_zero |= (d == 0);                 // :71-72, :377-379
                add_log(d);
            }
        }
        return need;
    };
    // Lean first level for one warp round (lean_ok sets): *need is set per lane; always returns true.
    auto lean_round = [&](int64_t i, bool in_range, bool* need, double x_loaded) {
        const LeanSet& L = gs.lean;
        const bool valid = in_range && !(a.cont && a.cont[i] != 0);
        const double x = valid ? x_loaded : 1.0;
        const int b = x > 0 ? 1 : 0;
        bool v_fits = true;
        const LeanTrial c = v_per_trial ? lean_begin_v<KT>(gs, L, b, trial_drift(valid ? i : begin), fabs(x) - gs.tn[0], &v_fits)
                                       : lean_begin(L, b);
        bool nd = false;
        int made = 0;
        double d = lean_first_level<KT>(gs, L, ebig, c, x, valid & v_fits, &nd, &made);
        nd |= valid & !v_fits;
        if (valid & !nd) {
            evals += made;
            d = fma(d, keep_mass, floor_mass);
            hit_zero |= (d == 0);
            add_log(d);
        }
        *need = nd;
        return true;
    };
    if (GRID && TWO_PASS) {
        // Pass 1 evaluates the first 5 x 5 level, which settles almost every trial; the few that need more are
        // queued per warp (ballot order: deterministic) and pass 2 runs them through the complete evaluator with
        // full warps.  Keeps the slow paths' code, registers and local-memory arrays out of the hot loop.
        constexpr int kMaxRounds = (kTrialsPerBigBlock + kLikeThreads - 1) / kLikeThreads;
        __shared__ unsigned short queue[kLikeThreads / 32][kMaxRounds * 32];
        static_assert(kTrialsPerBigBlock <= 65536, "queue entries are 16-bit offsets into the CTA's block of trials");
        static_assert(kBlocksPerDeferredCta * kTrialsPerBigBlock <= 65536, "16-bit offsets (second launch)");
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        int queued = 0;  // warp-uniform
        // Stage 1 over the warp's queue, 32 trials at a time; the few that need stage 2 are compacted into the
        // part of the queue already consumed (ballot order again) and run last.  Returns the new fill (0).
        auto run_queue = [&](int n_queued) {
            __syncwarp();
            if (TREE && tree_ok) {
                // Stage "tree" first: the whole outer recursion with lean nodes, lanes in step (wfpt_lean.cuh); what it
                // cannot settle is compacted, as below, for the complete evaluator
                int n_left = 0;
                for (int k0 = 0; k0 < n_queued; k0 += 32) {
                    const int k = k0 + lane;
                    const bool live = k < n_queued;
                    const unsigned short off = live ? queue[warp][k] : 0;
                    const bool on = live && !(a.cont && a.cont[begin + off] != 0);
                    const double x = on ? __ldg(a.x + begin + off) : 1.0;
                    bool fail = false, v_fits = true;
                    int made = 0;
                    const int bb = x > 0 ? 1 : 0;
                    const LeanTrial c = v_per_trial ? lean_begin_v<KT>(gs, gs.lean, bb, trial_drift(begin + (on ? off : 0)),
                                                                      fabs(x) - gs.tn[0], &v_fits)
                                                   : lean_begin(gs.lean, bb);
                    double d = lean_tree<KT>(gs, gs.lean, ebig, x, c, on & v_fits, &fail, &made);
                    fail = (fail | !v_fits) & on;
                    if (on & !fail) {
                        evals += made;
                        d = fma(d, keep_mass, floor_mass);
                        hit_zero |= (d == 0);
                        add_log(d);
                    }
                    const unsigned int votes = __ballot_sync(0xffffffffu, fail);
                    if (fail) queue[warp][n_left + __popc(votes & ((1u << lane) - 1u))] = off;
                    n_left += __popc(votes);
                    __syncwarp();
                }
                n_queued = n_left;
            }
            int n_fallback = 0;
            for (int k0 = 0; k0 < n_queued; k0 += 32) {
                const int k = k0 + lane;
                const unsigned short off = k < n_queued ? queue[warp][k] : 0;
                const bool need = k < n_queued ? score(begin + off, 1) : false;
                const unsigned int votes = __ballot_sync(0xffffffffu, need);
                if (need) queue[warp][n_fallback + __popc(votes & ((1u << lane) - 1u))] = off;
                n_fallback += __popc(votes);
                __syncwarp();
            }
            for (int k = lane; k < n_fallback; k += 32) score(begin + queue[warp][k], 2);
            __syncwarp();
            return 0;
        };
        if (PASSES == kSplitSecond) {
            // Second launch: the deferred trials of this CTA's blocks, compacted per warp in the same (block,
            // round, lane) order as the one-launch flavour queues them, scored whenever a queue is nearly full.
            auto drain = [&]() { queued = run_queue(queued); };
            for (int c = 0; c < kSpan; ++c) {
                const int64_t b0 = begin + (int64_t)c * a.trials_per_block;
                if (b0 >= end) break;
                const int64_t cnt = end - b0 < a.trials_per_block ? end - b0 : a.trials_per_block;
                const int rounds = rounds_for(cnt);
                const unsigned int* w = a.defer_bits + (p * a.chunks + (int64_t)chunk * kSpan + c) * words_per_block;
                // lane r fetches the warp's word of round r: one load per block instead of one per round
                static_assert(kMaxRounds <= 32, "a warp's words of one block fit one per lane");
                const unsigned int mine = lane < rounds ? w[lane * (kLikeThreads / 32) + warp] : 0u;
                for (int r = 0; r < rounds; ++r) {
                    const unsigned int votes = __shfl_sync(0xffffffffu, mine, r);
                    if (votes == 0) continue;  // warp-uniform
                    if ((votes >> lane) & 1u)
                        queue[warp][queued + __popc(votes & ((1u << lane) - 1u))] =
                            (unsigned short)(c * a.trials_per_block + trial_of(r, threadIdx.x));
                    queued += __popc(votes);
                    if (queued > kMaxRounds * 32 - 32) drain();
                }
            }
            drain();
        } else {
            const int rounds = rounds_for(end - begin);
            unsigned int* const bits = PASSES == kSplitFirst
                                           ? a.defer_bits + (p * a.chunks + chunk) * words_per_block : nullptr;
            unsigned int* const my_bits = bits ? bits + warp : nullptr;
            const bool vec = (reinterpret_cast<uintptr_t>(a.x + begin) & 15) == 0;  // CTA-uniform
            double2 xx = make_double2(1.0, 1.0);
            for (int r = 0; r < rounds; ++r) {
                const int off = trial_of(r, threadIdx.x);
                if (!kPairLoads) {
                    if (LEAN && lean_ok) xx.x = begin + off < end ? __ldg(a.x + begin + off) : 1.0;
                } else if (LEAN && lean_ok && !(r & 1)) {  // even round: the thread's pair of RTs
                    if (vec && begin + off + 1 < end) {
                        xx = __ldg(reinterpret_cast<const double2*>(a.x + begin + off));
                    } else {
                        xx.x = begin + off < end ? __ldg(a.x + begin + off) : 1.0;
                        xx.y = begin + off + 1 < end ? __ldg(a.x + begin + off + 1) : 1.0;
                    }
                }
                bool need = false;
                if (!(LEAN && lean_ok && lean_round(begin + off, begin + off < end, &need, (kPairLoads && (r & 1)) ? xx.y : xx.x)))
                    need = (begin + off < end) ? score(begin + off, 0) : false;
                const unsigned int votes = __ballot_sync(0xffffffffu, need);
                if (PASSES == kSplitFirst) {
                    if (lane == 0) my_bits[r * (kLikeThreads / 32)] = votes;
                } else {
                    if (need) queue[warp][queued + __popc(votes & ((1u << lane) - 1u))] = (unsigned short)off;
                    queued += __popc(votes);
                }
            }
            if (PASSES == kTwoPass) run_queue(queued);
        }
    } else {
        for (int64_t i = begin + threadIdx.x; i < end; i += kLikeThreads) score(i, 0);
    }
    acc += fma((double)expo, 0.693147180559945309417, log(mant));

    __shared__ double warp_part[kLikeThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    acc = warp_sum(acc);
    if (lane == 0) warp_part[warp] = acc;
    const int any_zero = __syncthreads_or(hit_zero ? 1 : 0);
    if (a.evals) {
        evals = warp_sum_u64(evals);
        if (lane == 0) atomicAdd(a.evals, evals);
    }
    if (threadIdx.x == 0) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kLikeThreads / 32; ++w) s += warp_part[w];
        if (!multi && any_zero) s = -INFINITY;
        *slot = s;
        if (single) a.logp[p] = s;  // single chunk: this IS the set's log-likelihood (no k_like_final)
    }
}

// Sweeps whose sets ALL have neither sz nor st (the simple DDM of BASELINE config 1, with or without sv; the host
// entry points see the parameters and say so in LikeArgs::all_simple): one pdf_sv per trial through lean_simple, one
// pass, no bitmap and no second launch -- the few trials outside the comfortable range (or a set the lean constants do
// not cover) take the complete evaluator in place.  Same grid, partial sums and reduction order as k_like_partial.
template <int KMAX>
__global__ void __launch_bounds__(kLikeThreads, 2) k_like_simple(const __grid_constant__ LikeArgs a)
{
    __shared__ GridSet<KMAX> gs;
    __shared__ double ebig[kExpBig];
    const int64_t p = blockIdx.x;
    const int chunk = blockIdx.y;
    const int64_t node = a.set_node ? a.set_node[p] : p;
    const int64_t lo = a.seg ? a.seg[node] : 0;
    const int64_t hi = a.seg ? a.seg[node + 1] : a.n;
    const int64_t begin = lo + (int64_t)chunk * a.trials_per_block;
    double* const slot = a.partial + p * a.chunks + chunk;
    const bool single = a.chunks == 1;
    const double p_out = load_par(a, 7, p, 0), w_out = load_par(a, 8, p, 0);
    const bool range_ok = (p_out >= 0) & (p_out <= 1);  // src/wfpt.pyx:63-64
    if (begin >= hi || !range_ok) {
        if (threadIdx.x == 0) {
            *slot = range_ok ? 0.0 : -INFINITY;
            if (single) a.logp[p] = range_ok ? 0.0 : -INFINITY;
        }
        return;
    }
    const int64_t end = begin + a.trials_per_block < hi ? begin + a.trials_per_block : hi;
    double par[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) par[k] = load_par(a, k, p, 0);
    if (a.set_tables) {
        const uint4* src = reinterpret_cast<const uint4*>(a.set_tables) + p * (sizeof(GridSet<KMAX>) / sizeof(uint4));
        uint4* dst = reinterpret_cast<uint4*>(&gs);
        for (int k = threadIdx.x; k < (int)(sizeof(GridSet<KMAX>) / sizeof(uint4)); k += kLikeThreads) dst[k] = __ldg(src + k);
    } else {
        build_gridset<KMAX>(gs, par, a.tab, (a.mode != kModeContaminant) ? w_out * p_out : 0.0);
        if (threadIdx.x == 0) build_leanset(gs, gs.lean, par[0], a.nm, true);
    }
    for (int k = threadIdx.x; k < kExpBig; k += kLikeThreads) ebig[k] = g_exp2_big[k];
    __syncthreads();
    const LeanSet& L = gs.lean;
    const bool simple_ok = (L.ok1 & 4) != 0 && L.uniform != 0;
    const double keep_mass = 1 - p_out, floor_mass = __dmul_rn(w_out, p_out);  // src/wfpt.pyx:70
    double acc = 0.0, mant = 1.0;
    int expo = 0;
    bool hit_zero = false;
    unsigned int made_fast = 0;          // pdf_sv evaluations settled in the loop (<= 32 per thread)
    unsigned long long evals_slow = 0;   // ... and by the complete evaluator (its counter is passed by reference:
                                         // kept apart so that nothing of the loop lives in local memory)
    static_assert(kTrialsPerBigBlock / kLikeThreads <= 256, "mantissa product must stay finite");
    auto add_log = [&](double d) {       // sum of logs = log of the mantissa product + ln2 * exponents
        d = fma(d, keep_mass, floor_mass);
        hit_zero |= (d == 0);            // src/wfpt.pyx:71-72, :377-379
        const int hw = __double2hiint(d);
        if ((unsigned)(hw - 0x00100000) < 0x7fe00000u) {
            expo += (hw >> 20) - 1023;
            mant *= __hiloint2double((hw & 0x000fffff) | 0x3ff00000, __double2loint(d));
        } else {
            acc += log(d);
        }
    };
    // trials the lean evaluator does not take are queued per warp (ballot order) and scored after the loop
    constexpr int kQueue = 256;
    __shared__ unsigned short queue[kLikeThreads / 32][kQueue];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int queued = 0;  // warp-uniform
    auto drain = [&]() {
        __syncwarp();
        for (int k = lane; k < queued; k += 32) {
            const int64_t i = begin + queue[warp][k];
            add_log(full_pdf_grid<KMAX, 0>(__ldg(a.x + i), par[0], gs, a.nm, &a.tab, evals_slow));
        }
        __syncwarp();
        queued = 0;
    };
    const int rounds = (int)((end - begin + kLikeThreads - 1) / kLikeThreads);
#pragma unroll 1
    for (int r = 0; r < rounds; ++r) {
        const int off = r * kLikeThreads + threadIdx.x;
        const int64_t i = begin + off;
        const bool valid = i < end && !(a.cont && a.cont[i] != 0);
        const double x = valid ? __ldg(a.x + i) : 1.0;
        bool need = valid;
        int made = 0;
        double d = 0.0;
        if (simple_ok) d = lean_simple<KMAX>(gs, L, ebig, lean_begin(L, x > 0 ? 1 : 0), x, valid, &need, &made);
        if (valid & !need) {
            made_fast += made;
            add_log(d);
        }
        const unsigned int votes = __ballot_sync(0xffffffffu, need);
        if (votes) {  // warp-uniform, rare
            if (need) queue[warp][queued + __popc(votes & ((1u << lane) - 1u))] = (unsigned short)off;
            queued += __popc(votes);
            if (queued > kQueue - 32) drain();
        }
    }
    drain();
    unsigned long long evals = evals_slow + made_fast;
    acc += fma((double)expo, 0.693147180559945309417, log(mant));
    __shared__ double warp_part[kLikeThreads / 32];
    acc = warp_sum(acc);
    if (lane == 0) warp_part[warp] = acc;
    const int any_zero = __syncthreads_or(hit_zero ? 1 : 0);
    if (a.evals) {
        evals = warp_sum_u64(evals);
        if (lane == 0) atomicAdd(a.evals, evals);
    }
    if (threadIdx.x == 0) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kLikeThreads / 32; ++w) s += warp_part[w];
        if (any_zero) s = -INFINITY;
        *slot = s;
        if (single) a.logp[p] = s;
    }
}

// The per-set tables of the grid evaluator, once per set (split sweeps: every CTA of the set, in both
// launches, copies them instead of rebuilding them).
template <int KMAX>
__global__ void __launch_bounds__(kLikeThreads) k_build_tables(const __grid_constant__ LikeArgs a, void* out)
{
    __shared__ GridSet<KMAX> gs;
    const int64_t p = blockIdx.x;
    double par[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) par[k] = load_par(a, k, p, 0);
    const double p_out = load_par(a, 7, p, 0), w_out = load_par(a, 8, p, 0);
    build_gridset<KMAX>(gs, par, a.tab,
                        (a.mode != kModeContaminant && p_out >= 0 && p_out <= 1) ? w_out * p_out : 0.0);
    if (threadIdx.x == 0) {
        bool uniform = a.n_reg == 0;
        for (int k = 0; k < 7; ++k) uniform &= a.trial_stride[k] == 0;
        if (KMAX == 4) build_leanset(gs, gs.lean, par[0], a.nm, uniform);
        else gs.lean.ok = 0;
    }
    __syncthreads();
    const uint4* src = reinterpret_cast<const uint4*>(&gs);
    uint4* dst = reinterpret_cast<uint4*>(out) + p * (sizeof(GridSet<KMAX>) / sizeof(uint4));
    for (int k = threadIdx.x; k < (int)(sizeof(GridSet<KMAX>) / sizeof(uint4)); k += kLikeThreads) dst[k] = src[k];
}

// The same tables, one WARP per set (a CTA builds kSetsPerCta sets side by side): for launches of many short
// segments, where a CTA per set spends its life in the table build's serial sections (~40 us against ~2 us of
// likelihood for a 200-trial node: profiles/r2_v9_bench_paths.jsonl).  Same arithmetic as k_build_tables.
constexpr int kWarpSets = 4;        // sets per CTA of the warp-per-set kernels
constexpr int kWarpOneLaunchSets = 1024;  // k_like_warp: up to this many sets run all phases in one launch (latency path)
#ifndef HDDM_WARP_CTA_TRIALS
#define HDDM_WARP_CTA_TRIALS 512
#endif
constexpr int kWarpCtaTrials = HDDM_WARP_CTA_TRIALS;  // k_like_warp: segments this long get a CTA per set whatever the number of sets
constexpr int kWarpMaxTrials = 1024;  // longest segment k_like_warp takes (its deferred-trial queue is 16-bit x 1024)
template <int KMAX>
__global__ void __launch_bounds__(32 * kWarpSets) k_build_tables_warp(const __grid_constant__ LikeArgs a, void* out)
{
    __shared__ GridSet<KMAX> gs[kWarpSets];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t p = (int64_t)blockIdx.x * kWarpSets + warp;
    if (p >= a.n_sets) return;
    GridSet<KMAX>& G = gs[warp];
    double par[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) par[k] = load_par(a, k, p, 0);
    const double p_out = load_par(a, 7, p, 0), w_out = load_par(a, 8, p, 0);
    build_gridset<KMAX, WarpGroup>(G, par, a.tab,
                                   (a.mode != kModeContaminant && p_out >= 0 && p_out <= 1) ? w_out * p_out : 0.0, WarpGroup());
    if (lane == 0) {
        if (KMAX == 4) build_leanset(G, G.lean, par[0], a.nm, true);
        else G.lean.ok = 0;
    }
    __syncwarp();
    const uint4* src = reinterpret_cast<const uint4*>(&G);
    uint4* dst = reinterpret_cast<uint4*>(out) + p * (sizeof(GridSet<KMAX>) / sizeof(uint4));
    for (int k = lane; k < (int)(sizeof(GridSet<KMAX>) / sizeof(uint4)); k += 32) dst[k] = src[k];
}

// Short segments: every trial's mixed density goes to a scratch row dens[p][.] -- from the lean first level, or, for
// the trials it defers, from the lean tree, or from the complete one-pass evaluator full_pdf_grid<.., 0> for what the
// tree cannot settle (and for sets the lean evaluators do not take) -- and ONE warp then reduces the row in a fixed
// order (lane l takes trials l, l + 32, ...; sum of logs = log of the lane's mantissa product + ln2 * exponents; fixed
// shuffle tree).  So a set's sum does not depend on who produced the densities: the launch gives a set ONE warp when
// there are many sets (kWarpSets sets per CTA) and the whole CTA when there are few (its four warps take rounds
// round-robin: a quarter of the latency), and a scalar wiener_like call and any batch give a set the same sum, bit
// for bit.  Tables come from `set_tables`.  No partials, no final kernel: logp[p] is written here.
// PHASE 0: everything in one launch (a handful of sets: the latency path of a single chain).  Otherwise two launches,
// so that the warps of an SM are not spread over the lean level, the tree and the complete evaluator at once (the
// one-launch flavour of a 4100-set batch spent 4.7 cycles per issue waiting for instructions,
// profiles/r2_v12_ncu_digest_nodes4096.txt): PHASE 1 runs the lean level, leaves kDeferred in the row where a trial
// needs more, reduces the sets it completed and flags the others in `pending`; PHASE 2 looks only at flagged sets,
// rebuilds each warp's queue from the marks (same trials, same order as PHASE 0) and finishes them.
constexpr double kDeferred = -1.0;  // never a mixed density (those are >= 0 or NaN)
// SIMPLE: the host saw that every set is a simple DDM (LikeArgs::all_simple): the lean level of this instantiation is
// lean_simple and nothing else; the other instantiation does not carry it (its sets without sz and st go to the
// complete evaluator, as every set the lean constants do not cover).
template <int KMAX, int PHASE, bool SIMPLE>
__global__ void __launch_bounds__(32 * kWarpSets, 4) k_like_warp(const __grid_constant__ LikeArgs a, int cta_per_set,
                                                               double* __restrict__ dens_all, int64_t dens_stride,
                                                               int* __restrict__ pending)
{
    __shared__ GridSet<KMAX> gsw[kWarpSets];
    __shared__ double ebig[kExpBig];
    __shared__ unsigned short queue[kWarpSets][kWarpMaxTrials];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t p = cta_per_set ? (int64_t)blockIdx.x : (int64_t)blockIdx.x * kWarpSets + warp;
    const bool live = p < a.n_sets && (PHASE != 2 || pending[p] != 0);
    if (PHASE == 2 && !__syncthreads_or(live)) return;  // nothing left over in this CTA's sets
    for (int k = threadIdx.x; k < kExpBig; k += 32 * kWarpSets) ebig[k] = g_exp2_big[k];
    GridSet<KMAX>& G = gsw[cta_per_set ? 0 : warp];
    if (live) {
        const uint4* src = reinterpret_cast<const uint4*>(a.set_tables) + p * (sizeof(GridSet<KMAX>) / sizeof(uint4));
        uint4* dst = reinterpret_cast<uint4*>(&G);
        const int first = cta_per_set ? threadIdx.x : lane, step = cta_per_set ? 32 * kWarpSets : 32;
        for (int k = first; k < (int)(sizeof(GridSet<KMAX>) / sizeof(uint4)); k += step) dst[k] = __ldg(src + k);
    }
    __syncthreads();
    if (!live) return;
    const LeanSet& L = G.lean;
    const int64_t node = a.set_node ? a.set_node[p] : p;
    const int64_t lo = a.seg ? a.seg[node] : 0;
    const int64_t hi = a.seg ? a.seg[node + 1] : a.n;
    const int n = (int)(hi - lo);
    const double v = load_par(a, 0, p, 0);
    const double p_out = load_par(a, 7, p, 0), w_out = load_par(a, 8, p, 0);
    if (!((p_out >= 0) & (p_out <= 1))) {  // src/wfpt.pyx:63-64 (CTA-uniform when cta_per_set)
        if (lane == 0 && (!cta_per_set || warp == 0)) {
            a.logp[p] = -INFINITY;
            if (PHASE == 1) pending[p] = 0;
        }
        return;
    }
    double* __restrict__ dens = dens_all + p * dens_stride;
    const double* __restrict__ xs = a.x + lo;
    const int32_t* __restrict__ skip = a.cont ? a.cont + lo : nullptr;
    // src/wfpt.pyx:70 (contaminant: p_out = 0).  Spelled out: which of the two products the compiler fuses into the
    // addition must not depend on the instantiation (the one- and two-launch flavours give a set the same bits)
    const double keep = 1 - p_out, floor_wp = __dmul_rn(w_out, p_out);
    auto mixed = [&](double d) { return fma(d, keep, floor_wp); };
    const bool lean_ok = !SIMPLE && KMAX == 4 && L.ok != 0 && L.uniform != 0;
    const bool tree_ok = lean_ok && a.nm.n_st == 2;
    const bool simple_ok = SIMPLE && KMAX == 4 && (L.ok1 & 4) != 0 && L.uniform != 0;  // neither sz nor st: lean_simple
    unsigned long long evals = 0;
    int queued = 0;  // warp-uniform
    DBG(10, lane == 0 && (!cta_per_set || warp == 0));
    DBG(11, lane == 0 && (!cta_per_set || warp == 0) && !lean_ok);
    DBG(12, lane == 0 && (!cta_per_set || warp == 0) && !lean_ok && !G.support);
    DBG(13, lane == 0 && (!cta_per_set || warp == 0) && !lean_ok && G.support && !G.ok);
    DBG(14, lane == 0 && (!cta_per_set || warp == 0) && !lean_ok && G.support && G.ok && (G.st == 0 || G.sz == 0));
#pragma unroll 1
    for (int i0 = cta_per_set ? 32 * warp : 0; i0 < n; i0 += cta_per_set ? 32 * kWarpSets : 32) {
        const int i = i0 + lane;
        const bool in_range = i < n;
        const bool valid = in_range && !(skip && skip[i] != 0);
        const double x = valid ? __ldg(xs + i) : 1.0;
        bool need = false;
        if (PHASE == 2) {
            need = in_range && dens[i] == kDeferred;
        } else if (lean_ok || simple_ok) {
            int made = 0;
            double d;
            if (SIMPLE) d = lean_simple<KMAX>(G, L, ebig, lean_begin(L, x > 0 ? 1 : 0), x, valid, &need, &made);
            else d = lean_first_level<KMAX>(G, L, ebig, lean_begin(L, x > 0 ? 1 : 0), x, valid, &need, &made);
            evals += made;
            if (in_range & !need) dens[i] = valid ? mixed(d) : 1.0;  // (a skipped trial adds log 1)
        } else {
            need = valid;
            if (in_range & !valid) dens[i] = 1.0;
        }
        DBG(15, need);
        const unsigned int votes = __ballot_sync(0xffffffffu, need);
        if (PHASE == 1) {
            if (need) dens[i] = kDeferred;
        } else if (need) {
            queue[warp][queued + __popc(votes & ((1u << lane) - 1u))] = (unsigned short)i;
        }
        queued += __popc(votes);
    }
    __syncwarp();
    if (PHASE == 1) {  // anything left over in this set?  (CTA-wide when the CTA shares the set)
        const bool some = cta_per_set ? __syncthreads_or(queued > 0) : queued > 0;
        if (lane == 0 && (!cta_per_set || warp == 0)) pending[p] = some ? 1 : 0;
        if (some) {
            if (a.evals) {
                evals = warp_sum_u64(evals);
                if (lane == 0) atomicAdd(a.evals, evals);
            }
            return;
        }
    }
    int left = PHASE == 1 ? 0 : queued;
    if (PHASE != 1 && tree_ok) {  // the lean tree over the warp's queue; what it cannot settle is compacted in place
        left = 0;
#pragma unroll 1
        for (int k0 = 0; k0 < queued; k0 += 32) {
            const int k = k0 + lane;
            const bool on = k < queued;
            const unsigned short off = on ? queue[warp][k] : 0;
            const double x = on ? __ldg(xs + off) : 1.0;
            bool fail = false;
            int made = 0;
            const double d = lean_tree<KMAX>(G, L, ebig, x, lean_begin(L, x > 0 ? 1 : 0), on, &fail, &made);
            fail &= on;
            if (on & !fail) {
                evals += made;
                dens[off] = mixed(d);
            }
            const unsigned int votes = __ballot_sync(0xffffffffu, fail);
            __syncwarp();
            if (fail) queue[warp][left + __popc(votes & ((1u << lane) - 1u))] = off;
            left += __popc(votes);
            __syncwarp();
        }
    }
    if (PHASE != 1)
    for (int k = lane; k < left; k += 32) {  // the complete evaluator (refinement walks, out-of-range fallback)
        const unsigned short off = queue[warp][k];
        dens[off] = mixed(full_pdf_grid<KMAX, 0>(__ldg(xs + off), v, G, a.nm, &a.tab, evals));
    }
    if (a.evals) {
        evals = warp_sum_u64(evals);
        if (lane == 0) atomicAdd(a.evals, evals);
    }
    // the row is complete: one warp reduces it
    if (cta_per_set) {
        __syncthreads();
        if (warp != 0) return;
    } else {
        __syncwarp();
    }
    double acc = 0.0, mant = 1.0;
    int expo = 0;
    bool hit_zero = false;
    static_assert(kWarpMaxTrials / 32 <= 256, "mantissa product must stay finite");
    for (int i = lane; i < n; i += 32) {
        const double d = dens[i];
        hit_zero |= (d == 0);  // src/wfpt.pyx:71-72, :377-379
        const int hw = __double2hiint(d);
        if ((unsigned)(hw - 0x00100000) < 0x7fe00000u) {
            expo += (hw >> 20) - 1023;
            mant *= __hiloint2double((hw & 0x000fffff) | 0x3ff00000, __double2loint(d));
        } else {
            acc += log(d);
        }
    }
    acc += fma((double)expo, 0.693147180559945309417, log(mant));
    acc = warp_sum(acc);
    const bool any_zero = __any_sync(0xffffffffu, hit_zero);
    if (lane == 0) a.logp[p] = any_zero ? -INFINITY : acc;
}

__global__ void __launch_bounds__(128) k_like_final(const LikeArgs a)
{
    const int64_t p = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (p >= a.n_sets) return;
    const int lane = threadIdx.x & 31;
    const double* row = a.partial + p * a.chunks;
    double s = 0.0;
    bool neg_inf = false;
    for (int c = lane; c < a.chunks; c += 32) {
        const double v = row[c];
        neg_inf |= (v == -INFINITY);
        s += v;
    }
    s = warp_sum(s);
    if (a.partial2) {  // split sweep: the deferred trials' sums, after the first launch's
        const double* row2 = a.partial2 + p * a.chunks2;
        double s2 = 0.0;
        for (int c = lane; c < a.chunks2; c += 32) {
            const double v = row2[c];
            neg_inf |= (v == -INFINITY);
            s2 += v;
        }
        s += warp_sum(s2);
    }
    neg_inf = __any_sync(0xffffffffu, neg_inf);
    // wiener_like returns -inf at the FIRST zero density whatever the running sum holds
    // (src/wfpt.pyx:71-72): a -inf partial (only produced by that rule) overrides NaN.
    if (a.mode != kModeMulti && a.mode != kModeMultiRL && neg_inf) s = -INFINITY;
    if (lane == 0) a.logp[p] = s;
}

// Freed blocks stay in the stream-ordered pool instead of going back to the OS (launch_like's scratch).
static void keep_pool_memory()
{
    static bool done = false;
    if (done) return;
    int dev = 0;
    cudaMemPool_t pool;
    unsigned long long keep = ~0ull;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess)
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    done = true;
}

// 2^(i/2048) for exp_lean, correctly rounded (long-double exp2l, then one rounding), uploaded once per process
// (one process per GPU: hddm_b200_init).
static cudaError_t ensure_exp_table()
{
    static std::once_flag once;
    static cudaError_t status = cudaSuccess;
    std::call_once(once, [] {
        std::vector<double> t(kExpBig);
        for (int i = 0; i < kExpBig; ++i) t[i] = (double)exp2l((long double)i / (long double)kExpBig);
        status = cudaMemcpyToSymbol(g_exp2_big, t.data(), sizeof(double) * kExpBig);
    });
    return status;
}

cudaError_t launch_like(const LikeArgs& a, cudaStream_t stream)
{
    if (a.n_sets <= 0) return cudaSuccess;
    {
        const cudaError_t e = ensure_exp_table();
        if (e != cudaSuccess) return e;
    }
    LikeArgs b = a;
    b.tab = series_table_for(a.nm.err);
    b.trials_per_block = kTrialsPerBlock;
    {
        // The per-set tables are rebuilt by every CTA of a set (~15k warp-instructions, mostly serial): with
        // plenty of CTAs to fill the GPU anyway, give each one four times the trials.  (Pass 2 stays at the end
        // of the CTA: running it in instalments as the queue fills was measured 20 % slower on onset-mixed data,
        // the two code regions evict each other from the instruction cache.)
        const int64_t longest = a.max_seg_len > 0 ? a.max_seg_len : a.n;
        const int64_t big_chunks = (longest + kTrialsPerBigBlock - 1) / kTrialsPerBigBlock;
        if (a.chunks >= 4 && (int64_t)a.n_sets * big_chunks >= 148 * 2 * 8) {
            b.trials_per_block = kTrialsPerBigBlock;
            b.chunks = (int)(big_chunks < 1 ? 1 : big_chunks);
        }
    }
    dim3 grid((unsigned)b.n_sets, (unsigned)b.chunks);
    // z (3) and sz (4) must be uniform per set for the shared start-point tables
    bool varies[7];  // per-trial columns and regression outcomes
    for (int k = 0; k < 7; ++k) varies[k] = a.trial_stride[k] != 0;
    for (int r = 0; r < a.n_reg; ++r) varies[a.reg[r].outcome] = true;
    const bool fast = g_kernel_choice != 1 && a.nm.use_adaptive != 0 && b.tab.n_thr >= 0 &&
                      b.tab.max_large_k <= 8 && !varies[3] && !varies[4];
    bool uniform = true;  // everything but the drift is uniform per set
    for (int k = 1; k < 7; ++k) uniform &= !varies[k];
    const bool gridk = fast && uniform && g_kernel_choice != 2;
    // two passes when threads own several trials each (segments longer than two rounds of the CTA)
    const int64_t longest = a.max_seg_len > 0 ? a.max_seg_len : a.n;
    const bool two_pass = gridk && g_kernel_choice != 3 && longest > 2 * kLikeThreads;
    // Every set is a simple DDM (no sz, no st): the one-pass kernel of its own.
    if (gridk && a.all_simple && !varies[0] && g_kernel_choice == 0 && b.tab.max_large_k <= 4 && a.n_reg == 0 &&
        (a.mode == kModeWienerLike || a.mode == kModeContaminant) && longest > kWarpMaxTrials) {
        // (shorter segments: the warp-per-set kernels below, which take such sets through lean_simple too)
        void* tables = nullptr;
        if (b.chunks >= 2 && (int64_t)b.n_sets * b.chunks >= 148 * 2 * 2) {  // tables once per set, not per CTA
            keep_pool_memory();
            if (cudaMallocAsync(&tables, (size_t)b.n_sets * sizeof(GridSet<4>), stream) == cudaSuccess) {
                k_build_tables<4><<<(unsigned)b.n_sets, kLikeThreads, 0, stream>>>(b, tables);
                b.set_tables = tables;
            } else {
                (void)cudaGetLastError();
                tables = nullptr;
            }
        }
        k_like_simple<4><<<grid, kLikeThreads, 0, stream>>>(b);
        if (b.chunks > 1) {
            const int sets_per_block = 128 / 32;
            k_like_final<<<(unsigned)((a.n_sets + sets_per_block - 1) / sets_per_block), 128, 0, stream>>>(b);
        }
        const cudaError_t e = cudaGetLastError();
        if (tables) cudaFreeAsync(tables, stream);
        return e;
    }
    // Big sweeps run the two passes as two launches: the first-pass code and the complete evaluator
    // together overflow the instruction cache when warps of one SM are in different passes (onset-mixed
    // data spent half of its cycles waiting for instructions).  One bit per trial travels between them.
    unsigned int* bits = nullptr;
    size_t map_bytes = 0;
    if (two_pass && b.trials_per_block == kTrialsPerBigBlock && g_kernel_choice != 4) {
        keep_pool_memory();
        map_bytes = (size_t)b.n_sets * b.chunks * (kTrialsPerBigBlock / 32) * sizeof(unsigned int);
        const size_t table_bytes = (size_t)b.n_sets * (b.tab.max_large_k <= 4 ? sizeof(GridSet<4>) : sizeof(GridSet<8>));
        if (cudaMallocAsync((void**)&bits, map_bytes + table_bytes, stream) != cudaSuccess) {
            (void)cudaGetLastError();  // no room for the bitmap: the one-launch flavour needs none
            bits = nullptr;
        }
    }
    if (bits) {
        b.defer_bits = bits;
        b.chunks2 = (b.chunks + kBlocksPerDeferredCta - 1) / kBlocksPerDeferredCta;
        b.partial2 = b.partial + (size_t)b.n_sets * b.chunks;  // callers sized partial for 2048-trial chunks
        const dim3 grid2((unsigned)b.n_sets, (unsigned)b.chunks2);
        void* tables = reinterpret_cast<char*>(bits) + map_bytes;  // map_bytes is a multiple of 1024
        if (b.tab.max_large_k <= 4)
            k_build_tables<4><<<(unsigned)b.n_sets, kLikeThreads, 0, stream>>>(b, tables);
        else
            k_build_tables<8><<<(unsigned)b.n_sets, kLikeThreads, 0, stream>>>(b, tables);
        b.set_tables = tables;
        if (b.tab.max_large_k <= 4) {
            k_like_partial<4, true, kSplitFirst><<<grid, kLikeThreads, 0, stream>>>(b);
            k_like_partial<4, true, kSplitSecond><<<grid2, kLikeThreads, 0, stream>>>(b);
        } else {
            k_like_partial<8, true, kSplitFirst><<<grid, kLikeThreads, 0, stream>>>(b);
            k_like_partial<8, true, kSplitSecond><<<grid2, kLikeThreads, 0, stream>>>(b);
        }
        const int sets_per_block = 128 / 32;
        k_like_final<<<(unsigned)((a.n_sets + sets_per_block - 1) / sets_per_block), 128, 0, stream>>>(b);
        const cudaError_t e = cudaGetLastError();
        cudaFreeAsync(bits, stream);
        return e;
    }
    // Short segments (one chunk each): a warp per set on tables built by a warp per set.  Whatever the number of
    // sets -- a scalar wiener_like call and a batch give a set the same sum, bit for bit.
    if (gridk && !varies[0] && g_kernel_choice == 0 && b.tab.max_large_k <= 4 && a.n_reg == 0 && longest <= kWarpMaxTrials &&
        (a.mode == kModeWienerLike || a.mode == kModeContaminant)) {
        keep_pool_memory();
        void* tables = nullptr;
        const size_t table_bytes = ((size_t)b.n_sets * sizeof(GridSet<4>) + 255) & ~(size_t)255;
        const int64_t dens_stride = longest > 0 ? longest : 1;
        const size_t dens_bytes = (((size_t)b.n_sets * dens_stride * sizeof(double)) + 255) & ~(size_t)255;
        if (cudaMallocAsync(&tables, table_bytes + dens_bytes + (size_t)b.n_sets * sizeof(int), stream) == cudaSuccess) {
            const unsigned blocks = (unsigned)((b.n_sets + kWarpSets - 1) / kWarpSets);
            k_build_tables_warp<4><<<blocks, 32 * kWarpSets, 0, stream>>>(b, tables);
            b.set_tables = tables;
            // few sets: a CTA per set (its four warps share the rounds); many: a warp per set.  Same sums either way.
            // (long segments too: a CTA's four warps then share >= 16 rounds, and the GPU sees four times as many,
            //  four times shorter units -- 1e4 sets of 1000 trials are 4.2 waves of warp-per-set CTAs)
            const int cta_per_set = (b.n_sets <= 148 * 16 || longest >= kWarpCtaTrials) ? 1 : 0;
            const unsigned grid_w = cta_per_set ? (unsigned)b.n_sets : blocks;
            double* dens = reinterpret_cast<double*>(reinterpret_cast<char*>(tables) + table_bytes);
            int* pending = reinterpret_cast<int*>(reinterpret_cast<char*>(tables) + table_bytes + dens_bytes);
            if (b.n_sets <= kWarpOneLaunchSets) {  // (the same sums, bit for bit, either way)
                if (a.all_simple) k_like_warp<4, 0, true><<<grid_w, 32 * kWarpSets, 0, stream>>>(b, cta_per_set, dens, dens_stride, pending);
                else k_like_warp<4, 0, false><<<grid_w, 32 * kWarpSets, 0, stream>>>(b, cta_per_set, dens, dens_stride, pending);
            } else if (a.all_simple) {
                k_like_warp<4, 1, true><<<grid_w, 32 * kWarpSets, 0, stream>>>(b, cta_per_set, dens, dens_stride, pending);
                k_like_warp<4, 2, true><<<grid_w, 32 * kWarpSets, 0, stream>>>(b, cta_per_set, dens, dens_stride, pending);
            } else {
                k_like_warp<4, 1, false><<<grid_w, 32 * kWarpSets, 0, stream>>>(b, cta_per_set, dens, dens_stride, pending);
                k_like_warp<4, 2, false><<<grid_w, 32 * kWarpSets, 0, stream>>>(b, cta_per_set, dens, dens_stride, pending);
            }
            const cudaError_t e = cudaGetLastError();
            cudaFreeAsync(tables, stream);
            return e;
        }
        (void)cudaGetLastError();  // no room for the tables: the CTA-per-set kernels build their own
    }
    // Several CTAs per set and more CTAs than the GPU holds at once: build each set's tables once here too.
    void* shared_tables = nullptr;
    if (gridk && b.chunks >= 2 && (int64_t)b.n_sets * b.chunks >= 148 * 2 * 2) {
        keep_pool_memory();
        const size_t bytes = (size_t)b.n_sets * (b.tab.max_large_k <= 4 ? sizeof(GridSet<4>) : sizeof(GridSet<8>));
        if (cudaMallocAsync(&shared_tables, bytes, stream) != cudaSuccess) {
            (void)cudaGetLastError();
            shared_tables = nullptr;
        } else {
            if (b.tab.max_large_k <= 4)
                k_build_tables<4><<<(unsigned)b.n_sets, kLikeThreads, 0, stream>>>(b, shared_tables);
            else
                k_build_tables<8><<<(unsigned)b.n_sets, kLikeThreads, 0, stream>>>(b, shared_tables);
            b.set_tables = shared_tables;
        }
    }
    if (two_pass && b.tab.max_large_k <= 4)
        k_like_partial<4, true, kTwoPass><<<grid, kLikeThreads, 0, stream>>>(b);
    else if (two_pass)
        k_like_partial<8, true, kTwoPass><<<grid, kLikeThreads, 0, stream>>>(b);
    else if (gridk && b.tab.max_large_k <= 4)
        k_like_partial<4, true, kOnePass><<<grid, kLikeThreads, 0, stream>>>(b);
    else if (gridk)
        k_like_partial<8, true><<<grid, kLikeThreads, 0, stream>>>(b);
    else if (fast && b.tab.max_large_k <= 4)
        k_like_partial<4, false><<<grid, kLikeThreads, 0, stream>>>(b);
    else if (fast)
        k_like_partial<8, false><<<grid, kLikeThreads, 0, stream>>>(b);
    else
        k_like_partial<0, false><<<grid, kLikeThreads, 0, stream>>>(b);
    if (b.chunks == 1) return cudaGetLastError();  // short nodes: one launch, k_like_partial wrote logp
    const int sets_per_block = 128 / 32;
    k_like_final<<<(unsigned)((a.n_sets + sets_per_block - 1) / sets_per_block), 128, 0, stream>>>(b);
    const cudaError_t e = cudaGetLastError();
    if (shared_tables) cudaFreeAsync(shared_tables, stream);
    return e;
}

// ---- allocation-free variant of the batched likelihood ------------------------------------------------------
// For callers that may not allocate (CUDA graph capture, serving loops): the [n_sets][chunks] partial sums that
// k_like_partial writes and k_like_final adds up live in this file-scope array instead of caller or pool memory,
// and the launch takes the route of launch_like that needs nothing else (every CTA builds its set's tables in
// shared memory; both passes in one launch).  Every GPU that loads the library holds the 2 MiB until exit.
// Each (set, chunk) slot is written by its CTA before k_like_final reads it, so nothing has to start at zero and
// the contents left by the previous call are never read; calls that overlap on the GPU (two streams) race on it.
static __device__ __align__(256) double g_static_partial[kStaticPartialDoubles];

// In host code `g_static_partial` is the host-side stand-in of the symbol, not a device address: the address is
// asked for once per device, with that device current.
static cudaError_t static_partial_address(double** out)
{
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static double* address[kMaxDevices] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!address[dev]) {
        double* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_static_partial);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        address[dev] = p;
    }
    *out = address[dev];
    return cudaSuccess;
}

cudaError_t launch_like_static(const LikeArgs& a, cudaStream_t stream)
{
    if (a.n_sets <= 0) return cudaSuccess;
    if (a.chunks < 1 || (int64_t)a.n_sets * a.chunks > kStaticPartialDoubles) return cudaErrorInvalidValue;
    {
        const cudaError_t e = ensure_exp_table();  // one upload per process, no allocation
        if (e != cudaSuccess) return e;
    }
    LikeArgs b = a;
    {
        const cudaError_t e = static_partial_address(&b.partial);
        if (e != cudaSuccess) return e;
    }
    b.tab = series_table_for(a.nm.err);
    b.trials_per_block = kTrialsPerBlock;  // chunks stays as the caller gave it: 2048 trials per CTA
    b.defer_bits = nullptr;
    b.partial2 = nullptr;
    b.chunks2 = 0;
    b.set_tables = nullptr;
    const dim3 grid((unsigned)b.n_sets, (unsigned)b.chunks);
    bool varies[7];
    for (int k = 0; k < 7; ++k) varies[k] = a.trial_stride[k] != 0;
    for (int r = 0; r < a.n_reg; ++r) varies[a.reg[r].outcome] = true;
    const bool fast = a.nm.use_adaptive != 0 && b.tab.n_thr >= 0 && b.tab.max_large_k <= 8 && !varies[3] && !varies[4];
    bool uniform = true;
    for (int k = 1; k < 7; ++k) uniform &= !varies[k];
    const bool gridk = fast && uniform;
    const int64_t longest = a.max_seg_len > 0 ? a.max_seg_len : a.n;
    const bool two_pass = gridk && longest > 2 * kLikeThreads;
    if (two_pass && b.tab.max_large_k <= 4)
        k_like_partial<4, true, kTwoPass><<<grid, kLikeThreads, 0, stream>>>(b);
    else if (two_pass)
        k_like_partial<8, true, kTwoPass><<<grid, kLikeThreads, 0, stream>>>(b);
    else if (gridk && b.tab.max_large_k <= 4)
        k_like_partial<4, true, kOnePass><<<grid, kLikeThreads, 0, stream>>>(b);
    else if (gridk)
        k_like_partial<8, true><<<grid, kLikeThreads, 0, stream>>>(b);
    else if (fast && b.tab.max_large_k <= 4)
        k_like_partial<4, false><<<grid, kLikeThreads, 0, stream>>>(b);
    else if (fast)
        k_like_partial<8, false><<<grid, kLikeThreads, 0, stream>>>(b);
    else
        k_like_partial<0, false><<<grid, kLikeThreads, 0, stream>>>(b);
    if (b.chunks == 1) return cudaGetLastError();  // k_like_partial wrote logp itself
    const int sets_per_block = 128 / 32;
    k_like_final<<<(unsigned)((a.n_sets + sets_per_block - 1) / sets_per_block), 128, 0, stream>>>(b);
    return cudaGetLastError();
}

// pdf_array (src/wfpt.pyx:32-48): per-trial density, outlier mix, optional log.
template <int KMAX, bool GRID>
__global__ void __launch_bounds__(kLikeThreads, KMAX > 0 ? 2 : 1)
    k_pdf_array(const __grid_constant__ PdfArgs a)
{
    constexpr int KT = KMAX > 0 ? KMAX : 1;
    __shared__ GridSet<KT> gs;
    WTable<KT>* wt = gs.wt;
    double par[kNumPar];
    const int64_t set = blockIdx.y;
#pragma unroll
    for (int k = 0; k < kNumPar; ++k) par[k] = a.par_sets ? __ldg(a.par_sets + k * a.n_sets + set) : a.par[k];
    const double* __restrict__ xs = a.x + set * a.x_set_stride;
    double* __restrict__ ys = a.y + set * a.n;
    if (GRID)
        build_gridset<KT>(gs, par, a.tab);
    else if (KMAX > 0)
        build_wtables<KT>(wt, par[3], par[4]);
    unsigned long long evals = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    // One decision time per point (st == 0: the density tables of gen_rts_from_cdf / gen_cdf_using_pdf): the lean
    // single-node evaluator, whole warps in step; a point that needs more than the first level takes the complete one.
    constexpr bool LEAN1 = GRID && KMAX == 4;
    __shared__ double ebig[LEAN1 ? kExpBig : 1];
    if (LEAN1) {
        if (threadIdx.x == 0) build_leanset(gs, gs.lean, par[0], a.nm, true);
        for (int k = threadIdx.x; k < kExpBig; k += kLikeThreads) ebig[k] = g_exp2_big[k];
        __syncthreads();
    }
    if (LEAN1 && (gs.lean.ok1 & 1)) {
        const LeanSet& L = gs.lean;
        constexpr int NP = kPdfIlp;  // ADJACENT points per thread and round: independent nodes for the FP64 pipe to
                                     // overlap, and (NP == 2) one 16-byte load of the RTs and one 16-byte store
        auto finish = [&](double x, double d, bool need, int made) {
            if (need) d = full_pdf_grid<KT>(x, par[0], gs, a.nm, &a.tab, evals);
            else evals += made;
            d = d * (1 - par[7]) + (par[8] * par[7]);
            return a.logp == 1 ? log(d) : d;
        };
        const bool vec = NP == 2 && ((reinterpret_cast<uintptr_t>(xs) | reinterpret_cast<uintptr_t>(ys)) & 15) == 0;
        for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x * NP; i0 < a.n; i0 += stride * NP) {
            const int64_t ib = i0 + (int64_t)threadIdx.x * NP;
            const bool pair = vec && ib + 1 < a.n;
            double x[NP], xt[NP], y[NP];
            if (pair) {
                const double2 t = __ldg(reinterpret_cast<const double2*>(xs + ib));
                x[0] = t.x;
                x[NP - 1] = t.y;
            } else {
#pragma unroll
                for (int u = 0; u < NP; ++u) x[u] = ib + u < a.n ? __ldg(xs + ib + u) : NAN;
            }
            bool all_large = true;
            int K = 1;
#pragma unroll
            for (int u = 0; u < NP; ++u) {
                xt[u] = fabs(x[u]) - gs.t;
                // large-time, inside the comfortable range of its boundary (signed compare: negative / NaN fail)
                all_large &= (__double_as_longlong(xt[u]) >= L.x_small[x[u] > 0 ? 1 : 0]) & (xt[u] <= L.xhi);
                K = max(K, 1 + (int)below(xt[u], L.x_k[0]) + (int)below(xt[u], L.x_k[1]) + (int)below(xt[u], L.x_k[2]));
            }
            if (__all_sync(0xffffffffu, all_large)) {  // the bulk of a density table: NP nodes in one basic block
                const int wK = __reduce_max_sync(0xffffffffu, K);
                double v[NP];
                int neg[NP];
                bool bad[NP];
#pragma unroll
                for (int u = 0; u < NP; ++u) {
                    neg[u] = 0;
                    bad[u] = false;
                    v[u] = lean_node_large<KT>(gs, L, lean_begin(L, x[u] > 0 ? 1 : 0), ebig, xt[u],
                                               exp_lean(gs.qk * xt[u], ebig), wK, neg[u], bad[u]);
                }
#pragma unroll
                for (int u = 0; u < NP; ++u) y[u] = finish(x[u], 12.0 * v[u], bad[u] | (neg[u] < 0), 5);
            } else {
#pragma unroll 1
                for (int u = 0; u < NP; ++u) {  // small-time nodes, the edge, the onset, the tail of the table
                    const bool valid = ib + u < a.n;
                    const double xu = valid ? x[u] : 1.0;
                    bool need = false;
                    int made = 0;
                    const double d = lean_single<KT>(gs, L, ebig, lean_begin(L, xu > 0 ? 1 : 0), xu, valid, &need, &made);
                    y[u] = valid ? finish(xu, d, need, made) : 0.0;
                }
            }
            if (pair) {
                *reinterpret_cast<double2*>(ys + ib) = make_double2(y[0], y[NP - 1]);
            } else {
#pragma unroll
                for (int u = 0; u < NP; ++u)
                    if (ib + u < a.n) ys[ib + u] = y[u];
            }
        }
    } else if (LEAN1 && (gs.lean.ok1 & 4)) {  // neither sz nor st: one pdf_sv per point
        const LeanSet& L = gs.lean;
        for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x; i0 < a.n; i0 += stride) {
            const int64_t i = i0 + threadIdx.x;
            const bool valid = i < a.n;
            const double x = valid ? __ldg(xs + i) : 1.0;
            bool need = false;
            int made = 0;
            double d = lean_simple<KT>(gs, L, ebig, lean_begin(L, x > 0 ? 1 : 0), x, valid, &need, &made);
            if (!valid) continue;
            if (need) d = full_pdf_grid<KT>(x, par[0], gs, a.nm, &a.tab, evals);
            else evals += made;
            d = d * (1 - par[7]) + (par[8] * par[7]);
            ys[i] = a.logp == 1 ? log(d) : d;
        }
    } else
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
        double d;
        if (GRID)
            d = full_pdf_grid<KT>(__ldg(xs + i), par[0], gs, a.nm, &a.tab, evals);
        else if (KMAX > 0)
            d = full_pdf_fast<(KMAX > 0 ? KMAX : 1)>(__ldg(xs + i), par[0], par[1], par[2], par[3],
                                                     par[4], par[5], par[6], a.nm, &a.tab, wt, evals);
        else
            d = full_pdf(__ldg(xs + i), par[0], par[1], par[2], par[3], par[4], par[5],
                         par[6], a.nm, &a.tab, evals);
        d = d * (1 - par[7]) + (par[8] * par[7]);
        ys[i] = a.logp == 1 ? log(d) : d;
    }
    if (a.evals) {
        evals = warp_sum_u64(evals);
        if ((threadIdx.x & 31) == 0) atomicAdd(a.evals, evals);
    }
}

cudaError_t launch_pdf_array(const PdfArgs& a, cudaStream_t stream)
{
    if (a.n <= 0) return cudaSuccess;
    PdfArgs b = a;
    if (!b.par_sets) {
        b.n_sets = 1;
        b.x_set_stride = 0;
    }
    if (b.n_sets <= 0) return cudaSuccess;
    if (b.n_sets > 65535) return cudaErrorInvalidValue;
    {
        const cudaError_t e = ensure_exp_table();
        if (e != cudaSuccess) return e;
    }
    // one CTA builds the per-set tables once, so with many sets give each CTA >= 2048 points
    int64_t blocks = (a.n + kLikeThreads - 1) / kLikeThreads;
    if (b.n_sets > 1) {
        // ... and with thousands of sets (the density tables of a posterior-predictive batch) no more CTAs per set
        // than fill the GPU a few times over: the table build costs about as much as 2048 points
        blocks = (a.n + kTrialsPerBlock - 1) / kTrialsPerBlock;
        const int64_t enough = (148 * 16 + b.n_sets - 1) / b.n_sets;
        if (blocks > enough) blocks = enough;
    }
    if (blocks > 148 * 64) blocks = 148 * 64;
    b.tab = series_table_for(a.nm.err);
    const bool fast = g_kernel_choice != 1 && a.nm.use_adaptive != 0 && b.tab.n_thr >= 0 &&
                      b.tab.max_large_k <= 8;
    const bool gridk = fast && g_kernel_choice != 2;
    if (gridk && b.tab.max_large_k <= 4)
        k_pdf_array<4, true><<<dim3((unsigned)blocks, (unsigned)b.n_sets), kLikeThreads, 0, stream>>>(b);
    else if (gridk)
        k_pdf_array<8, true><<<dim3((unsigned)blocks, (unsigned)b.n_sets), kLikeThreads, 0, stream>>>(b);
    else if (fast && b.tab.max_large_k <= 4)
        k_pdf_array<4, false><<<dim3((unsigned)blocks, (unsigned)b.n_sets), kLikeThreads, 0, stream>>>(b);
    else if (fast)
        k_pdf_array<8, false><<<dim3((unsigned)blocks, (unsigned)b.n_sets), kLikeThreads, 0, stream>>>(b);
    else
        k_pdf_array<0, false><<<dim3((unsigned)blocks, (unsigned)b.n_sets), kLikeThreads, 0, stream>>>(b);
    return cudaGetLastError();
}

// Register-resident dependent-FMA chains: the FP64 roofline denominator, measured on the box
// (MEASURED_PEAKS.json has no FP64 entry).  8 independent chains per thread.
__global__ void __launch_bounds__(256) k_fp64_peak(double* out, int iters)
{
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5,
           a6 = a0 + 6, a7 = a0 + 7;
    const double m = 0.999999, c = 1e-7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
            a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
        }
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
}

#ifdef HDDM_DEBUG_TREE
extern "C" int hddm_debug_counters(unsigned long long* out, int reset)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_dbg, sizeof(g_dbg));
    if (reset) {
        unsigned long long z[16] = {0};
        cudaMemcpyToSymbol(g_dbg, z, sizeof(z));
    }
    return 0;
}
#endif

// Returns achieved FP64 FLOP/s (2 per FMA) of the probe; *ms_out = kernel time.
double fp64_peak_probe(int iters, float* ms_out, cudaStream_t stream)
{
    const int blocks = 148 * 8, threads = 256;
    double* buf = nullptr;
    if (cudaMalloc(&buf, sizeof(double) * blocks * threads) != cudaSuccess) return -1.0;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k_fp64_peak<<<blocks, threads, 0, stream>>>(buf, 64);  // warm-up
    cudaEventRecord(e0, stream);
    k_fp64_peak<<<blocks, threads, 0, stream>>>(buf, iters);
    cudaEventRecord(e1, stream);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(buf);
    if (ms_out) *ms_out = ms;
    const double fmas = (double)blocks * threads * (double)iters * 64.0;
    return 2.0 * fmas / (ms * 1e-3);
}

}  // namespace hddm
