"""Batched slice-within-Gibbs sampler for the hierarchical HDDM of BASELINE.json configs[1]
(subjects x trials, include=('sv','sz','st')): the caller that turns the batched likelihood into
MCMC iterations per second.

The reference samples this model with PyMC 2.3 + kabuki (hddm/models/hddm_info.py:121-175): one
observed `wfpt` node per subject, slice steps of fixed width per stochastic
(`slice_widths`, hddm/models/hddm_info.py:103-105; `SliceStep(width, left, maxiter=5000)`, :163-175), each
proposal re-evaluating the children's logp one node at a time.  Neither package can be installed
here, so this module restates the model (priors of `_create_stochastic_knodes_info`,
hddm/models/hddm_info.py:121-141, families of hddm/models/base.py:578-681) with the update order
re-designed for a batched likelihood:

  * subject-level parameters (v_subj, a_subj, t_subj) of ALL subjects and ALL chains take their slice
    steps in lock-step: every stepping-out / shrinkage round is ONE launch of chains x subjects sets;
  * group-only parameters (sv, sz, st) touch every observed node: one launch of chains x subjects sets
    per proposal (all children of the node at once);
  * group means / spreads only see the subject-level priors: plain numpy, no likelihood.

The likelihood is a callable `loglik(set_node[P], params[9, P]) -> float64[P]`: on the product path it
is `hddm_b200.likelihoods.BatchedNodes.logp` (CUDA); the benchmark passes the CPU reference through the
same interface to time "the same sampler, reference likelihood".  This module never imports `oracle`.
"""
import math
import time

import numpy as np

SLICE_WIDTHS = {'a': 1, 't': 0.01, 'a_std': 1, 't_std': 0.15, 'sz': 1.1, 'v': 1.5, 'st': 0.1, 'sv': 3,
                'v_std': 1}  # hddm/models/hddm_info.py:103-105
SUBJ_WIDTHS = {'v': 1.5, 'a': 1, 't': 0.01}
MAXITER = 5000  # hddm/models/hddm_info.py:174


# ---- log-densities of the prior families (pymc 2.3 parameterisations) -----------------------------

def _norm_logpdf(x, mu, sd):
    return -0.5 * ((x - mu) / sd) ** 2 - np.log(sd) - 0.9189385332046727


def _halfnorm_logpdf(x, sd):
    return np.where(x > 0, -0.5 * (x / sd) ** 2 - np.log(sd) - 0.2257913526447274, -np.inf)


def _gamma_logpdf(x, shape, rate):
    with np.errstate(divide='ignore', invalid='ignore'):
        out = shape * np.log(rate) - _lgamma(shape) + (shape - 1) * np.log(x) - rate * x
    return np.where((x > 0) & (shape > 0) & (rate > 0), out, -np.inf)


def _beta_logpdf(x, a, b):
    with np.errstate(divide='ignore', invalid='ignore'):
        out = (a - 1) * np.log(x) + (b - 1) * np.log1p(-x) - (_lgamma(a) + _lgamma(b) - _lgamma(a + b))
    return np.where((x > 0) & (x < 1), out, -np.inf)


_lgamma = np.vectorize(math.lgamma, otypes=[np.float64])


def _gamma_from_mean_std(x, mean, std):
    """pm.Gamma(alpha=mean^2/std^2, beta=mean/std^2), hddm/models/base.py:649-668."""
    with np.errstate(divide='ignore', invalid='ignore'):
        return _gamma_logpdf(x, mean ** 2 / std ** 2, mean / std ** 2)


# ---- vectorised slice step (Neal 2003: stepping out + shrinkage), all units in lock-step -----------

def slice_step(x0, logp, width, rng, left=None, maxiter=MAXITER):
    """One slice update of every element of x0 (independent targets).  `logp(x) -> array like x` evaluates
    all targets at once (one batched launch).  Returns (x1, logp(x1), n_logp_calls)."""
    x0 = np.asarray(x0, dtype=np.float64)
    shape = x0.shape
    x0 = x0.ravel()
    n = x0.size
    calls = 1
    lp0 = np.asarray(logp(x0.reshape(shape)), dtype=np.float64).ravel()
    y = lp0 + np.log(rng.random(n))
    L = x0 - width * rng.random(n)
    R = L + width
    if left is not None:
        L = np.maximum(L, left)
    # stepping out, both ends per round
    grow_l = np.ones(n, dtype=bool) if left is None else (L > left)
    grow_r = np.ones(n, dtype=bool)
    it = 0
    while (grow_l.any() or grow_r.any()) and it < maxiter:
        lpl = np.asarray(logp(np.where(grow_l, L, x0).reshape(shape)), dtype=np.float64).ravel()
        lpr = np.asarray(logp(np.where(grow_r, R, x0).reshape(shape)), dtype=np.float64).ravel()
        calls += 2
        grow_l &= lpl > y
        grow_r &= lpr > y
        L = np.where(grow_l, L - width, L)
        if left is not None:
            hit = L <= left
            L = np.where(hit, left, L)
            grow_l &= ~hit
        R = np.where(grow_r, R + width, R)
        it += 1
    # shrinkage
    x1 = x0.copy()
    lp1 = lp0.copy()
    todo = np.ones(n, dtype=bool)
    it = 0
    while todo.any() and it < maxiter:
        cand = np.where(todo, L + (R - L) * rng.random(n), x1)
        lpc = np.asarray(logp(cand.reshape(shape)), dtype=np.float64).ravel()
        calls += 1
        ok = todo & (lpc > y)
        x1 = np.where(ok, cand, x1)
        lp1 = np.where(ok, lpc, lp1)
        rej = todo & ~ok
        L = np.where(rej & (cand < x0), cand, L)
        R = np.where(rej & (cand >= x0), cand, R)
        todo = rej
        it += 1
    return x1.reshape(shape), lp1.reshape(shape), calls


def slice_step_batched(x0, lp0, logp_many, width, rng, left=None, maxiter=MAXITER, m_out=3, m_shrink=6):
    """The same slice update as `slice_step` -- same random numbers consumed in the same order, same
    stepping-out and shrinkage decisions, hence the same result for the same generator state -- with the
    target evaluated in a few LARGE batches instead of once per probe (one launch per phase instead of one
    per probe: what makes a single chain fast on a GPU).

    Why speculation is exact: the points a slice step probes do not depend on the values it finds, only on
    how many probes it takes.  Stepping out visits L - k width and R + k width (k = 0, 1, ...) until the
    target drops below the level; shrinkage visits a candidate drawn from the current interval, and a
    REJECTED candidate narrows the interval towards x0 whatever its value was -- so, given the uniforms,
    the whole sequence of candidates is known in advance and the accepted one is the first above the level.

    `lp0`: the target at x0 (the caller keeps it from the previous update; None: evaluated here).
    `logp_many(X)`: X is [K, *x0.shape]; returns the target at every row, shape [K, *x0.shape].
    Returns (x1, lp1, n_batches, picks) with picks = (batch_id[n], row[n]): which batch and row of the
    shrinkage phase each unit's accepted point came from (row -1: kept x0), batch ids counting the calls of
    logp_many made here from 0."""
    x0 = np.asarray(x0, dtype=np.float64)
    shape = x0.shape
    x0 = x0.ravel()
    n = x0.size
    batches = 0

    def many(rows):  # rows: list of [n] vectors -> [K, n]
        X = np.stack(rows).reshape((len(rows),) + shape)
        return np.asarray(logp_many(X), dtype=np.float64).reshape(len(rows), n)

    if lp0 is None:
        lp0 = many([x0])[0]
        batches += 1
    else:
        lp0 = np.asarray(lp0, dtype=np.float64).ravel()
    y = lp0 + np.log(rng.random(n))
    L = x0 - width * rng.random(n)
    R = L + width
    if left is not None:
        L = np.maximum(L, left)
    grow_l = np.ones(n, dtype=bool) if left is None else (L > left)
    grow_r = np.ones(n, dtype=bool)
    it = 0
    while (grow_l.any() or grow_r.any()) and it < maxiter:
        # the next m_out rounds of both ends, as slice_step would probe them if every unit kept growing
        rows = []
        Lk, Rk, gl = L.copy(), R.copy(), grow_l.copy()
        for _ in range(m_out):
            rows.append(np.where(gl, Lk, x0)); rows.append(np.where(grow_r, Rk, x0))
            Lk = np.where(gl, Lk - width, Lk)
            if left is not None:
                hit = Lk <= left
                Lk = np.where(hit, left, Lk)
                gl = gl & ~hit
            Rk = np.where(grow_r, Rk + width, Rk)
        vals = many(rows)
        batches += 1
        for k in range(m_out):  # replay slice_step's loop on the table
            if not ((grow_l.any() or grow_r.any()) and it < maxiter):
                break
            lpl, lpr = vals[2 * k], vals[2 * k + 1]
            grow_l &= lpl > y
            grow_r &= lpr > y
            L = np.where(grow_l, L - width, L)
            if left is not None:
                hit = L <= left
                L = np.where(hit, left, L)
                grow_l &= ~hit
            R = np.where(grow_r, R + width, R)
            it += 1
    # shrinkage
    x1 = x0.copy()
    lp1 = lp0.copy()
    todo = np.ones(n, dtype=bool)
    pick_batch = np.full(n, -1, dtype=np.int64)
    pick_row = np.full(n, -1, dtype=np.int64)
    it = 0
    while todo.any() and it < maxiter:
        state = rng.bit_generator.state
        U = rng.random((m_shrink, n))  # row r = the draw of round r (the generator fills in C order)
        cands = []
        Lk, Rk = L.copy(), R.copy()
        for r in range(m_shrink):  # candidates of the next rounds if every pending unit keeps being rejected
            cand = np.where(todo, Lk + (Rk - Lk) * U[r], x1)
            cands.append(cand)
            Lk = np.where(todo & (cand < x0), cand, Lk)
            Rk = np.where(todo & (cand >= x0), cand, Rk)
        vals = many(cands)
        used = 0
        for r in range(m_shrink):
            if not (todo.any() and it < maxiter):
                break
            cand, lpc = cands[r], vals[r]
            ok = todo & (lpc > y)
            x1 = np.where(ok, cand, x1)
            lp1 = np.where(ok, lpc, lp1)
            pick_batch[ok] = batches
            pick_row[ok] = r
            rej = todo & ~ok
            L = np.where(rej & (cand < x0), cand, L)
            R = np.where(rej & (cand >= x0), cand, R)
            todo = rej
            it += 1
            used += 1
        batches += 1
        if used < m_shrink:  # give back the draws slice_step would not have made
            rng.bit_generator.state = state
            if used:
                rng.random((used, n))
    return x1.reshape(shape), lp1.reshape(shape), batches, (pick_batch, pick_row)


class HierarchicalDDM(object):
    """State + one-sweep update of the hierarchical full DDM, `chains` independent chains at once.

    loglik(set_node, params) evaluates P = chains * subjects sets in one call."""

    def __init__(self, n_subjects, loglik, chains=1, z=0.5, p_outlier=0.05, w_outlier=0.1, seed=0, batched=True):
        """batched=True: every slice step evaluates its probes speculatively in a few large launches
        (slice_step_batched) and the node likelihoods of the current state are carried from update to update;
        batched=False: one launch per probe (slice_step).  Both give the same chain for the same seed."""
        self.S, self.C = int(n_subjects), int(chains)
        self.loglik = loglik
        self.batched = bool(batched)
        self.node_ll = None  # [C, S] log-likelihood of every observed node at the current state (batched mode)
        self.z, self.p_outlier, self.w_outlier = z, p_outlier, w_outlier
        self.rng = np.random.default_rng(seed)
        C, S = self.C, self.S
        # starting values of hddm/models/hddm_info.py:124-141
        self.g = {'v': np.full(C, 2.0), 'v_std': np.full(C, 0.1), 'a': np.full(C, 1.5), 'a_std': np.full(C, 0.1),
                  't': np.full(C, 0.4), 't_std': np.full(C, 0.2), 'sv': np.full(C, 1.0), 'sz': np.full(C, 0.01),
                  'st': np.full(C, 0.001)}
        self.subj = {'v': np.full((C, S), 2.0), 'a': np.full((C, S), 1.0), 't': np.full((C, S), 0.001)}
        self.set_node = np.tile(np.arange(S, dtype=np.int64), C)
        self.loglik_calls = 0
        self.sets_evaluated = 0

    # -- likelihood of every (chain, subject) node for the current state with overrides
    def _node_loglik(self, **over):
        C, S = self.C, self.S
        get = lambda name: over[name] if name in over else (self.subj[name] if name in self.subj else self.g[name])
        bc = lambda a: np.broadcast_to(np.asarray(a, dtype=np.float64).reshape(C, -1), (C, S)).ravel()
        params = np.empty((9, C * S))
        params[0] = bc(get('v')); params[1] = bc(get('sv')); params[2] = bc(get('a')); params[3] = self.z
        params[4] = bc(get('sz')); params[5] = bc(get('t')); params[6] = bc(get('st'))
        params[7] = self.p_outlier; params[8] = self.w_outlier
        self.loglik_calls += 1
        self.sets_evaluated += C * S
        return np.asarray(self.loglik(self.set_node, params), dtype=np.float64).reshape(C, S)

    # -- the same for K candidate values of ONE parameter at once: X is [K, C, S] (subject-level) or [K, C]
    #    (group-only, shared by the chain's subjects); one launch of K * C * S sets -> [K, C, S]
    def _node_loglik_many(self, name, X):
        C, S = self.C, self.S
        K = X.shape[0]
        base = np.empty((9, C * S))
        bc = lambda a: np.broadcast_to(np.asarray(a, dtype=np.float64).reshape(C, -1), (C, S)).ravel()
        cur = lambda n: self.subj[n] if n in self.subj else self.g[n]
        base[0] = bc(cur('v')); base[1] = bc(cur('sv')); base[2] = bc(cur('a')); base[3] = self.z
        base[4] = bc(cur('sz')); base[5] = bc(cur('t')); base[6] = bc(cur('st'))
        base[7] = self.p_outlier; base[8] = self.w_outlier
        params = np.empty((9, K, C * S))
        params[:] = base[:, None, :]
        row = {'v': 0, 'sv': 1, 'a': 2, 'sz': 4, 't': 5, 'st': 6}[name]
        params[row] = np.broadcast_to(X.reshape(K, C, -1), (K, C, S)).reshape(K, C * S)
        self.loglik_calls += 1
        self.sets_evaluated += K * C * S
        out = self.loglik(np.tile(self.set_node, K), params.reshape(9, K * C * S))
        return np.asarray(out, dtype=np.float64).reshape(K, C, S)

    def _subj_prior(self, name, x, g=None, std=None):
        g = self.g[name][:, None] if g is None else g
        std = self.g[name + '_std'][:, None] if std is None else std
        if name == 'v':
            return _norm_logpdf(x, g, std)
        return _gamma_from_mean_std(x, g, std)

    def sweep(self):
        """One MCMC iteration: every stochastic of the model is updated once."""
        if self.batched:
            return self._sweep_batched()
        rng = self.rng
        # subject-level parameters: chains x subjects targets in lock-step
        for name in ('v', 'a', 't'):
            target = lambda x, name=name: self._node_loglik(**{name: x}) + self._subj_prior(name, x)
            self.subj[name], _, _ = slice_step(self.subj[name], target, SUBJ_WIDTHS[name], rng)
        # group-only parameters: every observed node is a child
        priors = {'sv': lambda x: _halfnorm_logpdf(x, 2.0), 'sz': lambda x: _beta_logpdf(x, 1.0, 3.0),
                  'st': lambda x: _halfnorm_logpdf(x, 0.3)}
        for name in ('sv', 'sz', 'st'):
            target = lambda x, name=name: self._node_loglik(**{name: x[:, None]}).sum(axis=1) + priors[name](x)
            self.g[name], _, _ = slice_step(self.g[name], target, SLICE_WIDTHS[name], rng, left=0.0)
        # group means and spreads: children are the subject-level priors only
        hyper = {'v': lambda x: _norm_logpdf(x, 2.0, 3.0), 'a': lambda x: _gamma_from_mean_std(x, 1.5, 0.75),
                 't': lambda x: _gamma_from_mean_std(x, 0.4, 0.2)}
        std_sd = {'v': 2.0, 'a': 2.0, 't': 1.0}
        for name in ('v', 'a', 't'):
            xs = self.subj[name]
            tg = lambda g, name=name, xs=xs: self._subj_prior(name, xs, g=g[:, None]).sum(axis=1) + hyper[name](g)
            self.g[name], _, _ = slice_step(self.g[name], tg, SLICE_WIDTHS[name], rng)
            ts = lambda s, name=name, xs=xs: (self._subj_prior(name, xs, std=s[:, None]).sum(axis=1) +
                                              _halfnorm_logpdf(s, std_sd[name]))
            self.g[name + '_std'], _, _ = slice_step(self.g[name + '_std'], ts, SLICE_WIDTHS[name + '_std'], rng,
                                                     left=0.0)

    def _sweep_batched(self):
        """The same iteration as sweep() with batched=False -- same update order, same random numbers, same
        decisions -- in about two launches per likelihood-bearing stochastic family instead of one per probe."""
        rng = self.rng
        C, S = self.C, self.S
        if self.node_ll is None:
            self.node_ll = np.ascontiguousarray(self._node_loglik())

        def take(payloads, picks, per_chain):
            """Carry the node likelihoods of the accepted points into the cache."""
            pb, pr = picks
            if per_chain:  # unit = chain: the whole row of subjects moves
                for c in np.nonzero(pr >= 0)[0]:
                    self.node_ll[c] = payloads[pb[c]][pr[c], c]
            else:          # unit = (chain, subject)
                pb, pr = pb.reshape(C, S), pr.reshape(C, S)
                cs, ss = np.nonzero(pr >= 0)
                for c, s_ in zip(cs, ss):
                    self.node_ll[c, s_] = payloads[pb[c, s_]][pr[c, s_], c, s_]

        for name in ('v', 'a', 't'):
            payloads = []

            def many(X, name=name, payloads=payloads):  # X: [K, C, S]
                ll = self._node_loglik_many(name, X)
                payloads.append(ll)
                return ll + self._subj_prior(name, X)

            lp0 = self.node_ll + self._subj_prior(name, self.subj[name])
            self.subj[name], _, _, picks = slice_step_batched(self.subj[name], lp0, many, SUBJ_WIDTHS[name], rng)
            take(payloads, picks, per_chain=False)
        priors = {'sv': lambda x: _halfnorm_logpdf(x, 2.0), 'sz': lambda x: _beta_logpdf(x, 1.0, 3.0),
                  'st': lambda x: _halfnorm_logpdf(x, 0.3)}
        for name in ('sv', 'sz', 'st'):
            payloads = []

            def many(X, name=name, payloads=payloads):  # X: [K, C]
                ll = self._node_loglik_many(name, X)
                payloads.append(ll)
                return ll.sum(axis=2) + priors[name](X)

            lp0 = self.node_ll.sum(axis=1) + priors[name](self.g[name])
            self.g[name], _, _, picks = slice_step_batched(self.g[name], lp0, many, SLICE_WIDTHS[name], rng, left=0.0)
            take(payloads, picks, per_chain=True)
        hyper = {'v': lambda x: _norm_logpdf(x, 2.0, 3.0), 'a': lambda x: _gamma_from_mean_std(x, 1.5, 0.75),
                 't': lambda x: _gamma_from_mean_std(x, 0.4, 0.2)}
        std_sd = {'v': 2.0, 'a': 2.0, 't': 1.0}
        for name in ('v', 'a', 't'):  # no likelihood below: numpy only, still fewer and larger evaluations
            xs = self.subj[name]
            tg = lambda G, name=name, xs=xs: self._subj_prior(name, xs, g=G[..., None]).sum(axis=-1) + hyper[name](G)
            self.g[name], _, _, _ = slice_step_batched(self.g[name], None, tg, SLICE_WIDTHS[name], rng)
            ts = lambda Sd, name=name, xs=xs: (self._subj_prior(name, xs, std=Sd[..., None]).sum(axis=-1) +
                                               _halfnorm_logpdf(Sd, std_sd[name]))
            self.g[name + '_std'], _, _, _ = slice_step_batched(self.g[name + '_std'], None, ts,
                                                                SLICE_WIDTHS[name + '_std'], rng, left=0.0)

    def sample_native(self, nodes, n_iter, trace=True):
        """n_iter sweeps in the native driver (hddm_hier_ddm_sample, csrc/abi_mcmc.cu): the same chain as
        sample() draws from the same generator state -- same model, update order and random numbers -- with no
        Python between the launches.  `nodes`: the likelihoods.BatchedNodes holding the subjects' trials (its
        numerics and w_outlier are used; self.loglik is not called).  Returns (trace, seconds) like sample()."""
        import ctypes as C
        from . import _lib
        st = self.rng.bit_generator.state
        if st.get('bit_generator') != 'PCG64':
            raise ValueError("sample_native replays numpy's PCG64 stream; the model's generator is %r" % st.get('bit_generator'))
        mask = (1 << 64) - 1
        words = (C.c_uint64 * 4)(st['state']['state'] >> 64, st['state']['state'] & mask,
                                 st['state']['inc'] >> 64, st['state']['inc'] & mask)
        Cn, S = self.C, self.S
        gnames = ('v', 'v_std', 'a', 'a_std', 't', 't_std', 'sv', 'sz', 'st')
        g = np.ascontiguousarray(np.stack([self.g[k] for k in gnames]), dtype=np.float64)
        subj = np.ascontiguousarray(np.stack([self.subj[k] for k in ('v', 'a', 't')]), dtype=np.float64)
        tg = np.empty((n_iter, 9, Cn)) if trace else None
        ts = np.empty((n_iter, 3, Cn, S)) if trace else None
        launches = C.c_int64(0)
        nm = nodes.numerics
        node_off = np.ascontiguousarray(nodes.node_off, dtype=np.int64)
        if node_off.shape[0] != S + 1:
            raise ValueError("nodes holds %d observed nodes, the model %d subjects" % (node_off.shape[0] - 1, S))
        t0 = time.perf_counter()
        _lib.check(_lib.lib().hddm_hier_ddm_sample(
            nodes._handle()._h, node_off.ctypes.data_as(_lib.c_lp), S, Cn, _lib.dptr(g), _lib.dptr(subj),
            float(self.z), float(self.p_outlier), float(nodes.w_outlier), nm['err'], nm['n_st'], nm['n_sz'],
            int(bool(nm['use_adaptive'])), nm['simps_err'], words, int(n_iter),
            _lib.dptr(tg) if trace else None, _lib.dptr(ts) if trace else None, C.byref(launches)))
        dt = time.perf_counter() - t0
        for k, name in enumerate(gnames):
            self.g[name] = g[k].copy()
        for k, name in enumerate(('v', 'a', 't')):
            self.subj[name] = subj[k].copy()
        st['state']['state'] = (int(words[0]) << 64) | int(words[1])
        st['has_uint32'], st['uinteger'] = 0, 0
        self.rng.bit_generator.state = st
        self.node_ll = None
        self.loglik_calls += int(launches.value)
        tr = None
        if trace:
            tr = {name: tg[:, k] for k, name in enumerate(gnames)}
            tr.update({name + '_subj': ts[:, k] for k, name in enumerate(('v', 'a', 't'))})
        return tr, dt

    def sample(self, n_iter, trace=True):
        """Run n_iter sweeps; returns (trace dict name -> [n_iter, chains(, subjects)], seconds)."""
        names = list(self.g) + [k + '_subj' for k in self.subj]
        tr = {k: [] for k in names} if trace else None
        t0 = time.perf_counter()
        for _ in range(n_iter):
            self.sweep()
            if trace:
                for k in self.g:
                    tr[k].append(self.g[k].copy())
                for k in self.subj:
                    tr[k + '_subj'].append(self.subj[k].copy())
        dt = time.perf_counter() - t0
        if trace:
            tr = {k: np.stack(v) for k, v in tr.items()}
        return tr, dt
