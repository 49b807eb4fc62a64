"""Generate the committed parity fixtures under tests/golden/ (run in the build container).

Sources:
  * matlab_wfpt.json  -- the reference's own 20 MATLAB golden vectors, read from
    /root/reference/hddm/tests/matlab_values.py:3-162 (used by test_likelihoods.py:49-56).
  * ref_*.npz         -- outputs of the compiled UNMODIFIED reference (oracle/_ref, built by
    oracle/build_ref.py from /root/reference/src) on seeded inputs; the inputs are stored next
    to the outputs so tests never regenerate them.

/root/reference does not exist on the GPU box; these files do.   python tests/golden/make_golden.py
"""
import importlib.util
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402

PNAMES = ("v", "sv", "a", "z", "sz", "t", "st")


def draw_params(rng, n, sv_on=True, sz_on=True, st_on=True):
    """Parameter ranges of hddm/generate.py:38-46 (gen_rand_params)."""
    sv = rng.uniform(0, 2.5, n) if sv_on else np.zeros(n)
    sz = rng.uniform(0, 0.4, n) if sz_on else np.zeros(n)
    st = rng.uniform(0, 0.35, n) if st_on else np.zeros(n)
    z = rng.uniform(0.4, 0.6, n)
    v = rng.uniform(-4, 4, n)
    t = rng.uniform(0.2, 0.5, n)
    a = rng.uniform(0.5, 2.0, n)
    return np.stack([v, sv, a, z, sz, t, st], axis=1)


def draw_rts(rng, n, onset):
    g = rng.gamma(2.0, 0.25, n)
    sgn = np.where(rng.random(n) < 0.6, 1.0, -1.0)
    return (onset + g) * sgn


def matlab():
    p = "/root/reference/hddm/tests/matlab_values.py"
    spec = importlib.util.spec_from_file_location("mv", p)
    mv = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mv)
    keys = ("v", "t", "a", "z", "z_nonorm", "rt", "err", "matlab_wfpt")
    rows = [dict(zip(keys, map(float, r))) for r in mv.vals]
    json.dump({"source": "hddm/tests/matlab_values.py:3-162", "vals": rows},
              open(os.path.join(HERE, "matlab_wfpt.json"), "w"), indent=0)
    return len(rows)


def pointwise(W):
    """full_pdf on every (sv,sz,st) quadrant x numerics; both boundaries; onset-straddling RTs."""
    rng = np.random.default_rng(7001)
    rows, outs = [], []
    numerics = [(1e-4, 2, 2, 1, 1e-3), (1e-4, 10, 10, 1, 1e-3), (1e-4, 10, 10, 1, 1e-8),
                (1e-6, 4, 6, 1, 1e-5), (1e-4, 4, 4, 0, 0.0), (1e-8, 10, 10, 0, 0.0), (1e-2, 2, 2, 1, 1e-3)]
    for q in range(8):
        for (err, n_st, n_sz, ad, se) in numerics:
            n = 24 if (se == 1e-8 and q == 7) else 60
            P = draw_params(rng, n, q & 1, q & 2, q & 4)
            x = draw_rts(rng, n, 0.0)
            # a third of the RTs hug the onset window [t - st/2, t + st/2] (x - t_node <= 0 nodes)
            near = np.where(rng.random(n) < 0.33, 0.05, 1.0)
            x = np.sign(x) * (np.abs(x) * near + P[:, 5] + P[:, 6] * (rng.random(n) - 0.55))
            for i in range(n):
                r = list(P[i]) + [err, n_st, n_sz, ad, se]
                rows.append([x[i]] + r)
                outs.append(W.full_pdf(x[i], *P[i], err, n_st, n_sz, ad, se))
    # support / failure-mode cases (hddm/tests/test_likelihoods.py:225-262) must give exact 0
    base = dict(v=1., sv=.3, a=2., z=.5, sz=.2, t=.3, st=.1)
    bad = [dict(z=-.1), dict(z=1.1), dict(z=.05, sz=.2), dict(z=.95, sz=.2), dict(a=-1.), dict(t=-.1),
           dict(st=-.1), dict(sv=-.1), dict(sz=-.1), dict(sz=1.1), dict(t=.02, st=.1), dict(t=1.5)]
    for b in bad:
        p = dict(base, **b)
        for x in (0.9, -0.9):
            rows.append([x] + [p[k] for k in PNAMES] + [1e-4, 2, 2, 1, 1e-3])
            outs.append(W.full_pdf(x, *[p[k] for k in PNAMES], 1e-4, 2, 2, 1, 1e-3))
    return np.array(rows), np.array(outs)


def summed(W):
    rng = np.random.default_rng(7002)
    out = {}
    # wiener_like: HDDM defaults, bare defaults, simple DDM, -inf cases
    cases = []
    for k, (onset, n, q, num, outl) in enumerate([
            (0.70, 400, 7, (1e-4, 2, 2, 1, 1e-3), (0.05, 0.1)),
            (0.30, 400, 7, (1e-4, 2, 2, 1, 1e-3), (0.05, 0.1)),
            (0.30, 300, 7, (1e-4, 10, 10, 1, 1e-3), (0.0, 0.1)),
            (0.55, 500, 0, (1e-4, 10, 10, 1, 1e-8), (0.0, 0.1)),
            (0.55, 500, 1, (1e-4, 10, 10, 1, 1e-8), (0.05, 0.1)),
            (0.75, 200, 3, (1e-4, 10, 10, 1, 1e-6), (0.0, 0.1)),
            (0.75, 200, 5, (1e-4, 4, 4, 0, 0.0), (0.1, 0.2)),
            (0.75, 150, 7, (1e-4, 6, 6, 0, 0.0), (0.0, 0.1)),
            (0.10, 100, 7, (1e-4, 2, 2, 1, 1e-3), (0.0, 0.1)),     # -inf: some RTs before onset
            (0.70, 50, 7, (1e-4, 2, 2, 1, 1e-3), (1.5, 0.1))]):    # -inf: p_outlier out of range
        P = draw_params(rng, 1, q & 1, q & 2, q & 4)[0]
        x = draw_rts(rng, n, onset)
        cases.append((x, P, num, outl))
        out["wl_x_%d" % k] = x
        out["wl_p_%d" % k] = np.array(list(P) + list(num) + list(outl))
        out["wl_y_%d" % k] = np.array(W.wiener_like(x, *P, num[0], num[1], num[2], num[3], num[4],
                                                    outl[0], outl[1]))
    out["wl_n"] = np.array(len(cases))
    # the reference's own -inf case, hddm/tests/test_likelihoods.py:97
    out["wl_zero_rt"] = np.array(W.wiener_like(np.array([1., 2., 3., 0.]), 1, 0, 2, .5, 0, 0, 0, 1e-4))

    # pdf_array (per-trial density and log-density surface)
    for k, (q, logp, outl) in enumerate([(7, 1, (0.05, 0.1)), (7, 0, (0.0, 0.0)), (0, 1, (0.05, 0.1)),
                                         (1, 1, (0.0, 0.0)), (6, 0, (0.05, 0.1))]):
        P = draw_params(rng, 1, q & 1, q & 2, q & 4)[0]
        x = draw_rts(rng, 300, 0.35)
        out["pa_x_%d" % k] = x
        out["pa_p_%d" % k] = np.array(list(P) + [1e-4, logp, 2, 2, 1, 1e-3, outl[0], outl[1]])
        with np.errstate(divide="ignore"):
            out["pa_y_%d" % k] = W.pdf_array(x, *P, 1e-4, logp, 2, 2, 1, 1e-3, outl[0], outl[1])
    out["pa_n"] = np.array(5)

    # wiener_like_multi: regression layout (v and/or a, t per trial), +-999 sentinels
    for k, multi in enumerate([("v",), ("v", "a"), ("v", "t", "z"), ("v", "sv", "a", "z", "sz", "t", "st")]):
        n = 250
        base = draw_params(rng, 1)[0]
        x = draw_rts(rng, n, 0.75)
        x[rng.choice(n, 6, replace=False)] = np.array([999., -999., 999., -999., 999., -999.])
        per = draw_params(rng, n)
        cols = np.tile(base, (n, 1))
        for nm in multi:
            j = PNAMES.index(nm)
            cols[:, j] = per[:, j]
        args = [cols[:, j].copy() if PNAMES[j] in multi else base[j] for j in range(7)]
        y = W.wiener_like_multi(x, *args, 1e-4, set(multi), 10, 10, 1, 1e-3, 0.05, 0.1)
        out["wm_x_%d" % k] = x
        out["wm_cols_%d" % k] = cols
        out["wm_multi_%d" % k] = np.array([PNAMES.index(nm) for nm in multi])
        out["wm_y_%d" % k] = np.array(y)
    out["wm_n"] = np.array(4)

    # wiener_like_contaminant
    for k in range(3):
        P = draw_params(rng, 1)[0]
        x = draw_rts(rng, 200, 0.75 if k < 2 else 0.1)
        cont = (rng.random(200) < 0.1).astype(np.int32)
        y = W.wiener_like_contaminant(x, cont, *P, 0.1, 5.0, 1e-4, 10, 10, 1, 1e-6)
        out["wc_x_%d" % k] = x
        out["wc_c_%d" % k] = cont
        out["wc_p_%d" % k] = P
        out["wc_y_%d" % k] = np.array(y)
    out["wc_n"] = np.array(3)
    return out


def cdfs(W, CW):
    rng = np.random.default_rng(7003)
    out = {}
    k = 0
    for q in (7, 7, 7, 0, 1, 2, 4, 3, 5, 6):
        P = draw_params(rng, 1, q & 1, q & 2, q & 4)[0]
        # RTs before the onset window, inside it and past it, both signs
        mag = np.concatenate([P[5] + P[6] * rng.uniform(-0.7, 0.7, 10), P[5] + P[6] / 2 + rng.gamma(2.0, 0.3, 20)])
        x = mag * np.where(rng.random(mag.size) < 0.5, 1.0, -1.0)
        po, wo = ((0.05, 0.1) if k % 2 == 0 else (0.0, 0.1))
        out["cd_x_%d" % k] = x
        out["cd_p_%d" % k] = np.array(list(P) + [po, wo])
        out["cd_y_%d" % k] = CW.dmat_cdf_array(x, *P, po, wo)
        k += 1
    # drift node numerically zero: v = 0 and sv = 0 puts every GH node within 1e-7 (src/cdfdif.c:185)
    P = np.array([0.0, 0.0, 1.5, 0.5, 0.2, 0.3, 0.2])
    x = np.array([0.25, -0.25, 0.3, -0.35, 0.39, 0.5, -0.8, 1.5])
    out["cd_x_%d" % k] = x
    out["cd_p_%d" % k] = np.array(list(P) + [0.0, 0.1])
    out["cd_y_%d" % k] = CW.dmat_cdf_array(x, *P, 0.0, 0.1)
    k += 1
    out["cd_n"] = np.array(k)

    # gen_cdf_using_pdf + split_cdf
    for j, q in enumerate((7, 0, 3)):
        P = draw_params(rng, 1, q & 1, q & 2, q & 4)[0]
        xg, c = W.gen_cdf_using_pdf(*P, 1e-4, 500, 5., 2, 2, 1, 1e-3, 0.05, 0.1)
        out["gc_p_%d" % j] = P
        out["gc_x_%d" % j] = xg
        out["gc_y_%d" % j] = c
        parts = W.split_cdf(xg.copy(), c.copy())
        for m, arr in enumerate(parts):
            out["gc_split%d_%d" % (m, j)] = arr
    out["gc_n"] = np.array(3)

    # gen_rts_from_cdf under numpy's global RNG (reference draw order: f, then delays if st != 0)
    for j, q in enumerate((7, 3)):
        P = draw_params(rng, 1, q & 1, q & 2, q & 4)[0]
        np.random.seed(4242 + j)
        rts = W.gen_rts_from_cdf(*P, 400, -4.0, 4.0, 1e-3)
        np.random.seed(4242 + j)
        f = np.random.rand(400)
        ud = np.random.rand(400) if P[6] != 0 else np.zeros(400)
        out["gr_p_%d" % j] = P
        out["gr_f_%d" % j] = f
        out["gr_u_%d" % j] = ud
        out["gr_y_%d" % j] = rts
    out["gr_n"] = np.array(2)
    return out


def rl_task(rng, n, n_cond, binary_feedback=True):
    """A two-armed-bandit session in the layout of hddm/models/hddm_rl.py:61-73: split_by = condition,
    response 0/1, feedback = reward received, signed RT by response."""
    split_by = rng.integers(0, n_cond, n).astype(np.int64)
    response = rng.integers(0, 2, n).astype(np.int64)
    feedback = (rng.random(n) < np.where(response == 1, 0.8, 0.3)).astype(np.float64) if binary_feedback \
        else rng.normal(0.5, 0.3, n)
    x = (0.35 + rng.gamma(2.0, 0.25, n)) * np.where(response == 1, 1.0, -1.0)
    return x, response, feedback, split_by


def rl(W):
    """wiener_like_rlddm / wiener_like_rl / wiener_like_multi_rlddm (src/wfpt.pyx:79-241,277-320): the
    reference has no known-answer tests for them; pinned by these outputs of the compiled reference."""
    rng = np.random.default_rng(7004)
    out = {}
    RLN = PNAMES + ("alpha",)
    for k in range(12):
        n, n_cond = (240, 3) if k % 3 else (90, 1)
        x, response, feedback, split_by = rl_task(rng, n, n_cond, binary_feedback=(k % 4 != 3))
        q = 0.5
        alpha = rng.normal(0, 1.5)
        pos_alpha = 100.0 if k % 2 == 0 else rng.normal(0, 1.5)
        v = rng.uniform(0.5, 6)
        P = draw_params(rng, 1, k & 1, k & 2, k & 4)[0]
        P[0], P[2], P[5] = v, rng.uniform(0.8, 2.5), rng.uniform(0.1, 0.3)
        num = (1e-4, 2, 2, 1, 1e-3) if k < 9 else (1e-4, 10, 10, 1, 1e-8)
        outl = (0.05, 0.1) if k != 5 else (0.0, 0.1)
        if k == 7:
            x[5] = np.sign(x[5]) * 0.01  # zero density -> -inf only with p_outlier = 0 ... exercised at k == 5 too
        if k == 5:
            x[11] = np.sign(x[11]) * 0.01
        out["rl_x_%d" % k] = x
        out["rl_resp_%d" % k] = response
        out["rl_fb_%d" % k] = feedback
        out["rl_split_%d" % k] = split_by
        out["rl_p_%d" % k] = np.array([q, alpha, pos_alpha] + list(P) + list(num) + list(outl))
        out["rl_ddm_%d" % k] = np.array(W.wiener_like_rlddm(x, response, feedback, split_by, q, alpha, pos_alpha,
                                                            *P, num[0], num[1], num[2], num[3], num[4], *outl))
        out["rl_choice_%d" % k] = np.array(W.wiener_like_rl(response, feedback, split_by, q, alpha, pos_alpha,
                                                            v, P[3], p_outlier=outl[0], w_outlier=outl[1]))
        # regression flavour: trials in the caller's order, split_by sorted so that resets are per block
        multi = [("v",), ("v", "alpha"), ("a", "t"), ("v", "sv", "a", "z", "sz", "t", "st", "alpha")][k % 4]
        sb = np.sort(split_by)
        per = np.concatenate([draw_params(rng, n), rng.normal(0, 1, (n, 1))], axis=1)
        per[:, 0] = np.abs(per[:, 0]) + 0.5
        cols = np.tile(np.array(list(P) + [alpha]), (n, 1))
        for nm in multi:
            j = RLN.index(nm)
            cols[:, j] = per[:, j]
        args = [cols[:, j].copy() if RLN[j] in multi else cols[0, j] for j in range(8)]
        y = W.wiener_like_multi_rlddm(x, response, feedback, sb, q, *args, 1e-4, list(multi), 2, 2, 1, 1e-3, *outl)
        out["rl_msplit_%d" % k] = sb
        out["rl_mcols_%d" % k] = cols
        out["rl_multi_%d" % k] = np.array([RLN.index(nm) for nm in multi])
        out["rl_my_%d" % k] = np.array(y)
    out["rl_n"] = np.array(12)
    return out


def main():
    if not ref.available():
        raise SystemExit("oracle/_ref not built: python oracle/build_ref.py")
    W, CW = ref.wfpt(), ref.cdfdif_wrapper()
    if "--only-rl" in sys.argv:  # added after the first fixtures were frozen: leaves them untouched
        np.savez_compressed(os.path.join(HERE, "ref_rl.npz"), **rl(W))
        return
    n = matlab()
    rows, outs = pointwise(W)
    np.savez_compressed(os.path.join(HERE, "ref_full_pdf.npz"), rows=rows, out=outs)
    np.savez_compressed(os.path.join(HERE, "ref_likelihoods.npz"), **summed(W))
    np.savez_compressed(os.path.join(HERE, "ref_cdf.npz"), **cdfs(W, CW))
    print("matlab vectors:", n, " full_pdf cases:", len(outs),
          " zero cases:", int((outs == 0).sum()), " nan:", int(np.isnan(outs).sum()))


if __name__ == "__main__":
    main()
