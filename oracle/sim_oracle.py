"""CPU restatement of the random-walk simulator, hddm/generate.py:208-319
(_gen_rts_from_simulated_drift).  TEST INFRASTRUCTURE ONLY.

hddm/generate.py imports the hddm package (kabuki, pymc: absent), so the loop is restated with the same
numpy operations -- np.cumsum over blocks of nn = 1000 steps, np.where for the crossing, the linear
interpolation of :304-315 -- taking its random numbers from the caller instead of the global RNG:
u_init[i] = (delay uniform, start-point uniform, standard-normal drift deviate) and u_steps[i, k] = the
k-th step uniform of trial i (scipy's uniform.rvs(loc, scale) is loc + scale*u, norm.rvs(loc, scale) is
loc + scale*n).  "parity unpinned" by any reference test; pinned here against the CDF sampler
statistically, as hddm/tests/test_generate.py:21-29 does.  The v_switch branch is left out."""
import numpy as np


def gen_rts_from_simulated_drift(params, u_init, u_steps, dt=1e-4, intra_sv=1.):
    samples = u_init.shape[0]
    nn = 1000
    a, v = params['a'], params['v']
    st, sz = params.get('st', 0.), params.get('sz', 0.)
    start_delay = (params['t'] + st * u_init[:, 0]) - st / 2.                 # :244-248
    starting_points = ((params['z'] + sz * u_init[:, 1]) - sz / 2.) * a       # :251-255
    rts = np.full(samples, np.nan)
    step_size = np.sqrt(dt) * intra_sv                                        # :258
    for i in range(samples):
        y_0 = starting_points[i]
        if 'sv' in params and params['sv'] != 0:                              # :266-269
            drift_rate = v + params['sv'] * u_init[i, 2]
        else:
            drift_rate = v
        prob_up = 0.5 * (1 + np.sqrt(dt) / intra_sv * drift_rate)             # :278
        it, crossed = 0, False
        while not crossed and it * nn < u_steps.shape[1]:
            it += 1
            u = u_steps[i, (it - 1) * nn: it * nn]
            position = ((u < prob_up) * 2 - 1) * step_size                    # :286
            position[0] += y_0
            position = np.cumsum(position)
            cross_idx = np.where((position < 0) | (position > a))[0]          # :290
            if cross_idx.shape[0] > 0:
                crossed = True
            else:
                y_0 = position[-1]
        if not crossed:
            continue
        y2 = position[cross_idx[0]]                                           # :304-315
        y1 = position[cross_idx[0] - 1] if cross_idx[0] != 0 else y_0
        m = (y2 - y1) / dt
        b = y2 - m * ((it - 1) * nn + cross_idx[0]) * dt
        rt = ((0 - b) / m) if y2 < 0 else ((a - b) / m)
        rts[i] = (rt + start_delay[i]) * np.sign(y2)
    return rts
