"""CPU restatement of the random-walk simulator, hddm/generate.py:208-319
(_gen_rts_from_simulated_drift), incl. its v_switch variant (:234-241,272-276,283-285).
TEST INFRASTRUCTURE ONLY.

The loop is restated with the same numpy operations -- np.cumsum over blocks of nn steps (nn = 1000, or
round(t_switch / dt) with a switch, :233-239), np.where for the crossing, the linear interpolation of
:304-315 -- taking its random numbers from the caller instead of the global RNG: u_init[i] = (delay
uniform, start-point uniform, standard-normal drift deviate[, standard-normal deviate of the switched
drift]) and u_steps[i, k] = the k-th step uniform of trial i (scipy's uniform.rvs(loc, scale) is
loc + scale*u, norm.rvs(loc, scale) is loc + scale*n).

PINNED: tests/test_oracle.py::test_simulator_oracle_reproduces_reference_runs requires this function to
return, bit for bit, what the reference's own function returned for the ten runs stored in
tests/golden/ref_simdrift.npz (made by tests/golden/make_golden_py.py, which executes
/root/reference/hddm/generate.py unmodified and records the random numbers it consumed)."""
import numpy as np


def gen_rts_from_simulated_drift(params, u_init, u_steps, dt=1e-4, intra_sv=1.):
    samples = u_init.shape[0]
    nn = 1000
    a, v = params['a'], params['v']
    switch = 'v_switch' in params
    if switch:
        nn = int(round(params['t_switch'] / dt))                              # :234-239
    if 'st' in params:                                                        # :244-248
        start_delay = (params['st'] * u_init[:, 0] + params['t']) - params['st'] / 2.
    else:
        start_delay = np.ones(samples) * params['t']
    if 'sz' in params:                                                        # :251-255
        starting_points = ((params['sz'] * u_init[:, 1] + params['z']) - params['sz'] / 2.) * a
    else:
        starting_points = np.ones(samples) * params['z'] * a
    rts = np.full(samples, np.nan)
    step_size = np.sqrt(dt) * intra_sv                                        # :258
    for i in range(samples):
        y_0 = starting_points[i]
        if 'sv' in params and params['sv'] != 0:                              # :266-269
            drift_rate = u_init[i, 2] * params['sv'] + v
        else:
            drift_rate = v
        if switch:                                                            # :271-276
            if 'V_switch' in params and params['V_switch'] != 0:
                drift_rate_switch = u_init[i, 3] * params['V_switch'] + params['v_switch']
            else:
                drift_rate_switch = params['v_switch']
        prob_up = 0.5 * (1 + np.sqrt(dt) / intra_sv * drift_rate)             # :278
        it, crossed = 0, False
        while not crossed and it * nn < u_steps.shape[1]:
            it += 1
            if it == 2 and switch:                                            # :283-285
                prob_up = 0.5 * (1 + np.sqrt(dt) / intra_sv * drift_rate_switch)
            u = u_steps[i, (it - 1) * nn: it * nn]
            if np.isnan(u).any():                                             # ran out of supplied uniforms
                break
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
