"""CPU restatement of the posterior-predictive summary statistics.  TEST INFRASTRUCTURE ONLY.

The fifteen lambdas of gen_ppc_stats (hddm/utils.py:241-267) restated one for one with the same numpy /
scipy calls; the simulated data sets themselves come from oracle.gen_rts_from_cdf (the C restatement of
src/wfpt.pyx:323-354) or from the compiled reference.

PINNED: tests/test_oracle.py::test_ppc_oracle_reproduces_reference_gen_ppc_stats requires this function to
return, bit for bit (NaN classes included), what the reference's own gen_ppc_stats() returned for the 19
data sets of tests/golden/ref_ppc_stats.npz (made by tests/golden/make_golden_py.py, which executes
/root/reference/hddm/utils.py unmodified with empty stand-ins for its unused kabuki / pymc / matplotlib
imports)."""
import warnings

import numpy as np
from scipy.stats import scoreatpercentile

QUANTILES = (10, 30, 50, 70, 90)


def ppc_stats(x):
    """The OrderedDict of hddm/utils.py:250-266 evaluated on one data set of signed RTs, as a vector in
    insertion order: accuracy, mean_ub, std_ub, 10..90q_ub, mean_lb, std_lb, 10..90q_lb."""
    x = np.asarray(x, dtype=np.float64)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")  # np.mean / np.std of an empty selection warn and give nan
        out = [np.mean(x > 0), np.mean(x[x > 0]), np.std(x[x > 0])]                      # :251-255
        out += [scoreatpercentile(x[x > 0], q) if np.any(x > 0) else np.nan for q in QUANTILES]  # :256-258
        out += [np.mean(x[x < 0]), np.std(x[x < 0])]                                    # :261-262
        out += [scoreatpercentile(np.abs(x[x < 0]), q) if np.any(x < 0) else np.nan for q in QUANTILES]  # :263-265
    return np.array(out, dtype=np.float64)
