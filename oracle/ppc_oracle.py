"""CPU restatement of the posterior-predictive summary statistics.  TEST INFRASTRUCTURE ONLY.

hddm/utils.py:241-267 (gen_ppc_stats) cannot be imported here (hddm/utils.py imports kabuki and pymc at
module level), so its fifteen lambdas are restated one for one with the same numpy / scipy calls; the
simulated data sets themselves come from oracle.gen_rts_from_cdf (the C restatement of
src/wfpt.pyx:323-354) or from the compiled reference."""
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
