"""Drop-in for the reference's `cdfdif_wrapper` extension module (hddm.cdfdif,
/root/reference/src/cdfdif_wrapper.pyx + src/cdfdif.c): the full-DDM CDF on the GPU."""
import numpy as np

from ._lib import as_rt_buffer, check, dptr, lib

__all__ = ["dmat_cdf_array"]


def dmat_cdf_array(x, v, sv, a, z, sz, t, st, p_outlier, w_outlier):
    """src/cdfdif_wrapper.pyx:16-52.  AssertionError when p_outlier > 0 and max|x| >= 1/(2 w_outlier);
    ValueError when the parameters are out of support."""
    x = as_rt_buffer(x)
    if w_outlier == 0:
        # the reference's mixture divides 1./(2*w_outlier) under Python division semantics
        # (src/cdfdif_wrapper.pyx:12, no cdivision directive in that module)
        raise ZeroDivisionError("float division")
    y = np.empty(x.shape[0], dtype=np.double)
    check(lib().hddm_dmat_cdf_array(dptr(x), x.shape[0], v, sv, a, z, sz, t, st, p_outlier, w_outlier,
                                    dptr(y)))
    return y
