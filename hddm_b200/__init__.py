"""hddm_b200 -- B200-native (sm_100a CUDA) implementation of HDDM's Wiener first-passage-time
hot path behind the reference's own module-function API.

    from hddm_b200 import wfpt, cdfdif_wrapper      # == hddm.wfpt / hddm.cdfdif of the reference
    import hddm_b200 as hddm; hddm.wfpt.wiener_like(...)

See include/hddm_b200.h for the C ABI, INTEGRATION.md for wiring it into the reference.
"""
from . import wfpt, cdfdif_wrapper  # noqa: F401

cdfdif = cdfdif_wrapper  # the reference re-exports cdfdif_wrapper as hddm.cdfdif (hddm/__init__.py:16-21)

__version__ = "0.1.0"
