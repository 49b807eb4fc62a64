"""Top-level `cdfdif_wrapper`, the module name the reference installs (setup.py:7, imported bare at
hddm/__init__.py:11-21)."""
from hddm_b200.cdfdif_wrapper import *  # noqa: F401,F403
from hddm_b200.cdfdif_wrapper import __all__  # noqa: F401
