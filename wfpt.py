"""Top-level `wfpt`, the module name the reference installs (setup.py:6, imported bare at
hddm/__init__.py:11-21): with this repository on sys.path `import wfpt` resolves to the CUDA path."""
from hddm_b200.wfpt import *  # noqa: F401,F403
from hddm_b200.wfpt import Trials, __all__  # noqa: F401
