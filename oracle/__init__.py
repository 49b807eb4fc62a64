"""CPU oracle for the WFPT hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  hddm_b200/ never does (tests/test_boundary.py enforces it).
"""
