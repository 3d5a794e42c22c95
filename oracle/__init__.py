"""ORACLE — CPU restatement of the reference's hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package, and only as the checker or the timed CPU baseline.  The product (nums_b200) never
does: its ops fail loudly when the CUDA library is missing.

Parity status: PINNED.  Every function here is checked (tests/test_oracle_golden.py) against outputs
of the reference itself, generated in the build container by tests/golden/make_golden.py, which
imports /root/reference through a NumPy-2 shim; the resulting vectors are committed under
tests/golden/.
"""
