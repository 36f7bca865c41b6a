"""Developer probes that use the oracle (checker code may only be called from tests/): run by hand, e.g. python tests/probes/config45_time.py.  Not collected by pytest (no test_ prefix)."""
