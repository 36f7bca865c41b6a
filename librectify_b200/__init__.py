"""librectify_b200 — B200-native hot path of librectify behind the reference's C API.

The product is the shared library librectify_b200/lib/librectify_b200.so (sm_100a CUDA kernels + host C++), built by
``librectify_b200.build``; ``librectify_b200.api`` is its ctypes mirror, ``librectify_b200.synth`` the synthetic
Manhattan-scene generator used by tests and the benchmark.
"""
