"""celerite_b200 — B200-native (sm_100a) semiseparable Cholesky GP solver.

Drop-in for celerite's ``CholeskySolver`` path.  The compute path is hand-written CUDA behind a
C ABI (include/celerite_b200.h, built to celerite_b200/libcelerite_b200.so); there is no CPU
fallback: using the solver without the built library, or without a GPU, raises.
"""
__version__ = "0.1.0"
