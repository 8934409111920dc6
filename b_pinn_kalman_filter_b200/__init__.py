"""B200-native (sm_100a) implementation of the reverse-SDE predictor-corrector sampling hot path of
XDzzzzzZyq/b-pinn-kalman-filter, behind that project's own Python API:

    from b_pinn_kalman_filter_b200 import sampling, sde_lib, op
    from b_pinn_kalman_filter_b200.models import utils as mutils

Everything that computes runs in hand-written CUDA kernels reached through the C ABI in include/b200_sde.h
(libb200sde.so).  There is no CPU fallback; see DESIGN.md.
"""
__version__ = "0.1.0"
