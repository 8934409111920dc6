"""Deterministic "trained-like" weights for parity fixtures (test infrastructure only).

The reference initialises every residual-branch output conv, attention NIN_3 and the final conv with scale 1e-10
(`init_scale=0.`, models/layers.py:83-86), so an as-initialised network outputs ~0 and would pass parity with most
kernels broken.  Fixtures therefore re-draw EVERY tensor at O(1) scale from a seeded CPU generator:
conv / NIN / linear weights ~ U(+-sqrt(3/fan_avg)), biases ~ 0.1 N(0,1), GroupNorm gamma ~ 1 + 0.1 N(0,1),
beta ~ 0.1 N(0,1), Fourier W ~ 16 N(0,1).  The same state dict is loaded into the reference / oracle and into the
B200 engine.
"""
import math

import numpy as np
import torch


def trained_like_state_dict(spec, config, seed=0):
    """spec: [(key, shape, init)] from b_pinn_kalman_filter_b200.models.engine.parameter_spec."""
    gen = torch.Generator().manual_seed(int(seed))
    sd = {}
    for key, shape, init in spec:
        if init == "ones":
            t = 1.0 + 0.1 * torch.randn(shape, generator=gen)
        elif init == "zeros":
            t = 0.1 * torch.randn(shape, generator=gen)
        elif init[0] == "randn":
            t = torch.randn(shape, generator=gen) * init[1]
        else:
            if len(shape) == 1:
                fan = shape[0]
            else:
                rf = float(np.prod(shape)) / shape[0] / shape[1]
                fan = (shape[0] + shape[1]) * rf / 2
            t = (torch.rand(shape, generator=gen) * 2 - 1) * math.sqrt(3.0 / fan)
        sd[key] = t.to(torch.float32)
    m = config.model
    sd["sigmas"] = torch.tensor(np.exp(np.linspace(np.log(m.sigma_max), np.log(m.sigma_min), m.num_scales)),
                                dtype=torch.float32)
    return sd
