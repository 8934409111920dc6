"""Navier-Stokes rollout, the loop of reference pinn_kalman/simulator.py:35-63 (`step`): velocity -> pressure ->
density, `dt = 0.0005 * 5`, `dx = 1/200`.  The netCDF loading and plotting around it are out of scope; `rollout` takes
the fields directly."""
import torch

from ..op import ns_step

dt = 0.0005 * 5
dx = 1 / 200


def rollout(f, v, p, t_range=(0, 100), stride=1, keep=True):
    """f, p [B,1,H,W], v [B,2,H,W] fp32 CUDA -> (result, vel, pres): lists of the fields after every step (keep=True,
    as simulator.step returns them) or only the final (f, v, p) (keep=False: no per-step retention)."""
    result, vel, pres = [], [], []
    bufs = None
    for _ in torch.arange(*t_range, stride):
        if keep:
            f, v, p = ns_step.step(f, v, p, dt, dx)
            result.append(f)
            vel.append(v)
            pres.append(p)
        else:
            # ping-pong between two buffer sets: no allocation inside the loop, the caller's inputs stay untouched
            if bufs is None:
                bufs = [tuple(torch.empty_like(a) for a in (f, v, p)) for _ in range(2)]
            f, v, p = ns_step.step(f, v, p, dt, dx, out=bufs[0])
            bufs.reverse()
    if keep:
        return result, vel, pres
    return f, v, p
