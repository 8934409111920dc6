"""Navier-Stokes simulator step with the reference's signatures (op/ns_step.py:19-26).

`update_velocity` follows the per-sample semantics the reference has at B = 1.  For B > 1 the reference mixes
velocity planes between samples (op/ns_step.cpp:70-72); pass `reference_bug_compat=True`
(or set `ns_step.REFERENCE_BUG_COMPAT = True`) to reproduce that bit of behaviour.
"""
import torch

from .. import _lib

REFERENCE_BUG_COMPAT = False


def _prep(*ts):
    _lib.require_cuda(*ts)
    out = []
    for t in ts:
        if t.dtype != torch.float32:
            raise RuntimeError("ns_step: fp32 fields only (got %s)" % t.dtype)
        out.append(t.contiguous())
    return out


def _shape(field, vel):
    B, C, H, W = field.shape
    if vel.shape != (B, 2, H, W) and not (C == 2 and vel.shape == (B, 1, H, W)):
        raise RuntimeError("ns_step: shape mismatch %s vs %s" % (tuple(field.shape), tuple(vel.shape)))
    return B, H, W


def update_density(dens, vel, dt, dx):
    dens, vel = _prep(dens, vel)
    B, H, W = _shape(dens, vel)
    out = torch.empty_like(dens)
    with torch.cuda.device(dens.device):
        rc = _lib.lib().b200_ns_update_density(_lib.ptr(out), _lib.ptr(dens), _lib.ptr(vel), B, H, W, float(dt),
                                               float(dx), _lib.stream())
    _lib.check(rc, "b200_ns_update_density")
    return out


def update_velocity(vel, pres, dt, dx, reference_bug_compat=None):
    vel, pres = _prep(vel, pres)
    B, H, W = _shape(vel, pres)
    compat = REFERENCE_BUG_COMPAT if reference_bug_compat is None else reference_bug_compat
    out = torch.empty_like(vel)
    scratch = torch.empty_like(vel) if compat else None
    with torch.cuda.device(vel.device):
        rc = _lib.lib().b200_ns_update_velocity(_lib.ptr(out), _lib.ptr(vel), _lib.ptr(pres), _lib.ptr(scratch), B, H,
                                                W, float(dt), float(dx), 1 if compat else 0, _lib.stream())
    _lib.check(rc, "b200_ns_update_velocity")
    return out


def update_pressure(pres, vel, dt, dx):
    pres, vel = _prep(pres, vel)
    B, H, W = _shape(pres, vel)
    out = torch.empty_like(pres)
    with torch.cuda.device(pres.device):
        rc = _lib.lib().b200_ns_update_pressure(_lib.ptr(out), _lib.ptr(pres), _lib.ptr(vel), B, H, W, float(dt),
                                                float(dx), _lib.stream())
    _lib.check(rc, "b200_ns_update_pressure")
    return out


def step(dens, vel, pres, dt, dx, out=None):
    """One simulator step velocity -> pressure -> density (pinn_kalman/simulator.py:55-57) through one C-ABI call.
    `out` = (dens, vel, pres) buffers to write into (ping-pong for rollouts), else fresh tensors.
    Returns (dens, vel, pres)."""
    dens, vel, pres = _prep(dens, vel, pres)
    B, H, W = _shape(dens, vel)
    _shape(pres, vel)
    if out is None:
        out = (torch.empty_like(dens), torch.empty_like(vel), torch.empty_like(pres))
    d1, v1, p1 = out
    for o, i in ((d1, dens), (v1, vel), (p1, pres)):
        if o.shape != i.shape or o.dtype != torch.float32 or not o.is_contiguous() or o.device != i.device:
            raise RuntimeError("ns_step.step: output buffers must match the inputs (fp32, contiguous, same device)")
    with torch.cuda.device(dens.device):
        rc = _lib.lib().b200_ns_step_fused(_lib.ptr(d1), _lib.ptr(v1), _lib.ptr(p1), _lib.ptr(dens), _lib.ptr(vel),
                                           _lib.ptr(pres), B, H, W, float(dt), float(dx), _lib.stream())
    _lib.check(rc, "b200_ns_step_fused")
    return d1, v1, p1
