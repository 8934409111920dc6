"""ctypes front end of ns_step_ref.c (CPU oracle for op.ns_step; test infrastructure only)."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libns_step_ref.so")
_lib = None


def build():
    subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _SO


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def update_density(dens, vel, dt, dx):
    dens, vel = _f32(dens), _f32(vel)
    B, _, H, W = dens.shape
    assert H == W, "the reference indexes correctly only for H == W"
    out, t0, t1 = np.empty_like(dens), np.empty_like(dens), np.empty_like(dens)
    _load().ref_update_density(_p(out), _p(dens), _p(vel), _p(t0), _p(t1), B, H, W, ctypes.c_float(dt), ctypes.c_float(dx))
    return out


def update_velocity(vel, pres, dt, dx, bug_compat=False):
    vel, pres = _f32(vel), _f32(pres)
    B, _, H, W = vel.shape
    assert H == W
    out, vel_n = np.empty_like(vel), np.empty_like(vel)
    # one spare plane: with bug_compat the reference reads plane index B of vel_n as "v" of the last sample, which is
    # still inside vel_n (2B planes) for every B >= 1.
    t = [np.empty_like(pres) for _ in range(6)]
    _load().ref_update_velocity(_p(out), _p(vel), _p(pres), _p(vel_n), _p(t[0]), _p(t[1]), _p(t[2]), _p(t[3]), _p(t[4]),
                                _p(t[5]), B, H, W, ctypes.c_float(dt), ctypes.c_float(dx), 1 if bug_compat else 0)
    return out


def update_pressure(pres, vel, dt, dx):
    pres, vel = _f32(pres), _f32(vel)
    B, _, H, W = pres.shape
    assert H == W
    out = np.empty_like(pres)
    _load().ref_update_pressure(_p(out), _p(pres), _p(vel), B, H, W, ctypes.c_float(dt), ctypes.c_float(dx))
    return out


def synthetic_fields(B, S, seed=0):
    """smooth fields, |v| in [0.05, 0.8], no exact zeros."""
    rng = np.random.default_rng(seed)
    y, x = np.meshgrid(np.linspace(0, 1, S, endpoint=False), np.linspace(0, 1, S, endpoint=False), indexing="ij")

    def modes(amp):
        f = np.zeros((B, S, S), dtype=np.float64)
        for _ in range(4):
            kx, ky = rng.integers(1, 5, size=2)
            ph = rng.uniform(0, 2 * np.pi, size=(B, 1, 1))
            a = rng.uniform(0.5, 1.0, size=(B, 1, 1))
            f += a * np.sin(2 * np.pi * (kx * x + ky * y)[None] + ph)
        return amp * f / 4.0

    dens = (0.5 + 0.25 * modes(1.0))[:, None]
    pres = (0.1 * modes(1.0))[:, None]
    vel = np.stack([modes(1.0), modes(1.0)], axis=1)
    mag = 0.05 + 0.75 * (np.abs(vel) / max(np.abs(vel).max(), 1e-9))
    vel = np.where(vel >= 0, 1.0, -1.0) * mag
    return dens.astype(np.float32), vel.astype(np.float32), pres.astype(np.float32)
