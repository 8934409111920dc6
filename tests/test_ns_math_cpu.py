"""The per-cell arithmetic of the ns_step CUDA kernels (csrc/ns_math.cuh: fp32 with reciprocals) compiled for the host
and pinned to the oracle (which keeps the reference's float/double promotions, op/ns_step_kernel.cu:114-158, 204-234).
Runs without a GPU; the -m gpu tests check the same thing through the C-ABI."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import ns_step_ref

HERE = os.path.dirname(os.path.abspath(__file__))
DT, DX = 0.0025, 1 / 200


@pytest.fixture(scope="module")
def host(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("nsm") / "libnsmath.so")
    subprocess.check_call(["g++", "-O2", "-shared", "-fPIC", "-x", "c++", os.path.join(HERE, "ns_math_host.cpp"), "-o", so])
    lib = ctypes.CDLL(so)
    for n in ("ns_host_density", "ns_host_velocity", "ns_host_pressure", "ns_host_density_strip",
              "ns_host_velocity_strip", "ns_host_pressure_strip"):
        getattr(lib, n).argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int] * 3 + [ctypes.c_float] * 2
        getattr(lib, n).restype = None
    return lib


def _call(fn, a, b, out_shape):
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    out = np.empty(out_shape, np.float32)
    B, _, H, W = out_shape
    fn(out.ctypes.data, a.ctypes.data, b.ctypes.data, B, H, W, DT, DX)
    return out


def _err(got, ref):
    return np.abs(got - ref).max() / np.abs(ref).max()


@pytest.mark.parametrize("B,S", [(1, 16), (3, 24), (2, 64), (2, 192)])
def test_fast_arithmetic_matches_oracle(host, B, S):
    dens, vel, pres = ns_step_ref.synthetic_fields(B, S, seed=B * 100 + S)
    e = _err(_call(host.ns_host_density, dens, vel, dens.shape), ns_step_ref.update_density(dens, vel, DT, DX))
    assert e <= 2e-6, e
    e = _err(_call(host.ns_host_velocity, vel, pres, vel.shape), ns_step_ref.update_velocity(vel, pres, DT, DX))
    assert e <= 2e-6, e
    e = _err(_call(host.ns_host_pressure, pres, vel, pres.shape), ns_step_ref.update_pressure(pres, vel, DT, DX))
    assert e <= 2e-6, e


@pytest.mark.parametrize("B,S", [(1, 4), (1, 8), (3, 24), (2, 64), (2, 192)])
def test_strip_form_matches_oracle(host, B, S):
    """The register-window form of the vectorised kernels (ghost cells instead of wall cases), W % 4 == 0."""
    dens, vel, pres = ns_step_ref.synthetic_fields(B, S, seed=B * 100 + S)
    e = _err(_call(host.ns_host_density_strip, dens, vel, dens.shape), ns_step_ref.update_density(dens, vel, DT, DX))
    assert e <= 2e-6, e
    e = _err(_call(host.ns_host_velocity_strip, vel, pres, vel.shape), ns_step_ref.update_velocity(vel, pres, DT, DX))
    assert e <= 2e-6, e
    e = _err(_call(host.ns_host_pressure_strip, pres, vel, pres.shape), ns_step_ref.update_pressure(pres, vel, DT, DX))
    assert e <= 2e-6, e


def test_zero_velocity_is_nan_like_reference(host):
    dens, vel, pres = ns_step_ref.synthetic_fields(1, 16, seed=3)
    vel[0, 0, 5, 7] = 0.0
    got = _call(host.ns_host_density, dens, vel, dens.shape)
    ref = ns_step_ref.update_density(dens, vel, DT, DX)
    assert np.isnan(got[0, 0, 5, 7]) and not np.isfinite(ref[0, 0, 5, 7])
    m = np.isfinite(ref)
    assert np.array_equal(m, np.isfinite(got))
    assert np.abs(got[m] - ref[m]).max() / np.abs(ref[m]).max() <= 2e-6
