"""Adaptive Runge-Kutta (RK45 Dormand-Prince, RK23 Bogacki-Shampine) with the state resident in HBM.

The reference integrates its probability-flow ODEs with `scipy.integrate.solve_ivp` (sampling.py:459-473,
likelihood.py:86-99, inverse/conditional_sampling.py:10-29): every evaluation flattens the state to numpy (D2H), and
lifts it back (H2D).  Here the state (fp64, as scipy holds it) and the stage derivatives (fp32, as the score network
produces them) stay on the device; the stage combinations and the scaled error sum are kernels of libb200sde.so
(csrc/ode.cu) and only the error norm (one 9.5 KB partial-sum read) reaches the host per step.

Step-size control restates scipy's published algorithm (scipy 1.18 `_ivp/rk.py` RungeKutta._step_impl, `_ivp/common.py`
select_initial_step; SAFETY 0.9, MIN_FACTOR 0.2, MAX_FACTOR 10, RMS error norm) so that the same right-hand side takes
the same steps and `nfev` matches `solve_ivp`'s; tests/test_ode_gpu.py checks both against scipy itself.
"""
import ctypes
import math

import numpy as np
import torch

from . import _lib

SAFETY = 0.9
MIN_FACTOR = 0.2
MAX_FACTOR = 10.0

_TABLEAUS = {
    'RK45': dict(
        order=5, error_estimator_order=4, n_stages=6,
        C=[0, 1 / 5, 3 / 10, 4 / 5, 8 / 9, 1],
        A=[[0, 0, 0, 0, 0],
           [1 / 5, 0, 0, 0, 0],
           [3 / 40, 9 / 40, 0, 0, 0],
           [44 / 45, -56 / 15, 32 / 9, 0, 0],
           [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729, 0],
           [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656]],
        B=[35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84],
        E=[-71 / 57600, 0, 71 / 16695, -71 / 1920, 17253 / 339200, -22 / 525, 1 / 40]),
    'RK23': dict(
        order=3, error_estimator_order=2, n_stages=3,
        C=[0, 1 / 2, 3 / 4],
        A=[[0, 0, 0],
           [1 / 2, 0, 0],
           [0, 3 / 4, 0]],
        B=[2 / 9, 1 / 3, 4 / 9],
        E=[5 / 72, -1 / 12, -1 / 9, 1 / 8]),
}


class OdeResult:
    """The fields of scipy's OdeResult that the reference reads: `y_final` stands for `sol.y[:, -1]` (a fp64 device
    tensor), plus nfev / t / status / success / message and the accepted step sizes."""

    def __init__(self, y_final, t, nfev, status, message, steps, rejected):
        self.y_final, self.t, self.nfev, self.status, self.message = y_final, t, nfev, status, message
        self.success = status >= 0
        self.steps, self.rejected = steps, rejected


def _doubles(v):
    return (ctypes.c_double * len(v))(*[float(x) for x in v])


def _rms(x):
    return float(torch.linalg.vector_norm(x)) / math.sqrt(x.numel())


def solve_ivp(fun, t_span, y0, method='RK45', rtol=1e-3, atol=1e-6, max_step=float('inf'), first_step=None):
    """fun(t: float, x: fp32 CUDA tensor [n]) -> fp32 CUDA tensor [n];  y0: CUDA tensor (any shape, flattened).

    Same defaults and semantics as scipy.integrate.solve_ivp restricted to what the reference uses (scalar rtol / atol,
    no events, no dense output, only the final state kept).  `fun` must not keep a reference to its argument between
    calls (the stage buffer is reused) and may return a view of an internal buffer (it is copied at once)."""
    if method not in _TABLEAUS:
        raise ValueError("ode.solve_ivp: method must be one of %s (got %r)" % (sorted(_TABLEAUS), method))
    _lib.require_cuda(y0)
    tab = _TABLEAUS[method]
    ns = tab['n_stages']
    lib = _lib.lib()
    dev = y0.device
    t0, t_bound = float(t_span[0]), float(t_span[1])
    direction = float(np.sign(t_bound - t0)) if t_bound != t0 else 1.0
    n = y0.numel()
    y = y0.detach().reshape(-1).to(torch.float64).clone()
    y_new = torch.empty_like(y)
    x32 = torch.empty(n, device=dev, dtype=torch.float32)
    K = torch.empty(ns + 1, n, device=dev, dtype=torch.float32)
    partial = torch.empty(1184, device=dev, dtype=torch.float64)
    A = [_doubles(row[:s]) if s else None for s, row in enumerate(tab['A'])]
    Bc, Ec = _doubles(tab['B']), _doubles(tab['E'])
    err_exp = -1.0 / (tab['error_estimator_order'] + 1)
    nfev = 0

    def call(t, x, slot):
        nonlocal nfev
        nfev += 1
        out = fun(float(t), x)
        if out.dtype != torch.float32 or out.device != dev or out.numel() != n:
            raise RuntimeError("ode.solve_ivp: fun must return %d fp32 values on %s" % (n, dev))
        K[slot].copy_(out.reshape(-1))

    def stage(dst32, dst64, coef, s, h):
        with torch.cuda.device(dev):
            _lib.check(lib.b200_rk_stage(_lib.ptr(dst32), _lib.ptr(dst64), _lib.ptr(y), _lib.ptr(K), n, s, coef, h, n,
                                         _lib.stream()), "b200_rk_stage")

    x32.copy_(y)
    call(t0, x32, 0)
    if n == 0 or t0 == t_bound:
        return OdeResult(y, t0, nfev, 0, "finished", [], 0)
    if first_step is None:
        # scipy/integrate/_ivp/common.py select_initial_step (order = error_estimator_order)
        interval = abs(t_bound - t0)
        scale = atol + y.abs() * rtol
        f0 = K[0].to(torch.float64)
        d0, d1 = _rms(y / scale), _rms(f0 / scale)
        h0 = 1e-6 if (d0 < 1e-5 or d1 < 1e-5) else 0.01 * d0 / d1
        h0 = min(h0, interval)
        stage(x32, None, _doubles([1.0]), 1, h0 * direction)
        call(t0 + h0 * direction, x32, 1)
        d2 = _rms((K[1].to(torch.float64) - f0) / scale) / h0
        if d1 <= 1e-15 and d2 <= 1e-15:
            h1 = max(1e-6, h0 * 1e-3)
        else:
            h1 = (0.01 / max(d1, d2)) ** (1.0 / (tab['error_estimator_order'] + 1))
        h_abs = min(100 * h0, h1, interval, max_step)
        del scale, f0
    else:
        h_abs = float(first_step)
    t = t0
    steps, rejected = [], 0
    status, message = None, "finished"
    while status is None:
        if t == t_bound:
            status = 0
            break
        # scipy/integrate/_ivp/rk.py RungeKutta._step_impl
        min_step = 10 * abs(np.nextafter(t, direction * np.inf) - t)
        h_abs = max_step if h_abs > max_step else (min_step if h_abs < min_step else h_abs)
        step_accepted = step_rejected = False
        while not step_accepted:
            if h_abs < min_step:
                status, message = -1, "Required step size is less than spacing between numbers."
                break
            h = h_abs * direction
            t_new = t + h
            if direction * (t_new - t_bound) > 0:
                t_new = t_bound
            h = t_new - t
            h_abs = abs(h)
            for s in range(1, ns):
                stage(x32, None, A[s], s, h)
                call(t + tab['C'][s] * h, x32, s)
            stage(x32, y_new, Bc, ns, h)
            call(t + h, x32, ns)
            with torch.cuda.device(dev):
                _lib.check(lib.b200_rk_error(_lib.ptr(partial), _lib.ptr(y), _lib.ptr(y_new), _lib.ptr(K), n, ns + 1, Ec,
                                             h, float(atol), float(rtol), n, _lib.stream()), "b200_rk_error")
            error_norm = math.sqrt(float(np.sum(partial.cpu().numpy())) / n)
            if not math.isfinite(error_norm):
                # scipy would spin on NaN (nan < 1 is False and the factor is nan); fail loudly instead
                raise FloatingPointError("ode.solve_ivp: non-finite error norm at t = %r" % t)
            if error_norm < 1:
                factor = MAX_FACTOR if error_norm == 0 else min(MAX_FACTOR, SAFETY * error_norm ** err_exp)
                if step_rejected:
                    factor = min(1.0, factor)
                h_abs *= factor
                step_accepted = True
            else:
                h_abs *= max(MIN_FACTOR, SAFETY * error_norm ** err_exp)
                step_rejected = True
                rejected += 1
        if status is not None:
            break
        steps.append(h)
        t = t_new
        y, y_new = y_new, y
        K[0].copy_(K[ns])                                     # first-same-as-last
        if direction * (t - t_bound) >= 0:
            status = 0
    return OdeResult(y, t, nfev, status, message, steps, rejected)
