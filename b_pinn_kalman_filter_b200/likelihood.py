"""Log-likelihood (bits/dim) under the probability-flow ODE with the reference's signature (likelihood.py:40-113).

The ODE is integrated by ode.solve_ivp with the state (image + accumulated log-det) resident on the device
(solver='device', RK45 / RK23) or by scipy as in the reference (solver='scipy'); each right-hand side needs the drift AND the Hutchinson-Skilling divergence
estimate eps^T (d drift / d x) eps (likelihood.py:27-37).  The reference gets the latter by autograd through a second
network evaluation; here ONE engine forward (activations kept) gives the drift and ONE input-gradient pass gives
J^T eps:   drift = f(t) x - g(t)^2 score / 2   =>   eps^T J_drift eps = f(t) |eps|^2 - g(t)^2 eps . (J_score^T eps) / 2.
"""
import numpy as np
import torch
from scipy import integrate

from .models import utils as mutils


def get_drift_div_fn(sde):
    """(model, x, t, eps) -> (drift of the probability-flow ODE, per-sample divergence estimate)."""

    def drift_div(model, x, t, eps):
        score, vjp = mutils.get_score_vjp_fn(sde, model, continuous=True)(x, t)
        f_coef, g = sde.coefficient(t)
        col = (slice(None),) + (None,) * (x.dim() - 1)
        drift = f_coef[col] * x - g[col] ** 2 * score * 0.5
        jt_eps = f_coef[col] * eps - g[col] ** 2 * vjp(eps) * 0.5
        return drift, torch.sum(jt_eps * eps, dim=tuple(range(1, x.dim())))

    return drift_div


def get_likelihood_fn(sde, inverse_scaler, hutchinson_type='Rademacher', rtol=1e-5, atol=1e-5, method='RK45', eps=1e-5,
                      solver='device'):
    """likelihood.py:40-113: returns likelihood_fn(model, data) -> (bpd [B], z, nfe)."""
    drift_div = get_drift_div_fn(sde)

    def likelihood_fn(model, data):
        with torch.no_grad():
            shape = data.shape
            if hutchinson_type == 'Gaussian':
                epsilon = torch.randn_like(data)
            elif hutchinson_type == 'Rademacher':
                epsilon = torch.randint_like(data, low=0, high=2).float() * 2 - 1.
            else:
                raise NotImplementedError(f"Hutchinson type {hutchinson_type} unknown.")

            nb = shape[0]
            if solver == 'device' and method in ('RK45', 'RK23'):
                from . import ode

                def ode_func(t, x):
                    sample = x[:-nb].reshape(shape)
                    vec_t = torch.ones(nb, device=sample.device) * t
                    drift, div = drift_div(model, sample, vec_t, epsilon)
                    return torch.cat([drift.reshape(-1), div.reshape(-1).to(torch.float32)])

                init = torch.cat([data.reshape(-1).to(torch.float32), torch.zeros(nb, device=data.device)])
                solution = ode.solve_ivp(ode_func, (eps, sde.T), init, rtol=rtol, atol=atol, method=method)
                zp = solution.y_final
                z = zp[:-nb].reshape(shape).type(torch.float32)
                delta_logp = zp[-nb:].type(torch.float32)
            elif solver in ('device', 'scipy'):
                def ode_func(t, x):
                    sample = mutils.from_flattened_numpy(x[:-nb], shape).to(data.device).type(torch.float32)
                    vec_t = torch.ones(sample.shape[0], device=sample.device) * t
                    drift, div = drift_div(model, sample, vec_t, epsilon)
                    return np.concatenate([mutils.to_flattened_numpy(drift), mutils.to_flattened_numpy(div)], axis=0)

                init = np.concatenate([mutils.to_flattened_numpy(data), np.zeros((nb,))], axis=0)
                solution = integrate.solve_ivp(ode_func, (eps, sde.T), init, rtol=rtol, atol=atol, method=method)
                zp = solution.y[:, -1]
                z = mutils.from_flattened_numpy(zp[:-nb], shape).to(data.device).type(torch.float32)
                delta_logp = mutils.from_flattened_numpy(zp[-nb:], (nb,)).to(data.device).type(torch.float32)
            else:
                raise ValueError("get_likelihood_fn: solver must be 'device' or 'scipy' (got %r)" % (solver,))
            nfe = solution.nfev
            prior_logp = sde.prior_logp(z)
            bpd = -(prior_logp + delta_logp) / np.log(2)
            N = np.prod(shape[1:])
            bpd = bpd / N
            offset = 7. - inverse_scaler(-1.)      # converts log-likelihoods to bits/dim (likelihood.py:108-110)
            bpd = bpd + offset
            return bpd, z, nfe

    return likelihood_fn
