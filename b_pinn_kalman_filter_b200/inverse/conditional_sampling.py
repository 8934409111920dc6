"""Diffusion posterior sampling (reference inverse/conditional_sampling.py:100-165) on the B200 engine.

Same structure and names as the reference (`get_solver`, `get_sampler`, `get_dps_sampler` with its `drift_fn`,
`x0_hat_fn`, `cond_grad_fn`, `ode_func`); the one change is how d logp / d x_t is obtained.  The reference calls
`torch.autograd.grad(logp, x_t)` through x0_hat = x_t / mean + std^2 score(x_t); the engine has no autograd graph, so
the chain rule is written out:   d logp/d x_t = v / mean + std^2 * J_score^T v,   v = d logp/d x0_hat
with v in closed form for a linear operator (logp = -||y - A x0_hat||^2 / var  =>  v = 2 A^T (y - A x0_hat) / var) and
J_score^T v from ScoreEngine.input_grad.
"""
from functools import partial

import torch
from scipy import integrate

from ..models.utils import from_flattened_numpy, get_score_vjp_fn, to_flattened_numpy


def get_solver(config, ode_func, x0, t1, shape, eps):
    """inverse/conditional_sampling.py:10-29.  Two calling conventions:
      * numpy x0 (the reference's): ode_func(t, ndarray) -> ndarray, scipy / host Euler loop as in the reference;
      * CUDA tensor x0: ode_func(t, fp32 tensor [n]) -> fp32 tensor [n]; 'RK45' / 'RK23' run on ode.solve_ivp and 'fixed'
        is an in-place device loop -- the state never leaves HBM."""
    on_device = torch.is_tensor(x0)
    if config.inverse.solver in ['RK45', 'RK23']:
        if on_device:
            from .. import ode
            solution = ode.solve_ivp(ode_func, (t1, eps), x0, rtol=1e-3, atol=1e-3, method=config.inverse.solver)
            return solution.y_final.reshape(shape).to(config.device).type(torch.float32)
        solution = integrate.solve_ivp(ode_func, (t1, eps), x0, rtol=1e-3, atol=1e-3, method=config.inverse.solver)
        return torch.tensor(solution.y[:, -1]).reshape(shape).to(config.device).type(torch.float32)
    elif config.inverse.solver == 'fixed':
        dt = -.00002  # inverse ODE
        if on_device:
            x = x0.detach().reshape(-1).to(torch.float32).clone()
            for t in torch.linspace(t1, eps, 5000).tolist():
                x.add_(ode_func(t, x), alpha=dt)
            return x.reshape(shape).to(config.device)
        x = x0
        for t in torch.linspace(t1, eps, 5000):
            x += ode_func(float(t), x) * dt
        return torch.tensor(x).reshape(shape).to(config.device).type(torch.float32)
    raise NotImplementedError(config.inverse.solver)


def get_sampler(config, obsv_sde, shape, eps=1e-3):
    if config.inverse.sampler == 'dps':
        return get_dps_sampler(config, obsv_sde, shape, eps=eps)
    raise NotImplementedError("sampler %r: only 'dps' is on the B200 path" % config.inverse.sampler)


def get_dps_functions(obsv_sde, observation, obsv_var, shape):
    """(x0_hat_fn, cond_grad_fn, drift_fn, rhs) of the reference's dps sampler as plain functions of (model, x, t)."""
    sde = obsv_sde.state_sde

    def drift_fn(score, score_cond, x, t):
        drift, diffusion = sde.sde(x, t)
        return drift - diffusion[:, None, None, None] ** 2 * (score + score_cond) * 0.5

    def x0_hat_fn(model, xt, t):
        score, vjp = get_score_vjp_fn(sde, model, continuous=True)(xt, t)
        mean, std = sde.marginal_coef(t)
        x0_hat = xt / mean[:, None, None, None] + std[:, None, None, None] ** 2 * score
        return x0_hat, score, vjp, mean, std

    def cond_grad_fn(x0_hat, vjp, mean, std, scale=True):
        difference = observation - obsv_sde.operator(x0_hat, keep_shape=False)
        norm = torch.linalg.norm(difference)
        v = obsv_sde.operator.adjoint(difference, x0_hat.shape) * (2.0 / obsv_var)      # d(-norm^2/var)/d x0_hat
        norm_grad = v / mean[:, None, None, None] + std[:, None, None, None] ** 2 * vjp(v)
        if scale is True:
            norm_grad = norm_grad / norm
        return norm_grad

    def rhs(model, x_hat, vec_t):
        x0_hat, score, vjp, mean, std = x0_hat_fn(model, x_hat, vec_t)
        score_cond = cond_grad_fn(x0_hat, vjp, mean, std)
        return drift_fn(score, score_cond, x_hat, vec_t)

    return x0_hat_fn, cond_grad_fn, drift_fn, rhs


def get_dps_sampler(config, obsv_sde, shape, eps=1e-3):
    '''diffusion posterior sampling (inverse/conditional_sampling.py:100-165)'''
    device = config.device
    obsv_var = config.inverse.variance
    observation = obsv_sde.y0 + torch.randn_like(obsv_sde.y0, device=device) * obsv_var ** .5
    _, _, _, rhs = get_dps_functions(obsv_sde, observation, obsv_var, shape)

    def dps_sampler(model, z=None, solver='device'):
        x = obsv_sde.state_sde.prior_sampling(shape).to(device) if z is None else z
        solve = partial(get_solver, config=config, shape=shape, eps=eps)
        if solver == 'device':
            def ode_func(t, x):
                vec_t = torch.ones(shape[0], device=x.device) * t
                return rhs(model, x.reshape(shape), vec_t).reshape(-1)

            return solve(ode_func=ode_func, x0=x.to(torch.float32), t1=obsv_sde.state_sde.T)

        def ode_func(t, x):
            x_hat = from_flattened_numpy(x, shape).to(device).type(torch.float32)
            vec_t = torch.ones(shape[0], device=device) * t
            return to_flattened_numpy(rhs(model, x_hat, vec_t))

        return solve(ode_func=ode_func, x0=to_flattened_numpy(x), t1=obsv_sde.state_sde.T)

    return dps_sampler
