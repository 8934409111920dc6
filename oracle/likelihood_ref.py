"""TEST INFRASTRUCTURE ONLY -- one right-hand side of the reference's likelihood ODE (likelihood.py:27-37, 61-69) on
the CPU by autograd through the oracle DDPM: drift of the probability-flow ODE and the Hutchinson-Skilling divergence
estimate sum(grad_x(sum(drift * eps)) * eps)."""
import torch


def drift_div(net, sde, x, t, eps):
    assert sde.kind == "vpsde" and net.m.name == "ddpm"
    x = x.detach().clone().to(net.dtype).requires_grad_(True)
    t = t.to(net.dtype)
    std = sde.marginal_std(t)
    score = -net.ddpm(x, t * 999) / std[:, None, None, None]
    f, g = sde.drift_diffusion(x, t)
    drift = f - g[:, None, None, None] ** 2 * score * 0.5        # rsde.sde with probability_flow=True (sde_lib.py:103-110)
    fn_eps = torch.sum(drift * eps.to(net.dtype))
    grad_fn_eps, = torch.autograd.grad(fn_eps, x)
    return drift.detach(), torch.sum(grad_fn_eps * eps, dim=(1, 2, 3))
