"""TEST INFRASTRUCTURE ONLY -- CPU restatement of one right-hand-side evaluation of the reference's DPS sampler
(inverse/conditional_sampling.py:100-165) with an inpainting mask, by torch autograd through the oracle network.

  x0_hat_fn    :116-125   score = -model(x, 999 t) / std(t);  x0_hat = x / mean + std^2 score
  cond_grad_fn :127-138   difference = y - A x0_hat (observed pixels only);  logp = -||difference||^2 / var;
                          grad = autograd(logp, x_t) / ||difference||
  drift_fn     :108-114   drift = f(x, t) - g(t)^2 (score + score_cond) / 2
The mean of the VP marginal is exp(-t^2 (b1 - b0)/4 - t b0/2) (sde_lib.py:171-177).
"""
import torch


def dps_rhs(net, sde, x, t, mask, observation_full, var):
    """-> (drift, score, score_cond).  `observation_full`: the observation laid out as an image (zeros where
    unobserved); the norm over observed pixels equals the norm of mask * (observation - x0_hat)."""
    assert sde.kind == "vpsde" and net.m.name == "ddpm"
    xt = x.detach().clone().to(net.dtype).requires_grad_(True)
    t = t.to(net.dtype)
    std = sde.marginal_std(t)
    log_mean = -0.25 * t ** 2 * (sde.beta_1 - sde.beta_0) - 0.5 * t * sde.beta_0
    mean = torch.exp(log_mean)
    score = -net.ddpm(xt, t * 999) / std[:, None, None, None]
    x0_hat = xt / mean[:, None, None, None] + std[:, None, None, None] ** 2 * score
    difference = mask * (observation_full - x0_hat)
    norm = torch.linalg.norm(difference)
    logp = -norm ** 2 / var
    norm_grad, = torch.autograd.grad(logp, xt)
    norm_grad = norm_grad / norm.detach()
    drift, diffusion = sde.drift_diffusion(xt.detach(), t)
    score = score.detach()
    drift = drift - diffusion[:, None, None, None] ** 2 * (score + norm_grad) * 0.5
    return drift, score, norm_grad
