"""CPU oracle for the SDEs, score_fn and the PC sampler (test infrastructure only).

Restates in plain PyTorch-on-CPU, as straight-line functions, what the reference spreads over
sde_lib.py:136-307 (VP / sub-VP / VE), models/utils.py:129-178 (score_fn), sampling.py:176-330 (predictors and
correctors) and sampling.py:390-409 (the PC loop).  Noise is taken from a caller-supplied iterator so that the
reference, this oracle and the CUDA path can be driven by the SAME noise tensors.
Pinned against the real reference in tests/test_oracle_pins.py and by tests/golden/pc_*.npz.
"""
import math

import numpy as np
import torch


class RefSDE:
    def __init__(self, kind, N, beta_min=0.1, beta_max=20., sigma_min=0.01, sigma_max=50.):
        self.kind, self.N, self.T = kind, N, 1
        self.beta_0, self.beta_1 = beta_min, beta_max
        self.sigma_min, self.sigma_max = sigma_min, sigma_max
        if kind == "vpsde":
            self.discrete_betas = torch.linspace(beta_min / N, beta_max / N, N)
            self.alphas = 1. - self.discrete_betas
            self.sqrt_1m_alphas_cumprod = torch.sqrt(1. - torch.cumprod(self.alphas, dim=0))
        if kind == "vesde":
            self.discrete_sigmas = torch.exp(torch.linspace(np.log(sigma_min), np.log(sigma_max), N))

    @staticmethod
    def from_config(config):
        m = config.model
        return RefSDE(config.training.sde.lower(), m.num_scales, m.beta_min, m.beta_max, m.sigma_min, m.sigma_max)

    @property
    def eps(self):
        return 1e-5 if self.kind == "vesde" else 1e-3   # run_lib.py:45-58

    def beta(self, t):
        return self.beta_0 + t * (self.beta_1 - self.beta_0)

    def marginal_std(self, t):
        if self.kind == "vesde":
            return self.sigma_min * (self.sigma_max / self.sigma_min) ** t
        lm = -0.25 * t ** 2 * (self.beta_1 - self.beta_0) - 0.5 * t * self.beta_0
        if self.kind == "vpsde":
            return torch.sqrt(1. - torch.exp(2. * lm))
        return 1 - torch.exp(2. * lm)

    def drift_diffusion(self, x, t):
        """forward sde(x,t): sde_lib.py:159-169, 220-231, 272-283"""
        if self.kind == "vesde":
            sigma = self.sigma_min * (self.sigma_max / self.sigma_min) ** t
            g = sigma * torch.sqrt(torch.tensor(2 * (np.log(self.sigma_max) - np.log(self.sigma_min))))
            return torch.zeros_like(t)[:, None, None, None], g
        b = self.beta(t)
        drift = (-0.5 * b)[:, None, None, None] * x
        if self.kind == "vpsde":
            return drift, torch.sqrt(b)
        discount = 1. - torch.exp(-2 * self.beta_0 * t - (self.beta_1 - self.beta_0) * t ** 2)
        return drift, torch.sqrt(b * discount)

    def discretize(self, x, t):
        """sde_lib.py:62-79 (default), :191-199 (VP), :299-307 (VE)"""
        i = (t * (self.N - 1) / self.T).long()
        if self.kind == "vpsde":
            beta, alpha = self.discrete_betas[i], self.alphas[i]
            return torch.sqrt(alpha)[:, None, None, None] * x - x, torch.sqrt(beta)
        if self.kind == "vesde":
            sigma = self.discrete_sigmas[i]
            adj = torch.where(i == 0, torch.zeros_like(t), self.discrete_sigmas[i - 1])
            return torch.zeros_like(x), torch.sqrt(sigma ** 2 - adj ** 2)
        dt = 1 / self.N
        drift, g = self.drift_diffusion(x, t)
        return drift * dt, g * torch.sqrt(torch.tensor(dt))


def make_score_fn(sde, net, continuous):
    """models/utils.py:144-173.  net(x, labels) -> model output (e.g. oracle.score_net_ref.RefScoreNet)."""
    def score_fn(x, t):
        if sde.kind in ("vpsde", "subvpsde"):
            if continuous or sde.kind == "subvpsde":
                labels = t * 999
                std = sde.marginal_std(t)
            else:
                labels = t * (sde.N - 1)
                std = sde.sqrt_1m_alphas_cumprod[labels.long()]
            return -net(x, labels) / std[:, None, None, None]
        if continuous:
            labels = sde.marginal_std(t)
        else:
            labels = torch.round((sde.T - t) * (sde.N - 1)).long()
        return net(x, labels)
    return score_fn


def col(v):
    return v[:, None, None, None]


def predictor_step(name, sde, score_fn, x, t, z, probability_flow=False):
    """One predictor update -> (x, x_mean).  sampling.py:181-187, 195-200, 213-233."""
    w = 0.5 if probability_flow else 1.
    if name == "none":
        return x, x
    if name == "euler_maruyama":
        dt = -1. / sde.N
        drift, g = sde.drift_diffusion(x, t)
        drift = drift - col(g) ** 2 * score_fn(x, t) * w
        g = torch.zeros_like(g) if probability_flow else g
        x_mean = x + drift * dt
        return x_mean + col(g) * np.sqrt(-dt) * z, x_mean
    if name == "reverse_diffusion":
        f, G = sde.discretize(x, t)
        f = f - col(G) ** 2 * score_fn(x, t) * w
        G = torch.zeros_like(G) if probability_flow else G
        x_mean = x - f
        return x_mean + col(G) * z, x_mean
    if name == "ancestral_sampling":
        i = (t * (sde.N - 1) / sde.T).long()
        if sde.kind == "vesde":
            sigma = sde.discrete_sigmas[i]
            adj = torch.where(i == 0, torch.zeros_like(t), sde.discrete_sigmas[i - 1])
            x_mean = x + score_fn(x, t) * col(sigma ** 2 - adj ** 2)
            std = torch.sqrt((adj ** 2 * (sigma ** 2 - adj ** 2)) / (sigma ** 2))
            return x_mean + col(std) * z, x_mean
        beta = sde.discrete_betas[i]
        x_mean = (x + col(beta) * score_fn(x, t)) / col(torch.sqrt(1. - beta))
        return x_mean + col(torch.sqrt(beta)) * z, x_mean
    raise KeyError(name)


def corrector_step(name, sde, score_fn, x, t, noises, snr, n_steps):
    """n_steps corrector updates -> (x, x_mean).  sampling.py:262-282, 299-319.  `noises`: iterator of z tensors."""
    if name == "none":
        return x, x
    if sde.kind in ("vpsde", "subvpsde") and hasattr(sde, "alphas"):
        alpha = sde.alphas[(t * (sde.N - 1) / sde.T).long()]
    else:
        alpha = torch.ones_like(t)
    x_mean = x
    for _ in range(n_steps):
        grad = score_fn(x, t)
        z = next(noises)
        if name == "langevin":
            grad_norm = torch.norm(grad.reshape(grad.shape[0], -1), dim=-1).mean()
            noise_norm = torch.norm(z.reshape(z.shape[0], -1), dim=-1).mean()
            step = (snr * noise_norm / grad_norm) ** 2 * 2 * alpha
        elif name == "ald":
            step = (snr * sde.marginal_std(t)) ** 2 * 2 * alpha
        else:
            raise KeyError(name)
        x_mean = x + col(step) * grad
        x = x_mean + col(torch.sqrt(step * 2)) * z
    return x, x_mean


def pc_sample(sde, net, shape, predictor, corrector, snr, n_steps, continuous, noises, denoise=True, eps=None,
              probability_flow=False, n_iters=None, return_trajectory=False):
    """The PC loop (sampling.py:390-409).  `noises` yields: prior z, then per step n_steps corrector z's followed by
    one predictor z (only for the stages that draw noise in the reference)."""
    eps = sde.eps if eps is None else eps
    score_fn = make_score_fn(sde, net, continuous)
    noises = iter(noises)
    x = next(noises) * (sde.sigma_max if sde.kind == "vesde" else 1.0)
    timesteps = torch.linspace(sde.T, eps, sde.N)
    x_mean = x
    traj = []
    for i in range(sde.N if n_iters is None else n_iters):
        vec_t = torch.ones(shape[0]) * timesteps[i]
        x, x_mean = corrector_step(corrector, sde, score_fn, x, vec_t, noises, snr, n_steps)
        z = next(noises) if predictor != "none" else None
        x, x_mean = predictor_step(predictor, sde, score_fn, x, vec_t, z, probability_flow)
        if return_trajectory:
            traj.append(x.clone())
    out = x_mean if denoise else x
    return (out, traj) if return_trajectory else out
