"""Forward SDEs (VP / sub-VP / VE) and their reverse-time counterparts, with the reference's interface
(sde_lib.py:7-119, 136-307): `T, N, sde, coefficient, marginal_coef, marginal_prob, prior_sampling, prior_logp,
discretize, reverse`.

Two differences from the reference, both deliberate:
  * `VESDE.marginal_coef` exists (the reference declares it abstract and never implements it for VE, so its VESDE
    cannot be instantiated: sde_lib.py:34-37 vs :253-307);
  * every SDE can also emit its per-step scalar schedule (`schedule(...)`) so that the fused sm_100a update kernel
    (csrc/sde_update.cu) replaces the per-step tensor arithmetic; the tensor-returning methods remain for user-defined
    predictors/correctors.
Only scalar / [B]-vector math lives here; everything [B,C,H,W]-sized runs in the CUDA kernels.
"""
import abc
import math

import numpy as np
import torch


class SDE(abc.ABC):
    """SDE abstract class. Functions are designed for a mini-batch of inputs."""

    def __init__(self, N):
        super().__init__()
        self.N = N

    @property
    @abc.abstractmethod
    def T(self):
        """End time of the SDE."""

    @abc.abstractmethod
    def sde(self, x, t):
        """(drift, diffusion) of the forward SDE at (x, t)."""

    @abc.abstractmethod
    def coefficient(self, t):
        """Linear SDEs only: (drift_coef, diffusion_coef) with drift = drift_coef * x."""

    @abc.abstractmethod
    def marginal_coef(self, t):
        """(mean scale, std) of p_t(x | x_0)."""

    @abc.abstractmethod
    def marginal_prob(self, x, t):
        """(mean, std) of p_t(x | x_0 = x)."""

    @abc.abstractmethod
    def prior_sampling(self, shape):
        """One sample from p_T."""

    @abc.abstractmethod
    def prior_logp(self, z):
        """log p_T(z)."""

    def discretize(self, x, t):
        """x_{i+1} = x_i + f_i(x_i) + G_i z_i; Euler-Maruyama by default (sde_lib.py:62-79)."""
        dt = 1 / self.N
        drift, diffusion = self.sde(x, t)
        return drift * dt, diffusion * math.sqrt(dt)

    def reverse(self, score_fn, probability_flow=False):
        """Reverse-time SDE/ODE (sde_lib.py:81-119).  The result is an instance of a subclass of type(self), so
        `isinstance(rsde, VESDE)`-style checks keep working."""
        fwd = self

        class RSDE(type(self)):
            def __init__(rs):
                rs.N = fwd.N
                rs.probability_flow = probability_flow

            @property
            def T(rs):
                return fwd.T

            def sde(rs, x, t):
                drift, diffusion = fwd.sde(x, t)
                weight = 0.5 if rs.probability_flow else 1.
                drift = drift - _col(diffusion) ** 2 * score_fn(x, t) * weight
                return drift, (0. if rs.probability_flow else diffusion)

            def discretize(rs, x, t):
                f, G = fwd.discretize(x, t)
                weight = 0.5 if rs.probability_flow else 1.
                rev_f = f - _col(G) ** 2 * score_fn(x, t) * weight
                return rev_f, (torch.zeros_like(G) if rs.probability_flow else G)

        return RSDE()


def _col(v):
    return v[:, None, None, None]


def _log_normal_prior(z, var):
    n = float(np.prod(z.shape[1:]))
    return -n / 2. * math.log(2 * math.pi * var) - torch.sum(z ** 2, dim=(1, 2, 3)) / (2 * var)


class _BetaSchedule(SDE):
    """Shared by VP and sub-VP: beta(t) = beta_0 + t (beta_1 - beta_0)."""

    def __init__(self, beta_min, beta_max, N):
        super().__init__(N)
        self.beta_0 = beta_min
        self.beta_1 = beta_max
        self.N = N

    @property
    def T(self):
        return 1

    def _beta(self, t):
        return self.beta_0 + t * (self.beta_1 - self.beta_0)

    def _log_mean(self, t):
        return -0.25 * t ** 2 * (self.beta_1 - self.beta_0) - 0.5 * t * self.beta_0

    def sde(self, x, t):
        drift_coef, diffusion_coef = self.coefficient(t)
        return _col(drift_coef) * x, diffusion_coef

    def marginal_prob(self, x, t):
        mean, std = self.marginal_coef(t)
        return _col(mean) * x, std

    def prior_sampling(self, shape):
        return torch.randn(*shape)

    def prior_logp(self, z):
        return _log_normal_prior(z, 1.0)


class VPSDE(_BetaSchedule):
    def __init__(self, beta_min=0.1, beta_max=20, N=1000):
        """Variance-preserving SDE (sde_lib.py:136-199)."""
        super().__init__(beta_min, beta_max, N)
        self.discrete_betas = torch.linspace(beta_min / N, beta_max / N, N)
        self.alphas = 1. - self.discrete_betas
        self.alphas_cumprod = torch.cumprod(self.alphas, dim=0)
        self.sqrt_alphas_cumprod = torch.sqrt(self.alphas_cumprod)
        self.sqrt_1m_alphas_cumprod = torch.sqrt(1. - self.alphas_cumprod)

    def coefficient(self, t):
        beta_t = self._beta(t)
        return -0.5 * beta_t, torch.sqrt(beta_t)

    def marginal_coef(self, t):
        lm = self._log_mean(t)
        return torch.exp(lm), torch.sqrt(1. - torch.exp(2. * lm))

    def discretize(self, x, t):
        """DDPM discretisation (sde_lib.py:191-199)."""
        i = (t * (self.N - 1) / self.T).long()
        beta = self.discrete_betas.to(x.device)[i]
        alpha = self.alphas.to(x.device)[i]
        return _col(torch.sqrt(alpha)) * x - x, torch.sqrt(beta)


class subVPSDE(_BetaSchedule):
    def __init__(self, beta_min=0.1, beta_max=20, N=1000):
        """Sub-VP SDE (sde_lib.py:202-250)."""
        super().__init__(beta_min, beta_max, N)

    def coefficient(self, t):
        beta_t = self._beta(t)
        discount = 1. - torch.exp(-2 * self.beta_0 * t - (self.beta_1 - self.beta_0) * t ** 2)
        return -0.5 * beta_t, torch.sqrt(beta_t * discount)

    def marginal_coef(self, t):
        lm = self._log_mean(t)
        return torch.exp(lm), 1 - torch.exp(2. * lm)


class VESDE(SDE):
    def __init__(self, sigma_min=0.01, sigma_max=50, N=1000):
        """Variance-exploding SDE (sde_lib.py:253-307)."""
        super().__init__(N)
        self.sigma_min = sigma_min
        self.sigma_max = sigma_max
        self.discrete_sigmas = torch.exp(torch.linspace(np.log(self.sigma_min), np.log(self.sigma_max), N))
        self.N = N

    @property
    def T(self):
        return 1

    def _sigma(self, t):
        return self.sigma_min * (self.sigma_max / self.sigma_min) ** t

    def sde(self, x, t):
        drift_coef, diffusion_coef = self.coefficient(t)
        return _col(drift_coef), diffusion_coef

    def coefficient(self, t):
        g = self._sigma(t) * math.sqrt(2 * (np.log(self.sigma_max) - np.log(self.sigma_min)))
        return torch.zeros_like(t), g

    def marginal_coef(self, t):
        return torch.ones_like(t), self._sigma(t)

    def marginal_prob(self, x, t):
        return x, self._sigma(t)

    def prior_sampling(self, shape):
        return torch.randn(*shape) * self.sigma_max

    def prior_logp(self, z):
        return _log_normal_prior(z, self.sigma_max ** 2)

    def discretize(self, x, t):
        """SMLD (NCSN) discretisation (sde_lib.py:299-307)."""
        i = (t * (self.N - 1) / self.T).long()
        table = self.discrete_sigmas.to(t.device)
        sigma = table[i]
        adjacent = torch.where(i == 0, torch.zeros_like(t), table[i - 1])
        return torch.zeros_like(x), torch.sqrt(sigma ** 2 - adjacent ** 2)


class OBSVSDE(SDE):
    """Observation process y_t of a hidden state following `state_sde` (reference sde_lib.py:122-133)."""

    def __init__(self, N, y0, operator):
        super().__init__(N)
        self.y0 = y0
        self.operator = operator


class LOBSVSDE(OBSVSDE):
    """Linear observation SDE (reference sde_lib.py:310-352): y_t = alpha(t) y_0 + beta(t) A z.  Only what the DPS
    sampler reads: `state_sde`, `y0`, `operator`, `observe_sampling`."""

    def __init__(self, state_sde, y0, operator):
        super().__init__(state_sde.N, y0, operator)
        self.state_sde = state_sde

    @property
    def T(self):
        return 1

    def observe_sampling(self, z, t):
        alpha, beta = self.state_sde.marginal_coef(t)
        return alpha[:, None, None] * self.y0 + beta[:, None, None] * self.operator(z, False)

    def sde(self, x, t):
        return self.state_sde.sde(x, t)

    def coefficient(self, t):
        return self.state_sde.coefficient(t)

    def marginal_coef(self, t):
        return self.state_sde.marginal_coef(t)

    def marginal_prob(self, x, t):
        return self.state_sde.marginal_prob(x, t)

    def prior_sampling(self, shape):
        return None

    def prior_logp(self, z):
        return None
