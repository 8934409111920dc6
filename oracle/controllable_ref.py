"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's inpainting / colorisation PC loops
(controllable_generation.py:8-84, 87-180) on top of oracle/sampling_ref.py, with injected noise.
Noise order: prior, then per step: corrector z (n_steps of them), mix z, predictor z, mix z."""
import torch

from . import sampling_ref as sr

_M = torch.tensor([[5.7735014e-01, -8.1649649e-01, 4.7008697e-08],
                   [5.7735026e-01, 4.0824834e-01, 7.0710671e-01],
                   [5.7735026e-01, 4.0824822e-01, -7.0710683e-01]])     # controllable_generation.py:109-111


def _marginal(sde, data, t):
    """sde.marginal_prob(data, t) for the VE SDE (sde_lib.py:285-288): mean = data, std = sigma(t)."""
    assert sde.kind == "vesde"
    return data, sde.marginal_std(t)


def pc_inpaint(sde, net, data, mask, predictor, corrector, snr, n_steps, continuous, noises, eps, denoise=True):
    score_fn = sr.make_score_fn(sde, net, continuous)
    noises = iter(noises)
    x = data * mask + next(noises) * sde.sigma_max * (1. - mask)
    timesteps = torch.linspace(sde.T, eps, sde.N)

    def mix(x, vec_t):
        mean, std = _marginal(sde, data, vec_t)
        noisy = mean + next(noises) * std[:, None, None, None]
        x = x * (1. - mask) + noisy * mask
        return x, x * (1. - mask) + mean * mask

    x_mean = x
    for i in range(sde.N):
        vec_t = torch.ones(data.shape[0]) * timesteps[i]
        x, _ = sr.corrector_step(corrector, sde, score_fn, x, vec_t, noises, snr, n_steps)
        x, x_mean = mix(x, vec_t)
        x, _ = sr.predictor_step(predictor, sde, score_fn, x, vec_t, next(noises), False)
        x, x_mean = mix(x, vec_t)
    return x_mean if denoise else x


def pc_colorize(sde, net, gray, predictor, corrector, snr, n_steps, continuous, noises, eps, denoise=True):
    score_fn = sr.make_score_fn(sde, net, continuous)
    noises = iter(noises)
    invM = torch.inverse(_M)
    dec = lambda v: torch.einsum('bihw,ij->bjhw', v, _M)
    cou = lambda v: torch.einsum('bihw,ij->bjhw', v, invM)
    mask = torch.cat([torch.ones_like(gray[:, :1]), torch.zeros_like(gray[:, 1:])], dim=1)
    x = cou(dec(gray) * mask + dec(next(noises) * sde.sigma_max * (1. - mask)))
    timesteps = torch.linspace(sde.T, eps, sde.N)

    def mix(x, vec_t):
        mean, std = _marginal(sde, dec(gray), vec_t)
        noisy = mean + next(noises) * std[:, None, None, None]
        x = cou(dec(x) * (1. - mask) + noisy * mask)
        return x, cou(dec(x) * (1. - mask) + mean * mask)

    x_mean = x
    for i in range(sde.N):
        vec_t = torch.ones(gray.shape[0]) * timesteps[i]
        x, _ = sr.corrector_step(corrector, sde, score_fn, x, vec_t, noises, snr, n_steps)
        x, x_mean = mix(x, vec_t)
        x, _ = sr.predictor_step(predictor, sde, score_fn, x, vec_t, next(noises), False)
        x, x_mean = mix(x, vec_t)
    return x_mean if denoise else x
