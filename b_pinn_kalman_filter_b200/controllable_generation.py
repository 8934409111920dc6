"""Inpainting and colorisation with PC samplers, reference signatures (controllable_generation.py:8-180).

After every predictor / corrector update the known part of the image is replaced by freshly noised data,
`x (1 - m) + (mean_t(data) + std_t z) m` (controllable_generation.py:44-54, 132-141).  With a B200 score model and the
built-in predictor / corrector classes the whole thing runs inside `sampling.FusedPCLoop`: the mix is one fused kernel
(`b200_mask_mix`, colour-basis change included) in the step's launch sequence, the schedule lives in device tables and
the step replays as a CUDA graph.  User-registered classes take the generic loop below (their `update_fn` + the same
arithmetic in torch).  Noise comes through `sampling._noise` so that tests can inject it."""
import functools

import torch

from . import sampling
from .sampling import shared_corrector_update_fn, shared_predictor_update_fn


def _randn_like(x):
    z = sampling._noise(x.shape, x.device)
    return torch.randn_like(x) if z is None else z


def _prior(sde, shape, device):
    z = sampling._noise(shape, device)
    if z is None:
        return sde.prior_sampling(shape).to(device)
    from . import sde_lib
    return z * (sde.sigma_max if isinstance(sde, sde_lib.VESDE) else 1.0)


def _fused(model, sde, predictor, corrector, snr, n_steps, probability_flow, continuous, eps, denoise, x0, mix):
    loop = sampling.FusedPCLoop(model, sde, tuple(x0.shape), predictor, corrector, snr, n_steps, probability_flow,
                                continuous, eps, mix=mix)
    with torch.cuda.device(loop.dev):
        loop.reset(x0.to(loop.dev))
        loop.run(sde.N)
        return (loop.x_mean if denoise else loop.x).clone()


def _update_fns(sde, predictor, corrector, snr, n_steps, probability_flow, continuous):
    p = functools.partial(shared_predictor_update_fn, sde=sde, predictor=predictor, probability_flow=probability_flow,
                          continuous=continuous)
    c = functools.partial(shared_corrector_update_fn, sde=sde, corrector=corrector, continuous=continuous, snr=snr,
                          n_steps=n_steps)
    return p, c


def get_pc_inpainter(sde, predictor, corrector, inverse_scaler, snr, n_steps=1, probability_flow=False,
                     continuous=False, denoise=True, eps=1e-5):
    """controllable_generation.py:8-84.  Returns pc_inpainter(model, data, mask); mask 1 = known pixel."""
    predictor_update_fn, corrector_update_fn = _update_fns(sde, predictor, corrector, snr, n_steps, probability_flow,
                                                           continuous)

    def get_inpaint_update_fn(update_fn):
        def inpaint_update_fn(model, data, mask, x, t):
            with torch.no_grad():
                vec_t = torch.ones(data.shape[0], device=data.device) * t
                x, x_mean = update_fn(x, vec_t, model=model)
                masked_data_mean, std = sde.marginal_prob(data, vec_t)
                masked_data = masked_data_mean + _randn_like(x) * std[:, None, None, None]
                x = x * (1. - mask) + masked_data * mask
                x_mean = x * (1. - mask) + masked_data_mean * mask
                return x, x_mean
        return inpaint_update_fn

    projector_inpaint_update_fn = get_inpaint_update_fn(predictor_update_fn)
    corrector_inpaint_update_fn = get_inpaint_update_fn(corrector_update_fn)

    def pc_inpainter(model, data, mask):
        with torch.no_grad():
            x = data * mask + _prior(sde, data.shape, data.device) * (1. - mask)
            if sampling.is_fusable(model, sde, predictor, corrector) and data.dim() == 4 and data.shape[1] <= 4:
                return inverse_scaler(_fused(model, sde, predictor, corrector, snr, n_steps, probability_flow, continuous,
                                             eps, denoise, x, dict(data=data, mask=mask)))
            timesteps = torch.linspace(sde.T, eps, sde.N)
            for i in range(sde.N):
                t = timesteps[i]
                x, x_mean = corrector_inpaint_update_fn(model, data, mask, x, t)
                x, x_mean = projector_inpaint_update_fn(model, data, mask, x, t)
            return inverse_scaler(x_mean if denoise else x)

    return pc_inpainter


def get_pc_colorizer(sde, predictor, corrector, inverse_scaler, snr, n_steps=1, probability_flow=False,
                     continuous=False, denoise=True, eps=1e-5):
    """controllable_generation.py:87-180.  Returns pc_colorizer(model, gray_scale_img)."""
    # orthonormal basis whose first axis is the gray level (controllable_generation.py:109-114)
    M = torch.tensor([[5.7735014e-01, -8.1649649e-01, 4.7008697e-08],
                      [5.7735026e-01, 4.0824834e-01, 7.0710671e-01],
                      [5.7735026e-01, 4.0824822e-01, -7.0710683e-01]])
    invM = torch.inverse(M)

    def decouple(inputs):
        return torch.einsum('bihw,ij->bjhw', inputs, M.to(inputs.device))

    def couple(inputs):
        return torch.einsum('bihw,ij->bjhw', inputs, invM.to(inputs.device))

    predictor_update_fn, corrector_update_fn = _update_fns(sde, predictor, corrector, snr, n_steps, probability_flow,
                                                           continuous)

    def get_mask(image):
        return torch.cat([torch.ones_like(image[:, :1, ...]), torch.zeros_like(image[:, 1:, ...])], dim=1)

    def get_colorization_update_fn(update_fn):
        def colorization_update_fn(model, gray_scale_img, x, t):
            mask = get_mask(x)
            vec_t = torch.ones(x.shape[0], device=x.device) * t
            x, x_mean = update_fn(x, vec_t, model=model)
            masked_data_mean, std = sde.marginal_prob(decouple(gray_scale_img), vec_t)
            masked_data = masked_data_mean + _randn_like(x) * std[:, None, None, None]
            x = couple(decouple(x) * (1. - mask) + masked_data * mask)
            x_mean = couple(decouple(x) * (1. - mask) + masked_data_mean * mask)
            return x, x_mean
        return colorization_update_fn

    predictor_colorize_update_fn = get_colorization_update_fn(predictor_update_fn)
    corrector_colorize_update_fn = get_colorization_update_fn(corrector_update_fn)

    def pc_colorizer(model, gray_scale_img):
        with torch.no_grad():
            shape = gray_scale_img.shape
            mask = get_mask(gray_scale_img)
            x = couple(decouple(gray_scale_img) * mask
                       + decouple(_prior(sde, shape, gray_scale_img.device) * (1. - mask)))
            if sampling.is_fusable(model, sde, predictor, corrector) and len(shape) == 4 and shape[1] == 3:
                return inverse_scaler(_fused(model, sde, predictor, corrector, snr, n_steps, probability_flow, continuous,
                                             eps, denoise, x, dict(data=gray_scale_img, mask=mask, M=M, Minv=invM)))
            timesteps = torch.linspace(sde.T, eps, sde.N)
            for i in range(sde.N):
                t = timesteps[i]
                x, x_mean = corrector_colorize_update_fn(model, gray_scale_img, x, t)
                x, x_mean = predictor_colorize_update_fn(model, gray_scale_img, x, t)
            return inverse_scaler(x_mean if denoise else x)

    return pc_colorizer
