"""Configurations of the BASELINE.json workloads, as attribute dictionaries with the reference's field names.

The reference builds `ml_collections.ConfigDict`s in configs/**/*.py; the hot path only READS attributes
(`config.model.nf`, `config.sampling.snr`, ...), so any attribute-style object works -- including the reference's own
ConfigDicts.  `get_config('ve/cifar10_ncsnpp_continuous')` restates the values of the reference file of the same
name (cited per entry) without needing `ml_collections`.
"""
import copy

import torch


class ConfigDict(dict):
    """Attribute dictionary (what the hot path needs from ml_collections.ConfigDict)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def __deepcopy__(self, memo):
        return ConfigDict({k: copy.deepcopy(v, memo) for k, v in self.items()})


def _cd(**kw):
    return ConfigDict(kw)


def _device():
    return torch.device("cuda:0") if torch.cuda.is_available() else torch.device("cpu")


def _base(image_size, channels, batch, snr, sigma_max, num_scales, dropout, dataset):
    return _cd(
        training=_cd(batch_size=batch, continuous=True, reduce_mean=False, sde="vesde", likelihood_weighting=False),
        sampling=_cd(method="pc", predictor="reverse_diffusion", corrector="langevin", n_steps_each=1,
                     noise_removal=True, probability_flow=False, snr=snr),
        eval=_cd(batch_size=1024),
        # configs/default_cifar10_configs.py:58-67 (identical in the lsun / nc defaults)
        optim=_cd(weight_decay=0, optimizer="Adam", lr=2e-4, beta1=0.9, eps=1e-8, warmup=5000, grad_clip=1.),
        data=_cd(dataset=dataset, image_size=image_size, num_channels=channels, centered=False, random_flip=True,
                 uniform_dequantization=False),
        model=_cd(sigma_min=0.01, sigma_max=sigma_max, num_scales=num_scales, beta_min=0.1, beta_max=20.,
                  dropout=dropout, embedding_type="fourier"),
        seed=42, device=_device())


def _ncsnpp_common(m):
    m.update(name="ncsnpp", ema_rate=0.999, normalization="GroupNorm", nonlinearity="swish", nf=128,
             attn_resolutions=(16,), resamp_with_conv=True, conditional=True, fir=True, fir_kernel=[1, 3, 3, 1],
             skip_rescale=True, resblock_type="biggan", progressive_combine="sum", attention_type="ddpm",
             init_scale=0., fourier_scale=16, conv_size=3)


def _cifar10_ve_ncsnpp():
    """configs/ve/cifar10_ncsnpp_continuous.py:21-56 over configs/default_cifar10_configs.py:22-82 (BASELINE config C3)."""
    c = _base(32, 3, 128, 0.16, 50, 1000, 0.1, "CIFAR10")
    _ncsnpp_common(c.model)
    c.model.update(scale_by_sigma=True, ch_mult=(1, 2, 2, 2), num_res_blocks=4, progressive="none",
                   progressive_input="residual")
    return c


def _cifar10_vp_ncsnpp():
    """configs/vp/cifar10_ncsnpp_continuous.py (BASELINE config C1, "DDPM++ ... Euler-Maruyama ... upfirdn2d" reading)."""
    c = _cifar10_ve_ncsnpp()
    c.training.update(sde="vpsde", continuous=True, reduce_mean=True)
    c.sampling.update(predictor="euler_maruyama", corrector="none")
    c.data.centered = True
    c.model.update(scale_by_sigma=False, ema_rate=0.9999, embedding_type="positional")
    return c


def _celebahq_256_ve_ncsnpp():
    """configs/ve/celebahq_256_ncsnpp_continuous.py:21-62 over configs/default_lsun_configs.py (BASELINE config C5)."""
    c = _base(256, 3, 64, 0.075, 348, 2000, 0., "CelebAHQ")
    _ncsnpp_common(c.model)
    c.model.update(scale_by_sigma=True, ch_mult=(1, 1, 2, 2, 2, 2, 2), num_res_blocks=2, progressive="output_skip",
                   progressive_input="input_skip")
    return c


def _nc_ddpm():
    """configs/vp/nc_ddpmpp.py:22-66 over configs/default_nc_configs.py (BASELINE config C1, literal reading)."""
    c = _base(64, 1, 64, 0.075, 378, 2000, 0., "NC")
    c.training.update(sde="vpsde", continuous=False, reduce_mean=True)
    c.sampling.update(predictor="ancestral_sampling", corrector="none")
    c.data.random_flip = False
    c.model.update(name="ddpm", scale_by_sigma=False, ema_rate=0.9999, normalization="GroupNorm",
                   nonlinearity="swish", nf=128, ch_mult=(1, 2, 2, 2), num_res_blocks=4, attn_resolutions=(16,),
                   resamp_with_conv=True, conditional=True, fir=False, fir_kernel=[1, 3, 3, 1], skip_rescale=True,
                   resblock_type="biggan", progressive="none", progressive_input="none", progressive_combine="sum",
                   attention_type="ddpm", init_scale=0., embedding_type="positional", fourier_scale=16, conv_size=3)
    return c


def _nc_ddpm_inpaint_dps():
    """configs/inverse/nc_ddpmpp_inpaint_dps.py (BASELINE config C4): the C1 network, continuous score, B = 16."""
    c = _nc_ddpm()
    c.training.continuous = True
    c.eval.batch_size = 16
    c.inverse = _cd(operator="inpaint", ratio=0.5, variance=0.1, solver="RK45", sampler="dps")
    return c


def _tiny_ncsnpp(fir=True, progressive="none", progressive_input="residual", embedding="fourier", image=16):
    """Not in the reference: a narrow NCSN++ (nf=32) exercising every block type, sized for golden fixtures."""
    c = _cifar10_ve_ncsnpp()
    c.data.image_size = image
    c.model.update(nf=32, ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=(image // 2,), fir=fir,
                   progressive=progressive, progressive_input=progressive_input, embedding_type=embedding,
                   num_scales=1000)
    return c


_REGISTRY = {
    "ve/cifar10_ncsnpp_continuous": _cifar10_ve_ncsnpp,
    "vp/cifar10_ncsnpp_continuous": _cifar10_vp_ncsnpp,
    "ve/celebahq_256_ncsnpp_continuous": _celebahq_256_ve_ncsnpp,
    "vp/nc_ddpmpp": _nc_ddpm,
    "inverse/nc_ddpmpp_inpaint_dps": _nc_ddpm_inpaint_dps,
    "test/tiny_ncsnpp": _tiny_ncsnpp,
}


def get_config(name):
    name = name.replace("configs/", "").replace(".py", "").replace(".", "/")
    if name not in _REGISTRY:
        raise KeyError("unknown config %r; have %s" % (name, sorted(_REGISTRY)))
    return _REGISTRY[name]()


def available():
    return sorted(_REGISTRY)
