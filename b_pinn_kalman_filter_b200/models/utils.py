"""Model registry, model creation and score functions with the reference's API (models/utils.py:24-187).

`create_model(config)` returns a `ScoreModel`: a callable `model(x, labels)` that runs the network on the B200
score engine.  It mimics what the sampling code touches on the reference's `nn.DataParallel(model)`:
`.eval()`, `.train()`, `.module`, `state_dict()`, `load_state_dict()`, `parameters()`.
"""
import math

import numpy as np
import torch

from .. import sde_lib
from .engine import ScoreEngine

_MODELS = {}


def register_model(cls=None, *, name=None):
    """A decorator for registering model builders (same contract as models/utils.py:27-43)."""

    def _register(cls):
        local_name = cls.__name__ if name is None else name
        if local_name in _MODELS:
            raise ValueError(f'Already registered model with name: {local_name}')
        _MODELS[local_name] = cls
        return cls

    if cls is None:
        return _register
    return _register(cls)


def get_model(name):
    return _MODELS[name]


def get_sigmas(config):
    """models/utils.py:50-60"""
    return np.exp(np.linspace(np.log(config.model.sigma_max), np.log(config.model.sigma_min), config.model.num_scales))


def get_ddpm_params(config):
    """models/utils.py:63-85"""
    num_diffusion_timesteps = 1000
    beta_start = config.model.beta_min / config.model.num_scales
    beta_end = config.model.beta_max / config.model.num_scales
    betas = np.linspace(beta_start, beta_end, num_diffusion_timesteps, dtype=np.float64)
    alphas = 1. - betas
    alphas_cumprod = np.cumprod(alphas, axis=0)
    return {
        'betas': betas, 'alphas': alphas, 'alphas_cumprod': alphas_cumprod,
        'sqrt_alphas_cumprod': np.sqrt(alphas_cumprod), 'sqrt_1m_alphas_cumprod': np.sqrt(1. - alphas_cumprod),
        'beta_min': beta_start * (num_diffusion_timesteps - 1), 'beta_max': beta_end * (num_diffusion_timesteps - 1),
        'num_diffusion_timesteps': num_diffusion_timesteps}


def _fans(shape):
    rf = float(np.prod(shape)) / shape[1] / shape[0]
    return shape[1] * rf, shape[0] * rf


def init_tensor(shape, init, gen):
    """Random initialisation with the reference's distributions (models/layers.py:53-86 `default_init` =
    variance_scaling(scale, 'fan_avg', 'uniform') with in_axis=1, out_axis=0; scale 0 -> 1e-10)."""
    if init == "zeros":
        return torch.zeros(shape)
    if init == "ones":
        return torch.ones(shape)
    kind, scale = init
    if kind == "randn":
        return torch.randn(shape, generator=gen) * scale
    scale = 1e-10 if scale == 0 else scale
    if len(shape) == 1:
        fan_in = fan_out = shape[0]
    else:
        fan_in, fan_out = _fans(shape)
    variance = scale / ((fan_in + fan_out) / 2)
    return (torch.rand(shape, generator=gen) * 2. - 1.) * math.sqrt(3 * variance)


class ParameterList(list):
    """`model.parameters()`: a list of tensors that knows which model it came from."""
    model = None


class ScoreModel:
    """The object `create_model` returns; stands where the reference has nn.DataParallel(NCSNpp(config))."""

    def __init__(self, config, device=None, precision=None, seed=None, init=True):
        name = config.model.name
        device = torch.device(device if device is not None else config.device)
        precision = precision or getattr(config.model, "b200_precision", "bf16")
        self.config = config
        self.engine = ScoreEngine(config, get_model(name), device, precision=precision)
        self.training = False
        self.module = self          # `.module` of DataParallel
        if init:
            self.init_weights(config.seed if seed is None else seed)

    # -- parameters --
    def init_weights(self, seed, zero_init_scale=None):
        """Random init with the reference's distributions.  `zero_init_scale`: replaces the 1e-10 scale of the
        `init_scale=0.` layers (models/layers.py:83-86) -- with it left at 1e-10 the score is ~0 and the Langevin
        step size (sampling.py:278) overflows, so throughput runs on random weights pass 1.0 here."""
        gen = torch.Generator().manual_seed(int(seed))

        def fix(init):
            if zero_init_scale is not None and isinstance(init, tuple) and init[0] == "default" and init[1] == 0:
                return ("default", zero_init_scale)
            return init

        sd = {key: init_tensor(shape, fix(init), gen) for key, shape, init in self.engine.spec()}
        sd["sigmas"] = torch.tensor(get_sigmas(self.config), dtype=torch.float32)
        self.engine.params["sigmas"] = sd["sigmas"].to(self.engine.device)
        self.engine.load_state_dict(sd)

    def state_dict(self):
        return {k: v.clone() for k, v in self.engine.state_dict().items()}

    def load_state_dict(self, sd, strict=True):
        """Accepts the reference's checkpoints: keys `[module.]all_modules.<i>.<name>` and `[module.]sigmas`."""
        for k in ("sigmas", "module.sigmas"):
            if k in sd:
                self.engine.params["sigmas"] = sd[k].detach().to(self.engine.device, torch.float32).contiguous()
        self.engine.load_state_dict(sd, strict=strict)

    def parameters(self):
        """The trainable parameters in the reference's `model.parameters()` order (what optimizers and
        ExponentialMovingAverage index by).  The returned list remembers its model (`.model`)."""
        tr = getattr(self, "trainer", None)
        vals = tr.parameters() if tr is not None else [self.engine.params[k] for k in trainable_parameter_order(self.config)]
        out = ParameterList(vals)
        out.model = self
        return out

    def parameters_changed(self):
        """Parameters were written in place (EMA copy_to / restore): re-derive the packed copies the plans point at."""
        tr = getattr(self, "trainer", None)
        if tr is not None:
            tr.refresh()
        else:
            self.engine.refresh_packed()
            self.engine.version += 1

    def to(self, *a, **k):
        return self

    def eval(self):
        self.training = False
        return self

    def train(self, mode=True):
        """Train mode only matters to `losses.get_step_fn` (dropout lives in the training plan, models/train.py);
        `model(x, labels)` itself is always the inference plan."""
        self.training = bool(mode)
        return self

    def __call__(self, x, labels):
        return self.engine.forward(x, labels)

    forward = __call__


def create_model(config, precision=None, seed=None):
    """models/utils.py:88-94"""
    import b_pinn_kalman_filter_b200.models  # noqa: F401  (registers the builders)
    return ScoreModel(config, precision=precision, seed=seed)


def get_model_fn(model, train=False):
    """models/utils.py:97-126"""

    def model_fn(x, labels):
        if not train:
            model.eval()
            return model(x, labels)
        model.train()
        return model(x, labels)

    return model_fn


def get_score_fn(sde, model, train=False, continuous=False):
    """models/utils.py:129-178: label construction + sign/scale of the network output."""
    model_fn = get_model_fn(model, train=train)

    if isinstance(sde, sde_lib.VPSDE) or isinstance(sde, sde_lib.subVPSDE):
        def score_fn(x, t):
            if continuous or isinstance(sde, sde_lib.subVPSDE):
                labels = t * 999
                score = model_fn(x, labels)
                std = sde.marginal_prob(torch.zeros_like(x), t)[1]
            else:
                labels = t * (sde.N - 1)
                score = model_fn(x, labels)
                std = sde.sqrt_1m_alphas_cumprod.to(labels.device)[labels.long()]
            score = -score / std[:, None, None, None]
            return score

    elif isinstance(sde, sde_lib.VESDE):
        def score_fn(x, t):
            if continuous:
                labels = sde.marginal_prob(torch.zeros_like(x), t)[1]
            else:
                labels = sde.T - t
                labels *= sde.N - 1
                labels = torch.round(labels).long()
            score = model_fn(x, labels)
            return score

    else:
        raise NotImplementedError(f"SDE class {sde.__class__.__name__} not yet supported.")

    return score_fn


def get_score_vjp_fn(sde, model, continuous=False):
    """score_and_vjp(x, t) -> (score, vjp) with vjp(v) = J_score(x)^T v (gradient w.r.t. x of <score(x, t), v>).

    The engine-side replacement of autograd through `get_score_fn`'s output (what DPS needs,
    inverse/conditional_sampling.py:116-136): same label / scale conventions as get_score_fn above, the Jacobian
    product comes from ScoreEngine.forward_for_grad + input_grad.  One vjp call per score evaluation."""
    if not isinstance(model, ScoreModel):
        raise TypeError("get_score_vjp_fn needs a ScoreModel (B200 engine)")
    eng = model.engine

    def run(x, labels, scale):
        out = eng.forward_for_grad(x, labels)
        score = out * scale[:, None, None, None]

        def vjp(v):
            return eng.input_grad(v * scale[:, None, None, None])
        return score, vjp

    if isinstance(sde, (sde_lib.VPSDE, sde_lib.subVPSDE)):
        def score_and_vjp(x, t):
            if continuous or isinstance(sde, sde_lib.subVPSDE):
                labels = t * 999
                std = sde.marginal_prob(torch.zeros_like(x), t)[1]
            else:
                labels = t * (sde.N - 1)
                std = sde.sqrt_1m_alphas_cumprod.to(labels.device)[labels.long()]
            return run(x, labels, -1.0 / std)
    elif isinstance(sde, sde_lib.VESDE):
        def score_and_vjp(x, t):
            if continuous:
                labels = sde.marginal_prob(torch.zeros_like(x), t)[1]
            else:
                labels = torch.round((sde.T - t) * (sde.N - 1)).long()
            return run(x, labels, torch.ones_like(t))
    else:
        raise NotImplementedError(f"SDE class {sde.__class__.__name__} not yet supported.")
    return score_and_vjp


def to_flattened_numpy(x):
    return x.detach().cpu().numpy().reshape((-1,))


def from_flattened_numpy(x, shape):
    return torch.from_numpy(x.reshape(shape))


# ---- on-disk format: the reference's checkpoints (utils.py:84-128 `save_checkpoint` / `load_checkpoint`) ---------------
# A checkpoint is `torch.save` of {'model': DataParallel state_dict (`module.all_modules.<i>.<name>`, `module.sigmas`),
# 'ema': {'decay', 'num_updates', 'shadow_params': [tensor, ...]}, 'optimizer'..., 'step'}.  `shadow_params` follows
# `model.parameters()` restricted to requires_grad (models/ema.py:28-29), i.e. nn.Module REGISTRATION order, which is
# not the order of the forward walk: inside a block the reference registers GroupNorm_0, Conv_0, Dense_0, GroupNorm_1,
# Dropout_0, Conv_1, then Conv_2 / NIN_0 (layerspp.py:166-192, 212-240, layers.py:611-634); attention registers
# GroupNorm_0, NIN_0..3 (layerspp.py:62-73).  The frozen Fourier projection `all_modules.0.W`
# (layerspp.py:37, requires_grad=False) has no shadow entry.
_SUBMODULE_RANK = {"GroupNorm_0": 0, "Conv_0": 1, "Conv2d_0": 1, "Dense_0": 2, "GroupNorm_1": 3, "Conv_1": 5, "Conv_2": 6,
                   "NIN_0": 6, "NIN_1": 7, "NIN_2": 8, "NIN_3": 9}
_LEAF_RANK = {"weight": 0, "W": 0, "bias": 1, "b": 1}


def trainable_parameter_order(config):
    """Keys of the network's trainable parameters in the order of the reference's `model.parameters()` (what
    `ExponentialMovingAverage.shadow_params` is indexed by).  Needs no device."""
    from .engine import parameter_spec
    spec, _ = parameter_spec(config, get_model(config.model.name))
    fourier = str(getattr(config.model, "embedding_type", "positional")).lower() == "fourier"
    keys = []
    for key, _, _ in spec:
        parts = key.split(".")          # all_modules.<i>[.<sub>].<leaf>
        if parts[0] != "all_modules":
            raise ValueError("unexpected parameter key %s" % key)
        if fourier and key == "all_modules.0.W":
            continue
        sub = parts[2] if len(parts) == 4 else None
        keys.append((int(parts[1]), _SUBMODULE_RANK[sub] if sub is not None else 0, _LEAF_RANK[parts[-1]], key))
    return [k[3] for k in sorted(keys)]


def ema_state_dict(config, ema_state, base=None):
    """{key: tensor} from an `ExponentialMovingAverage.state_dict()` of the reference (models/ema.py:91-93) -- the
    equivalent of `ema.copy_to(score_model.parameters())` (run_lib.py:172-175).  `base`: state dict providing the
    entries EMA does not track (`sigmas`, the frozen Fourier projection)."""
    order = trainable_parameter_order(config)
    shadow = ema_state["shadow_params"]
    if len(shadow) != len(order):
        raise RuntimeError("EMA holds %d tensors, the network has %d trainable parameters" % (len(shadow), len(order)))
    sd = {}
    if base is not None:
        for k, v in base.items():
            sd[k[len("module."):] if k.startswith("module.") else k] = v
    for k, v in zip(order, shadow):
        sd[k] = v
    return sd


def checkpoint_state_dict(ckpt, config, use_ema=False, map_location="cpu"):
    """State dict of a reference checkpoint (a path or the already loaded dict); use_ema=True returns the EMA weights
    the reference samples with (run_lib.py:172-175, 380-383)."""
    import os
    if isinstance(ckpt, (str, os.PathLike)):
        ckpt = torch.load(ckpt, map_location=map_location, weights_only=False)
    model_sd = ckpt["model"]
    if not use_ema:
        return {k[len("module."):] if k.startswith("module.") else k: v for k, v in model_sd.items()}
    return ema_state_dict(config, ckpt["ema"], base=model_sd)


def load_checkpoint(ckpt_dir, model, device=None, use_ema=False):
    """utils.load_checkpoint(ckpt_dir, model, device) (utils.py:96-106): warns and returns the model unchanged when the
    file does not exist, else loads checkpoint['model'] (or, with use_ema=True, the EMA weights) into it."""
    import logging
    import os
    if not os.path.exists(ckpt_dir):
        logging.warning(f"No checkpoint found at {ckpt_dir}. Returned the same state as input")
        return model
    model.load_state_dict(checkpoint_state_dict(ckpt_dir, model.config, use_ema=use_ema))
    return model
