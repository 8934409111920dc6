"""Our-side construction of the golden cases of oracle/make_golden.py (no reference needed at test time).

make_golden.py builds (our config, reference config) pairs with identical overrides and runs the reference; the
tests rebuild only our half, regenerate the seeded weights / inputs / noise and compare with the stored outputs.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import weights  # noqa: E402
from b_pinn_kalman_filter_b200 import configs  # noqa: E402
from b_pinn_kalman_filter_b200.models import utils as mutils  # noqa: E402
import b_pinn_kalman_filter_b200.models  # noqa: E402,F401
from b_pinn_kalman_filter_b200.models.engine import parameter_spec  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")

SCORE_CASES = {
    "tiny_fourier_residual": ("test/tiny_ncsnpp", {}, 3),
    "tiny_output_skip": ("test/tiny_ncsnpp", {"progressive": "output_skip", "progressive_input": "input_skip"}, 2),
    "tiny_positional_vp": ("test/tiny_ncsnpp", {"embedding_type": "positional", "scale_by_sigma": False,
                                                "_centered": True, "_sde": "vpsde"}, 2),
    "cifar10_ncsnpp_ve": ("ve/cifar10_ncsnpp_continuous", {}, 2),
    "nc_ddpm_vp": ("vp/nc_ddpmpp", {}, 1),
    "celebahq256_ncsnpp_ve": ("ve/celebahq_256_ncsnpp_continuous", {}, 1),
}
TINY = dict(nf=32, ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=(8,))

_POS = {"embedding_type": "positional", "scale_by_sigma": False, "_centered": True}
PC_CASES = {
    # name: (overrides, sde, predictor, corrector, snr, n_steps, continuous, N)
    "ve_rd_langevin": ({}, "vesde", "reverse_diffusion", "langevin", 0.16, 1, True, 12),
    "ve_ancestral_ald": ({}, "vesde", "ancestral_sampling", "ald", 0.16, 2, True, 8),
    "vp_em_none": (_POS, "vpsde", "euler_maruyama", "none", 0.16, 1, True, 10),
    "vp_ancestral_langevin_discrete": (dict(_POS, beta_max=5.0), "vpsde", "ancestral_sampling", "langevin", 0.2, 1, False, 10),
    "subvp_rd_none": (_POS, "subvpsde", "reverse_diffusion", "none", 0.1, 1, True, 8),
    # the full-length trajectory north_star asks for: N = 1000 PC steps (2000 network evaluations), fixed noise
    "ve_rd_langevin_1000": ({}, "vesde", "reverse_diffusion", "langevin", 0.16, 1, True, 1000),
    # the full C3 network (every conv on the tensor-core kernels), 50 PC steps = 100 forwards
    "c3_ve_rd_langevin": ({}, "vesde", "reverse_diffusion", "langevin", 0.16, 1, True, 50,
                          {"net": "ve/cifar10_ncsnpp_continuous"}),
    "vp_rd_none": (dict(_POS, beta_max=5.0), "vpsde", "reverse_diffusion", "none", 0.16, 1, True, 10),
    "ve_none_langevin": ({}, "vesde", "none", "langevin", 0.16, 1, True, 8),
    "ve_em_none": ({}, "vesde", "euler_maruyama", "none", 0.16, 1, True, 10),
    "subvp_em_none": (_POS, "subvpsde", "euler_maruyama", "none", 0.16, 1, True, 10),
    "vp_rd_probflow": (dict(_POS, beta_max=5.0), "vpsde", "reverse_diffusion", "none", 0.16, 1, True, 10, {"pf": True}),
    # (euler_maruyama + probability_flow=True raises in the reference itself: RSDE.sde returns the float 0. as
    # diffusion and sampling.py:186 subscripts it -- no anchor for that combination)
}


def our_config(our_name, over):
    c = configs.get_config(our_name)
    if our_name.startswith("test/"):
        c.model.update(TINY)
        c.data.image_size = 16
        c.training.sde = "vesde"
        c.training.continuous = True
        c.data.centered = False
        c.model.scale_by_sigma = True
        c.model.embedding_type = "fourier"
        c.model.progressive, c.model.progressive_input = "none", "residual"
        c.model.resblock_type = "biggan"
        c.model.num_scales = 1000
        c.model.sigma_max, c.model.sigma_min = 50, 0.01
    for k, v in over.items():
        if k == "_centered":
            c.data.centered = v
        elif k == "_sde":
            c.training.sde = v
        else:
            c.model[k] = v
    return c


def weight_checksum(sd):
    return sum(float(sd[k].double().abs().sum()) for k in sorted(sd))


def state_dict_for(cfg, seed):
    spec, _ = parameter_spec(cfg, mutils.get_model(cfg.model.name))
    return weights.trained_like_state_dict(spec, cfg, seed=seed)


def score_case(name):
    """-> (cfg, state_dict, x, labels, golden_out)"""
    our, over, B = SCORE_CASES[name]
    cfg = our_config(our, over)
    g = np.load(os.path.join(GOLDEN, "score_%s.npz" % name))
    sd = state_dict_for(cfg, int(g["weight_seed"]))
    assert abs(weight_checksum(sd) - float(g["weight_checksum"])) <= 1e-9 * float(g["weight_checksum"]), \
        "seeded weights drifted from the ones the golden vectors were made with"
    return cfg, sd, torch.from_numpy(g["x"]), torch.from_numpy(g["labels"]), torch.from_numpy(g["out"])


def score_fp64(name):
    """The reference network evaluated in fp64 on the same inputs and weights (oracle/make_golden.py gen_score64)."""
    return torch.from_numpy(np.load(os.path.join(GOLDEN, "score64_%s.npz" % name))["out64"])


def noise_stream(seed, shape):
    gen = torch.Generator().manual_seed(int(seed))
    while True:
        yield torch.randn(shape, generator=gen)


def pc_case(name):
    """-> dict(cfg, sd, shape, sde_name, pred, corr, snr, n_steps, continuous, N, eps, out, nfe, noise_seed)"""
    tup = PC_CASES[name]
    over, sde_name, pred, corr, snr, n_steps, continuous, N = tup[:8]
    extra = tup[8] if len(tup) > 8 else {}
    cfg = our_config(extra.get("net", "test/tiny_ncsnpp"), over)
    cfg.training.sde = sde_name
    cfg.training.continuous = continuous
    g = np.load(os.path.join(GOLDEN, "pc_%s.npz" % name))
    sd = state_dict_for(cfg, int(g["weight_seed"]))
    assert abs(weight_checksum(sd) - float(g["weight_checksum"])) <= 1e-9 * float(g["weight_checksum"])
    shape = (2, cfg.data.num_channels, cfg.data.image_size, cfg.data.image_size)
    return dict(cfg=cfg, sd=sd, shape=shape, sde_name=sde_name, pred=pred, corr=corr, snr=snr, n_steps=n_steps,
                continuous=continuous, N=N, eps=1e-5 if sde_name == "vesde" else 1e-3, pf=bool(extra.get("pf")),
                out=torch.from_numpy(g["out"]), nfe=int(g["nfe"]), noise_seed=int(g["noise_seed"]))


def rel_l2(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def max_rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def ctrl_case(kind):
    """Inpainting / colorisation golden (reference controllable_generation.py run by oracle/make_golden.py gen_ctrl)."""
    cfg = our_config("test/tiny_ncsnpp", {})
    g = np.load(os.path.join(GOLDEN, "ctrl_%s.npz" % kind))
    sd = state_dict_for(cfg, int(g["weight_seed"]))
    t = lambda k: torch.from_numpy(g[k])
    return dict(cfg=cfg, sd=sd, out=t("out"), data=t("data"), mask=t("mask"), gray=t("gray"), N=int(g["N"]),
                eps=float(g["eps"]), noise_seed=int(g["noise_seed"]), shape=tuple(g["data"].shape))


TRAIN_CASES = {
    # name: (score case it is built on, loss kind, reduce_mean, likelihood_weighting)   (oracle/make_golden.py gen_train)
    "ve_sde": ("tiny_fourier_residual", "sde", False, False),
    "ve_sde_lw_mean_skip": ("tiny_output_skip", "sde", True, True),
    "vp_ddpm": ("tiny_positional_vp", "ddpm", True, False),
}
TRAIN_OPT = dict(lr=2e-3, beta1=0.9, eps=1e-8, weight_decay=0.0, warmup=2, grad_clip=1.0, ema=0.999, steps=3)


def train_probe(i, shape):
    return torch.randn(shape, generator=torch.Generator().manual_seed(7000 + i))


def train_case(name):
    """-> dict(cfg, sd, kind, reduce_mean, lw, golden npz) for three reference training steps with injected (t, z)."""
    base, kind, reduce_mean, lw = TRAIN_CASES[name]
    our, over, _ = SCORE_CASES[base]
    cfg = our_config(our, over)
    cfg.model.dropout = 0.
    o = TRAIN_OPT
    cfg.optim.update(lr=o["lr"], beta1=o["beta1"], eps=o["eps"], weight_decay=o["weight_decay"], warmup=o["warmup"],
                     grad_clip=o["grad_clip"])
    cfg.model.ema_rate = o["ema"]
    g = np.load(os.path.join(GOLDEN, "train_%s.npz" % name))
    sd = state_dict_for(cfg, int(g["weight_seed"]))
    return dict(cfg=cfg, sd=sd, kind=kind, reduce_mean=reduce_mean, lw=lw, g=g)
