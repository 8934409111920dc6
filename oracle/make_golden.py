"""Generates tests/golden/*.npz by running the UNMODIFIED reference (/root/reference) in the authoring container.

TEST INFRASTRUCTURE ONLY.  Run:  python oracle/make_golden.py [--only NAME]
Large tensors that are a pure function of a seed (weights, noise) are NOT stored: both sides regenerate them with
`oracle/weights.py` / `noise_stream` and the fixture keeps a checksum to detect RNG drift.
"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_import, weights  # noqa: E402
from b_pinn_kalman_filter_b200 import configs  # noqa: E402
from b_pinn_kalman_filter_b200.models import utils as mutils  # noqa: E402
import b_pinn_kalman_filter_b200.models  # noqa: E402,F401
from b_pinn_kalman_filter_b200.models.engine import parameter_spec  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")

# name -> (our config name, reference config module, overrides on both, B, labels)
SCORE_CASES = {
    "tiny_fourier_residual": ("test/tiny_ncsnpp", "configs.ve.cifar10_ncsnpp_continuous", {}, 3),
    "tiny_output_skip": ("test/tiny_ncsnpp", "configs.ve.cifar10_ncsnpp_continuous",
                         {"progressive": "output_skip", "progressive_input": "input_skip"}, 2),
    "tiny_positional_vp": ("test/tiny_ncsnpp", "configs.vp.cifar10_ncsnpp_continuous",
                           {"embedding_type": "positional", "scale_by_sigma": False, "_centered": True,
                            "_sde": "vpsde"}, 2),
    # resblock_type='ddpm' inside NCSN++ has no reference anchor: its Upsample(fir=True, with_conv=True) hits the
    # broken upsample_conv_2d and Upsample(fir=False) calls F.interpolate(x, size, 'nearest') positionally, which
    # torch >= 2 rejects (models/layerspp.py:117).
    "cifar10_ncsnpp_ve": ("ve/cifar10_ncsnpp_continuous", "configs.ve.cifar10_ncsnpp_continuous", {}, 2),
    "nc_ddpm_vp": ("vp/nc_ddpmpp", "configs.vp.nc_ddpmpp", {}, 1),
    # full-size C5 (256x256, output_skip / input_skip pyramids, ncsnpp.py:289-292, 340-346), one image
    "celebahq256_ncsnpp_ve": ("ve/celebahq_256_ncsnpp_continuous", "configs.ve.celebahq_256_ncsnpp_continuous", {}, 1),
}
TINY = dict(nf=32, ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=(8,))


def tiny_pair(our_name, ref_module, over):
    """(our config, reference config) with identical overrides."""
    cfg = configs.get_config(our_name)
    rcfg = ref_import.get_config(ref_module)
    pairs = [cfg, rcfg]
    if our_name.startswith("test/"):
        for c in pairs:
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
        for c in pairs:
            if k == "_centered":
                c.data.centered = v
            elif k == "_sde":
                c.training.sde = v
            else:
                c.model[k] = v
    return cfg, rcfg


def weight_checksum(sd):
    s = 0.0
    for k in sorted(sd):
        s += float(sd[k].double().abs().sum())
    return s


def labels_for(cfg, B):
    if cfg.model.embedding_type == "fourier":
        return torch.tensor([0.05, 1.7, 40.0][:B], dtype=torch.float32)
    return torch.tensor([3.0, 512.0, 998.0][:B], dtype=torch.float32)


def noise_stream(seed, shape):
    gen = torch.Generator().manual_seed(int(seed))
    while True:
        yield torch.randn(shape, generator=gen)


def gen_score(ns, name):
    our, refm, over, B = SCORE_CASES[name]
    cfg, rcfg = tiny_pair(our, refm, over)
    spec, _ = parameter_spec(cfg, mutils.get_model(cfg.model.name))
    sd = weights.trained_like_state_dict(spec, cfg, seed=1234)
    model = ns.mutils.create_model(rcfg)
    model.load_state_dict({"module." + k: v for k, v in sd.items()})
    model.eval()
    g = torch.Generator().manual_seed(99)
    S, C = cfg.data.image_size, cfg.data.num_channels
    x = torch.rand(B, C, S, S, generator=g) * 1.5 - 0.25
    lab = labels_for(cfg, B)
    with torch.no_grad():
        y = model(x, lab)
    np.savez_compressed(os.path.join(GOLDEN, "score_%s.npz" % name), x=x.numpy(), labels=lab.numpy(), out=y.numpy(),
                        weight_checksum=np.float64(weight_checksum(sd)), weight_seed=1234)
    print("score", name, "out rms", float(y.pow(2).mean().sqrt()))


def gen_score64(ns, name):
    """The fp64 arbiter of SURVEY 7.3-H1: the UNMODIFIED reference network in double precision on the inputs and weights
    of score_<name>.npz (fp32 values, promoted).  err(ours, fp64) is judged against err(reference fp32, fp64)."""
    our, refm, over, B = SCORE_CASES[name]
    cfg, rcfg = tiny_pair(our, refm, over)
    spec, _ = parameter_spec(cfg, mutils.get_model(cfg.model.name))
    sd = weights.trained_like_state_dict(spec, cfg, seed=1234)
    model = ns.mutils.create_model(rcfg)
    model.load_state_dict({"module." + k: v for k, v in sd.items()})
    model = model.double().eval()
    g = torch.Generator().manual_seed(99)
    S, C = cfg.data.image_size, cfg.data.num_channels
    x = torch.rand(B, C, S, S, generator=g) * 1.5 - 0.25
    lab = labels_for(cfg, B)
    # harness shim (the reference is not edited): its FIR taps are built as fp32 tensors (up_or_down_sampling.py:221,
    # 254) and F.conv2d refuses mixed dtypes; the taps (1,3,3,1)^2/64 are exact in fp32, so promoting them is exact
    # and its sinusoidal embedding is computed through `.float()` (layers.py:500-514): promoted after the fact
    orig = ns.updown.upfirdn2d
    orig_emb = ns.layers.get_timestep_embedding
    ns.updown.upfirdn2d = lambda inp, kernel, *a, **kw: orig(inp, kernel.to(inp.dtype), *a, **kw)
    ns.layers.get_timestep_embedding = lambda *a, **kw: orig_emb(*a, **kw).double()
    try:
        with torch.no_grad():
            y = model(x.double(), lab.double())
    finally:
        ns.updown.upfirdn2d = orig
        ns.layers.get_timestep_embedding = orig_emb
    np.savez_compressed(os.path.join(GOLDEN, "score64_%s.npz" % name), out64=y.numpy())
    print("score64", name, "out rms", float(y.pow(2).mean().sqrt()))


VJP_CASES = ["nc_ddpm_vp", "tiny_fourier_residual", "tiny_output_skip", "tiny_positional_vp"]     # input-gradient (DPS, inverse/conditional_sampling.py:116-136) of the C1/C4 network


def gen_vjp(ns, name):
    """J^T v of the REFERENCE network w.r.t. its input by autograd: what torch.autograd.grad(logp, x_t) propagates
    through the score model in cond_grad_fn (inverse/conditional_sampling.py:125-131)."""
    our, refm, over, B = SCORE_CASES[name]
    cfg, rcfg = tiny_pair(our, refm, over)
    spec, _ = parameter_spec(cfg, mutils.get_model(cfg.model.name))
    sd = weights.trained_like_state_dict(spec, cfg, seed=1234)
    model = ns.mutils.create_model(rcfg)
    model.load_state_dict({"module." + k: v for k, v in sd.items()})
    model.eval()
    g = torch.Generator().manual_seed(99)
    S, C = cfg.data.image_size, cfg.data.num_channels
    x = (torch.rand(B, C, S, S, generator=g) * 1.5 - 0.25).requires_grad_(True)
    lab = labels_for(cfg, B)
    v = torch.randn(B, C, S, S, generator=torch.Generator().manual_seed(7))
    y = model(x, lab)
    gx, = torch.autograd.grad(y, x, v)
    np.savez_compressed(os.path.join(GOLDEN, "vjp_%s.npz" % name), v=v.numpy(), gx=gx.numpy(), out=y.detach().numpy())
    print("vjp", name, "gx rms", float(gx.pow(2).mean().sqrt()))


_POS = {"embedding_type": "positional", "scale_by_sigma": False, "_centered": True}
PC_CASES = {
    # name: (overrides, sde, predictor, corrector, snr, n_steps, continuous, N[, extra: net / pf])
    "ve_rd_langevin": ({}, "vesde", "reverse_diffusion", "langevin", 0.16, 1, True, 12),
    "ve_ancestral_ald": ({}, "vesde", "ancestral_sampling", "ald", 0.16, 2, True, 8),
    "vp_em_none": ({"embedding_type": "positional", "scale_by_sigma": False, "_centered": True}, "vpsde",
                   "euler_maruyama", "none", 0.16, 1, True, 10),
    # beta_max = 5: with N = 10 the default beta_max = 20 gives discrete betas > 1 and the reference itself returns NaN
    "vp_ancestral_langevin_discrete": ({"embedding_type": "positional", "scale_by_sigma": False, "_centered": True,
                                        "beta_max": 5.0},
                                       "vpsde", "ancestral_sampling", "langevin", 0.2, 1, False, 10),
    # (subVP + langevin/ald raises in the reference itself: subVPSDE has no `alphas`, sampling.py:269)
    "subvp_rd_none": ({"embedding_type": "positional", "scale_by_sigma": False, "_centered": True}, "subvpsde",
                      "reverse_diffusion", "none", 0.1, 1, True, 8),
    # north_star's "fixed-seed 1000-step trajectory": the whole schedule, 2000 network evaluations
    "ve_rd_langevin_1000": ({}, "vesde", "reverse_diffusion", "langevin", 0.16, 1, True, 1000),
    # the FULL C3 network (nf = 128: every conv on the tensor-core kernels) through 50 PC steps = 100 forwards
    "c3_ve_rd_langevin": ({}, "vesde", "reverse_diffusion", "langevin", 0.16, 1, True, 50,
                          {"net": ("ve/cifar10_ncsnpp_continuous", "configs.ve.cifar10_ncsnpp_continuous")}),
    # VPSDE.discretize (sde_lib.py:191-199) under reverse_diffusion; beta_max = 5 keeps the N = 10 betas < 1
    "vp_rd_none": (dict(_POS, beta_max=5.0), "vpsde", "reverse_diffusion", "none", 0.16, 1, True, 10),
    # NonePredictor returns (x, x): the sampler's final x_mean is the post-corrector NOISY x (sampling.py:242-250, 409)
    "ve_none_langevin": ({}, "vesde", "none", "langevin", 0.16, 1, True, 8),
    "ve_em_none": ({}, "vesde", "euler_maruyama", "none", 0.16, 1, True, 10),
    "subvp_em_none": (_POS, "subvpsde", "euler_maruyama", "none", 0.16, 1, True, 10),
    "vp_rd_probflow": (dict(_POS, beta_max=5.0), "vpsde", "reverse_diffusion", "none", 0.16, 1, True, 10, {"pf": True}),
    # (euler_maruyama + probability_flow=True raises in the reference itself: RSDE.sde returns the float 0. as
    # diffusion and sampling.py:186 subscripts it -- no anchor for that combination)
}


def pc_fields(name):
    t = PC_CASES[name]
    extra = t[8] if len(t) > 8 else {}
    return t[:8] + (extra,)


def gen_pc(ns, name):
    over, sde_name, pred, corr, snr, n_steps, continuous, N, extra = pc_fields(name)
    net = extra.get("net", ("test/tiny_ncsnpp", "configs.ve.cifar10_ncsnpp_continuous"))
    cfg, rcfg = tiny_pair(net[0], net[1], over)
    for c in (cfg, rcfg):
        c.training.sde = sde_name
        c.training.continuous = continuous
    # the network keeps num_scales = 1000 (its `sigmas` buffer is indexed by labels up to 999, ncsnpp.py:245);
    # only the SDE is shortened to N steps.
    spec, _ = parameter_spec(cfg, mutils.get_model(cfg.model.name))
    sd = weights.trained_like_state_dict(spec, cfg, seed=77)
    model = ns.mutils.create_model(rcfg)
    model.load_state_dict({"module." + k: v for k, v in sd.items()})
    model.eval()
    rcfg.model.num_scales = N
    sde, eps = ref_import.get_sde(ns, rcfg)
    B = 2
    shape = (B, cfg.data.num_channels, cfg.data.image_size, cfg.data.image_size)
    stream = noise_stream(4321, shape)
    torch_randn_like = torch.randn_like
    prior_orig = type(sde).prior_sampling
    scale = cfg.model.sigma_max if sde_name == "vesde" else 1.0
    try:
        torch.randn_like = lambda t, **kw: next(stream).to(t.dtype)
        type(sde).prior_sampling = lambda self, shp: next(stream) * scale
        fn = ns.sampling.get_pc_sampler(sde, shape, ns.sampling.get_predictor(pred), ns.sampling.get_corrector(corr),
                                        lambda v: v, snr, n_steps=n_steps, probability_flow=bool(extra.get("pf")),
                                        continuous=continuous, denoise=True, eps=eps, device="cpu")
        x, nfe = fn(model)
    finally:
        torch.randn_like = torch_randn_like
        type(sde).prior_sampling = prior_orig
    np.savez_compressed(os.path.join(GOLDEN, "pc_%s.npz" % name), out=x.numpy(), nfe=nfe,
                        weight_checksum=np.float64(weight_checksum(sd)), weight_seed=77, noise_seed=4321)
    print("pc", name, "out rms", float(x.pow(2).mean().sqrt()), "nfe", nfe)


TRAIN_CASES = {
    # name: (score case it is built on, loss kind, reduce_mean, likelihood_weighting)
    "ve_sde": ("tiny_fourier_residual", "sde", False, False),
    "ve_sde_lw_mean_skip": ("tiny_output_skip", "sde", True, True),
    "vp_ddpm": ("tiny_positional_vp", "ddpm", True, False),
}
TRAIN_OPT = dict(lr=2e-3, beta1=0.9, eps=1e-8, weight_decay=0.0, warmup=2, grad_clip=1.0, ema=0.999, steps=3)


def train_probe(i, shape):
    """Seeded direction a gradient / parameter is projected on (both sides build the same one)."""
    return torch.randn(shape, generator=torch.Generator().manual_seed(7000 + i))


def gen_train(ns, name):
    """Three steps of the reference's step_fn (losses.py:191-222) on the tiny network with dropout 0 and injected
    (t, z): loss_fn is restated from losses.py:94-110 / 149-160 around the reference's OWN model, get_score_fn,
    torch.optim.Adam and models/ema.py (losses.py itself imports the absent bayesian_torch); optimize_fn = warm-up +
    clip_grad_norm_ + step as losses.py:57-63."""
    import importlib
    ref_ema = importlib.import_module("models.ema")
    base, kind, reduce_mean, lw = TRAIN_CASES[name]
    our, refm, over, _ = SCORE_CASES[base]
    cfg, rcfg = tiny_pair(our, refm, over)
    rcfg.model.dropout = 0.
    spec, _ = parameter_spec(cfg, mutils.get_model(cfg.model.name))
    sd = weights.trained_like_state_dict(spec, cfg, seed=4321)
    model = ns.mutils.create_model(rcfg)
    model.load_state_dict({"module." + k: v for k, v in sd.items()})
    sde, _ = ref_import.get_sde(ns, rcfg)
    B, o = 2, TRAIN_OPT
    S, C = cfg.data.image_size, cfg.data.num_channels
    gen = torch.Generator().manual_seed(555)
    names = [n[len("module."):] for n, p in model.named_parameters() if p.requires_grad]
    params = [p for p in model.parameters() if p.requires_grad]
    optimizer = torch.optim.Adam(model.parameters(), lr=o["lr"], betas=(o["beta1"], 0.999), eps=o["eps"],
                                 weight_decay=o["weight_decay"])
    ema = ref_ema.ExponentialMovingAverage(model.parameters(), decay=o["ema"])
    reduce_op = torch.mean if reduce_mean else (lambda *a, **k: 0.5 * torch.sum(*a, **k))
    out = {"names": np.array(names)}
    losses, ts, zs, batches = [], [], [], []
    for step in range(o["steps"]):
        batch = torch.rand(B, C, S, S, generator=gen)
        z = torch.randn(B, C, S, S, generator=gen)
        optimizer.zero_grad()
        if kind == "sde":
            t = torch.rand(B, generator=gen) * (sde.T - 1e-5) + 1e-5
            score_fn = ns.mutils.get_score_fn(sde, model, train=True, continuous=True)
            mean, std = sde.marginal_prob(batch, t)
            score = score_fn(mean + std[:, None, None, None] * z, t)
            if not lw:
                l = torch.square(score * std[:, None, None, None] + z)
                l = reduce_op(l.reshape(B, -1), dim=-1)
            else:
                g2 = sde.sde(torch.zeros_like(batch), t)[1] ** 2
                l = torch.square(score + z / std[:, None, None, None])
                l = reduce_op(l.reshape(B, -1), dim=-1) * g2
        else:
            t = torch.randint(0, sde.N, (B,), generator=gen)
            model_fn = ns.mutils.get_model_fn(model, train=True)
            sa, s1 = sde.sqrt_alphas_cumprod, sde.sqrt_1m_alphas_cumprod
            score = model_fn(sa[t, None, None, None] * batch + s1[t, None, None, None] * z, t)
            l = reduce_op(torch.square(score - z).reshape(B, -1), dim=-1)
        loss = torch.mean(l)
        loss.backward()
        if step == 0:
            out["gnorm"] = np.array([float(p.grad.norm()) for p in params])
            out["gproj"] = np.array([float((p.grad * train_probe(i, p.shape)).sum()) for i, p in enumerate(params)])
            for tag, key in (("first", names[0]), ("mid", names[len(names) // 2]), ("last", names[-2])):
                out["gfull_" + tag] = params[names.index(key)].grad.numpy().copy()
                out["gfull_" + tag + "_key"] = np.array(key)
        lr = o["lr"] * np.minimum(step / o["warmup"], 1.0)
        for g in optimizer.param_groups:
            g["lr"] = lr
        torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=o["grad_clip"])
        optimizer.step()
        ema.update(model.parameters())
        losses.append(float(loss)); ts.append(t.numpy()); zs.append(z.numpy()); batches.append(batch.numpy())
    with torch.no_grad():
        p0 = [sd[n] for n in names]
        out["dnorm"] = np.array([float((p - q).norm()) for p, q in zip(params, p0)])
        out["dproj"] = np.array([float(((p - q) * train_probe(i, p.shape)).sum()) for i, (p, q) in enumerate(zip(params, p0))])
        out["eproj"] = np.array([float(((s - q) * train_probe(i, s.shape)).sum())
                                 for i, (s, q) in enumerate(zip(ema.shadow_params, p0))])
        out["enorm"] = np.array([float((s - q).norm()) for s, q in zip(ema.shadow_params, p0)])
    np.savez_compressed(os.path.join(GOLDEN, "train_%s.npz" % name), losses=np.array(losses), t=np.stack(ts), z=np.stack(zs),
                        batch=np.stack(batches), weight_seed=4321, **out)
    print("train", name, "losses", losses, "grad norm", float(np.sqrt((out["gnorm"] ** 2).sum())))


def gen_ctrl(ns, kind):
    """Inpainting / colorisation with the REFERENCE's controllable_generation.py (:8-180) on the tiny VE network, the
    same injected noise stream as gen_pc (prior, then per step: corrector, mix, predictor, mix)."""
    import controllable_generation as cg          # reference module (sys.path set by ref_import.load)
    cfg, rcfg = tiny_pair("test/tiny_ncsnpp", "configs.ve.cifar10_ncsnpp_continuous", {})
    spec, _ = parameter_spec(cfg, mutils.get_model(cfg.model.name))
    sd = weights.trained_like_state_dict(spec, cfg, seed=77)
    model = ns.mutils.create_model(rcfg)
    model.load_state_dict({"module." + k: v for k, v in sd.items()})
    model.eval()
    N = 8
    rcfg.model.num_scales = N
    sde, eps = ref_import.get_sde(ns, rcfg)
    B = 2
    shape = (B, cfg.data.num_channels, cfg.data.image_size, cfg.data.image_size)
    g = torch.Generator().manual_seed(21)
    data = torch.rand(shape, generator=g)
    mask = torch.zeros(shape)
    mask[:, :, :, : shape[3] // 2] = 1.0
    gray = data.mean(dim=1, keepdim=True).repeat(1, 3, 1, 1)
    stream = noise_stream(999, shape)
    torch_randn_like = torch.randn_like
    prior_orig = type(sde).prior_sampling
    try:
        torch.randn_like = lambda t, **kw: next(stream).to(t.dtype)
        type(sde).prior_sampling = lambda self, shp: next(stream) * cfg.model.sigma_max
        pred, corr = ns.sampling.get_predictor("reverse_diffusion"), ns.sampling.get_corrector("langevin")
        if kind == "inpaint":
            fn = cg.get_pc_inpainter(sde, pred, corr, lambda v: v, 0.16, n_steps=1, probability_flow=False,
                                     continuous=True, denoise=True, eps=eps)
            out = fn(model, data, mask)
        else:
            fn = cg.get_pc_colorizer(sde, pred, corr, lambda v: v, 0.16, n_steps=1, probability_flow=False,
                                     continuous=True, denoise=True, eps=eps)
            out = fn(model, gray)
    finally:
        torch.randn_like = torch_randn_like
        type(sde).prior_sampling = prior_orig
    np.savez_compressed(os.path.join(GOLDEN, "ctrl_%s.npz" % kind), out=out.numpy(), data=data.numpy(), mask=mask.numpy(),
                        gray=gray.numpy(), N=N, eps=eps, weight_seed=77, noise_seed=999)
    print("ctrl", kind, "out rms", float(out.pow(2).mean().sqrt()))


def gen_ops(ns):
    g = torch.Generator().manual_seed(5)
    cases = {}
    k1 = torch.tensor([1., 3., 3., 1.])
    k = torch.outer(k1, k1)
    k = k / k.sum()
    for tag, (shape, kk, up, down, pad) in {
        "down2": ((2, 3, 16, 16), k, 1, 2, (1, 1)),
        "up2": ((2, 3, 8, 8), k * 4, 2, 1, (2, 1)),
        "pad22": ((1, 4, 9, 9), k, 1, 1, (2, 2)),
        "odd": ((1, 2, 7, 5), torch.rand(3, 5, generator=g), 3, 2, (1, 2)),
        "crop": ((1, 2, 10, 10), k, 1, 1, (-1, 0)),
    }.items():
        x = torch.randn(shape, generator=g)
        y = ns.upfirdn2d_mod.upfirdn2d_native(x, kk, up, up, down, down, pad[0], pad[1], pad[0], pad[1])
        cases["ufd_%s_x" % tag], cases["ufd_%s_k" % tag], cases["ufd_%s_y" % tag] = x.numpy(), kk.numpy(), y.numpy()
        cases["ufd_%s_cfg" % tag] = np.array([up, down, pad[0], pad[1]])
    x = torch.randn(2, 6, 5, 5, generator=g)
    b = torch.randn(6, generator=g)
    cases["lrelu_x"], cases["lrelu_b"] = x.numpy(), b.numpy()
    cases["lrelu_y"] = ns.fused_act_mod.fused_leaky_relu(x, b).numpy()   # CPU branch: slope fixed at 0.2
    np.savez_compressed(os.path.join(GOLDEN, "ops.npz"), **cases)
    print("ops", len(cases))


def gen_ode(ns):
    """Reference get_ode_sampler (sampling.py:414-485) and get_likelihood_fn (likelihood.py:40-113) on the tiny VP
    network: scipy RK45 drives the reference on the CPU; fixed latent / data / Rademacher probe."""
    import importlib
    ref_likelihood = importlib.import_module("likelihood")
    over = {"embedding_type": "positional", "scale_by_sigma": False, "_centered": True}
    cfg, rcfg = tiny_pair("test/tiny_ncsnpp", "configs.ve.cifar10_ncsnpp_continuous", over)
    for c in (cfg, rcfg):
        c.training.sde = "vpsde"
        c.training.continuous = True
    spec, _ = parameter_spec(cfg, mutils.get_model(cfg.model.name))
    sd = weights.trained_like_state_dict(spec, cfg, seed=77)
    model = ns.mutils.create_model(rcfg)
    model.load_state_dict({"module." + k: v for k, v in sd.items()})
    model.eval()
    sde, eps = ref_import.get_sde(ns, rcfg)
    shape = (2, cfg.data.num_channels, cfg.data.image_size, cfg.data.image_size)
    gen = torch.Generator().manual_seed(2468)
    z = torch.randn(shape, generator=gen)
    fn = ns.sampling.get_ode_sampler(sde, shape, lambda v: v, denoise=False, rtol=1e-4, atol=1e-4, method="RK45",
                                     eps=eps, device="cpu")
    x, nfe = fn(model, z=z.clone())
    data = torch.rand(shape, generator=gen) * 2 - 1
    probe = torch.randint(0, 2, shape, generator=gen).float() * 2 - 1
    orig = torch.randint_like
    try:
        torch.randint_like = lambda t, **kw: (probe + 1) / 2
        lfn = ref_likelihood.get_likelihood_fn(sde, lambda v: (v + 1.) / 2., hutchinson_type="Rademacher", rtol=1e-4,
                                               atol=1e-4, method="RK45", eps=1e-5)
        bpd, zl, lnfe = lfn(model, data.clone())
    finally:
        torch.randint_like = orig
    np.savez_compressed(os.path.join(GOLDEN, "ode_tiny_vp.npz"), z=z.numpy(), out=x.numpy(), nfe=nfe, data=data.numpy(),
                        probe=probe.numpy(), bpd=bpd.numpy(), latent=zl.numpy(), lnfe=lnfe, rtol=1e-4, atol=1e-4,
                        weight_checksum=np.float64(weight_checksum(sd)), weight_seed=77)
    print("ode out rms", float(x.pow(2).mean().sqrt()), "nfe", nfe, "| bpd", bpd.tolist(), "nfe", lnfe)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default=None)
    a = ap.parse_args()
    os.makedirs(GOLDEN, exist_ok=True)
    ns = ref_import.load()
    torch.manual_seed(0)
    if a.only in (None, "ops"):
        gen_ops(ns)
    for n in SCORE_CASES:
        if a.only in (None, "score", n):
            gen_score(ns, n)
    for n in SCORE_CASES:
        if a.only in (None, "score64", "score64_" + n) and n != "celebahq256_ncsnpp_ve":
            gen_score64(ns, n)
    for n in PC_CASES:
        if a.only in (None, "pc", n):
            gen_pc(ns, n)
    for n in TRAIN_CASES:
        if a.only in (None, "train", "train_" + n):
            gen_train(ns, n)
    for kind in ("inpaint", "colorize"):
        if a.only in (None, "ctrl", "ctrl_" + kind):
            gen_ctrl(ns, kind)
    for n in VJP_CASES:
        if a.only in (None, "vjp", "vjp_" + n):
            gen_vjp(ns, n)
    if a.only in (None, "ode"):
        gen_ode(ns)
    if a.only in (None, "param_order"):
        gen_param_order(ns)


CKPT_CONFIGS = {  # our config name -> reference config module
    "ve/cifar10_ncsnpp_continuous": "configs.ve.cifar10_ncsnpp_continuous",
    "vp/nc_ddpmpp": "configs.vp.nc_ddpmpp",
    "ve/celebahq_256_ncsnpp_continuous": "configs.ve.celebahq_256_ncsnpp_continuous",
    "vp/cifar10_ncsnpp_continuous": "configs.vp.cifar10_ncsnpp_continuous",
}


def gen_param_order(ns):
    """Names of the REFERENCE networks' trainable parameters in `model.parameters()` order -- what the shadow_params
    list of a checkpoint's EMA state is indexed by (models/ema.py:28-29, utils.py:109-128)."""
    import hashlib
    import json
    out = {}
    for ours, refm in CKPT_CONFIGS.items():
        rcfg = ref_import.get_config(refm)
        rcfg.device = torch.device("cpu")
        m = ns.mutils.get_model(rcfg.model.name)(rcfg)
        names = [n for n, p in m.named_parameters() if p.requires_grad]
        out[ours] = {"count": len(names), "sha1": hashlib.sha1("\n".join(names).encode()).hexdigest(),
                     "first": names[:12], "last": names[-4:]}
    with open(os.path.join(GOLDEN, "param_order.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("param_order", {k: v["count"] for k, v in out.items()})


if __name__ == "__main__":
    main()
