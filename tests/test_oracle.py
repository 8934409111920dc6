"""Pins the CPU oracle (oracle/) against the golden vectors that oracle/make_golden.py produced by running the
UNMODIFIED reference (tests/golden/*.npz).  CPU only."""
import numpy as np
import pytest
import torch

import cases
from conftest import golden_path
from oracle import fused_act_ref, ns_step_ref, sampling_ref, score_net_ref, upfirdn2d_ref


def test_upfirdn2d_oracle_matches_reference_native():
    g = np.load(golden_path("ops.npz"))
    tags = sorted({k.split("_")[1] for k in g.files if k.startswith("ufd_")})
    assert len(tags) == 5
    for tag in tags:
        up, down, p0, p1 = (int(v) for v in g["ufd_%s_cfg" % tag])
        y = upfirdn2d_ref.upfirdn2d(g["ufd_%s_x" % tag], g["ufd_%s_k" % tag], up, down, (p0, p1))
        ref = g["ufd_%s_y" % tag]
        assert y.shape == ref.shape, tag
        np.testing.assert_allclose(y, ref, rtol=2e-6, atol=2e-6, err_msg=tag)


def test_fused_act_oracle_matches_reference_cpu_branch():
    g = np.load(golden_path("ops.npz"))
    y = fused_act_ref.fused_leaky_relu(g["lrelu_x"], g["lrelu_b"])
    np.testing.assert_allclose(y, g["lrelu_y"], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("name", ["tiny_fourier_residual", "tiny_output_skip", "tiny_positional_vp"])
def test_score_oracle_matches_reference(name):
    cfg, sd, x, labels, out = cases.score_case(name)
    net = score_net_ref.RefScoreNet(cfg, sd)
    with torch.no_grad():
        y = net(x, labels)
    assert cases.max_rel(y, out) < 2e-5, cases.max_rel(y, out)


def test_score_oracle_matches_reference_cifar10():
    cfg, sd, x, labels, out = cases.score_case("cifar10_ncsnpp_ve")
    net = score_net_ref.RefScoreNet(cfg, sd)
    with torch.no_grad():
        y = net(x[:1], labels[:1])
    assert cases.max_rel(y, out[:1]) < 5e-5


def test_score_oracle_matches_reference_celebahq256():
    """Full-size C5 (256 x 256, input_skip / output_skip pyramids): one image."""
    cfg, sd, x, labels, out = cases.score_case("celebahq256_ncsnpp_ve")
    net = score_net_ref.RefScoreNet(cfg, sd)
    with torch.no_grad():
        y = net(x, labels)
    assert cases.max_rel(y, out) < 5e-5, cases.max_rel(y, out)


@pytest.mark.parametrize("name", ["tiny_fourier_residual", "tiny_positional_vp"])
def test_fp64_arbiter_fixture_is_consistent(name):
    """score64_*.npz = the reference network in fp64 on the inputs of score_*.npz: the fp32 golden sits within fp32
    round-off of it, and the oracle restatement run in fp64 reproduces it to ~1e-12."""
    cfg, sd, x, labels, out = cases.score_case(name)
    out64 = cases.score_fp64(name)
    assert out64.dtype == torch.float64 and out64.shape == out.shape
    assert cases.rel_l2(out, out64) < 1e-5
    net = score_net_ref.RefScoreNet(cfg, sd, dtype=torch.float64)
    with torch.no_grad():
        y = net(x, labels)
    assert cases.rel_l2(y, out64) < 1e-6, cases.rel_l2(y, out64)


@pytest.mark.parametrize("name", sorted(cases.PC_CASES))
def test_pc_oracle_matches_reference(name):
    c = cases.pc_case(name)
    cfg = c["cfg"]
    net = score_net_ref.RefScoreNet(cfg, c["sd"])
    m = cfg.model
    sde = sampling_ref.RefSDE(c["sde_name"], c["N"], m.beta_min, m.beta_max, m.sigma_min, m.sigma_max)
    with torch.no_grad():
        x = sampling_ref.pc_sample(sde, net, c["shape"], c["pred"], c["corr"], c["snr"], c["n_steps"], c["continuous"],
                                   cases.noise_stream(c["noise_seed"], c["shape"]), denoise=True, eps=c["eps"],
                                   probability_flow=c["pf"])
    assert c["nfe"] == c["N"] * (c["n_steps"] + 1)
    assert cases.rel_l2(x, c["out"]) < 1e-4, cases.rel_l2(x, c["out"])


def test_ns_step_oracle_properties():
    """No reference CPU code exists for ns_step (parity unpinned, see oracle/ns_step_ref.c): check invariants of
    the restated scheme instead -- a uniform field is a fixed point and B>1 equals per-sample runs."""
    S = 24
    dens, vel, pres = ns_step_ref.synthetic_fields(3, S, seed=1)
    dt, dx = 0.0025, 1 / 200
    d1 = ns_step_ref.update_density(np.full_like(dens, 0.7), vel, dt, dx)
    np.testing.assert_allclose(d1, 0.7, rtol=1e-6)
    v = ns_step_ref.update_velocity(vel, pres, dt, dx)
    p = ns_step_ref.update_pressure(pres, vel, dt, dx)
    d = ns_step_ref.update_density(dens, vel, dt, dx)
    for b in range(3):
        np.testing.assert_array_equal(v[b:b + 1], ns_step_ref.update_velocity(vel[b:b + 1], pres[b:b + 1], dt, dx))
        np.testing.assert_array_equal(p[b:b + 1], ns_step_ref.update_pressure(pres[b:b + 1], vel[b:b + 1], dt, dx))
        np.testing.assert_array_equal(d[b:b + 1], ns_step_ref.update_density(dens[b:b + 1], vel[b:b + 1], dt, dx))
    assert np.isfinite(v).all() and np.isfinite(p).all() and np.isfinite(d).all()
    # bug-compat mode equals the per-sample semantics at B = 1 (op/ns_step.cpp:70-72 is only wrong for B > 1)
    np.testing.assert_array_equal(ns_step_ref.update_velocity(vel[:1], pres[:1], dt, dx, bug_compat=True), v[:1])


# ---------------------------------------------------------------- grid_sample
@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("padding_mode", ["zeros", "border"])
@pytest.mark.parametrize("align_corners", [True, False])
def test_grid_sample_restatement_matches_aten(dim, padding_mode, align_corners):
    """The differentiable restatement the grad2 oracle differentiates twice equals ATen's grid_sample (the
    reference's own forward / backward, op/grid_sample.py:38-39, 54-60) in value and first derivative."""
    import torch
    from oracle import grid_sample_ref as gs
    inp, grid = gs.synthetic(dim, seed=dim * 10 + align_corners)
    inp, grid = inp.double(), grid.double()
    y_ref = gs.forward(inp, grid, padding_mode, align_corners)
    y = gs.sample(inp, grid, padding_mode, align_corners)
    torch.testing.assert_close(y, y_ref, rtol=1e-12, atol=1e-12)
    go = torch.randn(y.shape, generator=torch.Generator().manual_seed(3)).double()
    gi_ref, gg_ref = gs.backward(go, inp, grid, padding_mode, align_corners)
    x = inp.clone().requires_grad_(True)
    g = grid.clone().requires_grad_(True)
    gi, gg = torch.autograd.grad(gs.sample(x, g, padding_mode, align_corners), (x, g), go)
    torch.testing.assert_close(gi, gi_ref, rtol=1e-12, atol=1e-12)
    torch.testing.assert_close(gg, gg_ref, rtol=1e-10, atol=1e-10)


# ---------------------------------------------------------------- input gradient (DPS)
@pytest.mark.parametrize("name", ["nc_ddpm_vp", "tiny_fourier_residual", "tiny_output_skip", "tiny_positional_vp"])
def test_oracle_input_grad_matches_reference_autograd(name):
    """J^T v of the oracle networks equals the golden vectors made by autograd through the REFERENCE models
    (oracle/make_golden.py gen_vjp; inverse/conditional_sampling.py:125-131 is the consumer)."""
    cfg, sd, x, labels, out = cases.score_case(name)
    g = np.load(golden_path("vjp_%s.npz" % name))
    net = score_net_ref.RefScoreNet(cfg, sd)
    y, gx = net.input_grad(x, labels, torch.from_numpy(g["v"]))
    assert cases.max_rel(y, torch.from_numpy(g["out"])) < 2e-5
    assert cases.max_rel(gx, torch.from_numpy(g["gx"])) < 5e-5, cases.max_rel(gx, torch.from_numpy(g["gx"]))


# ---------------------------------------------------------------- controllable generation
@pytest.mark.parametrize("kind", ["inpaint", "colorize"])
def test_oracle_controllable_generation_matches_reference(kind):
    from oracle import controllable_ref
    c = cases.ctrl_case(kind)
    net = score_net_ref.RefScoreNet(c["cfg"], c["sd"])
    m = c["cfg"].model
    sde = sampling_ref.RefSDE("vesde", c["N"], sigma_min=m.sigma_min, sigma_max=m.sigma_max)
    noises = cases.noise_stream(c["noise_seed"], c["shape"])
    if kind == "inpaint":
        out = controllable_ref.pc_inpaint(sde, net, c["data"], c["mask"], "reverse_diffusion", "langevin", 0.16, 1, True,
                                          noises, c["eps"])
    else:
        out = controllable_ref.pc_colorize(sde, net, c["gray"], "reverse_diffusion", "langevin", 0.16, 1, True, noises,
                                           c["eps"])
    assert cases.rel_l2(out, c["out"]) < 1e-4, cases.rel_l2(out, c["out"])
