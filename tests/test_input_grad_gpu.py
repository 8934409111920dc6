"""Input gradient of the score engine (J_model^T v w.r.t. the image: what DPS differentiates,
inverse/conditional_sampling.py:116-136) against the golden vector made by autograd through the reference model."""
import numpy as np
import pytest
import torch

import cases
from conftest import golden_path

pytestmark = pytest.mark.gpu


def _model(cfg, sd, dev, precision):
    from b_pinn_kalman_filter_b200.models import utils as mutils
    m = mutils.ScoreModel(cfg, device=dev, precision=precision, init=False)
    m.load_state_dict(sd)
    return m


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("bf16", 3e-2)])
def test_ddpm_input_grad_matches_reference(cuda, precision, tol):
    cfg, sd, x, labels, out = cases.score_case("nc_ddpm_vp")
    g = np.load(golden_path("vjp_nc_ddpm_vp.npz"))
    m = _model(cfg, sd, cuda, precision)
    y = m.engine.forward_for_grad(x.to(cuda), labels.to(cuda)).clone()
    gx = m.engine.input_grad(torch.from_numpy(g["v"]).to(cuda))
    assert cases.rel_l2(y.cpu(), torch.from_numpy(g["out"])) < (2e-5 if precision == "fp32" else 2e-2)
    err = cases.rel_l2(gx.cpu(), torch.from_numpy(g["gx"]))
    assert err < tol, err


@pytest.mark.parametrize("name", ["tiny_fourier_residual", "tiny_output_skip", "tiny_positional_vp"])
@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("bf16", 3e-2)])
def test_ncsnpp_input_grad_matches_reference(cuda, name, precision, tol):
    """NCSN++ blocks: BigGAN resblocks with FIR up/down, FIR conv-downsample of the input pyramid ('residual'),
    Combine conv1x1 ('input_skip') and the output pyramid ('output_skip')."""
    cfg, sd, x, labels, out = cases.score_case(name)
    g = np.load(golden_path("vjp_%s.npz" % name))
    m = _model(cfg, sd, cuda, precision)
    y = m.engine.forward_for_grad(x.to(cuda), labels.to(cuda)).clone()
    gx = m.engine.input_grad(torch.from_numpy(g["v"]).to(cuda))
    assert cases.rel_l2(y.cpu(), torch.from_numpy(g["out"])) < (2e-5 if precision == "fp32" else 2e-2)
    err = cases.rel_l2(gx.cpu(), torch.from_numpy(g["gx"]))
    assert err < tol, err


def test_input_grad_is_linear_in_v_and_batch_independent(cuda):
    """Size-independent properties at the C4 batch (B = 16): linear in the cotangent; samples do not mix."""
    cfg, sd, x, labels, _ = cases.score_case("nc_ddpm_vp")
    m = _model(cfg, sd, cuda, "bf16")
    B = 16
    gen = torch.Generator().manual_seed(3)
    xb = torch.rand(B, 1, 64, 64, generator=gen).to(cuda)
    lb = torch.linspace(5, 900, B).to(cuda)
    v1 = torch.randn(B, 1, 64, 64, generator=gen).to(cuda)
    v2 = torch.randn(B, 1, 64, 64, generator=gen).to(cuda)
    def vjp(v):
        m.engine.forward_for_grad(xb, lb)            # one backward per forward: the pass consumes the activations
        return m.engine.input_grad(v).clone()

    g1, g2, g12 = vjp(v1), vjp(v2), vjp(v1 + 2 * v2)
    assert torch.isfinite(g12).all()
    assert cases.rel_l2(g12.cpu(), (g1 + 2 * g2).cpu()) < 2e-2
    with pytest.raises(RuntimeError):
        m.engine.input_grad(v1)
    v3 = v1.clone()
    v3[8:] = 0
    g3 = vjp(v3)
    assert cases.rel_l2(g3[:8].cpu(), g1[:8].cpu()) < 1e-6
    assert float(g3[8:].abs().max()) == 0.0


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-4), ("bf16", 3e-2)])
def test_dps_rhs_matches_oracle(cuda, precision, tol):
    """One right-hand-side evaluation of the DPS sampler (inverse/conditional_sampling.py:100-165), inpainting with a
    half mask, C4's network: the engine path (closed-form v + input_grad) against autograd through the oracle."""
    from oracle import dps_ref, sampling_ref, score_net_ref
    from b_pinn_kalman_filter_b200 import sde_lib
    from b_pinn_kalman_filter_b200.inverse import conditional_sampling as cs
    from b_pinn_kalman_filter_b200.inverse.operators import InpaintOperator
    cfg, sd, x1, _, _ = cases.score_case("nc_ddpm_vp")
    B, shape = 2, (2, 1, 64, 64)
    gen = torch.Generator().manual_seed(11)
    x = torch.randn(shape, generator=gen) * 0.7
    t = torch.tensor([0.63, 0.21])
    mask = torch.zeros(shape)
    mask[:, :, :, :32] = 1.0                                   # left half observed (ratio 0.5)
    truth = torch.rand(shape, generator=gen) * 2 - 1
    var = 0.1
    # oracle
    net = score_net_ref.RefScoreNet(cfg, sd)
    rsde = sampling_ref.RefSDE("vpsde", cfg.model.num_scales, cfg.model.beta_min, cfg.model.beta_max)
    ref_drift, ref_score, ref_cond = dps_ref.dps_rhs(net, rsde, x, t, mask, mask * truth, var)
    # engine
    m = _model(cfg, sd, cuda, precision)
    op = InpaintOperator(mask.to(cuda))
    state = sde_lib.VPSDE(cfg.model.beta_min, cfg.model.beta_max, cfg.model.num_scales)
    y0 = op(truth.to(cuda), keep_shape=False)
    obsv = sde_lib.LOBSVSDE(state, y0, op)
    _, _, _, rhs = cs.get_dps_functions(obsv, y0, var, shape)
    drift = rhs(m, x.to(cuda), t.to(cuda)).cpu()
    assert drift.shape == ref_drift.shape
    err = cases.rel_l2(drift, ref_drift)
    assert err < tol, err
    # the conditional term is not negligible in this comparison
    assert float(ref_cond.abs().max()) > 1e-3 * float(ref_score.abs().max())


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-4), ("bf16", 3e-2)])
def test_likelihood_rhs_matches_oracle(cuda, precision, tol):
    """Drift + Hutchinson divergence of the likelihood ODE (likelihood.py:27-37, 61-69): one engine forward + one
    input-gradient pass against autograd through the oracle."""
    from oracle import likelihood_ref, sampling_ref, score_net_ref
    from b_pinn_kalman_filter_b200 import likelihood, sde_lib
    cfg, sd, _, _, _ = cases.score_case("nc_ddpm_vp")
    shape = (2, 1, 64, 64)
    gen = torch.Generator().manual_seed(5)
    x = torch.randn(shape, generator=gen) * 0.8
    t = torch.tensor([0.71, 0.12])
    eps = torch.randint(0, 2, shape, generator=gen).float() * 2 - 1
    net = score_net_ref.RefScoreNet(cfg, sd)
    rsde = sampling_ref.RefSDE("vpsde", cfg.model.num_scales, cfg.model.beta_min, cfg.model.beta_max)
    ref_drift, ref_div = likelihood_ref.drift_div(net, rsde, x, t, eps)
    m = _model(cfg, sd, cuda, precision)
    sde = sde_lib.VPSDE(cfg.model.beta_min, cfg.model.beta_max, cfg.model.num_scales)
    drift, div = likelihood.get_drift_div_fn(sde)(m, x.to(cuda), t.to(cuda), eps.to(cuda))
    assert cases.rel_l2(drift.cpu(), ref_drift) < tol
    assert cases.rel_l2(div.cpu(), ref_div) < 5 * tol, (div.cpu(), ref_div)


@pytest.mark.parametrize("B", [1, 5])
def test_attention_backward_tensor_core_path(cuda, B):
    """b200_attention_bwd in bf16 at L = C = 256 (warp-MMA kernels, csrc/attn_bwd_mma.cu) against torch autograd in fp32
    on the same bf16-rounded operands."""
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import backward  # noqa: F401  (registers the signature)
    L = C = 256
    g = torch.Generator().manual_seed(B)
    qkv = torch.randn(B, L, 3 * C, generator=g)
    qkv[:, :, :C] *= 1.5
    qkv = qkv.to(cuda, torch.bfloat16)
    dout = torch.randn(B, L, C, generator=g).to(cuda, torch.bfloat16)
    lib = _lib.lib()
    assert lib.b200_attention_bwd_mma_supported(L, C) == 1
    dqkv = torch.full((B, L, 3 * C), float("nan"), device=cuda, dtype=torch.bfloat16)
    scratch = torch.empty(2 * B * L * L, device=cuda)
    _lib.check(lib.b200_attention_bwd(_lib.ptr(qkv), _lib.ptr(dout), _lib.ptr(dqkv), _lib.ptr(scratch), 1, B, L, C,
                                      C ** -0.5, _lib.stream()), "b200_attention_bwd")
    torch.cuda.synchronize()
    x = qkv.float().requires_grad_(True)
    q, k, v = x.split(C, dim=-1)
    o = torch.softmax(q @ k.transpose(1, 2) * C ** -0.5, dim=-1) @ v
    ref, = torch.autograd.grad(o, x, dout.float())
    assert torch.isfinite(dqkv.float()).all()
    for name, sl in (("dq", slice(0, C)), ("dk", slice(C, 2 * C)), ("dv", slice(2 * C, 3 * C))):
        err = cases.rel_l2(dqkv[..., sl].float().cpu(), ref[..., sl].cpu())
        assert err < 1.5e-2, (name, err)
