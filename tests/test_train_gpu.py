"""The training step (b_pinn_kalman_filter_b200/losses.py + models/train.py + csrc/train.cu) against three steps of the
reference's step_fn run by oracle/make_golden.py (reference model / get_score_fn / torch.optim.Adam / models/ema.py,
dropout 0, injected t and z): loss values, the gradient of EVERY parameter after the first backward (norm and a seeded
projection, three tensors in full), and the parameter / EMA-shadow movement after three optimizer steps."""
import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu


def _setup(c, dev, precision):
    from b_pinn_kalman_filter_b200 import losses, sde_lib
    from b_pinn_kalman_filter_b200.models import utils as mutils
    from b_pinn_kalman_filter_b200.models.ema import ExponentialMovingAverage
    cfg = c["cfg"]
    m = mutils.ScoreModel(cfg, device=dev, precision=precision, init=False)
    m.load_state_dict(c["sd"])
    mm = cfg.model
    if cfg.training.sde == "vesde":
        sde = sde_lib.VESDE(mm.sigma_min, mm.sigma_max, mm.num_scales)
    else:
        sde = sde_lib.VPSDE(mm.beta_min, mm.beta_max, mm.num_scales)
    state = dict(optimizer=losses.get_optimizer(cfg, m.parameters()), model=m,
                 ema=ExponentialMovingAverage(m.parameters(), decay=mm.ema_rate), step=0)
    step_fn = losses.get_step_fn(sde, train=True, optimize_fn=losses.optimization_manager(cfg), reduce_mean=c["reduce_mean"],
                                 continuous=c["kind"] == "sde", likelihood_weighting=c["lw"])
    return m, sde, state, step_fn


@pytest.mark.parametrize("name", sorted(cases.TRAIN_CASES))
@pytest.mark.parametrize("precision,gtol,ltol", [("fp32", 2e-3, 1e-4), ("bf16", 6e-2, 2e-2)])
def test_training_steps_match_reference(cuda, name, precision, gtol, ltol):
    from b_pinn_kalman_filter_b200.models import utils as mutils
    c = cases.train_case(name)
    g = c["g"]
    m, sde, state, step_fn = _setup(c, cuda, precision)
    names = [str(n) for n in g["names"]]
    assert names == mutils.trainable_parameter_order(c["cfg"])          # the reference's model.parameters() order
    p0 = [c["sd"][n].to(cuda) for n in names]
    losses_got = []
    for step in range(cases.TRAIN_OPT["steps"]):
        t = torch.from_numpy(g["t"][step]).to(cuda)
        loss = step_fn(state, torch.from_numpy(g["batch"][step]).to(cuda), t=t, z=torch.from_numpy(g["z"][step]).to(cuda))
        losses_got.append(float(loss))
        if step == 0:
            tr = m.trainer
            grads = tr.grads()
            gn = np.array([float(grads[k].norm()) for k in names])
            gp = np.array([float((grads[k].cpu() * cases.train_probe(i, grads[k].shape)).sum()) for i, k in enumerate(names)])
            total = float(np.sqrt((g["gnorm"] ** 2).sum()))
            # per parameter, relative to that parameter's own gradient norm (floor: 1e-4 of the total norm)
            scale = np.maximum(g["gnorm"], 1e-4 * total)
            bad = np.nonzero((np.abs(gn - g["gnorm"]) > gtol * scale) | (np.abs(gp - g["gproj"]) > 3 * gtol * scale))[0]
            assert len(bad) == 0, [(names[i], gn[i], g["gnorm"][i], gp[i], g["gproj"][i]) for i in bad[:6]]
            for tag in ("first", "mid", "last"):
                key = str(g["gfull_%s_key" % tag])
                ref = torch.from_numpy(g["gfull_" + tag])
                assert cases.rel_l2(grads[key].cpu(), ref) < 5 * gtol, (key, cases.rel_l2(grads[key].cpu(), ref))
    assert state["step"] == 3
    np.testing.assert_allclose(losses_got, g["losses"], rtol=ltol)
    # parameter and EMA movement after three steps (step 0 has lr = 0 by the warm-up rule: losses.py:57-59)
    params = m.parameters()
    shadow = state["ema"].shadow_params
    dtol = 3e-2 if precision == "fp32" else 0.35          # Adam's early updates are ~ lr * sign(g): sign flips of tiny g
    dn = np.array([float((p - q).norm()) for p, q in zip(params, p0)])
    en = np.array([float((s - q).norm()) for s, q in zip(shadow, p0)])
    # Parameters whose TRUE gradient is zero (the attention key bias NIN_1.b: softmax is invariant to it) receive pure
    # round-off noise, which Adam's m / sqrt(v) turns into full-size steps of arbitrary sign -- in the reference as well
    # as here.  They are recognised by a first-step gradient norm 1e-5 below the median and left out.
    live = g["gnorm"] > 1e-5 * np.median(g["gnorm"])
    assert live.sum() >= len(names) - 8, [n for n, l in zip(names, live) if not l]
    bad = np.nonzero(live & (np.abs(dn - g["dnorm"]) > dtol * np.maximum(g["dnorm"], 1e-12)))[0]
    assert len(bad) == 0, [(names[i], dn[i], g["dnorm"][i], g["gnorm"][i]) for i in bad[:6]]
    bad = np.nonzero(live & (np.abs(en - g["enorm"]) > dtol * np.maximum(g["enorm"], 1e-12)))[0]
    assert len(bad) == 0, [(names[i], en[i], g["enorm"][i]) for i in bad[:6]]
    if precision == "fp32":
        dp = np.array([float(((p - q).cpu() * cases.train_probe(i, p.shape)).sum()) for i, (p, q) in enumerate(zip(params, p0))])
        bad = np.nonzero(live & (np.abs(dp - g["dproj"]) > 0.1 * g["dnorm"] + 1e-12))[0]
        assert len(bad) == 0, [(names[i], dp[i], g["dproj"][i], g["dnorm"][i]) for i in bad[:6]]
    # the model the sampler sees has moved with the parameters (packed weights refreshed in place)
    x = torch.from_numpy(g["batch"][0]).to(cuda)
    lab = torch.full((x.shape[0],), 1.0, device=cuda)
    y1 = m(x, lab)
    from b_pinn_kalman_filter_b200.models import utils as mu
    m2 = mu.ScoreModel(c["cfg"], device=cuda, precision=precision, init=False)
    sd2 = dict(c["sd"])
    sd2.update({k: v.clone() for k, v in zip(names, params)})
    m2.load_state_dict(sd2)
    assert cases.rel_l2(y1.cpu(), m2(x, lab).cpu()) < (1e-5 if precision == "fp32" else 2e-2)


def test_eval_step_uses_ema_weights_and_restores(cuda):
    from b_pinn_kalman_filter_b200 import losses
    c = cases.train_case("ve_sde")
    g = c["g"]
    m, sde, state, step_fn = _setup(c, cuda, "fp32")
    batch, t, z = (torch.from_numpy(g[k][0]).to(cuda) for k in ("batch", "t", "z"))
    for _ in range(2):
        step_fn(state, batch, t=torch.from_numpy(g["t"][1]).to(cuda), z=z)
    eval_fn = losses.get_step_fn(sde, train=False, reduce_mean=False, continuous=True, likelihood_weighting=False)
    before = [p.clone() for p in m.parameters()]
    le = float(eval_fn(state, batch, t=t, z=z))
    after = m.parameters()
    assert all(torch.equal(a, b) for a, b in zip(before, after))          # ema.restore put the raw parameters back
    assert np.isfinite(le) and le > 0
    # the same loss from a fresh model that HOLDS the EMA shadow as its weights, and from one holding the raw weights
    from b_pinn_kalman_filter_b200.models import utils as mutils
    from b_pinn_kalman_filter_b200.models.ema import ExponentialMovingAverage
    names = mutils.trainable_parameter_order(c["cfg"])

    def loss_with(weights):
        sd = dict(c["sd"])
        sd.update({k: v.clone() for k, v in zip(names, weights)})
        m2 = mutils.ScoreModel(c["cfg"], device=cuda, precision="fp32", init=False)
        m2.load_state_dict(sd)
        st2 = dict(model=m2, ema=ExponentialMovingAverage(m2.parameters(), decay=0.999), step=0, optimizer=None)
        return float(eval_fn(st2, batch, t=t, z=z))

    l_ema, l_raw = loss_with(state["ema"].shadow_params), loss_with(after)
    assert abs(le - l_ema) <= 1e-5 * l_ema, (le, l_ema)
    assert abs(le - l_raw) > 1e-4 * l_raw, (le, l_raw)


def test_dropout_mask_statistics_and_backward(cuda):
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import backward  # noqa: F401  (registers b200_dropout)
    lib = _lib.lib()
    n, p = 1 << 20, 0.1
    x = torch.ones(n, device=cuda)
    step = torch.zeros(1, device=cuda, dtype=torch.int32)
    ys = []
    for s in (0, 0, 1):
        step.fill_(s)
        y = torch.empty_like(x)
        _lib.check(lib.b200_dropout(_lib.ptr(y), _lib.ptr(x), 0, n, p, 11, 5 << 32, _lib.ptr(step), _lib.stream()), "dropout")
        ys.append(y)
    torch.cuda.synchronize()
    assert torch.equal(ys[0], ys[1]) and not torch.equal(ys[0], ys[2])           # same step: same mask; next step: new one
    frac = float((ys[0] == 0).float().mean())
    assert abs(frac - p) < 5 * (p * (1 - p) / n) ** 0.5
    kept = ys[0][ys[0] != 0]
    assert torch.allclose(kept, torch.full_like(kept, 1 / (1 - p)))
    assert abs(float(ys[0].mean()) - 1.0) < 2e-3
    # a training step with dropout 0.1 runs, is finite, and differs from the dropout-free loss
    c = cases.train_case("ve_sde")
    c["cfg"].model.dropout = 0.1
    m, sde, state, step_fn = _setup(c, cuda, "bf16")
    g = c["g"]
    l = float(step_fn(state, torch.from_numpy(g["batch"][0]).to(cuda), t=torch.from_numpy(g["t"][0]).to(cuda),
                      z=torch.from_numpy(g["z"][0]).to(cuda)))
    assert np.isfinite(l) and abs(l - float(g["losses"][0])) > 1e-3 * float(g["losses"][0])
    assert all(torch.isfinite(v).all() for v in m.trainer.grads().values())


@pytest.mark.parametrize("shape", [(16, 32, 32, 128, 0, 128, 3, 1), (8, 16, 16, 256, 256, 256, 3, 1), (8, 16, 16, 512, 0, 256, 1, 1),
                                   (4, 17, 17, 128, 0, 256, 3, 2)])
def test_weight_gradient_on_the_tensor_core_conv(cuda, shape):
    """dW through b200_transpose2d + b200_im2col_t + b200_conv2d_tc against torch.nn.grad.conv2d_weight (fp32 on the same
    bf16-rounded operands)."""
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import backward, tc  # noqa: F401
    B, H, W, C1, C2, Cout, k, stride = shape
    lib = _lib.lib()
    gen = torch.Generator().manual_seed(sum(shape))
    pad = k // 2 if stride == 1 else 0
    Ho = (H + 2 * pad - k) // stride + 1
    x1 = torch.randn(B, H, W, C1, generator=gen).to(cuda, torch.bfloat16)
    x2 = torch.randn(B, H, W, C2, generator=gen).to(cuda, torch.bfloat16) if C2 else None
    gy = torch.randn(B, Ho, Ho, Cout, generator=gen).to(cuda, torch.bfloat16)
    P, Ct = B * Ho * Ho, C1 + C2
    gyT = torch.empty(Cout, P, device=cuda, dtype=torch.bfloat16)
    xT = torch.empty(k * k * Ct, P, device=cuda, dtype=torch.bfloat16)
    gw = torch.empty(Cout, k * k * Ct, device=cuda, dtype=torch.bfloat16)
    st = _lib.stream()
    _lib.check(lib.b200_transpose2d(_lib.ptr(gy), _lib.ptr(gyT), 1, P, Cout, st), "transpose")
    _lib.check(lib.b200_im2col_t(_lib.ptr(x1), C1, _lib.ptr(x2), C2, _lib.ptr(xT), 1, B, H, W, Ho, Ho, k, stride, pad, st), "im2col_t")
    assert lib.b200_conv2d_tc_supported(P, 0, 1, 1, Cout, k * k * Ct, 1, 1, 0) == 1
    _lib.check(lib.b200_conv2d_tc(_lib.ptr(gyT), P, None, 0, _lib.ptr(xT), None, None, 0, None, 1.0, _lib.ptr(gw), 0, None,
                                  None, 1, 1, Cout, k * k * Ct, 1, st), "conv2d_tc (wgrad)")
    torch.cuda.synchronize()
    assert torch.equal(gyT, gy.reshape(P, Cout).t())
    xx = (torch.cat([x1, x2], -1) if C2 else x1).float().permute(0, 3, 1, 2)
    ref = torch.nn.grad.conv2d_weight(xx, (Cout, Ct, k, k), gy.float().permute(0, 3, 1, 2), stride=stride, padding=pad)
    got = gw.float().view(Cout, k, k, Ct).permute(0, 3, 1, 2)
    assert cases.rel_l2(got.cpu(), ref.cpu()) < 4e-3, cases.rel_l2(got.cpu(), ref.cpu())


@pytest.mark.parametrize("name,precision", [("ve_sde", "bf16"), ("vp_ddpm", "fp32")])
def test_descriptor_table_refresh_equals_torch_refresh(cuda, name, precision):
    """Gradient collection and the re-packing of every derived weight copy after optimizer.step() run as ONE
    b200_copy_batch launch each (models/copylist.py); the torch slice / permute / cast form they replaced
    (B200_FAST_REFRESH=0) must give the same bits: parameters, EMA and the next step's loss after three steps."""
    import os
    c = cases.train_case(name)
    g = c["g"]
    results = []
    old = os.environ.get("B200_FAST_REFRESH")
    try:
        for mode in ("0", "1"):
            os.environ["B200_FAST_REFRESH"] = mode
            m, sde, state, step_fn = _setup(c, cuda, precision)
            losses_ = []
            for i in range(3):
                batch, t, z = (torch.from_numpy(g[k][i]).to(cuda) for k in ("batch", "t", "z"))
                losses_.append(float(step_fn(state, batch, t=t, z=z)))
            results.append((losses_, [p.clone() for p in m.parameters()], [p.clone() for p in state["ema"].shadow_params]))
            if mode == "1":
                assert m.trainer._collect_cl is not None and len(m.trainer._collect_cl) > 100
                assert m.engine._refresh_cl[0].nblocks > 100
    finally:
        if old is None:
            os.environ.pop("B200_FAST_REFRESH", None)
        else:
            os.environ["B200_FAST_REFRESH"] = old
    (l0, p0, e0), (l1, p1, e1) = results
    assert l0 == l1, (l0, l1)
    assert all(torch.equal(a, b) for a, b in zip(p0, p1))
    assert all(torch.equal(a, b) for a, b in zip(e0, e1))


@pytest.mark.parametrize("shape", [  # B, H, C1, C2, Cout, k, S
    (16, 32, 128, 0, 128, 3, 16),      # K block = 2 rows of 32
    (8, 16, 256, 256, 256, 3, 4),      # two sources, K block = 4 rows of 16, 256-wide N tiles
    (8, 16, 256, 128, 256, 3, 4),      # two sources of different width -> 128-wide N tiles
    (32, 8, 256, 0, 256, 3, 8),        # K block = one 8 x 8 image
    (64, 4, 256, 0, 256, 3, 2),        # K block = four 4 x 4 images
    (8, 16, 512, 0, 256, 1, 4),        # 1x1
    (2, 64, 128, 0, 128, 3, 8),        # K block = one row of 64
])
def test_direct_weight_gradient_without_im2col(cuda, shape):
    """b200_wgrad_tc (MN-major operands straight from the channels-last tensors, TMA tap shifts, S pixel chunks) + b200_sum_partials against
    torch.nn.grad.conv2d_weight on the same bf16-rounded operands, and against the split-K GEMM over the materialised
    transposed im2col (b200_im2col_t + b200_gemm_tc_splitk), which must give the same bits."""
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import backward, tc  # noqa: F401
    B, H, C1, C2, Cout, k, S = shape
    lib = _lib.lib()
    gen = torch.Generator().manual_seed(sum(shape))
    x1 = torch.randn(B, H, H, C1, generator=gen).to(cuda, torch.bfloat16)
    x2 = torch.randn(B, H, H, C2, generator=gen).to(cuda, torch.bfloat16) if C2 else None
    gy = torch.randn(B, H, H, Cout, generator=gen).to(cuda, torch.bfloat16)
    P, Ct = B * H * H, C1 + C2
    N = k * k * Ct
    bf = dict(device=cuda, dtype=torch.bfloat16)
    gyT = torch.empty(Cout, P, **bf)
    st = _lib.stream()
    P_ = _lib.ptr
    _lib.check(lib.b200_transpose2d(P_(gy), P_(gyT), 1, P, Cout, st), "transpose")
    assert lib.b200_wgrad_tc_supported(Cout, C1, C2, B, H, H, k, S) == 1
    part = torch.full((S, Cout, N), float("nan"), **bf)
    _lib.check(lib.b200_wgrad_tc(P_(gy), P_(x1), C1, P_(x2), C2, P_(part), Cout, B, H, H, k, S, st), "b200_wgrad_tc")
    gw = torch.empty(Cout, N, device=cuda)
    _lib.check(lib.b200_sum_partials(P_(part), P_(gw), 1, S, Cout * N, st), "b200_sum_partials")
    # the same split through the materialised transposed im2col
    xT = torch.empty(N, P, **bf)
    _lib.check(lib.b200_im2col_t(P_(x1), C1, P_(x2), C2, P_(xT), 1, B, H, H, H, H, k, 1, k // 2, st), "im2col_t")
    part2 = torch.full((S, Cout, N), float("nan"), **bf)
    _lib.check(lib.b200_gemm_tc_splitk(P_(gyT), P, P_(xT), P, P_(part2), Cout, N, P, S, st), "b200_gemm_tc_splitk")
    torch.cuda.synchronize()
    assert torch.isfinite(part.float()).all()
    assert torch.equal(part, part2), (part.float() - part2.float()).abs().max().item()
    xx = (torch.cat([x1, x2], -1) if C2 else x1).float().permute(0, 3, 1, 2)
    ref = torch.nn.grad.conv2d_weight(xx, (Cout, Ct, k, k), gy.float().permute(0, 3, 1, 2), stride=1, padding=k // 2)
    got = gw.view(Cout, k, k, Ct).permute(0, 3, 1, 2)
    assert cases.rel_l2(got.cpu(), ref.cpu()) < 4e-3, cases.rel_l2(got.cpu(), ref.cpu())
