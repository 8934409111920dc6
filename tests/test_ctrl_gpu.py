"""Inpainting / colorisation PC samplers on the engine (b_pinn_kalman_filter_b200/controllable_generation.py) against
golden outputs of the reference's own controllable_generation.py with the same weights and injected noise."""
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind", ["inpaint", "colorize"])
@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("bf16", 2e-2)])
def test_controllable_generation_matches_reference(cuda, kind, precision, tol):
    from b_pinn_kalman_filter_b200 import controllable_generation as cg, sampling, sde_lib
    from b_pinn_kalman_filter_b200.models import utils as mutils
    c = cases.ctrl_case(kind)
    m = mutils.ScoreModel(c["cfg"], device=cuda, precision=precision, init=False)
    m.load_state_dict(c["sd"])
    mm = c["cfg"].model
    sde = sde_lib.VESDE(mm.sigma_min, mm.sigma_max, c["N"])
    stream = cases.noise_stream(c["noise_seed"], c["shape"])
    sampling.NOISE_FN = lambda shape, device: next(stream)
    try:
        args = (sde, sampling.ReverseDiffusionPredictor, sampling.LangevinCorrector, lambda v: v, 0.16)
        kw = dict(n_steps=1, probability_flow=False, continuous=True, denoise=True, eps=c["eps"])
        if kind == "inpaint":
            out = cg.get_pc_inpainter(*args, **kw)(m, c["data"].to(cuda), c["mask"].to(cuda))
        else:
            out = cg.get_pc_colorizer(*args, **kw)(m, c["gray"].to(cuda))
    finally:
        sampling.NOISE_FN = None
    err = cases.rel_l2(out.cpu(), c["out"])
    assert err < tol, err


def test_controllable_generation_generic_loop_matches_reference(cuda):
    """A user-registered predictor class keeps the inpainter off the fused loop: the generic path (the class's own
    `update_fn` + the mix in torch) must give the same golden result."""
    from b_pinn_kalman_filter_b200 import controllable_generation as cg, sampling, sde_lib
    from b_pinn_kalman_filter_b200.models import utils as mutils
    c = cases.ctrl_case("inpaint")
    m = mutils.ScoreModel(c["cfg"], device=cuda, precision="fp32", init=False)
    m.load_state_dict(c["sd"])
    sde = sde_lib.VESDE(c["cfg"].model.sigma_min, c["cfg"].model.sigma_max, c["N"])
    stream = cases.noise_stream(c["noise_seed"], c["shape"])
    sampling.NOISE_FN = lambda shape, device: next(stream)
    try:
        user = type("UserPredictor", (sampling.ReverseDiffusionPredictor,), {})
        assert not sampling.is_fusable(m, sde, user, sampling.LangevinCorrector)
        fn = cg.get_pc_inpainter(sde, user, sampling.LangevinCorrector, lambda v: v, 0.16, n_steps=1,
                                 probability_flow=False, continuous=True, denoise=True, eps=c["eps"])
        out = fn(m, c["data"].to(cuda), c["mask"].to(cuda))
    finally:
        sampling.NOISE_FN = None
    assert cases.rel_l2(out.cpu(), c["out"]) < 1e-4


def test_fused_inpainter_replays_as_a_graph_and_keeps_known_pixels(cuda):
    """Production mode (Philox noise on device): the step with its two mask-mix launches is captured into a CUDA graph;
    known pixels of the result carry data noised at the LAST time step (x_mean: exactly the data for VE, a = 1)."""
    from b_pinn_kalman_filter_b200 import sampling, sde_lib
    from b_pinn_kalman_filter_b200.models import utils as mutils
    c = cases.ctrl_case("inpaint")
    m = mutils.ScoreModel(c["cfg"], device=cuda, precision="bf16", init=False)
    m.load_state_dict(c["sd"])
    sde = sde_lib.VESDE(c["cfg"].model.sigma_min, c["cfg"].model.sigma_max, 12)
    data, mask = c["data"].to(cuda), c["mask"].to(cuda)
    loop = sampling.FusedPCLoop(m, sde, c["shape"], sampling.ReverseDiffusionPredictor, sampling.LangevinCorrector, 0.16, 1,
                                False, True, 1e-5, mix=dict(data=data, mask=mask))
    loop.reset(torch.randn(c["shape"], device=cuda) * 50)
    loop.run(12)
    torch.cuda.synchronize()
    assert loop.graph is not None
    assert torch.isfinite(loop.x).all()
    known = mask.bool()
    assert torch.equal(loop.x_mean[known], data[known])
    assert (loop.x[known] - data[known]).abs().max() < 1.0          # data + sigma(t_last) z, sigma ~ 0.01
    assert not torch.equal(loop.x[~known], torch.zeros_like(loop.x[~known]))
