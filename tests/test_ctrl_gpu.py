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
