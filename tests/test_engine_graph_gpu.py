"""ScoreEngine replays a plan's launch list as a CUDA graph from its third run on (the ODE / likelihood / DPS solvers call
the network hundreds of times at small batches, where the host's launch calls set the pace).  Replay must give the bits
of the eager launches, for the forward plan and for both halves of the input-gradient plan, with fresh inputs each time."""
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,precision", [("nc_ddpm_vp", "bf16"), ("tiny_fourier_residual", "fp32")])
def test_graph_replay_equals_eager(cuda, name, precision, monkeypatch):
    from b_pinn_kalman_filter_b200.models import utils as mutils
    cfg, sd, x, labels, out = cases.score_case(name)
    g = torch.Generator().manual_seed(5)
    xs = [(x + 0.1 * i * torch.randn(x.shape, generator=g)).to(cuda) for i in range(5)]
    vs = [torch.randn(x.shape, generator=g).to(cuda) for _ in range(5)]
    lab = labels.to(cuda)
    res = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("B200_ENGINE_GRAPH", mode)
        m = mutils.ScoreModel(cfg, device=cuda, precision=precision, init=False)
        m.load_state_dict(sd)
        ys, gs = [], []
        for xi, vi in zip(xs, vs):
            ys.append(m(xi, lab).clone())
            m.engine.forward_for_grad(xi, lab)
            gs.append(m.engine.input_grad(vi).clone())
        res[mode] = (ys, gs)
        if mode == "1":
            graphs = m.engine.plan(x.shape[0]).__dict__.get("_graphs", {})
            assert graphs.get("all", [0, None])[1] is not None          # the forward plan was captured and replayed
            gp = m.engine.plan(x.shape[0], grad=True).__dict__.get("_graphs", {})
            assert gp["fwd"][1] is not None and gp["bwd"][1] is not None
    for a, b in zip(res["0"][0] + res["0"][1], res["1"][0] + res["1"][1]):
        assert torch.isfinite(a).all() and torch.equal(a, b)
