"""Callers of ns_step / grid_sample: patch / unpatch host logic on the CPU, rollout / UKF dynamics /
project on the GPU against the oracle."""
import numpy as np
import pytest
import torch

from oracle import ns_step_ref


def _ref_patch(x, p):
    # reference ukf_utils.py:8-15 restated with explicit loops
    B, C, H, W = x.shape
    rows = []
    for c in range(C):
        for b in range(B):
            for i in range(H // p):
                for j in range(W // p):
                    rows.append(x[b, c, i * p:(i + 1) * p, j * p:(j + 1) * p].reshape(-1))
    return torch.stack(rows)


def test_patch_unpatch_roundtrip_cpu():
    from b_pinn_kalman_filter_b200.pinn_kalman import ukf_utils
    x = torch.arange(2 * 4 * 8 * 8, dtype=torch.float32).reshape(2, 4, 8, 8)
    p = ukf_utils.patch(x, 4)
    assert p.shape == (4 * 2 * 4, 16)
    assert torch.equal(p, _ref_patch(x, 4))
    assert torch.equal(ukf_utils.unpatch(p, 4, 8, 4), x)


@pytest.mark.gpu
def test_rollout_and_dynamics_vs_oracle(cuda):
    from b_pinn_kalman_filter_b200.pinn_kalman import simulator, ukf_utils
    dens, vel, pres = ns_step_ref.synthetic_fields(2, 32, seed=6)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda)
    res, vs, ps = simulator.rollout(t(dens), t(vel), t(pres), t_range=(0, 5))
    assert len(res) == len(vs) == len(ps) == 5
    d, v, p = dens, vel, pres
    for _ in range(5):
        v = ns_step_ref.update_velocity(v, p, simulator.dt, simulator.dx)
        p = ns_step_ref.update_pressure(p, v, simulator.dt, simulator.dx)
        d = ns_step_ref.update_density(d, v, simulator.dt, simulator.dx)
    for got, ref in ((res[-1], d), (vs[-1], v), (ps[-1], p)):
        assert np.abs(got.cpu().numpy() - ref).max() / np.abs(ref).max() <= 5e-5
    # UKF dynamics: patch -> (unpatch, one step, patch)
    state0 = ukf_utils.patch(torch.cat([t(dens), t(vel), t(pres)], 1), 8)
    state1, uncer = ukf_utils.ns_dynamics_forward(state0, 8, 32)
    v1 = ns_step_ref.update_velocity(vel, pres, simulator.dt, simulator.dx)
    p1 = ns_step_ref.update_pressure(pres, v1, simulator.dt, simulator.dx)
    d1 = ns_step_ref.update_density(dens, v1, simulator.dt, simulator.dx)
    ref = ukf_utils.patch(torch.from_numpy(np.concatenate([d1, v1, p1], 1)), 8)
    assert float((state1.cpu() - ref).abs().max() / ref.abs().max()) <= 1e-5
    assert uncer.shape == (state1.shape[0], 64, 64)


@pytest.mark.gpu
def test_project_matches_aten(cuda):
    from b_pinn_kalman_filter_b200.models import flownet
    g = torch.Generator().manual_seed(2)
    f = torch.randn(2, 1, 64, 64, generator=g)
    u = torch.randn(2, 2, 64, 64, generator=g) * 3
    got = flownet.project(f.to(cuda), u.to(cuda), 0.5).cpu()
    # reference formula (models/flownet.py:8-25) on the CPU with ATen's grid_sample in fp64
    gh = torch.linspace(-1.0, 1.0, 64).view(1, 1, 1, 64).expand(2, -1, 64, -1)
    gv = torch.linspace(-1.0, 1.0, 64).view(1, 1, 64, 1).expand(2, -1, -1, 64)
    grid = torch.cat([gh, gv], 1)
    un = torch.cat([u[:, 1:2] / ((64 - 1.0) / 2.0), u[:, 0:1] / ((64 - 1.0) / 2.0)], 1)
    ref = torch.nn.functional.grid_sample(f.double(), (grid - un * 0.5).permute(0, 2, 3, 1).double(), mode='bilinear',
                                          padding_mode='border', align_corners=True)
    assert float((got.double() - ref).abs().max() / ref.abs().max()) <= 5e-5
