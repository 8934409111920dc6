"""The fused Philox4x32-10 generator (csrc/common.cuh, csrc/sde_update.cu) and the stream bookkeeping around it
(sampling.reserve_streams): the parity tests inject the reference's noise tensors, so THIS file is what checks the
noise a production run actually draws -- moments, independence between streams / steps / shards, and that successive
sampler calls do not replay one another's noise (the reference draws fresh `torch.randn_like` every call,
sampling.py:197, 275)."""
import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu


def _fill(dev, n, seed, offset, elem_base=0, scale=1.0):
    from b_pinn_kalman_filter_b200 import _lib
    out = torch.empty(n, device=dev, dtype=torch.float32)
    _lib.check(_lib.lib().b200_fill_normal(_lib.ptr(out), n, float(scale), seed, offset, elem_base, _lib.stream()),
               "b200_fill_normal")
    return out


def test_fill_normal_moments_and_tails(cuda):
    n = 1 << 22
    z = _fill(cuda, n, seed=123, offset=1).double()
    se = 1.0 / np.sqrt(n)
    assert abs(z.mean().item()) < 5 * se
    assert abs(z.var().item() - 1.0) < 5 * np.sqrt(2.0) * se
    assert abs((z ** 3).mean().item()) < 5 * np.sqrt(15.0) * se            # skewness 0
    assert abs((z ** 4).mean().item() - 3.0) < 5 * np.sqrt(96.0) * se      # kurtosis 3
    # tail mass P(|z| > 3) = 2.6998e-3
    p = (z.abs() > 3).double().mean().item()
    assert abs(p - 2.6998e-3) < 5 * np.sqrt(2.6998e-3 / n)
    assert torch.isfinite(z).all() and z.abs().max().item() < 7.0


def test_streams_steps_and_shards_are_independent(cuda):
    n = 1 << 20
    a = _fill(cuda, n, 7, 1)
    se = 1.0 / np.sqrt(n)
    # another stream id, the stream 16 ids later (= the next device step of a fused loop), another seed
    for other in (_fill(cuda, n, 7, 2), _fill(cuda, n, 7, 1 + 16), _fill(cuda, n, 8, 1)):
        assert not torch.equal(a, other)
        assert abs((a.double() * other.double()).mean().item()) < 5 * se
    # lag correlations inside one stream (the 4 normals of one counter, neighbouring counters)
    for lag in (1, 2, 3, 4, 1024):
        assert abs((a[:-lag].double() * a[lag:].double()).mean().item()) < 5 * se
    # a shard starting at global element e draws exactly the slice [e, e + n) of the unsharded field
    whole = _fill(cuda, n, 7, 5)
    part = _fill(cuda, n // 4, 7, 5, elem_base=n // 2)
    assert torch.equal(part, whole[n // 2: n // 2 + n // 4])
    assert torch.equal(_fill(cuda, n, 7, 1), a)                      # and a (seed, stream, element) triple is reproducible


def test_fused_update_noise_has_unit_variance_and_matches_fill(cuda):
    """The update kernel's in-register noise is the same field b200_fill_normal writes (x = 0, a = 1, b = 0, c = 1)."""
    from b_pinn_kalman_filter_b200 import sampling
    B, shape = 8, (8, 3, 32, 32)
    x = torch.zeros(shape, device=cuda)
    s = torch.zeros(shape, device=cuda)
    one, zero = torch.ones(B, device=cuda), torch.zeros(B, device=cuda)
    xn, xm = sampling.fused_update(x, s, one, zero, one, seed=99, offset=42)
    assert torch.equal(xm, torch.zeros_like(xm))
    assert torch.equal(xn.flatten(), _fill(cuda, xn.numel(), 99, 42))
    assert abs(xn.double().var().item() - 1.0) < 0.03


def _sampler(dev, N=6, seed=None, **kw):
    from b_pinn_kalman_filter_b200 import sampling, sde_lib
    from b_pinn_kalman_filter_b200.models import utils as mutils
    c = cases.pc_case("ve_rd_langevin")
    m = mutils.ScoreModel(c["cfg"], device=dev, precision="bf16", init=False)
    m.load_state_dict(c["sd"])
    sde = sde_lib.VESDE(c["cfg"].model.sigma_min, c["cfg"].model.sigma_max, N)
    x0 = torch.full(c["shape"], 0.5)
    sde.prior_sampling = lambda shape: x0.clone()                     # identical x_T for every call: only z differs
    fn = sampling.get_pc_sampler(sde, c["shape"], sampling.ReverseDiffusionPredictor, sampling.LangevinCorrector,
                                 lambda v: v, 0.16, n_steps=1, continuous=True, denoise=False, eps=1e-5, device=dev,
                                 seed=seed, **kw)
    return fn, m


def test_successive_sampler_calls_draw_fresh_noise(cuda):
    """every pc_sampler(model) call used to replay the identical z sequence."""
    from b_pinn_kalman_filter_b200 import sampling
    sampling.set_seed(2024)
    fn, m = _sampler(cuda)
    a, _ = fn(m)
    b, _ = fn(m)
    assert torch.isfinite(a).all() and torch.isfinite(b).all()
    assert not torch.equal(a, b)
    d = (a - b).flatten().double()
    assert d.std().item() > 1e-3 * a.flatten().double().std().item()
    # rewinding the seed reproduces the first call bit for bit (what torch.manual_seed does for the reference)
    sampling.set_seed(2024)
    fn2, m2 = _sampler(cuda)
    a2, _ = fn2(m2)
    assert torch.equal(a, a2)
    # an explicit per-sampler seed is a different key
    sampling.set_seed(2024)
    fn3, m3 = _sampler(cuda, seed=5)
    a3, _ = fn3(m3)
    assert not torch.equal(a, a3)


def test_generic_and_fused_paths_never_share_a_stream(cuda):
    from b_pinn_kalman_filter_b200 import sampling
    sampling.set_seed(1)
    ids = [sampling.reserve_streams(1)]
    base = sampling.reserve_streams(sampling.STREAMS_PER_STEP * 11)
    ids.append(sampling.reserve_streams(1))
    block = range(base, base + sampling.STREAMS_PER_STEP * 11)
    assert ids[0] not in block and ids[1] not in block and ids[0] != ids[1]


def test_device_prior_is_the_global_field_slice(cuda):
    """prior='device': a shard's x_T is its slice of ONE global N(0, sigma_max^2) field, independent of the host RNG."""
    from b_pinn_kalman_filter_b200 import sampling, sde_lib
    from b_pinn_kalman_filter_b200.models import utils as mutils
    c = cases.pc_case("ve_rd_langevin")
    m = mutils.ScoreModel(c["cfg"], device=cuda, precision="bf16", init=False)
    m.load_state_dict(c["sd"])
    sde = sde_lib.VESDE(0.01, 50.0, 4)
    n_per = int(np.prod(c["shape"][1:]))
    outs = []
    for rank_offset in (0, 2):
        sampling.set_seed(77)
        loop = sampling.FusedPCLoop(m, sde, c["shape"], sampling.ReverseDiffusionPredictor, None, 0.16, 1, False, True,
                                    1e-5, rank_offset=rank_offset, total_batch=4)
        loop.reset()
        loop.prior_on_device(50.0)
        outs.append(loop.x.clone())
        off = loop.offset_base + sampling.STREAMS_PER_STEP * sde.N
    whole = _fill(cuda, 4 * n_per, sampling._SEED[0], off, scale=50.0).view(4, *c["shape"][1:])
    assert torch.equal(outs[0], whole[:2]) and torch.equal(outs[1], whole[2:])
    assert abs(whole.double().std().item() / 50.0 - 1.0) < 0.05
