"""b200_langevin_norms_static: the allocation-free form of b200_sample_sumsq + b200_langevin_norm_sums (per-sample scratch
in a file-scope static device array; csrc/sde_update.cu).  No counterpart in the reference.
Called through the C ABI and compared with the existing two-call entry points on two calls in a
row (the second smaller than the first: stale scratch contents of the previous call must not matter).

Collected last on purpose (file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

# the same kernels run in the same order on the same values: the results agree to the bit; the assertion uses the
# tolerance the sampler tests use for fp32 update arithmetic (1e-6 relative)
RTOL = 1e-6


def _existing(lib, _lib, s, z, B, n_per, seed, off, step):
    ss = torch.empty(2, B, device=s.device, dtype=torch.float32)
    sums = torch.empty(2, device=s.device, dtype=torch.float32)
    _lib.check(lib.b200_sample_sumsq(_lib.ptr(ss[0]), _lib.ptr(ss[1]), _lib.ptr(s), _lib.ptr(z), B, n_per, seed, off, 0,
                                     _lib.ptr(step), _lib.stream()), "b200_sample_sumsq")
    _lib.check(lib.b200_langevin_norm_sums(_lib.ptr(sums), _lib.ptr(ss[0]), _lib.ptr(ss[1]), B, _lib.stream()),
               "b200_langevin_norm_sums")
    return sums


def test_static_scratch_norms_match_existing_entry_points(cuda):
    from b_pinn_kalman_filter_b200 import _lib
    lib = _lib.lib()
    assert lib.b200_langevin_norms_static_max_batch() == 16384
    g = torch.Generator().manual_seed(11)
    n_per = 3 * 32 * 32
    step = torch.tensor([3], device=cuda, dtype=torch.int32)
    # two calls in a row: B = 64 with supplied noise, then B = 5 with in-kernel Philox noise and a device step counter
    for B, given_z, stp in [(64, True, None), (5, False, step)]:
        s = (torch.randn(B, n_per, generator=g) * 3).to(cuda)
        z = torch.randn(B, n_per, generator=g).to(cuda) if given_z else None
        out = torch.full((2,), float("nan"), device=cuda)          # the caller passes inputs and outputs only
        before = torch.cuda.memory_allocated(cuda)
        rc = lib.b200_langevin_norms_static(_lib.ptr(out), _lib.ptr(s), _lib.ptr(z), B, n_per, 99, 7, 0, _lib.ptr(stp),
                                            _lib.stream())
        _lib.check(rc, "b200_langevin_norms_static")
        assert torch.cuda.memory_allocated(cuda) == before          # nothing went through torch's allocator either
        ref = _existing(lib, _lib, s, z, B, n_per, 99, 7, stp)
        torch.cuda.synchronize()
        assert torch.isfinite(out).all()
        torch.testing.assert_close(out, ref, rtol=RTOL, atol=0)
        # and against plain torch where the noise is known
        want = s.double().norm(dim=1).sum().item()
        assert abs(out[0].item() - want) <= 1e-5 * want
        if given_z:
            wz = z.double().norm(dim=1).sum().item()
            assert abs(out[1].item() - wz) <= 1e-5 * wz


def test_static_scratch_norms_reject_oversize_and_bad_arguments(cuda):
    from b_pinn_kalman_filter_b200 import _lib
    lib = _lib.lib()
    out = torch.zeros(2, device=cuda)
    s = torch.zeros(4, 8, device=cuda)
    big = lib.b200_langevin_norms_static_max_batch() + 1
    rc = lib.b200_langevin_norms_static(_lib.ptr(out), _lib.ptr(s), None, big, 8, 0, 0, 0, None, _lib.stream())
    assert rc == -4
    with pytest.raises(RuntimeError, match="static scratch capacity"):
        _lib.check(rc, "b200_langevin_norms_static")
    assert lib.b200_langevin_norms_static(None, _lib.ptr(s), None, 4, 8, 0, 0, 0, None, _lib.stream()) == -1
    assert lib.b200_langevin_norms_static(_lib.ptr(out), _lib.ptr(s), None, 4, 8, 0, 0, 2, None, _lib.stream()) == -1
    torch.cuda.synchronize()
