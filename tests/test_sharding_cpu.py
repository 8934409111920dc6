"""Host-side logic of the batch-sharded sampler (b_pinn_kalman_filter_b200/sharding.py) with world_size 2 over gloo:
the Langevin step computed from all-reduced partial sums equals the step the oracle takes on the unsharded batch
(reference sampling.py:276-278), shards gather back in rank order, and timing is max over ranks.  No GPU needed."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from b_pinn_kalman_filter_b200 import sharding
from oracle import sampling_ref


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fields(total, shape=(3, 8, 8), seed=5):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn((total,) + shape, generator=g)
    grad = torch.randn((total,) + shape, generator=g) * torch.linspace(0.2, 3.0, total).view(-1, 1, 1, 1)
    z = torch.randn((total,) + shape, generator=g)
    return x, grad, z


def _worker(rank, world, port, total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pg = dist.group.WORLD
        x, grad, z = _fields(total)
        a, b = sharding.shard_range(total, rank, world)
        sums = sharding.local_norm_sums(grad[a:b], z[a:b])
        sharding.allreduce_norm_sums(sums, pg)
        snr, alpha = 0.16, 1.0
        step = sharding.langevin_step_size(sums, total, snr, alpha)
        x_mean = x[a:b] + step * grad[a:b]
        x_new = x_mean + (2 * step) ** 0.5 * z[a:b]
        full = sharding.gather_samples(x_new, pg, total)
        full_auto = sharding.gather_samples(x_new, pg)            # total discovered by an all-reduce
        t = sharding.max_over_ranks(1.0 + rank, pg)
        q.put((rank, step, full, full_auto, t))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [8, 7])      # even and ragged shards
def test_sharded_langevin_step_equals_unsharded_oracle(total):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=120) for _ in range(world)), key=lambda r: r[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # the oracle's Langevin step on the WHOLE batch with the same gradient and noise
    x, grad, z = _fields(total)
    sde = sampling_ref.RefSDE("vesde", 10)
    ref_x, _ = sampling_ref.corrector_step("langevin", sde, lambda xx, tt: grad, x, torch.ones(total), iter([z]), 0.16, 1)
    gn = grad.reshape(total, -1).norm(dim=-1).mean()
    zn = z.reshape(total, -1).norm(dim=-1).mean()
    ref_step = float((0.16 * zn / gn) ** 2 * 2)
    for rank, step, full, full_auto, t in res:
        assert abs(step - ref_step) <= 1e-6 * ref_step
        assert full.shape == x.shape and torch.equal(full, full_auto)
        torch.testing.assert_close(full, ref_x, rtol=1e-5, atol=1e-6)
        assert t == 2.0                                  # max over ranks of (1 + rank)
    assert torch.equal(res[0][2], res[1][2])             # every rank ends with the same gathered batch


def test_shard_range_partitions_the_batch():
    for total in (0, 1, 7, 128, 1024):
        for world in (1, 2, 3, 8):
            cuts = [sharding.shard_range(total, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == total
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(8, 2, 2)


def test_philox_element_base_alignment():
    assert sharding.philox_element_base(128, 3 * 32 * 32) == 128 * 3072
    assert sharding.philox_element_base(0, 3) == 0
    with pytest.raises(ValueError):
        sharding.philox_element_base(1, 3)               # shard would start inside a Philox 4-tuple


def test_single_process_is_identity():
    x = torch.arange(6.).view(3, 2)
    assert sharding.gather_samples(x, None) is x
    s = torch.tensor([1.0, 2.0])
    assert sharding.allreduce_norm_sums(s, None) is s
    assert sharding.max_over_ranks(3.5, None) == 3.5
