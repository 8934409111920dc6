"""Batch sharding of the PC sampler over one process per GPU.

Samples are independent through the score network and every predictor, so the batch is cut into contiguous slices
and the weights are replicated.  Two exchange points exist on the whole path:

* `LangevinCorrector` sizes its step from BATCH-MEAN norms (reference sampling.py:276-278).  Each rank reduces its own
  samples to two partial sums on the device (`b200_langevin_norm_sums`); `allreduce_norm_sums` adds them over the
  group (8 bytes) and the update kernel divides by the GLOBAL batch -- the step is the one the unsharded batch gets.
* the final samples are concatenated in rank order (`gather_samples`).

Everything here is backend-agnostic torch.distributed (NCCL on the GPUs, gloo in tests/test_sharding_cpu.py).
"""
import torch


def shard_range(total_batch, rank, world):
    """Contiguous slice [start, stop) of the global batch owned by `rank`; sizes differ by at most one."""
    if world < 1 or not 0 <= rank < world or total_batch < 0:
        raise ValueError("shard_range: bad (total_batch, rank, world) = (%r, %r, %r)" % (total_batch, rank, world))
    base, rem = divmod(total_batch, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def local_norm_sums(grad, noise):
    """[sum_b ||grad_b||_2, sum_b ||noise_b||_2] of this rank's samples: the host restatement of
    b200_sample_sumsq + b200_langevin_norm_sums (used by the generic path and by the tests)."""
    g = grad.reshape(grad.shape[0], -1).to(torch.float32)
    z = noise.reshape(noise.shape[0], -1).to(torch.float32)
    return torch.stack([g.norm(dim=-1).sum(), z.norm(dim=-1).sum()])


def allreduce_norm_sums(sums, group):
    """In-place sum of the two partial norm sums over the group (no-op for group None / a single rank)."""
    if group is None:
        return sums
    import torch.distributed as dist
    if dist.get_world_size(group) > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    return sums


class PeerExchange:
    """Slot tables of b200_langevin_norm_sums_xchg for one process group (all ranks on ONE node, <= 8 GPUs): every
    rank allocates its table, the CUDA IPC handles travel through `all_gather_object`, every rank maps its peers'
    tables.  `ok` is a COLLECTIVE verdict (all ranks succeeded) so that the ranks never disagree on which path they
    take; on False the sampler keeps the NCCL all-reduce.  The 1 KB table is deliberately never freed: a peer's kernel
    may still be writing into it when this object dies."""

    def __init__(self, group, device):
        import ctypes
        import torch.distributed as dist
        from . import _lib
        self.lib = _lib.lib()
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.ok, self.tables, self.why = False, None, ""
        device = torch.device(device)
        rc, raw, local = -1, b"", ctypes.c_void_p()
        if device.type == "cuda" and self.world <= self.lib.b200_xchg_max_ranks():
            with torch.cuda.device(device):
                buf = ctypes.create_string_buffer(64)
                rc = self.lib.b200_xchg_alloc(ctypes.byref(local), buf)
                raw = buf.raw
        handles = [None] * self.world
        dist.all_gather_object(handles, (rc, raw), group=group)
        ok = all(h[0] == 0 for h in handles)
        ptrs = []
        if ok:
            with torch.cuda.device(device):
                for r, (_, h) in enumerate(handles):
                    if r == self.rank:
                        ptrs.append(local.value)
                        continue
                    p = ctypes.c_void_p()
                    rc = self.lib.b200_xchg_open(h, ctypes.byref(p))
                    if rc != 0:
                        ok, self.why = False, "cudaIpcOpenMemHandle failed for rank %d (code %d)" % (r, rc)
                        break
                    ptrs.append(p.value)
        else:
            self.why = "slot table allocation failed or more than 8 ranks"
        flag = torch.tensor([1 if ok else 0], device=device if device.type == "cuda" else "cpu", dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        self.ok = bool(flag.item())
        if self.ok:
            self.tables = (ctypes.c_void_p * self.world)(*ptrs)


def langevin_step_size(sums, total_batch, snr, alpha):
    """step = 2 alpha (snr * mean||z|| / mean||grad||)^2 from the GLOBAL sums (sampling.py:276-278); the same
    expression the update kernel evaluates (csrc/sde_update.cu)."""
    gn = float(sums[0]) / total_batch
    zn = float(sums[1]) / total_batch
    return 2.0 * alpha * (snr * zn / gn) ** 2


def gather_samples(x, group, total_batch=None):
    """Concatenate the shards in rank order on every rank: [B_local, ...] -> [total_batch, ...].  Shards may differ
    in size by one sample (shard_range); they are padded to the largest for the collective and cut afterwards."""
    if group is None:
        return x
    import torch.distributed as dist
    world = dist.get_world_size(group)
    if world == 1:
        return x
    if total_batch is None:
        n = torch.tensor([x.shape[0]], device=x.device, dtype=torch.int64)
        dist.all_reduce(n, group=group)
        total_batch = int(n.item())
    sizes = [shard_range(total_batch, r, world) for r in range(world)]
    sizes = [b - a for a, b in sizes]
    rank = dist.get_rank(group)
    if x.shape[0] != sizes[rank]:
        raise ValueError("gather_samples: rank %d holds %d samples, shard_range says %d" % (rank, x.shape[0], sizes[rank]))
    big = max(sizes)
    pad = x if x.shape[0] == big else torch.cat([x, x.new_zeros((big - x.shape[0],) + tuple(x.shape[1:]))])
    outs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(outs, pad.contiguous(), group=group)
    return torch.cat([o[:s] for o, s in zip(outs, sizes)])


def max_over_ranks(value, group, device="cpu"):
    """max of a python float over the group (how bench.py turns per-rank device times into the job's time)."""
    if group is None:
        return float(value)
    import torch.distributed as dist
    t = torch.tensor([float(value)], device=device, dtype=torch.float64)
    if dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def philox_element_base(rank_offset, elems_per_sample):
    """First global element index of a shard: the kernels draw noise for element e of the GLOBAL batch from Philox
    counter (e / 4) of the step's stream, so shards of any size reproduce the unsharded noise field."""
    if rank_offset < 0 or (rank_offset * elems_per_sample) % 4 != 0:
        raise ValueError("shards must start on a multiple of 4 elements (Philox4x32 draws 4 normals per counter)")
    return rank_offset * elems_per_sample


__all__ = ["shard_range", "local_norm_sums", "allreduce_norm_sums", "langevin_step_size", "gather_samples",
           "max_over_ranks", "philox_element_base", "PeerExchange"]
