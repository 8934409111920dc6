"""Two GPUs, one process each (NCCL): the batch-sharded fused PC loop against the unsharded one.

* b200_langevin_norm_sums_xchg (peer-memory exchange fused into the reducing kernel) returns on every rank, bit for
  bit, what b200_langevin_norm_sums returns for the concatenated batch on one GPU;
* after 5 PC steps (reverse diffusion + Langevin, fused Philox noise, device-side prior) each rank's x equals the
  unsharded run's slice bit for bit, in fp32 and in bf16, with the sharded step captured into a CUDA graph.
Needs >= 2 GPUs (`gpurun --gpus 2`); skipped on a single-GPU box."""
import os
import socket

import pytest
import torch

import cases

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def sde_lib_mod():
    from b_pinn_kalman_filter_b200 import sde_lib
    return sde_lib


def _loop(m, c, N, B, **kw):
    from b_pinn_kalman_filter_b200 import sampling, sde_lib
    sampling.set_seed(31)          # same Philox key AND the same block of stream ids for every loop built here
    sde = sde_lib.VESDE(c["cfg"].model.sigma_min, c["cfg"].model.sigma_max, N)
    shape = (B,) + tuple(c["shape"][1:])
    return sampling.FusedPCLoop(m, sde, shape, sampling.ReverseDiffusionPredictor, sampling.LangevinCorrector, 0.16, 1,
                                False, True, 1e-5, **kw)


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import utils as mutils
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    out = {"rank": rank}
    try:
        pg = dist.group.WORLD
        lib = _lib.lib()
        # ---- the exchange kernel alone: ragged shards (5 + 3 samples), three epochs in a row
        from b_pinn_kalman_filter_b200 import sharding
        xc = sharding.PeerExchange(pg, dev)
        out["xchg_ok"] = xc.ok
        g = torch.Generator().manual_seed(3)
        ok = True
        for it in range(3):
            ss = (torch.rand(2, 8, generator=g) * 10 ** (it + 1)).to(dev)
            lo, hi = (0, 5) if rank == 0 else (5, 8)
            mine = ss[:, lo:hi].contiguous()
            got = torch.zeros(2, device=dev)
            _lib.check(lib.b200_langevin_norm_sums_xchg(_lib.ptr(got), _lib.ptr(mine[0]), _lib.ptr(mine[1]), hi - lo,
                                                        xc.tables, world, rank, _lib.stream()), "xchg")
            ref = torch.zeros(2, device=dev)
            _lib.check(lib.b200_langevin_norm_sums(_lib.ptr(ref), _lib.ptr(ss[0]), _lib.ptr(ss[1]), 8, _lib.stream()), "ns")
            torch.cuda.synchronize()
            ok = ok and torch.equal(got, ref) and bool(torch.isfinite(got).all())
        out["xchg_equal"] = ok
        # ---- the sampler: 4 samples unsharded on every rank vs 2 + 2 sharded
        c = cases.pc_case("ve_rd_langevin")
        for precision in ("fp32", "bf16"):
            m = mutils.ScoreModel(c["cfg"], device=dev, precision=precision, init=False)
            m.load_state_dict(c["sd"])
            N, B = 7, 4
            whole = _loop(m, c, N, B)
            whole.reset(); whole.prior_on_device(50.0); whole.run(5)
            torch.cuda.synchronize()
            xw = whole.x.clone()
            part = _loop(m, c, N, B // world, process_group=pg, rank_offset=rank * (B // world), total_batch=B)
            part.reset(); part.prior_on_device(50.0); part.run(5)
            torch.cuda.synchronize()
            sl = xw[rank * (B // world):(rank + 1) * (B // world)]
            out[precision] = (bool(torch.equal(part.x, sl)), float((part.x - sl).abs().max()), part.graph is not None,
                              part.xchg is not None, bool(torch.isfinite(part.x).all()))
        # ---- data-parallel training: 2 ranks x 2 samples (bucketed NCCL all-reduce of the flat gradient, mean folded
        # into the optimizer kernel) against one rank x 4 samples -- the same parameters after one optimizer step
        import numpy as np
        from b_pinn_kalman_filter_b200 import losses
        from b_pinn_kalman_filter_b200.models import train as train_mod
        from b_pinn_kalman_filter_b200.models.ema import ExponentialMovingAverage
        tc_ = cases.train_case("ve_sde")
        tc_["cfg"].optim.warmup = 0
        g = tc_["g"]
        xb = torch.from_numpy(np.concatenate([g["batch"][0], g["batch"][1]])).to(dev)
        zb = torch.from_numpy(np.concatenate([g["z"][0], g["z"][1]])).to(dev)
        tb = torch.from_numpy(np.concatenate([g["t"][0], g["t"][1]])).to(dev)
        train_mod.BUCKET_BYTES = 256 << 10          # several buckets even for the tiny network
        res = []
        for group, sl in ((None, slice(0, 4)), (pg, slice(2 * rank, 2 * rank + 2))):
            mt = mutils.ScoreModel(tc_["cfg"], device=dev, precision="fp32", init=False)
            mt.load_state_dict(tc_["sd"])
            sde_t = sde_lib_mod().VESDE(tc_["cfg"].model.sigma_min, tc_["cfg"].model.sigma_max, tc_["cfg"].model.num_scales)
            state = dict(optimizer=losses.get_optimizer(tc_["cfg"], mt.parameters()), model=mt,
                         ema=ExponentialMovingAverage(mt.parameters(), decay=0.999), step=0)
            fn = losses.get_step_fn(sde_t, train=True, optimize_fn=losses.optimization_manager(tc_["cfg"]),
                                    reduce_mean=False, continuous=True, likelihood_weighting=False, process_group=group)
            fn(state, xb[sl], t=tb[sl], z=zb[sl])
            res.append(torch.cat([p.reshape(-1) for p in mt.parameters()]).clone())
        moved = float((res[0] - torch.cat([tc_["sd"][k].reshape(-1) for k in mt.trainer.keys]).to(dev)).norm())
        out["ddp"] = (float((res[0] - res[1]).norm()) / moved, moved, mt.trainer.world)
        dist.barrier()
    finally:
        q.put(out)
        dist.destroy_process_group()


def test_sharded_loop_equals_unsharded_slice_on_two_gpus(cuda):
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = [q.get(timeout=600) for _ in ps]
    for p in ps:
        p.join(timeout=120)
        assert p.exitcode == 0
    for r in res:
        assert r["xchg_ok"], r
        assert r["xchg_equal"], r
        for precision in ("fp32", "bf16"):
            equal, diff, graphed, xchg, finite = r[precision]
            assert finite and xchg and graphed, (precision, r)
            assert equal, (precision, diff, r)
        rel, moved, world = r["ddp"]
        assert world == 2 and moved > 0
        assert rel < 2e-3, r["ddp"]      # Adam turns 1e-6 gradient differences (summation order) into 1e-3 of one step
