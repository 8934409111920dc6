"""Training step on the score engine: what `step_fn` of the reference does per batch (losses.py:165-222) --
`loss_fn` (perturb the data, run the network in train mode, score-matching loss, losses.py:69-163), `loss.backward()`,
`optimize_fn` (warm-up, clip_grad_norm_, Adam; losses.py:48-66) and `ema.update` (models/ema.py:36-56).

Layout: every trainable parameter lives in ONE flat fp32 buffer in the reference's `model.parameters()` order (so the
EMA shadow list of a checkpoint is the same order); the engine's `params` are views into it, gradients / Adam moments /
EMA shadow are flat buffers of the same shape.  One training plan per batch size (`engine.plan(B, train=True)`): forward
with dropout, loss cotangent, the input-gradient chain of models/backward.py and, per layer, its parameter gradients
(weight gradients on the tensor-core convolution kernel over K = pixels, see csrc/train.cu).  The optimizer is one fused
kernel over the flat buffers; with a process group the flat gradient is all-reduced in buckets over NCCL first.
"""
import math

import os

import torch

from .. import _lib, sde_lib
from .._lib import c_float, c_int, c_ll, c_vp
from . import backward, utils as mutils
from .engine import Workspace

_lib.register_signature("b200_score_loss", [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_float, c_int, c_ll, c_vp])
_lib.register_signature("b200_grad_sumsq", [c_vp, c_vp, c_ll, c_vp])
_lib.register_signature("b200_adam_ema", [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_float, c_ll, c_vp])

BUCKET_BYTES = 32 << 20      # gradient all-reduce bucket (launch-latency sized: NVSwitch bandwidth is not per-link bound)


class Trainer:
    """Owns the flat parameter / gradient / optimizer buffers of one ScoreModel and its training plan."""

    def __init__(self, model, sde, batch_size, continuous=True, reduce_mean=False, likelihood_weighting=False,
                 dropout=None, seed=0, process_group=None, eps=1e-5):
        self.model, self.sde, self.B = model, sde, int(batch_size)
        self.continuous, self.reduce_mean, self.lw, self.eps_t = continuous, reduce_mean, likelihood_weighting, eps
        if not continuous and likelihood_weighting:
            raise AssertionError("Likelihood weighting is not supported for original SMLD/DDPM training.")
        if not continuous and not isinstance(sde, (sde_lib.VESDE, sde_lib.VPSDE)):
            raise ValueError(f"Discrete training for {sde.__class__.__name__} is not recommended.")
        eng = self.eng = model.engine
        cfg = self.cfg = model.config
        self.dev = eng.device
        self.lib = _lib.lib()
        self.pg = process_group
        self.world = 1
        if process_group is not None:
            import torch.distributed as dist
            self.world = dist.get_world_size(process_group)
        # ---- flat buffers, reference parameter order ----
        self.keys = mutils.trainable_parameter_order(cfg)
        shapes = {k: tuple(eng.params[k].shape) for k in self.keys}
        self.offsets, n = {}, 0
        for k in self.keys:
            self.offsets[k] = n
            n += int(math.prod(shapes[k]))
        self.n = n
        mk = lambda: torch.zeros(n, device=self.dev, dtype=torch.float32)
        self.flat_p, self.flat_g, self.flat_m, self.flat_v = mk(), mk(), mk(), mk()
        self.views, self.gviews = {}, {}
        for k in self.keys:
            o, sz = self.offsets[k], int(math.prod(shapes[k]))
            self.views[k] = self.flat_p[o:o + sz].view(shapes[k])
            self.gviews[k] = self.flat_g[o:o + sz].view(shapes[k])
            self.views[k].copy_(eng.params[k])
            eng.params[k] = self.views[k]             # the engine now reads the flat buffer (kernels see updates in place)
        self.flat_ema = self.flat_p.clone()
        self.ema_views = {k: self.flat_ema[self.offsets[k]:self.offsets[k] + self.views[k].numel()].view(shapes[k])
                          for k in self.keys}
        eng._packed.clear(); eng._recipes.clear(); eng._splits.clear(); eng._refresh_cl = None; eng.plans.clear(); eng.graphs.clear()
        eng.workspace = Workspace(self.dev)
        eng.version += 1
        # ---- training plan ----
        eng.train_dropout = float(cfg.model.dropout if dropout is None else dropout)
        eng.train_seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        eng.train_step = torch.zeros(1, device=self.dev, dtype=torch.int32)
        eng.dropout_layers = 0
        self.plan = eng.plan(self.B, train=True)
        shape = tuple(self.plan.x_in.shape)
        self.shape, self.n_per = shape, int(math.prod(shape[1:]))
        f = lambda *s: torch.zeros(*s, device=self.dev, dtype=torch.float32)
        self.z, self.losses, self.loss = f(shape), f(self.B), f(1)
        self.alpha, self.beta, self.weight, self.zeros_b = f(self.B), f(self.B), f(self.B), f(self.B)
        self.hyper, self.sumsq = f(9), f(1)
        self.opt_steps = 0            # Adam's own step count (bias correction)
        self.ema_updates = 0
        self.pending = None           # optimizer step recorded by `adam_step`, executed by `flush` (so EMA can ride along)
        covered = self._check_coverage()
        missing = [k for k in self.keys if k not in covered]
        if missing:
            raise RuntimeError("no gradient is produced for %s" % missing[:5])

    # ---- pieces of loss_fn (losses.py:69-163) ---------------------------------------------------------------------
    def _schedule(self, t, labels_int):
        """Per-sample (network labels, mean scale a, std s, alpha, beta, weight): residual = alpha * out + beta * z."""
        sde = self.sde
        ones = torch.ones_like(t)
        if self.continuous:
            a, s = sde.marginal_coef(t)
            a = a * ones
            if isinstance(sde, sde_lib.VESDE):
                labels, m = s, ones                                   # score = model(x, sigma_t)
            else:
                labels, m = t * 999, -1.0 / s                         # score = -model(x, 999 t) / std
            if not self.lw:
                return labels, a, s, m * s, ones, ones                # (score * std + z)^2
            g2 = sde.sde(torch.zeros(self.shape, device=t.device), t)[1] ** 2
            return labels, a, s, m, 1.0 / s, g2                       # (score + z / std)^2 * g^2
        if isinstance(sde, sde_lib.VESDE):                            # SMLD, losses.py:117-137
            sig = torch.flip(sde.discrete_sigmas.to(t.device), dims=(0,))[labels_int]
            return labels_int.to(torch.float32), ones, sig, ones, 1.0 / sig, sig ** 2
        sa = sde.sqrt_alphas_cumprod.to(t.device)[labels_int]         # DDPM, losses.py:140-162
        s1 = sde.sqrt_1m_alphas_cumprod.to(t.device)[labels_int]
        return labels_int.to(torch.float32), sa, s1, ones, -ones, ones

    def draw(self, generator=None):
        """(t or integer labels, None): the random draws of loss_fn that are not the noise field."""
        if self.continuous:
            return torch.rand(self.B, device=self.dev, generator=generator) * (self.sde.T - self.eps_t) + self.eps_t
        return torch.randint(0, self.sde.N, (self.B,), device=self.dev, generator=generator)

    def _run(self, ops, st):
        for fn, args, name in ops:
            rc = fn(*args, st)
            if rc != 0:
                _lib.check(rc, name)

    def loss_and_grads(self, batch, t=None, z=None):
        """optimizer.zero_grad(); loss = loss_fn(model, batch); loss.backward()  (losses.py:207-209).  `t` (or integer
        labels for the discrete losses) and `z` may be injected (parity tests); otherwise they are drawn on the device.
        Returns the loss as a device scalar; the gradients are in `self.flat_g` (reference layout)."""
        self.flush()
        _lib.require_cuda(batch)
        p, lib = self.plan, self.lib
        if tuple(batch.shape) != self.shape:
            raise RuntimeError("Trainer was planned for batches of shape %s, got %s" % (self.shape, tuple(batch.shape)))
        with torch.cuda.device(self.dev):
            st = _lib.stream()
            if t is None:
                t = self.draw()
            t = t.to(self.dev)
            labels_int = t.long() if not self.continuous else None
            tf = t.to(torch.float32)
            labels, a, s, al, be, w = self._schedule(tf, labels_int)
            if z is None:
                # a fresh Philox stream per step (b200_fill_normal keyed by (seed, stream, element))
                self._z_stream = getattr(self, "_z_stream", 0) + 1
                _lib.check(lib.b200_fill_normal(_lib.ptr(self.z), self.z.numel(), 1.0, self.eng.train_seed ^ 0x5EED,
                                                0x7A000000 + self._z_stream, 0, st), "b200_fill_normal")
            else:
                self.z.copy_(z)
            self.alpha.copy_(al); self.beta.copy_(be); self.weight.copy_(w)
            # perturbed = a * x + s * z  (mean + std z): the fused update kernel with score := x, b := 0
            p.x_in.copy_(batch)
            a32, s32 = a.to(torch.float32).contiguous(), s.to(torch.float32).contiguous()
            _lib.check(lib.b200_sde_update(_lib.ptr(p.x_in), None, _lib.ptr(p.x_in), _lib.ptr(self.z), _lib.ptr(a32),
                                           _lib.ptr(self.zeros_b), _lib.ptr(s32), None, 0, None, 0.0, 0.0, self.B,
                                           self.n_per, 0, 0, 0, st), "b200_sde_update")
            p.labels.copy_(labels.to(torch.float32).reshape(self.B))
            self._run(p.ops[:p.n_fwd], st)
            r = 1.0 / self.n_per if self.reduce_mean else 0.5
            _lib.check(lib.b200_score_loss(_lib.ptr(self.losses), _lib.ptr(self.loss), _lib.ptr(p.cot_in), _lib.ptr(p.out),
                                           _lib.ptr(self.z), _lib.ptr(self.alpha), _lib.ptr(self.beta),
                                           _lib.ptr(self.weight), _lib.ptr(p.sample_mul_t), r, self.B, self.n_per, st),
                       "b200_score_loss")
            self._run(p.ops[p.n_fwd:], st)
            self._collect()
            self.eng.train_step += 1                  # next step: new dropout masks
        return self.loss[0]

    def eval_loss(self, batch, t=None, z=None):
        """loss_fn in eval mode (losses.py:214-220): no dropout, no gradients -- the inference plan + the loss kernel."""
        self.flush()
        _lib.require_cuda(batch)
        with torch.cuda.device(self.dev), torch.no_grad():
            st = _lib.stream()
            t = self.draw() if t is None else t.to(self.dev)
            labels_int = t.long() if not self.continuous else None
            labels, a, s, al, be, w = self._schedule(t.to(torch.float32), labels_int)
            if z is None:
                self._z_stream = getattr(self, "_z_stream", 0) + 1
                _lib.check(self.lib.b200_fill_normal(_lib.ptr(self.z), self.z.numel(), 1.0, self.eng.train_seed ^ 0x5EED,
                                                     0x7A000000 + self._z_stream, 0, st), "b200_fill_normal")
            else:
                self.z.copy_(z)
            col = lambda v: v.to(torch.float32).view(-1, 1, 1, 1)
            out = self.eng.forward(col(a) * batch + col(s) * self.z, labels.to(torch.float32))
            self.alpha.copy_(al); self.beta.copy_(be); self.weight.copy_(w)
            scratch = torch.empty_like(out)
            r = 1.0 / self.n_per if self.reduce_mean else 0.5
            _lib.check(self.lib.b200_score_loss(_lib.ptr(self.losses), _lib.ptr(self.loss), _lib.ptr(scratch), _lib.ptr(out),
                                                _lib.ptr(self.z), _lib.ptr(self.alpha), _lib.ptr(self.beta),
                                                _lib.ptr(self.weight), None, r, self.B, self.n_per, st), "b200_score_loss")
        return self.loss[0].clone()

    # ---- packed-layout gradients -> the flat buffer in the reference layout ------------------------------------------
    def _targets(self, g):
        """[(key, tensor expression)] of one plan.pgrads entry."""
        out = []
        # every entry is a VIEW of a gradient buffer the backward launches fill (plus a scale): the same list serves the
        # descriptor table of _collect and, through .to() / mul, the torch form it replaced
        if g["kind"] == "conv":
            k, cs, sc = g["k"], g["cin_s"], g["scale"]
            gw = g["gw"][:, :g["kout"]].unflatten(1, (k, k, cs))
            gb = g["gb"]
            if g["nin"] == "qkv":
                C = g["cout"] // 3
                for i in range(3):
                    out.append((g["key"] + ".NIN_%d.W" % i, gw[i * C:(i + 1) * C, 0, 0, :C].t(), sc))
                    out.append((g["key"] + ".NIN_%d.b" % i, gb[i * C:(i + 1) * C], 1.0))
            elif g["nin"]:
                out.append((g["key"] + ".W", gw[:g["cout"], 0, 0, :g["cin"]].t(), sc))
                out.append((g["key"] + ".b", gb[:g["cout"]], 1.0))
            else:
                out.append((g["key"] + ".weight", gw[:g["cout"], :, :, :g["cin"]].permute(0, 3, 1, 2), sc))
                out.append((g["key"] + ".bias", gb[:g["cout"]], 1.0))
        elif g["kind"] == "gn":
            buf = g["buf"].view(g["C"], 2)
            out += [(g["key"] + ".weight", buf[:, 0], 1.0), (g["key"] + ".bias", buf[:, 1], 1.0)]
        elif g["kind"] == "dense":
            H4, row = self.plan.temb["hidden"], 0
            gw = g["gw"].view(-1, H4)
            for prefix, oc in self.eng.dense_layout:
                out += [(prefix + ".weight", gw[row:row + oc], 1.0), (prefix + ".bias", g["gb"][row:row + oc], 1.0)]
                row += oc
        elif g["kind"] == "linear":
            out += [(g["key"] + ".weight", g["gw"].view(g["shape"]), 1.0), (g["key"] + ".bias", g["gb"], 1.0)]
        return out

    def _check_coverage(self):
        seen = set()
        for g in self.plan.pgrads:
            for key, _, _ in self._targets(g):
                if key in seen:
                    raise RuntimeError("parameter %s receives two gradients" % key)
                seen.add(key)
        return seen

    def _collect(self):
        """Per-layer gradients (GEMM layouts, bf16 or fp32, padded) -> the flat fp32 gradient buffer in the reference's
        parameter order: ONE b200_copy_batch launch over a descriptor table built at the first call (as ~330 torch
        slice / permute / cast / copy calls this was 6.4 ms of host time per step; B200_FAST_REFRESH=0 keeps that form)."""
        if os.environ.get("B200_FAST_REFRESH", "1") == "0":
            for g in self.plan.pgrads:
                for key, val, sc in self._targets(g):
                    self.gviews[key].copy_(val.to(torch.float32) * sc if sc != 1.0 else val)
            return
        if getattr(self, "_collect_cl", None) is None:
            from .copylist import CopyList
            cl = CopyList(self.flat_g.device)
            for g in self.plan.pgrads:
                for key, val, sc in self._targets(g):
                    cl.add(self.gviews[key], val.view(self.gviews[key].shape), scale=sc)
            self._collect_cl = cl.finalize()
        self._collect_cl.run()

    # ---- gradient all-reduce, optimizer, EMA ----------------------------------------------------------------------
    def allreduce_grads(self):
        """Sum of the flat gradient over the group, in buckets issued back to back (asynchronous NCCL operations that
        overlap one another); the 1 / world mean is folded into the optimizer kernel's grad_scale."""
        if self.world == 1:
            return
        import torch.distributed as dist
        per = BUCKET_BYTES // 4
        works = [dist.all_reduce(self.flat_g[o:min(o + per, self.n)], op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
                 for o in range(0, self.n, per)]
        for w in works:
            w.wait()

    def adam_step(self, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, max_norm=-1.0):
        """optimizer.step() after clip_grad_norm_(max_norm): RECORDED here, executed by flush() so that a following
        `ema_update` rides in the same pass over the parameters."""
        self.flush()
        self.opt_steps += 1
        t = self.opt_steps
        self.pending = [float(lr), float(beta1), float(beta2), float(eps), float(weight_decay), 1.0 - beta1 ** t,
                        1.0 - beta2 ** t, float(max_norm), 0.0]
        self._pending_ema = False

    def ema_update(self, decay):
        """ExponentialMovingAverage.update (models/ema.py:36-56): shadow -= (1 - decay_t) (shadow - param)."""
        self.ema_updates += 1
        d = min(decay, (1 + self.ema_updates) / (10 + self.ema_updates))
        if self.pending is not None:
            self.pending[8] = 1.0 - d
            self._pending_ema = True
            self.flush()
        else:
            self.flat_ema.sub_((1.0 - d) * (self.flat_ema - self.flat_p))

    def flush(self):
        if self.pending is None:
            return
        hyper, with_ema = self.pending, self._pending_ema
        self.pending = None
        with torch.cuda.device(self.dev):
            st = _lib.stream()
            self.allreduce_grads()
            self.hyper.copy_(torch.tensor(hyper, dtype=torch.float32))
            _lib.check(self.lib.b200_grad_sumsq(_lib.ptr(self.sumsq), _lib.ptr(self.flat_g), self.n, st), "b200_grad_sumsq")
            _lib.check(self.lib.b200_adam_ema(_lib.ptr(self.flat_p), _lib.ptr(self.flat_m), _lib.ptr(self.flat_v),
                                              _lib.ptr(self.flat_ema) if with_ema else None, _lib.ptr(self.flat_g),
                                              _lib.ptr(self.hyper), _lib.ptr(self.sumsq), 1.0 / self.world, self.n, st),
                       "b200_adam_ema")
            self.refresh()

    def refresh(self):
        """The parameters changed in place: re-derive every packed / transposed copy the plans point at."""
        self.eng.refresh_packed()
        for plan in self.eng.plans.values():
            backward.refresh_dgrad_weights(plan)
        self.eng.version += 1

    # ---- parameter access in the reference's order -------------------------------------------------------------------
    def parameters(self):
        self.flush()
        return [self.views[k] for k in self.keys]

    def grads(self):
        return {k: self.gviews[k] for k in self.keys}

    def load_flat(self, src):
        """Overwrite all parameters from a flat buffer in this order (EMA copy_to / restore)."""
        self.flush()
        self.flat_p.copy_(src)
        self.refresh()
