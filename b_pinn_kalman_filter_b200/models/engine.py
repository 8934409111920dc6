"""ScoreEngine: executes a score network (NCSN++ / DDPM) as a static plan of C-ABI kernel launches.

Replaces `model(x, labels)` = NCSNpp.forward / DDPM.forward (models/ncsnpp.py:232-381, models/ddpm.py:110-181).

Design:
  * activations are channels-last [B, H, W, C] in `bf16` (tensor-core mode) or `fp32` (strict mode); NCHW fp32 only
    at the API boundary;
  * the network is "compiled" once per batch size into a flat list of (entry point, argument tuple) pairs over a
    reusable workspace -- no allocation, no host sync and no Python tensor math while it runs, so the whole forward
    (and the whole PC step around it) can be captured into one CUDA graph;
  * elementwise work is folded into neighbours: bias + time-embedding + residual + 1/sqrt(2) into conv epilogues,
    skip-connection concat into two-source GroupNorm / conv loaders, `2x-1` into the input layout change and
    `h / sigma` into the output one.
"""
import ctypes
import math
import os

import torch

from .. import _lib
from .._lib import c_float, c_int, c_ll, c_vp

_lib.register_signature("b200_gn_nsplit", [c_int, c_int])
_lib.register_signature("b200_gn_partial", [c_vp, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp])
_lib.register_signature("b200_gn_apply", [c_vp, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp, c_vp,
                                          c_float, c_int, c_vp, c_vp])
_lib.register_signature("b200_nchw_to_nhwc", [c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_float, c_float, c_vp])
_lib.register_signature("b200_nhwc_to_nchw", [c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp])
_lib.register_signature("b200_temb_embed", [c_vp, c_vp, c_int, c_vp, c_int, c_vp, c_int, c_int, c_vp, c_vp, c_int, c_vp])
_lib.register_signature("b200_linear_f32", [c_vp, c_vp, c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp])
_lib.register_signature("b200_add_scale", [c_vp, c_vp, c_vp, c_int, c_ll, c_float, c_vp])
_lib.register_signature("b200_nearest_up2_nhwc", [c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp])
_lib.register_signature("b200_conv2d_simt", [c_vp, c_int, c_vp, c_int, c_vp, c_vp, c_vp, c_int, c_vp, c_float, c_vp, c_int,
                                             c_vp, c_int] + [c_int] * 11 + [c_vp])
_lib.register_signature("b200_attention_simt", [c_vp, c_vp, c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                                c_int, c_float, c_vp])

DT_F32, DT_BF16 = 0, 1
_TORCH_DT = {DT_F32: torch.float32, DT_BF16: torch.bfloat16}


class Act:
    """A channels-last activation living in the workspace."""
    __slots__ = ("t", "B", "H", "W", "C", "Ct", "producer", "stats")

    def __init__(self, t, B, H, W, C, Ct=None):
        self.t, self.B, self.H, self.W, self.C = t, B, H, W, C
        self.Ct = C if Ct is None else Ct   # channels that carry data (3-channel images live in 8-wide tensors)
        self.producer = None                # (launch index, arg position) of the tensor-core conv that wrote it
        self.stats = None                   # (pointer, slots) of its GroupNorm statistics by-product

    @property
    def ptr(self):
        return self.t.data_ptr()

    def at(self, byte_offset):
        """Pointer argument `byte_offset` bytes into this activation (resolved when the plan is finalised)."""
        return Ptr(self, byte_offset)


class LazyGN(Act):
    """GroupNorm + SiLU whose launch is deferred to its first consumer: a tensor-core 3x3 convolution that supports the
    GroupNorm-fused operand pipeline takes the RAW sources plus a coefficient table (no normalised tensor is ever
    written); any other consumer makes `Builder.emit` materialise it with b200_gn_apply_stats first."""
    __slots__ = ("src1", "src2", "key", "groups", "real")

    def __init__(self, src1, src2, key, groups):
        C = src1.C + (src2.C if src2 is not None else 0)
        super().__init__(None, src1.B, src1.H, src1.W, C)
        self.src1, self.src2, self.key, self.groups, self.real = src1, src2, key, groups, None


class Ptr:
    __slots__ = ("act", "off")

    def __init__(self, act, off):
        self.act, self.off = act, off


class Workspace:
    """Size-bucketed free list over torch allocations; buffers are reused in program order (single stream)."""

    def __init__(self, device):
        self.device = device
        self.free = {}
        self.all = []
        self.bytes = 0

    def take(self, numel, dtype):
        key = (numel, dtype)
        lst = self.free.get(key)
        if lst:
            return lst.pop()
        t = torch.empty(numel, device=self.device, dtype=dtype)
        self.all.append(t)
        self.bytes += t.numel() * t.element_size()
        return t

    def give(self, t):
        self.free.setdefault((t.numel(), t.dtype), []).append(t)


class Plan:
    """Flat launch list for one batch size."""

    def __init__(self, B):
        self.B = B
        self.ops = []          # (cfunc, args_without_stream, name)
        self.keep = []         # tensors that must outlive the plan
        self.x_in = None       # fp32 NCHW input staging   [B, C, H, W]
        self.labels = None     # fp32 [B]
        self.out = None        # fp32 NCHW output
        self.launches = 0
        self.tags = []         # per launch: shape tag (profiling aid)
        self.meta = []         # per launch: (entry point, algorithmic FLOPs) -- what bench.py's roofline is computed from

    def flops(self, prefix=None):
        return sum(f for n, f in self.meta if prefix is None or n.startswith(prefix))

    def run_timed(self, stream_ptr, by_tag=False):
        """Runs the plan with a CUDA-event pair around every launch (on the launching stream); returns
        {entry point: (launches, total ms, total FLOPs)}.  Measurement aid for bench.py, not used while sampling."""
        evs = []
        stream = torch.cuda.current_stream()
        for fn, args, name in self.ops:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            rc = fn(*args, stream_ptr)
            e1.record(stream)
            if rc != 0:
                _lib.check(rc, name)
            evs.append((e0, e1))
        stream.synchronize()
        out = {}
        for (e0, e1), (name, fl), tag in zip(evs, self.meta, self.tags):
            if by_tag:
                name = name + " " + tag
            n, ms, f = out.get(name, (0, 0.0, 0))
            out[name] = (n + 1, ms + e0.elapsed_time(e1), f + fl)
        return out

    def run(self, stream_ptr):
        for fn, args, name in self.ops:
            rc = fn(*args, stream_ptr)
            if rc != 0:
                _lib.check(rc, name)


class Builder:
    """Emits the plan (or, with dry=True, only walks the network to collect the parameter specification)."""

    def __init__(self, engine, B, dry):
        self.e = engine
        self.B = B
        self.dry = dry
        self.plan = Plan(B)
        self.ws = engine.workspace if not dry else None
        self.lib = _lib.lib() if not dry else None
        self.dt = engine.dt
        self.tdt = _TORCH_DT[engine.dt]
        self.specs = []        # (key, shape, init) in reference order
        self.dense = []        # (key_prefix, out_ch) of every Dense_0, in order
        self.temb_all = None   # [B, sum(out_ch)] fp32
        self.temb_stride = 0
        self.sample_mul = None
        self.tape = [] if (not dry and getattr(engine, "grad_mode", False)) else None   # primitives, for backward.py

    def rec(self, kind, **kw):
        """Records a primitive for the input-gradient pass (models/backward.py); no-op outside grad plans."""
        if self.tape is not None:
            kw["kind"] = kind
            self.tape.append(kw)

    # ---- parameters -------------------------------------------------------------------------------------------
    def param(self, key, shape, init):
        self.specs.append((key, tuple(shape), init))
        return key

    # ---- workspace --------------------------------------------------------------------------------------------
    def new(self, H, W, C, Ct=None):
        """A virtual activation; physical memory is assigned by `finalize` from first/last use (liveness)."""
        return Act(None, self.B, H, W, C, Ct)

    def finalize(self):
        """Linear-scan buffer assignment over the recorded launches, then bake the argument tuples."""
        first, last = {}, {}
        acts = {}
        for i, (_, args, _) in enumerate(self.plan.ops):
            for a in args:
                act = a if isinstance(a, Act) else (a.act if isinstance(a, Ptr) else None)
                if act is not None:
                    acts[id(act)] = act
                    first.setdefault(id(act), i)
                    last[id(act)] = i
        born, dies = {}, {}
        for k, i in first.items():
            born.setdefault(i, []).append(k)
        for k, i in last.items():
            dies.setdefault(i, []).append(k)
        baked = []
        for i, (fn, args, name) in enumerate(self.plan.ops):
            for k in born.get(i, ()):
                act = acts[k]
                act.t = self.ws.take(act.B * act.H * act.W * act.C, self.tdt)
            out = []
            for a in args:
                if isinstance(a, Act):
                    out.append(a.t.data_ptr())
                elif isinstance(a, Ptr):
                    out.append(a.act.t.data_ptr() + a.off)
                else:
                    out.append(a)
            baked.append((fn, tuple(out), name))
            for k in dies.get(i, ()):
                self.ws.give(acts[k].t)
        self.plan.ops = baked
        self.plan.launches = len(baked)

    def f32(self, numel):
        t = torch.empty(numel, device=self.e.device, dtype=torch.float32)
        self.plan.keep.append(t)
        return t

    def materialize(self, gn):
        """The normalised tensor of a deferred GroupNorm (emits its b200_gn_apply_stats launch on first use)."""
        if gn.real is None:
            from . import tc
            x, x2 = gn.src1, gn.src2
            out = self.new(x.H, x.W, gn.C)
            s1, t1 = tc.request_stats(self, x)
            s2, t2 = tc.request_stats(self, x2) if x2 is not None else (None, 0)
            p2, c2 = (x2, x2.C) if x2 is not None else (None, 0)
            HW = x.H * x.W
            gn.real = out
            self.emit("b200_gn_apply_stats", x, x.C, s1, t1, p2, c2, s2, t2, self.dt, self.B, HW, gn.groups,
                      self.lib.b200_gn_nsplit(self.B, HW), self.e.dev(gn.key + ".weight"), self.e.dev(gn.key + ".bias"),
                      1e-6, 1, out, tag="%dx%d %d" % (x.H, x.W, gn.C))
        return gn.real

    def emit(self, name, *args, flops=0, tag=""):
        if any(isinstance(a, LazyGN) or (isinstance(a, Ptr) and isinstance(a.act, LazyGN)) for a in args):
            args = tuple(self.materialize(a) if isinstance(a, LazyGN) else
                         (Ptr(self.materialize(a.act), a.off) if isinstance(a, Ptr) and isinstance(a.act, LazyGN) else a)
                         for a in args)
        fn = getattr(self.lib, name)
        self.plan.ops.append((fn, args, name))
        self.plan.meta.append((name, int(flops)))
        self.plan.tags.append(tag)

    # ---- primitives -------------------------------------------------------------------------------------------
    def gn(self, x, key, groups, act, x2=None):
        """GroupNorm(+SiLU) over one or two (concatenated) sources -> new activation."""
        C = x.C + (x2.C if x2 is not None else 0)
        self.param(key + ".weight", (C,), "ones")
        self.param(key + ".bias", (C,), "zeros")
        out = self.new(x.H, x.W, C)
        if self.dry:
            return out
        HW = x.H * x.W
        ns = self.lib.b200_gn_nsplit(self.B, HW)
        p2, c2 = (x2, x2.C) if x2 is not None else (None, 0)
        if self.e.use_tc and self.e.fuse_gn_stats and x.producer is not None and (x2 is None or x2.producer is not None):
            from . import tc
            if act and self.tape is None and self.e.fuse_gn_conv and tc.FUSE_GN:
                return LazyGN(x, x2, key, groups)          # resolved by its first consumer (see LazyGN)
            s1, t1 = tc.request_stats(self, x)
            s2, t2 = tc.request_stats(self, x2) if x2 is not None else (None, 0)
            self.emit("b200_gn_apply_stats", x, x.C, s1, t1, p2, c2, s2, t2, self.dt, self.B, HW, groups, ns,
                      self.e.dev(key + ".weight"), self.e.dev(key + ".bias"), 1e-6, 1 if act else 0, out,
                      tag="%dx%d %d" % (x.H, x.W, C))
            self.rec("gn", x=x, x2=x2, out=out, key=key, groups=groups, act=act)
            return out
        part = self.f32(self.B * ns * groups * 2)
        self.emit("b200_gn_partial", x, x.C, p2, c2, self.dt, self.B, HW, groups, ns, part.data_ptr())
        self.emit("b200_gn_apply", x, x.C, p2, c2, self.dt, self.B, HW, groups, ns, part.data_ptr(),
                  self.e.dev(key + ".weight"), self.e.dev(key + ".bias"), 1e-6, 1 if act else 0, out)
        self.rec("gn", x=x, x2=x2, out=out, key=key, groups=groups, act=act)
        return out

    def conv(self, x, key, cout, ksize, x2=None, stride=1, pad=None, out_hw=None, temb_off=None, res=None,
             out_scale=1.0, init_scale=1.0, nin=False, cout_store=None, nchw_out=None):
        """Convolution with fused epilogue.  `nin=True`: weight stored as NIN.W [Cin, Cout] (models/layers.py:540).
        `cout_store` > cout pads the output channels with zeros (3-channel image pyramids live in 8-wide tensors)."""
        cin = x.C + (x2.C if x2 is not None else 0)
        cin_true = x.Ct + (x2.Ct if x2 is not None else 0)
        if nin:
            self.param(key + ".W", (cin_true, cout), ("default", init_scale))
            self.param(key + ".b", (cout,), "zeros")
        else:
            self.param(key + ".weight", (cout, cin_true, ksize, ksize), ("default", init_scale))
            self.param(key + ".bias", (cout,), "zeros")
        if pad is None:
            pad = (ksize // 2, ksize // 2)
        Ho, Wo = out_hw if out_hw is not None else (x.H, x.W)
        cs = cout_store or cout
        if nchw_out is not None:
            out = None
        else:
            out = self.new(Ho, Wo, cs, cout)
        if self.dry:
            return out
        w = self.e.packed_conv(key, nin, cin, cs)
        bias = self.e.packed_bias(key, nin, cs)
        temb_ptr = None
        if temb_off is not None:
            temb_ptr = self.temb_all.data_ptr() + 4 * temb_off
        p2, c2 = (x2, x2.C) if x2 is not None else (None, 0)
        self.rec("conv", x=x, x2=x2, out=out, w=w, ksize=ksize, stride=stride, pad=tuple(pad), res=res,
                 out_scale=out_scale, head=nchw_out is not None, cout=cout, key=key, nin=nin, temb_off=temb_off,
                 out_hw=(Ho, Wo))
        if nchw_out is not None:
            from . import tc
            fl = 2 * self.B * Ho * Wo * cout * cin_true * ksize * ksize
            mul_ptr = self.sample_mul.data_ptr() if self.sample_mul is not None else None
            if (self.e.use_tc and x2 is None and temb_off is None and res is None and stride == 1
                    and tuple(pad) == (ksize // 2, ksize // 2) and tc.head_supported(self, x, cout, ksize)):
                tc.emit_head(self, x, w, bias, out_scale, nchw_out.data_ptr(), mul_ptr, cout, ksize, fl)
                self.e.tc_convs += 1
                return None
            self.e.simt_convs += 1
            self.emit("b200_conv2d_simt", x, x.C, p2, c2, w.data_ptr(), bias.data_ptr(), temb_ptr, self.temb_stride,
                      res, out_scale, nchw_out.data_ptr(), 1,
                      self.sample_mul.data_ptr() if self.sample_mul is not None else None, self.dt, self.B, x.H, x.W,
                      Ho, Wo, cout, ksize, ksize, stride, pad[0], pad[1],
                      flops=2 * self.B * Ho * Wo * cout * cin_true * ksize * ksize)
            return None
        self.e.emit_conv(self, x, x2, w, bias, temb_ptr, res, out_scale, out, ksize, stride, pad, key)
        return out

    def conv_sc(self, h, key1, key2, xs, xs2, cout, out_scale, init_scale, nin2=False, nin2_init=1.0):
        """Second convolution of a residual block: (conv3x3(h; key1) + shortcut1x1(cat(xs, xs2); key2)) * out_scale
        (models/layerspp.py:266-274, models/layers.py:648-655).  On the tensor-core path the shortcut rides Conv_1's K
        loop (b200_conv2d_tc_sc: one launch, no skip tensor); otherwise two convolutions, the second adding the first."""
        from . import tc
        fusable = (not self.dry and self.e.use_tc and self.tape is None and tc.ENABLED and tc.FUSE_SHORTCUT
                   and self.e.fuse_shortcut and (xs.H, xs.W) == (h.H, h.W) and xs.C % 64 == 0
                   and (xs2 is None or xs2.C % 64 == 0) and xs.Ct == xs.C and (xs2 is None or xs2.Ct == xs2.C))
        if fusable:
            probe = self.new(h.H, h.W, cout)
            fusable = tc.conv_supported(self, h, None, probe, 3, 1, (1, 1))
        if not fusable:
            skip = self.conv(xs, key2, cout, 1, x2=xs2, nin=nin2, init_scale=nin2_init)
            return self.conv(h, key1, cout, 3, res=skip, out_scale=out_scale, init_scale=init_scale)
        cx = xs.C + (xs2.C if xs2 is not None else 0)
        if nin2:                                    # same registration order as the two-launch form above
            self.param(key2 + ".W", (cx, cout), ("default", nin2_init))
            self.param(key2 + ".b", (cout,), "zeros")
        else:
            self.param(key2 + ".weight", (cout, cx, 1, 1), ("default", nin2_init))
            self.param(key2 + ".bias", (cout,), "zeros")
        self.param(key1 + ".weight", (cout, h.Ct, 3, 3), ("default", init_scale))
        self.param(key1 + ".bias", (cout,), "zeros")
        out = self.new(h.H, h.W, cout)
        w, bias = self.e.packed_conv_sc(key1, key2, nin2, h.C, cx, cout)
        flops = 2 * self.B * h.H * h.W * cout * (9 * h.Ct + cx)
        gn = None
        if isinstance(h, LazyGN) and h.real is None and tc.conv_gn_supported(self, h.src1, h.src2, out, 3, 1, (1, 1)):
            gn = h
        x = None if gn is not None else (self.materialize(h) if isinstance(h, LazyGN) else h)
        tc.emit_conv_sc(self, x, gn, xs, xs2, w, bias, None, out_scale, out, flops)
        self.e.tc_convs += 1
        self.e.sc_fused += 1
        if gn is not None:
            self.e.gn_fused += 1
        return out

    def dropout(self, x):
        """nn.Dropout (models/layerspp.py:265, layers.py:646): identity except in a training plan with p > 0."""
        p = getattr(self.e, "train_dropout", 0.0)
        if self.dry or self.tape is None or p <= 0.0:
            return x
        if isinstance(x, LazyGN):
            x = self.materialize(x)
        out = self.new(x.H, x.W, x.C)
        k = self.e.dropout_layers
        self.e.dropout_layers += 1
        n = self.B * x.H * x.W * x.C
        self.emit("b200_dropout", out, x, self.dt, n, float(p), self.e.train_seed, k << 32, self.e.train_step.data_ptr())
        self.rec("dropout", x=x, out=out, n=n, p=float(p), off=k << 32)
        return out

    def upfirdn(self, x, up, down, pad):
        Ho = (x.H * up + pad[0] + pad[1] - 4) // down + 1
        Wo = (x.W * up + pad[0] + pad[1] - 4) // down + 1
        out = self.new(Ho, Wo, x.C, x.Ct)
        if self.dry:
            return out
        k = self.e.fir_kernel(up)
        self.emit("b200_upfirdn2d_nhwc", x, k.data_ptr(), out, self.dt, self.B, x.H, x.W, x.C, 4, 4, up, down,
                  pad[0], pad[1])
        self.rec("upfirdn", x=x, out=out, up=up, down=down, pad=tuple(pad))
        return out

    def nearest_up2(self, x):
        out = self.new(x.H * 2, x.W * 2, x.C)
        if not self.dry:
            self.emit("b200_nearest_up2_nhwc", x, out, self.dt, self.B, x.H, x.W, x.C)
            self.rec("nearest_up2", x=x, out=out)
        return out

    def attention_core(self, qkv, C):
        """softmax(q k^T / sqrt(C)) v from a fused [B, L, 3C] projection."""
        out = self.new(qkv.H, qkv.W, C)
        if self.dry:
            return out
        L = qkv.H * qkv.W
        es = 2 if self.dt == DT_BF16 else 4
        self.e.emit_attention(self, qkv, out, L, C, es)
        self.rec("attention", qkv=qkv, out=out, L=L, C=C)
        return out


class _SpecOnly:
    """Stand-in engine for dry walks (parameter specification on machines without a GPU)."""

    def __init__(self, config, dt=DT_BF16):
        self.config, self.dt, self.workspace = config, dt, None


def parameter_spec(config, net_builder):
    """[(key, shape, init)] of the network in reference order, and the Dense_0 layout; needs no device."""
    b = Builder(_SpecOnly(config), 1, dry=True)
    net_builder(b, config)
    return b.specs, list(b.dense)


class ScoreEngine:
    """Holds packed weights + per-batch-size plans for one network on one device."""

    def __init__(self, config, net_builder, device, precision="bf16", use_tensor_cores=True):
        if torch.device(device).type != "cuda":
            raise RuntimeError("ScoreEngine needs a CUDA device; there is no CPU fallback")
        self.config = config
        self.net_builder = net_builder
        self.device = torch.device(device)
        self.precision = precision
        # "bf16": tensor-core mode (bf16 operands, fp32 accumulate).  "fp32": strict mode -- fp32 activations and weights;
        # convolutions on the tensor cores through the three-term tf32 split with grouped fp32 accumulation
        # (csrc/conv_tc32.cu: as close to the exact convolution as cuDNN's fp32 kernels), everything else on CUDA cores.
        # "tf32": the same with the single tf32 product (what cudnn.allow_tf32 gives the reference).  "fp32_simt": strict
        # mode with every convolution on CUDA cores (the reference point the split is tested against; B200_TC32=0).
        self.dt = {"bf16": DT_BF16, "fp32": DT_F32, "tf32": DT_F32, "fp32_simt": DT_F32}[precision]
        self.use_tc = bool(use_tensor_cores) and self.dt == DT_BF16
        self.tc32_terms = {"fp32": 3, "tf32": 1}.get(precision, 0) if bool(use_tensor_cores) else 0
        if os.environ.get("B200_TC32", "1") == "0":
            self.tc32_terms = 0
        self.tc32_convs = 0
        self.fuse_gn_stats = True           # GroupNorm statistics as a by-product of the producing conv's epilogue
        # GroupNorm + SiLU applied inside the consuming conv's operand pipeline: True = where it is the faster form
        # (b200_conv2d_tc_gn_preferred), "all" = wherever the fused kernel can run, False = never
        # (B200_FUSE_GN=0 / all: A/B measurements)
        self.fuse_gn_conv = {"0": False, "all": "all"}.get(os.environ.get("B200_FUSE_GN", "1"), True)
        self.gn_fused = 0
        # the 1x1 shortcut of a residual block folded into its second conv's K loop (B200_FUSE_SC=0: A/B measurements)
        self.fuse_shortcut = os.environ.get("B200_FUSE_SC", "1") != "0"
        self.sc_fused = 0
        self.workspace = Workspace(self.device)
        self.params = {}           # reference-layout fp32 tensors on device
        self._packed = {}
        self._splits = {}          # packed fp32 weight (by address) -> (w, tf32 high part, low part): strict-mode tensor cores
        self._recipes = {}         # cache key -> (method name, args): how to rebuild a packed tensor (refresh_packed)
        self.train_dropout, self.train_seed, self.train_step, self.dropout_layers = 0.0, 0, None, 0
        self._dev = {}
        self._fir = {}
        self.plans = {}
        self.graphs = {}
        self.version = 0
        self._spec = None
        self.dense_layout = []
        self.dense_total = 0
        self.tc_convs = 0
        self.simt_convs = 0

    # ---- parameters -------------------------------------------------------------------------------------------
    def spec(self):
        """Parameter specification [(key, shape, init)] + layout of the batched Dense_0 projection."""
        if self._spec is None:
            self._spec, self.dense_layout = parameter_spec(self.config, self.net_builder)
            self.dense_total = sum(c for _, c in self.dense_layout)
        return self._spec

    def packed_dense(self):
        """All Dense_0 layers stacked: W [sum(out_ch), 4nf], b [sum(out_ch)] (fp32)."""
        if "dense" not in self._packed:
            w = torch.cat([self.params[k + ".weight"] for k, _ in self.dense_layout], 0).contiguous()
            bb = torch.cat([self.params[k + ".bias"] for k, _ in self.dense_layout], 0).contiguous()
            self._packed["dense"] = (w, bb)
            self._recipes["dense"] = ("packed_dense", ())
        return self._packed["dense"]

    def packed_qkv(self, p, C):
        """NIN_0/1/2 stacked into one [3C][1][1][C] weight (+ [3C] bias): q, k, v in one GEMM."""
        pk = (p, "qkv")
        if pk not in self._packed:
            w = torch.cat([self.params[p + ".NIN_%d.W" % i].t() for i in range(3)], 0)   # [3C, C]
            bb = torch.cat([self.params[p + ".NIN_%d.b" % i] for i in range(3)], 0)
            self._packed[pk] = (w.reshape(3 * C, 1, 1, C).to(_TORCH_DT[self.dt]).contiguous(), bb.contiguous())
            self._recipes[pk] = ("packed_qkv", (p, C))
        return self._packed[pk]

    def load_state_dict(self, sd, strict=True):
        spec = self.spec()
        new = {}
        for key, shape, _ in spec:
            src = None
            for cand in (key, "module." + key):
                if cand in sd:
                    src = sd[cand]
                    break
            if src is None:
                if strict:
                    raise KeyError("missing parameter %s" % key)
                continue
            if tuple(src.shape) != tuple(shape):
                raise RuntimeError("shape mismatch for %s: %s vs %s" % (key, tuple(src.shape), shape))
            new[key] = src.detach().to(device=self.device, dtype=torch.float32).contiguous()
        self.params.update(new)
        self._packed.clear()
        self._recipes.clear()
        self._splits.clear()
        self._refresh_cl = None
        self._dev.clear()
        self.plans.clear()
        self.graphs.clear()
        self.workspace = Workspace(self.device)
        self.version += 1

    def state_dict(self):
        return dict(self.params)

    def dev(self, key):
        return self.params[key].data_ptr()

    def fir_kernel(self, up):
        """outer([1,3,3,1]) normalised, times up^2 (models/up_or_down_sampling.py:181-188, 221, 254)."""
        if up not in self._fir:
            k1 = torch.tensor(self.config.model.fir_kernel, dtype=torch.float32)
            k = torch.outer(k1, k1)
            k = k / k.sum() * (up ** 2)
            self._fir[up] = k.to(self.device).contiguous()
        return self._fir[up]

    def packed_conv(self, key, nin, cin_store, cout_store):
        """[Cout_store][KH][KW][Cin_store] in the activation dtype, zero padded."""
        pk = (key, cin_store, cout_store)
        if pk not in self._packed:
            if nin:
                w = self.params[key + ".W"].t().contiguous()[:, :, None, None]   # [Cout, Cin, 1, 1]
            else:
                w = self.params[key + ".weight"]
            cout, cin, kh, kw = w.shape
            p = torch.zeros(cout_store, kh, kw, cin_store, device=self.device, dtype=torch.float32)
            p[:cout, :, :, :cin] = w.permute(0, 2, 3, 1)
            self._packed[pk] = p.to(_TORCH_DT[self.dt]).contiguous()
            self._recipes[pk] = ("packed_conv", (key, nin, cin_store, cout_store))
        return self._packed[pk]

    def split_tf32(self, w):
        """(tf32(w), w - tf32(w)) of a packed fp32 weight tensor: the operands of b200_conv2d_tc32 (cached per tensor;
        `refresh_packed` recomputes them after the parameters have changed in place)."""
        ent = self._splits.get(w.data_ptr())
        if ent is None:
            ent = (w, torch.empty_like(w), torch.empty_like(w))
            self._splits[w.data_ptr()] = ent
            self._run_split(ent)
        return ent[1], ent[2]

    @staticmethod
    def _run_split(ent):
        w, hi, lo = ent
        _lib.check(_lib.lib().b200_split_tf32(_lib.ptr(w), _lib.ptr(hi), _lib.ptr(lo), w.numel(), _lib.stream()),
                   "b200_split_tf32")

    def packed_conv_sc(self, key1, key2, nin2, cin, cx, cout):
        """[W1 | W2]: [Cout][9 cin + cx] in the activation dtype (K contiguous) and the summed bias b1 + b2 (fp32)."""
        pk = (key1, key2, "sc")
        if pk not in self._packed:
            w1 = self.params[key1 + ".weight"]                                   # [Cout, Cin, 3, 3]
            if nin2:
                w2 = self.params[key2 + ".W"].t().contiguous()                   # [Cout, Cx]
                b2 = self.params[key2 + ".b"]
            else:
                w2 = self.params[key2 + ".weight"].reshape(cout, -1)             # [Cout, Cx, 1, 1] -> [Cout, Cx]
                b2 = self.params[key2 + ".bias"]
            assert w1.shape == (cout, cin, 3, 3) and w2.shape == (cout, cx), (w1.shape, w2.shape)
            w = torch.cat([w1.permute(0, 2, 3, 1).reshape(cout, 9 * cin), w2], 1)
            self._packed[pk] = (w.to(_TORCH_DT[self.dt]).contiguous(), (self.params[key1 + ".bias"] + b2).contiguous())
            self._recipes[pk] = ("packed_conv_sc", (key1, key2, nin2, cin, cx, cout))
        return self._packed[pk]

    # ---- refresh after an optimizer step: one descriptor table instead of ~600 torch copies ------------------------
    def _desc_packed_conv(self, cl, p, key, nin, cin_store, cout_store):
        if nin:
            w = self.params[key + ".W"]                                          # [Cin, Cout]
            cl.add(p[:w.shape[1], 0, 0, :w.shape[0]], w.t())
        else:
            w = self.params[key + ".weight"]                                     # [Cout, Cin, kh, kw]
            cl.add(p[:w.shape[0], :, :, :w.shape[1]], w.permute(0, 2, 3, 1))

    def _desc_packed_bias(self, cl, p, key, nin, cout_store):
        b = self.params[key + (".b" if nin else ".bias")]
        cl.add(p[:b.numel()], b)

    def _desc_packed_qkv(self, cl, packed, p, C):
        w, bb = packed
        for i in range(3):
            cl.add(w[i * C:(i + 1) * C, 0, 0, :], self.params[p + ".NIN_%d.W" % i].t())
            cl.add(bb[i * C:(i + 1) * C], self.params[p + ".NIN_%d.b" % i])

    def _desc_packed_dense(self, cl, packed):
        w, bb = packed
        row = 0
        for k, oc in self.dense_layout:
            cl.add(w[row:row + oc], self.params[k + ".weight"])
            cl.add(bb[row:row + oc], self.params[k + ".bias"])
            row += oc

    def _desc_packed_conv_sc(self, cl, packed, key1, key2, nin2, cin, cx, cout):
        w, bias = packed
        cl.add(w[:, :9 * cin].unflatten(1, (3, 3, cin)), self.params[key1 + ".weight"].permute(0, 2, 3, 1))
        if nin2:
            cl.add(w[:, 9 * cin:], self.params[key2 + ".W"].t())
            b2 = self.params[key2 + ".b"]
        else:
            cl.add(w[:, 9 * cin:], self.params[key2 + ".weight"].reshape(cout, cx))
            b2 = self.params[key2 + ".bias"]
        cl.add(bias, self.params[key1 + ".bias"], src2=b2)

    def _build_refresh_list(self):
        from .copylist import CopyList
        cl, slow = CopyList(self.device), []
        for pk, (name, args) in self._recipes.items():
            fn = getattr(self, "_desc_" + name, None)
            if fn is None:
                slow.append(pk)
            else:
                fn(cl, self._packed[pk], *args)
        self._refresh_cl = (cl.finalize(), slow, len(self._recipes), tuple(t.data_ptr() for t in self.params.values()))

    def refresh_packed(self):
        """Recomputes every cached packed tensor from `self.params` INTO ITS EXISTING STORAGE (plans hold raw pointers
        to them): called after an optimizer step has changed the parameters in place.  The re-packing recipes are all
        strided copies with a type change; they are compiled once into a descriptor table (models/copylist.py) and
        executed by ONE launch -- as separate torch calls they were 6 ms of host time per training step."""
        if os.environ.get("B200_FAST_REFRESH", "1") != "0":
            st = getattr(self, "_refresh_cl", None)
            if st is None or st[2] != len(self._recipes) or st[3] != tuple(t.data_ptr() for t in self.params.values()):
                self._build_refresh_list()
                st = self._refresh_cl
            st[0].run()
            self._refresh_torch(st[1])
        else:
            self._refresh_torch(list(self._recipes))
        for ent in self._splits.values():
            self._run_split(ent)

    def _refresh_torch(self, keys):
        for pk in keys:
            name, args = self._recipes[pk]
            old = self._packed.pop(pk)
            new = getattr(self, name)(*args)
            for o, n in zip(old if isinstance(old, tuple) else (old,), new if isinstance(new, tuple) else (new,)):
                o.copy_(n)
            self._packed[pk] = old

    def packed_bias(self, key, nin, cout_store):
        pk = (key, "bias", cout_store)
        if pk not in self._packed:
            b = self.params[key + (".b" if nin else ".bias")]
            p = torch.zeros(cout_store, device=self.device, dtype=torch.float32)
            p[:b.numel()] = b
            self._packed[pk] = p
            self._recipes[pk] = ("packed_bias", (key, nin, cout_store))
        return self._packed[pk]

    # ---- kernel selection -------------------------------------------------------------------------------------
    def emit_conv(self, b, x, x2, w, bias, temb_ptr, res, out_scale, out, ksize, stride, pad, key):
        from . import tc
        if (isinstance(x, LazyGN) and x.real is None and x2 is None and self.use_tc
                and tc.conv_gn_supported(b, x.src1, x.src2, out, ksize, stride, tuple(pad))):
            tc.emit_conv_gn(b, x, w, bias, temb_ptr, res, out_scale, out, ksize, self.conv_flops(b, x, x2, out, ksize))
            self.tc_convs += 1
            self.gn_fused += 1
            return
        if (self.use_tc and stride == 2 and ksize == 3 and tuple(pad) == (0, 0) and x2 is None and temb_ptr is None
                and tc.s2_supported(b, x, out)):
            tc.emit_s2(b, x, w, bias, res, out_scale, out, self.conv_flops(b, x, x2, out, ksize))
            self.tc_convs += 1
            return
        if (self.use_tc and stride == 1 and tuple(pad) == (ksize // 2, ksize // 2) and x2 is None and temb_ptr is None
                and tc.narrow_supported(b, x, out, ksize)):
            tc.emit_narrow(b, x, w, bias, res, out_scale, out, ksize, self.conv_flops(b, x, x2, out, ksize))
            self.tc_convs += 1
            return
        if self.use_tc and tc.conv_supported(b, x, x2, out, ksize, stride, tuple(pad)):
            tc.emit_conv(b, x, x2, w, bias, temb_ptr, res, out_scale, out, ksize, self.conv_flops(b, x, x2, out, ksize))
            self.tc_convs += 1
            return
        if self.tc32_terms and self.dt == DT_F32 and tc.conv32_supported(b, x, x2, out, ksize, stride, tuple(pad)):
            hi, lo = self.split_tf32(w)
            tc.emit_conv32(b, x, x2, hi, lo, bias, temb_ptr, res, out_scale, out, ksize, self.tc32_terms,
                           self.conv_flops(b, x, x2, out, ksize))
            self.tc32_convs += 1
            return
        self.simt_convs += 1
        p2, c2 = (x2, x2.C) if x2 is not None else (None, 0)
        b.emit("b200_conv2d_simt", x, x.C, p2, c2, w.data_ptr(), bias.data_ptr(), temb_ptr, b.temb_stride,
               res, out_scale, out, 0, None, self.dt, b.B, x.H, x.W, out.H, out.W,
               out.C, ksize, ksize, stride, pad[0], pad[1], flops=self.conv_flops(b, x, x2, out, ksize))

    @staticmethod
    def conv_flops(b, x, x2, out, ksize):
        """Algorithmic FLOPs (2 MAC) over the channels that carry data (padding lanes do not count)."""
        cin = x.Ct + (x2.Ct if x2 is not None else 0)
        return 2 * b.B * out.H * out.W * out.Ct * cin * ksize * ksize

    def emit_attention(self, b, qkv, out, L, C, es):
        from . import tc
        if self.use_tc and tc.attention_supported(L, C):
            tc.emit_attention(b, qkv, out, L, C)
            return
        if self.tc32_terms and self.dt == DT_F32 and tc.attention32_supported(L, C):
            tc.emit_attention32(b, qkv, out, L, C, self.tc32_terms)
            return
        b.emit("b200_attention_simt", qkv, qkv.at(C * es), qkv.at(2 * C * es), out, self.dt, b.B, L, C, 3 * C, 3 * C,
               3 * C, C, float(int(C) ** (-0.5)), flops=4 * b.B * L * L * C)

    # ---- plans ------------------------------------------------------------------------------------------------
    def plan(self, B, grad=False, uniform=False, train=False):
        """The launch list for batch size B.  grad=True: forward (activations retained) followed by the input-gradient
        pass (models/backward.py); `plan.n_fwd` splits the two.  uniform=True: every sample carries the same label
        (the fused PC loop), so the time-embedding MLP runs on one row."""
        # train=True: a grad plan that also emits parameter gradients (and dropout): models/train.py
        grad = grad or train
        key = (B, "train") if train else ((B, "grad") if grad else ((B, "uniform") if uniform else B))
        if key not in self.plans:
            if not self.params:
                raise RuntimeError("ScoreEngine has no weights; call load_state_dict / init_weights first")
            with torch.cuda.device(self.device):
                self.spec()
                self.grad_mode = grad
                self.param_grads = train
                self.uniform_mode = uniform and not grad
                try:
                    b = Builder(self, B, dry=False)
                    if grad:
                        # the retained activations of a grad plan must survive whatever other plan runs between
                        # forward_for_grad and input_grad (a `model(x, t)` call, a PC step): private buffers, not the
                        # engine-wide free list that every other plan is packed into
                        b.ws = Workspace(self.device)
                        b.plan.keep.append(b.ws)       # the plan owns it: its launch arguments are raw pointers into it
                    self.net_builder(b, self.config)
                    b.plan.n_fwd = len(b.plan.ops)
                    if grad:
                        from . import backward
                        backward.emit_backward(b)
                    b.finalize()
                finally:
                    self.grad_mode = False
                    self.param_grads = False
                    self.uniform_mode = False
                self.plans[key] = b.plan
        return self.plans[key]

    # ---- CUDA-graph replay of a plan's launch list ---------------------------------------------------------------------
    # The ODE / likelihood / DPS solvers call the network hundreds of times at small batches, where a plan is 200-700
    # launches of a few microseconds each and the host's ctypes calls, not the GPU, set the pace (C4 DPS right-hand side at
    # B = 16: 678 launches).  A plan touches only its own static buffers, so its launch list is captured once (on the
    # third run: the first two warm up attribute / descriptor caches) and replayed.  B200_ENGINE_GRAPH=0 turns it off.
    def _run_ops(self, p, key, ops):
        st = _lib.stream()
        if os.environ.get("B200_ENGINE_GRAPH", "1") == "0" or torch.cuda.is_current_stream_capturing():
            for fn, args, name in ops:
                rc = fn(*args, st)
                if rc != 0:
                    _lib.check(rc, name)
            return
        graphs = p.__dict__.setdefault("_graphs", {})
        ent = graphs.get(key)
        if ent is None:
            ent = graphs[key] = [0, None]
        if ent[1] is None:
            ent[0] += 1
            if ent[0] <= 2:
                for fn, args, name in ops:
                    rc = fn(*args, st)
                    if rc != 0:
                        _lib.check(rc, name)
                return
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                cst = _lib.stream()
                for fn, args, name in ops:
                    rc = fn(*args, cst)
                    if rc != 0:
                        _lib.check(rc, name)
            ent[1] = g
        ent[1].replay()

    def forward_for_grad(self, x, labels):
        """Forward pass that keeps every activation; returns the network output [B, C, H, W] fp32 (a view of the
        plan's output buffer).  Follow with `input_grad(v)`."""
        _lib.require_cuda(x, labels)
        B = x.shape[0]
        p = self.plan(B, grad=True)
        with torch.cuda.device(self.device):
            p.x_in.copy_(x)
            p.labels.copy_(labels.to(torch.float32).reshape(B))
            self._run_ops(p, "fwd", p.ops[:p.n_fwd])
        self._grad_B = B
        p.grad_ready = True
        return p.out

    def input_grad(self, v):
        """J^T v: gradient w.r.t. the input image of <model(x, labels), v> for the x of the last forward_for_grad."""
        _lib.require_cuda(v)
        p = self.plans[(self._grad_B, "grad")]
        if not getattr(p, "grad_ready", False):
            # the backward half reuses the buffers of activations it has consumed (liveness-packed workspace)
            raise RuntimeError("input_grad: one call per forward_for_grad (the pass consumes the retained activations)")
        p.grad_ready = False
        with torch.cuda.device(self.device):
            torch.mul(v.to(torch.float32), p.sample_mul_t.view(-1, 1, 1, 1), out=p.cot_in)
            self._run_ops(p, "bwd", p.ops[p.n_fwd:])
            return p.grad_out * p.grad_scale

    def forward(self, x, labels, out=None):
        """x: [B, C, H, W] fp32 CUDA; labels: [B] (sigma for Fourier nets, timestep index/continuous otherwise)."""
        _lib.require_cuda(x, labels)
        B = x.shape[0]
        p = self.plan(B)
        with torch.cuda.device(self.device):
            p.x_in.copy_(x)
            p.labels.copy_(labels.to(torch.float32).reshape(B))
            self._run_ops(p, "all", p.ops)
            if out is None:
                return p.out.clone()
            out.copy_(p.out)
            return out

    def run_inplace(self, B):
        """Runs the plan on whatever is already in plan.x_in / plan.labels (used by the fused sampler)."""
        p = self.plans[B]
        p.run(_lib.stream())
        return p.out

    __call__ = forward
