"""Input-gradient pass of the score engine: J_model(x)^T v as a second launch list behind the forward one.

What DPS needs (reference inverse/conditional_sampling.py:116-136): `torch.autograd.grad(logp, x_t)` through
x0_hat = x_t / mean + std^2 * score(x_t) is  v / mean + std^2 * J_score^T v  with  v = d logp / d x0_hat.  No weight
gradients are ever taken, so the pass is: walk the primitives the forward recorded (Builder.rec) in reverse and emit,
per primitive, its backward-data launches into the same plan.  Activations stay alive because the liveness scan in
Builder.finalize sees their uses in the backward half.

  conv            backward-data = the SAME convolution entry points (tcgen05 or SIMT) on flipped / transposed weights
                  packed here; `res` of a forward conv receives out_scale * dy (b200_axpby); stride-2 convs go through
                  b200_zero_stuff2_nhwc so that their backward-data is a stride-1 "same" convolution
  GroupNorm+SiLU  b200_gn_partial (statistics of x) -> b200_gn_bwd_partial -> b200_gn_bwd_apply (two-source aware)
  attention       b200_attention_bwd
  nearest x2      b200_nearest_up2_bwd_nhwc
  upfirdn2d       its adjoint is upfirdn2d with up <-> down, flipped (symmetric) kernel and re-derived padding
                  (op/upfirdn2d.py:24-36 computes the same numbers for the reference's backward)
  image boundary  cotangent NCHW fp32 -> channels-last; image gradient channels-last -> NCHW fp32
"""
import torch

from .. import _lib
from .._lib import c_float, c_int, c_ll, c_vp
from .engine import DT_BF16, _TORCH_DT, Act

_lib.register_signature("b200_gn_bwd_partial", [c_vp, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp, c_vp,
                                                c_float, c_int, c_vp, c_vp, c_vp])
_lib.register_signature("b200_gn_bwd_apply", [c_vp, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp, c_vp,
                                              c_vp, c_float, c_int, c_vp, c_vp, c_int, c_vp, c_int, c_vp])
_lib.register_signature("b200_axpby", [c_vp, c_float, c_vp, c_float, c_vp, c_int, c_ll, c_vp])
_lib.register_signature("b200_nearest_up2_bwd_nhwc", [c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_vp])
_lib.register_signature("b200_zero_stuff2_nhwc", [c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_vp])
_lib.register_signature("b200_attention_bwd", [c_vp, c_vp, c_vp, c_vp, c_int, c_int, c_int, c_int, c_float, c_vp])


class _Grads:
    """Gradient activations keyed by the forward activation they belong to; first contribution writes, later ones add."""

    def __init__(self, b):
        self.b = b
        self.g = {}

    def has(self, act):
        return id(act) in self.g

    def get(self, act):
        return self.g[id(act)]

    def target(self, act):
        """(gradient activation, accumulate?) for a new contribution to d(act).  Zero-padded 64-wide images (the
        tensor-core stem input and its FIR pyramid) get an 8-wide gradient: only their first channels carry one."""
        if id(act) in self.g:
            return self.g[id(act)], 1
        C = 8 if (act.Ct < act.C and act.C == 64) else act.C
        g = Act(None, act.B, act.H, act.W, C, act.Ct)
        self.g[id(act)] = g
        return g, 0


def _wt(b, wf, ksize, cin_off, cin_len, scale, rows, k_pad):
    """Backward-data weight [rows][k][k][k_pad] from the forward packed weight wf [cs][k][k][cin_total]:
    Wt[ci][kh][kw][co] = scale * wf[co][k-1-kh][k-1-kw][cin_off + ci]; rows >= cin_len and k_pad >= cs are zero padded."""
    cs = wf.shape[0]
    w = wf.view(cs, ksize, ksize, -1)[:, :, :, cin_off:cin_off + cin_len].to(torch.float32)
    w = w.flip(1, 2).permute(3, 1, 2, 0) * scale
    out = torch.zeros(rows, ksize, ksize, k_pad, device=wf.device, dtype=torch.float32)
    out[:cin_len, :, :, :cs] = w
    out = out.to(_TORCH_DT[b.dt]).contiguous()
    b.plan.keep.append(out)
    return out


def _dgrad(b, gy, wf, ksize, cin_off, cin_len, scale, gx, acc):
    """gx (+)= conv_same(gy, Wt): stride-1 backward-data of a ksize x ksize convolution."""
    from . import tc
    e = b.e
    flops = 2 * b.B * gy.H * gy.W * gx.C * gy.C * ksize * ksize
    if e.use_tc and gx.C == 8 and tc.narrow_supported(b, gy, gx, ksize):
        w = _wt(b, wf, ksize, cin_off, cin_len, scale, 8, gy.C)
        tc.emit_narrow(b, gy, w, _zero_bias(b, 8), gx if acc else None, 1.0, gx, ksize, flops)
        return
    w = _wt(b, wf, ksize, cin_off, cin_len, scale, gx.C, gy.C)
    res = gx if acc else None
    if e.use_tc and tc.conv_supported(b, gy, None, gx, ksize, 1, (ksize // 2, ksize // 2)):
        b.emit("b200_conv2d_tc", gy, gy.C, None, 0, w.data_ptr(), None, None, 0, res, 1.0, gx, 0, None, None, b.B, gy.H,
               gy.W, gx.C, ksize, flops=flops, tag="dgrad %dx%d %dx%d %d->%d" % (ksize, ksize, gy.H, gy.W, gy.C, gx.C))
        return
    b.emit("b200_conv2d_simt", gy, gy.C, None, 0, w.data_ptr(), None, None, 0, res, 1.0, gx, 0, None, b.dt, b.B, gy.H,
           gy.W, gx.H, gx.W, gx.C, ksize, ksize, 1, ksize // 2, ksize // 2, flops=flops)


def _zero_bias(b, n):
    z = getattr(b, "_zero_bias_t", None)
    if z is None or z.numel() < n:
        z = torch.zeros(max(n, 512), device=b.e.device, dtype=torch.float32)
        b._zero_bias_t = z
        b.plan.keep.append(z)
    return z


def _conv_bwd(b, G, t):
    x, x2, out, wf, k, stride, pad = t["x"], t["x2"], t["out"], t["w"], t["ksize"], t["stride"], t["pad"]
    if t["head"]:
        # network head: out = (conv + bias) * out_scale * sample_mul[b]  (NCHW fp32).  The cotangent arrives in
        # plan.cot_in already multiplied by sample_mul (engine.input_grad); out_scale is folded into the layout change.
        plan = b.plan
        cp = 64 if b.e.use_tc else 8
        gy = Act(None, b.B, x.H, x.W, cp, t["cout"])
        plan.cot_in = torch.zeros(b.B, t["cout"], x.H, x.W, device=b.e.device, dtype=torch.float32)
        b.emit("b200_nchw_to_nhwc", plan.cot_in.data_ptr(), gy, b.dt, b.B, t["cout"], x.H, x.W, cp, float(t["out_scale"]),
               0.0)
        gx, acc = G.target(x)
        _dgrad(b, gy, wf, k, 0, x.C, 1.0, gx, acc)
        return
    if not G.has(out):
        return                                         # nothing downstream depends on this convolution
    gy = G.get(out)
    scale = float(t["out_scale"])
    if t["res"] is not None:
        gr, acc = G.target(t["res"])
        n = gy.B * gy.H * gy.W * gy.C
        b.emit("b200_axpby", gy, scale, gr if acc else None, 1.0, gr, b.dt, n)
    if stride == 2:
        if k != 3 or tuple(pad) != (0, 0) or x2 is not None:
            raise NotImplementedError("backward of this strided convolution")
        z = Act(None, b.B, x.H, x.W, gy.C)
        b.emit("b200_zero_stuff2_nhwc", gy, z, b.dt, b.B, gy.H, gy.W, x.H, x.W, gy.C)
        gy = z
    elif stride != 1 or tuple(pad) != (k // 2, k // 2):
        raise NotImplementedError("backward of a convolution with stride %r pad %r" % (stride, pad))
    off = 0
    for src in (x, x2):
        if src is None:
            continue
        gx, acc = G.target(src)
        _dgrad(b, gy, wf, k, off, src.Ct if gx.C != src.C else src.C, scale, gx, acc)
        off += src.C


def _gn_bwd(b, G, t):
    x, x2, out = t["x"], t["x2"], t["out"]
    if not G.has(out):
        return
    gy = G.get(out)
    e = b.e
    HW = x.H * x.W
    ns = b.lib.b200_gn_nsplit(b.B, HW)
    groups = t["groups"]
    pf = b.f32(b.B * ns * groups * 2)
    pb = b.f32(b.B * ns * groups * 2)
    p2, c2 = (x2, x2.C) if x2 is not None else (None, 0)
    ga, be = e.dev(t["key"] + ".weight"), e.dev(t["key"] + ".bias")
    act = 1 if t["act"] else 0
    b.emit("b200_gn_partial", x, x.C, p2, c2, b.dt, b.B, HW, groups, ns, pf.data_ptr())
    b.emit("b200_gn_bwd_partial", x, x.C, p2, c2, b.dt, b.B, HW, groups, ns, pf.data_ptr(), ga, be, 1e-6, act, gy,
           pb.data_ptr())
    g1, a1 = G.target(x)
    g2, a2 = G.target(x2) if x2 is not None else (None, 0)
    b.emit("b200_gn_bwd_apply", x, x.C, p2, c2, b.dt, b.B, HW, groups, ns, pf.data_ptr(), pb.data_ptr(), ga, be, 1e-6,
           act, gy, g1, a1, g2, a2)


def _upfirdn_bwd(b, G, t):
    x, out, up, down, pad = t["x"], t["out"], t["up"], t["down"], t["pad"]
    if not G.has(out):
        return
    gy = G.get(out)
    # adjoint of upfirdn2d(up, down, pad) with a symmetric 4x4 kernel: upfirdn2d(up=down, down=up, pad=g_pad) with
    # g_pad0 = kw - pad0 - 1, g_pad1 = in*up - out*down + pad0 - up + 1 (op/upfirdn2d.py:24-36)
    gp0 = 4 - pad[0] - 1
    gp1 = x.H * up - out.H * down + pad[0] - up + 1
    gx, acc = G.target(x)
    k = b.e.fir_kernel(up)      # forward kernel carries the up^2 gain; the adjoint uses the same taps
    if acc:
        tmp = Act(None, x.B, x.H, x.W, gx.C, x.Ct)
        b.emit("b200_upfirdn2d_nhwc", gy, k.data_ptr(), tmp, b.dt, b.B, gy.H, gy.W, gy.C, 4, 4, down, up, gp0, gp1)
        b.emit("b200_axpby", tmp, 1.0, gx, 1.0, gx, b.dt, x.B * x.H * x.W * gx.C)
    else:
        b.emit("b200_upfirdn2d_nhwc", gy, k.data_ptr(), gx, b.dt, b.B, gy.H, gy.W, gy.C, 4, 4, down, up, gp0, gp1)


def emit_backward(b):
    """Appends the input-gradient launches to b.plan (call after the forward has been emitted, before finalize)."""
    plan, G = b.plan, _Grads(b)
    inp = None
    for t in reversed(b.tape):
        kind = t["kind"]
        if kind == "conv":
            _conv_bwd(b, G, t)
        elif kind == "gn":
            _gn_bwd(b, G, t)
        elif kind == "attention":
            if G.has(t["out"]):
                qkv, L, C = t["qkv"], t["L"], t["C"]
                scratch = b.f32(2 * b.B * L * L)
                gq, _ = G.target(qkv)
                b.emit("b200_attention_bwd", qkv, G.get(t["out"]), gq, scratch.data_ptr(), b.dt, b.B, L, C,
                       float(int(C) ** (-0.5)))
        elif kind == "nearest_up2":
            if G.has(t["out"]):
                x = t["x"]
                gx, acc = G.target(x)
                b.emit("b200_nearest_up2_bwd_nhwc", G.get(t["out"]), gx, b.dt, b.B, x.H, x.W, x.C, acc)
        elif kind == "upfirdn":
            _upfirdn_bwd(b, G, t)
        elif kind == "input":
            inp = t
        elif kind == "out_nhwc":
            # progressive='output_skip' head: out = pyramid[..., :channels] * sample_mul (NCHW fp32); the cotangent is
            # already multiplied by sample_mul (engine.input_grad)
            x = t["x"]
            plan.cot_in = torch.zeros(b.B, t["channels"], x.H, x.W, device=b.e.device, dtype=torch.float32)
            gx, acc = G.target(x)
            assert not acc
            b.emit("b200_nchw_to_nhwc", plan.cot_in.data_ptr(), gx, b.dt, b.B, t["channels"], x.H, x.W, gx.C, 1.0, 0.0)
        else:
            raise NotImplementedError("backward of primitive %r" % kind)
    if inp is None:
        raise RuntimeError("the network recorded no input primitive")
    x, xs, ch = inp["x"], inp["x_stem"], inp["channels"]
    srcs = [G.get(a) for a in ((xs,) if x is xs else (xs, x)) if G.has(a)]
    if not srcs:
        raise RuntimeError("no gradient reached the input image")
    g = srcs[0]
    for other in srcs[1:]:
        b.emit("b200_axpby", other, 1.0, g, 1.0, g, b.dt, g.B * g.H * g.W * g.C)
    plan.grad_out = torch.zeros(b.B, ch, x.H, x.W, device=b.e.device, dtype=torch.float32)
    b.emit("b200_nhwc_to_nchw", g, plan.grad_out.data_ptr(), b.dt, b.B, ch, x.H, x.W, g.C, None)
    plan.grad_scale = float(inp["scale"])
    plan.sample_mul_t = b.sample_mul
    _ = DT_BF16
