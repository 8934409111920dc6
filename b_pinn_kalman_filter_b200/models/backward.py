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
import ctypes

import os

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
# training (csrc/train.cu)
_lib.register_signature("b200_transpose2d", [c_vp, c_vp, c_int, c_ll, c_int, c_vp])
_lib.register_signature("b200_im2col_t", [c_vp, c_int, c_vp, c_int, c_vp, c_int] + [c_int] * 8 + [c_vp])
_lib.register_signature("b200_sum_partials", [c_vp, c_vp, c_int, c_int, c_ll, c_vp])
_lib.register_signature("b200_colsum", [c_vp, c_vp, c_int, c_int, c_ll, c_int, c_ll, c_float, c_int, c_vp])
_lib.register_signature("b200_gn_bwd_partial2", [c_vp, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp, c_vp,
                                                 c_float, c_int, c_vp, c_vp, c_vp, c_vp])
_lib.register_signature("b200_gn_param_grad", [c_vp, c_int, c_vp, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp,
                                               c_vp, c_float, c_int, c_vp, c_vp])
_lib.register_signature("b200_gemm_f32", [c_vp, c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_float, c_float, c_vp])
_lib.register_signature("b200_silu_f32", [c_vp, c_vp, c_ll, c_vp])
_lib.register_signature("b200_silu_bwd_f32", [c_vp, c_vp, c_vp, c_ll, c_vp])
_lib.register_signature("b200_dropout", [c_vp, c_vp, c_int, c_ll, c_float, ctypes.c_uint64, ctypes.c_uint64, c_vp, c_vp])


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
    out = _wt_compute(wf, ksize, cin_off, cin_len, scale, rows, k_pad, _TORCH_DT[b.dt])
    b.plan.keep.append(out)
    # training plans re-derive these from the (in place refreshed) forward weights after every optimizer step
    if not hasattr(b.plan, "wt_recipes"):
        b.plan.wt_recipes = []
    b.plan.wt_recipes.append((out, wf, ksize, cin_off, cin_len, scale, rows, k_pad))
    return out


def _wt_compute(wf, ksize, cin_off, cin_len, scale, rows, k_pad, dtype):
    cs = wf.shape[0]
    w = wf.view(cs, ksize, ksize, -1)[:, :, :, cin_off:cin_off + cin_len].to(torch.float32)
    w = w.flip(1, 2).permute(3, 1, 2, 0) * scale
    out = torch.zeros(rows, ksize, ksize, k_pad, device=wf.device, dtype=torch.float32)
    out[:cin_len, :, :, :cs] = w
    return out.to(dtype).contiguous()


def refresh_dgrad_weights(plan):
    """Re-derives the backward-data weights from the (already refreshed) forward packed weights: one b200_copy_batch
    launch over a descriptor table built once per plan (flip along kh, kw; (co, ci) transposed; scale; zero padding is
    never written).  The torch form below (`B200_FAST_REFRESH=0`) was 7.5 ms of host time per C3 training step."""
    recipes = getattr(plan, "wt_recipes", ())
    if not recipes:
        return
    if os.environ.get("B200_FAST_REFRESH", "1") == "0":
        for out, wf, ksize, cin_off, cin_len, scale, rows, k_pad in recipes:
            out.copy_(_wt_compute(wf, ksize, cin_off, cin_len, scale, rows, k_pad, out.dtype))
        return
    st = getattr(plan, "_wt_cl", None)
    if st is None or st[1] != len(recipes):
        from .copylist import CopyList
        cl = CopyList(recipes[0][0].device)
        for out, wf, ksize, cin_off, cin_len, scale, rows, k_pad in recipes:
            cs = wf.shape[0]
            src = wf.view(cs, ksize, ksize, -1)[:, :, :, cin_off:cin_off + cin_len].permute(3, 1, 2, 0)
            cl.add(out[:cin_len, :, :, :cs], src, scale=scale, flip=(1, 2))
        st = plan._wt_cl = (cl.finalize(), len(recipes))
    st[0].run()


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
        if getattr(b.e, "param_grads", False):
            _conv_param_grads(b, t, gy, 1.0)          # out_scale is already folded into gy by the layout change
        gx, acc = G.target(x)
        _dgrad(b, gy, wf, k, 0, x.C, 1.0, gx, acc)
        return
    if not G.has(out):
        return                                         # nothing downstream depends on this convolution
    gy = G.get(out)
    scale = float(t["out_scale"])
    if getattr(b.e, "param_grads", False):
        _conv_param_grads(b, t, gy, scale)
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
    if getattr(e, "param_grads", False):
        # d gamma / d beta are sums over the batch of per-channel sums the backward's first sweep forms anyway
        # (b200_gn_bwd_partial2 leaves them per block); a separate b200_gn_param_grad sweep was 3.7 ms of a C3 step
        C = x.C + c2
        part = b.f32(b.B * ns * C * 2)
        gsum = b.f32(C * 2)
        b.emit("b200_gn_bwd_partial2", x, x.C, p2, c2, b.dt, b.B, HW, groups, ns, pf.data_ptr(), ga, be, 1e-6, act, gy,
               pb.data_ptr(), part.data_ptr())
        b.emit("b200_colsum", part.data_ptr(), gsum.data_ptr(), 0, 1, b.B * ns, 2 * C, 2 * C, 1.0, 0)
        b.plan.pgrads.append(dict(kind="gn", key=t["key"], buf=gsum, C=C))
    else:
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


def _conv_param_grads(b, t, gy, scale):
    """Weight / bias / time-embedding gradients of one convolution (b.e.param_grads plans).

    dW[co][(tap, ci)] = sum_p gy[p][co] x[p + tap][ci]: gy^T [Cout][P] (b200_transpose2d) contracted over P with the
    transposed im2col [(tap, ci)][P] (b200_im2col_t) by the convolution kernels themselves -- a 1x1 "conv" whose fake image
    is [1, 1, Cout, P] -- so the result lands in the packed weight layout [Cout][kh][kw][Cin_store] (activation dtype).
    d bias = column sums of gy; the per-sample column sums are also the gradient of the block's Dense_0 output."""
    from . import tc
    x, x2, k, stride, pad = t["x"], t["x2"], t["ksize"], t["stride"], t["pad"]
    Ho, Wo = t["out_hw"]
    P = b.B * Ho * Wo
    cin_s = x.C + (x2.C if x2 is not None else 0)
    cout_s = gy.C
    kout = k * k * cin_s
    # column count padded to the tensor-core tile (stem-type shapes: 9 * 64 = 576 -> 640); the extra rows of the
    # transposed im2col stay zero (allocated zeroed, never written) and the extra gradient columns are ignored
    kpad = (kout + 127) // 128 * 128 if b.e.use_tc else kout
    gyT = Act(None, 1, 1, cout_s, P)
    if kpad != kout:
        xT_t = torch.zeros(kpad * P, device=b.e.device, dtype=_TORCH_DT[b.dt])
        b.plan.keep.append(xT_t)
        xT = xT_t.data_ptr()
    else:
        xT = Act(None, 1, 1, kout, P)
    # split K = pixels over S "images" of a batched 1x1 convolution (b200_gemm_tc_splitk) so that tiles x S CTAs, not 9-18,
    # share the loop; the bf16 partial products are summed in fp32 by b200_sum_partials
    split = 0
    if (b.e.use_tc and tc.ENABLED and os.environ.get("B200_WGRAD_SPLITK", "1") != "0" and cout_s % 128 == 0
            and kpad % 128 == 0 and P % 64 == 0):
        tiles = (cout_s // 128) * (kpad // (256 if kpad % 256 == 0 else 128))
        split = 1
        while split < 64 and tiles * split < 120 and P % (2 * split * 64) == 0 and P // (2 * split) >= 512:
            split *= 2
        if split == 1:
            split = 0
    gw = torch.zeros(cout_s, kpad, device=b.e.device, dtype=torch.float32 if split else _TORCH_DT[b.dt])
    b.plan.keep.append(gw)
    p2, c2 = (x2, x2.C) if x2 is not None else (None, 0)
    flops = 2 * P * cout_s * kout
    direct = (split and stride == 1 and tuple(pad) == (k // 2, k // 2) and (Ho, Wo) == (x.H, x.W) and kpad == kout
              and os.environ.get("B200_WGRAD_DIRECT", "1") != "0"
              and _lib.lib().b200_wgrad_tc_supported(cout_s, x.C, c2, b.B, x.H, x.W, k, split))
    if direct:
        # straight from the channels-last tensors: MN-major operands, tap shifts as TMA box coordinates (b200_wgrad_tc)
        part = Act(None, split, 1, cout_s, kpad)
        b.emit("b200_wgrad_tc", gy, x, x.C, p2, c2, part, cout_s, b.B, x.H, x.W, k, split, flops=flops,
               tag="wgrad %dx%d %dx%d %d->%d direct split %d" % (k, k, Ho, Wo, cin_s, cout_s, split))
        b.emit("b200_sum_partials", part, gw.data_ptr(), b.dt, split, cout_s * kpad)
    else:
        b.emit("b200_transpose2d", gy, gyT, b.dt, P, cout_s)
        b.emit("b200_im2col_t", x, x.C, p2, c2, xT, b.dt, b.B, x.H, x.W, Ho, Wo, k, stride, pad[0])
    if direct:
        pass
    elif split:
        part = Act(None, split, 1, cout_s, kpad)
        b.emit("b200_gemm_tc_splitk", gyT, P, xT, P, part, cout_s, kpad, P, split, flops=flops,
               tag="wgrad %dx%d %dx%d %d->%d split %d" % (k, k, Ho, Wo, cin_s, cout_s, split))
        b.emit("b200_sum_partials", part, gw.data_ptr(), b.dt, split, cout_s * kpad)
    elif b.e.use_tc and tc.ENABLED and _lib.lib().b200_conv2d_tc_supported(P, 0, 1, 1, cout_s, kpad, 1, 1, 0):
        b.emit("b200_conv2d_tc", gyT, P, None, 0, xT, None, None, 0, None, 1.0, gw.data_ptr(), 0, None, None, 1, 1, cout_s,
               kpad, 1, flops=flops, tag="wgrad %dx%d %dx%d %d->%d" % (k, k, Ho, Wo, cin_s, cout_s))
    else:
        b.emit("b200_conv2d_simt", gyT, P, None, 0, xT, None, None, 0, None, 1.0, gw.data_ptr(), 0, None, b.dt, 1, 1,
               cout_s, 1, cout_s, kpad, 1, 1, 1, 0, 0, flops=flops)
    # bias: per-sample column sums, then over the batch
    per = b.f32(b.B * cout_s)
    gb = b.f32(cout_s)
    b.emit("b200_colsum", gy, per.data_ptr(), b.dt, b.B, Ho * Wo, cout_s, cout_s, scale, 0)
    b.emit("b200_colsum", per.data_ptr(), gb.data_ptr(), 0, 1, b.B, cout_s, cout_s, 1.0, 0)
    if t.get("temb_off") is not None:
        tm = b.plan.temb
        if tm["rows"] != b.B:
            raise RuntimeError("training plans need per-sample time embeddings")
        dst = b.plan.d_temb_all.data_ptr() + 4 * t["temb_off"]
        b.emit("b200_colsum", gy, dst, b.dt, b.B, Ho * Wo, t["cout"], tm["total"], scale, 0)
    cin_true = x.Ct + (x2.Ct if x2 is not None else 0)
    if x2 is not None and (x.Ct != x.C or x2.Ct != x2.C):
        raise NotImplementedError("weight gradient of a two-source conv with padded channels")
    b.plan.pgrads.append(dict(kind="conv", key=t["key"], nin=t["nin"], gw=gw, gb=gb, scale=scale, cout=t["cout"],
                              cin=cin_true, cin_s=cin_s, k=k, kout=kout))


def _temb_param_grads(b):
    """Backward of the time-embedding MLP and the batched Dense_0 projection (blocks.Net.time_embedding):
    temb_all = Wcat silu(t2) + bcat, t2 = W1 silu(t1) + b1, t1 = W0 emb + b0 (the Fourier W is frozen in the reference)."""
    plan, e = b.plan, b.e
    tm = plan.temb
    B, H4, total, ed = b.B, tm["hidden"], tm["total"], tm["embed_dim"]
    wcat, _ = e.packed_dense()
    d_all = plan.d_temb_all
    s2, s1 = b.f32(B * H4), b.f32(B * H4)
    d_s2, d_t2, d_s1, d_t1 = b.f32(B * H4), b.f32(B * H4), b.f32(B * H4), b.f32(B * H4)
    gwcat, gbcat = b.f32(total * H4), b.f32(total)
    gw1, gb1, gw0, gb0 = b.f32(H4 * H4), b.f32(H4), b.f32(H4 * ed), b.f32(H4)
    p = lambda v: v.data_ptr()
    b.emit("b200_silu_f32", p(s2), p(tm["t2"]), B * H4)
    b.emit("b200_silu_f32", p(s1), p(tm["t1"]), B * H4)
    b.emit("b200_gemm_f32", p(gwcat), p(d_all), p(s2), total, H4, B, 1, 0, 1.0, 0.0)          # d_all^T s2
    b.emit("b200_colsum", p(d_all), p(gbcat), 0, 1, B, total, total, 1.0, 0)
    b.emit("b200_gemm_f32", p(d_s2), p(d_all), p(wcat), B, H4, total, 0, 0, 1.0, 0.0)          # d_all Wcat
    b.emit("b200_silu_bwd_f32", p(d_t2), p(d_s2), p(tm["t2"]), B * H4)
    b.emit("b200_gemm_f32", p(gw1), p(d_t2), p(s1), H4, H4, B, 1, 0, 1.0, 0.0)
    b.emit("b200_colsum", p(d_t2), p(gb1), 0, 1, B, H4, H4, 1.0, 0)
    b.emit("b200_gemm_f32", p(d_s1), p(d_t2), e.dev(tm["lin1"] + ".weight"), B, H4, H4, 0, 0, 1.0, 0.0)
    b.emit("b200_silu_bwd_f32", p(d_t1), p(d_s1), p(tm["t1"]), B * H4)
    b.emit("b200_gemm_f32", p(gw0), p(d_t1), p(tm["emb"]), H4, ed, B, 1, 0, 1.0, 0.0)
    b.emit("b200_colsum", p(d_t1), p(gb0), 0, 1, B, H4, H4, 1.0, 0)
    plan.pgrads.append(dict(kind="dense", gw=gwcat, gb=gbcat))
    plan.pgrads.append(dict(kind="linear", key=tm["lin1"], gw=gw1, gb=gb1, shape=(H4, H4)))
    plan.pgrads.append(dict(kind="linear", key=tm["lin0"], gw=gw0, gb=gb0, shape=(H4, ed)))


def emit_backward(b):
    """Appends the input-gradient launches to b.plan (call after the forward has been emitted, before finalize)."""
    plan, G = b.plan, _Grads(b)
    if getattr(b.e, "param_grads", False):
        plan.pgrads = []
        tm = plan.temb
        plan.d_temb_all = torch.zeros(b.B * tm["total"], device=b.e.device, dtype=torch.float32)
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
        elif kind == "dropout":
            if G.has(t["out"]):
                gx, acc = G.target(t["x"])
                assert not acc
                b.emit("b200_dropout", gx, G.get(t["out"]), b.dt, t["n"], t["p"], b.e.train_seed, t["off"],
                       b.e.train_step.data_ptr())
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
    if getattr(b.e, "param_grads", False):
        _temb_param_grads(b)
    plan.grad_scale = float(inp["scale"])
    plan.sample_mul_t = b.sample_mul
    _ = DT_BF16
