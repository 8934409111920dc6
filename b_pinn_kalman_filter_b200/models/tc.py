"""Tensor-core (tcgen05 / TMEM / TMA) launch selection for the score engine.  See csrc/conv_tc.cu."""
from .. import _lib
from .._lib import c_float, c_int, c_ll, c_vp

_lib.register_signature("b200_conv2d_tc_supported", [c_int] * 9)
_lib.register_signature("b200_conv2d_tc", [c_vp, c_int, c_vp, c_int, c_vp, c_vp, c_vp, c_int, c_vp, c_float, c_vp,
                                           c_int, c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp])

_lib.register_signature("b200_attention_tc_supported", [c_int, c_int])
_lib.register_signature("b200_attention_tc", [c_vp, c_vp, c_int, c_int, c_int, c_float, c_vp])

_lib.register_signature("b200_conv2d_tc_s2_supported", [c_int] * 5)
_lib.register_signature("b200_conv2d_tc_s2", [c_vp, c_int, c_vp, c_vp, c_vp, c_float, c_vp, c_vp] + [c_int] * 6 + [c_vp])
_lib.register_signature("b200_conv2d_tc_stat_slots", [c_int] * 3)
_lib.register_signature("b200_gn_apply_stats", [c_vp, c_int, c_vp, c_int, c_vp, c_int, c_vp, c_int, c_int, c_int, c_int,
                                                c_int, c_int, c_vp, c_vp, c_float, c_int, c_vp, c_vp])

_lib.register_signature("b200_gn_fold_stats", [c_vp, c_vp, c_int, c_int, c_int, c_vp])
_lib.register_signature("b200_gn_small_enable", [c_int])
_lib.register_signature("b200_gn_coef", [c_vp, c_int, c_int, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp, c_float,
                                        c_vp])
_lib.register_signature("b200_conv2d_tc_gn_supported", [c_int] * 7)
_lib.register_signature("b200_conv2d_tc_gn_preferred", [c_int] * 7)
_lib.register_signature("b200_conv2d_tc_gn", [c_vp, c_int, c_vp, c_int, c_vp, c_vp, c_vp, c_vp, c_int, c_vp, c_float,
                                              c_vp, c_vp, c_int, c_int, c_int, c_int, c_int])

_lib.register_signature("b200_conv2d_tc_sc", [c_vp, c_int, c_vp, c_int, c_vp, c_vp, c_int, c_vp, c_int, c_vp, c_vp, c_vp,
                                              c_int, c_float, c_vp, c_vp, c_int, c_int, c_int, c_int])

_lib.register_signature("b200_split_tf32", [c_vp, c_vp, c_vp, c_ll, c_vp])
_lib.register_signature("b200_conv2d_tc32_supported", [c_int] * 7)
_lib.register_signature("b200_conv2d_tc32", [c_vp, c_int, c_vp, c_int, c_vp, c_vp, c_vp, c_vp, c_int, c_vp, c_float, c_vp]
                        + [c_int] * 6 + [c_vp])

_lib.register_signature("b200_gemm_tc32_batched", [c_vp, c_ll, c_vp, c_vp, c_vp, c_int, c_int, c_int, c_int, c_float, c_int, c_vp])
_lib.register_signature("b200_softmax_rows_f32", [c_vp, c_ll, c_int, c_vp])
_lib.register_signature("b200_attn32_prep", [c_vp, c_vp, c_vp, c_vp, c_vp, c_int, c_int, c_int, c_vp])

_lib.register_signature("b200_gemm_tc_splitk", [c_vp, c_ll, c_vp, c_ll, c_vp, c_int, c_int, c_ll, c_int, c_vp])

_lib.register_signature("b200_wgrad_tc_supported", [c_int] * 8)
_lib.register_signature("b200_wgrad_tc", [c_vp, c_vp, c_int, c_vp, c_int, c_vp] + [c_int] * 6 + [c_vp])

ENABLED = True
FUSE_SHORTCUT = True  # Conv_2 (1x1 shortcut) folded into Conv_1's K loop (b200_conv2d_tc_sc)
FUSE_GN = True        # GroupNorm + SiLU inside the consumer conv's operand pipeline (b200_conv2d_tc_gn)


def conv_supported(b, x, x2, out, ksize, stride, pad):
    if not ENABLED or pad != (ksize // 2, ksize // 2) or (out.H, out.W) != (x.H, x.W):
        return False
    c2 = x2.C if x2 is not None else 0
    return bool(_lib.lib().b200_conv2d_tc_supported(x.C, c2, b.B, x.H, x.W, out.C, ksize, stride, 0))


def emit_conv(b, x, x2, w, bias, temb_ptr, res, out_scale, out, ksize, flops):
    p2, c2 = (x2, x2.C) if x2 is not None else (None, 0)
    b.emit("b200_conv2d_tc", x, x.C, p2, c2, w.data_ptr(), bias.data_ptr(), temb_ptr, b.temb_stride, res, out_scale,
           out, 0, None, None, b.B, x.H, x.W, out.C, ksize, flops=flops,
           tag="%dx%d %dx%d %d->%d" % (ksize, ksize, x.H, x.W, x.C + c2, out.C))
    out.producer = (len(b.plan.ops) - 1, 13)      # (launch index, position of the `stats` argument)


def conv32_supported(b, x, x2, out, ksize, stride, pad):
    """Strict-mode (fp32) convolution on the tensor cores: b200_conv2d_tc32."""
    if not ENABLED or stride != 1 or pad != (ksize // 2, ksize // 2) or (out.H, out.W) != (x.H, x.W):
        return False
    c2 = x2.C if x2 is not None else 0
    return bool(_lib.lib().b200_conv2d_tc32_supported(x.C, c2, b.B, x.H, x.W, out.C, ksize))


def emit_conv32(b, x, x2, w_hi, w_lo, bias, temb_ptr, res, out_scale, out, ksize, terms, flops):
    p2, c2 = (x2, x2.C) if x2 is not None else (None, 0)
    b.emit("b200_conv2d_tc32", x, x.C, p2, c2, w_hi.data_ptr(), w_lo.data_ptr(), bias.data_ptr(), temb_ptr, b.temb_stride,
           res, out_scale, out, b.B, x.H, x.W, out.C, ksize, terms, flops=flops,
           tag="%dx%d %dx%d %d->%d tf32x%d" % (ksize, ksize, x.H, x.W, x.C + c2, out.C, terms))


def attention32_supported(L, C):
    """Strict-mode attention as two batched tf32-split GEMMs around a row softmax."""
    return ENABLED and L % 128 == 0 and C % 128 == 0


def emit_attention32(b, qkv, out, L, C, terms):
    """softmax(q k^T / sqrt(C)) v in fp32 on the tensor cores (models/layerspp.py:82-86): operand preparation (k split, v
    transposed and split), q k^T -> scores [B, L, L], row softmax, p v."""
    kh, kl = b.new(qkv.H, qkv.W, C), b.new(qkv.H, qkv.W, C)
    vh, vl = b.new(qkv.H, qkv.W, C), b.new(qkv.H, qkv.W, C)
    sc = b.new(qkv.H, qkv.W, L)
    b.emit("b200_attn32_prep", qkv, kh, kl, vh, vl, b.B, L, C)
    b.emit("b200_gemm_tc32_batched", qkv, 3 * C, kh, kl, sc, b.B, L, L, C, float(int(C) ** (-0.5)), terms,
           flops=2 * b.B * L * L * C, tag="q k^T %dx%dx%d tf32x%d" % (L, L, C, terms))
    b.emit("b200_softmax_rows_f32", sc, b.B * L, L)
    b.emit("b200_gemm_tc32_batched", sc, L, vh, vl, out, b.B, L, C, L, 1.0, terms,
           flops=2 * b.B * L * L * C, tag="p v %dx%dx%d tf32x%d" % (L, C, L, terms))


def conv_gn_supported(b, x1, x2, out, ksize, stride, pad):
    """Can `conv(silu(groupnorm(cat(x1, x2))))` run with the normalisation inside the conv (raw sources x1 / x2)?"""
    if not (ENABLED and FUSE_GN) or ksize != 3 or stride != 1 or pad != (1, 1) or (out.H, out.W) != (x1.H, x1.W):
        return False
    c2 = x2.C if x2 is not None else 0
    # "all": every shape the fused kernel can run; otherwise only where it measured faster than conv + GroupNorm launch
    fn = _lib.lib().b200_conv2d_tc_gn_supported if b.e.fuse_gn_conv == "all" else _lib.lib().b200_conv2d_tc_gn_preferred
    return bool(fn(x1.C, c2, b.B, x1.H, x1.W, out.C, ksize))


def emit_conv_gn(b, gn, w, bias, temb_ptr, res, out_scale, out, ksize, flops):
    """`gn`: a deferred GroupNorm (engine.LazyGN).  Emits the coefficient fold + the GroupNorm-fused convolution."""
    x1, x2 = gn.src1, gn.src2
    s1, t1 = request_stats(b, x1)
    s2, t2 = request_stats(b, x2) if x2 is not None else (None, 0)
    p2, c2 = (x2, x2.C) if x2 is not None else (None, 0)
    coef = b.f32(b.B * (x1.C + c2) * 2)
    b.emit("b200_gn_coef", s1, t1, x1.C, s2, t2, c2, b.B, x1.H * x1.W, gn.groups, b.e.dev(gn.key + ".weight"),
           b.e.dev(gn.key + ".bias"), 1e-6, coef.data_ptr())
    b.emit("b200_conv2d_tc_gn", x1, x1.C, p2, c2, coef.data_ptr(), w.data_ptr(), bias.data_ptr(), temb_ptr,
           b.temb_stride, res, out_scale, out, None, b.B, x1.H, x1.W, out.C, ksize, flops=flops,
           tag="%dx%d %dx%d %d->%d gn" % (ksize, ksize, x1.H, x1.W, x1.C + c2, out.C))
    out.producer = (len(b.plan.ops) - 1, 12)      # position of the `stats` argument


def emit_conv_sc(b, x, gn, xs1, xs2, w, bias, temb_ptr, out_scale, out, flops):
    """Conv_1 + folded 1x1 shortcut.  `x`: the (materialised) conv input, or None when `gn` (a LazyGN) is fused in."""
    coef = None
    if gn is not None:
        x1, x2 = gn.src1, gn.src2
        s1, t1 = request_stats(b, x1)
        s2, t2 = request_stats(b, x2) if x2 is not None else (None, 0)
        coef_t = b.f32(b.B * gn.C * 2)
        b.emit("b200_gn_coef", s1, t1, x1.C, s2, t2, x2.C if x2 is not None else 0, b.B, x1.H * x1.W, gn.groups,
               b.e.dev(gn.key + ".weight"), b.e.dev(gn.key + ".bias"), 1e-6, coef_t.data_ptr())
        coef = coef_t.data_ptr()
    else:
        x1, x2 = x, None
    p2, c2 = (x2, x2.C) if x2 is not None else (None, 0)
    q2, d2 = (xs2, xs2.C) if xs2 is not None else (None, 0)
    b.emit("b200_conv2d_tc_sc", x1, x1.C, p2, c2, coef, xs1, xs1.C, q2, d2, w.data_ptr(), bias.data_ptr(), temb_ptr,
           b.temb_stride, out_scale, out, None, b.B, x1.H, x1.W, out.C, flops=flops,
           tag="3x3 %dx%d %d->%d +sc %d%s" % (x1.H, x1.W, x1.C + c2, out.C, xs1.C + d2, " gn" if gn is not None else ""))
    out.producer = (len(b.plan.ops) - 1, 15)      # position of the `stats` argument


def head_supported(b, x, cout, ksize):
    """Network head: conv -> NCHW fp32 * sample_mul (Cout <= 16 on one N = 16 tile)."""
    return ENABLED and bool(_lib.lib().b200_conv2d_tc_supported(x.C, 0, b.B, x.H, x.W, cout, ksize, 1, 1))


def emit_head(b, x, w, bias, out_scale, out_ptr, sample_mul_ptr, cout, ksize, flops):
    b.emit("b200_conv2d_tc", x, x.C, None, 0, w.data_ptr(), bias.data_ptr(), None, 0, None, out_scale, out_ptr, 1,
           sample_mul_ptr, None, b.B, x.H, x.W, cout, ksize, flops=flops)


def narrow_supported(b, x, out, ksize):
    """3-channel image pyramid stored 8 wide (out_mode 2 of b200_conv2d_tc)."""
    return ENABLED and out.C == 8 and (out.H, out.W) == (x.H, x.W) and bool(
        _lib.lib().b200_conv2d_tc_supported(x.C, 0, b.B, x.H, x.W, 8, ksize, 1, 2))


def emit_narrow(b, x, w, bias, res, out_scale, out, ksize, flops):
    b.emit("b200_conv2d_tc", x, x.C, None, 0, w.data_ptr(), bias.data_ptr(), None, 0, res, out_scale, out, 2, None, None,
           b.B, x.H, x.W, 8, ksize, flops=flops, tag="%dx%d %dx%d %d->8 pyramid" % (ksize, ksize, x.H, x.W, x.C))


def attention_supported(L, C):
    return ENABLED and bool(_lib.lib().b200_attention_tc_supported(L, C))


def emit_attention(b, qkv, out, L, C):
    b.emit("b200_attention_tc", qkv, out, b.B, L, C, float(int(C) ** (-0.5)), flops=4 * b.B * L * L * C)


def s2_supported(b, x, out):
    return ENABLED and bool(_lib.lib().b200_conv2d_tc_s2_supported(x.C, b.B, out.H, out.W, out.C))


def emit_s2(b, x, w, bias, res, out_scale, out, flops):
    b.emit("b200_conv2d_tc_s2", x, x.C, w.data_ptr(), bias.data_ptr(), res, out_scale, out, None, b.B, x.H, x.W, out.H,
           out.W, out.C, flops=flops, tag="3x3/2 %dx%d %d->%d" % (x.H, x.W, x.C, out.C))
    out.producer = (len(b.plan.ops) - 1, 7)


FOLD_SLOTS = 64       # images whose producer leaves at least this many statistics slots get them folded once (64x64 and up)


def request_stats(b, act):
    """Makes the tensor-core conv that produced `act` also write the GroupNorm statistics; returns (ptr, slots).
    Large images: the slots are summed ONCE by b200_gn_fold_stats (emitted here, i.e. in front of the first consumer)
    and every consumer sees one slot."""
    if act.stats is None:
        T = _lib.lib().b200_conv2d_tc_stat_slots(b.B, act.H, act.W)
        buf = b.f32(b.B * T * act.C * 2)
        idx, pos = act.producer
        fn, args, name = b.plan.ops[idx]
        b.plan.ops[idx] = (fn, args[:pos] + (buf.data_ptr(),) + args[pos + 1:], name)
        if T >= FOLD_SLOTS:
            folded = b.f32(b.B * act.C * 2)
            b.emit("b200_gn_fold_stats", buf.data_ptr(), folded.data_ptr(), b.B, T, act.C,
                   tag="%dx%d %d slots %d" % (act.H, act.W, act.C, T))
            act.stats = (folded.data_ptr(), 1)
        else:
            act.stats = (buf.data_ptr(), T)
    return act.stats
