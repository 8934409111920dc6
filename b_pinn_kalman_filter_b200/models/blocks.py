"""Network blocks expressed as fused launches on the score engine (used by ncsnpp.py and ddpm.py).

Each method cites the reference module it executes.  `idx` is the position in the reference's `all_modules` list,
which fixes the parameter names.
"""
import math

import torch

INV_SQRT2 = 1.0 / math.sqrt(2.0)


def gn_groups(c):
    """num_groups=min(C // 4, 32)  (models/layerspp.py:67,173,219)."""
    return min(c // 4, 32)


class Net:
    def __init__(self, b, config):
        self.b = b
        self.config = config
        self._idx = 0
        self.dense_off = 0
        self.dense_count = 0

    def next_module(self):
        i = self._idx
        self._idx += 1
        return i

    # ---- time embedding ------------------------------------------------------------------------------------
    def time_embedding(self, kind, nf, fourier_scale, scale_by_sigma):
        """Fourier (layerspp.py:32-41) or sinusoidal (layers.py:500-514) embedding, the 2-layer MLP
        (ncsnpp.py:86-91, 251-255) and -- in ONE batched projection -- act(temb) through every block's Dense_0
        (layerspp.py:227,263): [B, 4nf] x [4nf, sum(out_ch)]."""
        b, e = self.b, self.b.e
        if kind == "fourier":
            self.w_fourier = b.param("all_modules.%d.W" % self.next_module(), (nf,), ("randn", fourier_scale))
            embed_dim = 2 * nf
        else:
            self.w_fourier = None
            embed_dim = nf
        self.lin0 = "all_modules.%d" % self.next_module()
        b.param(self.lin0 + ".weight", (4 * nf, embed_dim), ("default", 1.0))
        b.param(self.lin0 + ".bias", (4 * nf,), "zeros")
        self.lin1 = "all_modules.%d" % self.next_module()
        b.param(self.lin1 + ".weight", (4 * nf, 4 * nf), ("default", 1.0))
        b.param(self.lin1 + ".bias", (4 * nf,), "zeros")
        self.temb_dim = 4 * nf
        if b.dry:
            return
        B = b.B
        plan = b.plan
        plan.labels = torch.zeros(B, device=e.device, dtype=torch.float32)
        plan.step_ptr = None
        emb = b.f32(B * embed_dim)
        # uniform-label plans (the fused PC loop: every sample of the batch sits at the same time step) run the
        # embedding MLP and the batched Dense_0 projection on ONE row and broadcast it (temb_stride = 0)
        Bt = 1 if getattr(e, "uniform_mode", False) else B
        t1 = b.f32(Bt * 4 * nf)
        t2 = b.f32(Bt * 4 * nf)
        total = e.dense_total
        b.temb_all = b.f32(Bt * total)
        b.temb_stride = total if Bt == B else 0
        b.sample_mul = b.f32(B)
        sig = e.params.get("sigmas")
        b.emit("b200_temb_embed", plan.labels.data_ptr(), None, 0 if kind == "fourier" else 1,
               e.dev(self.w_fourier) if self.w_fourier else None, nf, sig.data_ptr() if sig is not None else None,
               sig.numel() if sig is not None else 0, 1 if scale_by_sigma else 0, emb.data_ptr(),
               b.sample_mul.data_ptr(), B)
        plan.temb_op_index = len(plan.ops) - 1
        b.emit("b200_linear_f32", emb.data_ptr(), e.dev(self.lin0 + ".weight"), e.dev(self.lin0 + ".bias"),
               t1.data_ptr(), Bt, embed_dim, 4 * nf, 0, 0, flops=2 * Bt * embed_dim * 4 * nf)
        b.emit("b200_linear_f32", t1.data_ptr(), e.dev(self.lin1 + ".weight"), e.dev(self.lin1 + ".bias"),
               t2.data_ptr(), Bt, 4 * nf, 4 * nf, 1, 0, flops=2 * Bt * 16 * nf * nf)
        plan.temb = dict(emb=emb, t1=t1, t2=t2, all=b.temb_all, embed_dim=embed_dim, hidden=4 * nf, total=total, rows=Bt,
                         lin0=self.lin0, lin1=self.lin1)
        wcat, bcat = e.packed_dense()
        b.emit("b200_linear_f32", t2.data_ptr(), wcat.data_ptr(), bcat.data_ptr(), b.temb_all.data_ptr(), Bt, 4 * nf,
               total, 1, 0, flops=2 * Bt * 4 * nf * total)

    def dense(self, prefix, out_ch):
        """Registers a block's Dense_0 and returns its column offset in the batched projection."""
        b = self.b
        b.param(prefix + ".Dense_0.weight", (out_ch, self.temb_dim), ("default", 1.0))
        b.param(prefix + ".Dense_0.bias", (out_ch,), "zeros")
        b.dense.append((prefix + ".Dense_0", out_ch))
        off = self.dense_off
        self.dense_off += out_ch
        return off

    # ---- input / output ------------------------------------------------------------------------------------
    def input_image(self, channels, H, W, rescale):
        """NCHW fp32 -> channels-last (8-wide, zero padded), with `2x - 1` folded in (ncsnpp.py:259-261)."""
        b = self.b
        x = b.new(H, W, 8, channels)
        self.x_stem = x
        if b.dry:
            return x
        plan = b.plan
        plan.x_in = torch.zeros(b.B, channels, H, W, device=b.e.device, dtype=torch.float32)
        plan.out = torch.zeros(b.B, channels, H, W, device=b.e.device, dtype=torch.float32)
        if b.e.use_tc:
            # tensor-core path: the image lives in a 64-channel (one 128-byte TMA row) zero-padded tensor, which
            # feeds the stem convolution and the input pyramids alike; the 8-wide form is never read
            x = b.new(H, W, 64, channels)
        b.emit("b200_nchw_to_nhwc", plan.x_in.data_ptr(), x, b.dt, b.B, channels, H, W, x.C,
               2.0 if rescale else 1.0, -1.0 if rescale else 0.0)
        self.x_stem = x
        b.rec("input", x=x, x_stem=self.x_stem, channels=channels, scale=2.0 if rescale else 1.0)
        return x

    def output_conv3x3(self, a, idx, channels, init_scale):
        """Final conv3x3 -> NCHW fp32 with `h / used_sigmas` (ncsnpp.py:371-379) in the epilogue."""
        self.b.conv(a, "all_modules.%d" % idx, channels, 3, init_scale=init_scale, nchw_out=self.b.plan.out)

    def output_from_nhwc(self, pyramid, channels):
        b = self.b
        if b.dry:
            return
        b.emit("b200_nhwc_to_nchw", pyramid, b.plan.out.data_ptr(), b.dt, b.B, channels, pyramid.H, pyramid.W,
               pyramid.C, b.sample_mul.data_ptr())
        b.rec("out_nhwc", x=pyramid, channels=channels)

    def finish(self):
        pass

    # ---- convolutions --------------------------------------------------------------------------------------
    def conv3x3(self, x, idx, cout, **kw):
        return self.b.conv(x, "all_modules.%d" % idx if isinstance(idx, int) else idx, cout, 3, **kw)

    def conv1x1(self, x, key, cout, **kw):
        return self.b.conv(x, key, cout, 1, **kw)

    def gn_act(self, x, key, c, x2=None, act=True, groups=None):
        return self.b.gn(x, key, groups or gn_groups(c), act, x2=x2)

    # ---- FIR resampling (models/up_or_down_sampling.py:195-257) ---------------------------------------------
    def fir_up(self, x):
        return self.b.upfirdn(x, 2, 1, (2, 1))      # upsample_2d: pad=((p+1)//2+factor-1, p//2), p = 2

    def fir_down(self, x):
        return self.b.upfirdn(x, 1, 2, (1, 1))      # downsample_2d: pad=((p+1)//2, p//2), p = 2

    def fir_conv_down(self, x, key, cout, res=None, out_scale=1.0):
        """up_or_down_sampling.Conv2d(down=True) (:23-56) = conv_downsample_2d (:144-178): FIR with pad (2,2)
        then 3x3 stride-2 VALID conv, + bias."""
        f = self.b.upfirdn(x, 1, 1, (2, 2))         # p = (4-2)+(3-1) = 4 -> pad (2, 2): H+1 x W+1
        return self.b.conv(f, key, cout, 3, stride=2, pad=(0, 0), out_hw=(x.H // 2, x.W // 2), res=res,
                           out_scale=out_scale)

    # ---- residual blocks -----------------------------------------------------------------------------------
    def resblock_biggan(self, idx, x, x2, out_ch, up, down, fir, skip_rescale, init_scale):
        """ResnetBlockBigGANpp.forward (models/layerspp.py:242-274)."""
        b = self.b
        p = "all_modules.%d" % idx
        in_ch = x.C + (x2.C if x2 is not None else 0)
        out_ch = out_ch or in_ch
        h = self.gn_act(x, p + ".GroupNorm_0", in_ch, x2=x2)
        xs, xs2 = x, x2
        if up or down:
            if x2 is not None:
                raise NotImplementedError("up/down blocks never take concatenated inputs in the reference")
            if not fir:
                raise NotImplementedError("naive (non-FIR) resampling inside BigGAN blocks: no shipped config")
            h = self.fir_up(h) if up else self.fir_down(h)
            xs = self.fir_up(x) if up else self.fir_down(x)
        off = self.dense(p, out_ch)
        h1 = b.conv(h, p + ".Conv_0", out_ch, 3, temb_off=off)
        h2 = b.dropout(self.gn_act(h1, p + ".GroupNorm_1", out_ch))
        scale = INV_SQRT2 if skip_rescale else 1.0
        if in_ch != out_ch or up or down:
            return b.conv_sc(h2, p + ".Conv_1", p + ".Conv_2", xs, xs2, out_ch, scale, init_scale)
        if x2 is not None:
            raise NotImplementedError("identity skip over a concatenated input")
        return b.conv(h2, p + ".Conv_1", out_ch, 3, res=x, out_scale=scale, init_scale=init_scale)

    def resblock_ddpmpp(self, idx, x, x2, out_ch, skip_rescale, init_scale, groups=None, conv_shortcut=False):
        """ResnetBlockDDPMpp.forward (layerspp.py:193-209) / ResnetBlockDDPM.forward (layers.py:635-655, which is
        the same block with 32 groups, init_scale 0 and no rescale)."""
        b = self.b
        p = "all_modules.%d" % idx
        in_ch = x.C + (x2.C if x2 is not None else 0)
        out_ch = out_ch or in_ch
        h = self.gn_act(x, p + ".GroupNorm_0", in_ch, x2=x2, groups=groups)
        off = self.dense(p, out_ch)
        h1 = b.conv(h, p + ".Conv_0", out_ch, 3, temb_off=off)
        h2 = b.dropout(self.gn_act(h1, p + ".GroupNorm_1", out_ch, groups=groups))
        scale = INV_SQRT2 if skip_rescale else 1.0
        if in_ch != out_ch:
            if conv_shortcut:
                skip = b.conv(x, p + ".Conv_2", out_ch, 3, x2=x2)
                return b.conv(h2, p + ".Conv_1", out_ch, 3, res=skip, out_scale=scale, init_scale=init_scale)
            return b.conv_sc(h2, p + ".Conv_1", p + ".NIN_0", x, x2, out_ch, scale, init_scale, nin2=True, nin2_init=0.1)
        if x2 is not None:
            raise NotImplementedError("identity skip over a concatenated input")
        return b.conv(h2, p + ".Conv_1", out_ch, 3, res=x, out_scale=scale, init_scale=init_scale)

    # ---- attention -----------------------------------------------------------------------------------------
    def attn_block(self, idx, x, skip_rescale, init_scale, groups=None):
        """AttnBlockpp.forward (layerspp.py:75-91) / AttnBlock.forward (layers.py:560-573): GN -> q,k,v NIN ->
        softmax(q k^T C^-1/2) v -> NIN_3 -> (x + h)[/sqrt2].  q,k,v are ONE 1x1 conv with stacked weights."""
        b, e = self.b, self.b.e
        p = "all_modules.%d" % idx
        C = x.C
        hn = self.gn_act(x, p + ".GroupNorm_0", C, act=False, groups=groups)
        for i in range(3):
            b.param(p + ".NIN_%d.W" % i, (C, C), ("default", 0.1))
            b.param(p + ".NIN_%d.b" % i, (C,), "zeros")
        qkv = b.new(x.H, x.W, 3 * C)
        if not b.dry:
            w, bias = e.packed_qkv(p, C)
            e.emit_conv(b, hn, None, w, bias, None, None, 1.0, qkv, 1, 1, (0, 0), p + ".qkv")
            b.rec("conv", x=hn, x2=None, out=qkv, w=w, ksize=1, stride=1, pad=(0, 0), res=None, out_scale=1.0,
                  head=False, cout=3 * C, key=p, nin="qkv", temb_off=None, out_hw=(x.H, x.W))
        a = b.attention_core(qkv, C)
        scale = INV_SQRT2 if skip_rescale else 1.0
        return b.conv(a, p + ".NIN_3", C, 1, nin=True, init_scale=init_scale, res=x, out_scale=scale)

    # ---- DDPM-style resampling (models/layers.py:576-608, layerspp.py:94-163 with fir=False) -----------------
    def downsample(self, idx, x, fir, with_conv):
        p = "all_modules.%d" % idx
        if fir:
            if with_conv:
                return self.fir_conv_down(x, p + ".Conv2d_0", x.C)
            return self.fir_down(x)
        if not with_conv:
            raise NotImplementedError("avg_pool downsample: no shipped config")
        # F.pad(x, (0,1,0,1)) + conv3x3 stride 2 pad 0: zero fill to the right/bottom comes from bounds checks
        return self.b.conv(x, p + ".Conv_0", x.C, 3, stride=2, pad=(0, 0), out_hw=(x.H // 2, x.W // 2))

    def upsample(self, idx, x, fir, with_conv):
        p = "all_modules.%d" % idx
        if fir:
            if with_conv:
                raise NotImplementedError("Upsample(fir=True, with_conv=True) is broken in the reference (SURVEY 0.4)")
            return self.fir_up(x)
        h = self.b.nearest_up2(x)
        if with_conv:
            h = self.b.conv(h, p + ".Conv_0", x.C, 3)
        return h
