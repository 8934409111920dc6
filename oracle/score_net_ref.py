"""CPU oracle for the score networks: a plain-PyTorch (fp32 / fp64) functional restatement of the reference's
NCSNpp.forward (models/ncsnpp.py:232-381) and DDPM.forward (models/ddpm.py:110-181) driven by a reference-layout
state dict (`all_modules.<i>.<name>`).  TEST INFRASTRUCTURE ONLY.

Pinned against the real reference modules on shared weights in tests/test_oracle_pins.py (authoring container) and
by the golden vectors tests/golden/score_*.npz that oracle/make_golden.py generated from the reference itself.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

from . import upfirdn2d_ref


def _fir(x, k, up=1, down=1, pad=(0, 0)):
    """upfirdn2d on torch tensors through the numpy oracle (float64 accumulate), keeps dtype/device of x."""
    y = upfirdn2d_ref.upfirdn2d(x.detach().cpu().double().numpy(), np.asarray(k, dtype=np.float64), up, down, pad)
    return torch.from_numpy(np.asarray(y)).to(dtype=x.dtype)


def _fir_torch(x, k, up=1, down=1, pad=(0, 0)):
    """The same upfirdn2d written with differentiable torch ops (zero insertion, pad / crop, correlation with the
    flipped kernel, decimation: op/upfirdn2d.py:159-200) -- used only when the oracle must be differentiated."""
    B, C, H, W = x.shape
    k = torch.as_tensor(np.asarray(k), dtype=x.dtype, device=x.device)
    kh, kw = k.shape
    y = x.reshape(B * C, 1, H, 1, W, 1)
    y = F.pad(y, [0, up - 1, 0, 0, 0, up - 1])
    y = y.reshape(B * C, 1, H * up, W * up)
    p0, p1 = pad
    y = F.pad(y, [max(p0, 0), max(p1, 0), max(p0, 0), max(p1, 0)])
    y = y[:, :, max(-p0, 0):y.shape[2] - max(-p1, 0), max(-p0, 0):y.shape[3] - max(-p1, 0)]
    y = F.conv2d(y, torch.flip(k, [0, 1]).reshape(1, 1, kh, kw))
    y = y[:, :, ::down, ::down]
    return y.reshape(B, C, y.shape[2], y.shape[3])


class RefScoreNet:
    def __init__(self, config, state_dict, dtype=torch.float32, device="cpu"):
        """device: 'cpu' (the oracle proper) or a CUDA device -- the same torch ops then run on cuDNN / cuBLAS, which IS
        the reference's GPU path for these layers (tests: large-batch checks; bench.py: the "kernel to beat" leg)."""
        self.cfg = config
        self.m = config.model
        self.dtype = dtype
        self.device = torch.device(device)
        self.sd = {}
        for k, v in state_dict.items():
            k = k[len("module."):] if k.startswith("module.") else k
            self.sd[k] = v.detach().to(device=self.device, dtype=dtype)
        self.nf = self.m.nf
        self.fir_k = upfirdn2d_ref.setup_kernel(self.m.fir_kernel)

    # -- parameter access --
    def p(self, idx, name):
        return self.sd["all_modules.%d.%s" % (idx, name)]

    def act(self, x):
        return F.silu(x)   # every shipped config uses nonlinearity='swish' (layers.py:38-39)

    def gn(self, idx, name, x, groups):
        return F.group_norm(x, groups, self.p(idx, name + ".weight"), self.p(idx, name + ".bias"), eps=1e-6)

    def conv(self, idx, name, x, stride=1, padding=1):
        return F.conv2d(x, self.p(idx, name + ".weight"), self.p(idx, name + ".bias"), stride=stride, padding=padding)

    def nin(self, idx, name, x):
        W, b = self.p(idx, name + ".W"), self.p(idx, name + ".b")
        y = torch.tensordot(x.permute(0, 2, 3, 1), W, dims=1) + b      # layers.py:543-546
        return y.permute(0, 3, 1, 2)

    def dense(self, idx, temb):
        return F.linear(self.act(temb), self.p(idx, "Dense_0.weight"), self.p(idx, "Dense_0.bias"))[:, :, None, None]

    # -- FIR resampling (up_or_down_sampling.py:195-257, 144-178) --
    def fir(self, x, k, **kw):
        torch_form = getattr(self, "differentiable", False) or self.device.type != "cpu"
        return _fir_torch(x, k, **kw) if torch_form else _fir(x, k, **kw)

    def up2(self, x):
        return self.fir(x, self.fir_k * 4, up=2, pad=(2, 1))

    def down2(self, x):
        return self.fir(x, self.fir_k, down=2, pad=(1, 1))

    def conv_down(self, idx, x):
        w, b = self.p(idx, "Conv2d_0.weight"), self.p(idx, "Conv2d_0.bias")
        y = self.fir(x, self.fir_k, pad=(2, 2))
        return F.conv2d(y, w, stride=2, padding=0) + b.reshape(1, -1, 1, 1)

    # -- blocks --
    def res_biggan(self, idx, x, temb, up=False, down=False):
        in_ch = x.shape[1]
        out_ch = self.p(idx, "Conv_0.weight").shape[0]
        h = self.act(self.gn(idx, "GroupNorm_0", x, min(in_ch // 4, 32)))
        if up:
            h, x = self.up2(h), self.up2(x)
        elif down:
            h, x = self.down2(h), self.down2(x)
        h = self.conv(idx, "Conv_0", h) + self.dense(idx, temb)
        h = self.act(self.gn(idx, "GroupNorm_1", h, min(out_ch // 4, 32)))
        h = self.conv(idx, "Conv_1", h)                      # dropout is identity in eval mode
        if in_ch != out_ch or up or down:
            x = self.conv(idx, "Conv_2", x, padding=0)
        return (x + h) / np.sqrt(2.) if self.m.skip_rescale else x + h

    def res_ddpm(self, idx, x, temb, groups=None, rescale=False):
        in_ch = x.shape[1]
        out_ch = self.p(idx, "Conv_0.weight").shape[0]
        g0 = groups or min(in_ch // 4, 32)
        g1 = groups or min(out_ch // 4, 32)
        h = self.act(self.gn(idx, "GroupNorm_0", x, g0))
        h = self.conv(idx, "Conv_0", h) + self.dense(idx, temb)
        h = self.act(self.gn(idx, "GroupNorm_1", h, g1))
        h = self.conv(idx, "Conv_1", h)
        if in_ch != out_ch:
            x = self.nin(idx, "NIN_0", x)
        return (x + h) / np.sqrt(2.) if rescale else x + h

    def attn(self, idx, x, groups=None, rescale=False):
        B, C, H, W = x.shape
        h = self.gn(idx, "GroupNorm_0", x, groups or min(C // 4, 32))
        q, k, v = (self.nin(idx, "NIN_%d" % i, h) for i in range(3))
        w = torch.einsum('bchw,bcij->bhwij', q, k) * (int(C) ** (-0.5))
        w = F.softmax(w.reshape(B, H, W, H * W), dim=-1).reshape(B, H, W, H, W)
        h = self.nin(idx, "NIN_3", torch.einsum('bhwij,bcij->bchw', w, v))
        return (x + h) / np.sqrt(2.) if rescale else x + h

    def timestep_embedding(self, t, dim):
        half = dim // 2
        e = math.log(10000) / (half - 1)
        e = torch.exp(torch.arange(half, dtype=torch.float32, device=t.device) * -e).to(self.dtype)
        e = t.to(self.dtype)[:, None] * e[None, :]
        return torch.cat([torch.sin(e), torch.cos(e)], dim=1)

    # -- networks --
    def __call__(self, x, labels):
        x = x.detach().to(device=self.device, dtype=self.dtype)
        labels = labels.detach().to(self.device)
        with torch.no_grad():
            if self.m.name == "ddpm":
                return self.ddpm(x, labels)
            return self.ncsnpp(x, labels)

    def input_grad(self, x, labels, v):
        """(out, J^T v): vector-Jacobian product w.r.t. the image by autograd through this restatement -- the quantity
        DPS propagates through the score model (inverse/conditional_sampling.py:125-131).  FIR resampling switches
        to its differentiable torch form (_fir_torch) for the duration."""
        x = x.detach().cpu().to(self.dtype).requires_grad_(True)
        self.differentiable = True
        try:
            lab = labels.detach().cpu()
            out = self.ddpm(x, lab) if self.m.name == "ddpm" else self.ncsnpp(x, lab)
        finally:
            self.differentiable = False
        gx, = torch.autograd.grad(out, x, v.detach().cpu().to(self.dtype))
        return out.detach(), gx

    def ncsnpp(self, x, time_cond):
        m, nf = self.m, self.nf
        levels = len(m.ch_mult)
        i = 0
        if m.embedding_type.lower() == "fourier":
            used_sigmas = time_cond.to(self.dtype)
            proj = torch.log(used_sigmas)[:, None] * self.sd["all_modules.0.W"][None, :] * 2 * np.pi
            temb = torch.cat([torch.sin(proj), torch.cos(proj)], dim=-1)
            i = 1
        else:
            used_sigmas = self.sd["sigmas"][time_cond.long()]
            temb = self.timestep_embedding(time_cond, nf)
        temb = F.linear(temb, self.p(i, "weight"), self.p(i, "bias"))
        temb = F.linear(self.act(temb), self.p(i + 1, "weight"), self.p(i + 1, "bias"))
        i += 2
        if not self.cfg.data.centered:
            x = 2 * x - 1.
        pin_mode, pout_mode = m.progressive_input.lower(), m.progressive.lower()
        biggan = m.resblock_type.lower() == "biggan"
        res = (lambda idx, h, **kw: self.res_biggan(idx, h, temb, **kw)) if biggan else \
              (lambda idx, h, **kw: self.res_ddpm(idx, h, temb, rescale=m.skip_rescale))
        pyr_in = x if pin_mode != "none" else None
        hs = [F.conv2d(x, self.p(i, "weight"), self.p(i, "bias"), padding=1)]
        i += 1
        for lvl in range(levels):
            for _ in range(m.num_res_blocks):
                h = res(i, hs[-1]); i += 1
                if h.shape[-1] in m.attn_resolutions:
                    h = self.attn(i, h, rescale=m.skip_rescale); i += 1
                hs.append(h)
            if lvl != levels - 1:
                if biggan:
                    h = res(i, hs[-1], down=True); i += 1
                else:
                    h = self.conv_down(i, hs[-1]) if m.fir else F.conv2d(F.pad(hs[-1], (0, 1, 0, 1)), self.p(i, "Conv_0.weight"), self.p(i, "Conv_0.bias"), stride=2)
                    i += 1
                if pin_mode == "input_skip":
                    pyr_in = self.down2(pyr_in)
                    h = self.conv(i, "Conv_0", pyr_in, padding=0) + h; i += 1
                elif pin_mode == "residual":
                    pyr_in = self.conv_down(i, pyr_in); i += 1
                    pyr_in = (pyr_in + h) / np.sqrt(2.) if m.skip_rescale else pyr_in + h
                    h = pyr_in
                hs.append(h)
        h = hs[-1]
        h = res(i, h); i += 1
        h = self.attn(i, h, rescale=m.skip_rescale); i += 1
        h = res(i, h); i += 1
        pyramid = None
        for lvl in reversed(range(levels)):
            for _ in range(m.num_res_blocks + 1):
                h = res(i, torch.cat([h, hs.pop()], dim=1)); i += 1
            if h.shape[-1] in m.attn_resolutions:
                h = self.attn(i, h, rescale=m.skip_rescale); i += 1
            if pout_mode == "output_skip":
                g = min(h.shape[1] // 4, 32)
                ph = self.act(F.group_norm(h, g, self.p(i, "weight"), self.p(i, "bias"), eps=1e-6)); i += 1
                ph = F.conv2d(ph, self.p(i, "weight"), self.p(i, "bias"), padding=1); i += 1
                pyramid = ph if pyramid is None else self.up2(pyramid) + ph
            elif pout_mode != "none":
                raise NotImplementedError(pout_mode)
            if lvl != 0:
                if biggan:
                    h = res(i, h, up=True); i += 1
                else:
                    raise NotImplementedError
        assert not hs
        if pout_mode == "output_skip":
            h = pyramid
        else:
            g = min(h.shape[1] // 4, 32)
            h = self.act(F.group_norm(h, g, self.p(i, "weight"), self.p(i, "bias"), eps=1e-6)); i += 1
            h = F.conv2d(h, self.p(i, "weight"), self.p(i, "bias"), padding=1); i += 1
        if m.scale_by_sigma:
            h = h / used_sigmas.reshape(-1, 1, 1, 1)
        return h

    def ddpm(self, x, labels):
        m, nf = self.m, self.nf
        levels = len(m.ch_mult)
        temb = self.timestep_embedding(labels, nf)
        temb = F.linear(temb, self.p(0, "weight"), self.p(0, "bias"))
        temb = F.linear(self.act(temb), self.p(1, "weight"), self.p(1, "bias"))
        i = 2
        h = x if self.cfg.data.centered else 2 * x - 1.
        hs = [F.conv2d(h, self.p(i, "weight"), self.p(i, "bias"), padding=1)]; i += 1
        for lvl in range(levels):
            for _ in range(m.num_res_blocks):
                h = self.res_ddpm(i, hs[-1], temb, groups=32); i += 1
                if h.shape[-1] in m.attn_resolutions:
                    h = self.attn(i, h, groups=32); i += 1
                hs.append(h)
            if lvl != levels - 1:
                hs.append(F.conv2d(F.pad(hs[-1], (0, 1, 0, 1)), self.p(i, "Conv_0.weight"), self.p(i, "Conv_0.bias"), stride=2)); i += 1
        h = hs[-1]
        h = self.res_ddpm(i, h, temb, groups=32); i += 1
        h = self.attn(i, h, groups=32); i += 1
        h = self.res_ddpm(i, h, temb, groups=32); i += 1
        for lvl in reversed(range(levels)):
            for _ in range(m.num_res_blocks + 1):
                h = self.res_ddpm(i, torch.cat([h, hs.pop()], dim=1), temb, groups=32); i += 1
            if h.shape[-1] in m.attn_resolutions:
                h = self.attn(i, h, groups=32); i += 1
            if lvl != 0:
                h = F.interpolate(h, scale_factor=2, mode="nearest")
                h = F.conv2d(h, self.p(i, "Conv_0.weight"), self.p(i, "Conv_0.bias"), padding=1); i += 1
        assert not hs
        h = self.act(F.group_norm(h, 32, self.p(i, "weight"), self.p(i, "bias"), eps=1e-6)); i += 1
        h = F.conv2d(h, self.p(i, "weight"), self.p(i, "bias"), padding=1); i += 1
        if m.scale_by_sigma:
            h = h / self.sd["sigmas"][labels.long(), None, None, None]
        return h
