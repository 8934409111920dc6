"""NCSN++ (registered as 'ncsnpp') as a plan over the B200 score engine.

Same network, same parameter names and same module numbering as the reference's NCSNpp
(models/ncsnpp.py:34-381): the walk below visits modules in the order the reference constructor appends them
(:68-230), so `all_modules.<i>.<name>` keys of a reference checkpoint load unchanged.
What differs is how it executes: every call on `b` emits fused sm_100a kernel launches into a static plan.
"""
import math

from . import utils
from .blocks import Net


def build_ncsnpp(b, config):
    m = config.model
    net = Net(b, config)
    nf = m.nf
    ch_mult = tuple(m.ch_mult)
    nres = m.num_res_blocks
    levels = len(ch_mult)
    size = config.data.image_size
    res_at = [size // (2 ** i) for i in range(levels)]
    attn_res = tuple(m.attn_resolutions)
    progressive = m.progressive.lower()
    progressive_input = m.progressive_input.lower()
    embedding = m.embedding_type.lower()
    resblock_type = m.resblock_type.lower()
    combine = m.progressive_combine.lower()
    fir = bool(m.fir)
    skip_rescale = bool(m.skip_rescale)
    init_scale = m.init_scale
    channels = config.data.num_channels
    if not m.conditional:
        raise NotImplementedError("unconditional NCSN++ is not on the sampling hot path")
    if progressive not in ("none", "output_skip", "residual") or progressive_input not in ("none", "input_skip", "residual"):
        raise ValueError("bad progressive mode")
    if progressive == "residual":
        # needs Upsample(fir=True, with_conv=True) -> upsample_conv_2d, which raises in the reference itself
        # (models/up_or_down_sampling.py:126) and is used by no shipped config.
        raise NotImplementedError("progressive='residual' is unreachable in the reference (upsample_conv_2d is broken)")
    if not fir and (progressive != "none" or progressive_input != "none"):
        # the reference's pyramid resamplers follow `fir` (Down/Upsample(fir=fir, with_conv=False), ncsnpp.py:101, 110:
        # avg_pool2d / nearest when fir=False); only the FIR form is built here and no shipped config needs the other
        raise NotImplementedError("progressive pyramids with fir=False (avg-pool / nearest resamplers): no shipped config")
    if combine != "sum" and progressive_input == "input_skip":
        raise NotImplementedError("progressive_combine='cat' is used by no shipped config")

    # ---- time embedding (modules 0..2) ---------------------------------------------------------------------
    net.time_embedding(embedding, nf, m.fourier_scale if embedding == "fourier" else None,
                       scale_by_sigma=bool(m.scale_by_sigma))

    # ---- input ----------------------------------------------------------------------------------------------
    x = net.input_image(channels, size, size, rescale=not config.data.centered)
    pyramid_in = x if progressive_input != "none" else None
    if progressive_input in ("residual", "input_skip"):
        # 64-wide on the tensor-core path: the FIR conv-downsample ('residual') and the Combine conv1x1 ('input_skip')
        # then run on tcgen05 with K = 64 (61 zero lanes) instead of on CUDA cores with K = 3
        pyramid_in = net.x_stem

    hs = [net.conv3x3(net.x_stem, net.next_module(), nf)]
    hs_c = [nf]
    in_ch = nf

    def resblock(x, out_ch=None, x2=None, up=False, down=False):
        idx = net.next_module()
        if resblock_type == "biggan":
            return net.resblock_biggan(idx, x, x2, out_ch, up, down, fir, skip_rescale, init_scale)
        if up or down:
            raise AssertionError
        return net.resblock_ddpmpp(idx, x, x2, out_ch, skip_rescale, init_scale)

    # ---- down path ------------------------------------------------------------------------------------------
    for lvl in range(levels):
        for _ in range(nres):
            out_ch = nf * ch_mult[lvl]
            h = resblock(hs[-1], out_ch)
            in_ch = out_ch
            if res_at[lvl] in attn_res:
                h = net.attn_block(net.next_module(), h, skip_rescale, init_scale)
            hs.append(h)
            hs_c.append(in_ch)
        if lvl != levels - 1:
            if resblock_type == "ddpm":
                h = net.downsample(net.next_module(), hs[-1], fir, with_conv=bool(m.resamp_with_conv))
            else:
                h = resblock(hs[-1], down=True)
            if progressive_input == "input_skip":
                pyramid_in = net.fir_down(pyramid_in)
                # Combine(method='sum'): conv1x1(pyramid) + h   (layerspp.py:44-59)
                h = net.conv1x1(pyramid_in, "all_modules.%d.Conv_0" % net.next_module(), in_ch, res=h, out_scale=1.0)
            elif progressive_input == "residual":
                # FIR conv-downsample of the pyramid, then (pyramid + h)/sqrt(2)   (ncsnpp.py:294-301)
                idx = net.next_module()
                scale = 1.0 / math.sqrt(2.0) if skip_rescale else 1.0
                pyramid_in = net.fir_conv_down(pyramid_in, "all_modules.%d.Conv2d_0" % idx, in_ch, res=h, out_scale=scale)
                h = pyramid_in
            hs.append(h)
            hs_c.append(in_ch)

    # ---- middle ---------------------------------------------------------------------------------------------
    h = resblock(hs[-1])
    h = net.attn_block(net.next_module(), h, skip_rescale, init_scale)
    h = resblock(h)

    # ---- up path --------------------------------------------------------------------------------------------
    pyramid = None
    for lvl in reversed(range(levels)):
        for _ in range(nres + 1):
            out_ch = nf * ch_mult[lvl]
            hs_c.pop()
            h = resblock(h, out_ch, x2=hs.pop())
            in_ch = out_ch
        if res_at[lvl] in attn_res:
            h = net.attn_block(net.next_module(), h, skip_rescale, init_scale)
        if progressive == "output_skip":
            gn_idx = net.next_module()
            conv_idx = net.next_module()
            a = net.gn_act(h, "all_modules.%d" % gn_idx, in_ch)
            if pyramid is None:
                pyramid = net.conv3x3(a, conv_idx, channels, init_scale=init_scale, cout_store=8)
            else:
                pyramid = net.conv3x3(a, conv_idx, channels, init_scale=init_scale, cout_store=8,
                                      res=net.fir_up(pyramid), out_scale=1.0)
        if lvl != 0:
            if resblock_type == "ddpm":
                h = net.upsample(net.next_module(), h, fir, with_conv=bool(m.resamp_with_conv))
            else:
                h = resblock(h, up=True)
    assert not hs

    # ---- head -----------------------------------------------------------------------------------------------
    if progressive == "output_skip":
        net.output_from_nhwc(pyramid, channels)
    else:
        gn_idx = net.next_module()
        conv_idx = net.next_module()
        a = net.gn_act(h, "all_modules.%d" % gn_idx, in_ch)
        net.output_conv3x3(a, conv_idx, channels, init_scale)
    net.finish()


utils.register_model(build_ncsnpp, name="ncsnpp")
