"""DDPM U-Net (registered as 'ddpm') as a plan over the B200 score engine.

Same network and parameter names as the reference's DDPM (models/ddpm.py:39-181); in this fork every `*ddpmpp*`
VP config instantiates it.  Blocks: ResnetBlockDDPM (models/layers.py:611-655), AttnBlock
(:549-573), Downsample = pad(0,1,0,1) + 3x3 stride 2 (:592-608), Upsample = nearest x2 + 3x3 (:576-589).
"""
from . import utils
from .blocks import Net


def build_ddpm(b, config):
    m = config.model
    net = Net(b, config)
    nf = m.nf
    ch_mult = tuple(m.ch_mult)
    nres = m.num_res_blocks
    levels = len(ch_mult)
    size = config.data.image_size
    res_at = [size // (2 ** i) for i in range(levels)]
    attn_res = tuple(m.attn_resolutions)
    channels = config.data.num_channels
    with_conv = bool(m.resamp_with_conv)
    if not m.conditional:
        raise NotImplementedError("unconditional DDPM is not on the sampling hot path")

    # modules 0, 1: Linear(nf, 4nf), Linear(4nf, 4nf) on the sinusoidal embedding (ddpm.py:55-61, 114-120)
    net.time_embedding("positional", nf, None, scale_by_sigma=bool(m.scale_by_sigma))
    x = net.input_image(channels, size, size, rescale=not config.data.centered)

    def resblock(x, out_ch=None, x2=None):
        # ResnetBlockDDPM: 32 groups, Conv_1 init_scale 0, NIN shortcut, no 1/sqrt(2)
        return net.resblock_ddpmpp(net.next_module(), x, x2, out_ch, skip_rescale=False, init_scale=0., groups=32)

    hs = [net.conv3x3(net.x_stem, net.next_module(), nf)]
    for lvl in range(levels):
        for _ in range(nres):
            h = resblock(hs[-1], nf * ch_mult[lvl])
            if res_at[lvl] in attn_res:
                h = net.attn_block(net.next_module(), h, skip_rescale=False, init_scale=0., groups=32)
            hs.append(h)
        if lvl != levels - 1:
            hs.append(net.downsample(net.next_module(), hs[-1], fir=False, with_conv=with_conv))

    h = resblock(hs[-1])
    h = net.attn_block(net.next_module(), h, skip_rescale=False, init_scale=0., groups=32)
    h = resblock(h)

    in_ch = h.C
    for lvl in reversed(range(levels)):
        for _ in range(nres + 1):
            h = resblock(h, nf * ch_mult[lvl], x2=hs.pop())
        in_ch = nf * ch_mult[lvl]
        if res_at[lvl] in attn_res:
            h = net.attn_block(net.next_module(), h, skip_rescale=False, init_scale=0., groups=32)
        if lvl != 0:
            h = net.upsample(net.next_module(), h, fir=False, with_conv=with_conv)
    assert not hs

    gn_idx = net.next_module()
    conv_idx = net.next_module()
    a = net.gn_act(h, "all_modules.%d" % gn_idx, in_ch, groups=32)
    net.output_conv3x3(a, conv_idx, channels, init_scale=0.)
    net.finish()


utils.register_model(build_ddpm, name="ddpm")
