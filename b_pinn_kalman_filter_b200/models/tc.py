"""Tensor-core (tcgen05 / TMEM / TMA) launch selection for the score engine.  See csrc/conv_tc.cu."""

ENABLED = False


def conv_supported(x, x2, out, ksize, stride, pad):
    return False


def attention_supported(L, C):
    return False
