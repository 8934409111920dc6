"""Native operators, same exports as the reference's op/__init__.py:1-2; `ns_step` and `grid_sample` are
sub-modules imported explicitly (`from op import ns_step`), as in the reference."""
from .fused_act import FusedLeakyReLU, fused_leaky_relu
from .upfirdn2d import upfirdn2d
