"""Linear observation operators (reference inverse/operators.py).  Only what the DPS hot path touches: the operator
applied to an image batch and, because the engine replaces autograd, its adjoint."""
import torch


class LinearOperator:
    def __call__(self, x, keep_shape=True):
        raise NotImplementedError

    def adjoint(self, r, shape):
        """A^T r as an image batch of `shape` (r as returned by __call__(x, keep_shape=False))."""
        raise NotImplementedError


class InpaintOperator(LinearOperator):
    """Binary mask, 1 = observed (inverse/operators.py:133-166).  keep_shape=False returns the observed pixels only,
    [B, C, K] with K observed pixels per image (the reference's `bcmm(pL, x)`: a 0/1 selection matrix), which needs
    the same K for every image of the batch, as the reference's torch.stack of the per-image matrices does."""

    def __init__(self, mask):
        self.mask = mask

    def _idx(self, device):
        m = self.mask.to(device) > 0.5
        B, C = m.shape[:2]
        flat = m.reshape(B, C, -1)
        K = int(flat[0, 0].sum())
        if not bool((flat.sum(-1) == K).all()):
            raise ValueError("InpaintOperator: every image must observe the same number of pixels")
        order = torch.argsort((~flat).to(torch.int8), dim=-1, stable=True)[..., :K]      # observed pixels, in raster order
        return order, K

    def __call__(self, x, keep_shape=True, invert=False):
        assert self.mask.shape == x.shape
        m = self.mask.to(x.device)
        if keep_shape:
            return (1 - m) * x if invert else m * x
        if invert:
            raise NotImplementedError("invert + keep_shape=False is only used by the controlled sampler")
        idx, _ = self._idx(x.device)
        return torch.gather(x.reshape(x.shape[0], x.shape[1], -1), 2, idx)

    def adjoint(self, r, shape):
        idx, _ = self._idx(r.device)
        out = torch.zeros(shape[0], shape[1], shape[2] * shape[3], device=r.device, dtype=r.dtype)
        out.scatter_(2, idx, r)
        return out.reshape(shape)
