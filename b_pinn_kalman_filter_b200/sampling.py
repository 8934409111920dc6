"""Predictor-corrector sampling with the reference's plugin API (sampling.py:30-411) on the B200 kernels.

Public surface kept verbatim: `register_predictor/corrector`, `get_predictor/corrector`, `get_sampling_fn`,
`get_pc_sampler`, `shared_predictor_update_fn`, `shared_corrector_update_fn`, classes `Predictor`, `Corrector`,
the predictors `euler_maruyama`, `reverse_diffusion`, `ancestral_sampling`, `none` and the correctors `langevin`,
`ald`, `none`.

How it runs:
  * every built-in update is `x_mean = a x + b s ; x = x_mean + c z` with per-sample scalars, executed by ONE fused
    kernel (csrc/sde_update.cu) that also generates z (Philox) unless a noise source is injected;
  * `get_pc_sampler(...)(model)` with a B200 `ScoreModel` and built-in predictor/corrector classes takes the fused
    loop: per-step scalar tables live on the device, the score network reads its label from the table at a device
    step counter, so a whole PC step is a fixed launch sequence (CUDA-graph replayed N times, no host sync);
  * user-registered Predictor/Corrector subclasses still work: the loop then calls their `update_fn` around the engine.
"""
import abc
import functools
import math
import os

import numpy as np
import torch

from . import _lib, sde_lib, sharding
from .models import utils as mutils
from .models.utils import get_score_fn

_CORRECTORS = {}
_PREDICTORS = {}

# Optional noise source for parity tests: callable(shape, device) -> fp32 tensor; None = fused Philox on device.
NOISE_FN = None
# Fused generator: Philox4x32-10 keyed by `_SEED`, counter = (element group, stream id).  Stream ids are handed out by
# ONE process-wide counter: every update of the generic path takes one, every fused loop reserves a block of
# 16 * (N + 1) (update k of device step i draws stream base + 16 i + k), and the device-side prior takes one -- so no two
# draws under one seed ever share a stream, and successive `pc_sampler(model)` calls see fresh noise the way successive
# `torch.randn_like` calls do in the reference (sampling.py:197, 275).  `set_seed` rewinds the counter (reproducible runs).
_SEED = [0x5DEECE66D]
_STREAM = [0]
STREAMS_PER_STEP = 16      # csrc/sde_update.cu B200_OFFSETS_PER_STEP


def set_seed(seed):
    _SEED[0] = int(seed) & 0xFFFFFFFFFFFFFFFF
    _STREAM[0] = 0


def reserve_streams(n):
    """First id of a fresh block of `n` Philox stream ids."""
    base = _STREAM[0] + 1
    _STREAM[0] += int(n)
    return base


# Shard context of the generic (user-subclass) path: set by pc_sampler around its loop so that `_apply` /
# `langevin_update` draw the GLOBAL batch's noise field and all-reduce the Langevin norms like the fused loop does.
_SHARD = {"elem_base": 0, "group": None, "total_batch": None}


def _register(table, kind):
    def deco(cls=None, *, name=None):
        def _do(cls):
            local_name = cls.__name__ if name is None else name
            if local_name in table:
                raise ValueError(f'Already registered model with name: {local_name}')
            table[local_name] = cls
            return cls
        return _do if cls is None else _do(cls)
    deco.__doc__ = "A decorator for registering %s classes." % kind
    return deco


register_predictor = _register(_PREDICTORS, "predictor")
register_corrector = _register(_CORRECTORS, "corrector")


def get_predictor(name):
    return _PREDICTORS[name]


def get_corrector(name):
    return _CORRECTORS[name]


def get_sampling_fn(config, sde, shape, inverse_scaler, eps):
    """sampling.py:80-123"""
    sampler_name = config.sampling.method
    if sampler_name.lower() == 'ode':
        return get_ode_sampler(sde=sde, shape=shape, inverse_scaler=inverse_scaler,
                               denoise=config.sampling.noise_removal, eps=eps, device=config.device)
    if sampler_name.lower() == 'pc':
        return get_pc_sampler(sde=sde, shape=shape,
                              predictor=get_predictor(config.sampling.predictor.lower()),
                              corrector=get_corrector(config.sampling.corrector.lower()),
                              inverse_scaler=inverse_scaler, snr=config.sampling.snr,
                              n_steps=config.sampling.n_steps_each,
                              probability_flow=config.sampling.probability_flow,
                              continuous=config.training.continuous, denoise=config.sampling.noise_removal,
                              eps=eps, device=config.device)
    raise ValueError(f"Sampler name {sampler_name} unknown.")


# ------------------------------------------------------------------------------------------------------------------
# fused elementwise update (the only place [B,C,H,W]-sized sampler arithmetic happens)
# ------------------------------------------------------------------------------------------------------------------
def _noise(shape, device):
    return None if NOISE_FN is None else NOISE_FN(tuple(shape), device).to(device=device, dtype=torch.float32).contiguous()


def fused_update(x, score, a, b, c, z=None, seed=None, offset=0, elem_base=0):
    """Returns (x_new, x_mean) with x_mean = a x + b score, x_new = x_mean + c z; a/b/c are [B] tensors (a, c may be
    None = 1, 0).  z None -> Philox noise generated inside the kernel."""
    _lib.require_cuda(x, score)
    B = x.shape[0]
    n_per = x[0].numel()
    xn = x.detach().to(torch.float32).contiguous().clone()
    xm = torch.empty_like(xn)
    s = score.detach().to(torch.float32).contiguous()
    f = lambda v: None if v is None else v.detach().to(device=x.device, dtype=torch.float32).contiguous()
    a, b, c, z = f(a), f(b), f(c), f(z)
    with torch.cuda.device(x.device):
        rc = _lib.lib().b200_sde_update(_lib.ptr(xn), _lib.ptr(xm), _lib.ptr(s), _lib.ptr(z), _lib.ptr(a), _lib.ptr(b),
                                        _lib.ptr(c), None, 0, None, 0.0, 0.0, B, n_per,
                                        _SEED[0] if seed is None else seed, offset, elem_base, _lib.stream())
    _lib.check(rc, "b200_sde_update")
    return xn, xm


def _next_offset():
    return reserve_streams(1)


def _index(sde, t):
    return (t * (sde.N - 1) / sde.T).long()


class Predictor(abc.ABC):
    """The abstract class for a predictor algorithm."""

    def __init__(self, sde, score_fn, probability_flow=False):
        super().__init__()
        self.sde = sde
        self.rsde = sde.reverse(score_fn, probability_flow)
        self.score_fn = score_fn

    @abc.abstractmethod
    def update_fn(self, x, t):
        """One predictor update: returns (x, x_mean)."""


class Corrector(abc.ABC):
    """The abstract class for a corrector algorithm."""

    def __init__(self, sde, score_fn, snr, n_steps):
        super().__init__()
        self.sde = sde
        self.score_fn = score_fn
        self.snr = snr
        self.n_steps = n_steps

    @abc.abstractmethod
    def update_fn(self, x, t):
        """One corrector update: returns (x, x_mean)."""


def _apply(x, score, a, b, c):
    z = _noise(x.shape, x.device)
    return fused_update(x, score, a, b, c, z=z, offset=_next_offset(), elem_base=_SHARD["elem_base"])


@register_predictor(name='euler_maruyama')
class EulerMaruyamaPredictor(Predictor):
    """x_mean = x + (f(x,t) - g^2 s [/2]) dt, x = x_mean + g sqrt(-dt) z, dt = -1/N   (sampling.py:181-187)."""

    def coefficients(self, t):
        drift_coef, g = self.sde.coefficient(t)        # linear SDEs: f = drift_coef * x
        dt = -1. / self.rsde.N
        w = 0.5 if self.rsde.probability_flow else 1.
        a = 1. + drift_coef * dt
        b = -(g ** 2) * w * dt
        c = torch.zeros_like(g) if self.rsde.probability_flow else g * math.sqrt(-dt)
        return a, b, c

    def update_fn(self, x, t):
        a, b, c = self.coefficients(t)
        return _apply(x, self.score_fn(x, t), a, b, c)


@register_predictor(name='reverse_diffusion')
class ReverseDiffusionPredictor(Predictor):
    """x_mean = x - f + G^2 s [/2], x = x_mean + G z   (sampling.py:195-200 + RSDE.discretize sde_lib.py:112-117)."""

    def coefficients(self, t):
        sde = self.sde
        w = 0.5 if self.rsde.probability_flow else 1.
        if isinstance(sde, sde_lib.VESDE):
            i = _index(sde, t)
            table = sde.discrete_sigmas.to(t.device)
            sigma = table[i]
            adjacent = torch.where(i == 0, torch.zeros_like(t), table[i - 1])
            G = torch.sqrt(sigma ** 2 - adjacent ** 2)
            a = torch.ones_like(G)
        elif isinstance(sde, sde_lib.VPSDE):
            i = _index(sde, t)
            beta = sde.discrete_betas.to(t.device)[i]
            G = torch.sqrt(beta)
            a = 2. - torch.sqrt(sde.alphas.to(t.device)[i])     # x - (sqrt(alpha) x - x)
        else:                                                     # Euler-Maruyama discretisation (sde_lib.py:62-79)
            drift_coef, g = sde.coefficient(t)
            a = 1. - drift_coef / sde.N
            G = g * math.sqrt(1. / sde.N)
        b = G ** 2 * w
        c = torch.zeros_like(G) if self.rsde.probability_flow else G
        return a, b, c

    def update_fn(self, x, t):
        a, b, c = self.coefficients(t)
        return _apply(x, self.score_fn(x, t), a, b, c)


@register_predictor(name='ancestral_sampling')
class AncestralSamplingPredictor(Predictor):
    """The ancestral sampling predictor (VE / VP only; sampling.py:203-239)."""

    def __init__(self, sde, score_fn, probability_flow=False):
        super().__init__(sde, score_fn, probability_flow)
        if not isinstance(sde, sde_lib.VPSDE) and not isinstance(sde, sde_lib.VESDE):
            raise NotImplementedError(f"SDE class {sde.__class__.__name__} not yet supported.")
        assert not probability_flow, "Probability flow not supported by ancestral sampling"

    def coefficients(self, t):
        sde = self.sde
        i = _index(sde, t)
        if isinstance(sde, sde_lib.VESDE):
            table = sde.discrete_sigmas.to(t.device)
            sigma = table[i]
            adjacent = torch.where(i == 0, torch.zeros_like(t), table[i - 1])
            b = sigma ** 2 - adjacent ** 2
            c = torch.sqrt((adjacent ** 2 * (sigma ** 2 - adjacent ** 2)) / (sigma ** 2))
            return torch.ones_like(b), b, c
        beta = sde.discrete_betas.to(t.device)[i]
        inv = 1. / torch.sqrt(1. - beta)
        return inv, beta * inv, torch.sqrt(beta)

    def update_fn(self, x, t):
        a, b, c = self.coefficients(t)
        return _apply(x, self.score_fn(x, t), a, b, c)


@register_predictor(name='none')
class NonePredictor(Predictor):
    """An empty predictor that does nothing."""

    def __init__(self, sde, score_fn, probability_flow=False):
        pass

    def update_fn(self, x, t):
        return x, x


def _alpha(sde, t):
    if isinstance(sde, sde_lib.VPSDE) or isinstance(sde, sde_lib.subVPSDE):
        return sde.alphas.to(t.device)[_index(sde, t)]
    return torch.ones_like(t)


def _check_corrector_sde(sde):
    if not isinstance(sde, (sde_lib.VPSDE, sde_lib.VESDE, sde_lib.subVPSDE)):
        raise NotImplementedError(f"SDE class {sde.__class__.__name__} not yet supported.")


@register_corrector(name='langevin')
class LangevinCorrector(Corrector):
    """step = 2 alpha (snr |z| / |s|)^2 with batch-mean norms; x_mean = x + step s; x = x_mean + sqrt(2 step) z
    (sampling.py:262-282)."""

    def __init__(self, sde, score_fn, snr, n_steps):
        super().__init__(sde, score_fn, snr, n_steps)
        _check_corrector_sde(sde)

    def update_fn(self, x, t):
        alpha = _alpha(self.sde, t)
        x_mean = x
        for _ in range(self.n_steps):
            grad = self.score_fn(x, t)
            x, x_mean = langevin_update(x, grad, alpha, self.snr, process_group=_SHARD["group"],
                                        total_batch=_SHARD["total_batch"], elem_base=_SHARD["elem_base"])
        return x, x_mean


def langevin_update(x, grad, alpha, snr, z=None, process_group=None, total_batch=None, elem_base=0):
    """One fused Langevin step (norm pass + update pass).  With a process group the two batch-mean norms are
    all-reduced so that sharded batches reproduce the single-device step size exactly; `total_batch` is then the
    GLOBAL batch the means are taken over (shards may be ragged) and `elem_base` this shard's first global element."""
    _lib.require_cuda(x, grad)
    lib = _lib.lib()
    B = x.shape[0]
    n_per = x[0].numel()
    dev = x.device
    if z is None:
        z = _noise(x.shape, dev)
    xn = x.detach().to(torch.float32).contiguous().clone()
    xm = torch.empty_like(xn)
    g = grad.detach().to(torch.float32).contiguous()
    al = alpha.detach().to(device=dev, dtype=torch.float32).contiguous()
    ss = torch.empty(2, B, device=dev, dtype=torch.float32)
    sums = torch.empty(2, device=dev, dtype=torch.float32)
    seed, off = _SEED[0], _next_offset()
    total = B
    if process_group is not None:
        if total_batch is None:
            raise ValueError("langevin_update: a process group needs total_batch (the global batch the norms average over)")
        total = int(total_batch)
    with torch.cuda.device(dev):
        st = _lib.stream()
        _lib.check(lib.b200_sample_sumsq(_lib.ptr(ss[0]), _lib.ptr(ss[1]), _lib.ptr(g), _lib.ptr(z), B, n_per, seed, off,
                                         elem_base, None, st), "b200_sample_sumsq")
        _lib.check(lib.b200_langevin_norm_sums(_lib.ptr(sums), _lib.ptr(ss[0]), _lib.ptr(ss[1]), B, st),
                   "b200_langevin_norm_sums")
        if process_group is not None:
            sharding.allreduce_norm_sums(sums, process_group)
        _lib.check(lib.b200_sde_update(_lib.ptr(xn), _lib.ptr(xm), _lib.ptr(g), _lib.ptr(z), _lib.ptr(al), None, None,
                                       None, 1, _lib.ptr(sums), 1.0 / total, float(snr), B, n_per, seed, off, elem_base,
                                       st), "b200_sde_update")
    return xn, xm


@register_corrector(name='ald')
class AnnealedLangevinDynamics(Corrector):
    """step = 2 alpha (snr std(t))^2, no reductions (sampling.py:299-319)."""

    def __init__(self, sde, score_fn, snr, n_steps):
        super().__init__(sde, score_fn, snr, n_steps)
        _check_corrector_sde(sde)

    def coefficients(self, t):
        std = self.sde.marginal_coef(t)[1]
        step = (self.snr * std) ** 2 * 2 * _alpha(self.sde, t)
        return torch.ones_like(step), step, torch.sqrt(step * 2)

    def update_fn(self, x, t):
        a, b, c = self.coefficients(t)
        x_mean = x
        for _ in range(self.n_steps):
            x, x_mean = _apply(x, self.score_fn(x, t), a, b, c)
        return x, x_mean


@register_corrector(name='none')
class NoneCorrector(Corrector):
    """An empty corrector that does nothing."""

    def __init__(self, sde, score_fn, snr, n_steps):
        pass

    def update_fn(self, x, t):
        return x, x


def shared_predictor_update_fn(x, t, sde, model, predictor, probability_flow, continuous):
    """sampling.py:333-341"""
    score_fn = mutils.get_score_fn(sde, model, train=False, continuous=continuous)
    cls = NonePredictor if predictor is None else predictor
    return cls(sde, score_fn, probability_flow).update_fn(x, t)


def shared_corrector_update_fn(x, t, sde, model, corrector, continuous, snr, n_steps):
    """sampling.py:344-352"""
    score_fn = mutils.get_score_fn(sde, model, train=False, continuous=continuous)
    cls = NoneCorrector if corrector is None else corrector
    return cls(sde, score_fn, snr, n_steps).update_fn(x, t)


# ------------------------------------------------------------------------------------------------------------------
# PC sampler
# ------------------------------------------------------------------------------------------------------------------
_BUILTIN_PREDICTORS = (EulerMaruyamaPredictor, ReverseDiffusionPredictor, AncestralSamplingPredictor, NonePredictor)
_BUILTIN_CORRECTORS = (LangevinCorrector, AnnealedLangevinDynamics, NoneCorrector)


def _label_and_scale_tables(sde, timesteps, continuous):
    """Per-step network label and score multiplier (score = mult * model_out): models/utils.py:144-173."""
    t = timesteps
    if isinstance(sde, (sde_lib.VPSDE, sde_lib.subVPSDE)):
        if continuous or isinstance(sde, sde_lib.subVPSDE):
            labels = t * 999
            std = sde.marginal_coef(t)[1]
        else:
            labels = t * (sde.N - 1)
            std = sde.sqrt_1m_alphas_cumprod.to(t.device)[labels.long()]
        return labels, -1.0 / std
    if isinstance(sde, sde_lib.VESDE):
        if continuous:
            labels = sde.marginal_coef(t)[1]
        else:
            labels = torch.round((sde.T - t) * (sde.N - 1)).long().to(torch.float32)
        return labels, torch.ones_like(t)
    raise NotImplementedError(f"SDE class {sde.__class__.__name__} not yet supported.")


class FusedPCLoop:
    """State + launch sequence of the fused PC loop for one (model, batch) pair.  One `step()` = corrector
    (n_steps Langevin/ALD updates) + predictor, i.e. (n_steps + 1) score-network forwards."""

    def __init__(self, model, sde, shape, predictor, corrector, snr, n_steps, probability_flow, continuous, eps,
                 process_group=None, rank_offset=0, total_batch=None, use_graph=True, seed=None, mix=None):
        """mix: controllable generation (controllable_generation.py:44-54, 132-141) -- dict(data, mask[, M, Minv]); after
        the corrector and after the predictor of every step the known part of the state is replaced by freshly noised
        data (b200_mask_mix), inside the same launch sequence / CUDA graph."""
        eng = model.engine
        self.eng = eng
        self.sde = sde
        self.shape = tuple(shape)
        self.B = shape[0]
        self.n_per = int(np.prod(shape[1:]))
        self.dev = eng.device
        self.lib = _lib.lib()
        self.pg = process_group
        if process_group is not None and total_batch is None:
            raise ValueError("FusedPCLoop: a process group needs total_batch (the global batch the Langevin norms "
                             "average over; shards may be ragged)")
        self.total_batch = total_batch or self.B
        self.elem_base = sharding.philox_element_base(rank_offset, self.n_per)
        self.seed = _SEED[0] if seed is None else seed
        self.snr = float(snr)
        self.use_graph = use_graph and NOISE_FN is None
        # sharded Langevin: the two batch-mean norms cross the GPUs inside the reducing kernel over NVLink peer memory
        # (csrc/sde_update.cu norm_sums_xchg_kernel) -- no NCCL call in the step, so the step stays graph-capturable;
        # if the peers' memory cannot be mapped every rank falls back to an eager NCCL all-reduce of the 8 bytes
        self.xchg = None
        self.world = 1
        if process_group is not None:
            import torch.distributed as dist
            self.world = dist.get_world_size(process_group)
        if self.world > 1 and corrector is LangevinCorrector and n_steps > 0:
            if os.environ.get("B200_PEER_XCHG", "1") != "0":
                x = sharding.PeerExchange(process_group, self.dev)
                self.xchg = x if x.ok else None
        N = sde.N
        dev = self.dev
        t = torch.linspace(sde.T, eps, N, device=dev)
        labels, mult = _label_and_scale_tables(sde, t, continuous)
        score_stub = None
        # predictor tables
        self.pred = None
        if predictor is not None and predictor is not NonePredictor:
            # the class's own constructor runs its checks (ancestral sampling: VP / VE only, no probability flow)
            pobj = predictor(sde, None, probability_flow)
            a, b, c = pobj.coefficients(t)
            self.pred = [v.to(torch.float32).contiguous() for v in (a.expand_as(t) if a.dim() else a.repeat(N),
                                                                   b * mult, c)]
        # corrector tables
        self.corr = None
        self.n_steps = n_steps
        if corrector is not None and corrector is not NoneCorrector and n_steps > 0:
            corrector(sde, None, snr, n_steps)          # constructor checks (supported SDE classes)
            alpha = _alpha(sde, t).to(torch.float32)
            if corrector is LangevinCorrector:
                self.corr = ("langevin", alpha.contiguous())
                self.mult = mult.to(torch.float32).contiguous()
            else:
                std = sde.marginal_coef(t)[1]
                step = (snr * std) ** 2 * 2 * alpha
                self.corr = ("ald", torch.ones_like(step).contiguous(), (step * mult).to(torch.float32).contiguous(),
                             torch.sqrt(step * 2).to(torch.float32).contiguous())
        self.labels = labels.to(torch.float32).contiguous()
        self.step_dev = torch.zeros(1, device=dev, dtype=torch.int32)
        self.plan = eng.plan(self.B, uniform=True)     # every sample of the batch is at the same time step
        # network launches with the label read from the table at *step
        ops = list(self.plan.ops)
        fn, args, name = ops[self.plan.temb_op_index]
        args = (self.labels.data_ptr(), self.step_dev.data_ptr()) + tuple(args[2:])
        ops[self.plan.temb_op_index] = (fn, args, name)
        self.net_ops = ops
        self.x = self.plan.x_in            # the state IS the network's input buffer
        self.score = self.plan.out
        self.x_mean = torch.zeros_like(self.x)
        self.ss = torch.zeros(2, self.B, device=dev, dtype=torch.float32)
        self.sums = torch.zeros(2, device=dev, dtype=torch.float32)
        self.scaled = torch.zeros_like(self.x) if (self.corr and self.corr[0] == "langevin" and
                                                   not bool((mult == 1).all())) else None
        self.mix = None
        if mix is not None:
            import ctypes
            if len(self.shape) != 4 or self.shape[1] > 4:
                raise NotImplementedError("fused mask-mix handles [B, C <= 4, H, W] states")
            a_t, s_t = sde.marginal_coef(t)
            f32 = lambda v: v.detach().to(device=dev, dtype=torch.float32).expand(self.shape).contiguous()
            arr = lambda m: None if m is None else (ctypes.c_float * 9)(*[float(v) for v in m.reshape(-1).tolist()])
            self.mix = dict(data=f32(mix["data"]), mask=f32(mix["mask"]), M=arr(mix.get("M")), Minv=arr(mix.get("Minv")),
                            a=a_t.to(torch.float32).expand_as(t).contiguous(), s=s_t.to(torch.float32).contiguous())
        self.graph = None
        self.offset_base = reserve_streams(STREAMS_PER_STEP * (N + 1))
        self.launches_per_step = 0
        self.forwards_per_step = (n_steps if self.corr else 0) + (1 if self.pred else 0)

    # -- pieces --
    def _forward(self, st):
        for fn, args, name in self.net_ops:
            rc = fn(*args, st)
            if rc:
                _lib.check(rc, name)
        return len(self.net_ops)

    def _corrector(self, st, off):
        lib, B, n_per = self.lib, self.B, self.n_per
        n = self._forward(st)
        z = _noise(self.shape, self.dev)
        zp = _lib.ptr(z)
        if self.corr[0] == "langevin":
            score = self.score
            if self.scaled is not None:
                # VP: grad = -model/std; Langevin needs |grad| -> materialise the scaled score once (rare path)
                torch.mul(self.score, self.mult[self.step_host].item(), out=self.scaled)
                score = self.scaled
            _lib.check(lib.b200_sample_sumsq(_lib.ptr(self.ss[0]), _lib.ptr(self.ss[1]), _lib.ptr(score), zp, B, n_per,
                                             self.seed, off, self.elem_base, _lib.ptr(self.step_dev), st),
                       "b200_sample_sumsq")
            if self.xchg is not None:
                _lib.check(lib.b200_langevin_norm_sums_xchg(_lib.ptr(self.sums), _lib.ptr(self.ss[0]),
                                                            _lib.ptr(self.ss[1]), B, self.xchg.tables, self.xchg.world,
                                                            self.xchg.rank, st), "b200_langevin_norm_sums_xchg")
            else:
                _lib.check(lib.b200_langevin_norm_sums(_lib.ptr(self.sums), _lib.ptr(self.ss[0]), _lib.ptr(self.ss[1]),
                                                       B, st), "b200_langevin_norm_sums")
                sharding.allreduce_norm_sums(self.sums, self.pg)
            _lib.check(lib.b200_sde_update(_lib.ptr(self.x), _lib.ptr(self.x_mean), _lib.ptr(score), zp,
                                           _lib.ptr(self.corr[1]), None, None, _lib.ptr(self.step_dev), 1,
                                           _lib.ptr(self.sums), 1.0 / self.total_batch, self.snr, B, n_per, self.seed,
                                           off, self.elem_base, st), "b200_sde_update")
            return n + 3
        _, a, b, c = self.corr
        _lib.check(lib.b200_sde_update(_lib.ptr(self.x), _lib.ptr(self.x_mean), _lib.ptr(self.score), zp, _lib.ptr(a),
                                       _lib.ptr(b), _lib.ptr(c), _lib.ptr(self.step_dev), 0, None, 0.0, 0.0, B, n_per,
                                       self.seed, off, self.elem_base, st), "b200_sde_update")
        return n + 1

    def _predictor(self, st, off):
        n = self._forward(st)
        z = _noise(self.shape, self.dev)
        a, b, c = self.pred
        _lib.check(self.lib.b200_sde_update(_lib.ptr(self.x), _lib.ptr(self.x_mean), _lib.ptr(self.score), _lib.ptr(z),
                                            _lib.ptr(a), _lib.ptr(b), _lib.ptr(c), _lib.ptr(self.step_dev), 0, None,
                                            0.0, 0.0, self.B, self.n_per, self.seed, off, self.elem_base, st),
                   "b200_sde_update")
        return n + 1

    def _mix(self, st, off):
        m = self.mix
        z = _noise(self.shape, self.dev)
        _lib.check(self.lib.b200_mask_mix(_lib.ptr(self.x), _lib.ptr(self.x_mean), _lib.ptr(m["data"]), _lib.ptr(m["mask"]),
                                          _lib.ptr(z), _lib.ptr(m["a"]), _lib.ptr(m["s"]), _lib.ptr(self.step_dev), m["M"],
                                          m["Minv"], self.B, self.shape[1], self.shape[2] * self.shape[3], self.seed, off,
                                          self.elem_base, st), "b200_mask_mix")
        return 1

    def _step_body(self, st):
        """All launches of one PC step.  Noise streams: update k of device step i draws Philox stream
        base + i * 16 + k (the kernels add the device step counter themselves), so the launch arguments are the
        same every step and the sequence can be captured once and replayed."""
        n = 0
        k = 0
        if self.corr:
            for _ in range(self.n_steps):
                n += self._corrector(st, self.offset_base + k)
                k += 1
        if self.mix:            # the reference mixes after the corrector AND after the predictor, `none` ones included
            n += self._mix(st, self.offset_base + k)
            k += 1
        if self.pred:
            n += self._predictor(st, self.offset_base + k)
            k += 1
        if self.mix:
            n += self._mix(st, self.offset_base + k)
        _lib.check(self.lib.b200_step_advance(_lib.ptr(self.step_dev), 1, st), "b200_step_advance")
        return n + 1

    # -- API --
    def reset(self, x0=None, fresh_noise=True):
        """Rewinds the schedule to step 0.  fresh_noise: the next trajectory draws from a NEW block of Philox streams
        (the stream base is a launch argument, so a captured graph is dropped and re-captured); False replays the
        noise of the previous trajectory (A/B comparisons)."""
        self.step_dev.zero_()
        self.step_host = 0
        if fresh_noise and getattr(self, "_used", False):
            self.offset_base = reserve_streams(STREAMS_PER_STEP * (self.sde.N + 1))
            self.graph = None
        self._used = False
        if x0 is not None:
            self.x.copy_(x0)

    def prior_on_device(self, scale):
        """x_T ~ N(0, scale^2) for THIS shard of the global batch, drawn by b200_fill_normal from the loop's last
        reserved stream at the shard's global element offset: ranks of a sharded run start from disjoint slices of
        one noise field whatever their host RNG state (the reference's `sde.prior_sampling` is the host RNG)."""
        off = self.offset_base + STREAMS_PER_STEP * self.sde.N
        with torch.cuda.device(self.dev):
            _lib.check(self.lib.b200_fill_normal(_lib.ptr(self.x), self.x.numel(), float(scale), self.seed, off,
                                                 self.elem_base, _lib.stream()), "b200_fill_normal")

    def step(self):
        """One PC step launched eagerly on the current stream."""
        n = self._step_body(_lib.stream())
        self._used = True
        self.step_host += 1
        self.launches_per_step = n
        return n

    def _graphable(self):
        needs_nccl = self.world > 1 and self.corr is not None and self.corr[0] == "langevin" and self.xchg is None
        return (self.use_graph and NOISE_FN is None and not needs_nccl and self.scaled is None
                and self.n_steps + 3 <= STREAMS_PER_STEP)

    def run(self, n_iters):
        """n_iters PC steps.  After one eager step (lazy initialisation of kernels / TMA descriptors) the step is
        captured into a CUDA graph and replayed: no per-launch host cost inside the loop."""
        if n_iters <= 0:
            return 0
        if not self._graphable() or n_iters < 3:
            for _ in range(n_iters):
                self.step()
            return n_iters * self.launches_per_step
        if self.graph is None:
            self.step()
            n_iters -= 1
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._step_body(_lib.stream())
            # capture records the launches without running them
            self.graph = g
        for _ in range(n_iters):
            self.graph.replay()
        self._used = True
        self.step_host += n_iters
        return (n_iters + 1) * self.launches_per_step


def is_fusable(model, sde, predictor, corrector):
    """The fused device loop takes B200 score models with the BUILT-IN predictor / corrector classes; user-registered
    subclasses go through the generic loop that calls their `update_fn`."""
    return (isinstance(model, mutils.ScoreModel)
            and (predictor is None or predictor in _BUILTIN_PREDICTORS)
            and (corrector is None or corrector in _BUILTIN_CORRECTORS)
            and isinstance(sde, (sde_lib.VPSDE, sde_lib.subVPSDE, sde_lib.VESDE)))


def get_pc_sampler(sde, shape, predictor, corrector, inverse_scaler, snr,
                   n_steps=1, probability_flow=False, continuous=False,
                   denoise=True, eps=1e-3, device='cuda', process_group=None, rank_offset=0, total_batch=None,
                   gather_output=False, prior='auto', seed=None):
    """Create a Predictor-Corrector (PC) sampler (sampling.py:355-411).

    Beyond the reference's arguments: `process_group` / `rank_offset` / `total_batch` describe a batch-sharded run
    (one process per GPU): `shape[0]` is this rank's shard, `rank_offset` its first global sample index (decorrelates
    the Philox streams) and the Langevin batch-mean norms are all-reduced over the group so that the step size is the
    one the unsharded batch would get (sampling.py:276-278).  `gather_output=True` returns the samples of the whole
    group, concatenated in rank order, on every rank (the one data-path collective of the sharded run).
    `prior`: 'host' draws x_T with `sde.prior_sampling(shape)` (the host RNG, as the reference does, sampling.py:400) and
    copies it in; 'device' draws it on the GPU from the Philox stream at the shard's global element offset, so ranks that
    share a host seed still start from different x_T; 'auto' = 'host' unsharded, 'device' with a process group.
    `seed`: Philox key of this sampler's noise (default: the module seed, `set_seed`).  Every call of the returned
    sampler draws from fresh streams.

    Returns `pc_sampler(model) -> (samples, nfe)` with nfe = sde.N * (n_steps + 1) exactly as the reference reports
    it (sampling.py:409).
    """
    predictor_update_fn = functools.partial(shared_predictor_update_fn, sde=sde, predictor=predictor,
                                            probability_flow=probability_flow, continuous=continuous)
    corrector_update_fn = functools.partial(shared_corrector_update_fn, sde=sde, corrector=corrector,
                                            continuous=continuous, snr=snr, n_steps=n_steps)

    def fusable(model):
        return is_fusable(model, sde, predictor, corrector)

    if prior not in ('auto', 'host', 'device'):
        raise ValueError("get_pc_sampler: prior must be 'auto', 'host' or 'device' (got %r)" % (prior,))
    if process_group is not None and total_batch is None:
        raise ValueError("get_pc_sampler: a process group needs total_batch (the global batch size)")
    device_prior = prior == 'device' or (prior == 'auto' and process_group is not None)
    prior_scale = sde.sigma_max if isinstance(sde, sde_lib.VESDE) else 1.0
    n_per = int(np.prod(shape[1:]))

    def host_prior(dev):
        z = _noise(shape, dev)
        if z is not None:
            return z * prior_scale
        return sde.prior_sampling(shape).to(dev)

    def pc_sampler(model):
        with torch.no_grad():
            dev = torch.device(device) if not isinstance(model, mutils.ScoreModel) else model.engine.device
            if fusable(model):
                loop = FusedPCLoop(model, sde, shape, predictor, corrector, snr, n_steps, probability_flow,
                                   continuous, eps, process_group=process_group, rank_offset=rank_offset,
                                   total_batch=total_batch, seed=seed)
                with torch.cuda.device(dev):
                    if device_prior and NOISE_FN is None:
                        loop.reset()
                        loop.prior_on_device(prior_scale)
                    else:
                        loop.reset(host_prior(dev))
                    loop.run(sde.N)
                    # NonePredictor returns (x, x): the loop's last x_mean is then the post-corrector NOISY x
                    # (sampling.py:242-250, 407-409)
                    x = loop.x if loop.pred is None else (loop.x_mean if denoise else loop.x)
                    x = x.clone()
                    if gather_output:
                        x = sharding.gather_samples(x, process_group, total_batch)
                    return inverse_scaler(x), sde.N * (n_steps + 1)
            # generic path: user-defined predictor/corrector classes around the engine
            elem_base = sharding.philox_element_base(rank_offset, n_per)
            if device_prior and NOISE_FN is None:
                x = torch.empty(shape, device=dev, dtype=torch.float32)
                with torch.cuda.device(dev):
                    _lib.check(_lib.lib().b200_fill_normal(_lib.ptr(x), x.numel(), float(prior_scale),
                                                           _SEED[0] if seed is None else seed, _next_offset(),
                                                           elem_base, _lib.stream()), "b200_fill_normal")
            else:
                x = host_prior(dev)
            timesteps = torch.linspace(sde.T, eps, sde.N, device=dev)
            x_mean = x
            saved = dict(_SHARD)
            _SHARD.update(elem_base=elem_base, group=process_group, total_batch=total_batch)
            try:
                for i in range(sde.N):
                    vec_t = torch.ones(shape[0], device=dev) * timesteps[i]
                    x, x_mean = corrector_update_fn(x, vec_t, model=model)
                    x, x_mean = predictor_update_fn(x, vec_t, model=model)
            finally:
                _SHARD.update(saved)
            x = x_mean if denoise else x
            if gather_output:
                x = sharding.gather_samples(x, process_group, total_batch)
            return inverse_scaler(x), sde.N * (n_steps + 1)

    return pc_sampler


def get_ode_sampler(sde, shape, inverse_scaler, denoise=False, rtol=1e-5, atol=1e-5, method='RK45', eps=1e-3,
                    device='cuda', solver='device'):
    """Probability-flow ODE sampler (sampling.py:414-485).  solver='device' (RK45 / RK23): the adaptive integrator of
    ode.py keeps the state in HBM and reproduces scipy's step control; solver='scipy' (and every other `method`):
    scipy drives as in the reference, with a D2H + H2D of the state per evaluation."""

    def drift_fn(model, x, t):
        score_fn = get_score_fn(sde, model, train=False, continuous=True)
        return sde.reverse(score_fn, probability_flow=True).sde(x, t)[0]

    on_device = solver == 'device' and method in ('RK45', 'RK23')
    if solver not in ('device', 'scipy'):
        raise ValueError("get_ode_sampler: solver must be 'device' or 'scipy' (got %r)" % (solver,))

    def ode_sampler(model, z=None):
        with torch.no_grad():
            dev = model.engine.device if isinstance(model, mutils.ScoreModel) else torch.device(device)
            x = sde.prior_sampling(shape).to(dev) if z is None else z
            if on_device:
                from . import ode

                def ode_func(t, xf):
                    vec_t = torch.ones(shape[0], device=dev) * t
                    return drift_fn(model, xf.reshape(shape), vec_t).reshape(-1)

                sol = ode.solve_ivp(ode_func, (sde.T, eps), x.to(torch.float32), rtol=rtol, atol=atol, method=method)
                x = sol.y_final.reshape(shape).type(torch.float32)
            else:
                from scipy import integrate

                def ode_func(t, xf):
                    xt = mutils.from_flattened_numpy(xf, shape).to(dev).type(torch.float32)
                    vec_t = torch.ones(shape[0], device=dev) * t
                    return mutils.to_flattened_numpy(drift_fn(model, xt, vec_t))

                sol = integrate.solve_ivp(ode_func, (sde.T, eps), mutils.to_flattened_numpy(x), rtol=rtol, atol=atol,
                                          method=method)
                x = torch.tensor(sol.y[:, -1]).reshape(shape).to(dev).type(torch.float32)
            if denoise:
                score_fn = get_score_fn(sde, model, train=False, continuous=True)
                vec_eps = torch.ones(x.shape[0], device=dev) * eps
                _, x = ReverseDiffusionPredictor(sde, score_fn, probability_flow=False).update_fn(x, vec_eps)
            return inverse_scaler(x), sol.nfev

    return ode_sampler
