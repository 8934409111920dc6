"""Training losses, optimizer glue and the one-step function with the reference's interface (losses.py:28-222):
`get_optimizer`, `optimization_manager`, `get_sde_loss_fn`, `get_smld_loss_fn`, `get_ddpm_loss_fn`, `get_step_fn`.

    optimizer = losses.get_optimizer(config, model.parameters())
    ema = ExponentialMovingAverage(model.parameters(), decay=config.model.ema_rate)
    state = dict(optimizer=optimizer, model=model, ema=ema, step=0)
    train_step_fn = losses.get_step_fn(sde, train=True, optimize_fn=losses.optimization_manager(config), ...)
    loss = train_step_fn(state, batch)          # exactly the calls run_lib.train makes (run_lib.py:76-141)

Everything [B, C, H, W]- or parameter-sized runs in the kernels behind `models.train.Trainer`; this file is the calling
convention.  With `torch.distributed` initialised, pass `process_group=` to `get_step_fn` for data-parallel training:
the flat gradient is summed over the ranks in buckets (NCCL) before the optimizer kernel.
(The PINN / B-PINN step functions of the reference, losses.py:233-387, belong to another pipeline and are not here.)"""
import numpy as np
import torch

from . import sde_lib
from .models import utils as mutils
from .models.train import Trainer


class Adam:
    """torch.optim.Adam's hyper-parameters for the fused optimizer kernel (`param_groups[0]['lr']` is what the
    reference's warm-up writes, losses.py:57-59)."""

    def __init__(self, params, lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        self.model = getattr(params, "model", None)
        if self.model is None:
            raise TypeError("pass `model.parameters()` of a B200 ScoreModel")
        self.param_groups = [dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay)]
        self.max_norm = -1.0

    def zero_grad(self):
        pass                                   # the backward pass overwrites the flat gradient buffer

    def step(self):
        g = self.param_groups[0]
        tr = self.model.trainer
        tr.adam_step(g["lr"], g["betas"][0], g["betas"][1], g["eps"], g["weight_decay"], self.max_norm)
        self.max_norm = -1.0

    def state_dict(self):
        tr = getattr(self.model, "trainer", None)
        return dict(param_groups=self.param_groups, step=0 if tr is None else tr.opt_steps,
                    exp_avg=None if tr is None else tr.flat_m, exp_avg_sq=None if tr is None else tr.flat_v)


def get_optimizer(config, params, lr_mul=1.0):
    """losses.py:28-45"""
    if config.optim.optimizer == 'Adam':
        return Adam(params, lr=config.optim.lr * lr_mul, betas=(config.optim.beta1, 0.999), eps=config.optim.eps,
                    weight_decay=config.optim.weight_decay)
    raise NotImplementedError(f'Optimizer {config.optim.optimizer} not supported yet!')


def optimization_manager(config):
    """Returns an optimize_fn based on `config` (losses.py:48-66)."""
    lr = config.optim.lr

    def optimize_fn(optimizer, params, step, lr=lr, warmup=config.optim.warmup, grad_clip=config.optim.grad_clip):
        """Optimizes with warmup and gradient clipping (disabled if negative)."""
        if warmup > 0:
            for g in optimizer.param_groups:
                g['lr'] = lr * np.minimum(step / warmup, 1.0)
        if grad_clip >= 0:
            optimizer.max_norm = float(grad_clip)      # clip_grad_norm_ happens inside the fused optimizer kernel
        optimizer.step()

    return optimize_fn


def _trainer(model, sde, batch, **kw):
    tr = getattr(model, "trainer", None)
    sig = (int(batch.shape[0]), id(sde), tuple(sorted(kw.items(), key=lambda i: i[0])))
    if tr is None or getattr(tr, "_sig", None) != sig:
        if tr is not None:
            tr.flush()
        tr = Trainer(model, sde, batch.shape[0], **kw)
        tr._sig = sig
        model.trainer = tr
    return tr


def _loss_fn(sde, train, process_group=None, **kw):
    def loss_fn(model, batch, t=None, z=None):
        """loss = loss_fn(model, batch).  train=True also leaves d loss / d parameters in the trainer's gradient buffer
        (the reference gets them from `loss.backward()`); train=False evaluates without dropout."""
        tr = _trainer(model, sde, batch, process_group=process_group, **kw)
        if train:
            return tr.loss_and_grads(batch, t=t, z=z)
        return tr.eval_loss(batch, t=t, z=z)
    loss_fn.trainer_for = lambda model, batch: _trainer(model, sde, batch, process_group=process_group, **kw)
    return loss_fn


def get_sde_loss_fn(sde, train, reduce_mean=True, continuous=True, likelihood_weighting=True, eps=1e-5, process_group=None):
    """losses.py:69-112"""
    return _loss_fn(sde, train, process_group, continuous=True, reduce_mean=reduce_mean,
                    likelihood_weighting=likelihood_weighting, eps=eps)


def get_smld_loss_fn(vesde, train, reduce_mean=False, process_group=None):
    """losses.py:115-137"""
    assert isinstance(vesde, sde_lib.VESDE), "SMLD training only works for VESDEs."
    return _loss_fn(vesde, train, process_group, continuous=False, reduce_mean=reduce_mean)


def get_ddpm_loss_fn(vpsde, train, reduce_mean=True, process_group=None):
    """losses.py:140-162"""
    assert isinstance(vpsde, sde_lib.VPSDE), "DDPM training only works for VPSDEs."
    return _loss_fn(vpsde, train, process_group, continuous=False, reduce_mean=reduce_mean)


def get_step_fn(sde, train, optimize_fn=None, reduce_mean=False, continuous=True, likelihood_weighting=False,
                process_group=None):
    """Create a one-step training/evaluation function (losses.py:165-222)."""
    if continuous:
        loss_fn = get_sde_loss_fn(sde, train, reduce_mean=reduce_mean, continuous=True,
                                  likelihood_weighting=likelihood_weighting, process_group=process_group)
    else:
        assert not likelihood_weighting, "Likelihood weighting is not supported for original SMLD/DDPM training."
        if isinstance(sde, sde_lib.VESDE):
            loss_fn = get_smld_loss_fn(sde, train, reduce_mean=reduce_mean, process_group=process_group)
        elif isinstance(sde, sde_lib.VPSDE):
            loss_fn = get_ddpm_loss_fn(sde, train, reduce_mean=reduce_mean, process_group=process_group)
        else:
            raise ValueError(f"Discrete training for {sde.__class__.__name__} is not recommended.")

    def step_fn(state, batch, t=None, z=None):
        model = state['model']
        if train:
            optimizer = state['optimizer']
            optimizer.zero_grad()
            loss = loss_fn(model, batch, t=t, z=z)
            state['ema']._bind(model.trainer)
            # (`model.parameters` uncalled: neither callee reads the list, and reading it would force the recorded
            # optimizer step to execute before the EMA update could join its pass over the parameters)
            optimize_fn(optimizer, model.parameters, step=state['step'])
            state['step'] += 1
            state['ema'].update(model.parameters)
        else:
            with torch.no_grad():
                ema = state['ema']
                ema._bind(loss_fn.trainer_for(model, batch))
                ema.store(model.parameters())
                ema.copy_to(model.parameters())
                loss = loss_fn(model, batch, t=t, z=z)
                ema.restore(model.parameters())
        return loss

    return step_fn
