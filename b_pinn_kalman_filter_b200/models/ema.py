"""ExponentialMovingAverage with the reference's interface (models/ema.py:9-97): update / copy_to / store / restore /
state_dict / load_state_dict over `model.parameters()`.

For a B200 `ScoreModel` the shadow IS the trainer's flat EMA buffer (models/train.py) and `update` rides in the fused
optimizer kernel; for any other iterable of tensors the arithmetic of the reference is applied with torch."""
import torch


class ExponentialMovingAverage:
    def __init__(self, parameters, decay, use_num_updates=True):
        if decay < 0.0 or decay > 1.0:
            raise ValueError('Decay must be between 0 and 1')
        self.decay = decay
        self.num_updates = 0 if use_num_updates else None
        self.model = getattr(parameters, "model", None)        # set by ScoreModel.parameters()
        self.shadow_params = [p.clone().detach() for p in parameters]
        self.collected_params = []
        self._trainer = None

    # the trainer adopts the shadow into its flat buffer once it exists (first training step)
    def _bind(self, trainer):
        if self._trainer is trainer:
            return
        for k, s in zip(trainer.keys, self.shadow_params):
            trainer.ema_views[k].copy_(s)
        self.shadow_params = [trainer.ema_views[k] for k in trainer.keys]
        trainer.ema_updates = self.num_updates or 0
        self._trainer = trainer

    def update(self, parameters):
        tr = self._trainer
        if tr is not None:
            if self.num_updates is not None:
                self.num_updates += 1
                tr.ema_updates = self.num_updates - 1
                tr.ema_update(self.decay)
            else:
                tr.ema_updates = 10 ** 12          # (1 + n) / (10 + n) -> 1: the plain decay
                tr.ema_update(self.decay)
            return
        decay = self.decay
        if self.num_updates is not None:
            self.num_updates += 1
            decay = min(decay, (1 + self.num_updates) / (10 + self.num_updates))
        one_minus_decay = 1.0 - decay
        with torch.no_grad():
            for s_param, param in zip(self.shadow_params, parameters):
                s_param.sub_(one_minus_decay * (s_param - param))

    def copy_to(self, parameters):
        tr = self._trainer
        if tr is not None:
            tr.load_flat(tr.flat_ema)
            return
        for s_param, param in zip(self.shadow_params, parameters):
            param.data.copy_(s_param.data)
        if self.model is not None:
            self.model.parameters_changed()

    def store(self, parameters):
        self.collected_params = [param.clone() for param in parameters]

    def restore(self, parameters):
        tr = self._trainer
        if tr is not None:
            tr.load_flat(torch.cat([c.reshape(-1) for c in self.collected_params]))
            return
        for c_param, param in zip(self.collected_params, parameters):
            param.data.copy_(c_param.data)
        if self.model is not None:
            self.model.parameters_changed()

    def state_dict(self):
        return dict(decay=self.decay, num_updates=self.num_updates, shadow_params=self.shadow_params)

    def load_state_dict(self, state_dict):
        self.decay = state_dict['decay']
        self.num_updates = state_dict['num_updates']
        for s, v in zip(self.shadow_params, state_dict['shadow_params']):
            s.copy_(v)
