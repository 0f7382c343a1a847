"""One-launch optimizer and EMA steps over the flat arenas (utils/optimizer.py:8-16,
utils/model_ema.py:29-40) and the reference's LR policies (utils/scheduler.py:5-25).

Hyper-parameters follow the reference: SGD lr = base_lr*gpu_num, momentum, weight_decay; Adam/AdamW
lr = 0.1*base_lr*gpu_num with torch defaults.  ``FusedOptimizer`` IS a ``torch.optim.Optimizer``
(one param group over all parameters), so ``torch.optim.lr_scheduler.OneCycleLR`` - which the
reference steps every iteration and which also cycles momentum / beta1 - drives it unchanged.
"""
import copy
from math import ceil

import torch

from . import _lib as L
from ._lib import lib
from .flat import FlatBuffers, FlatParams
from .parallel import de_parallel


class FusedOptimizer(torch.optim.Optimizer):
    """SGD(momentum) / Adam / AdamW over the flat parameter arena: one kernel per step.

    * Accepts the model or its data-parallel wrapper (the arena belongs to the wrapped module).
    * ``param_groups[0]`` carries lr, momentum | betas, weight_decay like torch.optim.  Launched eagerly, ``step()``
      hands them to the kernel by value.  Under CUDA-graph capture scalars would be frozen into the graph, so there
      the kernel reads them from a 16-byte device array filled by a copy from a pinned host buffer; that copy is a
      graph node which re-reads the pinned buffer at every replay, and ``graph.capture_step`` rewrites the buffer
      (``refresh_hyper()``) before a replay whenever a scheduler has changed the values - an LR schedule keeps working.
    * ``zero_grad()`` does nothing: every backward OVERWRITES the arena gradients (flat.py).  A second backward
      between two steps therefore replaces, not accumulates - the reference never does that (core/seg_trainer.py:38).
    """

    def __init__(self, model, kind='adam', lr=1e-3, momentum=0.9, weight_decay=0.0, betas=(0.9, 0.999), eps=1e-8):
        kind = kind.lower()
        if kind not in ('sgd', 'adam', 'adamw'):
            raise NotImplementedError(f'Unsupported optimizer type: {kind}')
        model = de_parallel(model)
        self.kind, self.eps = kind, eps
        if kind == 'adamw' and weight_decay == 0.0:
            weight_decay = 0.01                       # torch.optim.AdamW default (utils/optimizer.py:16 passes none)
        self.flat = getattr(model, '_msx_flat', None) or FlatParams(model)
        defaults = dict(lr=lr, weight_decay=weight_decay)
        defaults.update(dict(momentum=momentum) if kind == 'sgd' else dict(betas=tuple(betas)))
        super().__init__(self.flat.params, defaults)
        dev = self.flat.data.device
        self.state1 = torch.zeros_like(self.flat.data)
        self.state2 = torch.zeros_like(self.flat.data) if kind != 'sgd' else None
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        self.steps = 0
        self._hyper_host = torch.zeros(4, dtype=torch.float32).pin_memory() if dev.type == 'cuda' else torch.zeros(4)
        self._hyper_dev = torch.zeros(4, dtype=torch.float32, device=dev)
        self._hyper_written = None

    # the scalars of this step, as torch.optim keeps them
    def _hyper(self):
        g = self.param_groups[0]
        if self.kind == 'sgd':
            return float(g['lr']), float(g['momentum']), 0.0, float(g['weight_decay'])
        return float(g['lr']), float(g['betas'][0]), float(g['betas'][1]), float(g['weight_decay'])

    def hyper_changed(self):
        return self._hyper() != self._hyper_written

    def refresh_hyper(self):
        """Write the current param_groups scalars into the pinned buffer a captured step copies from.  The caller must
        know that no earlier replay is still going to read the buffer (graph.CapturedStep waits for the previous one)."""
        h = self._hyper()
        for i in range(4):
            self._hyper_host[i] = h[i]
        self._hyper_written = h
        return h

    def zero_grad(self, set_to_none=False):
        pass                                      # gradients are overwritten by every backward (flat.py)

    @torch.no_grad()
    def step(self, closure=None, grad_scale=1.0):
        if closure is not None:
            raise NotImplementedError('FusedOptimizer.step does not take a closure')
        f = self.flat
        lr, m1, b2, wd = self._hyper()
        hyper = None
        if f.data.is_cuda and torch.cuda.is_current_stream_capturing():
            self.refresh_hyper()
            self._hyper_dev.copy_(self._hyper_host, non_blocking=True)
            hyper = self._hyper_dev
        if self.kind == 'sgd':
            L.check(lib.msx_sgd_step(L.ptr(f.data), L.ptr(f.grad), L.ptr(self.state1), f.numel, lr, m1, wd, grad_scale,
                                     int(self.steps == 0), L.ptr(hyper), L.stream()))
        else:
            L.check(lib.msx_adam_step(L.ptr(f.data), L.ptr(f.grad), L.ptr(self.state1), L.ptr(self.state2), f.numel, lr,
                                      m1, b2, self.eps, wd, int(self.kind == 'adamw'),
                                      grad_scale, L.ptr(self.step_dev), L.ptr(hyper), L.stream()))
        self.steps += 1

    def state_dict(self):
        sd = super().state_dict()
        sd['msx'] = {'kind': self.kind, 'steps': self.steps, 'state1': self.state1, 'state2': self.state2, 'step_dev': self.step_dev}
        return sd

    def load_state_dict(self, state_dict):
        state_dict = dict(state_dict)
        extra = state_dict.pop('msx', None)
        super().load_state_dict(state_dict)
        if extra is not None:
            if extra['kind'] != self.kind:
                raise ValueError(f"optimizer state is for '{extra['kind']}', this optimizer is '{self.kind}'")
            self.steps = int(extra['steps'])
            self.state1.copy_(extra['state1'])
            if self.state2 is not None:
                self.state2.copy_(extra['state2'])
            self.step_dev.copy_(extra['step_dev'])


def get_optimizer(config, model):
    """utils/optimizer.py:4-20 on the fused kernels."""
    if config.optimizer_type == 'sgd':
        config.lr = config.base_lr * config.gpu_num
        return FusedOptimizer(model, 'sgd', config.lr, momentum=config.momentum, weight_decay=config.weight_decay)
    if config.optimizer_type in ('adam', 'adamw'):
        config.lr = 0.1 * config.base_lr * config.gpu_num
        return FusedOptimizer(model, config.optimizer_type, config.lr)
    raise NotImplementedError(f'Unsupported optimizer type: {config.optimizer_type}')


def get_scheduler(config, optimizer):
    """utils/scheduler.py:5-25, unchanged host-side scalar math on top of ``FusedOptimizer.param_groups``."""
    from torch.optim.lr_scheduler import OneCycleLR, StepLR
    if config.DDP:
        config.iters_per_epoch = ceil(config.train_num / config.train_bs / config.gpu_num)
    else:
        config.iters_per_epoch = ceil(config.train_num / config.train_bs)
    config.total_itrs = int(config.total_epoch * config.iters_per_epoch)
    if config.lr_policy == 'cos_warmup':
        return OneCycleLR(optimizer, max_lr=config.lr, total_steps=config.total_itrs, pct_start=config.warmup_epochs / config.total_epoch)
    if config.lr_policy == 'linear':
        return OneCycleLR(optimizer, max_lr=config.lr, total_steps=config.total_itrs, pct_start=0., anneal_strategy='linear')
    if config.lr_policy == 'step':
        return StepLR(optimizer, step_size=config.step_size, gamma=0.1)
    raise NotImplementedError(f'Unsupported scheduler type: {config.lr_policy}')


class ModelEma:
    """utils/model_ema.py:17-40: an eval-mode deep copy refreshed every step - as three launches
    (parameter arena, float-buffer arena, int64 counters) instead of one per tensor (1 733 for DUCK-Net)."""

    def __init__(self, model, use_ema=False, total_itrs=1):
        model = de_parallel(model)
        if getattr(model, '_msx_flat', None) is None:
            FlatParams(model)
        if getattr(model, '_msx_flat_buffers', None) is None:
            FlatBuffers(model)
        self.src_p, self.src_b = model._msx_flat, model._msx_flat_buffers
        self.ema = _clone_model(model)
        self.ema.eval()
        self.dst_p, self.dst_b = FlatParams(self.ema), FlatBuffers(self.ema)
        for p in self.ema.parameters():
            p.requires_grad_(False)
        self.use_ema, self.total_itrs = use_ema, max(int(total_itrs), 1)

    def update(self, model=None, cur_itrs=0):
        # utils/model_ema.py:35-40: decay RAMPS UP from 0 to 1 over training; every state_dict tensor, the int64
        # num_batches_tracked counters included, goes through decay * e + (1 - decay) * m
        decay = float(min(max(cur_itrs / self.total_itrs, 0), 1)) if self.use_ema else 0.0
        L.check(lib.msx_ema_update(L.ptr(self.dst_p.data), L.ptr(self.src_p.data), self.src_p.numel, decay, L.stream()))
        if self.src_b.fdata.numel():
            L.check(lib.msx_ema_update(L.ptr(self.dst_b.fdata), L.ptr(self.src_b.fdata), self.src_b.fdata.numel(), decay, L.stream()))
        L.check(lib.msx_ema_update_i64(L.ptr(self.dst_b.idata), L.ptr(self.src_b.idata), self.src_b.idata.numel(), decay, L.stream()))


def _clone_model(model):
    """deepcopy that tolerates the per-parameter hooks / views installed by FlatParams and DDP."""
    saved = []
    for p in model.parameters():
        saved.append({k: p.__dict__.pop(k) for k in ('_msx_grad', '_msx_ready', '_msx_bucket') if k in p.__dict__})
    flat, fbuf = model.__dict__.pop('_msx_flat', None), model.__dict__.pop('_msx_flat_buffers', None)
    grads = [p.grad for p in model.parameters()]
    for p in model.parameters():
        p.grad = None
    try:
        clone = copy.deepcopy(model)
    finally:
        for p, s, g in zip(model.parameters(), saved, grads):
            p.__dict__.update(s)
            p.grad = g
        if flat is not None:
            model._msx_flat = flat
        if fbuf is not None:
            model._msx_flat_buffers = fbuf
    return clone


def get_ema_model(config, model, device=None):
    return ModelEma(model, use_ema=config.use_ema, total_itrs=getattr(config, 'total_itrs', 1))
