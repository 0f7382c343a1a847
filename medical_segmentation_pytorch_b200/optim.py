"""One-launch optimizer and EMA steps over the flat arenas (utils/optimizer.py:8-16,
utils/model_ema.py:29-40).  Hyper-parameters follow the reference: SGD lr = base_lr*gpu_num,
momentum, weight_decay; Adam/AdamW lr = 0.1*base_lr*gpu_num with torch defaults."""
import copy

import torch

from . import _lib as L
from ._lib import lib
from .flat import FlatBuffers, FlatParams


class FusedOptimizer:
    def __init__(self, model, kind='adam', lr=1e-3, momentum=0.9, weight_decay=0.0, betas=(0.9, 0.999), eps=1e-8):
        kind = kind.lower()
        if kind not in ('sgd', 'adam', 'adamw'):
            raise NotImplementedError(f'Unsupported optimizer type: {kind}')
        self.kind, self.lr, self.momentum, self.betas, self.eps = kind, lr, momentum, betas, eps
        self.weight_decay = 0.01 if (kind == 'adamw' and weight_decay == 0.0) else weight_decay
        self.flat = getattr(model, '_msx_flat', None) or FlatParams(model)
        z = lambda: torch.zeros_like(self.flat.data)
        self.state1 = z()
        self.state2 = z() if kind != 'sgd' else None
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=self.flat.data.device)
        self.steps = 0
        self.param_groups = [{'lr': lr}]        # scheduler-facing, like torch.optim

    def zero_grad(self, set_to_none=False):
        pass                                      # gradients are overwritten by every backward (flat.py)

    def step(self, grad_scale=1.0):
        f, lr = self.flat, self.param_groups[0]['lr']
        if self.kind == 'sgd':
            L.check(lib.msx_sgd_step(L.ptr(f.data), L.ptr(f.grad), L.ptr(self.state1), f.numel, lr, self.momentum,
                                     self.weight_decay, grad_scale, int(self.steps == 0), L.stream()))
        else:
            L.check(lib.msx_adam_step(L.ptr(f.data), L.ptr(f.grad), L.ptr(self.state1), L.ptr(self.state2), f.numel, lr,
                                      self.betas[0], self.betas[1], self.eps, self.weight_decay, int(self.kind == 'adamw'),
                                      grad_scale, L.ptr(self.step_dev), L.stream()))
        self.steps += 1


def get_optimizer(config, model):
    """utils/optimizer.py:4-20 on the fused kernels."""
    if config.optimizer_type == 'sgd':
        config.lr = config.base_lr * config.gpu_num
        return FusedOptimizer(model, 'sgd', config.lr, momentum=config.momentum, weight_decay=config.weight_decay)
    if config.optimizer_type in ('adam', 'adamw'):
        config.lr = 0.1 * config.base_lr * config.gpu_num
        return FusedOptimizer(model, config.optimizer_type, config.lr)
    raise NotImplementedError(f'Unsupported optimizer type: {config.optimizer_type}')


class ModelEma:
    """utils/model_ema.py:17-40: an eval-mode deep copy refreshed every step - as two launches
    (parameter arena, float-buffer arena) plus one small int64 copy instead of one per tensor."""

    def __init__(self, model, use_ema=False, total_itrs=1):
        if getattr(model, '_msx_flat', None) is None:
            FlatParams(model)
        if getattr(model, '_msx_flat_buffers', None) is None:
            FlatBuffers(model)
        self.src_p, self.src_b = model._msx_flat, model._msx_flat_buffers
        for p in model.parameters():            # deepcopy cannot copy bound-method attributes cleanly
            for a in ('_msx_ready',):
                if hasattr(p, a):
                    pass
        self.ema = _clone_model(model)
        self.ema.eval()
        self.dst_p, self.dst_b = FlatParams(self.ema), FlatBuffers(self.ema)
        for p in self.ema.parameters():
            p.requires_grad_(False)
        self.use_ema, self.total_itrs = use_ema, max(int(total_itrs), 1)

    def update(self, model=None, cur_itrs=0):
        decay = min(max(cur_itrs / self.total_itrs, 0), 1) if self.use_ema else 0.0
        L.check(lib.msx_ema_update(L.ptr(self.dst_p.data), L.ptr(self.src_p.data), self.src_p.numel, decay, L.stream()))
        if self.src_b.fdata.numel():
            L.check(lib.msx_ema_update(L.ptr(self.dst_b.fdata), L.ptr(self.src_b.fdata), self.src_b.fdata.numel(), decay, L.stream()))
        self.dst_b.idata.copy_(self.src_b.idata)


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
