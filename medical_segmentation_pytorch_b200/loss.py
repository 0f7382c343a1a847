"""Losses with the reference's call signatures (core/loss.py:6-50) on the CUDA library.

``get_loss_fn(config, device)`` returns a callable ``(logits, labels) -> 0-d tensor``;
``kd_loss_fn(config, outputs, outputsT)`` is the distillation loss.  Each forward computes the
loss AND the gradient w.r.t. the logits in one pass over the data; backward only applies the
upstream scalar (a no-op launch when it is 1).
"""
import math

import torch
import torch.nn as nn

from . import _lib as L
from ._lib import lib


def _logits(x):
    if x.dtype != torch.float32:
        x = x.float()
    return x.contiguous()


def _labels(t):
    if t.dtype != torch.int64:
        t = t.long()
    return t.contiguous()


def _apply(logits, run):
    """run(loss, grad) launches the fused kernel; grad is None when no gradient is wanted."""
    needs = logits.requires_grad and torch.is_grad_enabled()
    if not needs:
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        run(loss, None)
        return loss
    return _ScaledGradReq.apply(logits, run)


class _ScaledGradReq(torch.autograd.Function):
    """forward produced (loss, dlogits for upstream 1); backward scales dlogits in place by the
    upstream scalar on the device (early-exit when it is 1)."""

    @staticmethod
    def forward(ctx, logits, run):
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        grad = torch.empty_like(logits)
        run(loss, grad)
        ctx.save_for_backward(grad)
        return loss

    @staticmethod
    def backward(ctx, g):
        grad, = ctx.saved_tensors
        if getattr(ctx, 'scaled', False):
            # the saved gradient was scaled in place by the first backward; a second pass would scale it twice
            raise RuntimeError('backward through a fused loss a second time (retain_graph): call the loss again instead')
        ctx.scaled = True
        g = g.to(torch.float32).contiguous()
        L.check(lib.msx_scale_by_device_scalar(L.ptr(grad), grad.numel(), L.ptr(g), L.stream()))
        return grad, None


class CrossEntropyLoss(nn.Module):
    """nn.CrossEntropyLoss(ignore_index, reduction, weight) as configured at core/loss.py:29-31."""

    def __init__(self, weight=None, ignore_index=255, reduction='mean'):
        super().__init__()
        if reduction not in ('mean', 'sum', 'none'):
            raise ValueError(f'{reduction} is not a valid value for reduction')
        self.register_buffer('weight', None if weight is None else torch.as_tensor(weight, dtype=torch.float32))
        self.ignore_index, self.reduction = ignore_index, reduction

    def forward(self, logits, labels):
        logits, labels = _logits(logits), _labels(labels)
        n, c = logits.shape[:2]
        hw = logits[0, 0].numel()
        dev = logits.device
        w = None if self.weight is None else self.weight.to(dev)
        if self.reduction == 'none':
            return _CENone.apply(logits, labels, w, self.ignore_index)
        ws = L.workspace(lib.msx_ws_bytes_ce(n * hw), dev)
        red = L.REDUCE[self.reduction]

        def run(loss, grad):
            L.check(lib.msx_ce_loss_fwd_bwd(L.ptr(logits), L.ptr(labels), n, c, hw, L.ptr(w), self.ignore_index, red,
                                            L.ptr(loss), L.ptr(grad), L.ptr(ws), ws.numel(), L.stream()))
        return _apply(logits, run)


class _CENone(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, w, ignore):
        n, c = logits.shape[:2]
        hw = logits[0, 0].numel()
        out = torch.empty(labels.shape, dtype=torch.float32, device=logits.device)
        L.check(lib.msx_ce_loss_none_fwd(L.ptr(logits), L.ptr(labels), n, c, hw, L.ptr(w), ignore, L.ptr(out), L.stream()))
        ctx.ignore = ignore
        ctx.save_for_backward(logits, labels, w)
        return out

    @staticmethod
    def backward(ctx, g):
        logits, labels, w = ctx.saved_tensors
        n, c = logits.shape[:2]
        hw = logits[0, 0].numel()
        g = g.to(torch.float32).contiguous()
        d = torch.empty_like(logits)
        L.check(lib.msx_ce_loss_none_bwd(L.ptr(logits), L.ptr(labels), n, c, hw, L.ptr(w), ctx.ignore, L.ptr(g), L.ptr(d), L.stream()))
        return d, None, None, None


class OhemCELoss(nn.Module):
    """core/loss.py:6-20 with both branches decided on the device (no host sync, no .cuda() in the ctor)."""

    def __init__(self, thresh, ignore_index=255):
        super().__init__()
        self.thresh_prob = float(thresh)
        self.thresh = -math.log(thresh)
        self.ignore_index = ignore_index

    def forward(self, logits, labels):
        logits, labels = _logits(logits), _labels(labels)
        n, c = logits.shape[:2]
        hw = logits[0, 0].numel()
        dev = logits.device
        ws = L.workspace(lib.msx_ws_bytes_loss(), dev)
        pix = torch.empty(n * hw, dtype=torch.float32, device=dev)

        def run(loss, grad):
            L.check(lib.msx_ohem_loss_fwd_bwd(L.ptr(logits), L.ptr(labels), n, c, hw, self.thresh_prob, self.ignore_index,
                                              L.ptr(loss), L.ptr(grad), L.ptr(pix), L.ptr(ws), ws.numel(), L.stream()))
        return _apply(logits, run)


class DiceBCELoss(nn.Module):
    """north_star's sigmoid+Dice+BCE extension (NOT in the reference):
    BCEWithLogits(mean) + 1 - (2*sum(p*t)+eps)/(sum(p)+sum(t)+eps), eps = 1, sums over the batch."""

    def __init__(self, eps=1.0):
        super().__init__()
        self.eps = eps

    def forward(self, logits, target):
        logits = _logits(logits)
        is_float = target.is_floating_point()
        target = (target.float() if is_float else target.long()).contiguous()
        if target.numel() != logits.numel():
            raise ValueError('dice_bce: logits [N,1,H,W] and target must have the same number of elements')
        dev = logits.device
        ws = L.workspace(lib.msx_ws_bytes_loss(), dev)

        def run(loss, grad):
            L.check(lib.msx_dice_bce_loss_fwd_bwd(L.ptr(logits), L.ptr(target), int(is_float), logits.numel(), self.eps,
                                                  L.ptr(loss), L.ptr(grad), L.ptr(ws), ws.numel(), L.stream()))
        return _apply(logits, run)


def get_loss_fn(config, device):
    """core/loss.py:23-39 (+ 'dice_bce', the north_star extension)."""
    weights = None if config.class_weights is None else torch.tensor(config.class_weights, dtype=torch.float32, device=device)
    if config.loss_type == 'ce':
        criterion = CrossEntropyLoss(weight=weights, ignore_index=config.ignore_index, reduction=config.reduction)
    elif config.loss_type == 'ohem':
        criterion = OhemCELoss(thresh=config.ohem_thrs, ignore_index=config.ignore_index)
    elif config.loss_type == 'dice_bce':
        criterion = DiceBCELoss()
    else:
        raise NotImplementedError(f"Unsupport loss type: {config.loss_type}")
    return criterion.to(device)


def kd_loss_fn(config, outputs, outputsT):
    """core/loss.py:42-50: 'kl_div' = F.kl_div(log_softmax(s/T), softmax(t/T)) (default 'mean' over ALL
    elements) * T^2;  'mse' = F.mse_loss.  The teacher is detached."""
    if config.kd_loss_type not in ('kl_div', 'mse'):
        raise NotImplementedError(f'Unsupport kd loss type: {config.kd_loss_type}')
    s, t = _logits(outputs), _logits(outputsT.detach())
    if s.shape != t.shape:
        raise ValueError('student / teacher logits differ in shape')
    n, c = s.shape[:2]
    hw = s[0, 0].numel()
    ws = L.workspace(lib.msx_ws_bytes_loss(), s.device)
    kind = 0 if config.kd_loss_type == 'kl_div' else 1
    temp = float(getattr(config, 'kd_temperature', 1.0))

    def run(loss, grad):
        L.check(lib.msx_kd_loss_fwd_bwd(L.ptr(s), L.ptr(t), n, c, hw, kind, temp, L.ptr(loss), L.ptr(grad),
                                        L.ptr(ws), ws.numel(), L.stream()))
    return _apply(s, run)
