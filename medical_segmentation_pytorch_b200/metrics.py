"""Dice / mIoU evaluation with the call pattern the reference uses on torchmetrics objects
(utils/metrics.py:4-12; core/seg_trainer.py:22,118-125,144-145): ``.to(device)``, ``.update(preds,
target)``, ``.compute()``, ``.reset()``.

State is one C x C int64 confusion matrix on the device, accumulated by a single kernel per
update (argmax + count) without any host synchronisation; ``compute()`` does the tiny ratio
arithmetic in float32 after casting the counts, the order torchmetrics uses.  Under
torch.distributed the matrix is all-reduced (sum) once in ``compute()``.
"""
import torch
import torch.distributed as dist

from . import _lib as L
from ._lib import lib


class _ConfMatMetric:
    def __init__(self, num_classes, ignore_index):
        self.num_classes = int(num_classes)
        self.ignore_index = -1 if ignore_index is None else int(ignore_index)
        self.device = torch.device('cpu')
        self.confmat = None
        self._bad = None

    def to(self, device):
        self.device = torch.device(device)
        if self.device.type != 'cuda':
            raise L.MsxError('metrics run on the CUDA device only (no CPU fallback)')
        self.confmat = torch.zeros(self.num_classes, self.num_classes, dtype=torch.int64, device=self.device)
        self._bad = torch.zeros(1, dtype=torch.int32, device=self.device)
        return self

    def reset(self):
        if self.confmat is not None:
            self.confmat.zero_()
            self._bad.zero_()

    @torch.no_grad()
    def update(self, preds, target):
        if self.confmat is None:
            self.to(preds.device)
        preds = preds.detach()
        if preds.dtype != torch.float32:
            preds = preds.float()
        preds = preds.contiguous()
        target = target.long().contiguous() if target.dtype != torch.int64 or not target.is_contiguous() else target
        n, c = preds.shape[:2]
        if c != self.num_classes:
            raise ValueError(f'expected {self.num_classes} class channels, got {c}')
        hw = preds[0, 0].numel()
        L.check(lib.msx_confmat_update(L.ptr(preds), L.ptr(target), n, c, hw, self.ignore_index, L.ptr(self.confmat),
                                       L.ptr(self._bad), L.stream()))

    @torch.no_grad()
    def update_from_features(self, feat, head_weight, target):
        """Same counts as ``update(model(images), target)`` without the logits ever reaching HBM:
        ``feat`` is ``model.features(images)`` (NHWC ``Act``), ``head_weight`` is ``model.seg_head.weight``; the 1x1 head,
        the argmax and the count run as one kernel (msx_head_confmat_update)."""
        t = feat.t
        if self.confmat is None:
            self.to(t.device)
        nc = head_weight.shape[0]
        if nc != self.num_classes:
            raise ValueError(f'expected {self.num_classes} class channels, got {nc}')
        target = target.long().contiguous() if target.dtype != torch.int64 or not target.is_contiguous() else target
        n, h, w, _ = t.shape
        if target.numel() != n * h * w:
            raise ValueError(f'target has {target.numel()} pixels, features have {n * h * w}')
        wt = head_weight.detach()
        if wt.dtype != torch.float32 or not wt.is_contiguous():
            raise L.MsxError('seg_head weight must be contiguous fp32')
        L.check(lib.msx_head_confmat_update(L.dtype_code(t.dtype), L.ptr(t), t.stride(2), n * h * w, feat.c, L.ptr(wt), nc,
                                            L.ptr(target), self.ignore_index, L.ptr(self.confmat), L.ptr(self._bad), L.stream()))

    def _synced(self):
        cm = self.confmat
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            cm = cm.clone()
            dist.all_reduce(cm, op=dist.ReduceOp.SUM)
        if int(self._bad.item()):
            raise ValueError('target label outside [0, num_classes) (and not the ignore index)')
        return cm


class JaccardIndex(_ConfMatMetric):
    """torchmetrics.JaccardIndex(task='multiclass', average='none', ignore_index) (utils/metrics.py:6-7)."""

    def compute(self):
        cm = self._synced().to(torch.float32)
        num = torch.diag(cm)
        den = cm.sum(0) + cm.sum(1) - num
        return num / torch.where(den == 0, torch.ones_like(den), den)


class Dice(_ConfMatMetric):
    """Legacy torchmetrics.Dice(num_classes, average='macro') (utils/metrics.py:8-9): no ignore_index;
    classes with tp+fp+fn == 0 are left out of the mean."""

    def __init__(self, num_classes):
        super().__init__(num_classes, None)

    def compute(self):
        cm = self._synced()
        tp = torch.diag(cm)
        fp = cm.sum(0) - tp
        fn = cm.sum(1) - tp
        present = (tp + fp + fn) > 0
        num = (2 * tp).to(torch.float32)
        den = (2 * tp + fp + fn).to(torch.float32)
        score = torch.where(den == 0, torch.zeros_like(den), num / torch.where(den == 0, torch.ones_like(den), den))
        if not bool(present.any()):
            return torch.zeros((), dtype=torch.float32, device=cm.device)
        return score[present].mean()


@torch.no_grad()
def eval_sweep(model, images, masks, metrics, batch_size=16):
    """The validation loop of core/seg_trainer.py:100-125 for same-sized images: the reference feeds ``val_bs = 1`` image at
    a time (configs/my_config.py:27) and hands the logits to each metric; here ``batch_size`` images go through the
    eval-mode model together (per-image results do not depend on the batch in eval mode: BatchNorm uses running
    statistics) and each metric counts straight from the features.  ``images`` [N,C,H,W], ``masks`` [N,H,W] on the
    device; ``model`` in eval mode (the trainer passes ``ema_model.ema``).  Returns nothing: call ``metric.compute()``."""
    if model.training:
        raise ValueError('eval_sweep expects model.eval() (core/seg_trainer.py:114 validates the EMA copy in eval mode)')
    for i in range(0, images.shape[0], batch_size):
        feat = model.features(images[i:i + batch_size])
        for m in metrics:
            m.update_from_features(feat, model.seg_head.weight, masks[i:i + batch_size])


def get_seg_metrics(config, metric_name):
    if metric_name == 'iou':
        return JaccardIndex(config.num_class, config.ignore_index)
    if metric_name == 'dice':
        return Dice(config.num_class)
    raise ValueError(f'Unsupported metric: {metric_name}.\n')
