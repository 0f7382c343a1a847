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


def get_seg_metrics(config, metric_name):
    if metric_name == 'iou':
        return JaccardIndex(config.num_class, config.ignore_index)
    if metric_name == 'dice':
        return Dice(config.num_class)
    raise ValueError(f'Unsupported metric: {metric_name}.\n')
