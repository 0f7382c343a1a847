"""Oracle: numpy restatement of the reference's Dice / mIoU metrics (TEST INFRASTRUCTURE ONLY).

PARITY UNPINNED.  ``utils/metrics.py:4-12`` only instantiates torchmetrics==1.2.0 objects
(``requirements.txt:67``): ``JaccardIndex(task='multiclass', num_classes, ignore_index,
average='none')`` and the legacy ``Dice(num_classes, average='macro')``.  torchmetrics is not
vendored in the reference and is not installed in this image, so what follows restates its
published algorithm and is pinned by hand-computed known answers in
``tests/test_oracle_metrics.py`` rather than by outputs of the library itself.

Integer work (argmax, confusion-matrix counts) is exact; the ratios are formed in float32 after
casting the int64 counts, which is the order torchmetrics uses.
"""
import numpy as np


def argmax_first(logits):
    """torch.argmax(dim=1) semantics: first maximal index wins ties; NaN counts as maximal."""
    z = np.asarray(logits)
    n, c = z.shape[:2]
    best = np.zeros((n,) + z.shape[2:], dtype=np.int64)
    bv = z[:, 0].copy()
    for k in range(1, c):
        v = z[:, k]
        # strictly greater replaces; a NaN replaces any non-NaN; an earlier NaN is never replaced
        take = (v > bv) | (np.isnan(v) & ~np.isnan(bv))
        best = np.where(take, k, best)
        bv = np.where(take, v, bv)
    return best


def confusion_matrix(logits, target, num_classes, ignore_index=None):
    """C x C int64 counts, rows = target class, cols = predicted class.

    Pixels whose target equals ``ignore_index`` are dropped (JaccardIndex passes 255; the legacy
    Dice passes none - utils/metrics.py:6-9).  Raises on a label outside [0, C) that is not the
    ignore index, as torchmetrics' input validation does.
    """
    pred = argmax_first(logits).reshape(-1)
    t = np.asarray(target).reshape(-1).astype(np.int64)
    if ignore_index is not None:
        keep = t != ignore_index
        pred, t = pred[keep], t[keep]
    if t.size and (t.min() < 0 or t.max() >= num_classes):
        raise ValueError('target label outside [0, num_classes)')
    return np.bincount(t * num_classes + pred, minlength=num_classes * num_classes) \
             .reshape(num_classes, num_classes).astype(np.int64)


def iou_per_class(confmat):
    """JaccardIndex(average='none').compute(): diag / (rowsum + colsum - diag), 0 where 0/0."""
    cm = confmat.astype(np.float32)
    num = np.diag(cm)
    den = cm.sum(0) + cm.sum(1) - num
    return (num / np.where(den == 0, np.float32(1), den)).astype(np.float32)


def dice_macro(confmat):
    """Legacy Dice(average='macro', mdmc_average='global', zero_division=0).compute().

    Per class 2tp / (2tp + fp + fn); classes with tp+fp+fn == 0 are left out of the mean.
    """
    cm = confmat.astype(np.int64)
    tp = np.diag(cm)
    fp = cm.sum(0) - tp
    fn = cm.sum(1) - tp
    present = (tp + fp + fn) > 0
    if not present.any():
        return np.float32(0)
    num = (2 * tp).astype(np.float32)
    den = (2 * tp + fp + fn).astype(np.float32)
    score = np.where(den == 0, np.float32(0), num / np.where(den == 0, np.float32(1), den)).astype(np.float32)
    return np.float32(score[present].mean(dtype=np.float32))


class RefMetric:
    """update/compute/reset object with the call pattern of core/seg_trainer.py:118-125,144-145."""

    def __init__(self, name, num_classes, ignore_index=255):
        if name not in ('iou', 'dice'):
            raise ValueError(f'Unsupported metric: {name}.\n')           # utils/metrics.py:11
        self.name, self.c = name, num_classes
        self.ignore = ignore_index if name == 'iou' else None
        self.reset()

    def reset(self):
        self.cm = np.zeros((self.c, self.c), dtype=np.int64)

    def update(self, logits, target):
        self.cm += confusion_matrix(np.asarray(logits), np.asarray(target), self.c, self.ignore)

    def compute(self):
        return iou_per_class(self.cm) if self.name == 'iou' else dice_macro(self.cm)
