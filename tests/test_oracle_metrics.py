"""Known-answer tests for the restated torchmetrics arithmetic (parity unpinned: torchmetrics
1.2.0 is absent here; these values are computed by hand from its published definitions)."""
import numpy as np
import pytest

from oracle import ref_metrics as rm


def test_argmax_ties_and_nan():
    z = np.array([[[[1.0, 2.0, np.nan, 0.0]], [[1.0, 1.0, 5.0, np.nan]]]])     # [1,2,1,4]
    # tie -> class 0; 2>1 -> 0; NaN in class 0 wins; NaN in class 1 beats 0.0
    assert rm.argmax_first(z).reshape(-1).tolist() == [0, 0, 0, 1]


def test_confmat_iou_dice_known_answer():
    # target  : 0 0 0 1 1 255 ; pred : 0 1 0 1 0 1
    logits = np.array([[[2, 0, 3, 0, 1, 0]], [[1, 5, 0, 4, 0, 9]]], dtype=np.float32)[None]   # [1,2,1,6]
    target = np.array([[[0, 0, 0, 1, 1, 255]]])
    cm = rm.confusion_matrix(logits, target, 2, ignore_index=255)
    assert cm.tolist() == [[2, 1], [1, 1]]
    np.testing.assert_allclose(rm.iou_per_class(cm), [2 / 4, 1 / 3], rtol=1e-7)
    # dice per class: 2*2/(4+1+1)=2/3 ; 2*1/(2+1+1)=1/2 -> macro 7/12
    assert abs(float(rm.dice_macro(cm)) - 7 / 12) < 1e-7
    with pytest.raises(ValueError):
        rm.confusion_matrix(logits, target, 2, ignore_index=None)           # 255 not ignored by Dice


def test_absent_class_handling():
    cm = np.array([[10, 0], [0, 0]], dtype=np.int64)
    assert rm.iou_per_class(cm).tolist() == [1.0, 0.0]      # 0/0 -> 0
    assert float(rm.dice_macro(cm)) == 1.0                  # class 1 absent -> left out of the mean


def test_metric_object_accumulates_and_resets():
    rng = np.random.default_rng(0)
    m = rm.RefMetric('iou', 2)
    tot = np.zeros((2, 2), dtype=np.int64)
    for _ in range(3):
        z = rng.standard_normal((1, 2, 8, 8)).astype(np.float32)
        t = rng.integers(0, 2, size=(1, 8, 8))
        m.update(z, t)
        tot += rm.confusion_matrix(z, t, 2, 255)
    assert (m.cm == tot).all() and tot.sum() == 3 * 64
    m.reset()
    assert m.cm.sum() == 0
    with pytest.raises(ValueError):
        rm.RefMetric('f1', 2)
