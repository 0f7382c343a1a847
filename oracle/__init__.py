"""CPU oracle for the reference's training-step hot path.  TEST INFRASTRUCTURE ONLY.

This package is a CPU restatement (torch-CPU / numpy) of the algorithms on the hot path of
AjayGaneshC/medical-segmentation-pytorch: the UNet / DuckNet forward and backward, the
CE / OHEM / KD losses, the (north_star-named) sigmoid+Dice+BCE extension loss and the
torchmetrics-backed Dice / mIoU evaluation.  Each function cites the reference file:line it
follows.

Rules (tier ③): only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import anything from here, and only as the
checker or as the timed CPU baseline - never as the product path.  The product
(``medical_segmentation_pytorch_b200``) never imports ``oracle``.

Parity pin status
-----------------
* models, CE, KD-KL, KD-MSE: PINNED.  ``oracle/gen_golden.py`` imports the real reference modules
  from ``/root/reference`` (run in the build container, where the reference is mounted) and
  writes ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks this restatement against
  those vectors.  The reference itself ships no tests and no golden vectors.
* OHEM: pinned against the reference's ``OhemCELoss.forward`` body executed on CPU with the
  ``.cuda()`` call in its constructor bypassed (the constructor cannot run without a GPU,
  ``core/loss.py:9``); see gen_golden.py.
* Dice / mIoU metrics: PARITY UNPINNED.  The arithmetic lives in torchmetrics==1.2.0
  (``requirements.txt:67``), which is neither vendored in the reference nor installed here; the
  restatement in ``ref_metrics.py`` follows its published algorithm and is pinned only by
  hand-computed known answers.
* sigmoid+Dice+BCE: not in the reference at all; the oracle is the stated
  PyTorch composite.
"""
