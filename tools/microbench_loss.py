"""Kernel-only GB/s of the loss / metric kernels (24 B/px CE, 16 B/px confmat, 24 B/px KD, 28 B/px dice+bce)
at the bench batch (16 x 352^2) and at an asymptotic size (256 x 352^2, working set >> L2)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import medical_segmentation_pytorch_b200 as msx  # noqa: E402
from medical_segmentation_pytorch_b200 import _lib as L  # noqa: E402

dev = torch.device('cuda')


class Cfg:
    num_class, ignore_index, class_weights, loss_type, reduction, ohem_thrs = 2, 255, None, 'ce', 'mean', 0.7
    kd_loss_type, kd_temperature = 'kl_div', 4.0


flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for n in [int(v) for v in os.environ.get('MB_N', '16,256').split(',')]:
    h = 352
    px = n * h * h
    z = torch.randn(n, 2, h, h, device=dev, requires_grad=True)
    zt = torch.randn(n, 2, h, h, device=dev)
    t = (torch.rand(n, h, h, device=dev) > 0.8).long()
    z1 = torch.randn(n, 1, h, h, device=dev, requires_grad=True)
    ce = msx.get_loss_fn(Cfg, dev)
    dice = msx.DiceBCELoss()
    metric = msx.get_seg_metrics(Cfg, 'iou').to(dev)
    jobs = {
        'ce fwd+bwd (24 B/px)': (lambda: ce(z, t), 24),
        'kd kl fwd+bwd (24 B/px)': (lambda: msx.kd_loss_fn(Cfg, z, zt), 24),
        'dice+bce fwd+bwd (28 B/px)': (lambda: dice(z1, t), 28),
        'confmat update (16 B/px)': (lambda: metric.update(z.detach(), t), 16),
    }
    for name, (fn, bpp) in jobs.items():
        fn(); torch.cuda.synchronize()
        L.profile_collect()
        L.lib.msx_profile_enable(1)
        for _ in range(5):
            flush.sum()            # evict L2 with READS: a written flush buffer leaves ~126 MB of dirty lines whose
            fn()                   # write-back would be charged to the kernel under test
        prof = L.profile_collect()
        L.lib.msx_profile_enable(0)
        ms = sum(v['ms'] for v in prof.values()) / 5
        big = max(prof.values(), key=lambda v: v['ms'])
        ms_big = big['ms'] / big['launches'] * (big['launches'] / 5)
        print(f'N={n:3d}  {name:28s} all kernels {ms * 1e3:8.1f} us  {px * bpp / ms / 1e6:7.0f} GB/s | main kernels only {px * bpp / ms_big / 1e6:7.0f} GB/s')
