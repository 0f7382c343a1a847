"""Synthetic 'polyp-shaped' batches: ImageNet-normalised-range noise images and
int64 {0,1} masks made of 1-3 filled rotated ellipses (~10-20 % foreground)."""
import numpy as np
import torch


def synth_batch(n, h, w, seed=1, ignore_band=0.0, n_class=2):
    rng = np.random.Generator(np.random.PCG64(seed))
    img = np.clip(rng.standard_normal((n, 3, h, w), dtype=np.float32), -2.12, 2.64)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    mask = np.zeros((n, h, w), dtype=np.int64)
    for i in range(n):
        for _ in range(int(rng.integers(1, 4))):
            cy, cx = rng.uniform(0.2, 0.8) * h, rng.uniform(0.2, 0.8) * w
            a, b = rng.uniform(0.08, 0.25, size=2) * min(h, w)
            th = rng.uniform(0, np.pi)
            u = (yy - cy) * np.cos(th) + (xx - cx) * np.sin(th)
            v = -(yy - cy) * np.sin(th) + (xx - cx) * np.cos(th)
            mask[i][(u / a) ** 2 + (v / b) ** 2 <= 1.0] = n_class - 1
    if ignore_band > 0:
        mask[rng.uniform(size=mask.shape) < ignore_band] = 255
    return torch.from_numpy(img), torch.from_numpy(mask)
