"""Oracle: CPU restatement of the reference losses (TEST INFRASTRUCTURE ONLY).

Written as explicit formulas (log-sum-exp etc.) rather than by calling the same library
routine the reference calls, so that the golden-vector test actually pins the algorithm.
"""
import math

import torch
import torch.nn.functional as F


def ce_loss(logits, labels, class_weights=None, ignore_index=255, reduction='mean'):
    """nn.CrossEntropyLoss(ignore_index, reduction, weight) as built by core/loss.py:29-31.

    logits [N,C,H,W] float, labels [N,H,W] int64.  'mean' divides by the summed weight of the
    non-ignored pixels (NaN when there are none, like PyTorch); 'sum' / 'none' also supported.
    """
    n, c = logits.shape[:2]
    z = logits.permute(0, 2, 3, 1).reshape(-1, c)
    t = labels.reshape(-1)
    valid = t != ignore_index
    tc = torch.where(valid, t, torch.zeros_like(t))
    lse = torch.logsumexp(z, dim=1)
    nll = lse - z.gather(1, tc[:, None])[:, 0]
    w = torch.ones(c, dtype=z.dtype) if class_weights is None else torch.as_tensor(class_weights, dtype=z.dtype)
    wt = w[tc] * valid.to(z.dtype)
    per = nll * wt
    if reduction == 'none':
        return per.reshape(labels.shape)
    if reduction == 'sum':
        return per.sum()
    return per.sum() / wt.sum()


def ohem_loss(logits, labels, thresh=0.7, ignore_index=255):
    """OhemCELoss.forward, core/loss.py:13-20 (constructor's .cuda() at :9 dropped)."""
    th = -math.log(thresh)
    n_min = int((labels != ignore_index).sum().item()) // 16
    loss = ce_loss(logits, labels, None, ignore_index, 'none').reshape(-1)
    hard = loss[loss > th]
    if hard.numel() < n_min:
        hard, _ = loss.topk(n_min)
    return hard.mean()


def kd_loss(student, teacher, kind='kl_div', temperature=4.0):
    """kd_loss_fn, core/loss.py:42-50.

    'kl_div' uses F.kl_div's DEFAULT reduction ('mean' = mean over all N*C*H*W elements, not
    'batchmean'), times T^2.  'mse' is the plain element mean of squared differences.
    """
    teacher = teacher.detach()
    if kind == 'kl_div':
        T = temperature
        logp = torch.log_softmax(student / T, dim=1)
        q = torch.softmax(teacher / T, dim=1)
        logq = torch.log_softmax(teacher / T, dim=1)
        # xlogy semantics of F.kl_div: q * (log q - log p), 0 where q == 0
        el = torch.where(q > 0, q * (logq - logp), torch.zeros_like(q))
        return el.mean() * T ** 2
    if kind == 'mse':
        return ((student - teacher) ** 2).mean()
    raise NotImplementedError(kind)


def dice_bce_loss(logits, target, eps=1.0):
    """north_star's sigmoid+Dice+BCE; NOT in the reference.

    logits [N,1,H,W]; target {0,1} broadcastable to it.  loss = BCEWithLogits(mean) +
    (1 - (2*sum(p*t)+eps) / (sum(p)+sum(t)+eps)), p = sigmoid(logits), sums over the whole batch.
    """
    z = logits.reshape(-1)
    t = target.reshape(-1).to(z.dtype)
    bce = (F.softplus(z) - z * t).mean()
    p = torch.sigmoid(z)
    dice = 1.0 - (2.0 * (p * t).sum() + eps) / (p.sum() + t.sum() + eps)
    return bce + dice
