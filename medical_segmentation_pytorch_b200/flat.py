"""Flat parameter / gradient arenas.

All fp32 parameters of a model are re-pointed into ONE contiguous buffer, laid out in REVERSE
registration order (the order gradients become ready in backward), with a twin gradient buffer.
That gives:
  * wgrad / BN-backward kernels a fixed destination to write each gradient into (no autograd
    accumulation copies): ops.* check ``param._msx_grad`` and call ``param._msx_ready()``;
  * DDP buckets that are contiguous slices of the gradient arena (parallel.py);
  * one-launch optimizer and EMA steps over the arena (optim.py).
Gradients are OVERWRITTEN by each backward (the reference calls ``optimizer.zero_grad()`` every
step and uses every parameter exactly once per forward, core/seg_trainer.py:38).
"""
import torch


class FlatParams:
    def __init__(self, model, align=64):
        params = [p for p in model.parameters() if p.requires_grad]
        if not params:
            raise ValueError('model has no trainable parameters')
        if any(getattr(p, '_msx_grad', None) is not None for p in params):
            # a second arena would re-point p.data / p.grad while DDP buckets and the optimizer still hold the first
            raise RuntimeError('these parameters already live in a flat arena (model._msx_flat); pass the model itself, '
                               'not a wrapper, or reuse model._msx_flat')
        dev, dt = params[0].device, params[0].dtype
        assert all(p.dtype == dt and p.device == dev for p in params)
        self.params = params
        self.order = list(reversed(params))              # arena order == gradient-ready order
        self.offsets = {}
        off = 0
        for p in self.order:
            self.offsets[id(p)] = off
            off += (p.numel() + align - 1) // align * align
        self.numel = off
        self.data = torch.zeros(off, dtype=dt, device=dev)
        self.grad = torch.zeros(off, dtype=dt, device=dev)
        for p in self.order:
            o, n = self.offsets[id(p)], p.numel()
            self.data[o:o + n].copy_(p.data.reshape(-1))
            p.data = self.data[o:o + n].view(p.shape)
            p.grad = self.grad[o:o + n].view(p.shape)
            p._msx_grad = p.grad
        model._msx_flat = self

    def span(self, p):
        o = self.offsets[id(p)]
        return o, o + p.numel()

    def rebind_grads(self):
        """Re-attach p.grad views (e.g. after a zero_grad(set_to_none=True))."""
        for p in self.order:
            p.grad = p._msx_grad


class FlatBuffers:
    """Same idea for float buffers (BN running statistics): one flat fp32 arena, so that the
    per-step rank-0 broadcast (torch DDP broadcast_buffers=True) and the EMA state copy are one
    message / one launch.  int64 ``num_batches_tracked`` counters get their own small arena."""

    def __init__(self, model):
        fbufs = [(n, b) for n, b in model.named_buffers() if b is not None and b.is_floating_point()]
        ibufs = [(n, b) for n, b in model.named_buffers() if b is not None and not b.is_floating_point()]
        self.float_names = [n for n, _ in fbufs]
        dev = next(model.parameters()).device
        self.fdata = torch.zeros(sum(b.numel() for _, b in fbufs), dtype=torch.float32, device=dev)
        self.idata = torch.zeros(max(1, sum(b.numel() for _, b in ibufs)), dtype=torch.int64, device=dev)
        off = 0
        for _, b in fbufs:
            n = b.numel()
            self.fdata[off:off + n].copy_(b.reshape(-1))
            b.data = self.fdata[off:off + n].view(b.shape)
            off += n
        off = 0
        for _, b in ibufs:
            n = b.numel()
            self.idata[off:off + n].copy_(b.reshape(-1))
            b.data = self.idata[off:off + n].view(b.shape)
            off += n
        model._msx_flat_buffers = self
