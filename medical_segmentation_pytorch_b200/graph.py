"""Whole-step CUDA graph for users of the reference's training loop.

The reference's step body (core/seg_trainer.py:38-87) launches a few thousand kernels per iteration from Python.
``capture_step`` runs such a step function a few times eagerly (allocator warm-up, weight-pack plan, scratch
growth), captures it once - side-stream forks / joins and the DDP bucket all-reduces included - and returns a
callable that replays it.  The step function must read its inputs from tensors that stay at the same address:
copy each new batch into them (``x_static.copy_(x, non_blocking=True)``) before calling the replay.
"""
import torch


class CapturedStep:
    def __init__(self, fn, warmup=3, optimizers=()):
        if not torch.cuda.is_available():
            raise RuntimeError('capture_step needs a CUDA device')
        self.optimizers = list(optimizers)
        out = None
        for _ in range(max(int(warmup), 1)):
            out = fn()
            # a live eager result keeps its autograd graph - and the parameters' AccumulateGrad nodes, which remember
            # the stream they were created on - alive; the captured backward would then wait on uncaptured work
            out = _detach(out)
        del out
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = fn()
        self._done = None

    def __call__(self):
        """Replay the captured step; returns the (static) tensors the step function returned."""
        changed = [o for o in self.optimizers if o.hyper_changed()]
        if changed:
            if self._done is not None:
                self._done.synchronize()          # the previous replay has read the pinned hyper-parameter buffer
            for o in changed:
                o.refresh_hyper()
        self.graph.replay()
        if self.optimizers:
            self._done = torch.cuda.Event()
            self._done.record()
        return self.out


def _detach(out):
    if torch.is_tensor(out):
        return out.detach()
    if isinstance(out, (tuple, list)):
        return type(out)(_detach(o) for o in out)
    return out


def capture_step(fn, warmup=3, optimizers=()):
    """``replay = capture_step(step_fn, optimizers=[opt])``; ``loss = replay()``.

    ``optimizers``: the FusedOptimizers stepped inside ``fn`` whose learning rate / momentum a scheduler changes
    between iterations (their scalars are re-read at every replay).  Everything else the step does is frozen as
    captured: same shapes, same branches, inputs at the same addresses."""
    return CapturedStep(fn, warmup, optimizers)
