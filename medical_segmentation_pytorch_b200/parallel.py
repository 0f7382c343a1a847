"""Data-parallel glue with the reference's entry points (utils/parallel.py:7-54), re-implemented on
flat gradient arenas instead of torch's DistributedDataParallel.

Behaviour mirrored from torch DDP as the reference configures it (defaults): parameters and buffers
are broadcast from rank 0 at wrap time; float buffers (BN running stats) are re-broadcast from rank 0
before every training forward (broadcast_buffers=True) - here as ONE flat message; gradients are
averaged over ranks in buckets filled in reverse registration order (first bucket 1 MiB, then 25 MiB)
and each bucket's all-reduce is issued on a side stream as soon as its last gradient has been
written, overlapping the rest of backward.  BatchNorm statistics stay per-GPU (north_star; the
reference's ``synBN=False`` setting - SURVEY.md §0.2 D4) unless ``config.synBN`` asks for SyncBatchNorm
(the reference's default): then every BatchNorm reduces [sum, sum of squares, count] over the ranks (convert_sync_batchnorm).
"""
import torch
import torch.distributed as dist
import torch.nn as nn

from .flat import FlatBuffers, FlatParams

FIRST_BUCKET_BYTES = 1024 * 1024
BUCKET_CAP_BYTES = 25 * 1024 * 1024


def compute_buckets(sizes_bytes, first_cap=FIRST_BUCKET_BYTES, cap=BUCKET_CAP_BYTES):
    """torch's dist._compute_bucket_assignment_by_size over parameters taken in the given order
    (callers pass REVERSE registration order): a bucket is closed once its size reaches the cap.
    Returns a list of lists of indices into ``sizes_bytes``."""
    buckets, cur, cur_bytes = [], [], 0
    limit = first_cap
    for i, b in enumerate(sizes_bytes):
        cur.append(i)
        cur_bytes += b
        if cur_bytes >= limit:
            buckets.append(cur)
            cur, cur_bytes = [], 0
            limit = cap
    if cur:
        buckets.append(cur)
    return buckets


class DistributedDataParallelB200(nn.Module):
    """Drop-in for ``DDP(model, device_ids=[rank], output_device=rank)`` (utils/parallel.py:39)."""

    def __init__(self, module, process_group=None, broadcast_buffers=True):
        super().__init__()
        self.module = module
        self.pg = process_group
        self.world = dist.get_world_size(self.pg) if dist.is_initialized() else 1
        self.broadcast_buffers = broadcast_buffers
        self.flat = getattr(module, '_msx_flat', None) or FlatParams(module)
        self.fbuf = getattr(module, '_msx_flat_buffers', None) or FlatBuffers(module)
        dev = self.flat.data.device
        self.is_cuda = dev.type == 'cuda'
        self.comm_stream = torch.cuda.Stream(device=dev) if self.is_cuda else None
        # buckets over the arena order (== reverse registration order)
        order = self.flat.order
        idx = compute_buckets([p.numel() * p.element_size() for p in order])
        self.buckets = []
        for b, members in enumerate(idx):
            lo = self.flat.span(order[members[0]])[0]
            hi = self.flat.span(order[members[-1]])[1]
            self.buckets.append({'lo': lo, 'hi': hi, 'n': len(members), 'pending': len(members), 'work': None})
            for m in members:
                order[m]._msx_bucket = b
                order[m]._msx_ready = self._make_ready(order[m])
        self._callback_queued = False
        # NCCL has ReduceOp.AVG; gloo (CPU tensors, or the one-GPU multi-process tests) does not
        self._avg_native = self.is_cuda and dist.is_initialized() and dist.get_backend(self.pg) == 'nccl'
        if self.world > 1:
            dist.broadcast(self.flat.data, src=0, group=self.pg)
            dist.broadcast(self.fbuf.fdata, src=0, group=self.pg)
            dist.broadcast(self.fbuf.idata, src=0, group=self.pg)

    # -- gradient reduction -------------------------------------------------------------------
    def _make_ready(self, p):
        def ready():
            self.mark_ready(p)
        return ready

    def mark_ready(self, p):
        """Called by the backward kernels' Python shim right after p's gradient was enqueued."""
        if self.world == 1:
            return
        if not self._callback_queued:
            self._callback_queued = True
            try:
                torch.autograd.Variable._execution_engine.queue_callback(self.finish)
            except RuntimeError:
                pass                                   # not inside backward (tests drive finish() themselves)
        bk = self.buckets[p._msx_bucket]
        bk['pending'] -= 1
        if self.is_cuda:                               # gradients of one bucket may be produced on several streams
            cur = torch.cuda.current_stream()
            bk.setdefault('streams', {})[cur.cuda_stream] = cur
        if bk['pending'] == 0:
            self._launch(bk)

    def _launch(self, bk):
        view = self.flat.grad[bk['lo']:bk['hi']]
        if self.is_cuda:
            streams = bk.pop('streams', None) or {0: torch.cuda.current_stream()}
            for s in streams.values():                 # everything enqueued so far on each producing stream
                self.comm_stream.wait_event(s.record_event())
            with torch.cuda.stream(self.comm_stream):
                op = dist.ReduceOp.AVG if self._avg_native else dist.ReduceOp.SUM
                bk['work'] = dist.all_reduce(view, op=op, group=self.pg, async_op=True)
        else:
            bk['work'] = dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)

    def finish(self):
        """Wait for every bucket (end of backward); gradients in the arena are then the rank mean."""
        self._callback_queued = False
        if self.world == 1:
            return
        for bk in self.buckets:
            if bk['pending'] != 0:
                # parameters that received no gradient this step keep their previous value; reduce anyway
                self._launch(bk)
            bk['work'].wait()
            if not (self.is_cuda and self._avg_native):
                self.flat.grad[bk['lo']:bk['hi']].div_(self.world)
            bk['pending'], bk['work'] = bk['n'], None
        if self.is_cuda:
            torch.cuda.current_stream().wait_stream(self.comm_stream)

    # -- forward --------------------------------------------------------------------------------
    def forward(self, *args, **kwargs):
        if self.world > 1 and self.broadcast_buffers and self.module.training and self.fbuf.fdata.numel():
            dist.broadcast(self.fbuf.fdata, src=0, group=self.pg)       # torch DDP: _sync_buffers before forward
        return self.module(*args, **kwargs)

    def state_dict(self, *args, **kwargs):
        # torch DDP prefixes keys with 'module.'; the reference always saves de_parallel(model).state_dict()
        return self.module.state_dict(*args, **kwargs)

    def load_state_dict(self, *args, **kwargs):
        return self.module.load_state_dict(*args, **kwargs)


class SyncGroupRef:
    """Holder of the process group a BatchNorm synchronises over; survives copy.deepcopy of the model (ModelEma, the
    reference's get_ema_model) - a ProcessGroup itself cannot be copied."""

    def __init__(self, group):
        self.group = group

    def __deepcopy__(self, memo):
        return self


def convert_sync_batchnorm(model, process_group=None):
    """nn.SyncBatchNorm.convert_sync_batchnorm (utils/parallel.py:38) for these modules: every BatchNorm2d keeps its
    class, parameters and state_dict keys and is marked to take its training-mode statistics over all ranks of
    ``process_group`` (modules._sync_group -> ops.SyncBNActFn).  Eval mode and single-process runs are unchanged."""
    group = process_group if process_group is not None else (dist.group.WORLD if dist.is_initialized() else None)
    for m in model.modules():
        if isinstance(m, nn.BatchNorm2d):
            m._msx_sync_group = SyncGroupRef(group)
    return model


def is_parallel(model):
    return type(model) in (nn.parallel.DataParallel, nn.parallel.DistributedDataParallel, DistributedDataParallelB200)


def de_parallel(model):
    return model.module if is_parallel(model) else model


def set_device(config, rank):
    """utils/parallel.py:17-32 (DDP branch; the single-process nn.DataParallel branch is out of scope)."""
    if config.DDP:
        torch.cuda.set_device(rank)
        if not dist.is_initialized():
            dist.init_process_group(backend=dist.Backend.NCCL, init_method='env://')
        device = torch.device('cuda', rank)
        config.gpu_num = dist.get_world_size()
    else:
        if not torch.cuda.is_available():
            raise RuntimeError('medical_segmentation_pytorch_b200 needs a CUDA device (no CPU path)')
        device = torch.device('cuda')
        config.gpu_num = 1
    config.num_workers = config.gpu_num * getattr(config, 'base_workers', 8)
    return device


def parallel_model(config, model, rank, device):
    """utils/parallel.py:35-44."""
    if config.DDP:
        if getattr(config, 'synBN', False):
            convert_sync_batchnorm(model)
        return DistributedDataParallelB200(model.to(device))
    model.to(device)
    return model


def destroy_ddp_process(config):
    if config.DDP and getattr(config, 'destroy_ddp_process', True) and dist.is_initialized():
        dist.destroy_process_group()


def sampler_set_epoch(config, loader, cur_epochs):
    if config.DDP:
        loader.sampler.set_epoch(cur_epochs)
