"""Conv / BN / activation building blocks with the reference's constructors and state_dict keys
(models/modules.py:7-15, :73-131 - only the classes some model actually instantiates; the dead
depth-wise / pyramid-pooling / SegHead classes are out of scope).

Parameters live in ordinary ``nn.Conv2d`` / ``nn.ConvTranspose2d`` / ``nn.BatchNorm2d`` holders that
are constructed in the reference's order (same default init, same RNG consumption, same keys);
their own ``forward`` is never called - the module forwards below go to the CUDA library.
"""
import torch
import torch.nn as nn

from . import ops
from .ops import Act, ConvGeom


def conv3x3(in_channels, out_channels, stride=1, bias=False):
    return nn.Conv2d(in_channels, out_channels, kernel_size=3, stride=stride, padding=1, bias=bias)


def conv1x1(in_channels, out_channels, stride=1, bias=False):
    return nn.Conv2d(in_channels, out_channels, kernel_size=1, stride=stride, padding=0, bias=bias)


_FUSED_ACTS = {'relu': 1, 'none': 0}
_KNOWN_ACTS = ('relu', 'relu6', 'leakyrelu', 'prelu', 'celu', 'elu', 'hardswish', 'hardtanh', 'gelu', 'glu',
               'selu', 'silu', 'sigmoid', 'softmax', 'tanh', 'none')


class Activation(nn.Module):
    """Parameter-free placeholder keeping the reference's Sequential indices (models/modules.py:111-131).

    The activation itself is fused into the BatchNorm-apply kernel.  Only 'relu' (the sole type any
    reference model can reach, models/__init__.py:34) and 'none' have kernels.
    """

    def __init__(self, act_type, **kwargs):
        super().__init__()
        act_type = act_type.lower()
        if act_type not in _KNOWN_ACTS:
            raise NotImplementedError(f'Unsupport activation type: {act_type}')
        if act_type not in _FUSED_ACTS:
            raise NotImplementedError(f"activation '{act_type}' has no fused sm_100a kernel (only relu/none are reachable)")
        self.act_type = act_type
        self.relu = _FUSED_ACTS[act_type]


def _geom_of(conv):
    t = isinstance(conv, nn.ConvTranspose2d)
    return ConvGeom(conv.in_channels, conv.out_channels, conv.kernel_size[0], conv.kernel_size[1], conv.stride[0],
                    conv.dilation[0], conv.padding[0], conv.padding[1], transposed=t,
                    out_pad=conv.output_padding[0] if t else 0)


def _sync_group(bn):
    """The process group a BatchNorm synchronises its batch statistics over (parallel_model with synBN=True), or None."""
    ref = getattr(bn, '_msx_sync_group', None)
    if ref is None or not bn.training:
        return None
    import torch.distributed as dist
    if not dist.is_initialized():
        return None
    g = ref.group if ref.group is not None else dist.group.WORLD
    return g if dist.get_world_size(g) > 1 else None


def conv_bn_act(x, conv, bn, relu, slot=None):
    """Shared forward of ConvBNAct / DeConvBNAct."""
    if not hasattr(conv, '_msx_geom'):
        conv._msx_geom = _geom_of(conv)
    momentum = 0.1 if bn.momentum is None else bn.momentum
    group = _sync_group(bn)
    if group is not None:                       # SyncBatchNorm: conv, then statistics over all ranks (ops.SyncBNActFn)
        y = ops.Conv2dFn.apply(x.t, conv.weight, conv._msx_geom, conv.bias)
        z = ops.SyncBNActFn.apply(y, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.num_batches_tracked,
                                  conv.out_channels, bn.eps, momentum, relu, slot, group)
        return Act(z, conv.out_channels)
    # inference: eval mode and nothing that could be differentiated -> BN + ReLU run in the conv epilogue
    infer = not bn.training and not (torch.is_grad_enabled() and (x.t.requires_grad or conv.weight.requires_grad or bn.weight.requires_grad
                                                                  or bn.bias.requires_grad or (conv.bias is not None and conv.bias.requires_grad)))
    z = ops.ConvBNActFn.apply(x.t, conv.weight, conv.bias, bn.weight, bn.bias, bn.running_mean, bn.running_var,
                              bn.num_batches_tracked, conv._msx_geom, bn.eps, momentum, bn.training, relu, slot, infer)
    return Act(z, conv.out_channels)


def conv_plain(x, conv):
    if not hasattr(conv, '_msx_geom'):
        conv._msx_geom = _geom_of(conv)
    return Act(ops.Conv2dFn.apply(x.t, conv.weight, conv._msx_geom), conv.out_channels)


def sum_bn_act(xs, bn, relu=1, up2x=False, slot=None):
    """(sum of Acts) -> BN -> ReLU; xs[0] may be at half resolution when up2x."""
    c = xs[0].c
    momentum = 0.1 if bn.momentum is None else bn.momentum
    group = _sync_group(bn)
    if group is not None:
        s = ops.SumFn.apply(c, up2x, *[a.t for a in xs]) if (len(xs) > 1 or up2x) else xs[0].t
        return Act(ops.SyncBNActFn.apply(s, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.num_batches_tracked,
                                         c, bn.eps, momentum, relu, slot, group), c)
    infer = not bn.training and not (torch.is_grad_enabled() and (bn.weight.requires_grad or bn.bias.requires_grad
                                                                  or any(a.t.requires_grad for a in xs)))
    z = ops.SumBNActFn.apply(bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.num_batches_tracked, c, bn.eps,
                             momentum, bn.training, relu, up2x, slot, infer, *[a.t for a in xs])
    return Act(z, c)


def fanout(x, k):
    return [Act(t, x.c) for t in ops.FanoutFn.apply(x.t, k, x.c)]


def add(xs):
    return Act(ops.AddFn.apply(xs[0].c, *[a.t for a in xs]), xs[0].c)


class ConvBNAct(nn.Sequential):
    """Conv2d(bias=False) -> BatchNorm2d -> act; pad = (k-1)//2*dilation per axis (models/modules.py:73-85)."""

    def __init__(self, in_channels, out_channels, kernel_size=3, stride=1, dilation=1, groups=1,
                 bias=False, act_type='relu', **kwargs):
        if groups != 1:
            raise NotImplementedError('grouped convolutions are unreachable in the reference models')
        if isinstance(kernel_size, (list, tuple)):
            padding = ((kernel_size[0] - 1) // 2 * dilation, (kernel_size[1] - 1) // 2 * dilation)
        else:
            padding = (kernel_size - 1) // 2 * dilation
        super().__init__(nn.Conv2d(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, bias),
                         nn.BatchNorm2d(out_channels),
                         Activation(act_type))

    def forward(self, x, slot=None):
        return conv_bn_act(x, self[0], self[1], self[2].relu, slot)


class DeConvBNAct(nn.Module):
    """ConvTranspose2d(k=2s-1, stride=s, pad=(k-1)//2, output_padding=s-1, bias) -> BN -> act (models/modules.py:89-108)."""

    def __init__(self, in_channels, out_channels, scale_factor=2, kernel_size=None, padding=None, act_type='relu', **kwargs):
        super().__init__()
        if kernel_size is None:
            kernel_size = 2 * scale_factor - 1
        if padding is None:
            padding = (kernel_size - 1) // 2
        self.up_conv = nn.Sequential(
            nn.ConvTranspose2d(in_channels, out_channels, kernel_size=kernel_size, stride=scale_factor,
                               padding=padding, output_padding=scale_factor - 1),
            nn.BatchNorm2d(out_channels),
            Activation(act_type))

    def forward(self, x, slot=None):
        return conv_bn_act(x, self.up_conv[0], self.up_conv[1], self.up_conv[2].relu, slot)


def compute_dtype(model):
    """bf16 under torch.autocast (the reference's AMP switch, core/seg_trainer.py:61) or when the
    model's ``compute_dtype`` attribute says so; fp32 otherwise."""
    forced = getattr(model, 'compute_dtype', None)
    if forced is not None:
        return forced
    if torch.is_autocast_enabled():
        return torch.bfloat16
    return torch.float32
