"""DUCK-Net with the reference constructor, forward signature and state_dict (models/ducknet.py:15-179)."""
import torch.nn as nn

from . import ops, packplan
from .modules import (Act, Activation, ConvBNAct, add, compute_dtype, conv1x1, conv_plain, fanout, sum_bn_act)


def _bn_act(channels, act_type):
    return nn.Sequential(nn.BatchNorm2d(channels), Activation(act_type))


class ResidualBlock(nn.Module):
    """conv1x1(x) + ConvBNAct3x3(ConvBNAct3x3(x)) -> BN -> act (models/ducknet.py:90-110)."""

    def __init__(self, in_channels, out_channels, act_type):
        super().__init__()
        self.upper_branch = conv1x1(in_channels, out_channels)
        self.lower_branch = nn.Sequential(ConvBNAct(in_channels, out_channels, 3, act_type=act_type),
                                          ConvBNAct(out_channels, out_channels, 3, act_type=act_type))
        self.bn = _bn_act(out_channels, act_type)

    def forward(self, x_up, x_low=None):
        if x_low is None:
            x_up, x_low = fanout(x_up, 2)
        up, low = ops.run_branches([(lambda: conv_plain(x_up, self.upper_branch), [x_up]),
                                    (lambda: self.lower_branch[1](self.lower_branch[0](x_low)), [x_low])], main_job=1)
        return sum_bn_act([up, low], self.bn[0], self.bn[1].relu)        # add fused into the BN statistics pass


class MidscopeBlock(nn.Sequential):
    def __init__(self, in_channels, out_channels, act_type):
        super().__init__(ConvBNAct(in_channels, out_channels, 3, act_type=act_type),
                         ConvBNAct(out_channels, out_channels, 3, dilation=2, act_type=act_type))


class WidescopeBlock(nn.Sequential):
    def __init__(self, in_channels, out_channels, act_type):
        super().__init__(ConvBNAct(in_channels, out_channels, 3, act_type=act_type),
                         ConvBNAct(out_channels, out_channels, 3, dilation=2, act_type=act_type),
                         ConvBNAct(out_channels, out_channels, 3, dilation=3, act_type=act_type))


class SeparatedBlock(nn.Sequential):
    """Two DENSE convs, 1xk then kx1 (not depth-wise)."""

    def __init__(self, in_channels, out_channels, filter_size, act_type):
        super().__init__(ConvBNAct(in_channels, out_channels, (1, filter_size), act_type=act_type),
                         ConvBNAct(out_channels, out_channels, (filter_size, 1), act_type=act_type))


def _chain(seq, x):
    for m in seq:
        x = m(x)
    return x


class DUCK(nn.Module):
    def __init__(self, in_channels, out_channels, act_type, filter_size=6 + 1):
        super().__init__()
        self.in_bn = _bn_act(in_channels, act_type)
        self.branch1 = WidescopeBlock(in_channels, out_channels, act_type)
        self.branch2 = MidscopeBlock(in_channels, out_channels, act_type)
        self.branch3 = ResidualBlock(in_channels, out_channels, act_type)
        self.branch4 = nn.Sequential(ResidualBlock(in_channels, out_channels, act_type),
                                     ResidualBlock(out_channels, out_channels, act_type))
        self.branch5 = nn.Sequential(ResidualBlock(in_channels, out_channels, act_type),
                                     ResidualBlock(out_channels, out_channels, act_type),
                                     ResidualBlock(out_channels, out_channels, act_type))
        self.branch6 = SeparatedBlock(in_channels, out_channels, filter_size, act_type)
        self.out_bn = _bn_act(out_channels, act_type)

    def forward(self, xs, up2x=False):
        """xs: list of Acts whose sum is the block input (the reference adds them just before the
        call, models/ducknet.py:39-43,84); with up2x, xs[0] is nearest-upsampled 2x first."""
        x = sum_bn_act(xs, self.in_bn[0], self.in_bn[1].relu, up2x=up2x)
        f = fanout(x, 9)                                    # nine first-layer convs read the same tensor
        outs = ops.run_branches([
            (lambda: _chain(self.branch1, f[0]), f[0:1]),
            (lambda: _chain(self.branch2, f[1]), f[1:2]),
            (lambda: self.branch3(f[2], f[3]), f[2:4]),
            (lambda: self.branch4[1](self.branch4[0](f[4], f[5])), f[4:6]),
            (lambda: self.branch5[2](self.branch5[1](self.branch5[0](f[6], f[7]))), f[6:8]),     # longest chain: stays on the caller's stream
            (lambda: _chain(self.branch6, f[8]), f[8:9]),
        ], main_job=4)
        return sum_bn_act(outs, self.out_bn[0], self.out_bn[1].relu)


class DownsampleBlock(nn.Module):
    def __init__(self, in_channels, out_channels, act_type, fuse_channels=None):
        super().__init__()
        fuse_channels = in_channels if fuse_channels is None else fuse_channels
        self.duck = DUCK(in_channels, fuse_channels, act_type)
        self.conv1 = ConvBNAct(fuse_channels, out_channels, 3, 2, act_type=act_type)
        self.conv2 = ConvBNAct(in_channels, out_channels, 2, 2, act_type=act_type)

    def forward(self, x1_parts, x2):
        def main_path():
            skip = self.duck(x1_parts)
            s_conv, s_skip = fanout(skip, 2)
            return self.conv1(s_conv), s_skip
        x2n, (x1, s_skip) = ops.run_branches([(lambda: self.conv2(x2), [x2]), (main_path, [])], main_job=1)
        return x1, s_skip, x2n


class UpsampleBlock(nn.Module):
    def __init__(self, in_channels, out_channels, act_type):
        super().__init__()
        self.duck = DUCK(in_channels, out_channels, act_type)

    def forward(self, x, residual):
        if residual.t.shape[1] != 2 * x.t.shape[1] or residual.t.shape[2] != 2 * x.t.shape[2]:
            raise ValueError('DuckNet upsampling path expects an exact 2x size ratio (input H, W multiples of 32)')
        return self.duck([x, residual], up2x=True)      # F.interpolate(nearest) ; x += residual ; duck


class DuckNet(nn.Module):
    _features_only = False      # features(): stop before seg_head (fused head + argmax + confusion matrix in eval)

    def __init__(self, num_class=1, n_channel=3, base_channel=17, act_type='relu'):
        super().__init__()
        b = base_channel
        self.down_stage1 = DownsampleBlock(n_channel, b * 2, act_type, fuse_channels=b)
        self.down_stage2 = DownsampleBlock(b * 2, b * 4, act_type)
        self.down_stage3 = DownsampleBlock(b * 4, b * 8, act_type)
        self.down_stage4 = DownsampleBlock(b * 8, b * 16, act_type)
        self.down_stage5 = DownsampleBlock(b * 16, b * 32, act_type)
        self.mid_stage = nn.Sequential(ResidualBlock(b * 32, b * 32, act_type),
                                       ResidualBlock(b * 32, b * 32, act_type),
                                       ResidualBlock(b * 32, b * 16, act_type),
                                       ResidualBlock(b * 16, b * 16, act_type))
        self.up_stage5 = UpsampleBlock(b * 16, b * 8, act_type)
        self.up_stage4 = UpsampleBlock(b * 8, b * 4, act_type)
        self.up_stage3 = UpsampleBlock(b * 4, b * 2, act_type)
        self.up_stage2 = UpsampleBlock(b * 2, b, act_type)
        self.up_stage1 = UpsampleBlock(b, b, act_type)
        self.seg_head = conv1x1(b, num_class)

    def forward(self, x):
        if x.shape[2] % 32 or x.shape[3] % 32:
            raise ValueError('DuckNet input height/width must be multiples of 32')
        dt = compute_dtype(self)
        with packplan.active(self, dt, x.device, skip=(self.seg_head,)):     # all conv weights packed by one launch
            return self._forward(x, dt)

    def features(self, x):
        """The activation that feeds seg_head, as an NHWC ``Act`` (no logits are formed): input of
        ``metric.update_from_features`` / ``metrics.eval_sweep``."""
        self._features_only = True
        try:
            return self.forward(x)
        finally:
            self._features_only = False

    def _forward(self, x, dt):
        img = ops.to_nhwc(x, dt)
        x1, s1, xc = self.down_stage1([img], img)
        skips = [s1]
        for stage in (self.down_stage2, self.down_stage3, self.down_stage4, self.down_stage5):
            xa, xb = fanout(xc, 2)                       # x feeds the add and the next conv2
            x1, s, xc = stage([x1, xa], xb)
            skips.append(s)
        a = add([x1, xc])
        for blk in self.mid_stage:
            a = blk(a)
        for stage, s in zip((self.up_stage5, self.up_stage4, self.up_stage3, self.up_stage2, self.up_stage1), reversed(skips)):
            a = stage(a, s)
        if self._features_only:
            return a
        return ops.HeadFn.apply(a.t, self.seg_head.weight, a.c)
