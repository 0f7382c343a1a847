"""UNet with the reference constructor, forward signature and state_dict (models/unet.py:14-77)."""
import torch
import torch.nn as nn

from . import ops, packplan
from .modules import Act, ConvBNAct, DeConvBNAct, compute_dtype, conv1x1


class ConvBlock(nn.Sequential):
    def __init__(self, in_channels, out_channels, act_type):
        super().__init__(ConvBNAct(in_channels, out_channels, 3, act_type=act_type),
                         ConvBNAct(out_channels, out_channels, 3, act_type=act_type))

    def forward(self, x, slot=None):
        return self[1](self[0](x), slot)


class DownsampleBlock(nn.Module):
    def __init__(self, in_channels, out_channels, act_type):
        super().__init__()
        self.conv = ConvBlock(in_channels, out_channels, act_type)
        self.pool = nn.MaxPool2d(3, 2, 1)            # parameter-free; kept for module-tree parity

    def forward(self, x, slot):
        residual = self.conv(x, slot)
        # residual feeds the pool and (through the concat buffer) the decoder
        r_pool, r_skip = (Act(t, residual.c) for t in ops.FanoutFn.apply(residual.t, 2, residual.c))
        pooled = Act(ops.MaxPoolFn.apply(r_pool.t, r_pool.c), residual.c)
        return pooled, r_skip


class UpsampleBlock(nn.Module):
    def __init__(self, in_channels, out_channels, act_type):
        super().__init__()
        self.up = DeConvBNAct(in_channels, out_channels, act_type=act_type)
        self.conv = ConvBlock(in_channels, out_channels, act_type)

    def forward(self, x, residual, buf):
        c = residual.c
        up = self.up(x, ops.OutSlot(buf, 0))
        cat = Act(ops.ConcatFn.apply(up.t, residual.t, buf, c), 2 * c)      # torch.cat([x, residual], 1)
        return self.conv(cat)


class UNet(nn.Module):
    _features_only = False      # features(): stop before seg_head (fused head + argmax + confusion matrix in eval)

    def __init__(self, num_class=1, n_channel=3, base_channel=64, act_type='relu'):
        super().__init__()
        if base_channel % 8:
            raise NotImplementedError('zero-copy concat needs base_channel % 8 == 0 (reference configs use 32 / 64)')
        b = base_channel
        self.down_stage1 = DownsampleBlock(n_channel, b, act_type)
        self.down_stage2 = DownsampleBlock(b, b * 2, act_type)
        self.down_stage3 = DownsampleBlock(b * 2, b * 4, act_type)
        self.down_stage4 = DownsampleBlock(b * 4, b * 8, act_type)
        self.mid_stage = ConvBlock(b * 8, b * 16, act_type)
        self.up_stage4 = UpsampleBlock(b * 16, b * 8, act_type)
        self.up_stage3 = UpsampleBlock(b * 8, b * 4, act_type)
        self.up_stage2 = UpsampleBlock(b * 4, b * 2, act_type)
        self.up_stage1 = UpsampleBlock(b * 2, b, act_type)
        self.seg_head = conv1x1(b, num_class)
        self.base_channel = b

    def forward(self, x):
        if x.shape[2] % 16 or x.shape[3] % 16:
            raise ValueError('UNet input height/width must be multiples of 16')
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
        a = ops.to_nhwc(x, dt)
        n, _, h, w = x.shape
        skips, bufs = [], []
        for i, stage in enumerate((self.down_stage1, self.down_stage2, self.down_stage3, self.down_stage4)):
            c = self.base_channel << i
            buf = torch.empty(n, h >> i, w >> i, 2 * c, dtype=dt, device=x.device)   # [up | residual]
            a, r = stage(a, ops.OutSlot(buf, c))
            skips.append(r)
            bufs.append(buf)
        a = self.mid_stage(a)
        for i, stage in zip((3, 2, 1, 0), (self.up_stage4, self.up_stage3, self.up_stage2, self.up_stage1)):
            a = stage(a, skips[i], bufs[i])
        if self._features_only:
            return a
        return ops.HeadFn.apply(a.t, self.seg_head.weight, a.c)
