"""The distillation teacher: ``smp.Unet(encoder_name=<resnet>, encoder_weights=None, in_channels, classes)`` as the
reference constructs it (models/__init__.py:43-58) and calls it, frozen, under ``torch.no_grad()`` every KD step
(core/seg_trainer.py:70-74).  Inference only - the reference never trains it.

The arithmetic is segmentation-models-pytorch 0.3.2's ``Unet`` over torchvision's ResNet (both un-vendored
dependencies of the reference, requirements.txt:56); module and ``state_dict`` names follow them so that a
reference teacher checkpoint loads unchanged: ``encoder.conv1/bn1/layer{1..4}.{i}.conv{1,2,3}/bn{1,2,3}/downsample.{0,1}``,
``decoder.blocks.{i}.conv{1,2}.{0,1}``, ``segmentation_head.0``.

Everything runs on the library's inference path: eval-mode BN and ReLU in the conv epilogues, residual
``relu(out + identity)`` and the decoder's nearest-2x upsample as one ``msx_sum_act`` launch each, and
``torch.cat([x, skip])`` as two producers writing channel slices of one buffer (no copy).
"""
import ctypes as C

import torch
import torch.nn as nn

from . import _lib as L
from . import ops, packplan
from ._lib import lib
from .modules import Act, compute_dtype, conv_bn_act
from .ops import OutSlot

ENCODERS = {
    # name: (block, layers, feature channels)
    'resnet18': ('basic', (2, 2, 2, 2), (3, 64, 64, 128, 256, 512)),
    'resnet34': ('basic', (3, 4, 6, 3), (3, 64, 64, 128, 256, 512)),
    'resnet50': ('bottleneck', (3, 4, 6, 3), (3, 64, 256, 512, 1024, 2048)),
    'resnet101': ('bottleneck', (3, 4, 23, 3), (3, 64, 256, 512, 1024, 2048)),
    'resnet152': ('bottleneck', (3, 8, 36, 3), (3, 64, 256, 512, 1024, 2048)),
}
DECODER_CHANNELS = (256, 128, 64, 32, 16)


def sum_act(xs, c, relu=False, up2x=False, slot=None):
    """[relu](sum of Acts), xs[0] optionally nearest-2x upsampled first; forward only (msx_sum_act)."""
    ref = xs[1] if (up2x and len(xs) > 1) else xs[0]
    n, h, w, cp = ref.t.shape
    if up2x and len(xs) == 1:
        h, w = 2 * h, 2 * w
    out = ops._out_tensor(slot, n, h, w, cp, ref.t.dtype, ref.t.device)
    ptrs, lds = ops._ptr_arrays([a.t for a in xs])
    L.check(lib.msx_sum_act(L.dtype_code(out.dtype), len(xs), ptrs, lds, int(up2x), n, h, w, c, int(relu),
                            L.ptr(out), ops.ld_of(out), L.stream()))
    return Act(out, c)


def _strided_1x1_as_2x2(conv):
    """A 1x1 stride-2 conv reads pixel (2y, 2x): the same thing as a 2x2 stride-2 conv whose other three taps are
    zero, which is a geometry the tcgen05 engine has (the 2x2 s2 conv of DUCK-Net's down path)."""
    w = conv.weight
    cache = conv.__dict__.get('_msx_w2x2')
    if cache is None or cache[0] != (w.data_ptr(), w._version):
        w2 = torch.zeros(w.shape[0], w.shape[1], 2, 2, dtype=torch.float32, device=w.device)
        w2[:, :, 0, 0] = w.detach()[:, :, 0, 0]
        geom = ops.ConvGeom(conv.in_channels, conv.out_channels, 2, 2, 2, 1, 0, 0)
        cache = conv.__dict__['_msx_w2x2'] = ((w.data_ptr(), w._version), w2, geom)
    return cache[1], cache[2]


def _conv_bn(x, conv, bn, relu, slot=None):
    if conv.kernel_size == (1, 1) and conv.stride == (2, 2):
        w2, geom = _strided_1x1_as_2x2(conv)
        z = ops.ConvBNActFn.apply(x.t, w2, None, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.num_batches_tracked,
                                  geom, bn.eps, 0.1, False, relu, slot, True)
        return Act(z, conv.out_channels)
    return conv_bn_act(x, conv, bn, relu, slot)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin, planes, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, planes, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.downsample = None
        if stride != 1 or cin != planes:
            self.downsample = nn.Sequential(nn.Conv2d(cin, planes, 1, stride, bias=False), nn.BatchNorm2d(planes))

    def forward(self, x, slot=None):
        o = _conv_bn(x, self.conv1, self.bn1, 1)
        o = _conv_bn(o, self.conv2, self.bn2, 0)
        idt = x if self.downsample is None else _conv_bn(x, self.downsample[0], self.downsample[1], 0)
        return sum_act([o, idt], o.c, relu=True, slot=slot)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, planes, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, planes, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride, 1, bias=False)       # torchvision's v1.5: stride on the 3x3
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, planes * 4, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(planes * 4)
        self.downsample = None
        if stride != 1 or cin != planes * 4:
            self.downsample = nn.Sequential(nn.Conv2d(cin, planes * 4, 1, stride, bias=False), nn.BatchNorm2d(planes * 4))

    def forward(self, x, slot=None):
        o = _conv_bn(x, self.conv1, self.bn1, 1)
        o = _conv_bn(o, self.conv2, self.bn2, 1)
        o = _conv_bn(o, self.conv3, self.bn3, 0)
        idt = x if self.downsample is None else _conv_bn(x, self.downsample[0], self.downsample[1], 0)
        return sum_act([o, idt], o.c, relu=True, slot=slot)


class ResNetEncoder(nn.Module):
    def __init__(self, name, in_channels):
        super().__init__()
        kind, layers, self.out_channels = ENCODERS[name]
        block = Bottleneck if kind == 'bottleneck' else BasicBlock
        self.conv1 = nn.Conv2d(in_channels, 64, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        cin = 64
        for i, (planes, nblk) in enumerate(zip((64, 128, 256, 512), layers)):
            blocks = []
            for b in range(nblk):
                blocks.append(block(cin, planes, (1 if i == 0 else 2) if b == 0 else 1))
                cin = planes * block.expansion
            setattr(self, f'layer{i + 1}', nn.Sequential(*blocks))


class DecoderBlock(nn.Module):
    def __init__(self, cin, cskip, cout):
        super().__init__()
        self.conv1 = nn.Sequential(nn.Conv2d(cin + cskip, cout, 3, padding=1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(inplace=True))
        self.attention1 = nn.Identity()
        self.conv2 = nn.Sequential(nn.Conv2d(cout, cout, 3, padding=1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(inplace=True))
        self.attention2 = nn.Identity()
        self.cin, self.cskip = cin, cskip


class UnetDecoder(nn.Module):
    def __init__(self, encoder_channels):
        super().__init__()
        enc = encoder_channels[1:][::-1]
        ins = [enc[0]] + list(DECODER_CHANNELS[:-1])
        skips = list(enc[1:]) + [0]
        self.center = nn.Identity()
        self.blocks = nn.ModuleList(DecoderBlock(i, s, o) for i, s, o in zip(ins, skips, DECODER_CHANNELS))


class SmpUnet(nn.Module):
    """Inference-only ``smp.Unet`` over a ResNet encoder: ``forward(images NCHW fp32) -> logits NCHW fp32``."""

    def __init__(self, encoder_name='resnet101', encoder_weights=None, in_channels=3, classes=1):
        super().__init__()
        if encoder_name not in ENCODERS:
            raise ValueError(f'Unsupported teacher encoder: {encoder_name} (ResNet encoders only)')
        if encoder_weights is not None:
            raise ValueError('pretrained encoder weights cannot be downloaded here: load a checkpoint instead')
        self.encoder = ResNetEncoder(encoder_name, in_channels)
        self.decoder = UnetDecoder(self.encoder.out_channels)
        self.segmentation_head = nn.Sequential(nn.Conv2d(DECODER_CHANNELS[-1], classes, 3, padding=1), nn.Identity(), nn.Identity())
        for m in self.decoder.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_uniform_(m.weight, mode='fan_in', nonlinearity='relu')
        nn.init.xavier_uniform_(self.segmentation_head[0].weight)
        nn.init.constant_(self.segmentation_head[0].bias, 0)
        self.eval()

    def forward(self, x):
        if self.training:
            raise NotImplementedError('the smp teacher is inference-only (the reference never trains it): call .eval()')
        if x.shape[2] % 32 or x.shape[3] % 32:
            raise RuntimeError('input height/width must be divisible by 32')
        dt = compute_dtype(self)
        with torch.no_grad(), packplan.active(self, dt, x.device, skip=(self.segmentation_head[0],)):
            return self._forward(x, dt)

    def _forward(self, x, dt):
        e, blocks = self.encoder, self.decoder.blocks
        n, _, h, w = x.shape
        dev = x.device
        # concat buffers of decoder blocks 0..3: [ upsampled x | skip ]; the encoder writes the skips in place
        bufs = [torch.empty(n, h >> (4 - i), w >> (4 - i), b.cin + b.cskip, dtype=dt, device=dev) for i, b in enumerate(blocks[:4])]
        for b in blocks[:4]:
            if b.cin % 8 or b.cskip % 8:
                raise L.MsxError('decoder channel counts must be multiples of 8 for the zero-copy concat')
        a = ops.to_nhwc(x, dt)
        f1 = conv_bn_act(a, e.conv1, e.bn1, 1, OutSlot(bufs[3], blocks[3].cin))                # 64 @ H/2
        a = Act(ops.MaxPoolFn.apply(f1.t, f1.c), f1.c)
        for li, buf_i in ((1, 2), (2, 1), (3, 0), (4, None)):
            layer = getattr(e, f'layer{li}')
            for bi, blk in enumerate(layer):
                last = bi == len(layer) - 1 and buf_i is not None
                a = blk(a, OutSlot(bufs[buf_i], blocks[buf_i].cin) if last else None)
        for i, b in enumerate(blocks):
            if i < 4:
                sum_act([a], a.c, up2x=True, slot=OutSlot(bufs[i], 0))                           # F.interpolate(nearest, x2)
                a = Act(bufs[i], b.cin + b.cskip)
            else:
                a = sum_act([a], a.c, up2x=True)
            a = conv_bn_act(a, b.conv1[0], b.conv1[1], 1)
            a = conv_bn_act(a, b.conv2[0], b.conv2[1], 1)
        head = self.segmentation_head[0]
        return _conv_bias_nchw(a, head)


def _conv_bias_nchw(x, conv):
    """segmentation_head: Conv2d(16, classes, 3, padding=1) with bias -> fp32 NCHW logits (forward only)."""
    if not hasattr(conv, '_msx_geom'):
        conv._msx_geom = ops.ConvGeom(conv.in_channels, conv.out_channels, conv.kernel_size[0], conv.kernel_size[1], 1, 1,
                                      conv.padding[0], conv.padding[1])
    g = conv._msx_geom
    n, h, w, _ = x.t.shape
    cp = ops.rup(g.cout)
    y = torch.empty(n, h, w, cp, dtype=x.t.dtype, device=x.t.device)
    d = ops._desc(g, x.t, cp, h, w)
    pk = L.scratch('pack', lib.msx_conv_packed_bytes(C.byref(d)), x.t.device)
    L.check(lib.msx_conv2d_fwd(C.byref(d), L.ptr(x.t), L.ptr(conv.weight), L.ptr(conv.bias), L.ptr(pk), L.ptr(y), L.stream()))
    return Act(y, g.cout).nchw()
