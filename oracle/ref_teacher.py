"""TEST INFRASTRUCTURE ONLY (oracle): CPU restatement of the distillation teacher, ``smp.Unet(encoder_name=<resnet>,
encoder_weights=None, in_channels, classes)`` as the reference builds it in models/__init__.py:43-58 and calls it in
core/seg_trainer.py:70-74.

PARITY UNPINNED: the arithmetic lives in two dependencies that are not vendored in /root/reference and not
installed here - segmentation-models-pytorch 0.3.2 (requirements.txt:56) and its torchvision ResNet encoder.  The
encoder below IS torchvision's own ``resnet*`` (installed, 0.26) minus ``avgpool``/``fc``; the decoder and head are
restated from smp 0.3.2's published ``UnetDecoder`` / ``SegmentationHead``:
  features = [x, relu(bn1(conv1 x)), layer1(maxpool .), layer2, layer3, layer4]
  decoder blocks (in + skip -> out): nearest x2 -> cat([x, skip], 1) -> [Conv3x3(no bias) + BN + ReLU] x 2,
  decoder_channels (256, 128, 64, 32, 16), last block has no skip; head = Conv2d(16, classes, 3, padding=1).
Known answer it is pinned by: the resnet18 variant has 14,328,354 parameters = README.md:113's 14.33 M.
state_dict keys follow smp (``encoder.*``, ``decoder.blocks.i.conv{1,2}.{0,1}.*``, ``segmentation_head.0.*``).
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

ENCODER_CHANNELS = {
    'resnet18': (3, 64, 64, 128, 256, 512), 'resnet34': (3, 64, 64, 128, 256, 512),
    'resnet50': (3, 64, 256, 512, 1024, 2048), 'resnet101': (3, 64, 256, 512, 1024, 2048),
    'resnet152': (3, 64, 256, 512, 1024, 2048),
}
DECODER_CHANNELS = (256, 128, 64, 32, 16)


def _conv_bn_relu(cin, cout):
    return nn.Sequential(nn.Conv2d(cin, cout, 3, padding=1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(inplace=True))


class DecoderBlockRef(nn.Module):
    def __init__(self, cin, cskip, cout):
        super().__init__()
        self.conv1 = _conv_bn_relu(cin + cskip, cout)
        self.attention1 = nn.Identity()
        self.conv2 = _conv_bn_relu(cout, cout)
        self.attention2 = nn.Identity()

    def forward(self, x, skip=None):
        x = F.interpolate(x, scale_factor=2, mode='nearest')
        if skip is not None:
            x = torch.cat([x, skip], dim=1)
        return self.conv2(self.conv1(x))


class UnetDecoderRef(nn.Module):
    def __init__(self, encoder_channels):
        super().__init__()
        enc = encoder_channels[1:][::-1]
        ins = [enc[0]] + list(DECODER_CHANNELS[:-1])
        skips = list(enc[1:]) + [0]
        self.center = nn.Identity()
        self.blocks = nn.ModuleList(DecoderBlockRef(i, s, o) for i, s, o in zip(ins, skips, DECODER_CHANNELS))

    def forward(self, *features):
        features = features[1:][::-1]
        x, skips = self.center(features[0]), features[1:]
        for i, blk in enumerate(self.blocks):
            x = blk(x, skips[i] if i < len(skips) else None)
        return x


class SmpUnetRef(nn.Module):
    def __init__(self, encoder_name='resnet101', in_channels=3, classes=1):
        super().__init__()
        import torchvision
        if encoder_name not in ENCODER_CHANNELS:
            raise ValueError(f'unsupported encoder {encoder_name}')
        enc = getattr(torchvision.models, encoder_name)(weights=None)
        del enc.fc, enc.avgpool
        if in_channels != 3:
            enc.conv1 = nn.Conv2d(in_channels, 64, 7, 2, 3, bias=False)
        self.encoder = enc
        self.decoder = UnetDecoderRef(ENCODER_CHANNELS[encoder_name])
        self.segmentation_head = nn.Sequential(nn.Conv2d(DECODER_CHANNELS[-1], classes, 3, padding=1), nn.Identity(), nn.Identity())
        for m in self.decoder.modules():                       # smp.base.initialization.initialize_decoder
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_uniform_(m.weight, mode='fan_in', nonlinearity='relu')
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)
        nn.init.xavier_uniform_(self.segmentation_head[0].weight)      # initialize_head
        nn.init.constant_(self.segmentation_head[0].bias, 0)

    def features(self, x):
        e = self.encoder
        f = [x]
        x = e.relu(e.bn1(e.conv1(x)))
        f.append(x)
        x = e.layer1(e.maxpool(x))
        f.append(x)
        for layer in (e.layer2, e.layer3, e.layer4):
            x = layer(x)
            f.append(x)
        return f

    def forward(self, x):
        if x.shape[2] % 32 or x.shape[3] % 32:
            raise RuntimeError('input height/width must be divisible by 32')
        return self.segmentation_head(self.decoder(*self.features(x)))


def randomize_bn(model, seed=0):
    """Give every BatchNorm non-trivial affine parameters and running statistics (a fresh model has 1/0/0/1, which
    would hide errors in the eval-mode affine)."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for m in model.modules():
            if isinstance(m, nn.BatchNorm2d):
                m.weight.copy_(torch.rand(m.weight.shape, generator=g) * 0.5 + 0.75)
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
                m.running_mean.copy_(torch.randn(m.running_mean.shape, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(m.running_var.shape, generator=g) * 0.5 + 0.75)
