"""Oracle: functional CPU restatement of the reference UNet / DuckNet (TEST INFRASTRUCTURE ONLY).

The networks are evaluated straight from a reference-layout ``state_dict`` (same keys and
shapes as ``models/unet.py`` / ``models/ducknet.py`` produce), so that weights can be moved
between the reference, this oracle and the CUDA product without renaming anything.

Everything is expressed with explicit formulas on top of ``F.conv2d`` / ``F.conv_transpose2d``
(train-mode BatchNorm is written out; ``fused_bn=True`` switches to ``F.batch_norm`` - the ATen
kernel the reference's ``nn.BatchNorm2d`` calls - which is what the timed CPU baseline uses).
"""
from collections import OrderedDict

import numpy as np
import torch
import torch.nn.functional as F

BN_EPS = 1e-5        # nn.BatchNorm2d default, models/modules.py:83
BN_MOMENTUM = 0.1


# --------------------------------------------------------------------------------------------
# state_dict layouts (reference key order == registration order)
# --------------------------------------------------------------------------------------------
def _conv_bn_keys(sd, prefix, cin, cout, kh, kw):
    sd[prefix + '.0.weight'] = (cout, cin, kh, kw)
    _bn_keys(sd, prefix + '.1', cout)


def _bn_keys(sd, prefix, c):
    sd[prefix + '.weight'] = (c,)
    sd[prefix + '.bias'] = (c,)
    sd[prefix + '.running_mean'] = (c,)
    sd[prefix + '.running_var'] = (c,)
    sd[prefix + '.num_batches_tracked'] = ()


def unet_layout(num_class=1, n_channel=3, base_channel=64):
    """Key -> shape, in the order models/unet.py:15-27 registers them."""
    b = base_channel
    sd = OrderedDict()

    def conv_block(prefix, cin, cout):           # models/unet.py:71-77
        _conv_bn_keys(sd, prefix + '.0', cin, cout, 3, 3)
        _conv_bn_keys(sd, prefix + '.1', cout, cout, 3, 3)

    chans = [n_channel, b, 2 * b, 4 * b, 8 * b]
    for i in range(4):
        conv_block(f'down_stage{i + 1}.conv', chans[i], chans[i + 1])
    conv_block('mid_stage', 8 * b, 16 * b)
    for i, (cin, cout) in zip((4, 3, 2, 1), ((16 * b, 8 * b), (8 * b, 4 * b), (4 * b, 2 * b), (2 * b, b))):
        p = f'up_stage{i}'
        sd[p + '.up.up_conv.0.weight'] = (cin, cout, 3, 3)      # ConvTranspose2d: [Cin, Cout, k, k]
        sd[p + '.up.up_conv.0.bias'] = (cout,)
        _bn_keys(sd, p + '.up.up_conv.1', cout)
        conv_block(p + '.conv', cin, cout)
    sd['seg_head.weight'] = (num_class, b, 1, 1)
    return sd


def _resblock_keys(sd, p, cin, cout):            # models/ducknet.py:90-101
    sd[p + '.upper_branch.weight'] = (cout, cin, 1, 1)
    _conv_bn_keys(sd, p + '.lower_branch.0', cin, cout, 3, 3)
    _conv_bn_keys(sd, p + '.lower_branch.1', cout, cout, 3, 3)
    _bn_keys(sd, p + '.bn.0', cout)


def _duck_keys(sd, p, cin, cout, fs=7):          # models/ducknet.py:113-139
    _bn_keys(sd, p + '.in_bn.0', cin)
    _conv_bn_keys(sd, p + '.branch1.0', cin, cout, 3, 3)
    _conv_bn_keys(sd, p + '.branch1.1', cout, cout, 3, 3)
    _conv_bn_keys(sd, p + '.branch1.2', cout, cout, 3, 3)
    _conv_bn_keys(sd, p + '.branch2.0', cin, cout, 3, 3)
    _conv_bn_keys(sd, p + '.branch2.1', cout, cout, 3, 3)
    _resblock_keys(sd, p + '.branch3', cin, cout)
    _resblock_keys(sd, p + '.branch4.0', cin, cout)
    _resblock_keys(sd, p + '.branch4.1', cout, cout)
    _resblock_keys(sd, p + '.branch5.0', cin, cout)
    _resblock_keys(sd, p + '.branch5.1', cout, cout)
    _resblock_keys(sd, p + '.branch5.2', cout, cout)
    _conv_bn_keys(sd, p + '.branch6.0', cin, cout, 1, fs)
    _conv_bn_keys(sd, p + '.branch6.1', cout, cout, fs, 1)
    _bn_keys(sd, p + '.out_bn.0', cout)


def ducknet_layout(num_class=1, n_channel=3, base_channel=17):
    """Key -> shape, in the order models/ducknet.py:16-35 registers them."""
    b = base_channel
    sd = OrderedDict()
    downs = [(n_channel, 2 * b, b), (2 * b, 4 * b, 2 * b), (4 * b, 8 * b, 4 * b),
             (8 * b, 16 * b, 8 * b), (16 * b, 32 * b, 16 * b)]
    for i, (cin, cout, fuse) in enumerate(downs):
        p = f'down_stage{i + 1}'
        _duck_keys(sd, p + '.duck', cin, fuse)
        _conv_bn_keys(sd, p + '.conv1', fuse, cout, 3, 3)
        _conv_bn_keys(sd, p + '.conv2', cin, cout, 2, 2)
    for j, (cin, cout) in enumerate(((32 * b, 32 * b), (32 * b, 32 * b), (32 * b, 16 * b), (16 * b, 16 * b))):
        _resblock_keys(sd, f'mid_stage.{j}', cin, cout)
    for i, (cin, cout) in zip((5, 4, 3, 2, 1), ((16 * b, 8 * b), (8 * b, 4 * b), (4 * b, 2 * b), (2 * b, b), (b, b))):
        _duck_keys(sd, f'up_stage{i}.duck', cin, cout)
    sd['seg_head.weight'] = (num_class, b, 1, 1)
    return sd


def model_layout(name, **kw):
    return {'unet': unet_layout, 'ducknet': ducknet_layout}[name](**kw)


def synth_state_dict(layout, seed, dtype=torch.float32):
    """Deterministic, torch-version-independent weights (numpy PCG64), used by the golden fixtures.

    Conv weights ~ U(-1/sqrt(fan_in), 1/sqrt(fan_in)) like PyTorch's default init; BN gamma in
    [0.5, 1.5], beta in [-0.2, 0.2] (non-trivial so that affine terms are exercised), running
    stats at their defaults.
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    sd = OrderedDict()
    for k, shp in layout.items():
        if k.endswith('num_batches_tracked'):
            sd[k] = torch.zeros((), dtype=torch.int64)
        elif k.endswith('running_mean'):
            sd[k] = torch.zeros(shp, dtype=dtype)
        elif k.endswith('running_var'):
            sd[k] = torch.ones(shp, dtype=dtype)
        elif len(shp) == 4:
            fan_in = shp[1] * shp[2] * shp[3]
            bound = 1.0 / np.sqrt(fan_in)
            sd[k] = torch.from_numpy(rng.uniform(-bound, bound, size=shp)).to(dtype)
        elif k.endswith('.weight'):                                    # BN gamma (1-d '.weight')
            sd[k] = torch.from_numpy(rng.uniform(0.5, 1.5, size=shp)).to(dtype)
        else:                                                          # BN beta / ConvT bias
            sd[k] = torch.from_numpy(rng.uniform(-0.2, 0.2, size=shp)).to(dtype)
    return sd


# --------------------------------------------------------------------------------------------
# functional blocks
# --------------------------------------------------------------------------------------------
class FnNet:
    """Evaluates the reference architectures from a state_dict.

    ``params`` maps reference keys to tensors.  Float tensors that should receive gradients must
    have ``requires_grad=True`` (see :func:`make_trainable`).  Running statistics are updated in
    place in training mode, as ``nn.BatchNorm2d`` does.
    """

    def __init__(self, name, params, training=True, fused_bn=False):
        self.name = name
        self.p = params
        self.training = training
        self.fused_bn = fused_bn

    # -- primitives ---------------------------------------------------------------------------
    def bn_relu(self, x, prefix, relu=True):
        """BatchNorm2d (+ReLU). SURVEY Appendix B; nn.BatchNorm2d at models/modules.py:83."""
        p = self.p
        w, b = p[prefix + '.weight'], p[prefix + '.bias']
        rm, rv = p[prefix + '.running_mean'], p[prefix + '.running_var']
        if self.fused_bn:
            if self.training:
                p[prefix + '.num_batches_tracked'] += 1
            y = F.batch_norm(x, rm, rv, w, b, self.training, BN_MOMENTUM, BN_EPS)
        elif self.training:
            m = x.numel() // x.shape[1]
            mean = x.mean(dim=(0, 2, 3))
            var = ((x - mean[None, :, None, None]) ** 2).mean(dim=(0, 2, 3))        # biased
            with torch.no_grad():
                rm.mul_(1 - BN_MOMENTUM).add_(BN_MOMENTUM * mean.detach().to(rm.dtype))
                rv.mul_(1 - BN_MOMENTUM).add_(BN_MOMENTUM * (var.detach() * (m / max(m - 1, 1))).to(rv.dtype))
                p[prefix + '.num_batches_tracked'] += 1
            y = (x - mean[None, :, None, None]) * torch.rsqrt(var + BN_EPS)[None, :, None, None]
            y = y * w[None, :, None, None] + b[None, :, None, None]
        else:
            y = (x - rm[None, :, None, None]) * torch.rsqrt(rv + BN_EPS)[None, :, None, None]
            y = y * w[None, :, None, None] + b[None, :, None, None]
        return torch.relu(y) if relu else y

    def conv_bn_act(self, x, prefix, stride=1, dilation=1):
        """ConvBNAct, models/modules.py:73-85: pad = (k-1)//2*dilation per axis, bias=False."""
        w = self.p[prefix + '.0.weight']
        kh, kw = w.shape[2], w.shape[3]
        pad = ((kh - 1) // 2 * dilation, (kw - 1) // 2 * dilation)
        y = F.conv2d(x, w, None, stride, pad, dilation)
        return self.bn_relu(y, prefix + '.1')

    def deconv_bn_act(self, x, prefix):
        """DeConvBNAct, models/modules.py:89-108: ConvTranspose2d(k=3, s=2, p=1, op=1, bias)."""
        y = F.conv_transpose2d(x, self.p[prefix + '.up_conv.0.weight'], self.p[prefix + '.up_conv.0.bias'],
                               stride=2, padding=1, output_padding=1)
        return self.bn_relu(y, prefix + '.up_conv.1')

    # -- UNet (models/unet.py) ----------------------------------------------------------------
    def _unet_block(self, x, prefix):
        x = self.conv_bn_act(x, prefix + '.0')
        return self.conv_bn_act(x, prefix + '.1')

    def unet(self, x):
        skips = []
        for i in (1, 2, 3, 4):                                   # models/unet.py:29-33, :51-55
            r = self._unet_block(x, f'down_stage{i}.conv')
            skips.append(r)
            x = F.max_pool2d(r, 3, 2, 1)
        x = self._unet_block(x, 'mid_stage')
        for i in (4, 3, 2, 1):                                   # models/unet.py:64-69
            x = self.deconv_bn_act(x, f'up_stage{i}.up')
            x = torch.cat([x, skips[i - 1]], dim=1)
            x = self._unet_block(x, f'up_stage{i}.conv')
        return F.conv2d(x, self.p['seg_head.weight'])

    # -- DuckNet (models/ducknet.py) ----------------------------------------------------------
    def resblock(self, x, p):                                    # models/ducknet.py:103-110
        up = F.conv2d(x, self.p[p + '.upper_branch.weight'])
        low = self.conv_bn_act(self.conv_bn_act(x, p + '.lower_branch.0'), p + '.lower_branch.1')
        return self.bn_relu(up + low, p + '.bn.0')

    def duck(self, x, p):                                        # models/ducknet.py:141-154
        x = self.bn_relu(x, p + '.in_bn.0')
        x1 = self.conv_bn_act(x, p + '.branch1.0')
        x1 = self.conv_bn_act(x1, p + '.branch1.1', dilation=2)
        x1 = self.conv_bn_act(x1, p + '.branch1.2', dilation=3)
        x2 = self.conv_bn_act(x, p + '.branch2.0')
        x2 = self.conv_bn_act(x2, p + '.branch2.1', dilation=2)
        x3 = self.resblock(x, p + '.branch3')
        x4 = self.resblock(self.resblock(x, p + '.branch4.0'), p + '.branch4.1')
        x5 = self.resblock(self.resblock(self.resblock(x, p + '.branch5.0'), p + '.branch5.1'), p + '.branch5.2')
        x6 = self.conv_bn_act(self.conv_bn_act(x, p + '.branch6.0'), p + '.branch6.1')
        return self.bn_relu(x1 + x2 + x3 + x4 + x5 + x6, p + '.out_bn.0')

    def _duck_down(self, x1, x2, p):                             # models/ducknet.py:63-72
        x2 = self.conv_bn_act(x1 if x2 is None else x2, p + '.conv2', stride=2)
        skip = self.duck(x1, p + '.duck')
        x1 = self.conv_bn_act(skip, p + '.conv1', stride=2)
        return x1, skip, x2

    def ducknet(self, x):                                        # models/ducknet.py:37-52
        x1, s1, x = self._duck_down(x, None, 'down_stage1')
        x2, s2, x = self._duck_down(x1 + x, x, 'down_stage2')
        x3, s3, x = self._duck_down(x2 + x, x, 'down_stage3')
        x4, s4, x = self._duck_down(x3 + x, x, 'down_stage4')
        x5, s5, x = self._duck_down(x4 + x, x, 'down_stage5')
        x = x5 + x
        for j in range(4):
            x = self.resblock(x, f'mid_stage.{j}')
        for i, skip in zip((5, 4, 3, 2, 1), (s5, s4, s3, s2, s1)):
            x = F.interpolate(x, skip.shape[2:], mode='nearest')     # models/ducknet.py:80-85
            x = x + skip
            x = self.duck(x, f'up_stage{i}.duck')
        return F.conv2d(x, self.p['seg_head.weight'])

    # -- DuckNet cut at block boundaries (teacher-forcing protocol of the -m gpu model tests) -------
    def stage_fn(self, kind, p):
        """The function one DuckNet block computes from its input tensors (models/ducknet.py:38-49):
        'down1' (img) / 'down' (x_prev1, x) -> (x1, skip, x2); 'res' (x) -> y; 'up' (x, skip) -> y."""
        if kind == 'down1':
            return lambda img: self._duck_down(img, None, p)
        if kind == 'down':
            return lambda a, b: self._duck_down(a + b, b, p)
        if kind == 'res':
            return lambda x: (self.resblock(x, p),)
        if kind == 'up':
            return lambda x, skip: (self.duck(F.interpolate(x, skip.shape[2:], mode='nearest') + skip, p),)
        raise ValueError(kind)

    def ducknet_stages(self, x):
        """Run the DuckNet forward and return ([(prefix, kind, inputs)], logits): the inputs every block saw,
        detached, so that each block can be re-run alone from the same tensors."""
        rec = []

        def run(kind, p, *ins):
            rec.append((p, kind, tuple(t.detach().clone() for t in ins)))
            return self.stage_fn(kind, p)(*ins)
        x1, s1, x = run('down1', 'down_stage1', x)
        skips = [s1]
        for i in (2, 3, 4, 5):
            x1, s, x = run('down', f'down_stage{i}', x1, x)
            skips.append(s)
        x = x1 + x
        for j in range(4):
            x, = run('res', f'mid_stage.{j}', x)
        for i, skip in zip((5, 4, 3, 2, 1), reversed(skips)):
            x, = run('up', f'up_stage{i}.duck', x, skip)
        return rec, F.conv2d(x, self.p['seg_head.weight'])

    def __call__(self, x):
        return self.unet(x) if self.name == 'unet' else self.ducknet(x)


def stage_dout(stage_index, k, shape):
    """Upstream gradient of output ``k`` of DuckNet block ``stage_index`` in the teacher-forcing tests:
    N(0, 1) from numpy PCG64 (torch-version independent), rounded to bf16 so every precision sees the same values."""
    rng = np.random.Generator(np.random.PCG64(1000 + 10 * stage_index + k))
    return torch.from_numpy(rng.standard_normal(tuple(shape)).astype(np.float32)).bfloat16().float()


def make_trainable(sd, dtype=None):
    """Clone a state_dict; float non-buffer entries become leaf tensors with requires_grad."""
    out = OrderedDict()
    for k, v in sd.items():
        v = v.detach().clone()
        if v.is_floating_point() and dtype is not None:
            v = v.to(dtype)
        if v.is_floating_point() and not (k.endswith('running_mean') or k.endswith('running_var')):
            v.requires_grad_(True)
        out[k] = v
    return out


def param_keys(sd):
    return [k for k, v in sd.items()
            if v.is_floating_point() and not (k.endswith('running_mean') or k.endswith('running_var'))]
