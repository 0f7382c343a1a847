"""-m gpu: per-kernel parity of the CUDA path (through the package's public ops -> C-ABI) against
CPU float64 PyTorch restatements of the same op.  Tolerances: fp32 path 1e-5 rel-L2 per op
(north_star: fp32 within 1e-4); bf16 path 2e-2 rel-L2 with bf16-rounded inputs; integer work bit-exact."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from gpu_util import DEV, Cfg, from_act, rel, to_act  # noqa: E402


def _ops():
    from medical_segmentation_pytorch_b200 import ops
    return ops


CONV_CASES = [
    # cin, cout, kh, kw, stride, dil, H, W
    (3, 17, 3, 3, 1, 1, 20, 24),
    (17, 17, 3, 3, 1, 2, 16, 16),
    (17, 34, 3, 3, 1, 3, 16, 12),
    (34, 68, 3, 3, 2, 1, 16, 16),
    (3, 34, 2, 2, 2, 1, 16, 16),
    (24, 40, 1, 1, 1, 1, 9, 11),
    (17, 17, 1, 7, 1, 1, 12, 20),
    (17, 17, 7, 1, 1, 1, 20, 12),
    (64, 96, 3, 3, 1, 1, 11, 11),
    (136, 72, 3, 3, 1, 1, 8, 8),
]


@pytest.mark.parametrize('case', CONV_CASES)
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_conv2d_fwd_bwd(case, dtype):
    ops = _ops()
    cin, cout, kh, kw, stride, dil, H, W = case
    torch.manual_seed(sum(case))
    n = 2
    x = torch.randn(n, cin, H, W)
    w = torch.randn(cout, cin, kh, kw) / (cin * kh * kw) ** 0.5
    if dtype == torch.bfloat16:
        x, w = x.bfloat16().float(), w.bfloat16().float()
    pad = ((kh - 1) // 2 * dil, (kw - 1) // 2 * dil)
    xd, wd = x.double().requires_grad_(True), w.double().requires_grad_(True)
    yref = F.conv2d(xd, wd, None, stride, pad, dil)
    dy = torch.randn(yref.shape)
    if dtype == torch.bfloat16:
        dy = dy.bfloat16().float()
    yref.backward(dy.double())

    g = ops.ConvGeom(cin, cout, kh, kw, stride, dil, pad[0], pad[1], engine=1)     # SIMT engine
    a = to_act(x, dtype, requires_grad=True)
    wp = torch.nn.Parameter(w.to(DEV))
    y = ops.Conv2dFn.apply(a.t, wp, g)
    assert y.shape[3] == ops.rup(cout) and float(y.detach()[..., cout:].abs().sum()) == 0.0     # padded channels stay zero
    y.backward(to_act(dy, dtype).t)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert rel(from_act(y, cout), yref) < tol
    assert rel(from_act(a.t.grad, cin), xd.grad) < tol
    assert rel(wp.grad, wd.grad) < tol


TC_CASES = [
    # cin, cout, kh, kw, dil, H, W, n     (stride 1: the tcgen05 engine's domain, 97.7 % of DUCKNet MACs)
    (17, 17, 3, 3, 1, 32, 48, 2),       # Kp=24 -> 64 B swizzle rows, BN=32
    (3, 17, 3, 3, 1, 32, 32, 2),        # stem: Kp=8
    (34, 34, 3, 3, 2, 24, 24, 2),       # Kp=40 -> one 64-channel chunk, 3 k-steps
    (68, 68, 3, 3, 3, 16, 16, 2),       # two chunks (64 + 8)
    (136, 136, 3, 3, 1, 44, 44, 1),     # 11x11-friendly tiling
    (272, 136, 3, 3, 1, 22, 22, 2),
    (544, 272, 3, 3, 1, 11, 11, 2),     # BN=256+..: two n tiles, 121-row tiles
    (272, 544, 1, 1, 1, 11, 11, 3),     # 1x1, three n tiles
    (17, 17, 1, 7, 1, 20, 40, 2),
    (34, 34, 7, 1, 1, 40, 20, 2),
    (32, 32, 3, 3, 1, 64, 64, 1),
    (64, 32, 3, 3, 1, 40, 40, 2),       # UNet-style, 40x40 (ragged tiles)
    (128, 256, 3, 3, 1, 20, 20, 2),
    (17, 17, 1, 7, 1, 32, 64, 2),       # halo path, x-halo 6
    (17, 17, 7, 1, 1, 64, 32, 2),       # halo path, no x-halo (8-pixel-wide halo rows)
    (17, 17, 3, 3, 3, 48, 48, 2),       # halo path, dilation 3
    (64, 64, 3, 3, 1, 48, 40, 1),       # halo path, full 64-channel chunk
    (17, 17, 1, 1, 1, 32, 48, 2),       # 1x1 at <= 64 channels: conv_tc_halo_kernel<1,K> (resident weights, no halo), K = 2
    (34, 34, 1, 1, 1, 24, 24, 2),       # K = 3
    (64, 64, 1, 1, 1, 24, 40, 2),       # K = 4
    (3, 17, 1, 1, 1, 32, 32, 2),        # K = 1 (DUCK(3 -> 17) shortcuts)
    (34, 17, 1, 1, 1, 32, 32, 2),       # up-path shortcut 2b -> b
    (68, 68, 3, 3, 1, 24, 40, 2),       # halo kernel with a channel tail: 64 + a 16-wide SWIZZLE_32B patch (fprop and dgrad)
    (68, 68, 3, 3, 2, 32, 32, 2),
    (68, 34, 3, 3, 1, 24, 24, 2),       # fprop with the tail (K = 68), dgrad without (K = 34)
    (68, 68, 1, 7, 1, 16, 40, 2),
    (68, 68, 7, 1, 1, 40, 16, 2),
    (68, 68, 1, 1, 1, 24, 24, 3),
    (34, 34, 3, 3, 3, 48, 48, 2),       # wgrad: two x-tap groups of two, windows 3 pixels apart (17-pixel-wide patch)
    (3, 17, 7, 1, 1, 64, 32, 2),        # wgrad: one-column filter, the stacked windows coincide
    (17, 17, 1, 1, 1, 33, 47, 3),       # 1x1 walked as one row of 4 653 pixels: the last 128-pixel tile hangs over the end
    (17, 17, 3, 3, 1, 35, 45, 2),       # wgrad on 16 x 8 tiles with ragged right / bottom tiles
    (34, 17, 3, 3, 2, 24, 40, 2),       # wgrad 16 x 8 tiles, two x-tap groups, dilation 2
]


@pytest.mark.parametrize('case', TC_CASES)
def test_conv_tcgen05_fwd_dgrad(case):
    """tcgen05/TMEM/TMA implicit GEMM (engine forced) vs float64 conv on bf16-rounded operands."""
    ops = _ops()
    from medical_segmentation_pytorch_b200 import _lib
    cin, cout, kh, kw, dil, H, W, n = case
    torch.manual_seed(sum(case))
    x = torch.randn(n, cin, H, W).bfloat16().float()
    w = (torch.randn(cout, cin, kh, kw) / (cin * kh * kw) ** 0.5).bfloat16().float()
    pad = ((kh - 1) // 2 * dil, (kw - 1) // 2 * dil)
    xd, wd = x.double().requires_grad_(True), w.double().requires_grad_(True)
    yref = F.conv2d(xd, wd, None, 1, pad, dil)
    dy = torch.randn(yref.shape).bfloat16().float()
    yref.backward(dy.double())
    g = ops.ConvGeom(cin, cout, kh, kw, 1, dil, pad[0], pad[1], engine=2)
    a = to_act(x, torch.bfloat16, requires_grad=True)
    wp = torch.nn.Parameter(w.to(DEV))
    y = ops.Conv2dFn.apply(a.t, wp, g)
    assert _lib.lib.msx_last_conv_engine() == 2
    assert float(y.detach()[..., cout:].float().abs().sum()) == 0.0
    y.backward(to_act(dy, torch.bfloat16).t)
    torch.cuda.synchronize()
    e_y, e_dx, e_dw = rel(from_act(y, cout), yref), rel(from_act(a.t.grad, cin), xd.grad), rel(wp.grad, wd.grad)
    print(f'tc {case}: y {e_y:.2e} dx {e_dx:.2e} dw {e_dw:.2e}')
    assert e_y < 6e-3 and e_dx < 6e-3 and e_dw < 6e-3          # output rounding to bf16: ~2^-9 per element


S2_CASES = [
    # cin, cout, k, pad, H, W, n, transposed     (stride 2 through the four phase views of the fine side)
    (34, 68, 3, 1, 16, 16, 2, False),
    (3, 34, 2, 0, 32, 32, 2, False),
    (17, 34, 3, 1, 32, 48, 2, False),
    (136, 272, 3, 1, 22, 22, 2, False),
    (272, 544, 2, 0, 22, 22, 1, False),
    (64, 32, 3, 1, 20, 20, 2, True),         # ConvTranspose2d(3, s2, p1, op1): 20 -> 40
    (512, 256, 3, 1, 11, 11, 1, True),
    (17, 24, 3, 1, 12, 20, 2, True),
    (17, 34, 3, 1, 352, 352, 2, False),      # DownsampleBlock.conv1 of stage 1 at the benchmark's real 352 -> 176 size
    (3, 34, 2, 0, 352, 352, 2, False),       # DownsampleBlock.conv2 of stage 1 (2x2 s2 stem) at 352 -> 176
]


@pytest.mark.parametrize('case', S2_CASES)
def test_conv_tcgen05_stride2_and_transposed(case):
    """Stride-2 Conv2d (3x3 p1, 2x2 p0) and ConvTranspose2d(3, 2, 1, 1): fprop, dgrad and wgrad on tcgen05."""
    ops = _ops()
    from medical_segmentation_pytorch_b200 import _lib
    cin, cout, k, pad, H, W, n, tr = case
    torch.manual_seed(sum(case[:7]))
    x = torch.randn(n, cin, H, W).bfloat16().float()
    wshape = (cin, cout, k, k) if tr else (cout, cin, k, k)
    w = (torch.randn(wshape) / (cin * k * k) ** 0.5).bfloat16().float()
    xd, wd = x.double().requires_grad_(True), w.double().requires_grad_(True)
    yref = F.conv_transpose2d(xd, wd, None, 2, pad, 1) if tr else F.conv2d(xd, wd, None, 2, pad)
    dy = torch.randn(yref.shape).bfloat16().float()
    yref.backward(dy.double())
    g = ops.ConvGeom(cin, cout, k, k, 2, 1, pad, pad, transposed=tr, out_pad=1 if tr else 0, engine=2)
    a = to_act(x, torch.bfloat16, requires_grad=True)
    wp = torch.nn.Parameter(w.to(DEV))
    y = ops.Conv2dFn.apply(a.t, wp, g)
    assert _lib.lib.msx_last_conv_engine() == 2
    assert tuple(y.shape[1:3]) == tuple(yref.shape[2:])
    assert float(y.detach()[..., cout:].float().abs().sum()) == 0.0
    y.backward(to_act(dy, torch.bfloat16).t)
    torch.cuda.synchronize()
    e_y, e_dx, e_dw = rel(from_act(y, cout), yref), rel(from_act(a.t.grad, cin), xd.grad), rel(wp.grad, wd.grad)
    print(f'tc-s2 {case}: y {e_y:.2e} dx {e_dx:.2e} dw {e_dw:.2e}')
    assert e_y < 6e-3 and e_dx < 6e-3 and e_dw < 6e-3


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_conv_transpose_bn_relu(dtype):
    """DeConvBNAct: ConvTranspose2d(3, s2, p1, op1, bias) -> BN(train) -> ReLU, forward and all gradients."""
    import medical_segmentation_pytorch_b200 as msx
    torch.manual_seed(3)
    cin, cout, n, H = 16, 8, 2, 6
    m = msx.modules.DeConvBNAct(cin, cout).to(DEV)
    ref = torch.nn.Sequential(torch.nn.ConvTranspose2d(cin, cout, 3, 2, 1, 1), torch.nn.BatchNorm2d(cout), torch.nn.ReLU()).double()
    with torch.no_grad():
        m.up_conv[1].weight.uniform_(0.5, 1.5)
        m.up_conv[1].bias.uniform_(-0.3, 0.3)
        if dtype == torch.bfloat16:
            m.up_conv[0].weight.copy_(m.up_conv[0].weight.bfloat16().float())
    ref.load_state_dict({k.replace('up_conv.', ''): v.double().cpu() for k, v in m.state_dict().items()})
    x = torch.randn(n, cin, H, H)
    if dtype == torch.bfloat16:
        x = x.bfloat16().float()
    xd = x.double().requires_grad_(True)
    yref = ref(xd)
    dy = torch.randn(yref.shape)
    yref.backward(dy.double())
    a = to_act(x, dtype, requires_grad=True)
    z = m(a)
    z.t.backward(to_act(dy, dtype).t)
    tol = 2e-5 if dtype == torch.float32 else 3e-2
    assert rel(from_act(z.t, cout), yref) < tol
    assert rel(from_act(a.t.grad, cin), xd.grad) < tol * 5
    assert rel(m.up_conv[0].weight.grad, ref[0].weight.grad) < tol * 5
    assert rel(m.up_conv[1].weight.grad, ref[1].weight.grad) < tol * 5
    assert rel(m.up_conv[1].bias.grad, ref[1].bias.grad) < tol * 5
    # conv bias gradient is analytically zero when BN follows
    zero_tol = 1e-5 if dtype == torch.float32 else 1e-2        # bf16: dy is rounded before it is summed
    assert float(m.up_conv[0].bias.grad.abs().max()) < zero_tol * float(m.up_conv[1].bias.grad.abs().max()) + 1e-5
    if dtype == torch.float32:
        assert rel(m.up_conv[1].running_mean, ref[1].running_mean) < 1e-5
        assert rel(m.up_conv[1].running_var, ref[1].running_var) < 1e-5
        assert int(m.up_conv[1].num_batches_tracked) == 1


BN_FUSED_CASES = [
    # cin, cout, k, stride, dil, H, W, n
    (17, 17, 3, 1, 1, 40, 56, 2),       # halo kernel, one 32-column chunk (24 stored columns)
    (3, 17, 3, 1, 1, 33, 47, 3),        # ragged tiles: rows past the image must not enter the statistics
    (34, 34, 3, 1, 2, 24, 24, 2),       # halo kernel, 40 columns: one 32-chunk + one 16-chunk
    (64, 64, 1, 1, 1, 24, 40, 2),       # 1x1 halo kernel, two 32-column chunks
    (68, 68, 3, 1, 1, 24, 40, 2),       # halo kernel with the channel tail, 72 stored columns (32 + 32 + 16)
    (17, 17, (1, 7), 1, 1, 20, 40, 2),
    (136, 136, 3, 1, 1, 22, 22, 2),     # streaming kernel, 136 columns
    (272, 136, 3, 1, 1, 11, 11, 2),
    (17, 34, 3, 2, 1, 32, 48, 2),       # stride 2 (streaming kernel over phase views)
    (3, 34, 2, 2, 1, 32, 32, 2),
    (272, 544, 3, 1, 1, 11, 11, 2),     # several tiles of output channels: every CTA's statistics rows cover all 544 columns
]


@pytest.mark.parametrize('case', BN_FUSED_CASES)
def test_conv_bn_relu_train_bf16_statistics_from_the_epilogue(case):
    """ConvBNAct in training mode on the tcgen05 engine: the batch statistics come out of the convolution kernel's epilogue
    (sums of the bf16-rounded outputs, finalized by the launch's last CTA).  Output, running statistics (momentum 0.1,
    unbiased variance), num_batches_tracked and all gradients against float64 conv -> BatchNorm(train) -> ReLU, twice in
    a row (the second call re-uses a re-armed ticket and stale partial rows)."""
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200 import _lib
    cin, cout, k, stride, dil, H, W, n = case
    torch.manual_seed(cin * 1000 + cout + H)
    m = msx.modules.ConvBNAct(cin, cout, k, stride, dil).to(DEV)
    kk = k if isinstance(k, tuple) else (k, k)
    pad = ((kk[0] - 1) // 2 * dil, (kk[1] - 1) // 2 * dil)
    ref = torch.nn.Sequential(torch.nn.Conv2d(cin, cout, kk, stride, pad, dil, bias=False), torch.nn.BatchNorm2d(cout), torch.nn.ReLU()).double()
    with torch.no_grad():
        m[1].weight.uniform_(0.5, 1.5)
        m[1].bias.uniform_(-0.3, 0.3)
        m[0].weight.copy_(m[0].weight.bfloat16().float())
    ref.load_state_dict({k_: v.double().cpu() if v.is_floating_point() else v.cpu() for k_, v in m.state_dict().items()})
    for rep in range(2):
        x = (torch.randn(n, cin, H, W) * (1 + rep) + 0.3 * rep).bfloat16().float()
        xd = x.double().requires_grad_(True)
        zr = ref(xd)
        dz = torch.randn(zr.shape).bfloat16().float()
        ref.zero_grad()
        zr.backward(dz.double())
        a = to_act(x, torch.bfloat16, requires_grad=True)
        for p_ in m.parameters():
            p_.grad = None
        z = m(a)
        assert _lib.lib.msx_last_conv_engine() == 2
        z.t.backward(to_act(dz, torch.bfloat16).t)
        torch.cuda.synchronize()
        e_z = rel(from_act(z.t.detach(), cout), zr)
        e_rm, e_rv = rel(m[1].running_mean, ref[1].running_mean), rel(m[1].running_var, ref[1].running_var)
        e_dx, e_dw = rel(from_act(a.t.grad, cin), xd.grad), rel(m[0].weight.grad, ref[0].weight.grad)
        e_dg, e_db = rel(m[1].weight.grad, ref[1].weight.grad), rel(m[1].bias.grad, ref[1].bias.grad)
        print(f'fused-stats {case} rep {rep}: z {e_z:.2e} rm {e_rm:.2e} rv {e_rv:.2e} dx {e_dx:.2e} dw {e_dw:.2e} dgamma {e_dg:.2e} dbeta {e_db:.2e}')
        assert e_z < 1.5e-2 and e_rm < 5e-3 and e_rv < 5e-3
        assert e_dx < 6e-2 and e_dw < 6e-2 and e_dg < 3e-2 and e_db < 3e-2
        assert int(m[1].num_batches_tracked) == rep + 1 and float(z.t[..., cout:].float().abs().sum()) == 0.0


@pytest.mark.parametrize('c', [3, 17, 40, 272])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_sum_bn_relu_and_upsample(c, dtype):
    """(nearest-2x(a) + b) -> BN(train) -> ReLU vs F.interpolate/+/batch_norm/relu in float64."""
    import medical_segmentation_pytorch_b200 as msx
    torch.manual_seed(c)
    n, h, w = 2, 6, 10
    bn = torch.nn.BatchNorm2d(c).to(DEV)
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.3, 0.3)
    a, b = torch.randn(n, c, h // 2, w // 2) + 0.5, torch.randn(n, c, h, w)
    if dtype == torch.bfloat16:
        a, b = a.bfloat16().float(), b.bfloat16().float()
    ad, bd = a.double().requires_grad_(True), b.double().requires_grad_(True)
    gd, btd = bn.weight.detach().double().cpu().requires_grad_(True), bn.bias.detach().double().cpu().requires_grad_(True)
    s = F.interpolate(ad, (h, w), mode='nearest') + bd
    if dtype == torch.bfloat16:
        s = s + (s.detach().float().bfloat16().double() - s.detach())        # stored sum is bf16-rounded
    zref = F.relu(F.batch_norm(s, None, None, gd, btd, True, 0.1, 1e-5))
    dz = torch.randn(zref.shape)
    zref.backward(dz.double())
    aa, ba = to_act(a, dtype, True), to_act(b, dtype, True)
    z = msx.modules.sum_bn_act([aa, ba], bn, 1, up2x=True)
    z.t.backward(to_act(dz, dtype).t)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert rel(from_act(z.t, c), zref) < tol
    assert rel(from_act(aa.t.grad, c), ad.grad) < tol * 3
    assert rel(from_act(ba.t.grad, c), bd.grad) < tol * 3
    assert rel(bn.weight.grad, gd.grad) < tol * 3
    assert rel(bn.bias.grad, btd.grad) < tol * 3


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_maxpool_3s2(dtype):
    ops = _ops()
    torch.manual_seed(0)
    x = torch.randn(2, 24, 10, 14)
    if dtype == torch.bfloat16:
        x = x.bfloat16().float()
    xd = x.double().requires_grad_(True)
    zref = F.max_pool2d(xd, 3, 2, 1)
    dz = torch.randn(zref.shape)
    zref.backward(dz.double())
    a = to_act(x, dtype, True)
    z = ops.MaxPoolFn.apply(a.t, 24)
    z.backward(to_act(dz, dtype).t)
    assert rel(from_act(z, 24), zref) < (1e-7 if dtype == torch.float32 else 1e-7)
    assert rel(from_act(a.t.grad, 24), xd.grad) < (1e-6 if dtype == torch.float32 else 1e-2)


@pytest.mark.parametrize('case', [(17, 17, 3, 3, 1), (3, 17, 3, 3, 1), (17, 17, 1, 7, 1), (17, 17, 3, 3, 2), (17, 17, 3, 3, 3), (17, 17, 7, 1, 1)])
def test_wgrad_tcgen05_full_size(case):
    """Weight gradient at the benchmark's own size, 16 x 352 x 352 (tap-stacked halo mode, 148 pixel splits, 105 tiles
    accumulated in TMEM per CTA), against the float64 weight gradient of F.conv2d on the CPU."""
    ops = _ops()
    from medical_segmentation_pytorch_b200 import _lib
    cin, cout, kh, kw, dil = case
    n, H = 16, 352
    torch.manual_seed(sum(case))
    torch.set_num_threads(max(1, (__import__('os').cpu_count() or 2)))
    x = torch.randn(n, cin, H, H).bfloat16().float()
    dy = torch.randn(n, cout, H, H).bfloat16().float()
    pad = ((kh - 1) // 2 * dil, (kw - 1) // 2 * dil)
    dw_ref = torch.nn.grad.conv2d_weight(x.double(), (cout, cin, kh, kw), dy.double(), stride=1, padding=pad, dilation=dil)
    g = ops.ConvGeom(cin, cout, kh, kw, 1, dil, pad[0], pad[1], engine=2)
    a = to_act(x, torch.bfloat16)                       # no input gradient: only the wgrad kernels run in backward
    wp = torch.nn.Parameter((torch.randn(cout, cin, kh, kw) / (cin * kh * kw) ** 0.5).to(DEV))
    y = ops.Conv2dFn.apply(a.t, wp, g)
    y.backward(to_act(dy, torch.bfloat16).t)
    torch.cuda.synchronize()
    assert _lib.lib.msx_last_conv_engine() == 2
    e = rel(wp.grad, dw_ref)
    print(f'full-size wgrad {case}: rel-L2 {e:.2e}')
    assert e < 1e-4                                     # bf16-exact operands, fp32 accumulation over 2 M pixels


@pytest.mark.parametrize('nc', [1, 2, 5])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_head(nc, dtype):
    ops = _ops()
    torch.manual_seed(nc)
    c, n, h, w = 17, 2, 8, 12
    x, wt = torch.randn(n, c, h, w), torch.randn(nc, c, 1, 1)
    if dtype == torch.bfloat16:
        x = x.bfloat16().float()
    xd, wd = x.double().requires_grad_(True), wt.double().requires_grad_(True)
    ref = F.conv2d(xd, wd)
    dl = torch.randn(ref.shape)
    ref.backward(dl.double())
    a = to_act(x, dtype, True)
    wp = torch.nn.Parameter(wt.to(DEV))
    out = ops.HeadFn.apply(a.t, wp, c)
    assert out.dtype == torch.float32                    # logits leave the head as NCHW fp32 in either mode
    out.backward(dl.to(DEV))
    tol = 1e-6 if dtype == torch.float32 else 4e-3      # bf16: only dx is rounded to bf16 on store
    assert rel(out, ref) < 1e-6 and rel(wp.grad, wd.grad) < 1e-6
    assert rel(from_act(a.t.grad, c), xd.grad) < tol


def _arena_model():
    import medical_segmentation_pytorch_b200 as msx
    torch.manual_seed(2)
    m = msx.UNet(2, 3, 8).to(DEV)
    ref = [p.detach().clone().requires_grad_(True) for p in m.parameters()]
    return msx, m, ref


@pytest.mark.parametrize('kind', ['sgd', 'adam', 'adamw'])
def test_fused_optimizer_matches_torch_optim(kind):
    """msx_sgd_step / msx_adam_step over the flat arena vs torch.optim.{SGD, Adam, AdamW} with the reference's
    hyper-parameters (utils/optimizer.py:8-16: SGD momentum 0.9, weight decay 1e-4 on every parameter; Adam / AdamW
    torch defaults), four steps with fresh gradients, a changing learning rate and the 1/world gradient scale."""
    msx, m, ref = _arena_model()
    opt = msx.FusedOptimizer(m, kind, lr=0.05, momentum=0.9, weight_decay=1e-4 if kind == 'sgd' else 0.0)
    tkw = dict(lr=0.05)
    if kind == 'sgd':
        tkw.update(momentum=0.9, weight_decay=1e-4)
    topt = {'sgd': torch.optim.SGD, 'adam': torch.optim.Adam, 'adamw': torch.optim.AdamW}[kind](ref, **tkw)
    params = list(m.parameters())
    for step in range(4):
        lr = 0.05 * (1 + step) / 2
        opt.param_groups[0]['lr'] = lr
        topt.param_groups[0]['lr'] = lr
        for p, r in zip(params, ref):
            gr = torch.randn_like(r)
            p.grad.copy_(gr * 2.0)                        # arena gradient holds the SUM over 2 ranks ...
            r.grad = gr.clone()
        opt.step(grad_scale=0.5)                          # ... and the step applies 1/world
        topt.step()
        worst = max(rel(p, r) for p, r in zip(params, ref))
        assert worst < 2e-6, (kind, step, worst)


@pytest.mark.parametrize('use_ema', [False, True])
def test_model_ema_matches_reference_rule(use_ema):
    """ModelEmaV2.update (utils/model_ema.py:29-40) over the arenas: with use_ema the reference applies
    decay * e + (1 - decay) * m to EVERY state_dict tensor - the int64 num_batches_tracked counters included, whose
    float result is truncated back to int64 by copy_ - with decay = clamp(itr / total_itrs, 0, 1); otherwise a plain
    copy.  Bit-exact against the per-tensor torch expressions, over three updates with a training step in between."""
    msx, m, _ = _arena_model()
    from medical_segmentation_pytorch_b200.synth import synth_batch
    ema = msx.ModelEma(m, use_ema=use_ema, total_itrs=10)
    want = {k: v.detach().clone() for k, v in m.state_dict().items()}          # the deepcopy at construction
    x, t = synth_batch(2, 32, 32, seed=4)
    x, t = x.to(DEV), t.to(DEV)
    opt = msx.FusedOptimizer(m, 'sgd', lr=0.05)
    lf = msx.get_loss_fn(Cfg(), DEV)
    m.train()
    for itr in (3, 4, 7):
        for _ in range(itr):                                                # several forwards: num_batches_tracked moves by more than 1
            lf(m(x), t).backward()
        opt.step()
        ema.update(m, itr)
        decay = min(max(itr / 10, 0), 1)
        for k, mv in m.state_dict().items():
            want[k] = (decay * want[k] + (1. - decay) * mv).to(want[k].dtype) if use_ema else mv.detach().clone()
        got = ema.ema.state_dict()
        assert list(got.keys()) == list(want.keys())
        for k in want:
            assert got[k].dtype == want[k].dtype and torch.equal(got[k], want[k]), (k, itr)
    assert not ema.ema.training and all(not p.requires_grad for p in ema.ema.parameters())


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_eval_mode_bn_backward_uses_running_statistics(dtype):
    """A forward in eval mode with autograd on (frozen-BN fine-tuning, saliency maps): BatchNorm is a fixed affine map, so
    dx = gamma * invstd * g without the batch-statistics terms; checked for ConvBNAct and for a stand-alone sum -> BN ->
    ReLU against torch's eval-mode BatchNorm in float64."""
    import medical_segmentation_pytorch_b200 as msx
    torch.manual_seed(9)
    cin, cout, n, H = 8, 17, 2, 12
    m = msx.modules.ConvBNAct(cin, cout, 3).to(DEV)
    with torch.no_grad():
        m[1].running_mean.uniform_(-0.5, 0.5)
        m[1].running_var.uniform_(0.5, 1.5)
        m[1].weight.uniform_(0.5, 1.5)
        m[1].bias.uniform_(-0.3, 0.3)
        if dtype == torch.bfloat16:
            m[0].weight.copy_(m[0].weight.bfloat16().float())
    m.eval()
    ref = torch.nn.Sequential(torch.nn.Conv2d(cin, cout, 3, 1, 1, bias=False), torch.nn.BatchNorm2d(cout), torch.nn.ReLU()).double()
    ref.load_state_dict({k: v.double().cpu() if v.is_floating_point() else v.cpu() for k, v in m.state_dict().items()})
    ref.eval()
    x = torch.randn(n, cin, H, H)
    if dtype == torch.bfloat16:
        x = x.bfloat16().float()
    xd = x.double().requires_grad_(True)
    zr = ref(xd)
    dz = torch.randn(zr.shape)
    zr.backward(dz.double())
    a = to_act(x, dtype, requires_grad=True)
    z = m(a)
    z.t.backward(to_act(dz, dtype).t)
    tol = 2e-5 if dtype == torch.float32 else 3e-2
    assert rel(from_act(z.t.detach(), cout), zr) < tol
    assert rel(from_act(a.t.grad, cin), xd.grad) < tol
    assert rel(m[0].weight.grad, ref[0].weight.grad) < tol
    assert rel(m[1].weight.grad, ref[1].weight.grad) < tol and rel(m[1].bias.grad, ref[1].bias.grad) < tol
    # stand-alone BN after an add (ResidualBlock.bn in eval mode)
    bn = torch.nn.BatchNorm2d(cout).to(DEV)
    with torch.no_grad():
        bn.running_mean.uniform_(-0.5, 0.5)
        bn.running_var.uniform_(0.5, 1.5)
    bn.eval()
    u, v = torch.randn(n, cout, H, H), torch.randn(n, cout, H, H)
    if dtype == torch.bfloat16:
        u, v = u.bfloat16().float(), v.bfloat16().float()
    ud, vd = u.double().requires_grad_(True), v.double().requires_grad_(True)
    bref = torch.nn.BatchNorm2d(cout).double()
    bref.load_state_dict({k: t.double().cpu() if t.is_floating_point() else t.cpu() for k, t in bn.state_dict().items()})
    bref.eval()
    zr = torch.relu(bref(ud + vd))
    zr.backward(dz.double())
    ua, va = to_act(u, dtype, True), to_act(v, dtype, True)
    z = msx.modules.sum_bn_act([ua, va], bn, 1)
    z.t.backward(to_act(dz, dtype).t)
    assert rel(from_act(ua.t.grad, cout), ud.grad) < tol and rel(from_act(va.t.grad, cout), vd.grad) < tol
    assert rel(bn.weight.grad, bref.weight.grad) < tol


# ---- losses ------------------------------------------------------------------------------------
def _loss_and_grad(fn, z):
    zz = z.clone().to(DEV).requires_grad_(True)
    v = fn(zz)
    v.backward()
    return float(v), zz.grad.cpu()


def test_losses_match_golden(golden_dir):
    """CUDA losses vs outputs of the REAL reference (tests/golden/losses.npz)."""
    import os
    import medical_segmentation_pytorch_b200 as msx
    g = np.load(os.path.join(golden_dir, 'losses.npz'))
    z, zt, t = torch.from_numpy(g['logits']), torch.from_numpy(g['teacher']).to(DEV), torch.from_numpy(g['labels']).to(DEV)
    cases = {
        'ce': msx.get_loss_fn(Cfg(), DEV),
        'ce_weighted': msx.get_loss_fn(Cfg(class_weights=g['class_weights'].tolist()), DEV),
        'ohem_thresh_branch': msx.get_loss_fn(Cfg(loss_type='ohem', ohem_thrs=0.7), DEV),
        'ohem_topk_branch': msx.get_loss_fn(Cfg(loss_type='ohem', ohem_thrs=float(g['ohem_topk_thresh'])), DEV),
    }
    for key, fn in cases.items():
        v, gr = _loss_and_grad(lambda zz: fn(zz, t), z)
        assert abs(v - float(g[key])) < 1e-5 * max(1.0, abs(float(g[key]))), key
        assert rel(gr, g[key + '_grad']) < 1e-5, key
    for kind in ('kl_div', 'mse'):
        cfg = Cfg(kd_loss_type=kind)
        v, gr = _loss_and_grad(lambda zz: msx.kd_loss_fn(cfg, zz, zt), z)
        assert abs(v - float(g['kd_' + kind])) < 1e-5 * max(1.0, abs(float(g['kd_' + kind]))), kind
        assert rel(gr, g['kd_' + kind + '_grad']) < 1e-5, kind


@pytest.mark.parametrize('shape', [(2, 2, 16, 20), (3, 2, 7, 9), (1, 5, 8, 8), (2, 11, 6, 6)])
def test_ce_vs_oracle_shapes(shape):
    """vector (hw%4==0) and scalar paths, register-resident and generic class counts, 'sum', upstream scale."""
    import medical_segmentation_pytorch_b200 as msx
    from oracle import ref_loss
    torch.manual_seed(shape[1])
    n, c, h, w = shape
    z = torch.randn(shape) * 3
    t = torch.randint(0, c, (n, h, w))
    t[torch.rand(n, h, w) < 0.2] = 255
    for red in ('mean', 'sum'):
        fn = msx.get_loss_fn(Cfg(reduction=red), DEV)
        v, gr = _loss_and_grad(lambda zz: fn(zz, t.to(DEV)) * 0.5, z)
        zd = z.double().requires_grad_(True)
        ref = ref_loss.ce_loss(zd, t, None, 255, red) * 0.5
        ref.backward()
        assert abs(v - ref.item()) < 2e-6 * max(1, abs(ref.item()))
        assert rel(gr, zd.grad) < 1e-5
    fn = msx.get_loss_fn(Cfg(reduction='none'), DEV)
    zz = z.clone().to(DEV).requires_grad_(True)
    per = fn(zz, t.to(DEV))
    coef = torch.rand(n, h, w)
    (per * coef.to(DEV)).sum().backward()
    zd = z.double().requires_grad_(True)
    refp = ref_loss.ce_loss(zd, t, None, 255, 'none')
    (refp * coef.double()).sum().backward()
    assert rel(per, refp) < 1e-6 and rel(zz.grad, zd.grad) < 1e-5


@pytest.mark.parametrize('shape', [(2, 2, 8, 16), (3, 3, 6, 6), (2, 5, 5, 3), (1, 12, 4, 8)])
def test_ce_label_code_workspace_is_bit_identical(shape):
    """msx_ws_bytes_ce() workspace (one-byte label codes) vs the small workspace: same bits, incl. ignored,
    out-of-range and class-weighted pixels."""
    from medical_segmentation_pytorch_b200 import _lib as L
    torch.manual_seed(7)
    n, c, h, w = shape
    z = (torch.randn(shape) * 3).to(DEV)
    t = torch.randint(0, c, (n, h, w))
    t[torch.rand(n, h, w) < 0.2] = 255
    t[0, 0, 0] = c + 3                      # out of range: treated as ignored by both paths
    t = t.to(DEV)
    cw = (torch.rand(c) + 0.5).to(DEV)
    res = []
    for nbytes in (L.lib.msx_ws_bytes_loss(), L.lib.msx_ws_bytes_ce(n * h * w)):
        ws = torch.zeros(nbytes, dtype=torch.uint8, device=DEV)
        loss, grad = torch.empty((), device=DEV), torch.empty_like(z)
        L.check(L.lib.msx_ce_loss_fwd_bwd(L.ptr(z), L.ptr(t), n, c, h * w, L.ptr(cw), 255, L.REDUCE['mean'],
                                          L.ptr(loss), L.ptr(grad), L.ptr(ws), ws.numel(), L.stream()))
        res.append((loss.cpu(), grad.cpu()))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])


def test_ce_all_ignored_is_nan_with_zero_grad():
    import medical_segmentation_pytorch_b200 as msx
    fn = msx.get_loss_fn(Cfg(), DEV)
    z = torch.randn(1, 2, 4, 4, device=DEV, requires_grad=True)
    t = torch.full((1, 4, 4), 255, dtype=torch.int64, device=DEV)
    v = fn(z, t)
    v.backward()
    assert torch.isnan(v) and float(z.grad.abs().sum()) == 0.0


def test_dice_bce_vs_oracle():
    import medical_segmentation_pytorch_b200 as msx
    from oracle import ref_loss
    torch.manual_seed(1)
    z = torch.randn(2, 1, 12, 10) * 2
    for t in ((torch.rand(2, 12, 10) > 0.7).long(), (torch.rand(2, 1, 12, 10) > 0.7).float()):
        fn = msx.get_loss_fn(Cfg(loss_type='dice_bce'), DEV)
        v, gr = _loss_and_grad(lambda zz: fn(zz, t.to(DEV)), z)
        zd = z.double().requires_grad_(True)
        ref = ref_loss.dice_bce_loss(zd, t.double())
        ref.backward()
        assert abs(v - ref.item()) < 2e-6 and rel(gr, zd.grad) < 1e-5


# ---- metrics -----------------------------------------------------------------------------------
def test_confmat_bit_exact_with_ties_nan_ignore():
    import medical_segmentation_pytorch_b200 as msx
    from oracle import ref_metrics
    rng = np.random.default_rng(5)
    for c, shape in ((2, (3, 2, 16, 20)), (2, (1, 2, 7, 9)), (4, (2, 4, 8, 8)), (7, (2, 7, 8, 12))):
        z = rng.standard_normal(shape).astype(np.float32)
        z[rng.uniform(size=shape) < 0.1] = 0.5                  # many exact ties
        z[rng.uniform(size=shape) < 0.02] = np.nan
        t = rng.integers(0, c, size=(shape[0],) + shape[2:])
        t[rng.uniform(size=t.shape) < 0.1] = 255
        cfg = Cfg(num_class=c)
        m = msx.get_seg_metrics(cfg, 'iou').to(DEV)
        for _ in range(2):
            m.update(torch.from_numpy(z).to(DEV), torch.from_numpy(t).to(DEV))
        want = 2 * ref_metrics.confusion_matrix(z, t, c, 255)
        assert np.array_equal(m.confmat.cpu().numpy(), want)
        got = m.compute().cpu().numpy()
        assert np.array_equal(got, ref_metrics.iou_per_class(want))          # float32 ratios, same cast order
        m.reset()
        assert int(m.confmat.sum()) == 0
        # Dice has no ignore_index: a 255 label must raise at compute(), like torchmetrics' validation
        d = msx.get_seg_metrics(cfg, 'dice').to(DEV)
        d.update(torch.from_numpy(z).to(DEV), torch.from_numpy(t).to(DEV))
        with pytest.raises(ValueError):
            d.compute()
        d.reset()
        t2 = np.where(t == 255, 0, t)
        d.update(torch.from_numpy(z).to(DEV), torch.from_numpy(t2).to(DEV))
        cm2 = ref_metrics.confusion_matrix(z, t2, c, None)
        assert np.array_equal(d.confmat.cpu().numpy(), cm2)
        assert float(d.compute()) == float(ref_metrics.dice_macro(cm2))
    with pytest.raises(ValueError):
        msx.get_seg_metrics(Cfg(), 'f1')


def test_full_size_properties():
    """BASELINE config-2 size (16x2x352x352): size-independent properties instead of a CPU oracle."""
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200.synth import synth_batch
    n, h = 16, 352
    _, mask = synth_batch(n, h, h, seed=1, ignore_band=0.05)
    torch.manual_seed(0)
    z = torch.randn(n, 2, h, h, device=DEV, requires_grad=True)
    mask = mask.to(DEV)
    m = msx.get_seg_metrics(Cfg(), 'iou').to(DEV)
    m.update(z, mask)
    assert int(m.confmat.sum()) == int((mask != 255).sum())                        # every valid pixel counted once
    assert torch.equal(m.confmat.sum(1), torch.stack([(mask == 0).sum(), (mask == 1).sum()]))   # row sums = label histogram
    loss = msx.get_loss_fn(Cfg(), DEV)(z, mask)
    loss.backward()
    g = z.grad
    assert float(g.sum(1).abs().max()) < 1e-9                                       # softmax - onehot sums to 0 over classes
    assert float(g[:, 0][mask == 255].abs().max()) == 0.0                            # ignored pixels get no gradient
    want = torch.nn.functional.cross_entropy(z.detach(), mask, ignore_index=255)     # same-device library check of the scalar
    assert abs(float(loss) - float(want)) < 1e-5


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_pack_plan_matches_per_conv_pack(dtype):
    """msx_conv_pack_plan_* (one launch for many weights) writes the same bytes as msx_conv_pack_weights per conv and pass."""
    import ctypes as C
    from medical_segmentation_pytorch_b200 import _lib as L
    torch.manual_seed(3)
    geoms = [(17, 3, 3, 3, 0), (34, 17, 1, 7, 0), (5, 40, 7, 1, 0), (24, 24, 1, 1, 0), (16, 32, 3, 3, 1), (130, 68, 3, 3, 0), (8, 8, 2, 2, 0)]
    es = 2 if dtype == torch.bfloat16 else 4
    items, want = [], []
    for cout, cin, kh, kw, tr in geoms:
        w = torch.randn((cin, cout, kh, kw) if tr else (cout, cin, kh, kw), device=DEV)
        d = L.ConvDesc(n=1, h=8, w=8, cin=cin, oh=8 + 2 * ((kh - 1) // 2) - kh + 1, ow=8 + 2 * ((kw - 1) // 2) - kw + 1, cout=cout, kh=kh, kw=kw,
                       stride=1, dil=1, pad_h=(kh - 1) // 2, pad_w=(kw - 1) // 2,
                       ldx=(cin + 7) // 8 * 8, ldy=(cout + 7) // 8 * 8, transposed=0, dtype=L.dtype_code(dtype), engine=0)
        if tr:                       # stride-2 ConvTranspose2d 3x3 (DeConvBNAct): 8x8 -> 16x16
            d.stride, d.oh, d.ow, d.transposed = 2, 16, 16, 1
        nbytes = L.lib.msx_conv_packed_bytes(C.byref(d))
        for mode in (L.PACK_FPROP, L.PACK_DGRAD):
            ref = torch.full((nbytes,), 0x5a, dtype=torch.uint8, device=DEV)
            L.check(L.lib.msx_conv_pack_weights(C.byref(d), mode, L.ptr(w), L.ptr(ref), L.stream()))
            out = torch.full((nbytes,), 0x5a, dtype=torch.uint8, device=DEV)
            items.append(L.PackItem(w_master=w.data_ptr(), packed=out.data_ptr(), cout=cout, cin=cin, kh=kh, kw=kw, transposed=tr, mode=mode))
            want.append((ref, out, w))
    arr = (L.PackItem * len(items))(*items)
    nb = L.lib.msx_conv_pack_plan_bytes(arr, len(items))
    table = torch.empty(nb, dtype=torch.uint8, device=DEV)
    chunks = C.c_int32(0)
    L.check(L.lib.msx_conv_pack_plan_build(arr, len(items), L.ptr(table), nb, C.byref(chunks)))
    for _ in range(2):
        L.check(L.lib.msx_conv_pack_plan_run(L.dtype_code(dtype), L.ptr(table), len(items), chunks.value, L.stream()))
    for i, (ref, out, _) in enumerate(want):
        assert torch.equal(ref, out), (i, es)


def test_model_forward_repacks_changed_weights():
    """The per-forward pack plan must see in-place weight updates (optimizer steps, load_state_dict)."""
    import medical_segmentation_pytorch_b200 as msx
    torch.manual_seed(0)
    m = msx.UNet(num_class=2, base_channel=8).to(DEV).eval()
    x = torch.randn(1, 3, 32, 32, device=DEV)
    with torch.no_grad():
        y0 = m(x).clone()
        for p in m.parameters():
            if p.dim() == 4:
                p.mul_(0.5)
        y1 = m(x).clone()
        m2 = msx.UNet(num_class=2, base_channel=8).to(DEV).eval()
        m2.load_state_dict(m.state_dict())
        y2 = m2(x)
    assert not torch.equal(y0, y1)
    assert torch.equal(y1, y2)


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
@pytest.mark.parametrize('kind', ['conv3', 'conv17', 'conv_d2', 'conv_s2', 'deconv', 'conv136'])
def test_inference_epilogue_fusion_matches_unfused_eval(kind, dtype):
    """Eval-mode ConvBNAct / DeConvBNAct under no_grad (BN affine + ReLU in the conv epilogue, y_raw = NULL) vs the
    unfused eval path (taken when a gradient may be needed) and vs float64 conv -> BN(running stats) -> ReLU."""
    from medical_segmentation_pytorch_b200 import modules
    torch.manual_seed(11)
    n, H = 2, 16
    if kind == 'deconv':
        cin, cout = 16, 8
        m = modules.DeConvBNAct(cin, cout)
        conv, bn = m.up_conv[0], m.up_conv[1]
    else:
        cin, cout, k, s, d = {'conv3': (3, 17, 3, 1, 1), 'conv17': (17, 17, 3, 1, 1), 'conv_d2': (24, 24, 3, 1, 2),
                              'conv_s2': (17, 34, 3, 2, 1), 'conv136': (136, 136, 3, 1, 1)}[kind]
        m = modules.ConvBNAct(cin, cout, k, s, d)
        conv, bn = m[0], m[1]
    with torch.no_grad():
        bn.running_mean.copy_(torch.randn(cout) * 0.5)
        bn.running_var.copy_(torch.rand(cout) + 0.5)
        bn.weight.copy_(torch.randn(cout))
        bn.bias.copy_(torch.randn(cout))
        if dtype == torch.bfloat16:
            conv.weight.copy_(conv.weight.bfloat16().float())
    m = m.to(DEV).eval()
    x = torch.randn(n, cin, H, H)
    if dtype == torch.bfloat16:
        x = x.bfloat16().float()
    with torch.no_grad():
        fused = m(to_act(x, dtype))
    unfused = m(to_act(x, dtype, requires_grad=True))          # eval mode, but a backward could follow: keeps y_raw
    xd = x.double()
    cd = conv.double().cpu()
    yref = cd(xd)
    conv.float()
    zref = torch.relu((yref - bn.running_mean.cpu().double()[None, :, None, None]) / torch.sqrt(bn.running_var.cpu().double() + bn.eps)[None, :, None, None]
                      * bn.weight.detach().cpu().double()[None, :, None, None] + bn.bias.detach().cpu().double()[None, :, None, None])
    assert float(fused.t[..., cout:].float().abs().sum()) == 0.0          # padded channels stay zero
    e_f, e_u = rel(from_act(fused.t, cout), zref), rel(from_act(unfused.t.detach(), cout), zref)
    print(f'eval fusion {kind} {dtype}: fused {e_f:.2e} unfused {e_u:.2e}')
    tol = 1e-5 if dtype == torch.float32 else 8e-3
    assert e_f < tol and e_u < tol
    assert e_f <= 1.2 * e_u + 1e-6                                         # never less accurate than the two-pass path


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
@pytest.mark.parametrize('n_in,up2x', [(2, False), (6, False), (2, True)])
def test_inference_sum_bn_relu_single_kernel(n_in, up2x, dtype):
    """Eval-mode (sum of inputs [first nearest-2x upsampled]) -> BN -> ReLU under no_grad runs as ONE kernel (s = NULL);
    compared with the two-pass eval path and with float64."""
    from medical_segmentation_pytorch_b200 import modules
    torch.manual_seed(5)
    n, c, H = 2, 20, 12
    bn = torch.nn.BatchNorm2d(c)
    with torch.no_grad():
        bn.running_mean.copy_(torch.randn(c) * 0.5)
        bn.running_var.copy_(torch.rand(c) + 0.5)
        bn.weight.copy_(torch.randn(c))
        bn.bias.copy_(torch.randn(c))
    bn = bn.to(DEV).eval()
    xs = [torch.randn(n, c, H // 2 if (up2x and i == 0) else H, H // 2 if (up2x and i == 0) else H) for i in range(n_in)]
    if dtype == torch.bfloat16:
        xs = [x.bfloat16().float() for x in xs]
    with torch.no_grad():
        fused = modules.sum_bn_act([to_act(x, dtype) for x in xs], bn, 1, up2x=up2x)
    unfused = modules.sum_bn_act([to_act(x, dtype, requires_grad=True) for x in xs], bn, 1, up2x=up2x)
    tot = sum((F.interpolate(x.double(), scale_factor=2, mode='nearest') if (up2x and i == 0) else x.double()) for i, x in enumerate(xs))
    b = lambda t: t.detach().cpu().double()[None, :, None, None]          # noqa: E731
    zref = torch.relu((tot - b(bn.running_mean)) / torch.sqrt(b(bn.running_var) + bn.eps) * b(bn.weight) + b(bn.bias))
    e_f, e_u = rel(from_act(fused.t, c), zref), rel(from_act(unfused.t.detach(), c), zref)
    assert float(fused.t[..., c:].float().abs().sum()) == 0.0
    tol = 1e-6 if dtype == torch.float32 else 6e-3
    assert e_f < tol and e_u < tol and e_f <= 1.2 * e_u + 1e-7


def test_halo_conv_repeatable_under_concurrency():
    """The halo kernel's two MMA-issuing warps own disjoint stages / accumulators; a protocol slip shows up as a changed
    result (or a trap) only at full size and with other kernels competing for the SMs: 30 launches of a 17->17 3x3 conv at
    16x352x352 while a second stream runs bandwidth kernels, every result bit-identical to the first, and the first equal
    to the FFMA engine's within bf16 rounding."""
    ops = _ops()
    torch.manual_seed(0)
    n, c, H = 16, 17, 352
    x = torch.randn(n, H, H, ops.rup(c), device=DEV).bfloat16()
    x[..., c:] = 0
    w = torch.nn.Parameter((torch.randn(c, c, 3, 3) / (9 * c) ** 0.5).to(DEV))
    g_tc, g_simt = ops.ConvGeom(c, c, 3, 3, 1, 1, 1, 1, engine=2), ops.ConvGeom(c, c, 3, 3, 1, 1, 1, 1, engine=1)
    side = torch.cuda.Stream()
    noise = torch.randn(64 << 20, device=DEV)
    with torch.no_grad():
        ref = ops.Conv2dFn.apply(x, w, g_tc).clone()
        simt = ops.Conv2dFn.apply(x, w, g_simt)
        assert rel(ref[..., :c].float().cpu(), simt[..., :c].float().cpu()) < 6e-3
        for i in range(30):
            with torch.cuda.stream(side):
                noise.mul_(1.0001)
            y = ops.Conv2dFn.apply(x, w, g_tc)
            assert torch.equal(y, ref), i
    torch.cuda.synchronize()
