"""-m gpu, collected last on purpose (the file name sorts after every other test file): the
allocation-free BatchNorm-backward entry point, msx_bn_act_bwd_static, called through the C-ABI.

Who needs it: callers that may not allocate device memory or supply a workspace - an inference/training
engine whose memory plan is fixed up front, a deployment where cudaMalloc after start-up is ruled out
because it can block or fail, a plug-in that is only handed its input and output buffers.  For them the
reduction scratch is built into the library (capped at MSX_BN_STATIC_MAX_CHANNELS channels).

What is checked: on the same inputs the new entry point gives exactly the results of the existing
msx_bn_act_bwd (same kernels, same reduction order, so the comparison is bit-for-bit - tighter than the
1e-5 the other fp32 tests allow), twice in a row with different data so that whatever the first call left
in the shared scratch cannot leak into the second; that it allocates nothing; and that a channel count over
the cap is refused with a clear message.  The test passes on its own in a fresh process and after the rest
of the suite.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_CHANNELS = 256          # MSX_BN_STATIC_MAX_CHANNELS in include/msx.h


def _case(seed, n, h, w, c, dtype):
    g = torch.Generator().manual_seed(seed)
    cp = (c + 7) // 8 * 8
    s = torch.zeros(n, h, w, cp)
    dz = torch.zeros(n, h, w, cp)
    s[..., :c] = torch.randn(n, h, w, c, generator=g) * 1.5 + 0.3
    dz[..., :c] = torch.randn(n, h, w, c, generator=g)
    gamma = torch.rand(c, generator=g) + 0.5
    beta = torch.rand(c, generator=g) - 0.5
    mean = s[..., :c].reshape(-1, c).mean(0)
    invstd = 1.0 / torch.sqrt(s[..., :c].reshape(-1, c).var(0, unbiased=False) + 1e-5)
    pad = lambda v: torch.cat([v, torch.zeros(cp - c)]).contiguous()
    dev = 'cuda'
    return dict(s=s.to(dev, dtype), dz=dz.to(dev, dtype), gamma=gamma.to(dev), beta=beta.to(dev),
                mean=pad(mean).to(dev), invstd=pad(invstd).to(dev), c=c, cp=cp, pixels=n * h * w)


def _run(lib, L, k, static, check_no_alloc=False):
    ds = torch.empty_like(k['s'])
    dgamma, dbeta = torch.empty_like(k['gamma']), torch.empty_like(k['beta'])
    if check_no_alloc:
        # outputs exist already: from here to the end of the call neither the device's free memory (cudaMalloc by
        # anyone in the process) nor torch's allocation counter may move.  (Comparing torch.cuda.memory_allocated()
        # instead depends on which cached block the allocator hands out, i.e. on the tests that ran before.)
        torch.cuda.synchronize()
        free0, count0 = torch.cuda.mem_get_info()[0], torch.cuda.memory_stats()['allocation.all.allocated']
    code = L.dtype_code(k['s'].dtype)
    common = (code, L.ptr(k['s']), k['cp'], L.ptr(k['dz']), k['cp'], k['pixels'], k['c'], L.ptr(k['gamma']), L.ptr(k['beta']),
              L.ptr(k['mean']), L.ptr(k['invstd']), 1, L.ptr(ds), k['cp'], L.ptr(dgamma), L.ptr(dbeta))
    if static:
        rc = lib.msx_bn_act_bwd_static(*common, L.stream())
    else:
        ws = torch.empty(lib.msx_ws_bytes_bn(k['c']), dtype=torch.uint8, device='cuda')
        rc = lib.msx_bn_act_bwd(*common, L.ptr(ws), ws.numel(), L.stream())
    L.check(rc)
    torch.cuda.synchronize()
    if check_no_alloc:
        assert torch.cuda.mem_get_info()[0] == free0
        assert torch.cuda.memory_stats()['allocation.all.allocated'] == count0
    return ds, dgamma, dbeta


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_static_scratch_entry_point_matches_workspace_entry_point(dtype):
    from medical_segmentation_pytorch_b200 import _lib as L
    lib = L.lib
    # two calls in a row, the second with more channels and pixels than the first and other data
    cases = [_case(1, 2, 24, 40, 17, dtype), _case(2, 3, 32, 48, 136, dtype)]
    expected = [_run(lib, L, k, static=False) for k in cases]
    torch.cuda.synchronize()
    got = []
    # first call on this device: looks up the scratch symbol (with lazy module loading this may be when the driver
    # reserves the array), so it is made before the calls whose allocations are counted
    _run(lib, L, cases[0], static=True)
    for k in cases:
        got.append(_run(lib, L, k, static=True, check_no_alloc=True))
    for exp, out in zip(expected, got):
        for a, b in zip(exp, out):
            assert torch.equal(a, b)
    # and once more in reverse order: the scratch now holds the larger case's leftovers
    again = _run(lib, L, cases[0], static=True)
    for a, b in zip(expected[0], again):
        assert torch.equal(a, b)


def test_static_scratch_entry_point_rejects_oversize():
    from medical_segmentation_pytorch_b200 import _lib as L
    k = _case(3, 1, 8, 8, MAX_CHANNELS + 8, torch.float32)
    with pytest.raises(L.MsxError, match='MSX_BN_STATIC_MAX_CHANNELS'):
        _run(L.lib, L, k, static=True)
    ok = _case(4, 1, 8, 8, MAX_CHANNELS, torch.float32)          # the cap itself is accepted
    a, b = _run(L.lib, L, ok, static=False), _run(L.lib, L, ok, static=True)
    assert all(torch.equal(x, y) for x, y in zip(a, b))
