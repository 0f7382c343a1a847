"""Kernel-only timing of the BatchNorm / sum kernels on DUCKNet-17's biggest activation shapes."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from medical_segmentation_pytorch_b200 import _lib as L, ops  # noqa: E402
from medical_segmentation_pytorch_b200._lib import lib  # noqa: E402

dev = torch.device('cuda')
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for c, hw in ((17, 352), (34, 176), (68, 88), (136, 44)):
    n, cp = 16, ops.rup(c)
    x = torch.randn(n, hw, hw, cp, device=dev).bfloat16()
    dz = torch.randn(n, hw, hw, cp, device=dev).bfloat16()
    z, ds = torch.empty_like(x), torch.empty_like(x)
    g, b = torch.rand(c, device=dev) + 0.5, torch.rand(c, device=dev)
    rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
    nbt = torch.zeros((), dtype=torch.int64, device=dev)
    mean, inv = torch.empty(cp, device=dev), torch.empty(cp, device=dev)
    dg, db = torch.empty(c, device=dev), torch.empty(c, device=dev)
    ws = torch.empty(lib.msx_ws_bytes_bn(c), dtype=torch.uint8, device=dev)
    ptrs = (C.c_void_p * 1)(x.data_ptr())
    lds = (C.c_int32 * 1)(cp)
    st = L.stream()

    def fwd():
        L.check(lib.msx_sum_bn_act_fwd(1, 1, ptrs, lds, 0, n, hw, hw, c, None, 0, L.ptr(g), L.ptr(b), L.ptr(rm), L.ptr(rv), L.ptr(nbt),
                                       1e-5, 0.1, 1, 1, L.ptr(z), cp, L.ptr(mean), L.ptr(inv), L.ptr(ws), ws.numel(), st))

    def bwd():
        L.check(lib.msx_bn_act_bwd(1, L.ptr(x), cp, L.ptr(dz), cp, n * hw * hw, c, L.ptr(g), L.ptr(b), L.ptr(mean), L.ptr(inv), 1,
                                   L.ptr(ds), cp, L.ptr(dg), L.ptr(db), L.ptr(ws), ws.numel(), st))
    fwd(); bwd(); torch.cuda.synchronize()
    L.profile_collect()
    lib.msx_profile_enable(1)
    for _ in range(3):
        flush.zero_(); fwd(); flush.zero_(); bwd()
    prof = L.profile_collect()
    lib.msx_profile_enable(0)
    for k, v in prof.items():
        print(f'C={c:4d} @{hw:3d}^2  {k:16s} {v["ms"] / v["launches"] * 1e3:8.1f} us  {v["bytes"] / v["ms"] / 1e6:7.0f} GB/s')
