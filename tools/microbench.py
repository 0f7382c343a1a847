"""Per-layer-shape timing of the conv engines (CUDA events, L2 flushed between repeats).
python tools/microbench.py [--batch 16] [--what fprop,dgrad,wgrad]"""
import argparse
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from medical_segmentation_pytorch_b200 import _lib as L, ops  # noqa: E402
from medical_segmentation_pytorch_b200._lib import lib  # noqa: E402

SHAPES = [  # cin, cout, kh, kw, dil, HW     (DUCKNet-17 layer classes)
    (3, 17, 3, 3, 1, 352), (17, 17, 3, 3, 1, 352), (17, 17, 3, 3, 2, 352), (17, 17, 1, 7, 1, 352), (17, 17, 1, 1, 1, 352),
    (34, 34, 3, 3, 1, 176), (68, 68, 3, 3, 1, 88), (136, 136, 3, 3, 1, 44), (272, 272, 3, 3, 1, 22), (544, 544, 3, 3, 1, 11),
    (32, 32, 3, 3, 1, 352), (64, 64, 3, 3, 1, 176), (128, 128, 3, 3, 1, 88), (256, 256, 3, 3, 1, 44), (512, 512, 3, 3, 1, 22),
    (17, 17, 3, 3, 3, 352), (17, 17, 7, 1, 1, 352), (34, 34, 3, 3, 3, 176), (34, 34, 7, 1, 1, 176), (34, 34, 1, 7, 1, 176), (34, 34, 3, 3, 2, 176),   # 15-20
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--reps', type=int, default=5)
    ap.add_argument('--engine', type=int, default=0)
    ap.add_argument('--only', type=int, default=-1, help='index into SHAPES')
    ap.add_argument('--passes', default='fprop,dgrad,wgrad')
    a = ap.parse_args()
    dev = torch.device('cuda')
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    st = L.stream()
    print(f'{"shape":28s} {"pass":6s} {"us":>9s} {"GB/s":>8s} {"TF/s":>8s}')
    for idx, (cin, cout, kh, kw, dil, hw) in enumerate(SHAPES):
        if a.only >= 0 and idx != a.only:
            continue
        n = a.batch
        g = ops.ConvGeom(cin, cout, kh, kw, 1, dil, (kh - 1) // 2 * dil, (kw - 1) // 2 * dil, engine=a.engine)
        x = torch.randn(n, hw, hw, ops.rup(cin), device=dev).bfloat16()
        x[..., cin:] = 0
        dy = torch.randn(n, hw, hw, ops.rup(cout), device=dev).bfloat16()
        dy[..., cout:] = 0
        w = torch.randn(cout, cin, kh, kw, device=dev) * 0.05
        y, dx, dw = torch.empty_like(dy), torch.empty_like(x), torch.empty_like(w)
        d = g.desc(x, ops.rup(cout), hw, hw)
        d.engine = a.engine
        pk = torch.empty(lib.msx_conv_packed_bytes(C.byref(d)), dtype=torch.uint8, device=dev)
        ws = torch.empty(lib.msx_ws_bytes_conv(C.byref(d)), dtype=torch.uint8, device=dev)
        px = n * hw * hw
        by, fl = px * (cin + cout) * 2, 2.0 * px * cin * cout * kh * kw

        def fprop():
            L.check(lib.msx_conv2d_fwd(C.byref(d), L.ptr(x), L.ptr(w), None, L.ptr(pk), L.ptr(y), st))

        def dgrad():
            L.check(lib.msx_conv2d_bwd(C.byref(d), L.ptr(x), L.ptr(w), L.ptr(dy), L.ptr(pk), L.ptr(dx), None, None, L.ptr(ws), ws.numel(), st))

        def wgrad():
            L.check(lib.msx_conv2d_bwd(C.byref(d), L.ptr(x), L.ptr(w), L.ptr(dy), L.ptr(pk), None, L.ptr(dw), None, L.ptr(ws), ws.numel(), st))
        gam, bet = torch.rand(cout, device=dev) + 0.5, torch.rand(cout, device=dev)
        rm, rv = torch.zeros(cout, device=dev), torch.ones(cout, device=dev)
        nbt = torch.zeros((), dtype=torch.int64, device=dev)
        cpo = ops.rup(cout)
        mean, inv, z = torch.empty(cpo, device=dev), torch.empty(cpo, device=dev), torch.empty_like(dy)

        def cbr():      # ConvBNAct training forward: conv (+ statistics) -> finalize -> apply  (MSX_FUSED_STATS=0: separate statistics pass)
            L.check(lib.msx_conv_bn_act_fwd(C.byref(d), L.ptr(x), L.ptr(w), None, L.ptr(gam), L.ptr(bet), L.ptr(rm), L.ptr(rv), L.ptr(nbt),
                                            1e-5, 0.1, 1, 1, L.ptr(pk), L.ptr(y), L.ptr(z), cpo, L.ptr(mean), L.ptr(inv), L.ptr(ws), ws.numel(), st))
        for name, fn in (('fprop', fprop), ('dgrad', dgrad), ('wgrad', wgrad), ('cbr', cbr)):
            if name not in a.passes.split(','):
                continue
            fn()
            torch.cuda.synchronize()
            # kernel-only time from the library's per-launch CUDA events (host launch latency excluded)
            L.profile_collect()
            lib.msx_profile_enable(1)
            for _ in range(a.reps):
                flush.zero_()
                fn()
            prof = L.profile_collect()
            lib.msx_profile_enable(0)
            main_k = {k: v for k, v in prof.items() if k.startswith('conv_')}
            ms = sum(v['ms'] for v in main_k.values()) / a.reps
            other = sum(v['ms'] for k, v in prof.items() if not k.startswith('conv_')) / a.reps
            detail = ' '.join(f'{k}={v["ms"] / a.reps * 1e3:.1f}' for k, v in prof.items()) if name == 'cbr' else ''
            print(f'{cin:4d}->{cout:4d} {kh}x{kw} d{dil} @{hw:3d}^2     {name:6s} {ms * 1e3:9.1f} {by / ms / 1e6:8.0f} {fl / ms / 1e9:8.1f}   (+{other * 1e3:.1f} us pack/reduce/bn) {detail}')


if __name__ == '__main__':
    main()
