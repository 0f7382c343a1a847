"""Inference throughput (the reference's tools/test_speed.py: eval-mode forward, synthetic input), bf16,
one CUDA graph replay per forward.  Prints img/s for the epilogue-fused path (eval-mode BN + ReLU inside the
conv kernels) and for the unfused path (conv, BN-apply as separate passes), batch 16 and batch 1."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import medical_segmentation_pytorch_b200 as msx  # noqa: E402
from medical_segmentation_pytorch_b200 import ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--model', default='ducknet17')
ap.add_argument('--size', type=int, default=352)
ap.add_argument('--iters', type=int, default=30)
args = ap.parse_args()
dev = torch.device('cuda')
torch.manual_seed(1)
if args.model.startswith('teacher'):
    model = msx.SmpUnet('resnet' + args.model[7:], None, 3, 2)
elif args.model.startswith('ducknet'):
    model = msx.DuckNet(num_class=2, base_channel=int(args.model[7:]))
else:
    model = msx.UNet(num_class=2, base_channel=int(args.model[4:]))
model = model.to(dev).eval()
for batch in (16, 1):
    x = torch.randn(batch, 3, args.size, args.size, device=dev)
    ref = None
    for fused in (1, 0):
        ops.set_eval_fusion(fused)
        with torch.no_grad(), torch.autocast('cuda', dtype=torch.bfloat16):
            for _ in range(3):
                y = model(x)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                y = model(x)
            g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.iters):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.iters
        if ref is None:
            ref = y.float().clone()
        d = float((y.float() - ref).abs().max() / ref.abs().max())
        print(f'{args.model} {args.size}x{args.size} batch {batch:2d} bf16 eval  fused={fused}  {ms:7.3f} ms/forward  {batch / ms * 1e3:8.1f} img/s  max-rel-diff vs fused {d:.2e}')
ops.set_eval_fusion(1)
