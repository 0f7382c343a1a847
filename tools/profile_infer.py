"""Per-kernel-class time of one eval forward (library profiler, eager launches)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import medical_segmentation_pytorch_b200 as msx  # noqa: E402
from medical_segmentation_pytorch_b200 import _lib as L  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'teacher101'
dev = torch.device('cuda')
torch.manual_seed(1)
if name.startswith('teacher'):
    model = msx.SmpUnet('resnet' + name[7:], None, 3, 2)
elif name.startswith('ducknet'):
    model = msx.DuckNet(num_class=2, base_channel=int(name[7:]))
else:
    model = msx.UNet(num_class=2, base_channel=int(name[4:]))
model = model.to(dev).eval()
x = torch.randn(16, 3, 352, 352, device=dev)
with torch.no_grad(), torch.autocast('cuda', dtype=torch.bfloat16):
    for _ in range(2):
        model(x)
    torch.cuda.synchronize()
    L.profile_collect()
    L.lib.msx_profile_enable(1)
    model(x)
    prof = L.profile_collect()
    L.lib.msx_profile_enable(0)
tot = sum(v['ms'] for v in prof.values())
for k, v in sorted(prof.items(), key=lambda kv: -kv[1]['ms']):
    print(f"{k:24s} {v['ms']:8.3f} ms {v['ms'] / tot * 100:5.1f}% {v['launches']:5d} launches {v['bytes'] / v['ms'] / 1e6 if v['ms'] else 0:8.0f} GB/s {v['flops'] / v['ms'] / 1e9 if v['ms'] else 0:8.1f} TF/s")
print(f'total {tot:.3f} ms')
