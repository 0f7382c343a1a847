"""Bisect which part of the step breaks CUDA-graph capture."""
import os, sys, traceback
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import medical_segmentation_pytorch_b200 as msx
from medical_segmentation_pytorch_b200.flat import FlatBuffers, FlatParams
from medical_segmentation_pytorch_b200.synth import synth_batch

dev = torch.device('cuda')
class Cfg: num_class, ignore_index, class_weights, loss_type, reduction = 2, 255, None, 'ce', 'mean'
torch.manual_seed(1)
model = msx.DuckNet(2, 3, 17).to(dev).train()
FlatParams(model); FlatBuffers(model)
opt = msx.FusedOptimizer(model, 'adam', lr=1e-3)
ema = msx.ModelEma(model)
loss_fn = msx.get_loss_fn(Cfg, dev)
SZ = int(os.environ.get('SZ', 64))
x, t = synth_batch(2, SZ, SZ, seed=1)
x, t = x.to(dev), t.to(dev)

def fwd():
    with torch.no_grad(), torch.autocast('cuda', dtype=torch.bfloat16):
        return model(x)
def fwd_loss():
    with torch.no_grad(), torch.autocast('cuda', dtype=torch.bfloat16):
        return loss_fn(model(x), t)
def fwd_bwd():
    with torch.autocast('cuda', dtype=torch.bfloat16):
        l = loss_fn(model(x), t)
    l.backward()
    return l
def full():
    l = fwd_bwd(); opt.step(); ema.update(model, 1); return l

for name, fn in (('fwd', fwd), ('fwd_loss', fwd_loss), ('fwd_bwd', fwd_bwd), ('opt', lambda: opt.step()), ('ema', lambda: ema.update(model, 1)), ('full', full)):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode=os.environ.get('MODE', 'global')):
            out = fn()
        g.replay(); torch.cuda.synchronize()
        print(name, 'capture OK')
    except Exception as e:
        print(name, 'capture FAILED:', str(e)[:200])
        traceback.print_exc(limit=6)
        torch.cuda.synchronize()
