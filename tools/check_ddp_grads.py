"""torchrun --nproc-per-node 2 tools/check_ddp_grads.py
Numerical check of DistributedDataParallelB200 on real GPUs: gradients left in the flat arena after backward (bucketed
NCCL AVG issued from the comm stream while side streams are still producing other buckets) must equal, bit for bit, the
plain per-rank gradients averaged with one explicit all-reduce.  Several steps with different data, bf16 and fp32."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import medical_segmentation_pytorch_b200 as msx  # noqa: E402
from medical_segmentation_pytorch_b200.flat import FlatParams  # noqa: E402

rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)


class Cfg:
    num_class, ignore_index, class_weights, loss_type, reduction = 2, 255, None, 'ce', 'mean'


torch.manual_seed(1)
a = msx.DuckNet(num_class=2, n_channel=3, base_channel=8).to(dev).train()
b = msx.DuckNet(num_class=2, n_channel=3, base_channel=8).to(dev).train()
b.load_state_dict(a.state_dict())
FlatParams(a)
ddp = msx.DistributedDataParallelB200(a)
b.load_state_dict(a.state_dict())                       # after the rank-0 broadcast
loss_fn = msx.get_loss_fn(Cfg, dev)
bad = 0
for step in range(6):
    bf16 = step % 2 == 1
    g = torch.Generator(device='cpu').manual_seed(100 * step + rank)
    x = torch.randn(4, 3, 96, 96, generator=g).to(dev)
    t = torch.randint(0, 2, (4, 96, 96), generator=g).to(dev)
    for m in (a, b):
        for p in m.parameters():
            if getattr(p, '_msx_grad', None) is None:
                p.grad = None
    with torch.autocast('cuda', dtype=torch.bfloat16, enabled=bf16):
        loss_fn(ddp(x), t).backward()
    # (train-mode BN normalises with batch statistics, so the gradients do not depend on the running buffers)
    with torch.autocast('cuda', dtype=torch.bfloat16, enabled=bf16):
        loss_fn(b(x), t).backward()
    torch.cuda.synchronize()
    for (n, pa), pb in zip(a.named_parameters(), b.parameters()):
        ref = pb.grad.clone()
        dist.all_reduce(ref, op=dist.ReduceOp.AVG)
        if not torch.equal(pa.grad, ref):
            bad += 1
            if rank == 0 and bad < 5:
                print(f'step {step} {n}: max diff {(pa.grad - ref).abs().max().item():.3e}', flush=True)
    for ba, bb in zip(a.buffers(), b.buffers()):
        bb.copy_(ba)
tot = torch.tensor([bad], device=dev)
dist.all_reduce(tot)
if rank == 0:
    print('DDP gradient check:', 'OK' if int(tot) == 0 else f'{int(tot)} mismatching tensors', flush=True)
dist.barrier()
torch.cuda.synchronize()
os._exit(0 if int(tot) == 0 else 1)
