"""CPU tests of the host-side logic: the C-ABI library loads and exports every symbol include/msx.h
declares (no compute calls without a GPU), state_dict layouts, DDP bucket assignment against torch's
own algorithm, and the bucketed all-reduce over gloo with world_size 2."""
import os
import re

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import ctypes
    from medical_segmentation_pytorch_b200 import _lib
    hdr = open(os.path.join(ROOT, 'include', 'msx.h')).read()
    hdr = re.sub(r'/\*.*?\*/', '', hdr, flags=re.S)
    declared = set(re.findall(r'\b(msx_[a-z0-9_]+)\s*\(', hdr))
    assert len(declared) >= 35
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in sorted(declared) if not hasattr(lib, n)]
    assert not missing, f'declared in msx.h but not exported: {missing}'
    unbound = [n for n in sorted(declared) if n not in _lib.SIGNATURES]
    assert not unbound, f'declared but not bound in _lib.py: {unbound}'
    assert _lib.lib.msx_version() == 100
    assert _lib.lib.msx_launch_count() == 0                 # nothing ran: CPU box


def test_no_cpu_fallback():
    """The product path refuses CPU tensors instead of silently computing somewhere else."""
    import medical_segmentation_pytorch_b200 as msx
    m = msx.UNet(2, 3, 8)
    with pytest.raises(Exception):
        m(torch.zeros(1, 3, 32, 32))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, 'medical_segmentation_pytorch_b200')
    for fn in os.listdir(pkg):
        if fn.endswith('.py'):
            src = open(os.path.join(pkg, fn)).read()
            assert 'oracle' not in src.replace('# oracle', ''), fn


def test_state_dict_layout_matches_reference():
    import medical_segmentation_pytorch_b200 as msx
    from oracle import ref_models
    for name, cls, b in (('unet', msx.UNet, 32), ('ducknet', msx.DuckNet, 17)):
        sd = cls(num_class=2, n_channel=3, base_channel=b).state_dict()
        lay = ref_models.model_layout(name, num_class=2, n_channel=3, base_channel=b)
        assert list(sd.keys()) == list(lay.keys())
        assert all(tuple(sd[k].shape) == tuple(lay[k]) for k in lay)


def test_default_init_consumes_rng_like_torch_modules():
    """Same constructors in the same order as the reference => identical weights from the same seed."""
    import medical_segmentation_pytorch_b200 as msx
    torch.manual_seed(1)
    a = msx.DuckNet(2, 3, 2)
    torch.manual_seed(1)
    b = msx.DuckNet(2, 3, 2)
    assert all(torch.equal(x, y) for x, y in zip(a.state_dict().values(), b.state_dict().values()))
    w = a.down_stage1.duck.branch1[0][0].weight
    assert float(w.detach().abs().max()) <= 1.0 / (3 * 9) ** 0.5 + 1e-6        # kaiming_uniform(a=sqrt(5)) bound


@pytest.mark.parametrize('name,base,want', [('unet', 32, 3), ('ducknet', 17, 7)])
def test_bucket_assignment_matches_torch_ddp(name, base, want):
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200.parallel import BUCKET_CAP_BYTES, FIRST_BUCKET_BYTES, compute_buckets
    cls = msx.UNet if name == 'unet' else msx.DuckNet
    params = list(reversed(list(cls(2, 3, base).parameters())))
    mine = compute_buckets([p.numel() * 4 for p in params])
    ref, _ = dist._compute_bucket_assignment_by_size(params, [FIRST_BUCKET_BYTES, BUCKET_CAP_BYTES])
    assert mine == [list(b) for b in ref]
    assert len(mine) == want                                 # 3 / 7 buckets


def _ddp_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from medical_segmentation_pytorch_b200.parallel import DistributedDataParallelB200, compute_buckets, de_parallel, is_parallel
    import medical_segmentation_pytorch_b200.parallel as par
    torch.manual_seed(100 + rank)                                      # different init per rank: wrap must broadcast rank 0
    net = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.BatchNorm1d(32), torch.nn.Linear(32, 8))
    old = par.compute_buckets
    par.compute_buckets = lambda sizes: old(sizes, 64, 256)                 # force several buckets on a toy model
    ddp = DistributedDataParallelB200(net)
    assert is_parallel(ddp) and de_parallel(ddp) is net and len(ddp.buckets) >= 3
    assert list(ddp.state_dict().keys()) == list(net.state_dict().keys())       # no 'module.' prefix
    w0 = [p.detach().clone() for p in net.parameters()]
    gathered = [None] * world
    dist.all_gather_object(gathered, [w.tolist() for w in w0])
    assert gathered[0] == gathered[1]                                   # parameters broadcast from rank 0
    torch.manual_seed(7 + rank)
    x = torch.randn(4, 16)
    ddp.flat.grad.zero_()
    loss = ddp(x).pow(2).mean()
    loss.backward()                                                     # torch accumulates into the arena views
    local = ddp.flat.grad.clone()
    for p in ddp.flat.order:                                            # what ops.* do after each backward kernel
        ddp.mark_ready(p)
    ddp.finish()
    both = [None] * world
    dist.all_gather_object(both, local.tolist())
    want = (torch.tensor(both[0]) + torch.tensor(both[1])) / world
    ok = torch.allclose(ddp.flat.grad, want, atol=1e-7)
    # running statistics of rank 0 win at the next forward (broadcast_buffers=True)
    ddp(x)
    rm = [None] * world
    dist.all_gather_object(rm, net[1].running_mean.tolist())
    out.put((rank, bool(ok), all(b['pending'] == b['n'] for b in ddp.buckets)))
    dist.destroy_process_group()


def test_bucketed_allreduce_world2_gloo():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_ddp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert all(ok and reset for _, ok, reset in res)


def test_teacher_state_dict_matches_the_oracle_layout():
    """The product teacher's parameter holders carry smp/torchvision names, shapes and order, so a reference teacher
    checkpoint loads unchanged (models/__init__.py:52-53)."""
    import medical_segmentation_pytorch_b200 as msx
    from oracle.ref_teacher import SmpUnetRef
    for enc in ('resnet18', 'resnet101'):
        a, b = msx.SmpUnet(enc, None, 3, 2), SmpUnetRef(enc, 3, 2)
        sa, sb = a.state_dict(), b.state_dict()
        assert list(sa.keys()) == list(sb.keys())
        assert all(sa[k].shape == sb[k].shape and sa[k].dtype == sb[k].dtype for k in sb)
        a.load_state_dict(sb)


def test_optimizer_and_ema_built_from_the_ddp_wrapper_share_the_wrapped_arena():
    """The usual PyTorch order (wrap first, then build the optimizer): FusedOptimizer / ModelEma / get_optimizer must
    find the arena of the WRAPPED module instead of flattening the parameters a second time (which would leave DDP's
    buckets reducing a dead gradient arena)."""
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200.flat import FlatParams
    m = msx.UNet(2, 3, 8)
    ddp = msx.DistributedDataParallelB200(m)                # world 1 without a process group: host logic only
    opt = msx.FusedOptimizer(ddp, 'sgd', lr=0.1)
    assert opt.flat is ddp.flat and m._msx_flat is ddp.flat
    ema = msx.ModelEma(ddp)
    assert ema.src_p is ddp.flat

    class C:
        optimizer_type, base_lr, gpu_num, momentum, weight_decay = 'adamw', 0.01, 2, 0.9, 1e-4
    opt2 = msx.get_optimizer(C, ddp)
    assert opt2.flat is ddp.flat and abs(C.lr - 0.002) < 1e-12 and opt2.param_groups[0]['weight_decay'] == 0.01
    with pytest.raises(RuntimeError, match='already live in a flat arena'):
        FlatParams(m)
    for p, q in zip(m.parameters(), opt.param_groups[0]['params']):
        assert p is q


def test_reference_lr_policies_drive_the_fused_optimizer():
    """utils/scheduler.py:12-19: OneCycleLR stepped every iteration, which also cycles momentum (SGD) / beta1 (Adam);
    the fused optimizer is a torch.optim.Optimizer, so torch's scheduler runs on it unchanged and step() sees the values."""
    import medical_segmentation_pytorch_b200 as msx

    class C:
        DDP, train_num, train_bs, gpu_num, total_epoch, warmup_epochs, lr_policy = False, 80, 8, 1, 4, 1, 'cos_warmup'
        optimizer_type, base_lr, momentum, weight_decay = 'sgd', 0.01, 0.9, 1e-4
    m = msx.UNet(2, 3, 8)
    opt = msx.get_optimizer(C, m)
    sched = msx.get_scheduler(C, opt)
    ref_opt = torch.optim.SGD([torch.nn.Parameter(torch.zeros(1))], lr=C.lr, momentum=0.9, weight_decay=1e-4)
    ref_sched = torch.optim.lr_scheduler.OneCycleLR(ref_opt, max_lr=C.lr, total_steps=C.total_itrs, pct_start=0.25)
    assert C.total_itrs == 40
    for _ in range(5):
        assert opt._hyper()[:2] == (ref_opt.param_groups[0]['lr'], ref_opt.param_groups[0]['momentum'])
        sched.step()
        ref_opt.step()
        ref_sched.step()
    assert opt.hyper_changed()
    C.lr_policy = 'nope'
    with pytest.raises(NotImplementedError):
        msx.get_scheduler(C, opt)


def test_parallel_model_with_synbn_marks_every_batchnorm():
    """utils/parallel.py:35-39 with the reference's default synBN=True: BatchNorms are switched to cross-rank statistics
    (same classes, same state_dict keys), and nothing changes for synBN=False."""
    import medical_segmentation_pytorch_b200 as msx

    class C:
        DDP, synBN = True, True
    m = msx.UNet(2, 3, 8)
    keys = list(m.state_dict().keys())
    w = msx.parallel_model(C, m, 0, torch.device('cpu'))
    bns = [x for x in m.modules() if isinstance(x, torch.nn.BatchNorm2d)]
    assert msx.is_parallel(w) and bns and all(hasattr(b, '_msx_sync_group') for b in bns)
    assert list(w.state_dict().keys()) == keys
    C.synBN = False
    m2 = msx.UNet(2, 3, 8)
    msx.parallel_model(C, m2, 0, torch.device('cpu'))
    assert not any(hasattr(b, '_msx_sync_group') for b in m2.modules())
