"""-m gpu: whole-model parity of the CUDA path against vectors from the REAL reference modules
(tests/golden/*.npz; fp64 run = ground truth, fp32 run = the reference's own noise floor).

Protocol: an error of the new path against fp64 truth must not exceed
max(3 x the reference-fp32 error against the same truth, floor).  Loss within 1e-4 relative
(north_star fp32 tolerance).  bf16 (autocast) tolerance, stated.  UNet: loss within 5e-2 relative; logits rel-L2 against fp64 truth at most
1.5 x what the reference itself shows under torch.autocast(dtype=bfloat16) on the same inputs (golden
'logits_bf16': 0.06).  DUCK-Net: whole-network bf16 logits are NOT a usable yardstick - the unmodified reference
under bf16 autocast differs from its own fp64 twin by ~0.7 rel-L2 at every size and width tried (stacked
train-mode BN amplifies each rounding ~50x: the same network disagrees with itself by 4e-4 in fp32) - so the
bf16 gate is block-wise teacher forcing at the benchmark's own channel widths (17 * 2^k, 34 * 2^k): every
DuckNet block is run alone from the fp64-truth inputs of that block (rounded to bf16) and compared with the
fp64 oracle; the bound is a multiple of the error the REAL reference shows in the same protocol
(tests/golden/ducknet_b*_stages.npz, written by oracle/gen_golden.py --wide), every bound <= 0.2."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from gpu_util import DEV, Cfg, rel  # noqa: E402


def _build(name, base):
    import medical_segmentation_pytorch_b200 as msx
    from oracle import ref_models
    cls = msx.UNet if name == 'unet' else msx.DuckNet
    return cls(num_class=2, n_channel=3, base_channel=base).to(DEV), ref_models


@pytest.mark.parametrize('name,fname', [('unet', 'unet_b8_32.npz'), ('ducknet', 'ducknet_b3_64.npz'), ('ducknet', 'ducknet_b17_128.npz')])
def test_model_fp32_vs_reference_golden(golden_dir, name, fname):
    import medical_segmentation_pytorch_b200 as msx
    g = np.load(os.path.join(golden_dir, fname))
    base = int(g['base_channel'])
    model, ref_models = _build(name, base)
    sd = ref_models.synth_state_dict(ref_models.model_layout(name, num_class=2, n_channel=3, base_channel=base), int(g['weight_seed']))
    model.load_state_dict(sd)                               # reference key layout loads unchanged
    model.train()
    x, t = torch.from_numpy(g['images']).to(DEV), torch.from_numpy(g['masks']).to(DEV)
    logits = model(x)
    loss = msx.get_loss_fn(Cfg(), DEV)(logits, t)
    loss.backward()
    floor = 2e-5
    e_ref = rel(g['logits_f32'], g['logits_f64'])
    e_new = rel(logits, g['logits_f64'])
    print(f'{name}: logits err new {e_new:.2e} ref-fp32 {e_ref:.2e}')
    assert e_new <= max(3 * e_ref, floor)
    assert abs(float(loss) - float(g['loss_f64'])) <= 1e-4 * abs(float(g['loss_f64']))
    names = [str(s) for s in g['param_names']]
    params = dict(model.named_parameters())
    assert names == list(params.keys())
    gn = np.array([float(params[k].grad.norm()) for k in names])
    t64, r32 = g['grad_norms_f64'], g['grad_norms_f32']
    big = t64 > 1e-9 * t64.max()
    e_gn_new = np.abs(gn[big] / t64[big] - 1)
    e_gn_ref = np.abs(r32[big] / t64[big] - 1)
    print(f'{name}: grad-norm err median new {np.median(e_gn_new):.2e} ref {np.median(e_gn_ref):.2e}; max new {e_gn_new.max():.2e} ref {e_gn_ref.max():.2e}')
    assert np.median(e_gn_new) <= max(3 * np.median(e_gn_ref), 1e-4)
    assert e_gn_new.max() <= max(3 * e_gn_ref.max(), 1e-3)
    for k in g.files:
        if k.startswith('grad_f64::'):
            pn = k.split('::')[1]
            e_r = rel(g['grad_f32::' + pn], g[k])
            assert rel(params[pn].grad, g[k]) <= max(3 * e_r, 1e-4), pn
        if k.startswith('buf_f64::'):
            bn = k.split('::')[1]
            # statistics of late layers inherit the upstream activation noise (logits themselves differ by e_ref)
            assert rel(model.state_dict()[bn], g[k]) <= max(3 * rel(g['buf_f32::' + bn], g[k]), 0.1 * e_ref, 1e-5), bn
    assert int(model.state_dict()[[k for k in model.state_dict() if k.endswith('num_batches_tracked')][0]]) == 1
    # validation path: eval-mode BN with the just-updated running statistics, no autograd
    model.eval()
    with torch.no_grad():
        ev = model(x)
    assert rel(ev, g['eval_logits_f64']) <= max(3 * rel(g['eval_logits_f32'], g['eval_logits_f64']), floor)
    # state_dict round trip back into the reference layout
    out = model.state_dict()
    assert list(out.keys()) == list(sd.keys()) and all(tuple(out[k].shape) == tuple(sd[k].shape) for k in sd)


def test_unet_bf16_autocast(golden_dir):
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200 import _lib
    g = np.load(os.path.join(golden_dir, 'unet_b8_32.npz'))
    base = int(g['base_channel'])
    model, ref_models = _build('unet', base)
    model.load_state_dict(ref_models.synth_state_dict(ref_models.model_layout('unet', num_class=2, n_channel=3, base_channel=base), int(g['weight_seed'])))
    model.train()
    x, t = torch.from_numpy(g['images']).to(DEV), torch.from_numpy(g['masks']).to(DEV)
    with torch.autocast('cuda', dtype=torch.bfloat16):          # the reference's AMP switch (core/seg_trainer.py:61)
        logits = model(x)
        loss = msx.get_loss_fn(Cfg(), DEV)(logits, t)
    loss.backward()
    assert logits.dtype == torch.float32 and _lib.lib.msx_last_conv_engine() == 2
    e = rel(logits, g['logits_f64'])
    e_ref = rel(g['logits_bf16'], g['logits_f64'])           # the reference itself under bf16 autocast (CPU): 0.06
    print(f'unet bf16: logits rel-L2 {e:.3e} (reference bf16 autocast {e_ref:.3e}), loss {float(loss):.5f} vs {float(g["loss_f64"]):.5f}')
    assert e_ref < 0.1 and e <= 1.5 * e_ref
    assert abs(float(loss) - float(g['loss_f64'])) <= 5e-2 * abs(float(g['loss_f64']))
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.parameters())


@pytest.mark.parametrize('fname,loss_tol', [('ducknet_b3_64.npz', 5e-2), ('ducknet_b17_128.npz', 1e-2)])
def test_ducknet_bf16_autocast_whole_model(golden_dir, fname, loss_tol):
    """Whole DUCK-Net step under bf16 autocast on the tcgen05 engine: loss within 1e-2 of the fp64 reference at the
    benchmark's widths (the reference's own bf16 run is off by 6e-3 there; 5e-2 on the 3-channel toy whose deepest
    BatchNorms see 8 samples), all gradients finite.  The logits error is printed, not bounded (module docstring:
    the reference's own figure is 0.5-0.8); block-wise parity is the next test."""
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200 import _lib
    g = np.load(os.path.join(golden_dir, fname))
    base = int(g['base_channel'])
    model, ref_models = _build('ducknet', base)
    model.load_state_dict(ref_models.synth_state_dict(ref_models.model_layout('ducknet', num_class=2, n_channel=3, base_channel=base), int(g['weight_seed'])))
    model.train()
    x, t = torch.from_numpy(g['images']).to(DEV), torch.from_numpy(g['masks']).to(DEV)
    with torch.autocast('cuda', dtype=torch.bfloat16):
        logits = model(x)
        loss = msx.get_loss_fn(Cfg(), DEV)(logits, t)
    loss.backward()
    torch.cuda.synchronize()
    assert logits.dtype == torch.float32 and _lib.lib.msx_last_conv_engine() == 2
    e, e_ref = rel(logits, g['logits_f64']), rel(g['logits_bf16'], g['logits_f64'])
    print(f'ducknet-{base} bf16: logits rel-L2 {e:.3e} (reference bf16 autocast {e_ref:.3e}), loss {float(loss):.6f} vs {float(g["loss_f64"]):.6f} '
          f'(reference bf16 {float(g["loss_bf16"]):.6f})')
    assert abs(float(loss) - float(g['loss_f64'])) <= loss_tol * abs(float(g['loss_f64']))
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.parameters())


def _module_of(model, prefix):
    mod = model
    for part in prefix.replace('.duck', '').split('.'):
        mod = mod[int(part)] if part.isdigit() else getattr(mod, part)
    return mod


def _run_block(model, kind, prefix, acts):
    from medical_segmentation_pytorch_b200.modules import fanout
    mod = _module_of(model, prefix)
    if kind == 'down1':
        return mod([acts[0]], acts[0])
    if kind == 'down':                                   # DuckNet._forward: x feeds the add and the next conv2
        xa, xb = fanout(acts[1], 2)
        return mod([acts[0], xa], xb)
    if kind == 'res':
        return (mod(acts[0]),)
    return (mod(acts[0], acts[1]),)


@pytest.mark.parametrize('fname,backward', [('ducknet_b17_128', True), ('ducknet_b34_96', False)])
def test_ducknet_bf16_blocks_teacher_forced(golden_dir, fname, backward):
    """Every block of DuckNet-17 / -34 (5 DownsampleBlocks, 4 mid ResidualBlocks, 5 UpsampleBlocks: all 273 convs at
    17 * 2^k / 34 * 2^k channels) in bf16 on the tcgen05 engine, alone, from the fp64 oracle's inputs of that block
    rounded to bf16; outputs, input gradients and every parameter gradient of the block against the fp64 oracle.
    Bounds, all relative to what the REAL reference shows under bf16 autocast in the same protocol (fixture):
    forward <= 2 x (reference 0.005-0.016, so <= 0.032); input gradients, all parameter gradients of the block taken
    as one vector, and the per-tensor median <= 1.25 x (reference 0.06-0.18, so <= 0.22); worst single parameter
    tensor <= 1.25 x the reference's own worst tensor (0.07-0.42: a few weight tensors that feed a BatchNorm have
    gradients that are small differences of large terms, in the reference exactly as here)."""
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200 import _lib
    from gpu_util import from_act, to_act
    g = np.load(os.path.join(golden_dir, fname + '.npz'))
    st = np.load(os.path.join(golden_dir, fname + '_stages.npz'))
    base = int(g['base_channel'])
    model, ref_models = _build('ducknet', base)
    sd = ref_models.synth_state_dict(ref_models.model_layout('ducknet', num_class=2, n_channel=3, base_channel=base), int(g['weight_seed']))
    model.load_state_dict(sd)
    model.train()
    params = dict(model.named_parameters())
    torch.set_num_threads(max(1, (os.cpu_count() or 2)))
    p64 = ref_models.make_trainable(sd, torch.float64)
    net = ref_models.FnNet('ducknet', p64, training=True)
    with torch.no_grad():
        plan, _ = net.ducknet_stages(torch.from_numpy(g['images']).double())
    assert [pfx for pfx, _, _ in plan] == [str(s) for s in st['stage_names']]
    bad = []
    for idx, (prefix, kind, ins) in enumerate(plan):
        ins_r = [t.float().bfloat16().float() for t in ins]
        i64 = [t.double().requires_grad_(backward and kind != 'down1') for t in ins_r]
        o64 = net.stage_fn(kind, prefix)(*i64)
        acts = [to_act(t, torch.bfloat16, requires_grad=backward and kind != 'down1') for t in ins_r]
        outs = _run_block(model, kind, prefix, acts)
        assert _lib.lib.msx_last_conv_engine() == 2
        e_fwd = max(rel(from_act(o.t.detach(), o.c), r) for o, r in zip(outs, o64))
        ref_fwd = float(st['ref_bf16_fwd'][idx])
        msg = f'{fname} {prefix}: fwd {e_fwd:.3e} (reference bf16 {ref_fwd:.3e})'
        if e_fwd > 2.0 * ref_fwd:
            bad.append(msg)
        if backward:
            for v in list(p64.values()) + list(params.values()):
                v.grad = None
            douts = [ref_models.stage_dout(idx, k, o.shape) for k, o in enumerate(o64)]
            sum((o * d.double()).sum() for o, d in zip(o64, douts)).backward()
            torch.autograd.backward([o.t for o in outs], [to_act(d, torch.bfloat16).t for d in douts])
            torch.cuda.synchronize()
            e_din = max([rel(from_act(a.t.grad, a.c), r.grad) for a, r in zip(acts, i64) if r.grad is not None] or [0.0])
            keys = [k for k in params if k.startswith(prefix.replace('.duck', '') + '.')]
            gmax = max(float(p64[k].grad.norm()) for k in keys)
            errs = np.array([rel(params[k].grad, p64[k].grad) for k in keys if float(p64[k].grad.norm()) > 1e-9 * gmax])
            e_glob = rel(torch.cat([params[k].grad.flatten() for k in keys]), torch.cat([p64[k].grad.flatten() for k in keys]))
            r_din, r_med, r_max = float(st['ref_bf16_din'][idx]), float(st['ref_bf16_dparam_median'][idx]), float(st['ref_bf16_dparam_max'][idx])
            r_glob = float(st['ref_bf16_dparam_global'][idx])
            msg += (f' din {e_din:.3e} ({r_din:.3e}) dparam global {e_glob:.3e} ({r_glob:.3e}) median {np.median(errs):.3e} ({r_med:.3e}) '
                    f'max {errs.max():.3e} ({r_max:.3e})')
            bad.append(msg) if (e_din > 1.25 * r_din or e_glob > 1.25 * r_glob or np.median(errs) > 1.25 * r_med or errs.max() > 1.25 * r_max) else None
        print(msg)
    assert not bad, '\n'.join(bad)


def test_train_step_flat_arena_matches_autograd_path():
    """FlatParams + fused Adam + EMA copy: same gradients as the plain autograd path, parameters move
    exactly like torch.optim.Adam, EMA model equals the model (use_ema=False copies, utils/model_ema.py:39-40)."""
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200.flat import FlatParams
    from medical_segmentation_pytorch_b200.synth import synth_batch
    torch.manual_seed(1)
    a = msx.DuckNet(2, 3, 2).to(DEV)
    b = msx.DuckNet(2, 3, 2).to(DEV)
    b.load_state_dict(a.state_dict())
    x, t = synth_batch(2, 32, 32, seed=3)
    x, t = x.to(DEV), t.to(DEV)
    lf = msx.get_loss_fn(Cfg(), DEV)
    lf(a(x), t).backward()
    ref_grads = [p.grad.clone() for p in a.parameters()]
    opt_ref = torch.optim.Adam(a.parameters(), lr=1e-3)
    opt_ref.step()
    FlatParams(b)
    opt = msx.FusedOptimizer(b, 'adam', lr=1e-3)
    ema = msx.ModelEma(b)
    lf(b(x), t).backward()
    for p, gr in zip(b.parameters(), ref_grads):
        assert torch.equal(p.grad, gr)                       # same kernels, same order: bit-identical
    opt.step()
    ema.update(b, 1)
    for pa, pb in zip(a.parameters(), b.parameters()):
        assert rel(pb, pa) < 1e-6
    for (ka, va), (kb, vb) in zip(b.state_dict().items(), ema.ema.state_dict().items()):
        assert ka == kb and torch.equal(va, vb)


def test_unknown_names_raise_like_the_reference():
    import medical_segmentation_pytorch_b200 as msx
    with pytest.raises(NotImplementedError):
        msx.get_model(Cfg(model='resnet'))
    with pytest.raises(NotImplementedError):
        msx.get_loss_fn(Cfg(loss_type='focal'), DEV)
    with pytest.raises(ValueError):
        msx.get_model(Cfg(model='unet', use_aux=True))
    with pytest.raises(NotImplementedError):
        msx.modules.Activation('swishh')


def test_kd_training_step_and_eval_sweep():
    """BASELINE config 5 shape of work (teacher logits -> KD + CE student step, then a Dice/mIoU sweep over
    val_bs = 1 updates, core/seg_trainer.py:69-75,114-125) against the oracle on the same tensors."""
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200.synth import synth_batch
    from oracle import ref_loss, ref_metrics
    torch.manual_seed(5)
    cfg = Cfg(kd_loss_type='kl_div', kd_temperature=4.0)
    student = msx.UNet(2, 3, 8).to(DEV).train()
    x, t = synth_batch(2, 64, 64, seed=9, ignore_band=0.02)
    x, t = x.to(DEV), t.to(DEV)
    teacher_logits = torch.randn(2, 2, 64, 64, device=DEV) * 2          # stand-in for teacher(images).detach()
    preds = student(x)
    loss = msx.get_loss_fn(cfg, DEV)(preds, t)
    loss_kd = msx.kd_loss_fn(cfg, preds, teacher_logits)
    total = loss + 1.0 * loss_kd                                          # kd_loss_coefficient = 1.0 (base_config.py:103)
    total.backward()
    pd = preds.detach().double().cpu().requires_grad_(True)
    ref = ref_loss.ce_loss(pd, t.cpu()) + ref_loss.kd_loss(pd, teacher_logits.double().cpu(), 'kl_div', 4.0)
    ref.backward()
    assert abs(float(total) - float(ref)) < 1e-5 * abs(float(ref))
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in student.parameters())
    # eval sweep: one image per update into both metrics
    student.eval()
    dice, iou = msx.get_seg_metrics(cfg, 'dice').to(DEV), msx.get_seg_metrics(cfg, 'iou').to(DEV)
    cm_iou = np.zeros((2, 2), dtype=np.int64)
    cm_dice = np.zeros((2, 2), dtype=np.int64)
    t_clean = torch.where(t == 255, torch.zeros_like(t), t)
    with torch.no_grad():
        for i in range(2):
            lg = student(x[i:i + 1])
            iou.update(lg, t[i:i + 1])
            dice.update(lg, t_clean[i:i + 1])
            cm_iou += ref_metrics.confusion_matrix(lg.cpu().numpy(), t[i:i + 1].cpu().numpy(), 2, 255)
            cm_dice += ref_metrics.confusion_matrix(lg.cpu().numpy(), t_clean[i:i + 1].cpu().numpy(), 2, None)
    assert np.array_equal(iou.confmat.cpu().numpy(), cm_iou) and np.array_equal(dice.confmat.cpu().numpy(), cm_dice)
    assert np.array_equal(iou.compute().cpu().numpy(), ref_metrics.iou_per_class(cm_iou))
    assert float(dice.compute()) == float(ref_metrics.dice_macro(cm_dice))


def test_unet32_fp32_config0_shape_runs_and_is_consistent():
    """BASELINE config 0 geometry (UNet-32, 320x320, fp32) at batch 2: train step + eval; fp32 and bf16 paths of
    the same weights agree on the loss to the stated bf16 tolerance, and every gradient is finite."""
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200.synth import synth_batch
    torch.manual_seed(1)
    m = msx.UNet(2, 3, 32).to(DEV).train()
    x, t = synth_batch(2, 320, 320, seed=1)
    x, t = x.to(DEV), t.to(DEV)
    lf = msx.get_loss_fn(Cfg(), DEV)
    l32 = lf(m(x), t)
    l32.backward()
    g32 = [p.grad.clone() for p in m.parameters()]
    with torch.autocast('cuda', dtype=torch.bfloat16):
        l16 = lf(m(x), t)
    assert abs(float(l16) - float(l32)) < 5e-2 * abs(float(l32))
    assert all(torch.isfinite(g).all() for g in g32)
    want = torch.nn.functional.cross_entropy(m(x).detach(), t, ignore_index=255)
    assert torch.isfinite(want)


@pytest.mark.gpu
@pytest.mark.parametrize('bf16', [False, True])
def test_ducknet_branch_streams_bit_identical(bf16):
    """Sibling DUCK branches and weight gradients on side streams (ops.run_branches, ops._wgrad_stream) vs one stream:
    same loss, logits and gradients, bit for bit,
    over two consecutive steps (the second reuses freed blocks of the first)."""
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200 import ops
    dev = torch.device('cuda')
    res = []
    try:
        for on in (0, 1):
            ops.set_branch_streams(on)
            ops.set_wgrad_stream(on)
            torch.manual_seed(5)
            m = msx.DuckNet(num_class=2, base_channel=8).to(dev).train()
            x = torch.randn(2, 3, 64, 64, device=dev)
            t = torch.randint(0, 2, (2, 64, 64), device=dev)
            loss_fn = msx.get_loss_fn(Cfg(), dev)
            outs = []
            for _ in range(2):
                for p in m.parameters():
                    p.grad = None
                with torch.autocast('cuda', dtype=torch.bfloat16, enabled=bf16):
                    y = m(x)
                loss = loss_fn(y, t)
                loss.backward()
                torch.cuda.synchronize()
                outs.append((loss.detach().clone(), y.detach().clone(), [p.grad.clone() for p in m.parameters()],
                             [b.clone() for b in m.buffers()]))
            res.append(outs)
    finally:
        ops.set_branch_streams(1)
        ops.set_wgrad_stream(1)
    for a, b in zip(res[0], res[1]):
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
        assert all(torch.equal(g, h) for g, h in zip(a[2], b[2]))
        assert all(torch.equal(g, h) for g, h in zip(a[3], b[3]))


@pytest.mark.parametrize('arch,base,dtype', [('unet', 8, torch.float32), ('ducknet', 4, torch.float32), ('unet', 8, torch.bfloat16), ('ducknet', 17, torch.bfloat16)])
def test_eval_sweep_fused_head_argmax_confmat_is_bit_exact(arch, base, dtype):
    """core/seg_trainer.py:100-125 without logits in HBM: head 1x1 + argmax + confusion matrix in one kernel, images batched.
    The counts must equal - exactly - those of the reference's own order of work (one image per update, logits
    materialised, metric.update(logits, mask)) on the same eval-mode model, for both metrics, with ignore labels; and the
    CPU oracle's confusion matrix of the logits path."""
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200.synth import synth_batch
    from oracle import ref_metrics
    torch.manual_seed(11)
    model = (msx.UNet if arch == 'unet' else msx.DuckNet)(2, 3, base).to(DEV)
    with torch.no_grad():
        for m in model.modules():                      # non-trivial running statistics, as after training
            if isinstance(m, torch.nn.BatchNorm2d):
                m.running_mean.uniform_(-0.2, 0.2)
                m.running_var.uniform_(0.6, 1.4)
    model.eval()
    model.compute_dtype = dtype
    x, t = synth_batch(5, 64, 96, seed=8, ignore_band=0.04)
    x, t = x.to(DEV), t.to(DEV)
    t_clean = torch.where(t == 255, torch.zeros_like(t), t)
    cfg = Cfg()
    ref_iou, ref_dice = msx.get_seg_metrics(cfg, 'iou').to(DEV), msx.get_seg_metrics(cfg, 'dice').to(DEV)
    cm = np.zeros((2, 2), dtype=np.int64)
    with torch.no_grad():
        for i in range(x.shape[0]):                    # the reference's loop: val_bs = 1, logits -> metric.update
            lg = model(x[i:i + 1])
            ref_iou.update(lg, t[i:i + 1])
            ref_dice.update(lg, t_clean[i:i + 1])
            cm += ref_metrics.confusion_matrix(lg.cpu().numpy(), t[i:i + 1].cpu().numpy(), 2, 255)
    assert np.array_equal(ref_iou.confmat.cpu().numpy(), cm)
    for bs in (1, 2, 16):
        iou, dice = msx.get_seg_metrics(cfg, 'iou').to(DEV), msx.get_seg_metrics(cfg, 'dice').to(DEV)
        msx.eval_sweep(model, x, t, [iou], batch_size=bs)
        msx.eval_sweep(model, x, t_clean, [dice], batch_size=bs)
        assert torch.equal(iou.confmat, ref_iou.confmat), (bs, iou.confmat, ref_iou.confmat)
        assert torch.equal(dice.confmat, ref_dice.confmat), bs
        assert torch.equal(iou.compute(), ref_iou.compute()) and torch.equal(dice.compute(), ref_dice.compute())
    assert int(ref_iou.confmat.sum()) == int((t != 255).sum())
    bad = msx.get_seg_metrics(cfg, 'dice').to(DEV)      # Dice has no ignore_index: a 255 label is an error at compute()
    msx.eval_sweep(model, x, t, [bad])
    with pytest.raises(ValueError):
        bad.compute()
    with pytest.raises(ValueError):
        msx.eval_sweep(model.train(), x, t, [iou])


def _syncbn_worker(rank, world, port, q):
    """Two processes on ONE GPU (gloo carries the small all-reduces): each runs its shard through a SyncBatchNorm'ed
    ConvBNAct -> DeConvBNAct -> ResidualBlock chain and returns outputs, input gradients, parameter gradients and buffers."""
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        import medical_segmentation_pytorch_b200 as msx
        from medical_segmentation_pytorch_b200.ducknet import ResidualBlock
        from medical_segmentation_pytorch_b200.modules import ConvBNAct, DeConvBNAct
        from gpu_util import from_act, to_act
        torch.manual_seed(3)
        net = torch.nn.Sequential(ConvBNAct(5, 17, 3), DeConvBNAct(17, 12), ResidualBlock(12, 20, 'relu')).to(DEV)
        msx.convert_sync_batchnorm(net)
        net.train()
        g = torch.Generator().manual_seed(100 + rank)
        n = 2 + rank                                                    # unequal shard sizes: the count travels with the sums
        x = torch.randn(n, 5, 12, 16, generator=g)
        dz = torch.randn(n, 20, 24, 32, generator=g)
        a = to_act(x, torch.float32, requires_grad=True)
        out = net(a)
        out.t.backward(to_act(dz, torch.float32).t)
        torch.cuda.synchronize()
        # numpy (pickled by value): torch tensors travel as file descriptors that die with this process
        res = {'x': x.numpy(), 'dz': dz.numpy(), 'z': from_act(out.t.detach(), 20).numpy(), 'dx': from_act(a.t.grad, 5).numpy(),
               'grads': {k: p.grad.detach().cpu().numpy() for k, p in net.named_parameters()},
               'bufs': {k: b.detach().cpu().numpy() for k, b in net.named_buffers()}}
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


def test_sync_batchnorm_world2_matches_one_process_on_the_whole_batch():
    """synBN=True (the reference's default, configs/base_config.py:91; utils/parallel.py:37-38): batch statistics over the
    samples of all ranks.  Oracle: the same layers with plain nn.BatchNorm2d in float64 in ONE process on the concatenated
    batch - outputs, input gradients and running statistics must match per shard; weight / bias gradients of every layer
    are per-rank partial sums whose total is the single-process gradient (DDP then averages them, as it does for torch's
    SyncBatchNorm)."""
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + (os.getpid() * 7) % 2000
    procs = [ctx.Process(target=_syncbn_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    tt = lambda d: {k: (tt(v) if isinstance(v, dict) else torch.from_numpy(np.asarray(v))) for k, v in d.items()}   # noqa: E731
    got = {r: tt(v) for r, v in got.items()}
    torch.manual_seed(3)
    from medical_segmentation_pytorch_b200.ducknet import ResidualBlock
    from medical_segmentation_pytorch_b200.modules import ConvBNAct, DeConvBNAct
    mine = torch.nn.Sequential(ConvBNAct(5, 17, 3), DeConvBNAct(17, 12), ResidualBlock(12, 20, 'relu'))   # same seed: same weights
    sd = {k: v.double() if v.is_floating_point() else v for k, v in mine.state_dict().items()}

    def bn(x, pfx, relu=True):
        y = torch.nn.functional.batch_norm(x, sd[pfx + '.running_mean'], sd[pfx + '.running_var'], P[pfx + '.weight'], P[pfx + '.bias'], True, 0.1, 1e-5)
        return torch.relu(y) if relu else y
    P = {k: v.clone().requires_grad_(True) for k, v in sd.items() if v.is_floating_point() and 'running' not in k}
    x = torch.cat([got[0]['x'], got[1]['x']]).double().requires_grad_(True)
    F = torch.nn.functional
    h = bn(F.conv2d(x, P['0.0.weight'], None, 1, 1), '0.1')
    h = bn(F.conv_transpose2d(h, P['1.up_conv.0.weight'], P['1.up_conv.0.bias'], 2, 1, 1), '1.up_conv.1')
    up = F.conv2d(h, P['2.upper_branch.weight'])
    low = bn(F.conv2d(bn(F.conv2d(h, P['2.lower_branch.0.0.weight'], None, 1, 1), '2.lower_branch.0.1'), P['2.lower_branch.1.0.weight'], None, 1, 1), '2.lower_branch.1.1')
    z = bn(up + low, '2.bn.0')
    z.backward(torch.cat([got[0]['dz'], got[1]['dz']]).double())
    n0 = got[0]['x'].shape[0]
    for r, sl in ((0, slice(0, n0)), (1, slice(n0, None))):
        assert rel(got[r]['z'], z[sl]) < 1e-5 and rel(got[r]['dx'], x.grad[sl]) < 1e-5, r
        for k in got[r]['bufs']:
            if 'running' in k:
                assert rel(got[r]['bufs'][k], sd[k]) < 1e-5, (r, k)          # F.batch_norm updated sd in place
            else:
                assert int(got[r]['bufs'][k]) == 1
    for k, p in P.items():
        total = got[0]['grads'][k].double() + got[1]['grads'][k].double()
        if float(p.grad.norm()) > 1e-9:
            assert rel(total, p.grad) < 2e-5, k


def _ddp_worker(rank, world, port, q):
    """Two ranks on ONE GPU over gloo: the real CUDA backward (side streams, arena gradients, bucket-ready callbacks from
    the kernels' Python shims) with the bucketed all-reduce, against plain per-rank gradients averaged explicitly."""
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        import medical_segmentation_pytorch_b200 as msx
        import medical_segmentation_pytorch_b200.parallel as par
        from medical_segmentation_pytorch_b200.flat import FlatParams
        torch.manual_seed(1 + rank)                          # different init per rank: the wrap must broadcast rank 0's
        a = msx.DuckNet(num_class=2, n_channel=3, base_channel=4).to(DEV).train()
        b = msx.DuckNet(num_class=2, n_channel=3, base_channel=4).to(DEV).train()
        FlatParams(a)
        old = par.compute_buckets
        par.compute_buckets = lambda sizes: old(sizes, 16 * 1024, 128 * 1024)       # many buckets on a small model
        ddp = msx.DistributedDataParallelB200(a)
        par.compute_buckets = old
        b.load_state_dict(a.state_dict())
        loss_fn = msx.get_loss_fn(Cfg(), DEV)
        bad, nb = [], len(ddp.buckets)
        for step in range(4):
            bf16 = step % 2 == 1
            g = torch.Generator().manual_seed(100 * step + rank)
            x = torch.randn(2, 3, 64, 64, generator=g).to(DEV)
            t = torch.randint(0, 2, (2, 64, 64), generator=g).to(DEV)
            for p in b.parameters():
                p.grad = None
            with torch.autocast('cuda', dtype=torch.bfloat16, enabled=bf16):
                loss_fn(ddp(x), t).backward()
                loss_fn(b(x), t).backward()
            torch.cuda.synchronize()
            for (n, pa), pb in zip(a.named_parameters(), b.parameters()):
                ref = pb.grad.detach().cpu()
                dist.all_reduce(ref, op=dist.ReduceOp.SUM)
                ref = (ref / world).to(DEV)
                if not torch.allclose(pa.grad, ref, rtol=1e-6, atol=1e-9):
                    bad.append((step, n, float((pa.grad - ref).abs().max())))
            for ba, bb in zip(a.buffers(), b.buffers()):
                bb.copy_(ba)
        w = [p.detach().float().sum().item() for p in a.parameters()][:8]
        q.put((rank, bad[:5], nb, w))
    finally:
        dist.destroy_process_group()


def test_ddp_bucketed_allreduce_world2_on_cuda_kernels():
    """DistributedDataParallelB200 with the CUDA backward (the CPU test in test_host_cpu.py drives the bucket logic with
    torch CPU modules): two ranks, >= 5 buckets, fp32 and bf16 steps; the arena gradients after backward equal the
    per-rank gradients averaged by one explicit all-reduce (sum / world: same rounding as the gloo path, within 1 ulp
    of NCCL's AVG), parameters were broadcast from rank 0, every bucket is re-armed."""
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + (os.getpid() * 13 + 5) % 2000
    procs = [ctx.Process(target=_ddp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict((r, (bad, nb, w)) for r, bad, nb, w in (q.get(timeout=300) for _ in procs))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert got[0][0] == [] and got[1][0] == [], (got[0][0], got[1][0])
    assert got[0][1] >= 5 and got[0][2] == got[1][2]


@pytest.mark.parametrize('kind', ['sgd', 'adam'])
def test_capture_step_follows_the_lr_schedule(kind):
    """msx.capture_step: the reference's step body as one CUDA graph.  Scalars a kernel receives by value are frozen into a
    graph, so the fused optimizer reads lr / momentum / betas from a device array that a graph node refreshes from pinned
    host memory at every replay: a replay after `param_groups[0]['lr'] = 0` must leave the parameters untouched, and a
    replay with the lr restored must move them by exactly what an eager step with that lr does."""
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200.synth import synth_batch
    torch.manual_seed(4)
    a = msx.UNet(2, 3, 8).to(DEV).train()
    b = msx.UNet(2, 3, 8).to(DEV).train()
    b.load_state_dict(a.state_dict())
    x, t = synth_batch(2, 32, 32, seed=2)
    x, t = x.to(DEV), t.to(DEV)
    lf = msx.get_loss_fn(Cfg(), DEV)
    opts = [msx.FusedOptimizer(m, kind, lr=0.05) for m in (a, b)]

    def make_step(m, o):
        def step():
            o.zero_grad()
            loss = lf(m(x), t)
            loss.backward()
            o.step()
            return loss
        return step
    step_a, step_b = make_step(a, opts[0]), make_step(b, opts[1])
    replay = msx.capture_step(step_a, warmup=2, optimizers=[opts[0]])        # 2 eager steps + the capture pass (a no-op on data)
    for _ in range(2):
        step_b().detach()
    flat = lambda m: torch.cat([p.detach().flatten() for p in m.parameters()])   # noqa: E731
    assert rel(flat(a), flat(b)) < 1e-6                                      # capture itself does not execute the step
    for lr in (0.05, 0.0, 0.02):
        opts[0].param_groups[0]['lr'] = lr
        opts[1].param_groups[0]['lr'] = lr
        before = flat(a).clone()
        loss = replay()
        step_b().detach()
        torch.cuda.synchronize()
        assert torch.isfinite(loss).item()
        if lr == 0.0:
            assert torch.equal(flat(a), before)
        else:
            assert not torch.equal(flat(a), before)
        assert rel(flat(a), flat(b)) < 2e-5, lr
