"""Generate tests/golden/*.npz by running the REAL reference modules from /root/reference.

Run in the build container only (the reference is not present on the GPU box):

    python oracle/gen_golden.py            # UNet-8 / DuckNet-3 / losses / parameter counts
    python oracle/gen_golden.py --wide     # DuckNet-17 and -34 (the benchmark's channel widths)

The reference's ``models`` package imports ``segmentation_models_pytorch`` at module top
(models/__init__.py:2), which is not installed; a stub is placed in ``sys.modules`` so that
``models.unet`` / ``models.ducknet`` import unmodified.  ``core/loss.py`` is loaded by file path.
Weights come from ``oracle.ref_models.synth_state_dict`` (numpy PCG64 - independent of the torch
version) and are loaded into the reference modules through ``load_state_dict``; inputs are stored
in the fixture, so the tests never need the reference again.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get('MSX_REFERENCE', '/root/reference')
sys.path.insert(0, ROOT)

from oracle import ref_models  # noqa: E402


def load_reference():
    stub = types.ModuleType('segmentation_models_pytorch')
    for n in ('DeepLabV3', 'DeepLabV3Plus', 'FPN', 'Linknet', 'MAnet', 'PAN', 'PSPNet', 'Unet', 'UnetPlusPlus'):
        setattr(stub, n, None)
    sys.modules['segmentation_models_pytorch'] = stub
    sys.path.insert(0, REF)
    from models.unet import UNet            # noqa
    from models.ducknet import DuckNet      # noqa
    spec = importlib.util.spec_from_file_location('ref_core_loss', os.path.join(REF, 'core', 'loss.py'))
    loss_mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(loss_mod)
    sys.path.remove(REF)
    return UNet, DuckNet, loss_mod


def synth_batch(rng, n, h, w, n_class=2, ignore_frac=0.0):
    img = np.clip(rng.standard_normal((n, 3, h, w)), -2.12, 2.64).astype(np.float32)
    yy, xx = np.mgrid[0:h, 0:w]
    mask = np.zeros((n, h, w), dtype=np.int64)
    for i in range(n):
        cy, cx = rng.uniform(0.2, 0.8) * h, rng.uniform(0.2, 0.8) * w
        ay, ax = rng.uniform(0.15, 0.35) * h, rng.uniform(0.15, 0.35) * w
        mask[i] = (((yy - cy) / ay) ** 2 + ((xx - cx) / ax) ** 2 <= 1.0).astype(np.int64) * (n_class - 1)
    if ignore_frac > 0:
        mask[rng.uniform(size=mask.shape) < ignore_frac] = 255
    return img, mask


class Cfg:
    pass


def model_case(ref_cls, name, base, n, hw, seed, loss_mod, out, light=False):
    rng = np.random.Generator(np.random.PCG64(seed))
    img, mask = synth_batch(rng, n, hw, hw, 2, ignore_frac=0.03)
    layout = ref_models.model_layout(name, num_class=2, n_channel=3, base_channel=base)
    sd = ref_models.synth_state_dict(layout, seed + 100)
    res = {'images': img, 'masks': mask, 'base_channel': np.int64(base), 'weight_seed': np.int64(seed + 100)}
    cfg = Cfg()
    cfg.class_weights, cfg.loss_type, cfg.ignore_index, cfg.reduction = None, 'ce', 255, 'mean'
    for tag, dt in (('f32', torch.float32), ('f64', torch.float64)):
        torch.manual_seed(0)
        model = ref_cls(num_class=2, n_channel=3, base_channel=base)
        assert list(model.state_dict().keys()) == list(layout.keys()), 'oracle layout != reference state_dict'
        for k, v in model.state_dict().items():
            assert tuple(v.shape) == tuple(layout[k]), (k, v.shape, layout[k])
        model.load_state_dict(sd)
        model = model.to(dt).train()
        loss_fn = loss_mod.get_loss_fn(cfg, torch.device('cpu'))
        x = torch.from_numpy(img).to(dt)
        logits = model(x)
        loss = loss_fn(logits, torch.from_numpy(mask))
        res[f'logits_{tag}'] = logits.detach().numpy()
        res[f'loss_{tag}'] = loss.detach().numpy()
        if light:                      # forward only (the 160 M-parameter base-34 model)
            continue
        loss.backward()
        gn = []
        for k, p in model.named_parameters():
            gn.append(p.grad.norm().item())
        res[f'grad_norms_{tag}'] = np.asarray(gn, dtype=np.float64)
        named = dict(model.named_parameters())
        keep = ['seg_head.weight', list(named.keys())[0], list(named.keys())[1], list(named.keys())[2]]
        for k in keep:
            res[f'grad_{tag}::{k}'] = named[k].grad.detach().numpy()
        st = model.state_dict()
        bn_keys = [k for k in st if k.endswith('running_var')]
        for k in (bn_keys[0], bn_keys[len(bn_keys) // 2], bn_keys[-1]):
            res[f'buf_{tag}::{k}'] = st[k].numpy().copy()
            km = k.replace('running_var', 'running_mean')
            res[f'buf_{tag}::{km}'] = st[km].numpy().copy()
        # eval-mode forward with the updated running stats (validation path, core/seg_trainer.py:114)
        model.eval()
        with torch.no_grad():
            res[f'eval_logits_{tag}'] = model(x).numpy()
    # the reference's AMP switch (core/seg_trainer.py:61) with dtype=bfloat16 on CPU: the bf16 noise floor
    torch.manual_seed(0)
    model = ref_cls(num_class=2, n_channel=3, base_channel=base)
    model.load_state_dict(sd)
    model.train()
    loss_fn = loss_mod.get_loss_fn(cfg, torch.device('cpu'))
    with torch.autocast('cpu', dtype=torch.bfloat16):
        lg = model(torch.from_numpy(img))
        ls = loss_fn(lg, torch.from_numpy(mask))
    res['logits_bf16'] = lg.detach().float().numpy()
    res['loss_bf16'] = ls.detach().float().numpy()
    res['param_names'] = np.asarray([k for k, _ in model.named_parameters()])
    np.savez_compressed(out, **res)
    print('wrote', out, {k: getattr(v, 'shape', None) for k, v in res.items() if k.startswith('logits')})


def _rel(a, b):
    return float((a.detach().double() - b.detach().double()).norm() / max(float(b.detach().double().norm()), 1e-300))


def stage_case(ref_cls, base, n, hw, seed, out, backward=True):
    """Teacher-forced block-level yardstick for bf16: every DuckNet block of the REAL reference
    is run alone from the fp64-truth inputs of that block (rounded to bf16), once in fp64 (truth) and once under the
    reference's AMP switch with dtype=bfloat16 (core/seg_trainer.py:61; backward outside autocast, :82).  The
    recorded errors of the reference against its own fp64 twin are what the CUDA path is held to in
    tests/test_gpu_models.py - whole-network bf16 logits cannot serve (the reference disagrees with itself by
    ~0.7 rel-L2 there: stacked train-mode BN amplifies every rounding ~50x)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    img, _ = synth_batch(rng, n, hw, hw, 2, ignore_frac=0.03)
    layout = ref_models.model_layout('ducknet', num_class=2, n_channel=3, base_channel=base)
    sd = ref_models.synth_state_dict(layout, seed + 100)
    m64 = ref_cls(num_class=2, n_channel=3, base_channel=base)
    m64.load_state_dict(sd)
    m64 = m64.double().train()
    m32 = ref_cls(num_class=2, n_channel=3, base_channel=base)
    m32.load_state_dict(sd)
    m32.train()
    # block inputs of the fp64 forward, via the oracle restatement (pinned to the reference by the whole-model fixture)
    plan, _ = ref_models.FnNet('ducknet', ref_models.make_trainable(sd, torch.float64), training=True).ducknet_stages(
        torch.from_numpy(img).double())

    def module_of(model, prefix):
        mod = model
        for part in prefix.replace('.duck', '').split('.'):
            mod = mod[int(part)] if part.isdigit() else getattr(mod, part)
        return mod

    def call(mod, kind, ins):
        if kind == 'down1':
            return mod(ins[0])
        if kind == 'down':
            return mod(ins[0] + ins[1], ins[1])
        if kind == 'res':
            return (mod(ins[0]),)
        return (mod(ins[0], ins[1]),)

    names, fwd, din, dp_med, dp_max, dp_glob = [], [], [], [], [], []
    for idx, (prefix, kind, ins) in enumerate(plan):
        ins_r = [t.float().bfloat16().float() for t in ins]
        mod64, mod32 = module_of(m64, prefix), module_of(m32, prefix)
        i64 = [t.double().requires_grad_(kind != 'down1') for t in ins_r]
        o64 = call(mod64, kind, i64)
        douts = [ref_models.stage_dout(idx, k, o.shape) for k, o in enumerate(o64)]
        i32 = [t.clone().requires_grad_(kind != 'down1') for t in ins_r]
        with torch.autocast('cpu', dtype=torch.bfloat16):
            o32 = call(mod32, kind, i32)
        names.append(prefix)
        fwd.append(max(_rel(a, b) for a, b in zip(o32, o64)))
        if not backward:
            continue
        m64.zero_grad()
        m32.zero_grad()
        sum((o * d.double()).sum() for o, d in zip(o64, douts)).backward()
        sum((o.float() * d).sum() for o, d in zip(o32, douts)).backward()
        din.append(max([_rel(a.grad, b.grad) for a, b in zip(i32, i64) if b.grad is not None] or [0.0]))
        g64 = [p.grad for p in mod64.parameters()]
        g32 = [p.grad for p in mod32.parameters()]
        gmax = max(float(g.norm()) for g in g64)
        errs = np.array([_rel(a, b) for a, b in zip(g32, g64) if float(b.norm()) > 1e-9 * gmax])
        dp_med.append(float(np.median(errs)))
        dp_max.append(float(errs.max()))
        dp_glob.append(_rel(torch.cat([g.flatten() for g in g32]), torch.cat([g.flatten() for g in g64])))
        print(f'{prefix:18s} fwd {fwd[-1]:.3e} din {din[-1]:.3e} dparam global {dp_glob[-1]:.3e} med {dp_med[-1]:.3e} max {dp_max[-1]:.3e}', flush=True)
    res = {'stage_names': np.asarray(names), 'ref_bf16_fwd': np.asarray(fwd), 'base_channel': np.int64(base),
           'weight_seed': np.int64(seed + 100), 'data_seed': np.int64(seed), 'n': np.int64(n), 'hw': np.int64(hw)}
    if backward:
        res.update({'ref_bf16_din': np.asarray(din), 'ref_bf16_dparam_median': np.asarray(dp_med), 'ref_bf16_dparam_max': np.asarray(dp_max),
                    'ref_bf16_dparam_global': np.asarray(dp_glob)})
    np.savez_compressed(out, **res)
    print('wrote', out, 'fwd errors', np.round(fwd, 4))


def loss_case(loss_mod, out):
    rng = np.random.Generator(np.random.PCG64(7))
    n, c, h, w = 2, 3, 16, 24
    z = (rng.standard_normal((n, c, h, w)) * 2.0).astype(np.float32)
    zt = (rng.standard_normal((n, c, h, w)) * 2.0).astype(np.float32)
    t = rng.integers(0, c, size=(n, h, w)).astype(np.int64)
    t[rng.uniform(size=t.shape) < 0.1] = 255
    res = {'logits': z, 'teacher': zt, 'labels': t}
    cfg = Cfg()
    cfg.ignore_index, cfg.reduction, cfg.loss_type = 255, 'mean', 'ce'

    def run(fn, key):
        zz = torch.from_numpy(z).double().requires_grad_(True)
        val = fn(zz)
        val.backward()
        res[key] = val.detach().numpy()
        res[key + '_grad'] = zz.grad.numpy()

    cfg.class_weights = None
    ce = loss_mod.get_loss_fn(cfg, torch.device('cpu'))
    run(lambda zz: ce(zz, torch.from_numpy(t)), 'ce')
    cfg.class_weights = [0.3, 1.0, 2.5]
    res['class_weights'] = np.asarray(cfg.class_weights, dtype=np.float32)
    cew = loss_mod.get_loss_fn(cfg, torch.device('cpu')).double()
    run(lambda zz: cew(zz, torch.from_numpy(t)), 'ce_weighted')

    # OHEM: the reference forward body, constructed without the .cuda() of core/loss.py:9
    def make_ohem(thresh):
        o = loss_mod.OhemCELoss.__new__(loss_mod.OhemCELoss)
        torch.nn.Module.__init__(o)
        o.thresh = -torch.log(torch.tensor(thresh, dtype=torch.float))
        o.ignore_index = 255
        o.criteria = torch.nn.CrossEntropyLoss(ignore_index=255, reduction='none')
        return o
    run(lambda zz: make_ohem(0.7)(zz, torch.from_numpy(t)), 'ohem_thresh_branch')       # many hard pixels
    run(lambda zz: make_ohem(1e-4)(zz, torch.from_numpy(t)), 'ohem_topk_branch')        # thresh ~9.2: top-k
    res['ohem_topk_thresh'] = np.float64(1e-4)

    cfg.kd_temperature = 4.0
    for kind in ('kl_div', 'mse'):
        cfg.kd_loss_type = kind
        run(lambda zz: loss_mod.kd_loss_fn(cfg, zz, torch.from_numpy(zt).double()), 'kd_' + kind)
    np.savez_compressed(out, **res)
    print('wrote', out)


def main():
    torch.set_num_threads(1)
    UNet, DuckNet, loss_mod = load_reference()
    g = os.path.join(ROOT, 'tests', 'golden')
    os.makedirs(g, exist_ok=True)
    if '--wide' in sys.argv:
        # the benchmark's own channel widths (17 * 2^k and 34 * 2^k): whole model + per-block bf16 yardstick
        torch.set_num_threads(os.cpu_count() or 1)
        if '--stages-only' not in sys.argv:
            model_case(DuckNet, 'ducknet', 17, 2, 128, 13, loss_mod, os.path.join(g, 'ducknet_b17_128.npz'))
        stage_case(DuckNet, 17, 2, 128, 13, os.path.join(g, 'ducknet_b17_128_stages.npz'))
        if '--stages-only' in sys.argv:
            return
        model_case(DuckNet, 'ducknet', 34, 2, 96, 14, loss_mod, os.path.join(g, 'ducknet_b34_96.npz'), light=True)
        stage_case(DuckNet, 34, 2, 96, 14, os.path.join(g, 'ducknet_b34_96_stages.npz'), backward=False)
        return
    model_case(UNet, 'unet', 8, 2, 32, 11, loss_mod, os.path.join(g, 'unet_b8_32.npz'))
    model_case(DuckNet, 'ducknet', 3, 2, 64, 12, loss_mod, os.path.join(g, 'ducknet_b3_64.npz'))
    loss_case(loss_mod, os.path.join(g, 'losses.npz'))
    # parameter-count pins from README.md:110-112
    counts = {}
    for name, cls, b in (('unet32', UNet, 32), ('ducknet17', DuckNet, 17), ('ducknet34', DuckNet, 34)):
        counts[name] = sum(p.numel() for p in cls(num_class=2, n_channel=3, base_channel=b).parameters())
    np.savez(os.path.join(g, 'param_counts.npz'), **{k: np.int64(v) for k, v in counts.items()})
    print(counts)


if __name__ == '__main__':
    main()
