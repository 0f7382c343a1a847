"""Pins the CPU oracle (oracle/) against vectors produced by the real reference modules
(tests/golden/*.npz, written by oracle/gen_golden.py in the build container)."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_loss, ref_models


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.mark.parametrize('name,fname', [('unet', 'unet_b8_32.npz'), ('ducknet', 'ducknet_b3_64.npz')])
@pytest.mark.parametrize('fused_bn', [False, True])
def test_model_matches_reference_fp64(golden_dir, name, fname, fused_bn):
    g = np.load(os.path.join(golden_dir, fname))
    base = int(g['base_channel'])
    layout = ref_models.model_layout(name, num_class=2, n_channel=3, base_channel=base)
    sd = ref_models.synth_state_dict(layout, int(g['weight_seed']))
    p = ref_models.make_trainable(sd, torch.float64)
    net = ref_models.FnNet(name, p, training=True, fused_bn=fused_bn)
    x = torch.from_numpy(g['images']).double()
    logits = net(x)
    loss = ref_loss.ce_loss(logits, torch.from_numpy(g['masks']))
    loss.backward()
    assert rel(logits.detach().numpy(), g['logits_f64']) < 1e-10
    assert abs(loss.item() - float(g['loss_f64'])) < 1e-12 * max(1, abs(float(g['loss_f64'])))
    names = [str(s) for s in g['param_names']]
    assert names == ref_models.param_keys(sd)
    gn = np.array([p[k].grad.norm().item() for k in names])
    big = g['grad_norms_f64'] > 1e-9 * g['grad_norms_f64'].max()       # ConvT-bias grads are ~0
    assert np.max(np.abs(gn[big] / g['grad_norms_f64'][big] - 1)) < 1e-7
    for k in g.files:
        if k.startswith('grad_f64::'):
            assert rel(p[k.split('::')[1]].grad.numpy(), g[k]) < 1e-7, k
        if k.startswith('buf_f64::'):
            assert rel(p[k.split('::')[1]].detach().numpy(), g[k]) < 1e-10, k
    net.training = False
    with torch.no_grad():
        assert rel(net(x).numpy(), g['eval_logits_f64']) < 1e-10


def test_ducknet17_matches_reference_fp64(golden_dir):
    """The oracle at the benchmark's own channel widths (17 * 2^k): whole-model vectors from the real reference."""
    torch.set_num_threads(max(1, (os.cpu_count() or 2)))
    g = np.load(os.path.join(golden_dir, 'ducknet_b17_128.npz'))
    sd = ref_models.synth_state_dict(ref_models.model_layout('ducknet', num_class=2, n_channel=3, base_channel=17), int(g['weight_seed']))
    p = ref_models.make_trainable(sd, torch.float64)
    net = ref_models.FnNet('ducknet', p, training=True)
    x = torch.from_numpy(g['images']).double()
    plan, logits = net.ducknet_stages(x)                      # the block-cut forward IS the forward
    loss = ref_loss.ce_loss(logits, torch.from_numpy(g['masks']))
    loss.backward()
    assert rel(logits.detach().numpy(), g['logits_f64']) < 1e-10
    assert abs(loss.item() - float(g['loss_f64'])) < 1e-12
    names = [str(s) for s in g['param_names']]
    gn = np.array([p[k].grad.norm().item() for k in names])
    big = g['grad_norms_f64'] > 1e-9 * g['grad_norms_f64'].max()
    assert np.max(np.abs(gn[big] / g['grad_norms_f64'][big] - 1)) < 1e-6
    st = np.load(os.path.join(golden_dir, 'ducknet_b17_128_stages.npz'))
    assert [str(s) for s in st['stage_names']] == [pfx for pfx, _, _ in plan]
    # re-running one block alone from its recorded inputs reproduces what the whole forward fed the next block
    (p0, k0, ins0), (p1, k1, ins1) = plan[0], plan[1]
    with torch.no_grad():
        net2 = ref_models.FnNet('ducknet', ref_models.make_trainable(sd, torch.float64), training=True)
        x1, _, xc = net2.stage_fn(k0, p0)(*ins0)
    assert rel(x1.numpy(), ins1[0].numpy()) < 1e-12 and rel(xc.numpy(), ins1[1].numpy()) < 1e-12


def test_ducknet34_forward_matches_reference_fp64(golden_dir):
    torch.set_num_threads(max(1, (os.cpu_count() or 2)))
    g = np.load(os.path.join(golden_dir, 'ducknet_b34_96.npz'))
    sd = ref_models.synth_state_dict(ref_models.model_layout('ducknet', num_class=2, n_channel=3, base_channel=34), int(g['weight_seed']))
    with torch.no_grad():
        logits = ref_models.FnNet('ducknet', ref_models.make_trainable(sd, torch.float64), training=True)(torch.from_numpy(g['images']).double())
    assert rel(logits.numpy(), g['logits_f64']) < 1e-10


def test_param_counts_match_readme(golden_dir):
    g = np.load(os.path.join(golden_dir, 'param_counts.npz'))
    cases = {'unet32': ('unet', 32), 'ducknet17': ('ducknet', 17), 'ducknet34': ('ducknet', 34)}
    for key, (name, b) in cases.items():
        lay = ref_models.model_layout(name, num_class=2, n_channel=3, base_channel=b)
        n = sum(int(np.prod(s)) for k, s in lay.items()
                if not k.endswith(('running_mean', 'running_var', 'num_batches_tracked')))
        assert n == int(g[key])
    # README.md:110-112: 40.1 M / 160.28 M / 8.63 M
    assert round(int(g['ducknet17']) / 1e6, 1) == 40.1
    assert round(int(g['ducknet34']) / 1e6, 2) == 160.28
    assert round(int(g['unet32']) / 1e6, 2) == 8.63


def _grad(fn, z):
    zz = torch.from_numpy(z).double().requires_grad_(True)
    v = fn(zz)
    v.backward()
    return v.item(), zz.grad.numpy()


def test_losses_match_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, 'losses.npz'))
    z, zt, t = g['logits'], g['teacher'], torch.from_numpy(g['labels'])
    cases = {
        'ce': lambda zz: ref_loss.ce_loss(zz, t),
        'ce_weighted': lambda zz: ref_loss.ce_loss(zz, t, g['class_weights']),
        'ohem_thresh_branch': lambda zz: ref_loss.ohem_loss(zz, t, 0.7),
        'ohem_topk_branch': lambda zz: ref_loss.ohem_loss(zz, t, float(g['ohem_topk_thresh'])),
        'kd_kl_div': lambda zz: ref_loss.kd_loss(zz, torch.from_numpy(zt).double(), 'kl_div', 4.0),
        'kd_mse': lambda zz: ref_loss.kd_loss(zz, torch.from_numpy(zt).double(), 'mse'),
    }
    for key, fn in cases.items():
        v, gr = _grad(fn, z)
        assert abs(v - float(g[key])) < 1e-12 * max(1.0, abs(v)), key
        assert rel(gr, g[key + '_grad']) < 1e-10, key


def test_ce_all_ignored_is_nan():
    z = torch.zeros(1, 2, 4, 4)
    t = torch.full((1, 4, 4), 255, dtype=torch.int64)
    assert torch.isnan(ref_loss.ce_loss(z, t))


def test_dice_bce_is_the_stated_composite():
    torch.manual_seed(0)
    z = torch.randn(2, 1, 8, 8, dtype=torch.float64)
    t = (torch.rand(2, 1, 8, 8) > 0.7).double()
    want = torch.nn.functional.binary_cross_entropy_with_logits(z, t) + \
        (1 - (2 * (torch.sigmoid(z) * t).sum() + 1.0) / (torch.sigmoid(z).sum() + t.sum() + 1.0))
    assert abs(ref_loss.dice_bce_loss(z, t).item() - want.item()) < 1e-14


def test_teacher_oracle_param_counts_and_keys():
    """smp.Unet(resnet18) = 14.33 M parameters (reference README.md:113) - the one known answer the teacher restatement
    can be pinned by; resnet101 = 42,500,160 encoder + 9,012,928 decoder + 290 head."""
    import torch
    from oracle.ref_teacher import SmpUnetRef
    r18 = SmpUnetRef('resnet18', 3, 2)
    assert sum(p.numel() for p in r18.parameters()) == 14_328_354
    r101 = SmpUnetRef('resnet101', 3, 2)
    assert sum(p.numel() for p in r101.encoder.parameters()) == 42_500_160
    assert sum(p.numel() for p in r101.decoder.parameters()) == 9_012_928
    assert sum(p.numel() for p in r101.segmentation_head.parameters()) == 290
    with torch.no_grad():
        assert r18.eval()(torch.zeros(1, 3, 64, 64)).shape == (1, 2, 64, 64)
    keys = list(r101.state_dict().keys())
    assert keys[0] == 'encoder.conv1.weight' and 'decoder.blocks.0.conv1.0.weight' in keys and keys[-1] == 'segmentation_head.0.bias'
