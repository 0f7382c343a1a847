"""The distillation teacher (smp.Unet over a ResNet, inference only) on the CUDA path vs the CPU oracle
(oracle/ref_teacher.py: torchvision's ResNet + smp 0.3.2's decoder restated; parity unpinned, see its header)."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gpu_util import DEV, Cfg, rel  # noqa: E402

pytestmark = pytest.mark.gpu


def _pair(enc, seed=0):
    import medical_segmentation_pytorch_b200 as msx
    from oracle.ref_teacher import SmpUnetRef, randomize_bn
    torch.manual_seed(seed)
    ref = SmpUnetRef(enc, 3, 2).eval()
    randomize_bn(ref, seed)
    new = msx.SmpUnet(enc, None, 3, 2)
    new.load_state_dict(ref.state_dict())
    return ref, new.to(DEV).eval()


@pytest.mark.parametrize('enc,size', [('resnet18', 64), ('resnet50', 64), ('resnet101', 96)])
def test_teacher_fp32_matches_oracle(enc, size):
    ref, new = _pair(enc)
    x = torch.randn(2, 3, size, size)
    with torch.no_grad():
        y64 = ref.double()(x.double())
        y32 = ref.float()(x)
    y = new(x.to(DEV)).cpu()
    assert y.shape == y32.shape and y.dtype == torch.float32
    e_new, e_ref = rel(y, y64), rel(y32, y64)
    print(f'teacher {enc} fp32: new {e_new:.2e} reference-fp32 {e_ref:.2e}')
    assert e_new <= max(3 * e_ref, 2e-5)          # SURVEY 4.3 protocol: no worse than 3x the reference's own fp32 error


@pytest.mark.parametrize('enc,size', [('resnet18', 64), ('resnet101', 96)])
def test_teacher_bf16_autocast(enc, size):
    """bf16 (tcgen05 convs incl. the 7x7 s2 stem, 3x3 s2 and the 1x1 s2 shortcuts run as 2x2 s2) vs fp64, bounded by the
    error of the oracle's own bf16-autocast run on the CPU."""
    ref, new = _pair(enc, 1)
    x = torch.randn(2, 3, size, size)
    with torch.no_grad():
        y64 = ref.double()(x.double())
        ref.float()
        with torch.autocast('cpu', dtype=torch.bfloat16):
            yb = ref(x).float()
    with torch.autocast('cuda', dtype=torch.bfloat16):
        y = new(x.to(DEV)).cpu()
    e_new, e_ref = rel(y, y64), rel(yb, y64)
    print(f'teacher {enc} bf16: new {e_new:.2e} reference-bf16 {e_ref:.2e}')
    assert e_new <= max(1.5 * e_ref, 2e-2)


def test_teacher_is_inference_only_and_checks_shapes():
    import medical_segmentation_pytorch_b200 as msx
    m = msx.SmpUnet('resnet18', None, 3, 2).to(DEV)
    with pytest.raises(RuntimeError):
        m(torch.randn(1, 3, 40, 64, device=DEV))
    with pytest.raises(NotImplementedError):
        m.train()(torch.randn(1, 3, 64, 64, device=DEV))
    with pytest.raises(ValueError):
        msx.SmpUnet('mobilenet_v2')


def test_kd_step_with_real_teacher(tmp_path):
    """BASELINE config 5: teacher checkpoint -> get_teacher_model -> frozen forward under no_grad -> CE + KD(kl, T=4) on a
    UNet student, one training step; the KD term equals the oracle's kd loss on the same logits."""
    import medical_segmentation_pytorch_b200 as msx
    from oracle import ref_loss
    from oracle.ref_teacher import SmpUnetRef, randomize_bn
    torch.manual_seed(2)
    ref = SmpUnetRef('resnet18', 3, 2).eval()
    randomize_bn(ref, 2)
    ckpt = os.path.join(tmp_path, 'teacher.pth')
    torch.save({'state_dict': ref.state_dict()}, ckpt)
    cfg = Cfg(kd_training=True, teacher_ckpt=ckpt, teacher_encoder='resnet18', teacher_decoder='unet', num_channel=3,
              kd_loss_type='kl_div', kd_temperature=4.0)
    teacher = msx.get_teacher_model(cfg, DEV)
    student = msx.UNet(num_class=2, base_channel=8).to(DEV).train()
    x = torch.randn(2, 3, 64, 64, device=DEV)
    t = torch.randint(0, 2, (2, 64, 64), device=DEV)
    preds = student(x)
    loss = msx.get_loss_fn(cfg, DEV)(preds, t)
    with torch.no_grad():
        tp = teacher(x)
    kd = msx.kd_loss_fn(cfg, preds, tp.detach())
    (loss + kd).backward()
    with torch.no_grad():
        tp_ref = ref.float()(x.cpu())
    assert rel(tp.cpu(), tp_ref) < 1e-4
    kd_ref = ref_loss.kd_loss(preds.detach().cpu().double(), tp.cpu().double(), 'kl_div', 4.0)
    assert abs(float(kd) - float(kd_ref)) < 1e-5 * max(1.0, abs(float(kd_ref)))
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in student.parameters())
    assert all(p.grad is None for p in teacher.parameters())
