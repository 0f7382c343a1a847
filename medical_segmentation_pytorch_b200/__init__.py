"""B200-native (sm_100a) training-step path for the UNet / DUCK-Net segmentation models of
AjayGaneshC/medical-segmentation-pytorch: drop-in model, loss, metric and data-parallel entry
points (same names, arguments and error behaviour as the reference) over libmsx.so.

Importing this package loads the CUDA library; it raises if the library is missing - there is no
PyTorch-eager or CPU fallback."""
from ._lib import MsxError, lib  # noqa: F401  (fails loudly when libmsx.so is absent)
from .ducknet import DuckNet
from .loss import CrossEntropyLoss, DiceBCELoss, OhemCELoss, get_loss_fn, kd_loss_fn
from .metrics import Dice, JaccardIndex, eval_sweep, get_seg_metrics
from .graph import CapturedStep, capture_step
from .optim import FusedOptimizer, ModelEma, get_ema_model, get_optimizer, get_scheduler
from .parallel import (DistributedDataParallelB200, convert_sync_batchnorm, de_parallel, destroy_ddp_process, is_parallel, parallel_model,
                       sampler_set_epoch, set_device)
from .teacher import SmpUnet
from .unet import UNet


def get_model(config):
    """models/__init__.py:13-39 for the two native architectures ('smp' students are out of scope)."""
    model_hub = {'unet': UNet, 'ducknet': DuckNet}
    if config.model == 'smp':
        raise NotImplementedError("model 'smp' (segmentation_models_pytorch students) is outside the B200 hot path")
    if config.model in model_hub:
        if getattr(config, 'use_aux', False):
            raise ValueError(f'Model {config.model} does not support auxiliary heads.\n')
        return model_hub[config.model](num_class=config.num_class, n_channel=config.num_channel,
                                       base_channel=config.base_channel)
    raise NotImplementedError(f"Unsupport model type: {config.model}")


def get_teacher_model(config, device):
    """models/__init__.py:42-62: the frozen distillation teacher, ``smp.Unet`` over a ResNet encoder (the only
    decoder/encoder family built here), loaded from ``config.teacher_ckpt['state_dict']`` and put in eval mode."""
    import os
    import torch
    if not config.kd_training:
        return None
    if not os.path.isfile(config.teacher_ckpt):
        raise ValueError(f'Could not find teacher checkpoint at path {config.teacher_ckpt}.')
    if config.teacher_decoder != 'unet':
        raise ValueError(f"Unsupported teacher decoder type: {config.teacher_decoder}")
    model = SmpUnet(encoder_name=config.teacher_encoder, encoder_weights=None, in_channels=config.num_channel,
                    classes=config.num_class)
    ckpt = torch.load(config.teacher_ckpt, map_location=torch.device('cpu'))
    model.load_state_dict(ckpt['state_dict'])
    del ckpt
    return model.to(device).eval()


__all__ = ['UNet', 'DuckNet', 'SmpUnet', 'get_model', 'get_teacher_model', 'get_loss_fn', 'kd_loss_fn', 'get_seg_metrics', 'eval_sweep', 'parallel_model',
           'de_parallel', 'convert_sync_batchnorm', 'is_parallel', 'set_device', 'destroy_ddp_process', 'sampler_set_epoch', 'get_optimizer',
           'get_ema_model', 'get_scheduler', 'capture_step', 'CapturedStep', 'FusedOptimizer', 'ModelEma', 'CrossEntropyLoss', 'OhemCELoss', 'DiceBCELoss',
           'JaccardIndex', 'Dice', 'DistributedDataParallelB200', 'MsxError']
