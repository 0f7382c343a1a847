"""Helpers shared by the -m gpu parity tests (they call the product through its public Python API,
which goes through the C-ABI of libmsx.so)."""
import numpy as np
import torch

import medical_segmentation_pytorch_b200 as msx
from medical_segmentation_pytorch_b200 import ops
from medical_segmentation_pytorch_b200.ops import Act

DEV = 'cuda'


def rel(a, b):
    a = a.detach().double().cpu() if torch.is_tensor(a) else torch.as_tensor(np.asarray(a)).double()
    b = b.detach().double().cpu() if torch.is_tensor(b) else torch.as_tensor(np.asarray(b)).double()
    return float((a - b).norm() / max(float(b.norm()), 1e-300))


def to_act(x_nchw, dtype=torch.float32, requires_grad=False):
    a = ops.to_nhwc(x_nchw.to(DEV).float(), dtype)
    if requires_grad:
        a.t.requires_grad_(True)
    return a


def from_act(t, c):
    return Act(t, c).nchw().cpu()


class Cfg:
    def __init__(self, **kw):
        self.model, self.num_class, self.num_channel, self.base_channel = 'unet', 2, 3, 8
        self.use_aux, self.class_weights, self.loss_type, self.ignore_index = False, None, 'ce', 255
        self.reduction, self.ohem_thrs, self.kd_loss_type, self.kd_temperature = 'mean', 0.7, 'kl_div', 4.0
        self.DDP, self.synBN = False, False
        self.__dict__.update(kw)
