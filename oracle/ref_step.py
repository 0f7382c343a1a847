"""Oracle: the reference's training / eval step on CPU (TEST INFRASTRUCTURE + timed CPU baseline).

Restates the step body of core/seg_trainer.py:38-87 (zero_grad -> forward -> CrossEntropyLoss ->
backward -> Adam step -> EMA state copy) and the eval body :114-125 on top of oracle.ref_models /
ref_loss / ref_metrics, with the same ATen CPU kernels the reference's nn.Modules would call
(F.conv2d, F.batch_norm(training=True), F.max_pool2d, ...), so its wall time is a fair stand-in for
the reference's own CPU training step (kind = "port" in bench.py's cpu_baseline).
"""
import time

import torch

from . import ref_loss, ref_metrics, ref_models


class CpuTrainer:
    """``device='cuda'`` runs the very same torch ops through PyTorch eager on the GPU (cuDNN / cuBLAS library kernels):
    bench.py's same-box library comparator, never part of the product path."""

    def __init__(self, name, base_channel, num_class=2, seed=1, lr=1e-3, threads=None, device='cpu'):
        if threads:
            torch.set_num_threads(threads)
        layout = ref_models.model_layout(name, num_class=num_class, n_channel=3, base_channel=base_channel)
        sd = ref_models.synth_state_dict(layout, seed)
        if str(device) != 'cpu':
            sd = {k: v.to(device) for k, v in sd.items()}
        self.params = ref_models.make_trainable(sd)
        self.net = ref_models.FnNet(name, self.params, training=True, fused_bn=True)
        self.keys = ref_models.param_keys(self.params)
        self.opt = torch.optim.Adam([self.params[k] for k in self.keys], lr=lr)          # utils/optimizer.py:14-16
        self.ema = {k: v.detach().clone() for k, v in self.params.items()}               # utils/model_ema.py:21
        self.num_class = num_class

    def train_step(self, images, masks, autocast_dtype=None, sync=True):
        self.net.training = True
        self.opt.zero_grad()                                                             # core/seg_trainer.py:38
        with torch.autocast(images.device.type, dtype=autocast_dtype or torch.bfloat16, enabled=autocast_dtype is not None):   # :61
            logits = self.net(images)
            loss = torch.nn.functional.cross_entropy(logits, masks, ignore_index=255)    # core/loss.py:29-31
        loss.backward()                                                                  # :82
        self.opt.step()                                                                  # :83
        with torch.no_grad():                                                            # :87, use_ema=False -> plain copy
            for k, v in self.params.items():
                self.ema[k].copy_(v)
        return float(loss.detach()) if sync else loss.detach()

    @torch.no_grad()
    def eval_step(self, images, masks):
        net = ref_models.FnNet(self.net.name, self.ema, training=False, fused_bn=True)    # :114 validates the EMA copy
        logits = net(images).numpy()
        cm = ref_metrics.confusion_matrix(logits, masks.numpy(), self.num_class, 255)
        return ref_metrics.iou_per_class(cm), ref_metrics.dice_macro(ref_metrics.confusion_matrix(
            logits, masks.clamp(max=self.num_class - 1).numpy(), self.num_class, None))

    def time_steps(self, images, masks, steps, warmup):
        for _ in range(warmup):
            self.train_step(images, masks)
        t0 = time.perf_counter()
        for _ in range(steps):
            self.train_step(images, masks)
        return (time.perf_counter() - t0) / max(steps, 1)
