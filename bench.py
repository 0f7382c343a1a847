#!/usr/bin/env python
"""Headline benchmark: DUCKNet-17 training step at 352x352 (BASELINE.json metric "train imgs/sec").

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W     # the reference's CPU step (oracle port)

One "step" = zero_grad -> forward -> CrossEntropyLoss -> backward (+ bucketed gradient all-reduce
when N > 1) -> Adam -> EMA state copy, the body of core/seg_trainer.py:38-87, on one synthetic batch
per GPU.  Prints ONE JSON line on rank 0 (see the contract in the task statement):
  value      img/s with inputs already resident in HBM (CUDA events, max over ranks)
  e2e        same step driven from pinned HOST buffers (H2D of images+masks and D2H of the loss inside
             the timed region)
  roofline   the dominant kernel class of a profiled pass (live CUDA-event timings per launch,
             algorithmic bytes/flops booked by the library) against MEASURED_PEAKS.json
  cpu_baseline  the oracle port of the reference step timed on this box's host cores (rank 0, N=1)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MODELS = {  # name -> (arch, base_channel, train GFLOP/img @352 (2*MAC*3), conv activation GB/img bf16)  SURVEY.md §8d
    'ducknet17': ('ducknet', 17, 279.7, 2.197),
    'ducknet34': ('ducknet', 34, 1114.3, 4.370),
    'unet32': ('unet', 32, 144.6, 0.467),
}


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return {'hbm_gbs': d['hbm_gbs'], 'tf': d.get('bf16_tflops_sustained', d['bf16_tflops']), 'tf_burst': d['bf16_tflops'], 'src': 'measured'}
    return {'hbm_gbs': 6650.0, 'tf': 1400.0, 'tf_burst': 1590.0, 'src': 'fallback'}


class ClockSampler:
    """nvidia-smi clock / throttle sampling DURING the timed region (B200_PROFILING.md recipe)."""
    Q = 'index,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
        'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--query-gpu={self.Q}', '--format=csv,noheader,nounits', '-lms', '200',
                                          '-i', str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if not self.proc:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
                for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), r[3:7]):
                    if v.lower().startswith('active'):
                        reasons.add(name)
            except Exception:
                pass
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': mx, 'reasons': sorted(reasons), 'samples': len(sm)}


def _synth_batch_fn():
    """synth.py loaded by file path: the reference arm must not import the package, which would map libmsx.so."""
    import importlib.util
    spec = importlib.util.spec_from_file_location('msx_synth', os.path.join(ROOT, 'medical_segmentation_pytorch_b200', 'synth.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.synth_batch


def cpu_reference_leg(name, steps, warmup, batch=2, hw=352):
    """The reference's CPU training step (oracle port) on this box's host cores."""
    import torch
    from oracle.ref_step import CpuTrainer
    synth_batch = _synth_batch_fn()
    arch, base, _, _ = MODELS[name]
    cores = os.cpu_count() or 1
    threads = max(1, cores // 2) if cores > 16 else cores          # physical cores on SMT hosts
    tr = CpuTrainer(arch, base, 2, seed=1, lr=1e-3, threads=threads)
    x, t = synth_batch(batch, hw, hw, seed=1)
    sec = tr.time_steps(x, t, steps, warmup)
    return {'value': batch / sec, 'unit': 'img/s', 'cores': threads, 'kind': 'port',
            'sample': f'{steps} timed step(s) after {warmup} warm-up of {name} batch {batch} at {hw}x{hw} fp32 '
                      f'(oracle/ref_step.py: forward+CE+backward+Adam+EMA copy), {sec:.2f} s/step'}, sec


def gpu_eager_leg(name, batch, hw, dev, steps=5, warmup=10):
    """Same-box library comparator (tools/test_speed.py:29-58 discipline): the reference's step body on the
    reference's own torch ops (oracle restatement on F.conv2d / F.batch_norm / ...), NCHW, cudnn.benchmark off, under
    PyTorch eager on this GPU - (i) fp32 with TF32 convolutions (torch's default, which the reference never changes),
    (ii) torch.autocast(bfloat16).  Runs after the timed region; cuDNN / cuBLAS kernels, none of this repo's."""
    import torch
    from oracle.ref_step import CpuTrainer
    synth_batch = _synth_batch_fn()
    arch, base, _, _ = MODELS[name]
    torch.backends.cudnn.benchmark = False
    out = {'how': f'oracle/ref_step.py on cuda (PyTorch eager, cuDNN/cuBLAS), batch {batch} at {hw}x{hw}, {warmup} warm-up + {steps} timed steps, CUDA events',
           'cudnn_allow_tf32': bool(torch.backends.cudnn.allow_tf32)}
    x, t = synth_batch(batch, hw, hw, seed=1)
    x, t = x.to(dev), t.to(dev)
    for tag, dt in (('tf32', None), ('bf16', torch.bfloat16)):
        try:
            tr = CpuTrainer(arch, base, 2, seed=1, lr=1e-3, device=dev)
            for _ in range(warmup):
                tr.train_step(x, t, autocast_dtype=dt, sync=False)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                loss = tr.train_step(x, t, autocast_dtype=dt, sync=False)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            out[tag] = {'value': batch / (ms / 1e3), 'unit': 'img/s', 'ms_per_step': ms, 'loss': float(loss)}
            del tr
            torch.cuda.empty_cache()
        except Exception as e:                                   # noqa: BLE001 - a comparator must not sink the bench line
            out[tag] = {'error': f'{type(e).__name__}: {e}'[:200]}
    return out


def run_reference(args):
    rank = int(os.environ.get('RANK', 0))
    if rank != 0:
        return
    cb, sec = cpu_reference_leg(args.model, args.steps, args.warmup, batch=2, hw=args.size)
    arch, base, _, _ = MODELS[args.model]
    line = {'impl': 'reference', 'metric': 'train_imgs_per_sec', 'value': cb['value'], 'unit': 'img/s', 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': sec * 1e3, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': f'{args.model} training step {args.size}x{args.size}, CPU host cores, bounded sample: batch 2 per step',
                       'timing': 'host wall clock'},
            'cpu_baseline': cb, 'e2e': {'value': cb['value'], 'unit': 'img/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), flush=True)


def loss_metric_leg(msx, _lib, dev, pk, n=256, h=352, reps=5):
    """north_star's second figure: algorithmic GB/s of the loss / metric kernels (SURVEY 8d bytes per pixel) on a
    working set far beyond L2 (n x h x h pixels), all launches of one call summed, L2 evicted by reads between calls."""
    import torch
    class Cfg:
        num_class, ignore_index, class_weights, loss_type, reduction, ohem_thrs = 2, 255, None, 'ce', 'mean', 0.7
        kd_loss_type, kd_temperature = 'kl_div', 4.0
    px = n * h * h
    z = torch.randn(n, 2, h, h, device=dev, requires_grad=True)          # requires_grad: the fused kernels also emit the gradient
    z1 = torch.randn(n, 1, h, h, device=dev, requires_grad=True)
    zt = torch.randn(n, 2, h, h, device=dev)
    t = (torch.rand(n, h, h, device=dev) > 0.8).long()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ce, dice, metric = msx.get_loss_fn(Cfg, dev), msx.DiceBCELoss(), msx.get_seg_metrics(Cfg, 'iou').to(dev)
    jobs = {'ce_fwd_bwd': (lambda: ce(z, t), 24), 'kd_kl_fwd_bwd': (lambda: msx.kd_loss_fn(Cfg, z, zt), 24),
            'dice_bce_fwd_bwd': (lambda: dice(z1, t), 28), 'confmat_update': (lambda: metric.update(z.detach(), t), 16)}
    out = {'pixels': px, 'peak_GBps': pk['hbm_gbs'], 'how': f'{reps} calls each, per-launch CUDA events, every launch of the call counted'}
    for name, (fn, bpp) in jobs.items():
        fn()
        torch.cuda.synchronize()
        _lib.profile_collect()
        _lib.lib.msx_profile_enable(1)
        for _ in range(reps):
            flush.sum()
            fn()
        prof = _lib.profile_collect()
        _lib.lib.msx_profile_enable(0)
        ms = sum(v['ms'] for v in prof.values()) / reps
        out[name] = {'bytes_per_px': bpp, 'us': ms * 1e3, 'GBps': px * bpp / ms / 1e6, 'frac': px * bpp / ms / 1e6 / pk['hbm_gbs']}
    return out


class Workload:
    """One rank's training step of a BASELINE config: model + loss + fused optimizer + EMA on static input buffers."""

    def __init__(self, msx, name, batch, size, bf16, dev, world, rank, kd=None, sync_bn=False):
        import torch
        from medical_segmentation_pytorch_b200.flat import FlatBuffers, FlatParams
        from medical_segmentation_pytorch_b200.synth import synth_batch
        self.torch, self.msx, self.name, self.B, self.S, self.bf16, self.world = torch, msx, name, batch, size, bf16, world
        arch, base, self.gflop_img, self.gb_img = MODELS[name]

        class Cfg:
            num_class, ignore_index, class_weights, loss_type, reduction = 2, 255, None, 'ce', 'mean'
            kd_loss_type, kd_temperature = 'kl_div', 4.0               # configs/base_config.py:102-104
        self.cfg = Cfg
        torch.manual_seed(1)                                           # configs/base_config.py:75 random_seed = 1
        self.model = (msx.DuckNet if arch == 'ducknet' else msx.UNet)(num_class=2, n_channel=3, base_channel=base).to(dev)
        self.model.train()
        FlatParams(self.model)
        FlatBuffers(self.model)
        if sync_bn and world > 1:                                      # utils/parallel.py:37-38 (the reference's default synBN=True)
            msx.convert_sync_batchnorm(self.model)
        self.net = msx.DistributedDataParallelB200(self.model) if world > 1 else self.model
        self.opt = msx.FusedOptimizer(self.net, 'adam', lr=0.1 * 0.01 * world)   # utils/optimizer.py:15 with my_config.py:23
        self.ema = msx.ModelEma(self.net)
        self.loss_fn = msx.get_loss_fn(Cfg, dev)
        self.teacher = msx.SmpUnet(kd, None, 3, 2).to(dev).eval() if kd else None     # random-init (no checkpoints offline)
        xh, th = synth_batch(batch, size, size, seed=1 + rank)
        self.xh, self.th = xh.pin_memory(), th.pin_memory()
        self.sx, self.st = self.xh.to(dev), self.th.to(dev)             # static input buffers (graph replays read these)
        self.it = 0

    def step(self):
        torch, msx = self.torch, self.msx
        self.opt.zero_grad()
        with torch.autocast('cuda', dtype=torch.bfloat16, enabled=self.bf16):
            preds = self.net(self.sx)
            loss = self.loss_fn(preds, self.st)
            if self.teacher is not None:                                 # core/seg_trainer.py:69-75
                with torch.no_grad():
                    tp = self.teacher(self.sx)
                loss = loss + 1.0 * msx.kd_loss_fn(self.cfg, preds, tp.detach())
        loss.backward()
        self.opt.step()
        self.it += 1
        self.ema.update(self.model, self.it)
        return loss

    def fractions(self, value_per_gpu, pk):
        return {'tensor_frac': value_per_gpu * self.gflop_img / 1e3 / pk['tf'],
                'hbm_frac': value_per_gpu * self.gb_img * (1 if self.bf16 else 2) / pk['hbm_gbs']}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--model', default='ducknet17', choices=sorted(MODELS))
    ap.add_argument('--batch', type=int, default=16, help='images per GPU')
    ap.add_argument('--size', type=int, default=352)
    ap.add_argument('--dtype', default='bf16', choices=['bf16', 'fp32'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-extras', action='store_true', help='skip the legs that only add context to the line: GPU eager (cuDNN) comparator, '
                    'eager (no-graph) value, other BASELINE configs, loss/metric GB/s')
    ap.add_argument('--profile-steps', type=int, default=2)
    ap.add_argument('--ncu', action='store_true', help='short eager run for an ncu launch list: 1 warm-up + 1 step, no graph / e2e / profile pass')
    ap.add_argument('--kd', default=None, metavar='ENCODER', help="BASELINE config 5: add a frozen smp.Unet(<ENCODER>, e.g. resnet101) teacher forward "
                    "and the KL distillation term (T = 4, coefficient 1) to every step")
    ap.add_argument('--sync-bn', action='store_true', help='SyncBatchNorm (the reference default synBN=True) instead of per-GPU statistics (north_star); N > 1 only')
    ap.add_argument('--no-branch-streams', action='store_true', help='run sibling DUCK branches one after another on one stream')
    ap.add_argument('--no-graph', action='store_true', help='launch every kernel from Python instead of replaying one CUDA graph per step')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'b200' else args.warmup
    if args.ncu:
        args.warmup, args.steps, args.no_graph, args.profile_steps, args.no_cpu_baseline, args.no_extras = 1, 1, True, 0, True, True
    if args.impl == 'reference':
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import medical_segmentation_pytorch_b200 as msx
    from medical_segmentation_pytorch_b200 import _lib, ops as msx_ops
    msx_ops.set_branch_streams(not args.no_branch_streams)
    msx_ops.set_wgrad_stream(not args.no_branch_streams)

    rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    B, S = args.batch, args.size
    bf16 = args.dtype == 'bf16'
    extras = not args.no_extras and world == 1

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    def capture(wl, warmup):
        """Eager warm-up, then the whole step as ONE CUDA graph through the package's public helper (msx.capture_step:
        what a user of the reference's trainer calls to get this number).  Returns (replay | None, error, launches/step)."""
        for _ in range(warmup):
            loss = wl.step().detach()
        torch.cuda.synchronize()
        assert torch.isfinite(loss).item(), 'non-finite loss in warm-up'
        replay, err, per_step = None, None, None
        if not args.no_graph:
            try:
                l0 = _lib.lib.msx_launch_count()
                replay = msx.capture_step(wl.step, warmup=1, optimizers=[wl.opt])
                per_step = (_lib.lib.msx_launch_count() - l0) // 2           # one eager warm-up + the captured pass
                replay()
                torch.cuda.synchronize()
                assert torch.isfinite(replay.out).item(), 'non-finite loss from the graph replay'
            except Exception as e:                                   # noqa: BLE001 - report and fall back to eager launches
                replay, err = None, f'{type(e).__name__}: {e}'[:300]
                if os.environ.get('MSX_BENCH_DEBUG'):
                    import traceback
                    traceback.print_exc()
                torch.cuda.synchronize()
        if world > 1:                                              # every rank must take the same path
            flag = torch.tensor([1 if replay is not None else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag) == 0:
                replay = None
        return replay, err, per_step

    wl = Workload(msx, args.model, B, S, bf16, dev, world, rank, args.kd, args.sync_bn)
    replay, graph_err, launches_per_step = capture(wl, args.warmup)

    def run_step():
        return replay() if replay is not None else wl.step()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = _lib.lib.msx_launch_count()
    ms = timed(run_step, args.steps)
    launches = (_lib.lib.msx_launch_count() - l0) if replay is None else launches_per_step * args.steps
    clocks = sampler.stop() if rank == 0 else None

    # end to end: pinned host buffers -> device every step, loss read back every step
    def e2e_step():
        wl.sx.copy_(wl.xh, non_blocking=True)
        wl.st.copy_(wl.th, non_blocking=True)
        return float(run_step().detach())
    if args.ncu:
        ms_e2e = ms
    else:
        e2e_step()
        ms_e2e = timed(e2e_step, args.steps)

    # the same step WITHOUT the graph: what the reference's own Python loop gets when it drives these modules directly
    eager_value = None
    if extras and replay is not None:
        k = min(args.steps, 5)
        wl.step().detach()
        eager_value = B * world * k / (timed(lambda: wl.step().detach(), k) / 1e3)

    # profiled pass: per-launch CUDA events on the launching stream, algorithmic bytes/flops per launch
    prof = {}
    if args.profile_steps > 0:                                   # every rank steps (collectives), rank 0 records
        msx_ops.set_branch_streams(0)                            # kernels timed one at a time, not overlapped with siblings
        msx_ops.set_wgrad_stream(0)
        if rank == 0:
            _lib.lib.msx_profile_enable(1)
        for _ in range(args.profile_steps):
            wl.step().detach()
        if rank == 0:
            prof = _lib.profile_collect()
            _lib.lib.msx_profile_enable(0)
        msx_ops.set_branch_streams(not args.no_branch_streams)
        msx_ops.set_wgrad_stream(not args.no_branch_streams)
    barrier()

    def leave():
        # ranks leave through os._exit after a final barrier: tearing the NCCL communicator down while CUDA
        # graphs that captured its collectives are still alive can block in ncclCommDestroy
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
            sys.stdout.flush()
            os._exit(0)

    if rank != 0:
        leave()
        return
    pk = peaks()
    imgs = B * world * args.steps
    value = imgs / (ms / 1e3)
    tot_ms = sum(v['ms'] for v in prof.values()) or 1.0
    kernels = {k: {'ms_per_step': v['ms'] / args.profile_steps, 'share': v['ms'] / tot_ms, 'launches_per_step': v['launches'] // args.profile_steps,
                   'GBps': v['bytes'] / v['ms'] / 1e6 if v['ms'] else 0.0, 'TFLOPs': v['flops'] / v['ms'] / 1e9 if v['ms'] else 0.0}
               for k, v in sorted(prof.items(), key=lambda kv: -kv[1]['ms'])}
    roofline = None
    if prof:
        top, tv = max(prof.items(), key=lambda kv: kv[1]['ms'])
        t_tensor, t_hbm = tv['flops'] / (pk['tf'] * 1e12), tv['bytes'] / (pk['hbm_gbs'] * 1e9)
        if t_tensor >= t_hbm:
            ach = tv['flops'] / tv['ms'] / 1e9
            roofline = {'kernel': top, 'bound': 'tensor', 'achieved': ach, 'peak': pk['tf'], 'unit': 'TFLOP/s', 'frac': ach / pk['tf']}
        else:
            ach = tv['bytes'] / tv['ms'] / 1e6
            roofline = {'kernel': top, 'bound': 'hbm', 'achieved': ach, 'peak': pk['hbm_gbs'], 'unit': 'GB/s', 'frac': ach / pk['hbm_gbs']}
        traffic, traffic_note, traffic_source = None, None, None
        try:
            tr = json.load(open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json'))).get(top)
            if tr and args.model == 'ducknet17' and not args.kd:      # the capture is of the headline workload's top kernel
                traffic = tr['traffic_bytes']
                traffic_note = f"{tr['kernel']} {tr['shape']}: algorithmic (unpadded channels) {tr['algorithmic_bytes']} B per launch"
                traffic_source = f"committed ncu --set full capture ({tr['source']}), NOT measured in this run"
        except Exception:
            pass
        roofline.update({'traffic': traffic, 'traffic_note': traffic_note, 'traffic_source': traffic_source,
                         'peak_source': pk['src'] + (' (sustained bf16)' if roofline['bound'] == 'tensor' else ''),
                         'avg_launch_us': tv['ms'] / tv['launches'] * 1e3, 'launches_per_step': tv['launches'] // args.profile_steps,
                         'share_of_step': tv['ms'] / tot_ms,
                         'how': f'{args.profile_steps} extra step(s) after the timed region with per-launch CUDA events (msx_profile_*)'})
    loss_metric = loss_metric_leg(msx, _lib, dev, pk) if extras else None
    workload_name = wl.name
    model_roofline = {'train_gflop_per_img': wl.gflop_img, 'conv_act_gb_per_img_bf16': wl.gb_img, **wl.fractions(value / world, pk)}

    # the other single-GPU BASELINE configs, briefly (bounded: 3 warm-up + 5 graph replays each), for the driver's record
    other = None
    if extras and args.model == 'ducknet17' and not args.kd and bf16 and replay is not None:
        del replay, wl
        torch.cuda.empty_cache()
        other = {}
        for name, ob, okd in (('unet32', 16, None), ('ducknet34', 32, None), ('unet32', 16, 'resnet101')):
            key = name + (f'+kd_{okd}' if okd else '')
            try:
                w2 = Workload(msx, name, ob, S, True, dev, 1, 0, okd)
                r2, err2, _ = capture(w2, 3)
                if r2 is None:
                    other[key] = {'error': err2 or 'graph capture disabled'}
                else:
                    ms2 = timed(r2, 5)
                    v2 = ob * 5 / (ms2 / 1e3)
                    other[key] = {'value': v2, 'unit': 'img/s', 'ms_per_step': ms2 / 5, 'batch': ob, 'size': S, 'dtype': 'bf16', **w2.fractions(v2, pk)}
                    if okd:
                        other[key].pop('tensor_frac'), other[key].pop('hbm_frac')      # student-only flop/byte model does not cover the teacher
                del r2, w2
            except Exception as e:                               # noqa: BLE001
                other[key] = {'error': f'{type(e).__name__}: {e}'[:200]}
            torch.cuda.empty_cache()
    gpu_eager = gpu_eager_leg(workload_name, B, S, dev) if extras and not args.kd else None
    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        cpu_baseline, _ = cpu_reference_leg(workload_name, 2, 1, batch=2, hw=S)
    graph_on = launches_per_step is not None and graph_err is None and not args.no_graph
    line = {
        'metric': 'train_imgs_per_sec', 'value': value, 'unit': 'img/s', 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'bf16' if bf16 else 'f32', 'data': 'synthetic',
        'config': {'workload': f'{workload_name} training step (fwd+CE+' + (f'frozen smp.Unet({args.kd}) teacher fwd+KD(kl,T=4)+' if args.kd else '') + f'bwd+Adam+EMA copy), {S}x{S}, batch {B}/GPU, num_class 2, ' + ('SyncBatchNorm' if args.sync_bn and world > 1 else 'per-GPU BN'),
                   'global_batch': B * world, 'parallelism': f'dp{world}', 'l2': 'per-step working set >> 126 MB L2 (no flush needed)',
                   'seed': 'weights manual_seed(1); data PCG64(1+rank)',
                   'launch': 'one CUDA graph replay per step (msx.capture_step)' if graph_on else 'eager Python launches' + (f' (graph capture failed: {graph_err})' if graph_err else '')},
        'clocks': clocks,
        'e2e': {'value': imgs / (ms_e2e / 1e3), 'unit': 'img/s', 'h2d_bytes_per_step': B * 3 * S * S * 4 + B * S * S * 8, 'd2h_bytes_per_step': 4},
        'gpu_launches': int(launches),
        'eager_value': eager_value,
        'roofline': roofline,
        'cpu_baseline': cpu_baseline,
        'gpu_eager_baseline': gpu_eager,
        'model_roofline': model_roofline,
        'other_configs': other,
        'kernels': kernels,
        'loss_metric_gbs': loss_metric,
    }
    print(json.dumps(line), flush=True)
    leave()


if __name__ == '__main__':
    main()
