"""Autograd bindings: each Function is a thin shim over one C-ABI composite of libmsx.so.

Activations travel between modules as :class:`Act` - an NHWC tensor ``[N, H, W, round_up(C, 8)]``
(fp32 or bf16; padded channels hold zeros) plus the logical channel count.  A tensor may alias a
channel slice of a wider buffer (pixel stride ``ld`` > its own width), which is how UNet's
``torch.cat`` (models/unet.py:66) becomes zero-copy.

No op here has a PyTorch-eager or CPU implementation: everything calls the CUDA library.
"""
import ctypes as C

import torch

from . import _lib as L
from . import packplan
from ._lib import lib


def rup(x, m=8):
    return (x + m - 1) // m * m


class Act:
    """NHWC activation: tensor ``t`` [N,H,W,cp] (stride(3) == 1) and logical channels ``c``."""
    __slots__ = ('t', 'c')

    def __init__(self, t, c):
        self.t, self.c = t, c

    @property
    def shape(self):
        return self.t.shape

    def nchw(self):
        """Debug/test helper: fp32 NCHW copy of the logical channels."""
        n, h, w, _ = self.t.shape
        out = torch.empty(n, self.c, h, w, dtype=torch.float32, device=self.t.device)
        L.check(lib.msx_nhwc_to_nchw(L.dtype_code(self.t.dtype), L.ptr(self.t), ld_of(self.t), n, self.c, h, w,
                                     L.ptr(out), L.stream()))
        return out


def ld_of(t):
    """Pixel stride of an NHWC tensor; validates that it is a dense-pixel layout."""
    n, h, w, c = t.shape
    s = t.stride()
    ld = s[2]
    if s[3] != 1 or ld % 8 or ld < c or s[1] != w * ld or s[0] != h * w * ld or (t.data_ptr() % 16):
        raise L.MsxError(f'tensor with shape {tuple(t.shape)} stride {s} is not an NHWC activation layout')
    return ld


def as_act_layout(t):
    """Return ``t`` if it is a valid NHWC layout, else a contiguous copy (rare: autograd-made grads)."""
    n, h, w, c = t.shape
    s = t.stride()
    ok = s[3] == 1 and s[2] % 8 == 0 and s[2] >= c and s[1] == w * s[2] and s[0] == h * w * s[2] and t.data_ptr() % 16 == 0
    return t if ok else t.contiguous()


def alias_slice(buf, off, c):
    """Tensor aliasing channels [off, off+c) of NHWC ``buf`` WITHOUT being an autograd view of it."""
    n, h, w, ld = buf.shape
    return torch.empty(0, dtype=buf.dtype, device=buf.device).set_(
        buf.untyped_storage(), buf.storage_offset() + off, (n, h, w, c), (h * w * ld, w * ld, ld, 1))


def alias(t):
    """Tensor aliasing all of ``t`` (same size/stride) WITHOUT being an autograd view of it."""
    return torch.empty(0, dtype=t.dtype, device=t.device).set_(t.untyped_storage(), t.storage_offset(), t.shape, t.stride())


class OutSlot:
    """Destination for a producer's output: channels [off, off+cp) of a preallocated NHWC buffer."""
    __slots__ = ('buf', 'off')

    def __init__(self, buf, off):
        self.buf, self.off = buf, off


def _out_tensor(slot, n, h, w, cp, dtype, device):
    if slot is None:
        return torch.empty(n, h, w, cp, dtype=dtype, device=device)
    assert slot.buf.shape[:3] == (n, h, w) and slot.buf.dtype == dtype
    return alias_slice(slot.buf, slot.off, cp)


class ConvGeom:
    """Static geometry of a conv module (mirrors nn.Conv2d / nn.ConvTranspose2d arguments)."""
    __slots__ = ('cin', 'cout', 'kh', 'kw', 'stride', 'dil', 'pad_h', 'pad_w', 'transposed', 'out_pad', 'engine')

    def __init__(self, cin, cout, kh, kw, stride=1, dil=1, pad_h=0, pad_w=0, transposed=False, out_pad=0, engine=L.ENGINE_AUTO):
        self.cin, self.cout, self.kh, self.kw = cin, cout, kh, kw
        self.stride, self.dil, self.pad_h, self.pad_w = stride, dil, pad_h, pad_w
        self.transposed, self.out_pad, self.engine = transposed, out_pad, engine

    def out_hw(self, h, w):
        if not self.transposed:
            return ((h + 2 * self.pad_h - self.dil * (self.kh - 1) - 1) // self.stride + 1,
                    (w + 2 * self.pad_w - self.dil * (self.kw - 1) - 1) // self.stride + 1)
        return ((h - 1) * self.stride - 2 * self.pad_h + self.dil * (self.kh - 1) + self.out_pad + 1,
                (w - 1) * self.stride - 2 * self.pad_w + self.dil * (self.kw - 1) + self.out_pad + 1)

    def desc(self, x, ldy, oh, ow):
        n, h, w, _ = x.shape
        return L.ConvDesc(n, h, w, self.cin, oh, ow, self.cout, self.kh, self.kw, self.stride, self.dil,
                          self.pad_h, self.pad_w, ld_of(x), ldy, int(self.transposed), L.dtype_code(x.dtype), self.engine)


_DEFAULT_ENGINE = [L.ENGINE_AUTO]


def set_conv_engine(engine):
    """Force the conv engine for subsequently built descriptors (tests: SIMT as comparator)."""
    _DEFAULT_ENGINE[0] = {'auto': L.ENGINE_AUTO, 'simt': L.ENGINE_SIMT, 'tcgen05': L.ENGINE_TCGEN05}[engine]


def _desc(g, x, ldy, oh, ow):
    d = g.desc(x, ldy, oh, ow)
    if d.engine == L.ENGINE_AUTO:
        d.engine = _DEFAULT_ENGINE[0]
    return d


def _f32(t):
    if t is not None and (t.dtype != torch.float32 or not t.is_contiguous()):
        raise L.MsxError('parameters / statistics must be contiguous fp32')
    return t


def _pgrad(p):
    """Destination of a parameter gradient: its slot in the flat gradient arena when the model has
    been flattened (flat.py), else a fresh tensor handed back to autograd."""
    g = getattr(p, '_msx_grad', None)
    return g if g is not None else torch.empty_like(p)


def _pdone(p, g):
    """After the kernel writing ``g`` was enqueued: notify DDP bucket bookkeeping; arena-resident
    gradients are not returned to autograd (it would copy them)."""
    if g is not None and getattr(p, '_msx_grad', None) is g:
        ready = getattr(p, '_msx_ready', None)
        if ready is not None:
            ready()
        return None
    return g


def _grad_like_input(x):
    """dx buffer with the SAME pixel stride as x (the C-ABI uses desc.ldx for both)."""
    n, h, w, cp = x.shape
    ld = ld_of(x)
    if ld == cp:
        return torch.empty(n, h, w, cp, dtype=x.dtype, device=x.device)
    return alias_slice(torch.empty(n, h, w, ld, dtype=x.dtype, device=x.device), 0, cp)


class AddFn(torch.autograd.Function):
    """Plain n-way add (DuckNet `x5 + x` before mid_stage, models/ducknet.py:43) -> msx_sum_n."""

    @staticmethod
    def forward(ctx, c, *xs):
        ctx.n = len(xs)
        return sum_n(list(xs), c)

    @staticmethod
    def backward(ctx, g):
        return (None,) + (g,) * ctx.n


# ---------------------------------------------------------------------------------------------
class ConvBNActFn(torch.autograd.Function):
    """ConvBNAct / DeConvBNAct (models/modules.py:73-108) -> msx_conv_bn_act_{fwd,bwd}."""

    @staticmethod
    def forward(ctx, x, weight, bias, gamma, beta, rm, rv, nbt, g, eps, momentum, training, relu, slot, infer=False):
        n, h, w, _ = x.shape
        oh, ow = g.out_hw(h, w)
        cp = rup(g.cout)
        dev = x.device
        # inference (eval mode, nothing to differentiate): no raw conv output; BN affine + ReLU run in the conv epilogue
        # infer: the caller saw that no backward can follow (ctx.needs_input_grad ignores torch.no_grad())
        fused_eval = _EVAL_FUSION[0] and not training and infer
        y = None if fused_eval else torch.empty(n, oh, ow, cp, dtype=x.dtype, device=dev)
        z = _out_tensor(slot, n, oh, ow, cp, x.dtype, dev)
        mean = torch.empty(cp, dtype=torch.float32, device=dev)
        invstd = torch.empty(cp, dtype=torch.float32, device=dev)
        d = _desc(g, x, cp, oh, ow)
        wsb = lib.msx_ws_bytes_conv(C.byref(d))
        ws = L.workspace(wsb, dev)
        pre = packplan.lookup(_f32(weight))          # pre-packed by the model's plan: pass w_master = NULL
        pk = pre[0] if pre else L.scratch('pack', lib.msx_conv_packed_bytes(C.byref(d)), dev)
        ctx.pk_b = pre[1] if pre else None
        L.check(lib.msx_conv_bn_act_fwd(
            C.byref(d), L.ptr(x), None if pre else L.ptr(weight), L.ptr(_f32(bias)), L.ptr(_f32(gamma)), L.ptr(_f32(beta)),
            L.ptr(rm), L.ptr(rv), L.ptr(nbt), eps, momentum, int(training), int(relu),
            L.ptr(pk), L.ptr(y), L.ptr(z), ld_of(z), L.ptr(mean), L.ptr(invstd), L.ptr(ws), ws.numel(), L.stream()))
        ctx.g, ctx.d = g, d
        ctx.relu = int(relu) | (0 if training else L.BN_EVAL)      # MSX_BN_* bit set for the backward
        ctx.bias = bias                      # leaf parameter (or None); only its identity is needed
        ctx.save_for_backward(x, weight, gamma, beta, y, mean, invstd)
        return z

    @staticmethod
    def backward(ctx, dz):
        x, weight, gamma, beta, y, mean, invstd = ctx.saved_tensors
        dz = as_act_layout(dz)
        d, dev = ctx.d, x.device
        dy = torch.empty_like(y)
        dx = _grad_like_input(x) if ctx.needs_input_grad[0] else None
        bias = ctx.bias
        dw = _pgrad(weight)
        db = _pgrad(bias) if bias is not None else None
        dgamma, dbeta = _pgrad(gamma), _pgrad(beta)
        pk = ctx.pk_b if ctx.pk_b is not None else L.scratch('pack', lib.msx_conv_packed_bytes(C.byref(d)), dev)
        side = _wgrad_stream() if _WGRAD_STREAM[0] else None
        if side is None:
            ws = L.workspace(lib.msx_ws_bytes_conv(C.byref(d)), dev)
            L.check(lib.msx_conv_bn_act_bwd(
                C.byref(d), L.ptr(x), None if ctx.pk_b is not None else L.ptr(weight), L.ptr(y), L.ptr(dz), ld_of(dz), L.ptr(gamma), L.ptr(beta),
                L.ptr(mean), L.ptr(invstd), int(ctx.relu), L.ptr(pk), L.ptr(dy), L.ptr(dx), L.ptr(dw), L.ptr(db),
                L.ptr(dgamma), L.ptr(dbeta), L.ptr(ws), ws.numel(), L.stream()))
        else:
            # the same three steps as the composite, with the weight gradient (tensor-core bound, nothing downstream
            # in the backward chain needs it) on a side stream where it overlaps the data gradient and the next
            # layers' bandwidth-bound BN kernels
            n_, oh_, ow_, cpy = y.shape
            ws = L.workspace(lib.msx_ws_bytes_bn(ctx.g.cout), dev)
            L.check(lib.msx_bn_act_bwd(L.dtype_code(y.dtype), L.ptr(y), cpy, L.ptr(dz), ld_of(dz), n_ * oh_ * ow_, ctx.g.cout,
                                       L.ptr(gamma), L.ptr(beta), L.ptr(mean), L.ptr(invstd), int(ctx.relu), L.ptr(dy), cpy,
                                       L.ptr(dgamma), L.ptr(dbeta), L.ptr(ws), ws.numel(), L.stream()))
            main = torch.cuda.current_stream()
            side.wait_event(main.record_event())
            x.record_stream(side)
            dy.record_stream(side)
            with torch.cuda.stream(side):
                ws2 = L.workspace(lib.msx_ws_bytes_conv(C.byref(d)), dev)          # the side stream's own lane
                L.check(lib.msx_conv2d_bwd(C.byref(d), L.ptr(x), None, L.ptr(dy), L.ptr(pk), None, L.ptr(dw), None,
                                           L.ptr(ws2), ws2.numel(), L.stream()))
            if dx is not None or db is not None:
                ws = L.workspace(lib.msx_ws_bytes_conv(C.byref(d)), dev)
                L.check(lib.msx_conv2d_bwd(C.byref(d), L.ptr(x), None if ctx.pk_b is not None else L.ptr(weight), L.ptr(dy), L.ptr(pk),
                                           L.ptr(dx), None, L.ptr(db), L.ptr(ws), ws.numel(), L.stream()))
            if getattr(weight, '_msx_grad', None) is dw:
                _defer_join(side)                     # arena gradient: joined once, at the end of the backward pass
            else:
                main.wait_stream(side)                # handed back to autograd, which assumes this node's stream
        # notify in arena order (reverse registration): bn.bias, bn.weight, conv.bias, conv.weight
        dbeta, dgamma = _pdone(beta, dbeta), _pdone(gamma, dgamma)
        db = _pdone(bias, db) if bias is not None else None
        if side is not None:
            with torch.cuda.stream(side):             # DDP records the stream the gradient is produced on
                dw = _pdone(weight, dw)
        else:
            dw = _pdone(weight, dw)
        return (dx, dw, db, dgamma, dbeta) + (None,) * 10


class Conv2dFn(torch.autograd.Function):
    """Bias-free conv without BN (ResidualBlock.upper_branch, models/ducknet.py:93) -> msx_conv2d_{fwd,bwd}."""

    @staticmethod
    def forward(ctx, x, weight, g, *rest):             # rest: optionally the bias (ConvTranspose2d before a SyncBatchNorm)
        bias = rest[0] if rest else None
        ctx.nrest = len(rest)
        n, h, w, _ = x.shape
        oh, ow = g.out_hw(h, w)
        cp = rup(g.cout)
        y = torch.empty(n, oh, ow, cp, dtype=x.dtype, device=x.device)
        d = _desc(g, x, cp, oh, ow)
        pre = packplan.lookup(_f32(weight))
        pk = pre[0] if pre else L.scratch('pack', lib.msx_conv_packed_bytes(C.byref(d)), x.device)
        L.check(lib.msx_conv2d_fwd(C.byref(d), L.ptr(x), None if pre else L.ptr(weight), L.ptr(_f32(bias)), L.ptr(pk), L.ptr(y), L.stream()))
        ctx.d, ctx.pk_b = d, (pre[1] if pre else None)
        ctx.bias = bias
        ctx.save_for_backward(x, weight)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        d = ctx.d
        dy = as_act_layout(dy)
        if ld_of(dy) != d.ldy:
            d = L.ConvDesc.from_buffer_copy(d)
            d.ldy = ld_of(dy)
        dx = _grad_like_input(x) if ctx.needs_input_grad[0] else None
        dw = _pgrad(weight)
        bias = ctx.bias
        db = _pgrad(bias) if bias is not None else None
        ws = L.workspace(lib.msx_ws_bytes_conv(C.byref(d)), x.device)
        pk = ctx.pk_b if ctx.pk_b is not None else L.scratch('pack', lib.msx_conv_packed_bytes(C.byref(d)), x.device)
        L.check(lib.msx_conv2d_bwd(C.byref(d), L.ptr(x), None if ctx.pk_b is not None else L.ptr(weight), L.ptr(dy), L.ptr(pk), L.ptr(dx), L.ptr(dw), L.ptr(db),
                                   L.ptr(ws), ws.numel(), L.stream()))
        db = _pdone(bias, db) if bias is not None else None          # arena order: conv.bias before conv.weight
        return (dx, _pdone(weight, dw), None) + ((db,) if ctx.nrest else ())


# ---- sibling branches on side streams ---------------------------------------------------------
_BRANCH_STREAMS = [int(__import__('os').environ.get('MSX_BRANCH_STREAMS', '1'))]
_BRANCH_MAX_ELEMS = [int(__import__('os').environ.get('MSX_BRANCH_MAX_ELEMS', '0'))]      # 0: no limit
_STREAM_POOL = {}


def set_branch_streams(on):
    """Run independent branches of a block (DUCK's six, models/ducknet.py:66-74) on side streams (default on)."""
    _BRANCH_STREAMS[0] = int(bool(on))


def run_branches(jobs, main_job=0):
    """jobs: (callable returning an Act, [input Acts produced on the current stream]).  Job ``main_job``
    stays on the current stream, the others each get a pooled side stream forked from it and joined back
    before returning, so that small layers of sibling branches overlap instead of queueing behind each other.
    torch.autograd replays the same streams in backward.  Results are identical to the serial order: every
    kernel is deterministic and the branches share no buffers (scratch is per stream)."""
    big = _BRANCH_MAX_ELEMS[0] and any(a.t.numel() > _BRANCH_MAX_ELEMS[0] for _, ins in jobs for a in ins)
    if not _BRANCH_STREAMS[0] or big or not 2 <= len(jobs) <= 7 or not torch.cuda.is_available():
        return [j() for j, _ in jobs]
    main = torch.cuda.current_stream()
    key = (main.device.index, main.cuda_stream)
    pool = _STREAM_POOL.setdefault(key, [])                  # side streams of THIS stream: forks may nest
    while len(pool) < len(jobs) - 1:
        pool.append(torch.cuda.Stream(device=main.device))
        # scratch lane = position in the fork tree (lane 0 = a stream the library did not create; < 8 siblings per fork),
        # so that an eager warm-up and a graph capture on another stream land on the same buffers
        L.LANE_OF_STREAM[pool[-1].cuda_stream] = L.LANE_OF_STREAM.get(main.cuda_stream, 0) * 8 + len(pool)
    fork = main.record_event()
    outs, k = [None] * len(jobs), 0
    for i, (job, inputs) in enumerate(jobs):
        if i == main_job:
            continue
        s = pool[k]
        k += 1
        s.wait_event(fork)
        for a in inputs:
            a.t.record_stream(s)                     # allocated on main, read (also by the backward) on the side stream
        with torch.cuda.stream(s):
            outs[i] = job()
        outs[i].t.record_stream(main)                # consumed on the main stream after the join
    outs[main_job] = jobs[main_job][0]()
    for s in pool[:k]:
        main.wait_stream(s)
    return outs


_EVAL_FUSION = [int(__import__('os').environ.get('MSX_EVAL_FUSION', '1'))]


def set_eval_fusion(on):
    """Inference: apply eval-mode BN + ReLU in the conv epilogue (default on) instead of as separate passes."""
    _EVAL_FUSION[0] = int(bool(on))


# ---- weight gradients on a side stream ---------------------------------------------------------
_WGRAD_STREAM = [int(__import__('os').environ.get('MSX_WGRAD_STREAM', '1'))]
_PENDING_JOIN = {}


def set_wgrad_stream(on):
    """ConvBNAct backward: run the weight-gradient kernels on a side stream of the current stream (default on)."""
    _WGRAD_STREAM[0] = int(bool(on))


def _wgrad_stream():
    main = torch.cuda.current_stream()
    key = ('wgrad', main.device.index, main.cuda_stream)
    s = _STREAM_POOL.get(key)
    if s is None:
        s = _STREAM_POOL[key] = torch.cuda.Stream(device=main.device)
        L.LANE_OF_STREAM[s.cuda_stream] = L.LANE_OF_STREAM.get(main.cuda_stream, 0) * 8 + 7     # sibling slot 7: run_branches uses 1..6
    return s


def _join_wgrad_streams():
    cur = torch.cuda.current_stream()
    for s in _PENDING_JOIN.values():
        cur.wait_stream(s)
    _PENDING_JOIN.clear()


def _defer_join(side):
    """Make the stream that called backward() wait for ``side`` when the backward pass ends (engine callback);
    outside a backward pass (a Function's backward called by hand) join at once."""
    if not _PENDING_JOIN:
        try:
            torch.autograd.Variable._execution_engine.queue_callback(_join_wgrad_streams)
        except RuntimeError:
            torch.cuda.current_stream().wait_stream(side)
            return
    _PENDING_JOIN[side.cuda_stream] = side


def _ptr_arrays(ts):
    n = len(ts)
    ptrs = (C.c_void_p * n)(*[t.data_ptr() for t in ts])
    lds = (C.c_int32 * n)(*[ld_of(t) for t in ts])
    return ptrs, lds


class SumBNActFn(torch.autograd.Function):
    """(sum of inputs, first optionally nearest-2x upsampled) -> BN -> ReLU -> msx_sum_bn_act_fwd / msx_bn_act_bwd."""

    @staticmethod
    def forward(ctx, gamma, beta, rm, rv, nbt, c, eps, momentum, training, relu, up2x, slot, infer, *xs):
        ref = xs[1] if up2x else xs[0]
        n, h, w, cp = ref.shape
        dev, dt = ref.device, ref.dtype
        need_sum = len(xs) > 1 or up2x
        fused_eval = _EVAL_FUSION[0] and infer and not training and need_sum     # sum + BN affine + ReLU in one kernel
        s = torch.empty(n, h, w, cp, dtype=dt, device=dev) if need_sum and not fused_eval else None
        z = _out_tensor(slot, n, h, w, cp, dt, dev)
        mean = torch.empty(cp, dtype=torch.float32, device=dev)
        invstd = torch.empty(cp, dtype=torch.float32, device=dev)
        ptrs, lds = _ptr_arrays(xs)
        ws = L.workspace(lib.msx_ws_bytes_bn(c), dev)
        L.check(lib.msx_sum_bn_act_fwd(
            L.dtype_code(dt), len(xs), ptrs, lds, int(up2x), n, h, w, c, L.ptr(s), cp if s is not None else 0,
            L.ptr(_f32(gamma)), L.ptr(_f32(beta)), L.ptr(rm), L.ptr(rv), L.ptr(nbt), eps, momentum, int(training), int(relu),
            L.ptr(z), ld_of(z), L.ptr(mean), L.ptr(invstd), L.ptr(ws), ws.numel(), L.stream()))
        ctx.c, ctx.up2x, ctx.nx = c, up2x, len(xs)
        ctx.relu = int(relu) | (0 if training else L.BN_EVAL)
        ctx.x0_shape = xs[0].shape
        ctx.save_for_backward(s if need_sum else xs[0], gamma, beta, mean, invstd)
        return z

    @staticmethod
    def backward(ctx, dz):
        s, gamma, beta, mean, invstd = ctx.saved_tensors
        dz = as_act_layout(dz)
        n, h, w, cp = s.shape
        dev = s.device
        ds = torch.empty(n, h, w, cp, dtype=s.dtype, device=dev)
        dgamma, dbeta = _pgrad(gamma), _pgrad(beta)
        ws = L.workspace(lib.msx_ws_bytes_bn(ctx.c), dev)
        code = L.dtype_code(s.dtype)
        L.check(lib.msx_bn_act_bwd(code, L.ptr(s), ld_of(s), L.ptr(dz), ld_of(dz), n * h * w, ctx.c, L.ptr(gamma), L.ptr(beta),
                                   L.ptr(mean), L.ptr(invstd), int(ctx.relu), L.ptr(ds), cp, L.ptr(dgamma), L.ptr(dbeta),
                                   L.ptr(ws), ws.numel(), L.stream()))
        grads = [ds] * ctx.nx
        if ctx.up2x:
            da = torch.empty(ctx.x0_shape, dtype=s.dtype, device=dev)
            L.check(lib.msx_upsample2x_bwd(code, L.ptr(ds), cp, n, h, w, ctx.c, L.ptr(da), ld_of(da), L.stream()))
            grads[0] = da
        for i in range(ctx.nx):
            if not ctx.needs_input_grad[13 + i]:
                grads[i] = None
        dbeta, dgamma = _pdone(beta, dbeta), _pdone(gamma, dgamma)
        return (dgamma, dbeta) + (None,) * 11 + tuple(grads)


# ---- SyncBatchNorm (utils/parallel.py:37-38) -----------------------------------------------------
def _all_reduce_sum(buf, group):
    """Sum ``buf`` (a small double tensor on the device) over the ranks of ``group``.  NCCL reduces it in place on the
    current stream (capturable); a gloo group - the CPU-side tests - gets a host round trip."""
    import torch.distributed as dist
    if dist.get_backend(group) == 'nccl':
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    else:
        host = buf.cpu()
        dist.all_reduce(host, op=dist.ReduceOp.SUM, group=group)
        buf.copy_(host)


class SyncBNActFn(torch.autograd.Function):
    """BatchNorm over the samples of ALL ranks (+ReLU) of an already formed tensor ``s`` (a raw conv output or a sum):
    msx_bn_stats_sums -> all-reduce of [sum, sum of squares, count] -> msx_bn_finalize_sums -> msx_bn_apply, and in backward
    msx_bn_bwd_reduce_sums -> all-reduce of [sum g, sum g*xhat] -> msx_bn_bwd_apply_sums; the arithmetic of
    torch.nn.SyncBatchNorm (weight / bias gradients stay per-rank and are averaged by DDP like every other parameter)."""

    @staticmethod
    def forward(ctx, s, gamma, beta, rm, rv, nbt, c, eps, momentum, relu, slot, group):
        n, h, w, cp = s.shape
        dev = s.device
        code = L.dtype_code(s.dtype)
        z = _out_tensor(slot, n, h, w, cp, s.dtype, dev)
        mean = torch.empty(cp, dtype=torch.float32, device=dev)
        invstd = torch.empty(cp, dtype=torch.float32, device=dev)
        sums = torch.empty(2 * cp + 1, dtype=torch.float64, device=dev)
        ws = L.workspace(lib.msx_ws_bytes_bn(c), dev)
        L.check(lib.msx_bn_stats_sums(code, L.ptr(s), ld_of(s), n * h * w, c, L.ptr(sums), L.ptr(ws), ws.numel(), L.stream()))
        _all_reduce_sum(sums, group)
        L.check(lib.msx_bn_finalize_sums(L.ptr(sums), c, eps, momentum, L.ptr(mean), L.ptr(invstd), L.ptr(rm), L.ptr(rv), L.ptr(nbt), L.stream()))
        L.check(lib.msx_bn_apply(code, L.ptr(s), ld_of(s), L.ptr(z), ld_of(z), n * h * w, c, L.ptr(mean), L.ptr(invstd),
                                 L.ptr(_f32(gamma)), L.ptr(_f32(beta)), int(relu), L.stream()))
        ctx.c, ctx.relu, ctx.group = c, int(relu), group
        ctx.save_for_backward(s, gamma, beta, mean, invstd, sums)
        return z

    @staticmethod
    def backward(ctx, dz):
        s, gamma, beta, mean, invstd, fsums = ctx.saved_tensors
        dz = as_act_layout(dz)
        n, h, w, cp = s.shape
        dev = s.device
        code = L.dtype_code(s.dtype)
        ds = torch.empty(n, h, w, cp, dtype=s.dtype, device=dev)
        dgamma, dbeta = _pgrad(gamma), _pgrad(beta)
        bsums = torch.empty(2 * cp, dtype=torch.float64, device=dev)
        ws = L.workspace(lib.msx_ws_bytes_bn(ctx.c), dev)
        L.check(lib.msx_bn_bwd_reduce_sums(code, L.ptr(s), ld_of(s), L.ptr(dz), ld_of(dz), n * h * w, ctx.c, L.ptr(gamma), L.ptr(beta),
                                           L.ptr(mean), L.ptr(invstd), ctx.relu, L.ptr(dgamma), L.ptr(dbeta), L.ptr(bsums),
                                           L.ptr(ws), ws.numel(), L.stream()))
        _all_reduce_sum(bsums, ctx.group)
        L.check(lib.msx_bn_bwd_apply_sums(code, L.ptr(s), ld_of(s), L.ptr(dz), ld_of(dz), n * h * w, ctx.c, L.ptr(gamma), L.ptr(beta),
                                          L.ptr(mean), L.ptr(invstd), ctx.relu, L.ptr(bsums), L.ptr(fsums[2 * cp:]), L.ptr(ds), cp,
                                          L.ptr(ws), ws.numel(), L.stream()))
        dbeta, dgamma = _pdone(beta, dbeta), _pdone(gamma, dgamma)
        return (ds, dgamma, dbeta) + (None,) * 9


class SumFn(torch.autograd.Function):
    """sum of inputs (first optionally nearest-2x upsampled) as its own node: the SyncBatchNorm path of sum -> BN."""

    @staticmethod
    def forward(ctx, c, up2x, *xs):
        ref = xs[1] if up2x else xs[0]
        n, h, w, cp = ref.shape
        out = torch.empty(n, h, w, cp, dtype=ref.dtype, device=ref.device)
        ptrs, lds = _ptr_arrays(xs)
        L.check(lib.msx_sum_act(L.dtype_code(ref.dtype), len(xs), ptrs, lds, int(up2x), n, h, w, c, 0, L.ptr(out), cp, L.stream()))
        ctx.c, ctx.up2x, ctx.nx, ctx.x0_shape = c, up2x, len(xs), xs[0].shape
        return out

    @staticmethod
    def backward(ctx, g):
        g = as_act_layout(g)
        grads = [g] * ctx.nx
        if ctx.up2x:
            n, h, w, cp = g.shape
            da = torch.empty(ctx.x0_shape, dtype=g.dtype, device=g.device)
            L.check(lib.msx_upsample2x_bwd(L.dtype_code(g.dtype), L.ptr(g), ld_of(g), n, h, w, ctx.c, L.ptr(da), ld_of(da), L.stream()))
            grads[0] = da
        return (None, None) + tuple(grads)


def sum_n(ts, c):
    """out = sum(ts) with the library's n-way sum kernel (<= 8 inputs per launch)."""
    ts = [as_act_layout(t) for t in ts]
    out = torch.empty(ts[0].shape, dtype=ts[0].dtype, device=ts[0].device)
    n, h, w, cp = out.shape
    code = L.dtype_code(out.dtype)
    first = True
    while ts:
        chunk = ts[:8] if first else [out] + ts[:7]
        ts = ts[8:] if first else ts[7:]
        first = False
        ptrs, lds = _ptr_arrays(chunk)
        L.check(lib.msx_sum_n(code, len(chunk), ptrs, lds, n * h * w, c, L.ptr(out), cp, L.stream()))
    return out


class FanoutFn(torch.autograd.Function):
    """Identity with k outputs; backward adds the k gradients in ONE pass (msx_sum_n) instead of the
    k-1 pairwise adds autograd would issue (DUCK.in_bn output feeds nine convs, models/ducknet.py:144-149)."""

    @staticmethod
    def forward(ctx, x, k, c):
        ctx.c = c
        return tuple(alias(x) for _ in range(k))

    @staticmethod
    def backward(ctx, *gs):
        gs = [g for g in gs if g is not None]
        if len(gs) == 1:
            return gs[0], None, None
        return sum_n(gs, ctx.c), None, None


class MaxPoolFn(torch.autograd.Function):
    """nn.MaxPool2d(3, 2, 1) (models/unet.py:49) -> msx_maxpool3s2_{fwd,bwd}."""

    @staticmethod
    def forward(ctx, x, c):
        n, h, w, cp = x.shape
        oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
        z = torch.empty(n, oh, ow, cp, dtype=x.dtype, device=x.device)
        idx = torch.empty(n, oh, ow, cp, dtype=torch.int8, device=x.device)
        L.check(lib.msx_maxpool3s2_fwd(L.dtype_code(x.dtype), L.ptr(x), ld_of(x), n, h, w, c, L.ptr(z), cp, L.ptr(idx), cp, L.stream()))
        ctx.c, ctx.xshape = c, x.shape
        ctx.save_for_backward(idx)
        return z

    @staticmethod
    def backward(ctx, dz):
        idx, = ctx.saved_tensors
        dz = as_act_layout(dz)
        n, h, w, cp = ctx.xshape
        dx = torch.empty(ctx.xshape, dtype=dz.dtype, device=dz.device)
        L.check(lib.msx_maxpool3s2_bwd(L.dtype_code(dz.dtype), L.ptr(dz), ld_of(dz), L.ptr(idx), cp, n, h, w, ctx.c, L.ptr(dx), cp, L.stream()))
        return dx, None


class ConcatFn(torch.autograd.Function):
    """torch.cat([a, b], dim=1) (models/unet.py:66) with zero copies: both producers already wrote
    their channels into ``buf``; backward hands out channel slices of the incoming gradient."""

    @staticmethod
    def forward(ctx, a, b, buf, ca):
        assert a.data_ptr() == buf.data_ptr() and b.data_ptr() == buf.data_ptr() + ca * buf.element_size()
        ctx.ca = ca
        return alias_slice(buf, 0, buf.shape[3])

    @staticmethod
    def backward(ctx, g):
        g = as_act_layout(g)
        ca = ctx.ca
        return alias_slice(g, 0, ca), alias_slice(g, ca, g.shape[3] - ca), None, None


class HeadFn(torch.autograd.Function):
    """seg_head conv1x1 (models/unet.py:27,41) -> msx_head_{fwd,bwd}; returns NCHW fp32 logits."""

    @staticmethod
    def forward(ctx, x, weight, c):
        n, h, w, _ = x.shape
        nc = weight.shape[0]
        logits = torch.empty(n, nc, h, w, dtype=torch.float32, device=x.device)
        L.check(lib.msx_head_fwd(L.dtype_code(x.dtype), L.ptr(x), ld_of(x), n, h, w, c, L.ptr(_f32(weight)), nc, L.ptr(logits), L.stream()))
        ctx.c = c
        ctx.save_for_backward(x, weight)
        return logits

    @staticmethod
    def backward(ctx, dl):
        x, weight = ctx.saved_tensors
        n, h, w, cp = x.shape
        dl = dl.contiguous().float()
        dx = torch.empty(n, h, w, cp, dtype=x.dtype, device=x.device) if ctx.needs_input_grad[0] else None
        dw = _pgrad(weight)
        ws = L.workspace(lib.msx_ws_bytes_bn(ctx.c), x.device)
        L.check(lib.msx_head_bwd(L.dtype_code(x.dtype), L.ptr(x), ld_of(x), n, h, w, ctx.c, L.ptr(weight), weight.shape[0],
                                 L.ptr(dl), L.ptr(dx), cp, L.ptr(dw), L.ptr(ws), ws.numel(), L.stream()))
        return dx, _pdone(weight, dw), None


def to_nhwc(images, dtype):
    """NCHW fp32 images (core/seg_trainer.py:35) -> NHWC activation (no gradient to the images)."""
    if images.dtype != torch.float32:
        images = images.float()
    images = images.contiguous()
    n, c, h, w = images.shape
    out = torch.empty(n, h, w, rup(c), dtype=dtype, device=images.device)
    L.check(lib.msx_nchw_to_nhwc(L.dtype_code(dtype), L.ptr(images), n, c, h, w, L.ptr(out), rup(c), L.stream()))
    return Act(out, c)
