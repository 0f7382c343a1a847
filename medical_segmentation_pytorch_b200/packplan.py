"""One-launch weight packing for a whole model (msx_conv_pack_plan_*, include/msx.h).

The tcgen05 / SIMT conv kernels read a packed copy of the weight ([tap][round16(N)][round8(K)] in the compute
dtype), one for the forward operand and one for the data-gradient operand.  Packing inside every conv call cost
~540 tiny launches per DUCK-Net step; a model forward instead refreshes every conv's two packed copies with one
launch (``PackPlan.refresh``) and the layer calls pass ``w_master = NULL``.

The plan is only consulted while the owning model's ``forward`` runs (``active()``), so a layer called on its
own - the unit tests do that - still packs per call.  The reference has no such step (its convs read the master
weights, models/modules.py:73-108); weights must not change between a forward and its backward, which is the
rule torch.autograd enforces for the reference too (saved-tensor version check).
"""
import contextlib
import ctypes as C

import torch
import torch.nn as nn

from . import _lib as L
from ._lib import lib

_ACTIVE = [None]


def _rup(x, m):
    return (x + m - 1) // m * m


def lookup(weight):
    """(fprop_packed, dgrad_packed) of the conv owning ``weight`` from the plan of the model whose forward is running, or None."""
    plan = _ACTIVE[0]
    return None if plan is None else plan.packed.get(weight.data_ptr())


class PackPlan:
    def __init__(self, model, dtype, device, skip=()):
        self.dtype, self.device = dtype, device
        es = 2 if dtype == torch.bfloat16 else 4
        convs = [m for m in model.modules() if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)) and all(m is not s for s in skip)]
        sizes = []
        for m in convs:
            t = isinstance(m, nn.ConvTranspose2d)
            cin, cout, taps = m.in_channels, m.out_channels, m.kernel_size[0] * m.kernel_size[1]
            sizes.append((_rup(taps * _rup(cout, 16) * _rup(cin, 8) * es, 256), _rup(taps * _rup(cin, 16) * _rup(cout, 8) * es, 256), t))
        total = sum(a + b for a, b, _ in sizes)
        self.arena = torch.empty(max(total, 256), dtype=torch.uint8, device=device)
        self.packed, self.ptrs = {}, []
        items_f, items_b = [], []
        off = 0
        for m, (nf, nb, t) in zip(convs, sizes):
            pf, pb = self.arena[off:off + nf], self.arena[off + nf:off + nf + nb]
            off += nf + nb
            self.packed[m.weight.data_ptr()] = (pf, pb)
            self.ptrs.append((m, m.weight.data_ptr()))
            common = dict(w_master=m.weight.data_ptr(), cout=m.out_channels, cin=m.in_channels, kh=m.kernel_size[0],
                          kw=m.kernel_size[1], transposed=int(t))
            items_f.append(L.PackItem(packed=pf.data_ptr(), mode=L.PACK_FPROP, **common))
            items_b.append(L.PackItem(packed=pb.data_ptr(), mode=L.PACK_DGRAD, **common))
        # two tables: forward operands only (eval / no_grad), and both passes (training)
        self.tables = {False: self._build(items_f), True: self._build(items_f + items_b)}

    def _build(self, items):
        n = len(items)
        if n == 0:
            return None
        arr = (L.PackItem * n)(*items)
        nbytes = lib.msx_conv_pack_plan_bytes(arr, n)
        table = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        chunks = C.c_int32(0)
        L.check(lib.msx_conv_pack_plan_build(arr, n, L.ptr(table), nbytes, C.byref(chunks)))
        return table, n, chunks.value

    def __deepcopy__(self, memo):
        return None                      # a copied model (EMA) builds its own plan on first use

    def valid(self):
        return all(m.weight.data_ptr() == p and m.weight.dtype == torch.float32 for m, p in self.ptrs)

    def refresh(self, with_dgrad):
        t = self.tables[bool(with_dgrad)]
        if t is not None:
            table, n, chunks = t
            L.check(lib.msx_conv_pack_plan_run(L.dtype_code(self.dtype), L.ptr(table), n, chunks, L.stream()))


@contextlib.contextmanager
def active(model, dtype, device, skip=()):
    """Refresh (building on first use / after parameters moved) and expose the model's plan for the duration of its forward."""
    plans = model.__dict__.setdefault('_msx_packplans', {})
    key = (dtype, device)
    plan = plans.get(key)
    if plan is None or not plan.valid():
        if torch.cuda.is_current_stream_capturing():
            raise L.MsxError('pack plan must be built before stream capture: run one forward eagerly first')
        plan = plans[key] = PackPlan(model, dtype, device, skip)
    plan.refresh(torch.is_grad_enabled())
    prev, _ACTIVE[0] = _ACTIVE[0], plan
    try:
        yield plan
    finally:
        _ACTIVE[0] = prev
