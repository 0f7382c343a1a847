"""ctypes binding of libmsx.so (the C-ABI declared in include/msx.h).

The library is the product: if it is missing or fails to load this module raises - there is no
eager-PyTorch or CPU fallback anywhere in the package.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('MSX_LIB') or os.path.join(_HERE, 'libmsx.so')     # MSX_LIB: a measurement build (build.py --variant)

F32, BF16 = 0, 1
PACK_FPROP, PACK_DGRAD = 0, 1
ENGINE_AUTO, ENGINE_SIMT, ENGINE_TCGEN05 = 0, 1, 2
REDUCE = {'mean': 0, 'sum': 1, 'none': 2}
BN_RELU, BN_EVAL = 1, 2


class ConvDesc(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        'n', 'h', 'w', 'cin', 'oh', 'ow', 'cout', 'kh', 'kw', 'stride', 'dil', 'pad_h', 'pad_w',
        'ldx', 'ldy', 'transposed', 'dtype', 'engine')]


class PackItem(C.Structure):
    _fields_ = [('w_master', C.c_void_p), ('packed', C.c_void_p)] + [(n, C.c_int32) for n in (
        'cout', 'cin', 'kh', 'kw', 'transposed', 'mode')]


class MsxError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f'{LIB_PATH} not found: build it with `python medical_segmentation_pytorch_b200/build.py` '
            '(nvcc, sm_100a). There is no fallback path.')
    lib = C.CDLL(LIB_PATH)
    p, i32, i64, f32, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_size_t
    dp = C.POINTER(ConvDesc)
    sig = {
        'msx_last_error': (C.c_char_p, []),
        'msx_version': (C.c_int, []),
        'msx_last_conv_engine': (C.c_int, []),
        'msx_launch_count': (C.c_longlong, []),
        'msx_profile_enable': (C.c_int, [C.c_int]),
        'msx_profile_num_classes': (C.c_int, []),
        'msx_profile_class_name': (C.c_char_p, [C.c_int]),
        'msx_profile_collect': (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_longlong), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        'msx_conv_packed_bytes': (sz, [dp]),
        'msx_ws_bytes_conv': (sz, [dp]),
        'msx_ws_bytes_bn': (sz, [i32]),
        'msx_ws_bytes_loss': (sz, []),
        'msx_ws_bytes_ce': (sz, [i64]),
        'msx_conv_pack_weights': (C.c_int, [dp, C.c_int, p, p, p]),
        'msx_conv_pack_plan_bytes': (sz, [p, i32]),
        'msx_conv_pack_plan_build': (C.c_int, [p, i32, p, sz, p]),
        'msx_conv_pack_plan_run': (C.c_int, [C.c_int, p, i32, i32, p]),
        'msx_conv2d_fwd': (C.c_int, [dp, p, p, p, p, p, p]),
        'msx_conv2d_bwd': (C.c_int, [dp, p, p, p, p, p, p, p, p, sz, p]),
        'msx_conv_bn_act_fwd': (C.c_int, [dp, p, p, p, p, p, p, p, p, f32, f32, C.c_int, C.c_int,
                                          p, p, p, i32, p, p, p, sz, p]),
        'msx_conv_bn_act_bwd': (C.c_int, [dp, p, p, p, p, i32, p, p, p, p, C.c_int,
                                          p, p, p, p, p, p, p, p, sz, p]),
        'msx_sum_bn_act_fwd': (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(i32), C.c_int,
                                         i32, i32, i32, i32, p, i32, p, p, p, p, p, f32, f32, C.c_int, C.c_int,
                                         p, i32, p, p, p, sz, p]),
        'msx_bn_act_bwd': (C.c_int, [C.c_int, p, i32, p, i32, i64, i32, p, p, p, p, C.c_int, p, i32, p, p, p, sz, p]),
        'msx_bn_act_bwd_static': (C.c_int, [C.c_int, p, i32, p, i32, i64, i32, p, p, p, p, C.c_int, p, i32, p, p, p]),
        'msx_bn_stats_sums': (C.c_int, [C.c_int, p, i32, i64, i32, p, p, sz, p]),
        'msx_bn_finalize_sums': (C.c_int, [p, i32, f32, f32, p, p, p, p, p, p]),
        'msx_bn_apply': (C.c_int, [C.c_int, p, i32, p, i32, i64, i32, p, p, p, p, C.c_int, p]),
        'msx_bn_bwd_reduce_sums': (C.c_int, [C.c_int, p, i32, p, i32, i64, i32, p, p, p, p, C.c_int, p, p, p, p, sz, p]),
        'msx_bn_bwd_apply_sums': (C.c_int, [C.c_int, p, i32, p, i32, i64, i32, p, p, p, p, C.c_int, p, p, p, i32, p, sz, p]),
        'msx_sum_n': (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(i32), i64, i32, p, i32, p]),
        'msx_sum_act': (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(i32), C.c_int, i32, i32, i32, i32, C.c_int, p, i32, p]),
        'msx_upsample2x_bwd': (C.c_int, [C.c_int, p, i32, i32, i32, i32, i32, p, i32, p]),
        'msx_maxpool3s2_fwd': (C.c_int, [C.c_int, p, i32, i32, i32, i32, i32, p, i32, p, i32, p]),
        'msx_maxpool3s2_bwd': (C.c_int, [C.c_int, p, i32, p, i32, i32, i32, i32, i32, p, i32, p]),
        'msx_nchw_to_nhwc': (C.c_int, [C.c_int, p, i32, i32, i32, i32, p, i32, p]),
        'msx_nhwc_to_nchw': (C.c_int, [C.c_int, p, i32, i32, i32, i32, i32, p, p]),
        'msx_head_fwd': (C.c_int, [C.c_int, p, i32, i32, i32, i32, i32, p, i32, p, p]),
        'msx_head_bwd': (C.c_int, [C.c_int, p, i32, i32, i32, i32, i32, p, i32, p, p, i32, p, p, sz, p]),
        'msx_ce_loss_fwd_bwd': (C.c_int, [p, p, i32, i32, i64, p, i64, C.c_int, p, p, p, sz, p]),
        'msx_ce_loss_none_fwd': (C.c_int, [p, p, i32, i32, i64, p, i64, p, p]),
        'msx_ce_loss_none_bwd': (C.c_int, [p, p, i32, i32, i64, p, i64, p, p, p]),
        'msx_ohem_loss_fwd_bwd': (C.c_int, [p, p, i32, i32, i64, f32, i64, p, p, p, p, sz, p]),
        'msx_kd_loss_fwd_bwd': (C.c_int, [p, p, i32, i32, i64, C.c_int, f32, p, p, p, sz, p]),
        'msx_dice_bce_loss_fwd_bwd': (C.c_int, [p, p, C.c_int, i64, f32, p, p, p, sz, p]),
        'msx_scale_by_device_scalar': (C.c_int, [p, i64, p, p]),
        'msx_confmat_update': (C.c_int, [p, p, i32, i32, i64, i64, p, p, p]),
        'msx_head_confmat_update': (C.c_int, [C.c_int, p, i32, i64, i32, p, i32, p, i64, p, p, p]),
        'msx_adam_step': (C.c_int, [p, p, p, p, i64, f32, f32, f32, f32, f32, C.c_int, f32, p, p, p]),
        'msx_sgd_step': (C.c_int, [p, p, p, i64, f32, f32, f32, f32, C.c_int, p, p]),
        'msx_ema_update': (C.c_int, [p, p, i64, C.c_double, p]),
        'msx_ema_update_i64': (C.c_int, [p, p, i64, C.c_double, p]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)          # AttributeError here == header/library mismatch
        fn.restype, fn.argtypes = res, args
    return lib, sig


lib, SIGNATURES = _load()


def check(rc):
    if rc != 0:
        raise MsxError(f'libmsx error {rc}: {lib.msx_last_error().decode()}')


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def dtype_code(dt):
    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    raise MsxError(f'unsupported activation dtype {dt}')


# ---- per-device scratch ----------------------------------------------------------------------
_scratch = {}
LANE_OF_STREAM = {}


def scratch(kind, nbytes, device):
    """Grow-only byte buffers shared by all calls of one lane of a device.  Lane 0 is whatever stream the caller
    runs on (calls there are ordered); DUCK-Net's sibling branches run on side streams registered in LANE_OF_STREAM
    (ops.run_branches), each with its own buffers.  Keyed by lane, not stream, so that an eager warm-up and a later
    graph capture (which runs on another stream) find the same buffers."""
    lane = LANE_OF_STREAM.get(torch.cuda.current_stream().cuda_stream, 0) if LANE_OF_STREAM else 0
    key = (kind, device.index if device.index is not None else torch.cuda.current_device(), lane)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        if torch.cuda.is_current_stream_capturing():
            raise MsxError(f'scratch "{kind}" must grow to {nbytes} B during graph capture: run one eager step first')
        size = max(int(nbytes), 1 << 20)
        size = (size * 5 // 4 + 511) // 512 * 512
        buf = torch.empty(size, dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


def workspace(nbytes, device):
    return scratch('ws', nbytes, device)


def profile_collect():
    """{class_name: dict(ms, launches, bytes, flops)} of everything recorded since the last call."""
    k = lib.msx_profile_num_classes()
    ms, ln, by, fl = (C.c_double * k)(), (C.c_longlong * k)(), (C.c_double * k)(), (C.c_double * k)()
    check(lib.msx_profile_collect(ms, ln, by, fl))
    out = {}
    for i in range(k):
        if ln[i]:
            out[lib.msx_profile_class_name(i).decode()] = {'ms': ms[i], 'launches': int(ln[i]), 'bytes': by[i], 'flops': fl[i]}
    return out
