"""ctypes binding of libnr_sm100.so (C ABI: include/nr_sm100.h).

The library is the product: there is NO CPU or PyTorch fallback.  Import of this
module fails loudly when the shared library has not been built, and every op
raises when handed a non-CUDA tensor.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libnr_sm100.so')

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f'{LIB_PATH} not found: build it with `python -c "import __graft_entry__ as g; g.build()"` '
        f'or `make -C {os.path.join(_HERE, "csrc")}` (nvcc, sm_100a). There is no fallback path.')

lib = C.CDLL(LIB_PATH)

ROWS_DENSE, ROWS_TOKEN, ROWS_NEWS = 0, 1, 2
PREC_FP32, PREC_BF16 = 0, 1
STREAM_IN, STREAM_OUT = 1, 2
SCORE_CE_NOALLOC_MAX_B = 16384      # include/nr_sm100.h: NR_SCORE_CE_NOALLOC_MAX_B


class NrRows(C.Structure):
    _fields_ = [('mode', C.c_int32), ('ids_stride', C.c_int32), ('dense', C.c_void_p),
                ('table', C.c_void_p), ('ids', C.c_void_p), ('table_rows', C.c_int64),
                ('drop_p', C.c_float), ('seed', C.c_uint64)]


class NrTitleEncoder(C.Structure):
    _fields_ = [('kind', C.c_int32), ('S', C.c_int32), ('Dm', C.c_int32), ('H', C.c_int32), ('Q', C.c_int32),
                ('heads', C.c_int32), ('w', C.c_void_p * 3), ('b', C.c_void_p * 3), ('p_w1', C.c_void_p),
                ('p_b1', C.c_void_p), ('p_w2', C.c_void_p), ('p_b2', C.c_void_p), ('mask', C.c_void_p),
                ('out_drop_p', C.c_float)]


class NrTitleEncoderGrads(C.Structure):
    _fields_ = [('dw', C.c_void_p * 3), ('db', C.c_void_p * 3), ('dp_w1', C.c_void_p), ('dp_b1', C.c_void_p),
                ('dp_w2', C.c_void_p), ('dp_b2', C.c_void_p), ('dtable', C.c_void_p), ('dx_dense', C.c_void_p), ('dtable_ready', C.c_void_p)]


ENC_MHSA, ENC_CONV, ENC_POOL = 0, 1, 2

_P, _I32, _I64, _F, _U64, _SZ = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_uint64, C.c_size_t
_R = C.POINTER(NrRows)
_E = C.POINTER(NrTitleEncoder)
_G = C.POINTER(NrTitleEncoderGrads)

_SIGS = {
    'nr_abi_version': (C.c_int, []),
    'nr_last_error': (C.c_char_p, []),
    'nr_device_info': (C.c_int, [_P, _P, _P]),
    'nr_launch_count': (C.c_longlong, []),
    'nr_prof_enable': (C.c_int, [C.c_int]),
    'nr_prof_report': (_SZ, [C.c_char_p, _SZ]),
    'nr_set_seed_salt': (C.c_int, [_P]),
    'nr_seed_salt_bump': (C.c_int, [_P, _P]),
    'nr_zero_bytes': (C.c_int, [_P, _SZ, _P]),
    'nr_concat2': (C.c_int, [_P, _SZ, _P, _SZ, _P, _P]),
    'nr_dropout_keep_mask': (C.c_int, [_P, _I64, _F, _U64, _I32, _I32, _I32, _P]),
    'nr_gather_rows': (C.c_int, [_P, _I64, _P, _I64, _I32, _P, _P]),
    'nr_attention_pooling_workspace': (_SZ, [_I64, _I32, _I32, _I32, _I32]),
    'nr_attention_pooling_fwd': (C.c_int, [_P] * 9 + [_I64, _I32, _I32, _I32, _I32, _P, _SZ, _P]),
    'nr_attention_pooling_bwd': (C.c_int, [_P] * 11 + [_I64, _I32, _I32, _I32, _I32, _P, _SZ, _P]),
    'nr_mhsa_workspace': (_SZ, [_I64, _I32, _I32, _I32, _I32]),
    'nr_mhsa_fwd': (C.c_int, [_R] + [_P] * 9 + [_I64, _I32, _I32, _I32, _I32, _F, _I32, _P, _SZ, _P]),
    'nr_mhsa_bwd': (C.c_int, [_R] + [_P] * 14 + [_I64, _I32, _I32, _I32, _I32, _F, _I32, _P, _SZ, _P]),
    'nr_conv1d_k3_workspace': (_SZ, [_I64, _I32, _I32, _I32, _I32]),
    'nr_conv1d_k3_fwd': (C.c_int, [_R, _P, _P, _P, _I64, _I32, _I32, _I32, _I32, _P, _SZ, _P]),
    'nr_conv1d_k3_bwd': (C.c_int, [_R, _P, _P, _P, _P, _P, _P, _I64, _I32, _I32, _I32, _I32, _P, _SZ, _P]),
    'nr_embed_dense_workspace': (C.c_size_t, [_I64, _I32]),
    'nr_embed_dense_fwd': (C.c_int, [_P, _I64, _P, _I32, _P, _P, _P, _I64, _I64, _I32, _I32, _P, C.c_size_t, _P]),
    'nr_embed_dense_bwd': (C.c_int, [_P, _I64, _P, _I32, _P, _P, _I64, _P, _P, _P, _I64, _I32, _I32, _P, _SZ, _P]),
    'nr_pad_blend_fwd': (C.c_int, [_P, _P, _P, _P, _I64, _I32, _I32, _P]),
    'nr_pad_blend_bwd': (C.c_int, [_P, _P, _P, _P, _I64, _I32, _I32, _P]),
    'nr_score_ce_fwd': (C.c_int, [_P] * 7 + [_I64, _I32, _I32, _P]),
    'nr_score_ce_fwd_noalloc': (C.c_int, [_P] * 6 + [_I64, _I32, _I32, _P]),
    'nr_score_ce_bwd': (C.c_int, [_P] * 7 + [_I64, _I32, _I32, _P]),
    'nr_adam_step': (C.c_int, [_P, _P, _P, _P, _I64, _F, _F, _F, _F, _I32, _F, _P]),
    'nr_adam_step_dev': (C.c_int, [_P, _P, _P, _P, _P, _P, _I64, _I64, _F, _F, _F, _F, _P, _F, _P]),
    'nr_tc_gemm_nt': (C.c_int, [_P, _I64, _I32, _I64, _P, _I32, _I64, _P, _P, _I64, _I32, _I64, _I32, _P, _P]),
    'nr_tc_gemm_tn': (C.c_int, [_P, _I32, _I64, _P, _I32, _I64, _I64, _I32, _P, _P, _I64, _I64, _P]),
    'nr_title_encoder_save_bytes': (_SZ, [_E, _I64]),
    'nr_title_encoder_workspace': (_SZ, [_E, _I64]),
    'nr_title_encoder_infer_bytes': (_SZ, [_E, _I64]),
    'nr_title_encoder_fwd': (C.c_int, [_E, _R, _I64, _P, _P, _SZ, _P, _SZ, _P]),
    'nr_title_encoder_bwd': (C.c_int, [_E, _R, _I64, _P, _P, _SZ, _G, _P, _SZ, _P]),
    'nr_batch_train': (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _I64, _I64, _I32, _I32, _I32, _P, _P, _P, _P]),
    'nr_batch_eval_history': (C.c_int, [_P, _P, _P, _I64, _I64, _I32, _P, _P, _P]),
    'nr_eval_score': (C.c_int, [_P, _I64, _P, _P, _P, _P, _I64, _I64, _I32, _P]),
    'nr_eval_rank_metrics': (C.c_int, [_P, _P, _P, _P, _P, _I64, _P]),
    'nr_eval_reduce_metrics': (C.c_int, [_P, _P, _P, _I64, _P]),
}

for _name, (_res, _args) in _SIGS.items():
    _fn = getattr(lib, _name)      # AttributeError here = header/library mismatch: fail loudly
    _fn.restype = _res
    _fn.argtypes = _args

if lib.nr_abi_version() != 1:
    raise ImportError(f'libnr_sm100.so ABI version {lib.nr_abi_version()} != 1')

EXPORTS = tuple(_SIGS)


def last_error():
    return lib.nr_last_error().decode('utf-8', 'replace')


def check(rc, what):
    if rc != 0:
        raise RuntimeError(f'{what} failed (rc={rc}): {last_error()}')


def call(name, *args):
    check(getattr(lib, name)(*args), name)


def ptr(t):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError('newsrecommendation_b200: CUDA tensor required (no CPU path exists)')
    if not t.is_contiguous():
        raise RuntimeError('newsrecommendation_b200: contiguous tensor required')
    return t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


def f32(t):
    return t if (t.dtype == torch.float32 and t.is_contiguous()) else t.contiguous().float()


def make_rows(mode, *, dense=None, table=None, ids=None, ids_stride=0, drop_p=0.0, seed=0):
    r = NrRows()
    r.mode = mode
    r.ids_stride = int(ids_stride)
    r.dense = ptr(dense)
    r.table = ptr(table)
    r.ids = ptr(ids)
    r.table_rows = 0 if table is None else table.shape[0]
    r.drop_p = float(drop_p)
    r.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    return r


def prof_report():
    """{kernel name: (launches, total_ms)} recorded since nr_prof_enable(1); clears the log."""
    buf = C.create_string_buffer(1 << 16)
    lib.nr_prof_report(buf, len(buf))
    out = {}
    for line in buf.value.decode().splitlines():
        name, n, ms = line.rsplit(' ', 2)
        out[name] = (int(n), float(ms))
    return out


def workspace(nbytes, device):
    return torch.empty(max(256, int(nbytes)), dtype=torch.uint8, device=device)
