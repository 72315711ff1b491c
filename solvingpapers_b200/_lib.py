"""ctypes binding of libspb200.so (include/spb200.h).  Fails LOUDLY when the library is missing:
there is no CPU or eager-PyTorch fallback anywhere in this package."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SPB_LIB_PATH") or os.path.join(HERE, "libspb200.so")  # override: A/B experiments only

F32, BF16 = 0, 1
NORM_NONE, NORM_RMS, NORM_LAYER = 0, 1, 2
ROPE_NONE, ROPE_TABLE = 0, 1
W_IN_OUT, W_OUT_IN = 0, 1
IMPL_AUTO, IMPL_SIMT, IMPL_TCGEN05 = 0, 1, 2

vp = C.c_void_p
i32, i64, f32, sz = C.c_int32, C.c_int64, C.c_float, C.c_size_t


class BlockDesc(C.Structure):
    _fields_ = [("B", i32), ("T", i32), ("C", i32), ("Hq", i32), ("Hkv", i32), ("hd", i32), ("dtype", i32),
                ("norm_kind", i32), ("norm_eps", f32), ("norm_upcast", i32), ("rope_kind", i32), ("causal", i32),
                ("scale", f32), ("weight_layout", i32), ("qkv_bias", i32), ("out_bias", i32), ("attn_impl", i32),
                ("gemm_impl", i32), ("residual", i32)]


class BlockParams(C.Structure):
    _fields_ = [("norm_w", vp), ("norm_b", vp), ("wq", vp), ("wk", vp), ("wv", vp), ("ldw_q", i64), ("ldw_k", i64),
                ("ldw_v", i64), ("bq", vp), ("bk", vp), ("bv", vp), ("wo", vp), ("ldw_o", i64), ("bo", vp), ("rope", vp)]


class BlockGrads(C.Structure):
    _fields_ = [("dx", vp), ("dnorm_w", vp), ("dnorm_b", vp), ("dwq", vp), ("dwk", vp), ("dwv", vp), ("ldw_q", i64),
                ("ldw_k", i64), ("ldw_v", i64), ("dbq", vp), ("dbk", vp), ("dbv", vp), ("dwo", vp), ("ldw_o", i64),
                ("dbo", vp)]


class WsLayout(C.Structure):
    _fields_ = [("total", sz), ("xhat", sz), ("stats", sz), ("qkv", sz), ("attn_o", sz), ("lse", sz), ("delta", sz),
                ("dattn_o", sz), ("dqkv", sz), ("dq_acc", sz), ("dxhat", sz), ("partial", sz), ("partial_bytes", sz),
                ("ld_qkv", i64), ("attn_scratch", sz), ("attn_scratch_bytes", sz), ("gemm_scratch", sz),
                ("gemm_scratch_bytes", sz)]


class FfnDesc(C.Structure):
    _fields_ = [("N", i32), ("C", i32), ("H", i32), ("dtype", i32), ("norm_kind", i32), ("norm_eps", f32), ("norm_upcast", i32),
                ("act", i32), ("weight_layout", i32), ("residual", i32), ("gemm_impl", i32)]


class FfnParams(C.Structure):
    _fields_ = [("norm_w", vp), ("norm_b", vp), ("w1", vp), ("w3", vp), ("ldw_13", i64), ("w2", vp), ("ldw_2", i64), ("b1", vp),
                ("b2", vp)]


class FfnGrads(C.Structure):
    _fields_ = [("dx", vp), ("dnorm_w", vp), ("dnorm_b", vp), ("dw1", vp), ("dw3", vp), ("ldw_13", i64), ("dw2", vp),
                ("ldw_2", i64), ("db1", vp), ("db2", vp)]


class FfnWs(C.Structure):
    _fields_ = [("total", sz), ("xhat", sz), ("stats", sz), ("g13", sz), ("act", sz), ("dact", sz), ("dg13", sz), ("dxhat", sz),
                ("partial", sz), ("partial_bytes", sz)]


class DecodeWs(C.Structure):
    _fields_ = [("total", sz), ("xhat", sz), ("qkv", sz), ("attn_o", sz), ("gemm_part", sz), ("attn_part", sz)]


ACT_SWIGLU, ACT_GEGLU, ACT_GELU_TANH = 0, 1, 2

# name -> (restype, argtypes); every symbol include/spb200.h declares
SIGNATURES = {
    "spb_version": (C.c_char_p, []),
    "spb_last_error": (C.c_char_p, []),
    "spb_device_info": (C.c_int, [C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(sz)]),
    "spb_launch_count": (i64, []),
    "spb_prof_enable": (C.c_int, [C.c_int]),
    "spb_prof_steps": (C.c_int, []),
    "spb_prof_read": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_char_p), C.POINTER(f32)]),
    "spb_norm_fwd": (C.c_int, [C.c_int, C.c_int, C.c_int, f32, vp, vp, vp, vp, vp, i64, i32, vp]),
    "spb_norm_bwd_partial_bytes": (sz, [i64, i32]),
    "spb_norm_bwd": (C.c_int, [C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, vp]),
    "spb_rope_apply": (C.c_int, [C.c_int, vp, i64, i64, i32, i32, i32, vp, C.c_int, vp]),
    "spb_gemm": (C.c_int, [C.c_int, C.c_int, C.c_int, i64, i64, i64, vp, i64, C.c_int, vp, i64, C.c_int, vp, i64, vp,
                           C.c_int, vp]),
    "spb_colsum": (C.c_int, [C.c_int, vp, i64, i64, i32, vp, vp, vp]),
    "spb_colsum_static": (C.c_int, [C.c_int, vp, i64, i64, i32, vp, vp]),
    "spb_flash_fwd": (C.c_int, [C.c_int, C.c_int, i32, i32, i32, i32, i32, C.c_int, f32, vp, i64, vp, i64, vp, i64, vp,
                                i64, vp, vp]),
    "spb_flash_bwd": (C.c_int, [C.c_int, C.c_int, i32, i32, i32, i32, i32, C.c_int, f32, vp, i64, vp, i64, vp, i64, vp,
                                i64, vp, i64, vp, vp, vp, i64, vp, i64, vp, i64, vp, vp]),
    "spb_attn_wide_scratch_bytes": (sz, [i32, i32, i32, i32, i32]),
    "spb_attn_wide_fwd": (C.c_int, [i32, i32, i32, i32, i32, C.c_int, f32, vp, i64, vp, i64, vp, i64, vp, i64, vp, vp, vp]),
    "spb_attn_wide_bwd": (C.c_int, [i32, i32, i32, i32, i32, C.c_int, f32, vp, i64, vp, i64, vp, i64, vp, i64, vp, i64, vp, vp, vp, i64, vp, i64,
                                    vp, i64, vp, vp]),
    "spb_block_ws_layout": (C.c_int, [C.POINTER(BlockDesc), C.POINTER(WsLayout)]),
    "spb_block_fwd": (C.c_int, [C.POINTER(BlockDesc), C.POINTER(BlockParams), vp, vp, vp, vp]),
    "spb_block_bwd": (C.c_int, [C.POINTER(BlockDesc), C.POINTER(BlockParams), vp, vp, vp, C.POINTER(BlockGrads), vp, vp]),
    "spb_ffn_ws_layout": (C.c_int, [C.POINTER(FfnDesc), C.POINTER(FfnWs)]),
    "spb_ffn_fwd": (C.c_int, [C.POINTER(FfnDesc), C.POINTER(FfnParams), vp, vp, vp, vp]),
    "spb_ffn_bwd": (C.c_int, [C.POINTER(FfnDesc), C.POINTER(FfnParams), vp, vp, vp, C.POINTER(FfnGrads), vp, vp]),
    "spb_nll_fwd": (C.c_int, [C.c_int, vp, i64, vp, i64, i32, vp, vp, vp, vp]),
    "spb_nll_bwd": (C.c_int, [C.c_int, vp, i64, vp, i64, i32, vp, f32, vp, i64, vp]),
    "spb_loss_head_ws_bytes": (sz, [i64, i32, i32]),
    "spb_loss_head_fwd": (C.c_int, [C.c_int, vp, i64, vp, i64, vp, i64, i32, i32, vp, vp, vp, vp, vp]),
    "spb_loss_head_bwd": (C.c_int, [C.c_int, vp, i64, vp, i64, vp, i64, i32, i32, vp, f32, vp, i64, vp, i64, vp, vp]),
    "spb_attn_decode_scratch_bytes": (sz, [i32, i32, i32, i32, i32, i32]),
    "spb_attn_decode": (C.c_int, [C.c_int, i32, i32, i32, i32, i32, i32, f32, vp, i64, vp, i64, i64, vp, i64, i64, vp, i64, vp, vp, vp]),
    "spb_block_decode_ws_layout": (C.c_int, [C.POINTER(BlockDesc), i32, C.POINTER(DecodeWs)]),
    "spb_block_decode": (C.c_int, [C.POINTER(BlockDesc), C.POINTER(BlockParams), vp, vp, vp, i32, i32, vp, vp, vp]),
    "spb_block_decode_dev": (C.c_int, [C.POINTER(BlockDesc), C.POINTER(BlockParams), vp, vp, vp, vp, i32, vp, vp, vp]),
    "spb_comm_unique_id": (C.c_int, [vp]),
    "spb_comm_init_rank": (C.c_int, [vp, C.c_int, C.c_int, C.POINTER(vp)]),
    "spb_comm_allreduce_f32": (C.c_int, [vp, C.POINTER(vp), C.POINTER(sz), C.c_int, vp]),
    "spb_comm_destroy": (C.c_int, [vp]),
    "spb_comm_init_all": (C.c_int, [C.c_int, C.POINTER(vp)]),
    "spb_comm_alloc": (C.c_int, [vp, sz, C.POINTER(vp)]),
    "spb_comm_free": (C.c_int, [vp, vp]),
    "spb_comm_buffer_kind": (C.c_int, [vp, vp]),
}

_lib = None


class SpbError(RuntimeError):
    pass


def load():
    """dlopen the in-tree library and bind every entry point.  No fallback: a missing .so is an error."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SpbError(f"{LIB_PATH} is missing - build it with `python -m solvingpapers_b200.build` "
                       "(there is no CPU / PyTorch fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().spb_last_error().decode(errors="replace")
        raise SpbError(f"{what} failed with code {rc}: {msg}")
