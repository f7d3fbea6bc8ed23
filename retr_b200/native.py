"""ctypes binding of ``libretr_b200.so`` (the C ABI in ``include/retr_b200.h``).

The library is the only implementation of the hot path: if it is missing or a call fails,
this module raises — nothing falls back to PyTorch or the CPU.
"""
import ctypes as C
import os

import torch  # noqa: F401  (loads libcudart.so.12 before the native library needs it)

FP32, BF16 = 0, 1
_LIB = None
LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libretr_b200.so")


class ModelDesc(C.Structure):
    _fields_ = [("device", C.c_int32), ("precision", C.c_int32), ("blocks", C.c_int32 * 4),
                ("dilation", C.c_int32), ("n_images", C.c_int32), ("loc_in_dim", C.c_int32),
                ("n_loc", C.c_int32), ("image_size", C.c_int32), ("d_model", C.c_int32),
                ("n_heads", C.c_int32), ("d_ff", C.c_int32), ("enc_layers", C.c_int32),
                ("dec_layers", C.c_int32), ("max_pos", C.c_int32), ("vocab", C.c_int32),
                ("pre_norm", C.c_int32), ("embed_ln_eps", C.c_float), ("max_batch", C.c_int32)]


class GemmArgs(C.Structure):
    _fields_ = [("A", C.c_void_p), ("W", C.c_void_p), ("C", C.c_void_p),
                ("M", C.c_int64), ("N", C.c_int32), ("K", C.c_int32), ("lda", C.c_int64), ("ldc", C.c_int64),
                ("scale", C.c_void_p), ("bias", C.c_void_p), ("rowbias", C.c_void_p),
                ("rowbias_period", C.c_int32),
                ("residual", C.c_void_p), ("ldr", C.c_int64), ("res_fp32", C.c_int32), ("relu", C.c_int32),
                ("out_fp32", C.c_int32),
                ("kh", C.c_int32), ("kw", C.c_int32), ("stride", C.c_int32), ("pad", C.c_int32),
                ("dil", C.c_int32), ("H", C.c_int32), ("W_in", C.c_int32), ("C_in", C.c_int32),
                ("OH", C.c_int32), ("OW", C.c_int32),
                ("argmax_out", C.c_void_p),
                ("mn_major", C.c_int32), ("ldb", C.c_int64)]


# name -> (restype, argtypes); every symbol include/retr_b200.h declares
SIGNATURES = {
    "retr_last_error": (C.c_char_p, []),
    "retr_abi_version": (C.c_int, []),
    "retr_create": (C.c_int, [C.POINTER(ModelDesc), C.POINTER(C.c_void_p)]),
    "retr_destroy": (None, [C.c_void_p]),
    "retr_load_weight": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.POINTER(C.c_int64), C.c_int32, C.c_int32]),
    "retr_finalize_weights": (C.c_int, [C.c_void_p]),
    "retr_encode": (C.c_int, [C.c_void_p, C.c_int32] + [C.c_void_p] * 6),
    "retr_forward_teacher_forced": (C.c_int, [C.c_void_p] * 7),
    "retr_set_encoder_attention_capture": (C.c_int, [C.c_void_p, C.c_void_p]),
    "retr_decode_greedy": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p,
                                     C.POINTER(C.c_int32), C.c_void_p]),
    "retr_generate_host": (C.c_int, [C.c_void_p, C.c_int32] + [C.c_void_p] * 5 + [C.c_int32] * 3 +
                           [C.c_void_p, C.POINTER(C.c_int32), C.c_void_p]),
    "retr_prefetch_host": (C.c_int, [C.c_void_p, C.c_int32] + [C.c_void_p] * 5),
    "retr_encode_host": (C.c_int, [C.c_void_p, C.c_int32] + [C.c_void_p] * 5 + [C.c_void_p]),
    "retr_decode_host": (C.c_int, [C.c_void_p] + [C.c_int32] * 3 + [C.c_void_p, C.POINTER(C.c_int32), C.c_void_p]),
    "retr_set_encode_cta_cap": (C.c_int, [C.c_void_p, C.c_int32]),
    "retr_encoder_tokens": (C.c_int, [C.c_void_p]),
    "retr_read_state": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "retr_kernel_launches": (C.c_int64, [C.c_void_p]),
    "retr_last_phase_ms": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "retr_op_gemm": (C.c_int, [C.c_int32, C.POINTER(GemmArgs), C.c_void_p]),
    "retr_op_layernorm": (C.c_int, [C.c_void_p] * 4 + [C.c_int64, C.c_int32, C.c_float, C.c_int32, C.c_void_p]),
    "retr_op_decode_ffn": (C.c_int, [C.c_void_p] * 3 + [C.c_float] + [C.c_void_p] * 4 + [C.c_int32, C.c_int32, C.c_void_p]),
    "retr_op_decode_out_ln_proj": (C.c_int, [C.c_void_p] * 6 + [C.c_float] + [C.c_void_p] * 5 + [C.c_int32, C.c_void_p]),
    "retr_op_decode_cross_attention": (C.c_int, [C.c_void_p] * 3 + [C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                                 C.c_void_p, C.c_void_p]),
    "retr_preprocess": (C.c_int, [C.c_void_p] * 4 + [C.c_int32, C.c_int32] + [C.c_void_p] * 6),
    "retr_preprocess_max_side": (C.c_int, [C.c_int32]),
    "retr_op_small_gemm_noalloc": (C.c_int, [C.c_void_p] * 4 + [C.c_int32] * 4 + [C.c_void_p]),
    "retr_train_create": (C.c_int, [C.POINTER(ModelDesc), C.c_int32, C.POINTER(C.c_void_p)]),
    "retr_train_destroy": (None, [C.c_void_p]),
    "retr_train_bind": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64), C.c_int32]),
    "retr_train_set_arena": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]),
    "retr_train_finalize": (C.c_int, [C.c_void_p]),
    "retr_train_forward": (C.c_int, [C.c_void_p, C.c_int32] + [C.c_void_p] * 8 + [C.c_float, C.c_uint64, C.c_void_p]),
    "retr_train_backward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "retr_train_loss_and_grads": (C.c_int, [C.c_void_p, C.c_int32] + [C.c_void_p] * 8 +
                                  [C.c_float, C.c_uint64, C.c_void_p, C.c_void_p]),
    "retr_train_clip_adamw": (C.c_int, [C.c_void_p] + [C.c_float] * 7 + [C.c_int32, C.c_void_p, C.c_void_p]),
    "retr_train_kernel_launches": (C.c_int64, [C.c_void_p]),
    "retr_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "retr_nccl_comm_create": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "retr_nccl_comm_destroy": (None, [C.c_void_p]),
    "retr_train_set_comm": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32]),
    "retr_allreduce_grads": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "retr_train_grad_buckets": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.c_int32]),
    "retr_op_attention": (C.c_int, [C.c_int32] + [C.c_void_p] * 4 + [C.c_int32] * 4 + [C.c_int64] * 4 +
                          [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
}


class NativeError(RuntimeError):
    pass


def lib():
    """The loaded library; raises if it has not been built (``python -m retr_b200.build``)."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise NativeError(f"{LIB_PATH} is missing: build it with `python -m retr_b200.build` "
                              "(there is no PyTorch/CPU fallback for the hot path)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        _LIB = handle
    return _LIB


def check(status):
    if status != 0:
        raise NativeError(lib().retr_last_error().decode("utf-8", "replace"))


def ptr(t):
    """Device/host address of a tensor (None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
