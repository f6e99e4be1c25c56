"""ctypes binding of the C-ABI library (include/espnet_b200_ctc.h).

There is no CPU fallback: if the CUDA library is missing this module raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# (B200_CTC_LIB: a developer build of the same library, e.g. one with -DB200_SWEEP_CLOCKS)
LIB_PATH = os.environ.get("B200_CTC_LIB") or os.path.join(_HERE, "libespnet_b200_ctc.so")

F32, BF16, F16 = 0, 1, 2
LOGZERO = -10000000000.0

_i64, _int, _f32, _vp, _sz = (ctypes.c_int64, ctypes.c_int, ctypes.c_float, ctypes.c_void_p,
                              ctypes.c_size_t)

# name -> (restype, argtypes); mirrors include/espnet_b200_ctc.h one to one.
SIGNATURES = {
    "b200_ctc_abi_version": (_int, []),
    "b200_last_error": (ctypes.c_char_p, []),
    "b200_ctc_loss_workspace_bytes": (_sz, [_i64, _i64, _i64, _i64]),
    "b200_ctc_loss_forward": (_int, [_vp, _int, _i64, _i64, _i64, _i64, _i64, _vp, _vp, _vp, _vp,
                                     _i64, _int, _vp, _vp, _f32, _vp, _sz, _vp]),
    "b200_ctc_loss_backward": (_int, [_vp, _int, _i64, _i64, _i64, _i64, _i64, _vp, _vp, _vp, _vp,
                                      _i64, _int, _vp, _f32, _vp, _i64, _i64, _vp, _sz, _vp]),
    "b200_ctc_loss_fwd_bwd": (_int, [_vp, _int, _i64, _i64, _i64, _i64, _i64, _vp, _vp, _vp, _vp,
                                     _i64, _int, _vp, _f32, _vp, _vp, _f32, _vp, _i64, _i64, _vp, _sz, _vp]),
    "b200_ctc_loss_stream_pass": (_int, [_vp, _int, _i64, _i64, _i64, _i64, _i64, _vp, _vp, _vp, _vp,
                                         _i64, _int, _vp, _f32, _vp, _i64, _i64, _vp, _sz, _vp]),
    "b200_ctc_prefix_init": (_int, [_vp, _vp, _i64, _i64, _i64, _int, _vp, _vp, _vp]),
    "b200_ctc_prefix_work_floats": (_sz, [_i64, _i64, _i64]),
    "b200_ctc_prefix_score": (_int, [_vp, _vp, _vp, _i64, _i64, _i64, _i64, _int, _int, _vp, _vp,
                                     _vp, _i64, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp]),
    "b200_ctc_prefix_column_stride": (_i64, [_i64]),
    "b200_ctc_prefix_columns": (_int, [_vp, _i64, _i64, _i64, _vp, _vp]),
    "b200_ctc_prefix_score_columns": (_int, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _int, _int, _vp, _vp,
                                             _vp, _i64, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp]),
    "b200_ctc_prefix_select": (_int, [_vp, _i64, _i64, _i64, _i64, _int, _vp, _vp, _vp, _i64,
                                      _i64, _i64, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _vp]),
    "b200_ctc_debug_force_log_sweep": (_int, [_int]),
    "b200_beam_prune": (_int, [_vp, _vp, _vp, _i64, _i64, _i64, _i64, _f32, _f32, _int, _int, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200_pre_beam_candidates": (_int, [_vp, _i64, _i64, _i64, _f32, _int, _vp, _vp]),
    "b200_ctc_debug_prefix_fp32_score": (None, [_int]),
    "b200_ctc_debug_single_cta_sweep": (None, [_int]),
    "b200_ctc_stage_labels": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, _vp, _vp, _vp]),
    "b200_ctc_stage_labels_pinned": (_int, [_vp, _i64, _vp, _vp, _vp, _i64, _vp, _vp]),
    "b200_ctc_loss_status": (_int, [_vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp]),
    "b200_ctc_grad_rescale": (_int, [_vp, _int, _i64, _i64, _i64, _i64, _i64, _vp, _i64, _f32, _vp, _vp]),
    "b200_ctc_greedy": (_int, [_vp, _int, _i64, _i64, _i64, _i64, _i64, _vp, _vp]),
    "b200_ctc_viterbi_work_bytes": (_sz, [_i64, _i64]),
    "b200_ctc_viterbi": (_int, [_vp, _i64, _i64, _vp, _i64, _int, _vp, _vp, _sz, _vp]),
    "b200_peer_exchange_ring_floats": (_sz, [_i64]),
    "b200_peer_sum_exchange": (_int, [_vp, _i64, _i64, _i64, _vp, _f32, _f32, _vp, _vp, _vp]),
    "b200_ctc_viterbi_static_max_bytes": (_sz, []),
    "b200_ctc_viterbi_static": (_int, [_vp, _i64, _i64, _vp, _i64, _int, _vp, _vp]),
}

_lib = None


def load():
    """Loads the library (once) and binds every symbol of the header."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m espnet_b200.build` "
                "(there is no CPU or PyTorch fallback for the b200 CTC path)")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            try:
                fn = getattr(lib, name)
            except AttributeError:
                raise RuntimeError(f"{LIB_PATH} does not export {name}: rebuild it "
                                   "(`python -m espnet_b200.build --force`)")
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def last_error():
    return load().b200_last_error().decode("utf-8", "replace")


def check(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} failed with status {rc}: {last_error()}")


def dtype_code(torch_dtype):
    import torch

    try:
        return {torch.float32: F32, torch.bfloat16: BF16, torch.float16: F16}[torch_dtype]
    except KeyError:
        raise TypeError(f"b200 CTC kernels take float32, bfloat16 or float16, got {torch_dtype}")
