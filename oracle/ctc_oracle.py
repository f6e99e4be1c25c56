"""ctypes binding of oracle/libctc_oracle.so (see ctc_oracle.c for citations).

TEST INFRASTRUCTURE ONLY.  Every function takes and returns numpy arrays.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

LOGZERO = -10000000000.0


def build(force=False):
    so = os.path.join(_HERE, "libctc_oracle.so")
    src = os.path.join(_HERE, "ctc_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.oracle_num_threads.restype = ctypes.c_int
    return _LIB


def _p(a, ct):
    return None if a is None else a.ctypes.data_as(ctypes.POINTER(ct))


def num_threads():
    return int(lib().oracle_num_threads())


def set_num_threads(n):
    lib().oracle_set_num_threads(ctypes.c_int(int(n)))


def ctc_loss(logits_tbv, targets, tgt_offsets, ilen, olen, blank=0, gw=None,
             grad_scale=1.0, want_grad=True):
    """logits_tbv: float32 (T,B,V) (any strides with V contiguous).

    Returns (nll float64 (B,), grad float64 (T,B,V) or None).
    """
    x = np.asarray(logits_tbv)
    assert x.dtype == np.float32 and x.ndim == 3 and x.strides[2] == 4
    T, B, V = x.shape
    st, sb = x.strides[0] // 4, x.strides[1] // 4
    targets = np.ascontiguousarray(targets, dtype=np.int64)
    tgt_offsets = np.ascontiguousarray(tgt_offsets, dtype=np.int64)
    ilen = np.ascontiguousarray(ilen, dtype=np.int64)
    olen = np.ascontiguousarray(olen, dtype=np.int64)
    nll = np.empty(B, dtype=np.float64)
    grad = np.empty((T, B, V), dtype=np.float64) if want_grad else None
    gwa = None if gw is None else np.ascontiguousarray(gw, dtype=np.float64)
    rc = lib().oracle_ctc_loss(
        _p(x, ctypes.c_float), ctypes.c_int64(T), ctypes.c_int64(B), ctypes.c_int64(V),
        ctypes.c_int64(st), ctypes.c_int64(sb),
        _p(targets, ctypes.c_int64), _p(tgt_offsets, ctypes.c_int64),
        _p(ilen, ctypes.c_int64), _p(olen, ctypes.c_int64), ctypes.c_int(blank),
        _p(gwa, ctypes.c_double), ctypes.c_double(grad_scale),
        _p(nll, ctypes.c_double), _p(grad, ctypes.c_double))
    if rc != 0:
        raise ValueError("oracle_ctc_loss: invalid lengths")
    return nll, grad


def prefix_init(x_btv, xlens, blank=0):
    """Pads x in place (float32 (B,T,V) contiguous) and returns r0 (T,2,B)."""
    assert x_btv.dtype == np.float32 and x_btv.flags["C_CONTIGUOUS"]
    B, T, V = x_btv.shape
    xlens = np.ascontiguousarray(xlens, dtype=np.int64)
    r0 = np.empty((T, 2, B), dtype=np.float32)
    lib().oracle_prefix_init(_p(x_btv, ctypes.c_float), _p(xlens, ctypes.c_int64),
                             ctypes.c_int64(B), ctypes.c_int64(T), ctypes.c_int64(V),
                             ctypes.c_int(blank), _p(r0, ctypes.c_float))
    return r0


def prefix_score(x_btv, xlens, n_hyps, blank, eos, r_prev, s_prev, last_ids, output_len,
                 scoring_ids=None, start=None, end=None, want_r=False):
    """One CTCPrefixScoreTH.__call__.  Returns (scores, log_psi, r_full or None)."""
    B, T, V = x_btv.shape
    n_bh = B * n_hyps
    xlens = np.ascontiguousarray(xlens, dtype=np.int64)
    r_prev = np.ascontiguousarray(r_prev, dtype=np.float32)
    assert r_prev.shape == (T, 2, n_bh)
    sp = None if s_prev is None else np.ascontiguousarray(s_prev, dtype=np.float32)
    last_ids = np.ascontiguousarray(last_ids, dtype=np.int64)
    sid = None
    snum = 0
    if scoring_ids is not None:
        sid = np.ascontiguousarray(scoring_ids, dtype=np.int64)
        snum = sid.shape[-1]
    if start is None:
        start = max(output_len, 1)
    if end is None:
        end = T
    scores = np.empty((n_bh, V), dtype=np.float32)
    log_psi = np.empty((n_bh, V), dtype=np.float32)
    r_full = np.empty((T, 2, n_bh, snum if snum else V), dtype=np.float32) if want_r else None
    lib().oracle_prefix_score(
        _p(x_btv, ctypes.c_float), _p(xlens, ctypes.c_int64),
        ctypes.c_int64(B), ctypes.c_int64(T), ctypes.c_int64(V), ctypes.c_int64(n_hyps),
        ctypes.c_int(blank), ctypes.c_int(eos),
        _p(r_prev, ctypes.c_float), _p(sp, ctypes.c_float), _p(last_ids, ctypes.c_int64),
        ctypes.c_int64(output_len), _p(sid, ctypes.c_int64), ctypes.c_int64(snum),
        ctypes.c_int64(start), ctypes.c_int64(end),
        _p(scores, ctypes.c_float), _p(log_psi, ctypes.c_float), _p(r_full, ctypes.c_float))
    return scores, log_psi, r_full


def prefix_select(x_btv, n_hyps, blank, r_prev, log_psi, last_ids, output_len, parent, token,
                  start=None, end=None):
    """index_select_state by recomputation.  Returns (r_new (T,2,n_new), s_new (n_new,))."""
    B, T, V = x_btv.shape
    r_prev = np.ascontiguousarray(r_prev, dtype=np.float32)
    n_bh = r_prev.shape[2]
    log_psi = np.ascontiguousarray(log_psi, dtype=np.float32)
    last_ids = np.ascontiguousarray(last_ids, dtype=np.int64)
    parent = np.ascontiguousarray(parent, dtype=np.int64)
    token = np.ascontiguousarray(token, dtype=np.int64)
    n_new = parent.shape[0]
    if start is None:
        start = max(output_len, 1)
    if end is None:
        end = T
    r_new = np.empty((T, 2, n_new), dtype=np.float32)
    s_new = np.empty((n_new,), dtype=np.float32)
    lib().oracle_prefix_select(
        _p(x_btv, ctypes.c_float), ctypes.c_int64(B), ctypes.c_int64(T), ctypes.c_int64(V),
        ctypes.c_int64(n_hyps), ctypes.c_int(blank),
        _p(r_prev, ctypes.c_float), _p(log_psi, ctypes.c_float), _p(last_ids, ctypes.c_int64),
        ctypes.c_int64(output_len), ctypes.c_int64(start), ctypes.c_int64(end),
        ctypes.c_int64(n_bh), _p(parent, ctypes.c_int64), _p(token, ctypes.c_int64),
        ctypes.c_int64(n_new), _p(r_new, ctypes.c_float), _p(s_new, ctypes.c_float))
    return r_new, s_new


def viterbi(lpz_tv, y, blank=0):
    """CTC.forced_align: returns the per-frame label sequence (T,) int64."""
    lpz = np.ascontiguousarray(lpz_tv, dtype=np.float32)
    T, V = lpz.shape
    y = np.ascontiguousarray(y, dtype=np.int64)
    align = np.empty((T,), dtype=np.int64)
    lib().oracle_ctc_viterbi(_p(lpz, ctypes.c_float), ctypes.c_int64(T), ctypes.c_int64(V),
                             _p(y, ctypes.c_int64), ctypes.c_int64(y.shape[0]),
                             ctypes.c_int(blank), _p(align, ctypes.c_int64))
    return align
