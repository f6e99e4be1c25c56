"""Greedy path and forced alignment on the B200 kernels.

``ctc_argmax`` replaces ``torch.argmax(ctc_lo(hs_pad), dim=2)`` (``espnet2/asr/ctc.py:103-111``,
``espnet/nets/pytorch_backend/ctc.py:144-151``); ``ctc_forced_align`` replaces the numpy double loop of
``CTC.forced_align`` (``espnet/nets/pytorch_backend/ctc.py:153-216``).  Both go through the C ABI
(``b200_ctc_greedy``, ``b200_ctc_viterbi``); there is no CPU fallback.
"""
import torch

from . import _lib


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


def ctc_argmax(logits):
    """(B, T, V) activations (any strides with a contiguous vocabulary axis) -> (B, T) int64."""
    if logits.dim() != 3:
        raise RuntimeError(f"logits must be (B, T, V), got {tuple(logits.shape)}")
    if not logits.is_cuda:
        raise RuntimeError("the b200 CTC path runs on CUDA tensors only (there is no CPU fallback)")
    if logits.stride(2) != 1 or logits.stride(0) < 0 or logits.stride(1) < 0:
        logits = logits.contiguous()
    B, T, V = logits.shape
    ids = torch.empty((B, T), dtype=torch.int64, device=logits.device)
    if B == 0 or T == 0:
        return ids
    with torch.cuda.device(logits.device):
        rc = _lib.load().b200_ctc_greedy(logits.data_ptr(), _lib.dtype_code(logits.dtype), B, T, V,
                                         logits.stride(0), logits.stride(1), ids.data_ptr(),
                                         _stream(logits.device))
    _lib.check(rc, "b200_ctc_greedy")
    return ids


def ctc_forced_align(lpz, y, blank_id=0):
    """lpz (T, V) log-probabilities, y (L,) label ids -> list of T ids (labels and blanks), the best
    alignment exactly as ``CTC.forced_align`` returns it."""
    if not lpz.is_cuda:
        raise RuntimeError("the b200 CTC path runs on CUDA tensors only (there is no CPU fallback)")
    if lpz.dim() != 2:
        raise RuntimeError(f"lpz must be (T, V), got {tuple(lpz.shape)}")
    lpz = lpz.detach().to(torch.float32).contiguous()
    T, V = lpz.shape
    yd = torch.as_tensor(y).to(device=lpz.device, dtype=torch.int64).contiguous().view(-1)
    L = yd.numel()
    lib = _lib.load()
    nbytes = lib.b200_ctc_viterbi_work_bytes(T, L)
    work = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=lpz.device)
    align = torch.empty(T, dtype=torch.int64, device=lpz.device)
    with torch.cuda.device(lpz.device):
        rc = lib.b200_ctc_viterbi(lpz.data_ptr(), T, V, yd.data_ptr() if L else None, L, int(blank_id),
                                  align.data_ptr(), work.data_ptr(), nbytes, _stream(lpz.device))
    _lib.check(rc, "b200_ctc_viterbi")
    return align.cpu().tolist()
