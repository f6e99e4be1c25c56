"""CTC loss on the B200 kernels: the autograd boundary.

Replaces what ``espnet2/asr/ctc.py:46-52`` (``CTC.loss_fn``) obtains from
``th_pred.log_softmax(2)`` + ``torch.nn.CTCLoss`` and its backward: the log-softmax is
fused into the kernels, so the functions here take PRE-softmax logits.

All device work goes through the C ABI of ``libespnet_b200_ctc.so``
(``include/espnet_b200_ctc.h``); torch only owns the memory and the stream.
"""
import torch

from . import _lib


def _as_i64(x, device):
    if not torch.is_tensor(x):
        x = torch.as_tensor(x)
    return x.to(device=device, dtype=torch.int64, non_blocking=True).contiguous()


def _prep_logits(logits):
    if logits.dim() != 3:
        raise RuntimeError(f"logits must be (T, B, V), got {tuple(logits.shape)}")
    if not logits.is_cuda:
        raise RuntimeError("the b200 CTC path runs on CUDA tensors only (there is no CPU fallback)")
    if logits.stride(2) != 1 or logits.stride(0) < 0 or logits.stride(1) < 0:
        logits = logits.contiguous()
    return logits


def _empty_like_strided(x):
    """An uninitialised tensor with x's shape and (dense) strides, e.g. the (T,B,V) view of a
    (B,T,V)-contiguous buffer that ``ctc_lo(...).transpose(0, 1)`` is."""
    T, B, V = x.shape
    st, sb = x.stride(0), x.stride(1)
    if sb == T * V and st == V:
        return torch.empty((B, T, V), dtype=x.dtype, device=x.device).transpose(0, 1)
    return torch.empty((T, B, V), dtype=x.dtype, device=x.device)


class _Plan(object):
    """Device-side arguments shared by forward and backward of one call."""

    def __init__(self, logits, targets, tgt_offsets, ilen, olen, lmax, blank):
        dev = logits.device
        self.logits = logits
        self.T, self.B, self.V = logits.shape
        self.targets = _as_i64(targets, dev)
        self.tgt_offsets = _as_i64(tgt_offsets, dev)
        self.ilen = _as_i64(ilen, dev)
        self.olen = _as_i64(olen, dev)
        if self.ilen.numel() != self.B or self.olen.numel() != self.B or self.tgt_offsets.numel() != self.B:
            raise RuntimeError("ilen, olen and tgt_offsets must have one entry per utterance")
        self.lmax = int(lmax)
        self.blank = int(blank)
        self.dtype = _lib.dtype_code(logits.dtype)
        lib = _lib.load()
        nbytes = lib.b200_ctc_loss_workspace_bytes(self.T, self.B, self.V, self.lmax)
        self.workspace = torch.empty(nbytes + 256, dtype=torch.uint8, device=dev)
        base = self.workspace.data_ptr()
        self.ws_ptr = (base + 255) // 256 * 256
        self.ws_bytes = nbytes

    def common_args(self):
        x = self.logits
        return (x.data_ptr(), self.dtype, self.T, self.B, self.V, x.stride(0), x.stride(1),
                self.targets.data_ptr(), self.tgt_offsets.data_ptr(), self.ilen.data_ptr(),
                self.olen.data_ptr(), self.lmax, self.blank)


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


class _CTCLossFunction(torch.autograd.Function):
    """nll (B,) fp32 = -log p(y_b | x_b); backward gives d/d logits."""

    @staticmethod
    def forward(ctx, logits, targets, tgt_offsets, ilen, olen, lmax, blank):
        logits = _prep_logits(logits.detach())
        plan = _Plan(logits, targets, tgt_offsets, ilen, olen, lmax, blank)
        nll = torch.empty(plan.B, dtype=torch.float32, device=logits.device)
        with torch.cuda.device(logits.device):
            rc = _lib.load().b200_ctc_loss_forward(
                *plan.common_args(), nll.data_ptr(), None, plan.ws_ptr, plan.ws_bytes,
                _stream(logits.device))
        _lib.check(rc, "b200_ctc_loss_forward")
        ctx.plan = plan
        return nll

    @staticmethod
    def backward(ctx, grad_nll):
        plan = ctx.plan
        x = plan.logits
        gw = grad_nll.to(dtype=torch.float32).contiguous()
        grad = _empty_like_strided(x)
        with torch.cuda.device(x.device):
            rc = _lib.load().b200_ctc_loss_backward(
                *plan.common_args(), gw.data_ptr(), 1.0, grad.data_ptr(), grad.stride(0),
                grad.stride(1), plan.ws_ptr, plan.ws_bytes, _stream(x.device))
        _lib.check(rc, "b200_ctc_loss_backward")
        return grad, None, None, None, None, None, None


def upload_valid_frames(x_host, x_dev, ilen):
    """Host -> device copy of a (B,T,V) batch of logits that moves only the frames t < ilen[b] of
    every utterance (one contiguous block per utterance).  The kernels never read padded frames --
    as ``torch.nn.CTCLoss`` ignores them through ``input_lengths`` -- and write exact zeros to their
    gradient rows, so the padding does not have to cross PCIe.  x_host should be pinned; the copies
    are asynchronous on the current stream.  Returns the number of bytes copied."""
    if x_host.shape != x_dev.shape or x_host.dtype != x_dev.dtype or not x_dev.is_cuda:
        raise RuntimeError("upload_valid_frames: x_host and x_dev must be (B,T,V) tensors of one dtype, x_dev on CUDA")
    if not (x_host.is_contiguous() and x_dev.is_contiguous()):
        raise RuntimeError("upload_valid_frames: both buffers must be (B,T,V) contiguous")
    B, T, V = x_host.shape
    nbytes = 0
    for b, n in enumerate(ilen.tolist() if torch.is_tensor(ilen) else list(ilen)):
        n = max(0, min(int(n), T))
        if n:
            x_dev[b, :n].copy_(x_host[b, :n], non_blocking=True)
            nbytes += n * V * x_host.element_size()
    return nbytes


def ctc_nll(logits, targets, tgt_offsets, ilen, olen, lmax, blank=0):
    """Per-utterance negative log-likelihood (B,) fp32 from pre-softmax logits (T,B,V).

    targets: 1-D label buffer; utterance b's labels start at tgt_offsets[b] and there are
    olen[b] of them (so both the concatenated form of ``torch.nn.CTCLoss`` and a padded
    (B, Lmax) matrix are accepted without copying).  lmax: host-side bound on olen.
    """
    return _CTCLossFunction.apply(logits, targets, tgt_offsets, ilen, olen, lmax, blank)


def ctc_loss_fwd_bwd(logits, targets, tgt_offsets, ilen, olen, lmax, blank=0, grad_weight=None,
                     grad_scale=1.0, grad_out=None, plan=None, nll_out=None, loss_sum_out=None):
    """One fused call: returns (nll (B,), loss_sum (1,), grad (T,B,V) in logits' dtype).

    grad = grad_scale * grad_weight[b] * d nll_b / d logits.  This is the unit of work of
    ``bench.py`` (CTC loss fwd+bwd) and what a trainer with a known upstream scale uses.
    nll_out (B,) / loss_sum_out (1,) fp32: caller-owned outputs, e.g. a slot of
    ``sharding.LossLedger`` (no allocation and no copy kernel per call).
    """
    logits = _prep_logits(logits)
    if plan is None:
        plan = _Plan(logits, targets, tgt_offsets, ilen, olen, lmax, blank)
    dev = logits.device
    nll = nll_out if nll_out is not None else torch.empty(plan.B, dtype=torch.float32, device=dev)
    loss_sum = loss_sum_out if loss_sum_out is not None else torch.empty(1, dtype=torch.float32, device=dev)
    for name, o, n in (("nll_out", nll, plan.B), ("loss_sum_out", loss_sum, 1)):
        if o.dtype != torch.float32 or o.device != dev or o.numel() != n or not o.is_contiguous():
            raise ValueError(f"{name} must be a contiguous fp32 tensor of {n} element(s) on {dev}")
    grad = grad_out if grad_out is not None else _empty_like_strided(logits)
    gw = None if grad_weight is None else grad_weight.to(device=dev, dtype=torch.float32).contiguous()
    with torch.cuda.device(dev):
        rc = _lib.load().b200_ctc_loss_fwd_bwd(
            *plan.common_args(), None if gw is None else gw.data_ptr(), float(grad_scale),
            nll.data_ptr(), loss_sum.data_ptr(), grad.data_ptr(), grad.stride(0), grad.stride(1),
            plan.ws_ptr, plan.ws_bytes, _stream(dev))
    _lib.check(rc, "b200_ctc_loss_fwd_bwd")
    return nll, loss_sum, grad
