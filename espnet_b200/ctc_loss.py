"""CTC loss on the B200 kernels: the autograd boundary.

Replaces what ``espnet2/asr/ctc.py:46-52`` (``CTC.loss_fn``) obtains from
``th_pred.log_softmax(2)`` + ``torch.nn.CTCLoss`` and its backward: the log-softmax is
fused into the kernels, so the functions here take PRE-softmax logits.

All device work goes through the C ABI of ``libespnet_b200_ctc.so``
(``include/espnet_b200_ctc.h``); torch only owns the memory and the stream.

How autograd reaches the ONE-PASS kernels.  ``b200_ctc_loss_fwd_bwd`` reads the logits once and
writes the gradient in the same call, but needs the upstream gradient d loss / d nll_b at forward
time.  ``ctc_nll`` therefore runs it in ``forward`` with an EXPECTED upstream gradient
(``grad_scale_hint``: what the caller will multiply nll_b by, 1/B in ``CTC.loss_fn``), keeps the
finished gradient, and ``backward`` hands it out after ``b200_ctc_grad_rescale`` -- a kernel that
compares the real upstream gradient with the expectation ON THE DEVICE (no host synchronisation)
and touches only the rows whose ratio is not exactly 1.  When the trainer's scale equals the hint,
``loss.backward()`` costs one launch of a kernel that reads B floats.
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


def pack_labels(targets, tgt_offsets, ilen, olen, device):
    """ONE host->device copy for the label side of a call: [targets | tgt_offsets | ilen | olen] as int64
    in one pinned staging buffer (a step otherwise issues four small pageable copies, each a host
    synchronisation).  Inputs that already live on `device` are used as they are.
    Returns (targets, tgt_offsets, ilen, olen) as int64 device tensors."""
    parts = [targets, tgt_offsets, ilen, olen]
    parts = [p if torch.is_tensor(p) else torch.as_tensor(p) for p in parts]
    host = [k for k, p in enumerate(parts) if not p.is_cuda]
    out = list(parts)
    if host:
        sizes = [parts[k].numel() for k in host]
        stage = torch.empty(sum(sizes), dtype=torch.int64, pin_memory=torch.device(device).type == "cuda")
        o = 0
        for k, n in zip(host, sizes):
            stage[o:o + n].copy_(parts[k].reshape(-1))
            o += n
        dev = stage.to(device, non_blocking=True)
        o = 0
        for k, n in zip(host, sizes):
            out[k] = dev[o:o + n]
            o += n
    return tuple(_as_i64(p, device) for p in out)


class _Plan(object):
    """Device-side arguments shared by the kernels of one call."""

    def __init__(self, logits, targets, tgt_offsets, ilen, olen, lmax, blank):
        dev = logits.device
        self.logits = logits
        self.T, self.B, self.V = logits.shape
        self.targets, self.tgt_offsets, self.ilen, self.olen = pack_labels(targets, tgt_offsets, ilen, olen, dev)
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

    def invalid_utterances(self):
        """Number of utterances of the last call on this plan whose targets were invalid (a label outside
        [0,V) or equal to blank, or olen > lmax).  Synchronises the stream."""
        import ctypes

        n = ctypes.c_int32(0)
        with torch.cuda.device(self.logits.device):
            rc = _lib.load().b200_ctc_loss_status(self.ws_ptr, self.T, self.B, self.V, self.lmax, ctypes.byref(n),
                                                  _stream(self.logits.device))
        _lib.check(rc, "b200_ctc_loss_status")
        return int(n.value)


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


def _check_strict(plan):
    n = plan.invalid_utterances()
    if n:
        raise ValueError(f"{n} utterance(s) with invalid CTC targets: labels must lie in [0, {plan.V}) and differ "
                         f"from blank={plan.blank}, and olen must not exceed lmax={plan.lmax}")


class _CTCLossFunction(torch.autograd.Function):
    """nll (B,) fp32 = -log p(y_b | x_b); backward gives d/d logits.

    With a gradient expected for the logits, forward runs the one-pass fused call with
    grad_scale = `hint` and keeps the finished gradient; backward rescales it on the device where
    the upstream gradient differs from the hint.  Without one (evaluation), forward only runs
    the row statistics and the lattice sweeps."""

    @staticmethod
    def forward(ctx, logits, targets, tgt_offsets, ilen, olen, lmax, blank, hint, strict):
        logits = _prep_logits(logits.detach())
        plan = _Plan(logits, targets, tgt_offsets, ilen, olen, lmax, blank)
        dev = logits.device
        nll = torch.empty(plan.B, dtype=torch.float32, device=dev)
        ctx.hint = float(hint)
        if ctx.needs_input_grad[0]:
            grad = _empty_like_strided(logits)
            with torch.cuda.device(dev):
                rc = _lib.load().b200_ctc_loss_fwd_bwd(
                    *plan.common_args(), None, ctx.hint, nll.data_ptr(), None, grad.data_ptr(),
                    grad.stride(0), grad.stride(1), plan.ws_ptr, plan.ws_bytes, _stream(dev))
            _lib.check(rc, "b200_ctc_loss_fwd_bwd")
            ctx.save_for_backward(grad, plan.ilen)
        else:
            with torch.cuda.device(dev):
                rc = _lib.load().b200_ctc_loss_forward(
                    *plan.common_args(), nll.data_ptr(), None, plan.ws_ptr, plan.ws_bytes, _stream(dev))
            _lib.check(rc, "b200_ctc_loss_forward")
        if strict:
            _check_strict(plan)
        # the workspace goes back to the allocator here: reuse is ordered on this stream
        return nll

    @staticmethod
    def backward(ctx, grad_nll):
        grad, ilen = ctx.saved_tensors
        T, B, V = grad.shape
        gw = grad_nll.to(dtype=torch.float32).reshape(-1)
        if gw.numel() > 1 and gw.stride(0) == 0:  # the expanded scalar that `nll.sum()` sends back
            gw = gw[:1]
        gw = gw.contiguous()
        with torch.cuda.device(grad.device):
            rc = _lib.load().b200_ctc_grad_rescale(
                grad.data_ptr(), _lib.dtype_code(grad.dtype), T, B, V, grad.stride(0), grad.stride(1),
                gw.data_ptr(), gw.numel(), ctx.hint, ilen.data_ptr(), _stream(grad.device))
        _lib.check(rc, "b200_ctc_grad_rescale")
        return grad, None, None, None, None, None, None, None, None


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


def ctc_nll(logits, targets, tgt_offsets, ilen, olen, lmax, blank=0, grad_scale_hint=1.0, strict=False):
    """Per-utterance negative log-likelihood (B,) fp32 from pre-softmax logits (T,B,V).

    targets: 1-D label buffer; utterance b's labels start at tgt_offsets[b] and there are
    olen[b] of them (so both the concatenated form of ``torch.nn.CTCLoss`` and a padded
    (B, Lmax) matrix are accepted without copying).  lmax: host-side bound on olen.
    grad_scale_hint: the upstream gradient d loss / d nll_b the caller expects (see the module
    docstring); any other upstream gradient is still honoured, at the price of one more pass over
    the gradient.  strict: raise ValueError for invalid targets (synchronises; without it such
    utterances get a NaN loss and NaN gradient rows).
    """
    return _CTCLossFunction.apply(logits, targets, tgt_offsets, ilen, olen, lmax, blank, grad_scale_hint, strict)


def ctc_loss_fwd_bwd(logits, targets, tgt_offsets, ilen, olen, lmax, blank=0, grad_weight=None,
                     grad_scale=1.0, grad_out=None, plan=None, nll_out=None, loss_sum_out=None):
    """One fused call: returns (nll (B,), loss_sum (1,), grad (T,B,V) in logits' dtype).

    grad = grad_scale * grad_weight[b] * d nll_b / d logits.  The raw-ABI form of the unit of
    work of ``bench.py`` (CTC loss fwd+bwd) for a trainer that knows its upstream scale.
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


def ctc_loss_split(logits, targets, tgt_offsets, ilen, olen, lmax, blank=0, grad_weight=None, grad_scale=1.0):
    """The separate forward / backward entry points (``b200_ctc_loss_forward`` then
    ``b200_ctc_loss_backward`` on the same workspace: the logits are read twice).  Kept for callers that
    cannot hold a gradient-sized buffer between forward and backward; the parity tests run it next to the
    one-pass call.  Returns (nll, grad)."""
    logits = _prep_logits(logits)
    plan = _Plan(logits, targets, tgt_offsets, ilen, olen, lmax, blank)
    dev = logits.device
    nll = torch.empty(plan.B, dtype=torch.float32, device=dev)
    grad = _empty_like_strided(logits)
    gw = None if grad_weight is None else grad_weight.to(device=dev, dtype=torch.float32).contiguous()
    lib = _lib.load()
    with torch.cuda.device(dev):
        rc = lib.b200_ctc_loss_forward(*plan.common_args(), nll.data_ptr(), None, plan.ws_ptr, plan.ws_bytes,
                                       _stream(dev))
        _lib.check(rc, "b200_ctc_loss_forward")
        rc = lib.b200_ctc_loss_backward(*plan.common_args(), None if gw is None else gw.data_ptr(), float(grad_scale),
                                        grad.data_ptr(), grad.stride(0), grad.stride(1), plan.ws_ptr, plan.ws_bytes,
                                        _stream(dev))
        _lib.check(rc, "b200_ctc_loss_backward")
    return nll, grad
