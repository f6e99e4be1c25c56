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
import numpy as np
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


class _PinnedRing(object):
    """A few pinned host buffers per device for the label side of a call, reused round-robin.  A slot is reused
    only after the copy that last read it has completed (one CUDA event per slot, normally long done)."""

    SLOTS = 16

    def __init__(self):
        n = self.SLOTS
        self.buf, self.np, self.ptr, self.ev, self.i = [None] * n, [None] * n, [0] * n, [None] * n, 0

    def take(self, n):
        k = self.i
        self.i = (k + 1) % self.SLOTS
        if self.ev[k] is not None:
            self.ev[k].synchronize()
        if self.buf[k] is None or self.buf[k].numel() < n:
            self.buf[k] = torch.empty(max(n, 16384), dtype=torch.int64, pin_memory=True)
            self.np[k] = self.buf[k].numpy()
            self.ptr[k] = self.buf[k].data_ptr()
        return k

    def sent(self, k, stream):
        if self.ev[k] is None:
            self.ev[k] = torch.cuda.Event()
        self.ev[k].record(stream)


_rings = {}


def _ring(device):
    r = _rings.get(device)
    if r is None:
        r = _rings[device] = _PinnedRing()
    return r


def _h64(x):
    """A host-side label / length vector as a contiguous int64 CPU tensor (no copy when it already is one)."""
    if not torch.is_tensor(x):
        x = torch.as_tensor(x)
    if x.dtype != torch.int64:
        x = x.to(torch.int64)
    return x if x.is_contiguous() else x.contiguous()


def _is_host(x):
    return not (torch.is_tensor(x) and x.is_cuda)


def _host_np(x):
    """int64 numpy view / copy of a host-side length or label vector (tensor, list, numpy)."""
    if torch.is_tensor(x):
        x = x.detach().numpy()
    return np.asarray(x).reshape(-1)


def pack_labels(targets, tgt_offsets, ilen, olen, device):
    """The label side of a call as int64 device memory: (targets, tgt_offsets, ilen, olen) tensors.

    Everything that lives on the host travels in ONE pinned staging buffer and ONE asynchronous copy (a step
    otherwise issues four small pageable copies, each a host synchronisation); inputs that already live on
    `device` are used as they are.  tgt_offsets=None means "targets are concatenated": the offsets are the
    exclusive prefix sum of olen (computed on the host when olen is there)."""
    device = torch.device(device)
    parts = [targets, tgt_offsets, ilen, olen]
    on_dev = [torch.is_tensor(p) and p.is_cuda for p in parts]
    if tgt_offsets is None:
        if on_dev[3]:
            o64 = olen.to(torch.int64)
            parts[1], on_dev[1] = torch.cumsum(o64, 0) - o64, True
        else:
            o = _host_np(olen).astype(np.int64, copy=False)
            parts[1] = np.cumsum(o) - o
    out = [None] * 4
    host = [k for k in range(4) if not on_dev[k]]
    if host and device.type != "cuda":   # (CPU tensors: only the argument plumbing is exercised, by the CPU tests)
        for k in host:
            out[k] = torch.from_numpy(np.ascontiguousarray(_host_np(parts[k]), dtype=np.int64))
        host = []
    if host:
        arrs = [_host_np(parts[k]) for k in host]
        n = sum(a.size for a in arrs)
        ring = _ring(device)
        slot = ring.take(n)
        stage = ring.np[slot]
        o = 0
        for a in arrs:
            stage[o:o + a.size] = a
            o += a.size
        dev = torch.empty(n, dtype=torch.int64, device=device)
        dev.copy_(ring.buf[slot][:n], non_blocking=True)
        ring.sent(slot, torch.cuda.current_stream(device))
        o = 0
        for k, a in zip(host, arrs):
            out[k] = dev[o:o + a.size]
            o += a.size
    for k in range(4):
        if out[k] is None:
            out[k] = parts[k].to(device=device, dtype=torch.int64).contiguous()
    return tuple(out)


_ws_bytes_cache = {}


class _Plan(object):
    """Device-side arguments shared by the kernels of one call (raw pointers; `keep` owns the memory)."""

    def __init__(self, logits, targets, tgt_offsets, ilen, olen, lmax, blank):
        dev = logits.device
        self.logits = logits
        self.T, self.B, self.V = T, B, V = logits.shape
        self.lmax = lmax = int(lmax)
        self.blank = int(blank)
        self.dtype = _lib.dtype_code(logits.dtype)
        self.stream = stream = torch._C._cuda_getCurrentRawStream(dev.index) if dev.type == "cuda" else 0
        lib = _lib.load()
        key = (T, B, V, lmax)
        nbytes = _ws_bytes_cache.get(key)
        if nbytes is None:
            nbytes = _ws_bytes_cache[key] = lib.b200_ctc_loss_workspace_bytes(*key)
        self.ws_bytes = nbytes
        if (dev.type == "cuda" and _is_host(targets) and _is_host(ilen) and _is_host(olen)
                and (tgt_offsets is None or _is_host(tgt_offsets))):
            # the usual trainer-side call: labels and lengths arrive from the data loader on the host.  The library
            # lays them out in one of its pinned staging buffers and issues ONE asynchronous copy; the device copy,
            # the nll vector and the workspace share ONE allocation
            tg, il, ol = _h64(targets), _h64(ilen), _h64(olen)
            off = None if tgt_offsets is None else _h64(tgt_offsets)
            if il.numel() != B or ol.numel() != B or (off is not None and off.numel() != B):
                raise RuntimeError("ilen, olen and tgt_offsets must have one entry per utterance")
            nt = tg.numel()
            lab = (8 * (nt + 3 * B) + 255) & ~255
            nlb = (4 * B + 255) & ~255
            self.keep = pool = torch.empty(lab + nlb + nbytes + 256, dtype=torch.uint8, device=dev)
            base = (pool.data_ptr() + 255) & ~255
            with _on(dev):
                rc = lib.b200_ctc_stage_labels_pinned(tg.data_ptr(), nt, None if off is None else off.data_ptr(),
                                                      il.data_ptr(), ol.data_ptr(), B, base, stream)
            if rc:
                _lib.check(rc, "b200_ctc_stage_labels_pinned")
            self.targets_ptr, self.offsets_ptr = base, base + 8 * nt
            self.ilen_ptr, self.olen_ptr = base + 8 * (nt + B), base + 8 * (nt + 2 * B)
            self.nll_ptr = base + lab
            self.ws_ptr = base + lab + nlb
        else:
            t = pack_labels(targets, tgt_offsets, ilen, olen, dev)
            if t[2].numel() != B or t[3].numel() != B or t[1].numel() != B:
                raise RuntimeError("ilen, olen and tgt_offsets must have one entry per utterance")
            self.targets_ptr, self.offsets_ptr, self.ilen_ptr, self.olen_ptr = (x.data_ptr() for x in t)
            nlb = (4 * B + 255) & ~255
            workspace = torch.empty(nlb + nbytes + 256, dtype=torch.uint8, device=dev)
            self.keep = (t, workspace)
            self.nll_ptr = (workspace.data_ptr() + 255) // 256 * 256
            self.ws_ptr = self.nll_ptr + nlb

    def common_args(self):
        x = self.logits
        return (x.data_ptr(), self.dtype, self.T, self.B, self.V, x.stride(0), x.stride(1),
                self.targets_ptr, self.offsets_ptr, self.ilen_ptr, self.olen_ptr, self.lmax, self.blank)

    def invalid_utterances(self):
        """Number of utterances of the last call on this plan whose targets were invalid (a label outside
        [0,V) or equal to blank, or olen > lmax).  Synchronises the stream."""
        import ctypes

        return self.status()[0]

    def status(self):
        """(invalid utterances, utterances the log-domain sweeps had to redo) of the last call on this plan."""
        import ctypes

        n, r = ctypes.c_int32(0), ctypes.c_int32(0)
        with torch.cuda.device(self.logits.device):
            rc = _lib.load().b200_ctc_loss_status(self.ws_ptr, self.T, self.B, self.V, self.lmax, ctypes.byref(n),
                                                  ctypes.byref(r), self.stream)
        _lib.check(rc, "b200_ctc_loss_status")
        return int(n.value), int(r.value)


def _stream(device):
    return torch._C._cuda_getCurrentRawStream(device.index)


class _on(object):
    """`with _on(device)`: torch.cuda.device(device) only when it is not already the current device."""

    __slots__ = ("guard",)

    def __init__(self, device):
        self.guard = None if torch.cuda.current_device() == device.index else torch.cuda.device(device)

    def __enter__(self):
        if self.guard is not None:
            self.guard.__enter__()

    def __exit__(self, *a):
        if self.guard is not None:
            self.guard.__exit__(*a)


def _check_strict(plan):
    n = plan.invalid_utterances()
    if n:
        raise ValueError(f"{n} utterance(s) with invalid CTC targets: labels must lie in [0, {plan.V}) and differ "
                         f"from blank={plan.blank}, and olen must not exceed lmax={plan.lmax}")


class _CTCLossFunction(torch.autograd.Function):
    """reduce_scale is None:  nll (B,) fp32 = -log p(y_b | x_b);
    reduce_scale = s:         the 0-dim loss s * sum_b nll_b, summed inside the lattice kernel (CTC.loss_fn's
                              CTCLoss(reduction="sum") / B is s = 1/B: no reduction or division kernel follows).
    backward gives d/d logits.

    With a gradient expected for the logits, forward runs the one-pass fused call with the EXPECTED upstream
    gradient `hint` and keeps the finished gradient; backward rescales it on the device where the real upstream
    gradient differs.  Without one (evaluation), forward only runs the row statistics and the lattice sweeps."""

    @staticmethod
    def forward(ctx, logits, targets, tgt_offsets, ilen, olen, lmax, blank, hint, strict, reduce_scale):
        logits = _prep_logits(logits)
        plan = _Plan(logits, targets, tgt_offsets, ilen, olen, lmax, blank)
        dev = logits.device
        reduced = reduce_scale is not None
        if reduced:   # the per-utterance nll is not returned: it lives in the plan's allocation
            loss = torch.empty((), dtype=torch.float32, device=dev)
            nll_ptr, loss_ptr, s = plan.nll_ptr, loss.data_ptr(), float(reduce_scale)
        else:
            nll = torch.empty(plan.B, dtype=torch.float32, device=dev)
            nll_ptr, loss_ptr, s = nll.data_ptr(), None, 1.0
        ctx.hint = float(hint)
        lib = _lib.load()
        if ctx.needs_input_grad[0]:
            grad = _empty_like_strided(logits)
            with _on(dev):
                rc = lib.b200_ctc_loss_fwd_bwd(
                    *plan.common_args(), None, ctx.hint * s, nll_ptr, loss_ptr, s, grad.data_ptr(),
                    grad.stride(0), grad.stride(1), plan.ws_ptr, plan.ws_bytes, plan.stream)
            if rc:
                _lib.check(rc, "b200_ctc_loss_fwd_bwd")
            ctx.save_for_backward(grad)
            ctx.labels, ctx.ilen_ptr = plan.keep, plan.ilen_ptr
        else:
            with _on(dev):
                rc = lib.b200_ctc_loss_forward(
                    *plan.common_args(), nll_ptr, loss_ptr, s, plan.ws_ptr, plan.ws_bytes, plan.stream)
            _lib.check(rc, "b200_ctc_loss_forward")
        if strict:
            _check_strict(plan)
        # the workspace goes back to the allocator here: reuse is ordered on this stream
        return loss if reduced else nll

    @staticmethod
    def backward(ctx, gout):
        grad, = ctx.saved_tensors
        T, B, V = grad.shape
        gw = gout
        if gw.dtype != torch.float32:
            gw = gw.to(dtype=torch.float32)
        if gw.dim() != 0:
            gw = gw.reshape(-1)
            if gw.numel() > 1 and gw.stride(0) == 0:  # the expanded scalar that `nll.sum()` sends back
                gw = gw[:1]
            gw = gw.contiguous()
        dev = grad.device
        with _on(dev):
            rc = _lib.load().b200_ctc_grad_rescale(
                grad.data_ptr(), _lib.dtype_code(grad.dtype), T, B, V, grad.stride(0), grad.stride(1),
                gw.data_ptr(), gw.numel(), ctx.hint, ctx.ilen_ptr, _stream(dev))
        if rc:
            _lib.check(rc, "b200_ctc_grad_rescale")
        return grad, None, None, None, None, None, None, None, None, None


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


def ctc_nll(logits, targets, tgt_offsets, ilen, olen, lmax, blank=0, grad_scale_hint=1.0, strict=False,
            reduce_scale=None):
    """Per-utterance negative log-likelihood (B,) fp32 from pre-softmax logits (T,B,V).

    targets: 1-D label buffer; utterance b's labels start at tgt_offsets[b] and there are
    olen[b] of them (so both the concatenated form of ``torch.nn.CTCLoss`` and a padded
    (B, Lmax) matrix are accepted without copying).  lmax: host-side bound on olen.
    grad_scale_hint: the upstream gradient d loss / d nll_b the caller expects (see the module
    docstring); any other upstream gradient is still honoured, at the price of one more pass over
    the gradient.  strict: raise ValueError for invalid targets (synchronises; without it such
    utterances get a NaN loss and NaN gradient rows).  tgt_offsets=None: targets are concatenated.
    reduce_scale=s: return the 0-dim loss s * sum_b nll_b instead (grad_scale_hint is then the upstream
    gradient expected for THAT loss).
    """
    return _CTCLossFunction.apply(logits, targets, tgt_offsets, ilen, olen, lmax, blank, grad_scale_hint, strict,
                                  reduce_scale)


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
            nll.data_ptr(), loss_sum.data_ptr(), 1.0, grad.data_ptr(), grad.stride(0), grad.stride(1),
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
        rc = lib.b200_ctc_loss_forward(*plan.common_args(), nll.data_ptr(), None, 1.0, plan.ws_ptr, plan.ws_bytes,
                                       _stream(dev))
        _lib.check(rc, "b200_ctc_loss_forward")
        rc = lib.b200_ctc_loss_backward(*plan.common_args(), None if gw is None else gw.data_ptr(), float(grad_scale),
                                        grad.data_ptr(), grad.stride(0), grad.stride(1), plan.ws_ptr, plan.ws_bytes,
                                        _stream(dev))
        _lib.check(rc, "b200_ctc_loss_backward")
    return nll, grad
