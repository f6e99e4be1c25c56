"""Drop-in CTC modules with ``ctc_type="b200"``.

``CTC`` mirrors ``espnet2/asr/ctc.py:6-111`` and ``CTCEspnet1`` mirrors
``espnet/nets/pytorch_backend/ctc.py:12-151`` (same constructor arguments, methods, return
values and error behaviour); the only addition is the new ``ctc_type`` value ``"b200"``,
which sends ``loss_fn`` through the fused sm_100a kernels instead of
``log_softmax`` + ``torch.nn.CTCLoss``.  ``"builtin"`` is kept exactly as the reference has it
(it is the reference's own branch and what the parity tests compare against); ``"warpctc"``
needs the ``warpctc_pytorch`` wheel as in the reference.
"""
import logging

import numpy as np
import torch
import torch.nn.functional as F

from .ctc_align import ctc_argmax, ctc_forced_align
from .ctc_loss import ctc_nll


def _b200_loss(th_pred, targets, tgt_offsets, th_ilen, th_olen, lmax, reduce, hint=1.0, strict=False):
    """(T,B,V) pre-softmax logits -> CTCLoss(sum|none)(log_softmax(.)) / B.

    The one-pass kernels write the gradient during the forward call, scaled by the upstream gradient they
    expect: `hint` (what the trainer multiplies this loss by: ctc_weight / accum_grad x the AMP loss scale;
    1.0 for a bare ``loss.backward()``).  Any other upstream gradient is honoured in backward by one more
    pass over the gradient.  The reduced loss (sum over utterances, / B) comes out of the lattice kernel
    itself."""
    B = th_pred.size(1)
    if reduce:
        return ctc_nll(th_pred, targets, tgt_offsets, th_ilen, th_olen, lmax, blank=0, grad_scale_hint=hint,
                       strict=strict, reduce_scale=1.0 / B)
    nll = ctc_nll(th_pred, targets, tgt_offsets, th_ilen, th_olen, lmax, blank=0, grad_scale_hint=hint / B,
                  strict=strict)
    return nll / B


def _host_max(lens):
    """max of a length vector without touching the device when the lengths are host values."""
    if torch.is_tensor(lens):
        if lens.numel() == 0:
            return 0
        return int(lens.numpy().max()) if not lens.is_cuda else int(lens.max())
    return int(max(lens)) if len(lens) else 0


def _viterbi_numpy(lpz, y, blank_id):
    """Best CTC alignment on the host, the reference's recursion (pytorch_backend/ctc.py:153-216) with the inner
    loop over states vectorised: float64 scores, candidates [s, s-1, s-2] with the first maximum kept, and the
    reference's wrap-around reads (state -1 is the last state)."""
    y = np.asarray(y, dtype=np.int64).reshape(-1)
    ext = np.full(2 * len(y) + 1, blank_id, dtype=np.int64)
    ext[1::2] = y
    T, S = lpz.shape[0], len(ext)
    sidx = np.arange(S)
    two_only = (ext == blank_id) | (sidx < 2) | (ext == np.roll(ext, 2))
    delta = np.full(S, -100000000000.0)
    delta[0] = lpz[0, ext[0]]
    if S > 1:
        delta[1] = lpz[0, ext[1]]
    back = np.full((T, S), -1, dtype=np.int16)
    for t in range(1, T):
        c1, c2 = np.roll(delta, 1), np.roll(delta, 2)
        best, arg = delta.copy(), sidx.copy()
        m = c1 > best
        best[m], arg[m] = c1[m], sidx[m] - 1
        m = (c2 > best) & ~two_only
        best[m], arg[m] = c2[m], sidx[m] - 2
        delta = best + lpz[t, ext].astype(np.float64)
        back[t] = arg
    st = S - 1 if delta[S - 1] >= delta[S - 2] else S - 2
    states = np.empty(T, dtype=np.int64)
    states[T - 1] = st
    for t in range(T - 2, -1, -1):
        states[t] = back[t + 1, states[t + 1]]
    return [int(ext[k]) for k in states]


class CTC(torch.nn.Module):
    """CTC module (espnet2 interface).

    Args:
        odim: dimension of outputs
        encoder_output_sizse: number of encoder projection units
        dropout_rate: dropout rate (0.0 ~ 1.0)
        ctc_type: b200, builtin or warpctc
        reduce: reduce the CTC loss into a scalar
    """

    def __init__(self, odim: int, encoder_output_sizse: int, dropout_rate: float = 0.0,
                 ctc_type: str = "b200", reduce: bool = True, grad_scale_hint: float = 1.0,
                 strict_targets: bool = False):
        super().__init__()
        # b200 only: the upstream gradient the trainer will send into this loss (see _b200_loss), and
        # whether invalid targets raise (costs a host synchronisation per call) or give NaN
        self.grad_scale_hint = grad_scale_hint
        self.strict_targets = strict_targets
        eprojs = encoder_output_sizse
        self.dropout_rate = dropout_rate
        self.ctc_lo = torch.nn.Linear(eprojs, odim)
        self.ctc_type = ctc_type

        if self.ctc_type == "b200":
            self.ctc_loss = None
        elif self.ctc_type == "builtin":
            reduction_type = "sum" if reduce else "none"
            self.ctc_loss = torch.nn.CTCLoss(reduction=reduction_type)
        elif self.ctc_type == "warpctc":
            import warpctc_pytorch as warp_ctc

            self.ctc_loss = warp_ctc.CTCLoss(size_average=True, reduce=reduce)
        else:
            raise ValueError(f'ctc_type must be "b200", "builtin" or "warpctc": {self.ctc_type}')

        self.reduce = reduce

    def loss_fn(self, th_pred, th_target, th_ilen, th_olen) -> torch.Tensor:
        """th_pred (T,B,V) logits, th_target 1-D concatenated labels (espnet2/asr/ctc.py:46)."""
        if self.ctc_type == "b200":
            # (lengths on the host, as espnet1 passes them, keep this free of device synchronisation)
            return _b200_loss(th_pred, th_target, None, th_ilen, th_olen, _host_max(th_olen), self.reduce,
                              self.grad_scale_hint, self.strict_targets)
        elif self.ctc_type == "builtin":
            th_pred = th_pred.log_softmax(2)
            loss = self.ctc_loss(th_pred, th_target, th_ilen, th_olen)
            # Batch-size average
            loss = loss / th_pred.size(1)
            return loss
        elif self.ctc_type == "warpctc":
            th_pred = th_pred.to(dtype=torch.float32)
            th_target = th_target.cpu().int()
            th_ilen = th_ilen.cpu().int()
            th_olen = th_olen.cpu().int()
            loss = self.ctc_loss(th_pred, th_target, th_ilen, th_olen)
            if self.reduce:
                loss = loss.sum()
            return loss
        else:
            raise NotImplementedError

    def forward(self, hs_pad, hlens, ys_pad, ys_lens):
        """Calculate CTC loss.

        Args:
            hs_pad: batch of padded hidden state sequences (B, Tmax, D)
            hlens: batch of lengths of hidden state sequences (B)
            ys_pad: batch of padded character id sequence tensor (B, Lmax)
            ys_lens: batch of lengths of character sequence (B)
        """
        # hs_pad: (B, L, NProj) -> ys_hat: (B, L, Nvocab) -> (L, B, Nvocab) view
        ys_hat = self.ctc_lo(F.dropout(hs_pad, p=self.dropout_rate))
        ys_hat = ys_hat.transpose(0, 1)

        if self.ctc_type == "b200":
            # The padded label matrix is consumed in place (row b starts at b * Lmax): no
            # per-utterance slicing, concatenation or host synchronisation.
            B, lmax = ys_pad.size(0), ys_pad.size(1)
            ys_flat = ys_pad.contiguous().view(-1)
            off = torch.arange(B, device=ys_pad.device, dtype=torch.int64) * lmax
            loss = _b200_loss(ys_hat, ys_flat, off, hlens, ys_lens, lmax, self.reduce, self.grad_scale_hint,
                              self.strict_targets)
        else:
            # (B, L) -> (BxL,)
            ys_true = torch.cat([ys_pad[i, :l] for i, l in enumerate(ys_lens)])
            loss = self.loss_fn(ys_hat, ys_true, hlens, ys_lens)
        return loss.to(device=hs_pad.device, dtype=hs_pad.dtype)

    def log_softmax(self, hs_pad):
        """log_softmax of frame activations: (B, Tmax, eprojs) -> (B, Tmax, odim)."""
        return F.log_softmax(self.ctc_lo(hs_pad), dim=2)

    def argmax(self, hs_pad):
        """argmax of frame activations: (B, Tmax, eprojs) -> (B, Tmax)."""
        if self.ctc_type == "b200" and hs_pad.is_cuda:
            return ctc_argmax(self.ctc_lo(hs_pad))
        return torch.argmax(self.ctc_lo(hs_pad), dim=2)


class CTCEspnet1(torch.nn.Module):
    """CTC module (espnet1 interface, espnet/nets/pytorch_backend/ctc.py:12).

    :param int odim: dimension of outputs
    :param int eprojs: number of encoder projection units
    :param float dropout_rate: dropout rate (0.0 ~ 1.0)
    :param str ctc_type: b200, builtin or warpctc
    :param bool reduce: reduce the CTC loss into a scalar
    """

    def __init__(self, odim, eprojs, dropout_rate, ctc_type="b200", reduce=True, grad_scale_hint=1.0,
                 strict_targets=False):
        super().__init__()
        self.grad_scale_hint = grad_scale_hint  # b200 only, see _b200_loss
        self.strict_targets = strict_targets
        self.dropout_rate = dropout_rate
        self.loss = None
        self.ctc_lo = torch.nn.Linear(eprojs, odim)
        self.probs = None  # for visualization
        # the reference forces "builtin" for torch >= 1.7 when asked for warpctc
        # (pytorch_backend/ctc.py:30-36); "b200" is never rewritten.
        self.ctc_type = "builtin" if ctc_type == "warpctc" else ctc_type
        if ctc_type != self.ctc_type:
            logging.warning(f"CTC was set to {self.ctc_type} due to PyTorch version.")
        if self.ctc_type == "b200":
            self.ctc_loss = None
        elif self.ctc_type == "builtin":
            reduction_type = "sum" if reduce else "none"
            self.ctc_loss = torch.nn.CTCLoss(reduction=reduction_type)
        else:
            raise ValueError('ctc_type must be "b200", "builtin" or "warpctc": {}'.format(self.ctc_type))
        self.ignore_id = -1
        self.reduce = reduce

    def loss_fn(self, th_pred, th_target, th_ilen, th_olen):
        if self.ctc_type == "b200":
            return _b200_loss(th_pred, th_target, None, th_ilen, th_olen, _host_max(th_olen), self.reduce,
                              self.grad_scale_hint, self.strict_targets)
        elif self.ctc_type == "builtin":
            th_pred = th_pred.log_softmax(2)
            with torch.backends.cudnn.flags(deterministic=True):
                loss = self.ctc_loss(th_pred, th_target, th_ilen, th_olen)
            loss = loss / th_pred.size(1)
            return loss
        else:
            raise NotImplementedError

    def forward(self, hs_pad, hlens, ys_pad):
        """CTC forward: hs_pad (B,Tmax,D), hlens (B), ys_pad (B,Lmax) padded with -1."""
        ys = [y[y != self.ignore_id] for y in ys_pad]  # parse padded ys
        self.loss = None
        hlens = torch.from_numpy(np.fromiter(hlens, dtype=np.int32))
        olens = torch.from_numpy(np.fromiter((x.size(0) for x in ys), dtype=np.int32))
        ys_hat = self.ctc_lo(F.dropout(hs_pad, p=self.dropout_rate))
        ys_true = torch.cat(ys).cpu().int()  # batch x olen
        logging.info(self.__class__.__name__ + " input lengths:  " + "".join(str(hlens).split("\n")))
        logging.info(self.__class__.__name__ + " output lengths: " + "".join(str(olens).split("\n")))
        dtype = ys_hat.dtype
        ys_hat = ys_hat.transpose(0, 1)
        if self.ctc_type == "builtin" and dtype == torch.float16:
            # torch.ctc does not support float16 (#1751); the b200 kernels read fp16 directly
            ys_hat = ys_hat.to(dtype=torch.float32)
        ys_true = ys_true.to(hs_pad.device)
        self.loss = self.loss_fn(ys_hat, ys_true, hlens, olens).to(device=hs_pad.device, dtype=dtype)
        if self.reduce:
            self.loss = self.loss.sum()
            logging.info("ctc loss:" + str(float(self.loss.detach())))
        return self.loss

    def softmax(self, hs_pad):
        """softmax of frame activations (kept for PlotCTCReport)."""
        self.probs = F.softmax(self.ctc_lo(hs_pad), dim=2)
        return self.probs

    def log_softmax(self, hs_pad):
        return F.log_softmax(self.ctc_lo(hs_pad), dim=2)

    def argmax(self, hs_pad):
        if self.ctc_type == "b200" and hs_pad.is_cuda:
            return ctc_argmax(self.ctc_lo(hs_pad))
        return torch.argmax(self.ctc_lo(hs_pad), dim=2)

    def forced_align(self, h, y, blank_id=0):
        """Best alignment of the id sequence y (L,) to the hidden states h (T, D): a list of T ids
        (espnet/nets/pytorch_backend/ctc.py:153-216; on the b200 path the Viterbi recursion runs
        in ``b200_ctc_viterbi`` instead of a numpy double loop)."""
        lpz = self.log_softmax(h)
        lpz = lpz.squeeze(0)
        if self.ctc_type != "b200" or not lpz.is_cuda:
            return _viterbi_numpy(lpz.detach().cpu().numpy(), torch.as_tensor(y).cpu().numpy(), blank_id)
        return ctc_forced_align(lpz, y, blank_id)


def ctc_for(args, odim, reduce=True):
    """espnet/nets/pytorch_backend/ctc.py:219-259 with ctc_type="b200" accepted."""
    num_encs = getattr(args, "num_encs", 1)
    if num_encs == 1:
        return CTCEspnet1(odim, args.eprojs, args.dropout_rate, ctc_type=args.ctc_type, reduce=reduce)
    elif num_encs >= 1:
        ctcs_list = torch.nn.ModuleList()
        if args.share_ctc:
            ctcs_list.append(CTCEspnet1(odim, args.eprojs, args.dropout_rate[0],
                                        ctc_type=args.ctc_type, reduce=reduce))
        else:
            for idx in range(num_encs):
                ctcs_list.append(CTCEspnet1(odim, args.eprojs, args.dropout_rate[idx],
                                            ctc_type=args.ctc_type, reduce=reduce))
        return ctcs_list
    else:
        raise ValueError("Number of encoders needs to be more than one. {}".format(num_encs))
