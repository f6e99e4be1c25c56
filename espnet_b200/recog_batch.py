"""The multi-utterance joint CTC/attention search loop around the B200 prefix scorer.

Restates the CTC-relevant glue of ``Decoder.recognize_beam_batch``
(``espnet/nets/pytorch_backend/rnn/decoders.py:641-974``) -- per-utterance top-k over beam x V (:836-846), the
scorer call with the pre-beam candidates and the attention weights (:818-833), ``index_select_state`` (:940-944),
ended-hypothesis bookkeeping (:896-928) and end detection (:931-937) -- around ``CTCPrefixScoreB200``.  The
attention decoder itself (embedding, attention, RNN, output layer: :779-812, :853-893) is NOT part of the path: it is
a callback object, see ``DecoderStep``.

What differs from the reference is where the beam lives: ``yseq`` is one (batch*beam, maxlen+1) int64 device tensor
gathered with one indexed read per step (the reference keeps batch*beam Python lists and rebuilds them with two list
comprehensions per step), the scorer is given that tensor directly, and the host sees ONE device->host copy per step
(the rows the end conditions need: eos scores, the beam threshold, the survivors' ids), because "which hypotheses have
ended" and "may the search stop" are host decisions in the reference as well.
"""
import math

import numpy as np
import torch

from . import _lib
from .ctc_prefix_score import CTCPrefixScoreB200

LOGZERO = -10000000000.0          # decoders.py:84
END_DETECT_D = math.log(1 * math.exp(-10))   # e2e_asr_common.py:21
CTC_SCORING_RATIO = 1.5           # decoders.py:25


class DecoderStep(object):
    """What the loop needs from an attention decoder (decoders.py:779-812, :853-893).

    start(n_bb) -> state                    initial decoder state for batch*beam hypotheses
    step(vy, state) -> (logp, att_w, state) vy (n_bb,) int64 last tokens; logp (n_bb, odim) log-softmax of the output
                                            layer; att_w (n_bb, T) attention weights of this step or None
    select(state, vidx) -> state            survivors' states, vidx (n_bb,) int64 indices into the old n_bb
    """

    def start(self, n_bb):
        raise NotImplementedError

    def step(self, vy, state):
        raise NotImplementedError

    def select(self, state, vidx):
        raise NotImplementedError


def beam_prune(logp, ctc_scores, vscores, eos, att_weight, ctc_weight, first_position, packed=False):
    """decoders.py:812-852 in ONE kernel (``b200_beam_prune``, csrc/beam_prune.cu): the joint local scores, the eos
    scores, the accumulated scores and the per-utterance top-k over beam x V.

    logp (batch*beam, V) fp32 log-softmax of the attention decoder; ctc_scores the same shape or None;
    vscores (batch, beam).  Returns (best_scores (batch, beam), best_ids (batch, beam) in hyp * V + token space,
    eos_vscores (batch, beam), tokens (batch*beam,), parent_rows (batch*beam,)); packed=True: the two buffers they
    are views of, (2, batch, beam) float32 [best_scores, eos_vscores] and (3, batch, beam) int64 [ids, tokens, rows]."""
    batch, beam = vscores.shape
    V = logp.size(-1)
    dev = logp.device
    logp = logp.contiguous()
    if ctc_scores is not None:
        ctc_scores = ctc_scores.contiguous()
    vscores = vscores.contiguous()
    sc = torch.empty((2, batch, beam), dtype=torch.float32, device=dev)  # best_scores, eos_vscores
    ids = torch.empty((3, batch, beam), dtype=torch.int64, device=dev)   # best_ids, tokens, parent rows
    nb = batch * beam * 8
    sp = sc.data_ptr()
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    args = (logp.data_ptr(), None if ctc_scores is None else ctc_scores.data_ptr(), vscores.data_ptr(), batch, beam, V, beam,
            float(att_weight), float(ctc_weight), int(eos), 1 if first_position else 0, sp,
            ids.data_ptr(), sp + 4 * batch * beam, ids.data_ptr() + nb, ids.data_ptr() + 2 * nb,
            torch._C._cuda_getCurrentRawStream(idx))
    fn = _lib.load().b200_beam_prune
    if torch.cuda.current_device() == idx:   # (the device guard only where the current device is another)
        rc = fn(*args)
    else:
        with torch.cuda.device(dev):
            rc = fn(*args)
    _lib.check(rc, "b200_beam_prune")
    if packed:
        return sc, ids
    return sc[0], ids[0], sc[1], ids[1].view(-1), ids[2].view(-1)


def pre_beam_candidates(logp, k, att_weight, blank=0):
    """decoders.py:816-822 in ONE kernel (``b200_pre_beam_candidates``): the k best tokens of every hypothesis by
    ``att_weight * logp`` with the blank masked; logp (n, V) fp32 -> ids (n, k) int64, best first."""
    n, V = logp.shape
    logp = logp.contiguous()
    ids = torch.empty((n, k), dtype=torch.int64, device=logp.device)
    dev = logp.device
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    args = (logp.data_ptr(), n, V, int(k), float(att_weight), int(blank), ids.data_ptr(), torch._C._cuda_getCurrentRawStream(idx))
    fn = _lib.load().b200_pre_beam_candidates
    if torch.cuda.current_device() == idx:
        rc = fn(*args)
    else:
        with torch.cuda.device(dev):
            rc = fn(*args)
    _lib.check(rc, "b200_pre_beam_candidates")
    return ids


def _add_ended(ended, summary, yseq, v):
    """An ended hypothesis, and what end detection needs of it: the best score overall and per length."""
    score = np.float32(v)
    ended.append({"yseq": yseq, "vscore": v, "score": score})
    sc, n = float(score), len(yseq)
    if summary[0] is None or sc > summary[0]:
        summary[0] = sc
    if n not in summary[1] or sc > summary[1][n]:
        summary[1][n] = sc


def _end_detect(summary, i, M=3, D_end=END_DETECT_D):
    """``end_detect`` (e2e_asr_common.py:21-51) from the running summary instead of a pass over all ended hypotheses."""
    best, by_len = summary
    if best is None:
        return False
    return all((i - m) in by_len and by_len[i - m] - best < D_end for m in range(M))


def recognize_beam_batch(decoder, lpz, hlens, beam, sos, eos, ctc_weight=0.3, maxlenratio=0.0, minlenratio=0.0,
                         penalty=0.0, nbest=1, ctc_margin=0, normalize_score=True, pre_beam=None, scorer=None,
                         stats=None, fused=True):
    """Batched joint CTC/attention beam search (decoders.py:641-974 with one encoder, no LM).

    decoder: a DecoderStep; lpz (batch, T, odim) CTC log-probs on the device (or None: attention only);
    hlens: per-utterance encoder lengths; pre_beam: None = the reference's rule (decoders.py:759-765: 1.5 x beam
    candidates from the attention scores on CPU inputs, all tokens on CUDA inputs -- here the inputs are always on
    CUDA, so None means all tokens), or the number of candidates.
    stats: a dict that receives {"steps": output positions the loop ran}.  fused=False: the score accumulation and the
    top-k as eager torch operations (the reference's lines) instead of ``b200_beam_prune``.
    Returns the n-best list per utterance: [{"yseq": [...], "score": float}, ...] sorted by score.
    """
    hl = [int(v) for v in (hlens.tolist() if torch.is_tensor(hlens) else hlens)]
    batch = len(hl)
    n_bb = batch * beam
    att_weight = 1.0 - ctc_weight
    max_hlen = max(hl)
    maxlen = max_hlen if maxlenratio == 0 else max(1, int(maxlenratio * max_hlen))   # :693-697
    minlen = int(minlenratio * max_hlen)
    dev = lpz.device if lpz is not None else torch.device("cuda")
    odim = None
    own_scorer = False
    if lpz is not None:
        odim = lpz.size(-1)
        if scorer is None:
            scorer = CTCPrefixScoreB200(lpz, torch.as_tensor(hl), 0, eos, margin=ctc_margin)
            own_scorer = True
    own_scorer = own_scorer or isinstance(scorer, CTCPrefixScoreB200)
    scoring_num = 0
    if lpz is not None and pre_beam:
        scoring_num = min(int(pre_beam), odim)
    pad_b = (torch.arange(batch, device=dev) * beam).view(-1, 1)
    yseq = torch.full((n_bb, maxlen + 2), eos, dtype=torch.int64, device=dev)
    yseq[:, 0] = sos
    yseq_other = torch.empty_like(yseq)
    vscores = torch.zeros(batch, beam, device=dev)
    ended = [[] for _ in range(batch)]
    summary = [[None, {}] for _ in range(batch)]   # per utterance: best ended score, best ended score per length
    live_len = np.asarray(hl).reshape(-1, 1)
    min_hl = min(hl)
    stop = [False] * batch
    state = decoder.start(n_bb)
    ctc_state = None
    # Without a pre-beam and without an attention window the CTC scores of a position depend on the prefixes alone:
    # they are issued BEFORE the decoder step, so that the score kernel runs while the host issues the decoder's kernels
    early = own_scorer and scoring_num == 0 and not scorer.margin
    tok = None
    for i in range(maxlen):
        vy = yseq.select(1, i)
        ctc_scores = None
        if early:
            ctc_scores, ctc_state = scorer(yseq.narrow(1, 0, i + 1), ctc_state, None, None, last_ids=tok)
        logp, att_w, state = decoder.step(vy, state)
        if odim is None:
            odim = logp.size(-1)
        if scorer is not None and not early:
            part_ids = None
            if scoring_num > 0:   # pre-beam on the attention scores with the blank masked (:816-822)
                if fused and logp.dtype == torch.float32 and scoring_num <= 128 and odim <= 51200:
                    part_ids = pre_beam_candidates(logp, scoring_num, att_weight, 0)
                else:
                    local = att_weight * logp
                    local[:, 0] = LOGZERO
                    part_ids = torch.topk(local, scoring_num, dim=-1)[1]
            if own_scorer:   # (the last labels are at hand: the previous position's tokens)
                ctc_scores, ctc_state = scorer(yseq.narrow(1, 0, i + 1), ctc_state, part_ids, att_w, last_ids=tok)
            else:
                ctc_scores, ctc_state = scorer(yseq.narrow(1, 0, i + 1), ctc_state, part_ids, att_w)
        if fused and logp.dtype == torch.float32 and beam <= 128 and beam * odim < 2 ** 31:
            # joint local scores, eos scores, accumulated scores and the per-utterance top-k in one kernel
            sc_pack, id_pack = beam_prune(logp, ctc_scores, vscores, eos, att_weight, ctc_weight, i == 0, packed=True)
            best_scores, eos_v = sc_pack[0], sc_pack[1]
            best_ids, tok, vidx = id_pack[0], id_pack[1].view(-1), id_pack[2].view(-1)
        else:
            sc_pack = None
            local = att_weight * logp
            if scorer is not None:
                local[:, 0] = LOGZERO  # avoid choosing blank (:816)
                local = local + ctc_weight * ctc_scores
            local = local.view(batch, beam, odim)
            if i == 0:
                local[:, 1:, :] = LOGZERO
            # accumulate scores, global pruning (:840-846)
            eos_v = local[:, :, eos] + vscores
            acc = vscores.view(batch, beam, 1) + local
            acc[:, :, eos] = LOGZERO
            best_scores, best_ids = torch.topk(acc.view(batch, -1), beam, 1)
            tok = torch.fmod(best_ids, odim).view(-1)
            vidx = (torch.div(best_ids, odim, rounding_mode="floor") + pad_b).view(-1)
        if scorer is not None:   # (before the host waits for this position's scores: the kernel runs under the wait)
            ctc_state = scorer.index_select_state(ctc_state, best_ids)
        y_prev = yseq
        yseq = torch.index_select(y_prev, 0, vidx, out=yseq_other)   # (two buffers: the survivors' rows, no allocation)
        yseq_other = y_prev
        yseq.select(1, i + 1).copy_(tok)
        vscores = best_scores
        state = decoder.select(state, vidx)
        # pick ended hypotheses (:896-928): ONE device->host copy with everything the host decides on; the comparison
        # runs over the whole (batch, beam) array at once and Python only visits the hypotheses that did end
        if i >= minlen:
            if sc_pack is not None:
                packed = sc_pack.cpu().numpy()
                best_h, eos_h = packed[0], packed[1]
            else:
                packed = torch.cat([eos_v, best_scores], dim=1).cpu().numpy()
                eos_h, best_h = packed[:, :beam], packed[:, beam:]
            pen = (i + 1) * penalty
            hit = eos_h > best_h[:, -1:]
            if i + 1 > min_hl:
                hit &= live_len >= i + 1          # len(yk) <= hlens (:908-911); y_prev rows hold i + 1 tokens
            visit = hit if i < maxlen - 1 else np.ones_like(hit)
            if any(stop):
                visit = visit & ~np.asarray(stop)[:, None]
            if visit.any():
                need_prev = need_cur = None
                for s, j in zip(*np.nonzero(visit)):
                    k = s * beam + j
                    if eos_h[s, j] > best_h[s, -1]:
                        if hit[s, j]:
                            if need_prev is None:
                                need_prev = y_prev[:, :i + 1].cpu().numpy()
                            v = float(eos_h[s, j]) + pen
                            if v:                  # (`if _vscore:` in the reference: a score of exactly 0 is dropped)
                                _add_ended(ended[s], summary[s], need_prev[k].tolist() + [eos], v)
                    else:                          # i == maxlen - 1: the survivors end here
                        if need_cur is None:
                            need_cur = yseq[:, :i + 2].cpu().numpy()
                        v = float(best_h[s, j]) + pen
                        if v:
                            _add_ended(ended[s], summary[s], need_cur[k].tolist() + [eos], v)
        # end detection (:931-937)
        stop = [stop[s] or _end_detect(summary[s], i) for s in range(batch)]
        if all(stop):
            break
    if stats is not None:
        stats["steps"] = i + 1
    out = []
    for s in range(batch):
        hyps = ended[s] if ended[s] else [{"yseq": [sos, eos], "score": np.float32(-np.inf)}]
        if normalize_score:
            for h in hyps:
                h["score"] = h["score"] / len(h["yseq"])
        hyps = sorted(hyps, key=lambda h: h["score"], reverse=True)[:min(len(hyps), nbest)]
        out.append(hyps)
    return out
