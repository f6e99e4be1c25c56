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
import numpy as np
import torch

from . import _lib
from .batch_beam_search import end_detect
from .ctc_prefix_score import CTCPrefixScoreB200

LOGZERO = -10000000000.0          # decoders.py:84
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


def beam_prune(logp, ctc_scores, vscores, eos, att_weight, ctc_weight, first_position):
    """decoders.py:812-852 in ONE kernel (``b200_beam_prune``, csrc/beam_prune.cu): the joint local scores, the eos
    scores, the accumulated scores and the per-utterance top-k over beam x V.

    logp (batch*beam, V) fp32 log-softmax of the attention decoder; ctc_scores the same shape or None;
    vscores (batch, beam).  Returns (best_scores (batch, beam), best_ids (batch, beam) in hyp * V + token space,
    eos_vscores (batch, beam), tokens (batch*beam,), parent_rows (batch*beam,))."""
    batch, beam = vscores.shape
    V = logp.size(-1)
    dev = logp.device
    logp = logp.contiguous()
    if ctc_scores is not None:
        ctc_scores = ctc_scores.contiguous()
    vscores = vscores.contiguous()
    best_scores = torch.empty((batch, beam), dtype=torch.float32, device=dev)
    eos_v = torch.empty((batch, beam), dtype=torch.float32, device=dev)
    ids = torch.empty((3, batch, beam), dtype=torch.int64, device=dev)   # best_ids, tokens, parent rows
    nb = batch * beam * 8
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    args = (logp.data_ptr(), None if ctc_scores is None else ctc_scores.data_ptr(), vscores.data_ptr(), batch, beam, V, beam,
            float(att_weight), float(ctc_weight), int(eos), 1 if first_position else 0, best_scores.data_ptr(),
            ids.data_ptr(), eos_v.data_ptr(), ids.data_ptr() + nb, ids.data_ptr() + 2 * nb,
            torch._C._cuda_getCurrentRawStream(idx))
    fn = _lib.load().b200_beam_prune
    if torch.cuda.current_device() == idx:   # (the device guard only where the current device is another)
        rc = fn(*args)
    else:
        with torch.cuda.device(dev):
            rc = fn(*args)
    _lib.check(rc, "b200_beam_prune")
    return best_scores, ids[0], eos_v, ids[1].view(-1), ids[2].view(-1)


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
    if lpz is not None:
        odim = lpz.size(-1)
        if scorer is None:
            scorer = CTCPrefixScoreB200(lpz, torch.as_tensor(hl), 0, eos, margin=ctc_margin)
    scoring_num = 0
    if lpz is not None and pre_beam:
        scoring_num = min(int(pre_beam), odim)
    pad_b = (torch.arange(batch, device=dev) * beam).view(-1, 1)
    yseq = torch.full((n_bb, maxlen + 2), eos, dtype=torch.int64, device=dev)
    yseq[:, 0] = sos
    vscores = torch.zeros(batch, beam, device=dev)
    ended = [[] for _ in range(batch)]
    stop = [False] * batch
    state = decoder.start(n_bb)
    ctc_state = None
    for i in range(maxlen):
        vy = yseq[:, i]
        logp, att_w, state = decoder.step(vy, state)
        if odim is None:
            odim = logp.size(-1)
        ctc_scores = None
        if scorer is not None:
            part_ids = None
            if scoring_num > 0:   # pre-beam on the attention scores with the blank masked (:816-822)
                local = att_weight * logp
                local[:, 0] = LOGZERO
                part_ids = torch.topk(local, scoring_num, dim=-1)[1]
            ctc_scores, ctc_state = scorer(yseq[:, :i + 1], ctc_state, part_ids, att_w)
        if fused and logp.dtype == torch.float32 and beam <= 128 and beam * odim < 2 ** 31:
            # joint local scores, eos scores, accumulated scores and the per-utterance top-k in one kernel
            best_scores, best_ids, eos_v, tok, vidx = beam_prune(logp, ctc_scores, vscores, eos, att_weight, ctc_weight, i == 0)
        else:
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
        y_prev = yseq
        yseq = yseq.index_select(0, vidx)
        yseq[:, i + 1] = tok
        vscores = best_scores
        state = decoder.select(state, vidx)
        # pick ended hypotheses (:896-928): ONE device->host copy with everything the host decides on
        if i >= minlen:
            packed = torch.cat([eos_v, best_scores], dim=1).cpu().numpy()
            eos_h, best_h = packed[:, :beam], packed[:, beam:]
            pen = (i + 1) * penalty
            need_prev = need_cur = None
            for s in range(batch):
                if stop[s]:
                    continue
                thr = best_h[s, -1]
                for j in range(beam):
                    k = s * beam + j
                    if eos_h[s, j] > thr:
                        if i + 1 <= hl[s]:   # len(yk) <= hlens (:908-911); y_prev rows hold i + 1 tokens
                            if need_prev is None:
                                need_prev = y_prev[:, :i + 1].cpu().numpy()
                            v = float(eos_h[s, j]) + pen
                            if v:                      # (`if _vscore:` in the reference: a score of exactly 0 is dropped)
                                ended[s].append({"yseq": need_prev[k].tolist() + [eos], "vscore": v, "score": np.float32(v)})
                    elif i == maxlen - 1:
                        if need_cur is None:
                            need_cur = yseq[:, :i + 2].cpu().numpy()
                        v = float(best_h[s, j]) + pen
                        if v:
                            ended[s].append({"yseq": need_cur[k].tolist() + [eos], "vscore": v, "score": np.float32(v)})
        # end detection (:931-937)
        stop = [stop[s] or end_detect([(len(h["yseq"]), float(h["score"])) for h in ended[s]], i) for s in range(batch)]
        if all(stop):
            break
        if scorer is not None:
            ctc_state = scorer.index_select_state(ctc_state, best_ids)
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
