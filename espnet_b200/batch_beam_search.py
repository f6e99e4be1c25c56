"""Batched beam search around the B200 prefix scorer ("the batched search glue").

Drop-in for ``espnet/nets/batch_beam_search.py:31-348`` (BatchBeamSearch) over the same scorer
interfaces (``espnet/nets/scorer_interface.py``): full scorers expose ``batch_score`` /
``select_state``, partial scorers ``batch_score_partial`` / ``select_state(state, i, new_id)``.
Same constructor, same ``forward(x, maxlenratio, minlenratio)``, same n-best list.

What is different from the reference is where the hypotheses live.  The reference rebuilds a
Python ``Hypothesis`` per survivor every step (``batch_beam_search.py:249-284``: per hypothesis one
tensor index for the score, one ``torch.cat`` for the tokens, one dict per scorer) and converts
lists to batches and back (``batchfy`` / ``unbatchfy``).  Here the beam is a handful of device
tensors (tokens, total score, one score vector per scorer) gathered with one indexed read per
tensor, the surviving (parent, token) pairs stay on the device for scorers that implement
``batch_select_state`` (the CTC prefix scorer: one gather kernel for the whole beam), and the host
sees ONE small device->host copy per step (the top-k ids and scores: the loop's end conditions
are host decisions, ``beam_search.py:364-377``).
"""
import logging
import math
from typing import Any
from typing import Dict
from typing import List
from typing import NamedTuple
from typing import Union

import torch


class Hypothesis(NamedTuple):
    """Hypothesis data type (espnet/nets/beam_search.py:19-33)."""

    yseq: torch.Tensor
    score: Union[float, torch.Tensor] = 0
    scores: Dict[str, Union[float, torch.Tensor]] = dict()
    states: Dict[str, Any] = dict()

    def asdict(self) -> dict:
        """Convert data to JSON-friendly dict."""
        return self._replace(
            yseq=self.yseq.tolist(),
            score=float(self.score),
            scores={k: float(v) for k, v in self.scores.items()},
        )._asdict()


def end_detect(ended, i, M=3, D_end=math.log(1 * math.exp(-10))):
    """End detection, Eq. (50) of Watanabe et al. 2017 (espnet/nets/e2e_asr_common.py:21-51).

    ended: list of (length, score) of the ended hypotheses; i: current output position.
    True when, for each of the last M lengths, a hypothesis of that length exists and the best
    of them is more than -D_end below the best ended hypothesis.
    """
    if len(ended) == 0:
        return False
    best = max(s for _, s in ended)
    count = 0
    for m in range(M):
        same = [s for n, s in ended if n == i - m]
        if len(same) > 0 and max(same) - best < D_end:
            count += 1
    return count == M


class _Beam(object):
    """The running hypotheses as tensors: tokens (n, len), total score (n,), per-scorer score (n,)."""

    __slots__ = ("yseq", "score", "scores", "states")

    def __init__(self, yseq, score, scores, states):
        self.yseq, self.score, self.scores, self.states = yseq, score, scores, states

    def __len__(self):
        return self.yseq.shape[0]


class BatchBeamSearchB200(torch.nn.Module):
    """Batch beam search with device-resident hypotheses."""

    def __init__(self, scorers, weights, beam_size, vocab_size, sos, eos, token_list=None, pre_beam_ratio=1.5,
                 pre_beam_score_key=None):
        """Same arguments as BeamSearch.__init__ (espnet/nets/beam_search.py:39-112)."""
        super().__init__()
        self.weights = weights
        self.scorers = dict()
        self.full_scorers = dict()
        self.part_scorers = dict()
        self.nn_dict = torch.nn.ModuleDict()  # for the recursive `self.to(device, dtype)` of recog.py
        for k, v in scorers.items():
            w = weights.get(k, 0)
            if w == 0 or v is None:
                continue
            if hasattr(v, "batch_score_partial"):
                self.part_scorers[k] = v
            elif hasattr(v, "batch_score"):
                self.full_scorers[k] = v
            else:
                raise TypeError(f"{k} ({type(v)}) implements neither batch_score nor batch_score_partial")
            self.scorers[k] = v
            if isinstance(v, torch.nn.Module):
                self.nn_dict[k] = v
        self.sos = sos
        self.eos = eos
        self.token_list = token_list
        self.pre_beam_size = int(pre_beam_ratio * beam_size)
        self.beam_size = beam_size
        self.n_vocab = vocab_size
        if (pre_beam_score_key is not None and pre_beam_score_key != "full"
                and pre_beam_score_key not in self.full_scorers):
            raise KeyError(f"{pre_beam_score_key} is not found in {self.full_scorers}")
        self.pre_beam_score_key = pre_beam_score_key
        self.do_pre_beam = (self.pre_beam_score_key is not None and self.pre_beam_size < self.n_vocab
                            and len(self.part_scorers) > 0)

    # ------------------------------------------------------------------ one step
    def init_hyp(self, x):
        """One hypothesis [sos] with every scorer's initial state (batch_beam_search.py:110-135)."""
        states = {k: [d.batch_init_state(x)] for k, d in self.scorers.items()}
        zero = torch.zeros(1, dtype=x.dtype, device=x.device)
        return _Beam(torch.full((1, 1), self.sos, dtype=torch.int64, device=x.device), zero,
                     {k: zero.clone() for k in self.scorers}, states)

    def search(self, beam, x):
        """Extend every running hypothesis by one token and keep the best `beam_size`
        (batch_beam_search.py:206-284).  Returns the new beam and its host-side (token, score) lists."""
        n, V = len(beam), self.n_vocab
        weighted = torch.zeros(n, V, dtype=x.dtype, device=x.device)
        scores, states = {}, {}
        xs = x.expand(n, *x.shape)
        for k, d in self.full_scorers.items():
            scores[k], states[k] = d.batch_score(beam.yseq, beam.states[k], xs)
            weighted += self.weights[k] * scores[k]
        part_ids = None  # no pre-beam
        if self.do_pre_beam:
            pre = weighted if self.pre_beam_score_key == "full" else scores[self.pre_beam_score_key]
            part_ids = torch.topk(pre, self.pre_beam_size, dim=-1)[1]
        # partial scorers return full-size matrices (logzero outside part_ids)
        part_scores, part_states = {}, {}
        for k, d in self.part_scorers.items():
            part_scores[k], part_states[k] = d.batch_score_partial(beam.yseq, part_ids, beam.states[k], x)
            weighted += self.weights[k] * part_scores[k].to(dtype=x.dtype, device=x.device)
        weighted += beam.score.unsqueeze(1)

        flat = weighted.view(-1)
        top_score, top = flat.topk(self.beam_size)       # hyp * V + token, best first
        prev = torch.div(top, V, rounding_mode="floor")
        tok = top - prev * V
        new_scores = {}
        for k, v in scores.items():
            new_scores[k] = beam.scores[k][prev] + v.reshape(-1)[top]
        for k, v in part_scores.items():
            new_scores[k] = beam.scores[k][prev] + v.reshape(-1)[top].to(x.dtype)
        yseq = torch.cat((beam.yseq[prev], tok.unsqueeze(1)), dim=1)
        # the one device->host copy of the step (ids are exact in float64)
        host = torch.cat((top.to(torch.float64), top_score.to(torch.float64))).tolist()
        ids = [int(v) for v in host[: self.beam_size]]
        prev_l, tok_l = [i // V for i in ids], [i % V for i in ids]
        new_states = {}
        for k, d in self.full_scorers.items():
            new_states[k] = [d.select_state(states[k], i) for i in prev_l]
        for k, d in self.part_scorers.items():
            if hasattr(d, "batch_select_state"):
                new_states[k] = d.batch_select_state(part_states[k], prev, tok)
            else:
                new_states[k] = [d.select_state(part_states[k], i, t) for i, t in zip(prev_l, tok_l)]
        return _Beam(yseq, top_score, new_scores, new_states), tok_l, host[self.beam_size:]

    def _take(self, beam, i):
        """Hypothesis i of the beam as the reference's data type (batch_beam_search.py:60-68)."""
        states = {}
        for k, v in beam.states.items():
            states[k] = v[i] if isinstance(v, (list, tuple)) else self.scorers[k].select_state(v, i)
        return Hypothesis(yseq=beam.yseq[i], score=beam.score[i], scores={k: v[i] for k, v in beam.scores.items()},
                          states=states)

    def _keep(self, beam, ids, dev):
        idx = torch.tensor(ids, dtype=torch.int64, device=dev)
        states = {}
        for k, v in beam.states.items():
            if isinstance(v, (list, tuple)):
                states[k] = [v[i] for i in ids]
            else:
                states[k] = self.scorers[k].batch_keep_state(v, idx)
        return _Beam(beam.yseq[idx], beam.score[idx], {k: v[idx] for k, v in beam.scores.items()}, states)

    # ------------------------------------------------------------------ the loop
    def forward(self, x, maxlenratio=0.0, minlenratio=0.0):
        """Perform beam search (espnet/nets/beam_search.py:336-410).

        x: encoded speech feature (T, D).  maxlenratio 0.0 ends by the end-detect rule.
        Returns the ended hypotheses, best first.
        """
        if maxlenratio == 0:
            maxlen = x.shape[0]
        else:
            maxlen = max(1, int(maxlenratio * x.size(0)))
        minlen = int(minlenratio * x.size(0))
        logging.info("decoder input length: " + str(x.shape[0]))
        logging.info("max output length: " + str(maxlen))
        logging.info("min output length: " + str(minlen))

        beam = self.init_hyp(x)
        ended: List[Hypothesis] = []
        ended_ls = []  # (length, score) of `ended`, host side, for end detection and the final sort
        for i in range(maxlen):
            beam, tok_l, score_l = self.search(beam, x)
            # post-processing of one iteration (batch_beam_search.py:286-348)
            if i == maxlen - 1:
                logging.info("adding <eos> in the last position in the loop")
                eos_col = torch.full((len(beam), 1), self.eos, dtype=torch.int64, device=beam.yseq.device)
                beam.yseq = torch.cat((beam.yseq, eos_col), dim=1)
                tok_l = [self.eos] * len(tok_l)
            length = beam.yseq.shape[1]
            remained = []
            for b, t in enumerate(tok_l):
                if t == self.eos:
                    ended.append(self._take(beam, b))
                    ended_ls.append((length, score_l[b]))
                else:
                    remained.append(b)
            if len(remained) < len(tok_l):
                beam = self._keep(beam, remained, x.device)
            if maxlenratio == 0.0 and end_detect(ended_ls, i):
                logging.info(f"end detected at {i}")
                break
            if len(beam) == 0:
                logging.info("no hypothesis. Finish decoding.")
                break

        order = sorted(range(len(ended)), key=lambda j: ended_ls[j][1], reverse=True)
        nbest = [ended[j] for j in order]
        if len(nbest) == 0:
            logging.warning("there is no N-best results, perform recognition again with smaller minlenratio.")
            return [] if minlenratio < 0.1 else self.forward(x, maxlenratio, max(0.0, minlenratio - 0.1))
        best = nbest[0]
        for k, v in best.scores.items():
            logging.info(f"{float(v):6.2f} * {self.weights[k]:3} = {float(v) * self.weights[k]:6.2f} for {k}")
        logging.info(f"total log probability: {float(best.score):.2f}")
        logging.info(f"normalized log probability: {float(best.score) / len(best.yseq):.2f}")
        logging.info(f"total number of ended hypotheses: {len(nbest)}")
        if self.token_list is not None:
            logging.info("best hypo: " + "".join([self.token_list[t] for t in best.yseq[1:-1]]) + "\n")
        return nbest
