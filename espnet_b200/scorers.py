"""CTCPrefixScorer with the B200 backend (drop-in for espnet/nets/scorers/ctc.py:11-127).

Implements the reference's ``BatchPartialScorerInterface`` methods (scorer_interface.py:
164-188) -- ``init_state``, ``select_state``, ``score_partial``, ``batch_init_state``,
``batch_score_partial`` -- on top of :class:`CTCPrefixScoreB200`.  The per-hypothesis state
that travels through ``BeamSearch`` / ``BatchBeamSearch`` is a small token
``(step_state, hypothesis index, new label)``; the forward variables of all surviving
hypotheses are materialised together, by one kernel launch, when the next scoring step needs
them (the reference slices them one hypothesis at a time out of a (T,2,n_hyps,snum) tensor).
"""
import torch

from .ctc_prefix_score import CTCPrefixScoreB200

try:  # inside an ESPnet installation the scorer IS a BatchPartialScorerInterface: BeamSearch.__init__ asserts
    # isinstance(v, ScorerInterface) and sorts scorers by isinstance(v, PartialScorerInterface)
    # (espnet/nets/beam_search.py:80-88); asr_inference.py:115 / recog.py:105 test BatchScorerInterface
    from espnet.nets.scorer_interface import BatchPartialScorerInterface as _ScorerBase
except ImportError:  # stand-alone use (this repository's own BatchBeamSearchB200 duck-types its scorers)
    _ScorerBase = object


class _Survivor(object):
    """A hypothesis that survived pruning: (state of the step, parent index, appended label)."""

    __slots__ = ("step_state", "parent", "token")

    def __init__(self, step_state, parent, token):
        self.step_state = step_state
        self.parent = parent  # int, or a 0-dim device tensor (read only if the hypothesis is scored again)
        self.token = token


class _BeamSurvivors(object):
    """All survivors of one step, ids still on the device: (state of the step, parent (n,), token (n,))."""

    __slots__ = ("step_state", "parent", "token")

    def __init__(self, step_state, parent, token):
        self.step_state, self.parent, self.token = step_state, parent, token

    def __len__(self):
        return self.parent.numel()


class CTCPrefixScorer(_ScorerBase):
    """Decoder interface wrapper for CTCPrefixScoreB200."""

    def __init__(self, ctc: torch.nn.Module, eos: int):
        """Initialize class.

        Args:
            ctc (torch.nn.Module): The CTC implementation (anything with ``log_softmax``), e.g.
                :class:`espnet_b200.ctc.CTC`
            eos (int): The end-of-sequence id.
        """
        self.ctc = ctc
        self.eos = eos
        self.impl = None

    # ---- shared helpers
    def _build(self, x):
        logp = self.ctc.log_softmax(x.unsqueeze(0)).detach()  # assuming batch_size = 1
        if not logp.is_cuda and torch.cuda.is_available():
            logp = logp.cuda()   # encoder output left on the host by the caller: the scorer itself is CUDA only
        xlen = torch.tensor([logp.size(1)])
        self.impl = CTCPrefixScoreB200(logp, xlen, 0, self.eos)

    def _resolve(self, states):
        """list of per-hypothesis states -> batch state (r_prev, s_prev, f_min, f_max) or None."""
        if isinstance(states, _BeamSurvivors):
            return self.impl.select(states.step_state, states.parent, states.token)
        if states[0] is None:
            return None
        step_state = states[0].step_state
        dev = self.impl.device
        parent = torch.tensor([int(s.parent) for s in states], dtype=torch.int64, device=dev)
        token = torch.tensor([int(s.token) for s in states], dtype=torch.int64, device=dev)
        return self.impl.select(step_state, parent, token)

    # ---- BatchPartialScorerInterface
    def batch_init_state(self, x: torch.Tensor):
        """Get an initial state for decoding (scorers/ctc.py:88-100)."""
        self._build(x)
        return None

    def batch_score_partial(self, y, ids, state, x):
        """Score new tokens for a batch of hypotheses of one utterance (scorers/ctc.py:102-127).

        Args:
            y (torch.Tensor): (n_hyps, len) prefix tokens
            ids (torch.Tensor): (n_hyps, snum) next tokens to score, or None for all
            state: list of per-hypothesis states (from select_state), [None, ...], or the survivors
                object of ``batch_select_state``
            x (torch.Tensor): 2D encoder feature that generates ys
        Returns:
            (scores (n_hyps, n_vocab), batch state for select_state)
        """
        batch_state = self._resolve(state)
        scores, new_state = self.impl(y, batch_state, ids)
        if x is not None and scores.device != x.device:
            # a host-side search (the reference's BatchBeamSearch keeps its hypotheses in CPU tensors) around
            # the CUDA scorer: only the (n_hyps, n_vocab) score matrix crosses, the states stay on the device
            scores = scores.to(x.device)
        return scores, new_state

    def batch_select_state(self, state, prev_hyp_ids, new_token_ids):
        """All survivors of a step at once, ids as device tensors (no host round trip): what
        ``select_state(state, i, new_id)`` does per hypothesis in the reference (scorers/ctc.py:41-64).
        The result is accepted by ``batch_score_partial`` in place of the list of per-hypothesis states."""
        return _BeamSurvivors(state, prev_hyp_ids, new_token_ids)

    def batch_keep_state(self, survivors, idx):
        """Subset (device index tensor) of what ``batch_select_state`` returned."""
        return _BeamSurvivors(survivors.step_state, survivors.parent[idx], survivors.token[idx])

    def select_state(self, state, i, new_id=None):
        """Select state with relative ids in the main beam search (scorers/ctc.py:41-64)."""
        if isinstance(state, _BeamSurvivors):
            return _Survivor(state.step_state, state.parent[i], state.token[i])
        if type(state) == tuple:
            if len(state) == 2:  # single-hypothesis path: (scores over ids, (step state, ids))
                sc, (step_state, ids) = state
                return sc[i], _Survivor(step_state, 0, ids[i])
            return _Survivor(state, i, new_id)  # batch path (new_id > 0)
        return None if state is None else state[i]

    # ---- PartialScorerInterface (non-batch BeamSearch)
    def init_state(self, x: torch.Tensor):
        """Get an initial state for decoding (scorers/ctc.py:27-39)."""
        self._build(x)
        return 0.0, None

    def score_partial(self, y, ids, state, x):
        """Score new tokens for one hypothesis (scorers/ctc.py:66-86).

        Returns (scores (len(ids),), state) where the scores are log_psi(ids) - log_psi(prefix).
        """
        prev_score, st = state
        batch_state = self._resolve([st])
        _, new_state = self.impl([[int(t) for t in y]], batch_state, ids.view(1, -1))
        presub_score = new_state[1][0, ids.to(new_state[1].device)]
        tscore = (presub_score - prev_score).to(device=x.device, dtype=x.dtype)
        return tscore, (presub_score, (new_state, [int(t) for t in ids]))
