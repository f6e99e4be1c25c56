"""Small deterministic stand-ins for the callers on either side of the prefix scorer in beam-search tests.

Shared by tests/golden/make_golden_nbest.py (where they are mixed into the REFERENCE's scorer
interfaces and driven by the reference's BatchBeamSearch) and by the GPU tests (where the same
math is driven by espnet_b200.batch_beam_search).
"""
import torch


class LogSoftmaxCTC(torch.nn.Module):
    """What CTCPrefixScorer needs of a CTC module: log_softmax(hs_pad) (espnet2/asr/ctc.py:93-101).
    The "encoder output" of these tests already is the (T, V) logit matrix."""

    def log_softmax(self, hs_pad):
        return torch.log_softmax(hs_pad, dim=2)


class Bigram(object):
    """A full (whole-vocabulary) scorer: log p(next | last token) from a fixed table."""

    def __init__(self, table):
        self.table = torch.log_softmax(table, dim=1)

    def init_state(self, x):
        return None

    def batch_init_state(self, x):
        return None

    def select_state(self, state, i, new_id=None):
        return None if state is None else state[i]

    def final_score(self, state):
        return 0.0

    def score(self, y, state, x):
        return self.table.to(x.device)[y[-1]], None

    def batch_score(self, ys, states, xs):
        return self.table.to(xs.device)[ys[:, -1]], [None] * ys.shape[0]


def make_search_input(T, V, L, seed, twin=0.0, peak=7.0, tboost=3.0, **_):
    """Peaky (T, V) logits around a seeded label sequence (blank 0, eos V-1 never a label), and the
    bigram table of the toy decoder, which prefers the same sequence.

    twin > 0 builds a NEAR-TIE: token V-2 becomes a copy of the second label (same logit column, same decoder
    row and column), except that its logit is `twin` lower on ONE frame inside that label's segment.  Every
    hypothesis through the label then has a twin through V-2 whose score is lower by a few 1e-6; the n-best
    list holds such pairs next to each other.  Both implementations evaluate the twins with the same
    operations on inputs that differ in one number, so the order is fixed by that number and not by
    rounding luck."""
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(1, V - 1, (L,), generator=g)
    x = torch.randn(T, V, generator=g)
    seg = T // (2 * L + 1)
    for k in range(L):
        t0 = (2 * k + 1) * seg
        x[t0:t0 + seg, y[k]] += peak
        x[t0 + seg:t0 + 2 * seg, 0] += peak - 1.0
    x[: seg, 0] += peak - 1.0
    x[(2 * L + 1) * seg:, 0] += peak - 1.0
    table = torch.randn(V, V, generator=g)
    prev = V - 1                                   # sos == eos
    for k in range(L):
        table[prev, y[k]] += tboost
        prev = int(y[k])
    table[prev, V - 1] += tboost
    if twin > 0.0:
        a, tw = int(y[1]), V - 2
        assert tw not in y.tolist()
        x[:, tw] = x[:, a]
        table[:, tw] = table[:, a]
        table[tw, :] = table[a, :]
        # (no hypothesis runs through BOTH twins: such pairs would tie up to rounding, in no defined order)
        table[a, tw] = table[tw, a] = table[tw, tw] = table[a, a] = -40.0
        x[3 * seg + seg // 2, tw] -= twin
    return x, table, y


class LengthBonus(object):
    """Length bonus in beam search: score 1 for every token (espnet/nets/scorers/length_bonus.py:11-61)."""

    def __init__(self, n_vocab: int):
        self.n = n_vocab

    def init_state(self, x):
        return None

    def batch_init_state(self, x):
        return None

    def select_state(self, state, i, new_id=None):
        return None if state is None else state[i]

    def final_score(self, state):
        return 0.0

    def score(self, y, state, x):
        return torch.tensor([1.0], device=x.device, dtype=x.dtype).expand(self.n), None

    def batch_score(self, ys, states, xs):
        return torch.tensor([1.0], device=xs.device, dtype=xs.dtype).expand(ys.shape[0], self.n), None
