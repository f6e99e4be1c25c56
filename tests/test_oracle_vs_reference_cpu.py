"""The C oracle against the REFERENCE itself (oracle/_ref, staged from /root/reference) on the input
distributions of tests/_cases.KINDS: the oracle is what the GPU tests use at sizes where the reference is slow,
so it is pinned on every distribution, not only on the near-uniform one of the golden fixtures.  CPU only."""
import numpy as np
import pytest
import torch

from oracle import ctc_oracle as O
from oracle import stage_ref
from tests import _cases
from tests._walk import walk_vs_reference_th

needs_ref = pytest.mark.skipif(stage_ref.path() is None, reason="oracle/_ref is not staged")


def _close(got, ref, step=None, what="scores", mag=None, atol=1e-4):
    got, ref = np.asarray(got, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    real = np.abs(ref) < 1e9
    assert (got[~real] < -1e9).all() and (got[real] > -1e9).all()
    if real.any():
        assert (np.abs(got[real] - ref[real]) <= atol + (2.4e-7 if what == "scores" else 7.2e-7) * np.abs(ref[real])).all()


class _OracleScorer(object):
    def __init__(self, lp, xlens, eos, margin):
        assert margin == 0
        self.x = lp.numpy().copy()
        self.xlens, self.eos = xlens.numpy(), eos
        self.r0 = O.prefix_init(self.x, self.xlens)

    def step(self, y, state, ids, att):
        n = len(y)
        H = n // self.x.shape[0]
        out_len = len(y[0]) - 1
        last = np.array([yi[-1] for yi in y], dtype=np.int64)
        r_prev, s_prev = (np.repeat(self.r0, H, axis=2), None) if state is None else state
        scores, psi, _ = O.prefix_score(self.x, self.xlens, H, 0, self.eos, r_prev, s_prev, last, out_len,
                                        scoring_ids=None if ids is None else ids.numpy())
        return scores, dict(H=H, r_prev=r_prev, psi=psi, last=last, out_len=out_len)

    def select(self, st, best):
        B, H = best.shape
        V = self.x.shape[2]
        parent = (best // V + torch.arange(B).view(-1, 1) * H).view(-1).numpy()
        token = (best % V).view(-1).numpy()
        r, s = O.prefix_select(self.x, st["H"], 0, st["r_prev"], st["psi"], st["last"], st["out_len"], parent, token)
        return (r, s), r, s


@needs_ref
@pytest.mark.parametrize("kind", _cases.KINDS)
@pytest.mark.parametrize("snum", [0, 5])
def test_oracle_prefix_walk_matches_reference_th(kind, snum):
    c = _cases.make_prefix_case(2, 40, 30, 3, seed=500 + len(kind), nsteps=5, kind=kind, L=6)
    walk_vs_reference_th(c, 5, _OracleScorer, snum=snum, seed=3, close=_close)


@needs_ref
@pytest.mark.parametrize("kind", _cases.KINDS)
def test_oracle_loss_matches_reference_module(kind):
    """espnet2 CTC.loss_fn (builtin) in float64 vs the oracle, loss and gradient."""
    stage_ref.activate()
    from espnet2.asr.ctc import CTC

    c = _cases.make_loss_case_kind(kind, 3, 50, 24, 8, seed=510)
    x, ilen, olen, ys = c["x"], c["ilen"], c["olen"], c["ys"]
    B = x.shape[0]
    ctc = CTC(x.shape[2], 4, ctc_type="builtin", reduce=False)
    xt = x.double().transpose(0, 1).detach().requires_grad_()
    tg = torch.cat([ys[i, : olen[i]] for i in range(B)])
    loss = ctc.loss_fn(xt, tg, ilen, olen)       # nll / B
    loss.sum().backward()
    onll, ograd = O.ctc_loss(np.ascontiguousarray(x.numpy().transpose(1, 0, 2)), ys.numpy().reshape(-1),
                             np.arange(B) * ys.shape[1], ilen.numpy(), olen.numpy(), grad_scale=1.0 / B)
    np.testing.assert_allclose(onll / B, loss.detach().numpy(), rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(ograd, xt.grad.numpy(), atol=1e-9 / B)
