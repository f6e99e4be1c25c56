"""The multi-utterance search loop (espnet_b200.recog_batch.recognize_beam_batch around CTCPrefixScoreB200) against
the reference's Decoder.recognize_beam_batch (espnet/nets/pytorch_backend/rnn/decoders.py:631-974).

Both sides run the SAME attention decoder (the reference's modules with seeded weights, tests/_recog.py); what differs
is the glue and the CTC prefix scorer.  Bar: the n-best token sequences are identical; scores (length-normalised sums
of ~20-30 float32 terms) agree to 5e-4."""
import json
import os

import pytest
import torch

from tests import _recog
from tests._util import ROOT
from tests.golden.make_golden_recog import CASES

pytestmark = pytest.mark.gpu
ATOL = 5e-4


def _staged():
    from oracle import stage_ref

    if not stage_ref.activate():
        pytest.skip("oracle/_ref is not staged (python oracle/stage_ref.py in the build container)")


def _ours(name, pre_beam):
    from espnet_b200.recog_batch import recognize_beam_batch

    ckw, dseed, beam, rkw = CASES[name]
    case = _recog.make_case(**ckw)
    dev = torch.device("cuda")
    dec = _recog.build_reference_decoder(case["odim"], seed=dseed).to(dev)
    step = _recog.ReferenceDecoderStep(dec, case["h"].to(dev), case["hlens"], beam)
    out = recognize_beam_batch(step, case["lpz"].to(dev), case["hlens"], beam, case["eos"], case["eos"],
                               ctc_weight=rkw["ctc_weight"], maxlenratio=rkw.get("maxlenratio", 0.0),
                               minlenratio=rkw.get("minlenratio", 0.0), penalty=rkw.get("penalty", 0.0),
                               nbest=rkw["nbest"], ctc_margin=rkw.get("margin", 0),
                               pre_beam=int(beam * 1.5) if pre_beam else None)
    return case, dec, beam, rkw, out


def _same(got, ref):
    assert len(got) == len(ref)
    for g_hyps, r_hyps in zip(got, ref):
        assert [h["yseq"] for h in g_hyps] == [h["yseq"] for h in r_hyps]
        for g, r in zip(g_hyps, r_hyps):
            assert abs(float(g["score"]) - float(r["score"])) <= ATOL, (float(g["score"]), float(r["score"]))


@pytest.mark.parametrize("name", sorted(CASES))
def test_matches_golden_reference_cpu_pre_beam(name):
    """The stored n-best of the reference's CPU run (pre-beam: 1.5 x beam candidates per hypothesis)."""
    _staged()   # (the decoder modules are the reference's)
    with open(os.path.join(ROOT, "tests", "golden", "recog_batch.json")) as f:
        ref = json.load(f)[name]
    _, _, _, _, out = _ours(name, pre_beam=True)
    _same(out, ref)


@pytest.mark.parametrize("name", sorted(CASES))
def test_matches_reference_on_the_same_gpu_full_vocabulary(name):
    """The reference's loop run live on the GPU, where it scores every token (decoders.py:759-765)."""
    _staged()
    case, dec, beam, rkw, out = _ours(name, pre_beam=False)
    ref = _recog.reference_nbest(dec, case, _recog.recog_args(beam, **rkw), torch.device("cuda"))
    _same(out, ref)


def test_attention_only_and_empty_results():
    """ctc_weight = 0 without lpz runs the glue alone; an utterance with no ended hypothesis gets the dummy entry."""
    from espnet_b200.recog_batch import DecoderStep, recognize_beam_batch

    class Never(DecoderStep):   # a decoder that never proposes eos
        def start(self, n_bb):
            return None

        def step(self, vy, state):
            lp = torch.full((vy.numel(), 6), -2.0, device=vy.device)
            lp[:, 5] = -1e4
            return lp, None, None

        def select(self, state, vidx):
            return None

    out = recognize_beam_batch(Never(), None, [3, 2], 2, 5, 5, ctc_weight=0.0, nbest=2)
    assert len(out) == 2 and all(len(h) >= 1 for h in out)


def _prune_eager(logp, ctc, vscores, eos, aw, cw, first):
    """decoders.py:812-852 as eager torch operations."""
    batch, beam = vscores.shape
    V = logp.size(-1)
    local = aw * logp
    if ctc is not None:
        local[:, 0] = -10000000000.0
        local = local + cw * ctc
    local = local.view(batch, beam, V)
    if first:
        local[:, 1:, :] = -10000000000.0
    eos_v = local[:, :, eos] + vscores
    acc = vscores.view(batch, beam, 1) + local
    acc[:, :, eos] = -10000000000.0
    s, ids = torch.topk(acc.view(batch, -1), beam, 1)
    return s, ids, eos_v, acc.view(batch, -1)


@pytest.mark.parametrize("batch,beam,V", [(3, 4, 50), (32, 10, 5000), (2, 60, 1000), (1, 1, 7), (5, 7, 33)])
@pytest.mark.parametrize("with_ctc,first", [(True, False), (True, True), (False, False)])
def test_beam_prune_kernel_equals_the_eager_formula(batch, beam, V, with_ctc, first):
    """b200_beam_prune against the reference's own lines: scores bitwise (same float32 operation order), the selected
    ids identical wherever the k-th value is not tied, eos scores bitwise, tokens / parent rows consistent."""
    from espnet_b200.recog_batch import beam_prune

    g = torch.Generator(device="cuda").manual_seed(1000 * batch + beam + V)
    dev = torch.device("cuda")
    logp = torch.log_softmax(torch.randn(batch * beam, V, generator=g, device=dev) * 3, dim=1)
    ctc = (torch.randn(batch * beam, V, generator=g, device=dev) * 8 - 10) if with_ctc else None
    if with_ctc:
        ctc[:, V // 2] = -10000000000.0   # an unscored token, as the prefix scorer returns it
    vs = torch.randn(batch, beam, generator=g, device=dev) * 5
    eos = V - 1
    s_ref, id_ref, eos_ref, acc = _prune_eager(logp.clone(), None if ctc is None else ctc.clone(), vs, eos, 0.7, 0.3, first)
    s, ids, eos_v, tok, vidx = beam_prune(logp, ctc, vs, eos, 0.7, 0.3, first)
    assert torch.equal(s, s_ref)                      # the same values, sorted
    assert torch.equal(eos_v, eos_ref)
    assert torch.equal(acc.gather(1, ids), s)         # every id holds the value reported for it
    assert all(len(set(r)) == beam for r in ids.tolist())
    untied = s_ref[:, -1] != torch.topk(acc, min(beam + 1, acc.size(1)), 1)[0][:, -1] if acc.size(1) > beam else None
    strictly = (s_ref[:, 1:] != s_ref[:, :-1]).all(1) if beam > 1 else torch.ones(batch, dtype=torch.bool, device=dev)
    rows = strictly if untied is None else (strictly & untied)
    assert torch.equal(ids[rows], id_ref[rows])
    assert torch.equal(tok.view(batch, beam), ids % V)
    assert torch.equal(vidx.view(batch, beam), ids // V + torch.arange(batch, device=dev).view(-1, 1) * beam)


def test_beam_prune_ties_are_taken_in_index_order():
    from espnet_b200.recog_batch import beam_prune

    dev = torch.device("cuda")
    logp = torch.full((4, 10), -1.0, device=dev)      # batch 2 x beam 2, every candidate equal
    vs = torch.zeros(2, 2, device=dev)
    s, ids, _, _, _ = beam_prune(logp, None, vs, 9, 1.0, 0.0, False)
    assert ids.tolist() == [[0, 1], [0, 1]] and (s == -1.0).all()


@pytest.mark.parametrize("n,V,k", [(320, 5000, 15), (7, 33, 5), (3, 50, 50), (1, 2, 1), (60, 12000, 90), (4, 300, 128)])
def test_pre_beam_candidates_equal_the_eager_lines(n, V, k):
    """b200_pre_beam_candidates against decoders.py:816-822: the same values in the same order (the ids wherever the
    values are distinct), blank never chosen unless k forces it."""
    from espnet_b200.recog_batch import LOGZERO, pre_beam_candidates

    g = torch.Generator(device="cuda").manual_seed(31 * n + V + k)
    logp = torch.log_softmax(torch.randn(n, V, generator=g, device="cuda") * 4, dim=1)
    local = 0.7 * logp
    local[:, 0] = LOGZERO
    v_ref, id_ref = torch.topk(local, k, dim=-1)
    ids = pre_beam_candidates(logp, k, 0.7, 0)
    assert ids.shape == (n, k) and ids.dtype == torch.int64
    assert torch.equal(local.gather(1, ids), v_ref)
    assert all(len(set(r)) == k for r in ids.tolist())
    distinct = (v_ref[:, 1:] != v_ref[:, :-1]).all(1) if k > 1 else torch.ones(n, dtype=torch.bool, device="cuda")
    if k < V:
        distinct &= v_ref[:, -1] != torch.topk(local, k + 1, dim=-1)[0][:, -1]
    assert torch.equal(ids[distinct], id_ref[distinct])


def test_pre_beam_candidates_take_ties_in_index_order():
    from espnet_b200.recog_batch import pre_beam_candidates

    logp = torch.full((2, 40), -2.0, device="cuda")
    logp[1, 30] = -1.0
    ids = pre_beam_candidates(logp, 4, 1.0, 0)
    assert ids.tolist() == [[1, 2, 3, 4], [30, 1, 2, 3]]
