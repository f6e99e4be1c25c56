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
