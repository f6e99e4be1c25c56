"""Developer check: the near-tie search case under the reference's non-batch BeamSearch, with the reference's
own scorer and with the b200 scorer, side by side (run on the GPU box)."""
import os
import sys

import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
from oracle import stage_ref  # noqa: E402

assert stage_ref.activate()
from espnet.nets.beam_search import BeamSearch  # noqa: E402
from espnet.nets.scorer_interface import BatchScorerInterface  # noqa: E402
from espnet.nets.scorers.ctc import CTCPrefixScorer as RefScorer  # noqa: E402
from espnet.nets.scorers.length_bonus import LengthBonus  # noqa: E402

from espnet_b200.scorers import CTCPrefixScorer  # noqa: E402
from tests import _cases, _toy  # noqa: E402


class RefBigram(_toy.Bigram, BatchScorerInterface):
    pass


name = sys.argv[1] if len(sys.argv) > 1 else "v5000_near_tie"
cfg = _cases.nbest_cases()[name]
x, table, y = _toy.make_search_input(**cfg)
V = cfg["V"]
for dev in ("cpu", "cuda"):
    for tag, mk in (("ref", RefScorer), ("b200", CTCPrefixScorer)):
        if tag == "b200" and dev == "cpu":
            continue
        scorers = {"decoder": RefBigram(table.to(dev)), "ctc": mk(_toy.LogSoftmaxCTC(), V - 1), "length_bonus": LengthBonus(V)}
        bs = BeamSearch(scorers=scorers, weights=cfg["weights"], beam_size=cfg["beam"], vocab_size=V, sos=V - 1, eos=V - 1,
                        pre_beam_score_key=cfg["pre_beam_score_key"], pre_beam_ratio=cfg["pre_beam_ratio"])
        with torch.no_grad():
            nb = bs(x.to(dev), maxlenratio=cfg["maxlenratio"], minlenratio=0.0)
        print(dev, tag)
        for h in nb[:4]:
            print("   ", h.yseq.tolist()[:8], "%.7f" % float(h.score), {k: "%.7f" % float(v) for k, v in h.scores.items()})
