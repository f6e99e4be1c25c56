"""nbest.npz: n-best lists of the REFERENCE's BatchBeamSearch (espnet/nets/batch_beam_search.py) with the
reference's CTCPrefixScorer (espnet/nets/scorers/ctc.py), LengthBonus and a toy bigram decoder, CPU, fixed seeds.

    PYTHONPATH=/root/reference:tests/golden/stubs python tests/golden/make_golden_nbest.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, "stubs"))

from tests import _cases, _toy  # noqa: E402


def main():
    from espnet.nets.batch_beam_search import BatchBeamSearch
    from espnet.nets.beam_search import BeamSearch
    from espnet.nets.scorer_interface import BatchScorerInterface
    from espnet.nets.scorers.ctc import CTCPrefixScorer
    from espnet.nets.scorers.length_bonus import LengthBonus

    class RefBigram(_toy.Bigram, BatchScorerInterface):
        pass

    out = {}
    for name, cfg in _cases.nbest_cases().items():
        x, table, y = _toy.make_search_input(**cfg)
        V = cfg["V"]
        res = {}
        for tag, cls in (("batch", BatchBeamSearch), ("single", BeamSearch)):
            scorers = {"decoder": RefBigram(table), "ctc": CTCPrefixScorer(_toy.LogSoftmaxCTC(), V - 1),
                       "length_bonus": LengthBonus(V)}
            bs = cls(scorers=scorers, weights=cfg["weights"], beam_size=cfg["beam"], vocab_size=V, sos=V - 1,
                     eos=V - 1, pre_beam_score_key=cfg["pre_beam_score_key"], pre_beam_ratio=cfg["pre_beam_ratio"])
            with torch.no_grad():
                nbest = bs(x, maxlenratio=cfg["maxlenratio"], minlenratio=0.0)
            res[tag] = nbest
        nbest = res["batch"]
        n = len(nbest)
        maxlen = max(len(h.yseq) for h in nbest)
        ys = np.full((n, maxlen), -1, dtype=np.int64)
        for i, h in enumerate(nbest):
            ys[i, : len(h.yseq)] = h.yseq.numpy()
        out[f"{name}_yseq"] = ys
        out[f"{name}_score"] = np.array([float(h.score) for h in nbest], dtype=np.float64)
        for k in nbest[0].scores:
            out[f"{name}_scores_{k}"] = np.array([float(h.scores[k]) for h in nbest], dtype=np.float64)
        out[f"{name}_label"] = y.numpy()
        # the non-batch BeamSearch of the reference (per-hypothesis numpy CTCPrefixScore): best hypothesis, and its
        # whole n-best list
        out[f"{name}_single_best"] = res["single"][0].yseq.numpy()
        sn = res["single"]
        sy = np.full((len(sn), max(len(h.yseq) for h in sn)), -1, dtype=np.int64)
        for i, h in enumerate(sn):
            sy[i, : len(h.yseq)] = h.yseq.numpy()
        out[f"{name}_single_yseq"] = sy
        out[f"{name}_single_score"] = np.array([float(h.score) for h in sn], dtype=np.float64)
        gap = np.diff(out[f"{name}_score"])
        print(name, "n-best", n, "best", ys[0][ys[0] >= 0].tolist(), "label", y.tolist(),
              "min score gap %.3g" % (np.abs(gap).min() if n > 1 else 0.0),
              "single==batch", res["single"][0].yseq.tolist() == nbest[0].yseq.tolist())
    np.savez_compressed(os.path.join(HERE, "nbest.npz"), **out)


if __name__ == "__main__":
    main()
