"""Host-side logic that needs no GPU: module construction, the reference-interface mirror,
the absence of any CPU fallback on the b200 path, and the oracle quarantine."""
import os
import re

import numpy as np
import pytest
import torch

from tests._util import ROOT, golden


def test_unknown_ctc_type_raises_value_error():
    from espnet_b200.ctc import CTC, CTCEspnet1

    with pytest.raises(ValueError):
        CTC(5, 10, ctc_type="nope")
    with pytest.raises(ValueError):
        CTCEspnet1(5, 10, 0.0, ctc_type="nope")


@pytest.mark.parametrize("name", ["tiny", "v64"])
def test_builtin_branch_reproduces_reference_loss(name):
    """ctc_type="builtin" is the reference's own branch: same loss as the recorded one."""
    from espnet_b200.ctc import CTC

    g = golden(f"loss_{name}.npz")
    x = torch.from_numpy(g["x"]).double()
    ilen, olen, ys = (torch.from_numpy(g[k]) for k in ("ilen", "olen", "ys"))
    ctc = CTC(x.shape[2], 4, ctc_type="builtin")
    tg = torch.cat([ys[i, : olen[i]] for i in range(x.shape[0])])
    loss = ctc.loss_fn(x.transpose(0, 1), tg, ilen, olen)
    np.testing.assert_allclose(loss.item(), g["loss_f64"], rtol=1e-12)


def test_b200_path_has_no_cpu_fallback():
    from espnet_b200.ctc import CTC

    ctc = CTC(7, 4, ctc_type="b200")
    hs = torch.randn(2, 9, 4)
    with pytest.raises(RuntimeError, match="CUDA"):
        ctc(hs, torch.tensor([9, 7]), torch.tensor([[1, 2], [3, 0]]), torch.tensor([2, 1]))


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from espnet_b200 import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _lib.load()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "espnet_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"\boracle\b", text), f"{f} mentions the oracle"


def test_empty_like_strided_matches_ctc_lo_view():
    from espnet_b200.ctc_loss import _empty_like_strided

    x = torch.empty(3, 5, 8).transpose(0, 1)  # (T,B,V) view of (B,T,V)
    g = _empty_like_strided(x)
    assert g.shape == x.shape and g.stride() == x.stride()
    y = torch.empty(5, 3, 8)
    assert _empty_like_strided(y).stride() == y.stride()


def test_forced_align_of_the_builtin_type_runs_on_the_host():
    """ctc_type="builtin" keeps the reference's numpy forced_align (pytorch_backend/ctc.py:153-216)."""
    from espnet_b200.ctc import CTCEspnet1

    g = golden("viterbi.npz")
    for k in range(3):
        h, y = torch.from_numpy(g[f"h_{k}"]), torch.from_numpy(g[f"y_{k}"])
        V = h.shape[1]
        ctc = CTCEspnet1(V, V, 0.0, ctc_type="builtin")
        with torch.no_grad():
            ctc.ctc_lo.weight.copy_(torch.eye(V))
            ctc.ctc_lo.bias.zero_()
            assert [int(t) for t in ctc.forced_align(h.unsqueeze(0), y)] == g[f"align_{k}"].tolist()


def test_scorer_derives_from_the_reference_interfaces_when_they_are_importable():
    import importlib

    from oracle import stage_ref

    if not stage_ref.activate():
        pytest.skip("oracle/_ref is not staged")
    import espnet_b200.scorers as S

    S = importlib.reload(S)
    from espnet.nets.scorer_interface import BatchPartialScorerInterface, ScorerInterface

    assert issubclass(S.CTCPrefixScorer, BatchPartialScorerInterface) and issubclass(S.CTCPrefixScorer, ScorerInterface)


def test_pack_labels_host_and_mixed_inputs():
    from espnet_b200.ctc_loss import pack_labels

    out = pack_labels([3, 4, 5], torch.tensor([0, 2]), torch.tensor([7, 9], dtype=torch.int32), [2, 1], torch.device("cpu"))
    assert [o.tolist() for o in out] == [[3, 4, 5], [0, 2], [7, 9], [2, 1]]
    assert all(o.dtype == torch.int64 for o in out)


def test_running_end_detection_equals_the_list_form():
    """recog_batch keeps (best score, best score per length) instead of re-reading every ended hypothesis at every
    position: the same decisions as end_detect (e2e_asr_common.py:21-51) on a random stream of ended hypotheses."""
    import random

    from espnet_b200.batch_beam_search import end_detect
    from espnet_b200.recog_batch import _add_ended, _end_detect

    rng = random.Random(5)
    seen = {True: 0, False: 0}
    for trial in range(200):
        ended, summary = [], [None, {}]
        for i in range(12):
            for _ in range(rng.randrange(0, 3)):
                v = -rng.random() * (2.0 if rng.random() < 0.5 else 40.0) - 0.1 * i
                _add_ended(ended, summary, [0] * (i + 2), v)      # hypotheses that end at position i hold i + 2 tokens
            want = end_detect([(len(h["yseq"]), float(h["score"])) for h in ended], i)
            assert _end_detect(summary, i) == want
            seen[want] += 1
    assert seen[True] > 0 and seen[False] > 0
