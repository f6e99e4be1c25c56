"""Greedy path (row argmax) and Viterbi forced alignment on the GPU:
bit-exact against torch.argmax, the golden alignments recorded from the reference's
CTC.forced_align, and the C oracle on larger seeded cases."""
import numpy as np
import pytest
import torch

from tests._util import golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("B,T,V", [(3, 17, 30), (2, 40, 1000), (4, 9, 129), (1, 5, 5000)])
def test_argmax_matches_torch(dtype, B, T, V):
    from espnet_b200.ctc_align import ctc_argmax

    g = torch.Generator().manual_seed(B * 1000 + V)
    x = torch.randn(B, T, V, generator=g).to(dtype).cuda()   # bf16: many exact ties -> first index wins
    assert (ctc_argmax(x) == torch.argmax(x, dim=2)).all()
    # the (T,B,V) buffer seen as its (B,T,V) view: strided rows
    xt = torch.randn(T, B, V, generator=g).to(dtype).cuda().transpose(0, 1)
    assert (ctc_argmax(xt) == torch.argmax(xt, dim=2)).all()


def test_argmax_ties_nan_and_inf():
    from espnet_b200.ctc_align import ctc_argmax

    x = torch.zeros(1, 4, 64, device="cuda")
    x[0, 1, 40] = 1.0
    x[0, 1, 50] = 1.0
    x[0, 2, 33] = float("nan")
    x[0, 2, 3] = float("inf")
    x[0, 3] = -float("inf")
    got = ctc_argmax(x)
    assert got.tolist() == [[0, 40, 33, 0]] == torch.argmax(x, dim=2).tolist()


def test_argmax_full_size_bf16():
    """C2 shape: B=64, T=500, V=5000 bf16 (the one-pass row loader at full size)."""
    from espnet_b200.ctc_align import ctc_argmax

    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(64, 500, 5000, generator=g, device="cuda").to(torch.bfloat16)
    assert (ctc_argmax(x) == torch.argmax(x, dim=2)).all()


def test_module_argmax_uses_kernel():
    from espnet_b200.ctc import CTC

    torch.manual_seed(3)
    ctc = CTC(50, 16, ctc_type="b200").cuda()
    hs = torch.randn(2, 11, 16, device="cuda")
    assert (ctc.argmax(hs) == torch.argmax(ctc.ctc_lo(hs), dim=2)).all()


def _lpz(h):
    return torch.from_numpy(h).log_softmax(-1)


def test_forced_align_matches_golden_reference():
    from espnet_b200.ctc_align import ctc_forced_align

    g = golden("viterbi.npz")
    k = 0
    while f"h_{k}" in g.files:
        ali = ctc_forced_align(_lpz(g[f"h_{k}"]).cuda(), g[f"y_{k}"])
        assert ali == g[f"align_{k}"].tolist()
        k += 1
    assert k == 3


@pytest.mark.parametrize("T,V,L,seed", [(300, 40, 60, 1), (500, 5000, 150, 2), (64, 7, 0, 3), (90, 12, 44, 4),
                                         (4000, 300, 1000, 5)])
def test_forced_align_matches_oracle(T, V, L, seed):
    from espnet_b200.ctc_align import ctc_forced_align
    from oracle import ctc_oracle as O

    g = torch.Generator().manual_seed(seed)
    lpz = torch.randn(T, V, generator=g).log_softmax(-1)
    y = torch.randint(1, V, (L,), generator=g)
    if L >= 4:
        y[1] = y[0]
        y[3] = y[2]
    ali = ctc_forced_align(lpz.cuda(), y)
    ref = O.viterbi(lpz.numpy().astype(np.float32), y.numpy())
    assert ali == ref.tolist()


def test_espnet1_module_forced_align():
    from espnet_b200.ctc import CTCEspnet1

    g = golden("viterbi.npz")
    h, y = torch.from_numpy(g["h_1"]), g["y_1"]
    V = h.shape[1]
    ctc = CTCEspnet1(V, V, 0.0, ctc_type="b200").cuda()
    with torch.no_grad():
        ctc.ctc_lo.weight.copy_(torch.eye(V))
        ctc.ctc_lo.bias.zero_()
        ali = ctc.forced_align(h.cuda().unsqueeze(0), y)
    assert ali == g["align_1"].tolist()
