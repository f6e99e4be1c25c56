"""Seeded synthetic inputs shared by the golden generator and the tests."""
import torch


def make_loss_case(B, T, V, L, seed, dtype=torch.float32, full=False, forced_repeat=True,
                   min_ratio=0.6, scale=1.0):
    """logits (B,T,V) ~ N(0, scale), ilen in [min_ratio*T, T] with ilen[0] = T,
    olen in [L/2, L], labels in [1, V-1], one forced-repeat utterance; feasible by construction."""
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(B, T, V, generator=g) * scale).to(dtype)
    ilen = torch.randint(int(min_ratio * T), T + 1, (B,), generator=g)
    ilen[0] = T
    if full:
        ilen[:] = T
    olen = torch.randint(max(L // 2, 1), L + 1, (B,), generator=g) if L > 0 else torch.zeros(B, dtype=torch.long)
    ys = torch.randint(1, V, (B, max(L, 1)), generator=g)
    if forced_repeat and L >= 4:
        ys[0, 1] = ys[0, 0]
        ys[0, 3] = ys[0, 0]
    # keep every utterance feasible: T_b >= L_b + #repeats
    for b in range(B):
        lb = int(olen[b])
        rep = int((ys[b, 1:lb] == ys[b, : lb - 1]).sum()) if lb > 1 else 0
        need = lb + rep
        if int(ilen[b]) < need:
            olen[b] = max(1, int(ilen[b]) // 2)
    return {"x": x, "ilen": ilen, "olen": olen, "ys": ys}


def loss_cases():
    """Small cases recorded as golden fixtures (reference outputs stored next to the inputs)."""
    cases = {
        # C1: mini_an4-shaped
        "c1": make_loss_case(8, 200, 30, 40, seed=1235),
        # reference test shapes (test/test_loss.py:17-63): in_length [11,17,15], out_length [4,2,3]
        "tiny": None,
        # V multiple of 4 (vectorised path in fp32), ragged
        "v64": make_loss_case(4, 50, 64, 10, seed=11),
        # many repeated labels (V small)
        "repeats": make_loss_case(3, 40, 4, 12, seed=5),
    }
    g = torch.Generator().manual_seed(3)
    x = torch.randn(3, 17, 7, generator=g)
    cases["tiny"] = {"x": x, "ilen": torch.tensor([11, 17, 15]), "olen": torch.tensor([4, 2, 3]),
                     "ys": torch.tensor([[1, 2, 2, 3], [4, 5, 1, 1], [6, 6, 2, 1]])}
    # empty target, single frame, and an infeasible utterance (T < L + repeats) in one batch
    g = torch.Generator().manual_seed(9)
    x = torch.randn(4, 6, 5, generator=g)
    cases["edge"] = {"x": x, "ilen": torch.tensor([6, 1, 5, 3]), "olen": torch.tensor([0, 1, 4, 3]),
                     "ys": torch.tensor([[1, 1, 1, 1], [2, 1, 1, 1], [1, 1, 1, 1], [1, 2, 3, 1]])}
    return cases


def make_prefix_case(B, T, V, H, seed, nsteps, snum=0):
    g = torch.Generator().manual_seed(seed)
    lp = torch.log_softmax(torch.randn(B, T, V, generator=g), dim=-1)
    xlens = torch.randint(int(0.6 * T), T + 1, (B,), generator=g)
    xlens[0] = T
    return {"lp": lp, "xlens": xlens, "H": H, "eos": V - 1, "nsteps": nsteps, "snum": snum, "seed": seed}


def prefix_cases():
    return {
        "full": make_prefix_case(2, 30, 12, 3, seed=21, nsteps=5),
        "partial": make_prefix_case(2, 30, 12, 3, seed=22, nsteps=5, snum=4),
        "single": make_prefix_case(1, 25, 8, 4, seed=23, nsteps=6),
        "wide": make_prefix_case(3, 40, 260, 2, seed=24, nsteps=4),
    }


def viterbi_cases():
    # (T, V, L, seed)
    return [(20, 6, 4, 31), (35, 9, 7, 32), (12, 5, 1, 33)]


def nbest_cases():
    """Beam-search cases: name -> config of tests/_toy.make_search_input and of the search itself."""
    return {
        # CTC prefix scores alone (plus the length bonus), every token scored
        "ctc_only": dict(T=48, V=24, L=9, seed=41, beam=5, weights={"ctc": 1.0, "length_bonus": 0.1},
                         pre_beam_score_key=None, pre_beam_ratio=1.5, maxlenratio=0.0),
        # joint decoding: bigram "decoder" 0.7 + CTC 0.3, pre-beam on the full scorers' sum
        "joint_full": dict(T=60, V=40, L=12, seed=42, beam=6, weights={"decoder": 0.7, "ctc": 0.3, "length_bonus": 0.1},
                           pre_beam_score_key="full", pre_beam_ratio=1.5, maxlenratio=0.0),
        # pre-beam on the decoder alone, fixed maximum length (the <eos>-at-the-end branch)
        "joint_decoder": dict(T=40, V=32, L=8, seed=43, beam=4, weights={"decoder": 0.5, "ctc": 0.5},
                              pre_beam_score_key="decoder", pre_beam_ratio=2.0, maxlenratio=0.3),
        # wide beam against a small vocabulary: fewer live candidates than beam slots at step 0
        "wide_beam": dict(T=30, V=12, L=6, seed=44, beam=10, weights={"ctc": 1.0, "length_bonus": 0.1},
                          pre_beam_score_key=None, pre_beam_ratio=1.5, maxlenratio=0.0),
    }
