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


def random_alignment(T, ys, g):
    """A random valid CTC alignment (T,) of the label sequence ys: every label gets >= 1 frame, repeated labels
    are separated by >= 1 blank frame, the remaining frames go to random segments (labels or the L+1 blank gaps)."""
    L = len(ys)
    need_blank = [0] + [1 if ys[k] == ys[k - 1] else 0 for k in range(1, L)] + [0]   # gap before label k / after the last
    base = L + sum(need_blank)
    assert T >= base, (T, base)
    # segments in order: gap0, y0, gap1, y1, ..., gapL
    seg_len = []
    for k in range(L):
        seg_len += [need_blank[k], 1]
    seg_len.append(need_blank[L])
    extra = torch.randint(0, len(seg_len), (T - base,), generator=g).tolist()
    for e in extra:
        seg_len[e] += 1
    a = []
    for k, n in enumerate(seg_len):
        a += [0 if k % 2 == 0 else int(ys[k // 2])] * n
    return torch.tensor(a, dtype=torch.long)


def make_logits(kind, B, T, V, ilen, olen, ys, g):
    """(B,T,V) float32 logits of one of the input distributions the parity tests cover.

      randn      N(0,1): near-uniform posteriors (what bench.py measures)
      scale8/30  N(0,1) * 8 or * 30: saturated posteriors, log-probs spanning hundreds of nats per frame
      trained    a converged model: the frame's class on a random valid alignment of the utterance's labels at 0,
                 every other class at -U(20,40), plus N(0,1) noise
      mismatch   the same, but the alignment spells a DIFFERENT label sequence than the targets (wrong transcript)
      level      first half of every utterance `trained`, second half `randn` (an abrupt level change)
    """
    if kind == "randn":
        return torch.randn(B, T, V, generator=g)
    if kind in ("scale8", "scale30"):
        return torch.randn(B, T, V, generator=g) * float(kind[5:])
    x = -(20.0 + 20.0 * torch.rand(B, T, V, generator=g)) + torch.randn(B, T, V, generator=g)
    for b in range(B):
        Tb, Lb = int(ilen[b]), int(olen[b])
        if Tb <= 0:
            continue
        yb = ys[b, :Lb].clone()
        if kind == "mismatch" and Lb > 2:
            yb = yb[torch.randperm(Lb, generator=g)]            # same labels, different order
        rep = int((yb[1:] == yb[:-1]).sum()) if Lb > 1 else 0
        if Tb < Lb + rep:                                       # infeasible anyway: leave the noise
            continue
        a = random_alignment(Tb, yb.tolist(), g)
        x[b, torch.arange(Tb), a] = torch.randn(Tb, generator=g) * 0.1
    if kind == "level":
        r = torch.randn(B, T, V, generator=g)
        for b in range(B):
            h = int(ilen[b]) // 2
            x[b, h:] = r[b, h:]
    return x


KINDS = ["randn", "scale8", "scale30", "trained", "mismatch", "level"]


def make_loss_case_kind(kind, B, T, V, L, seed, dtype=torch.float32, min_ratio=0.6):
    """make_loss_case with the logits drawn from `kind` (lengths and labels as in make_loss_case)."""
    c = make_loss_case(B, T, V, L, seed, dtype=torch.float32, min_ratio=min_ratio)
    g = torch.Generator().manual_seed(seed + 9000)
    c["x"] = make_logits(kind, B, T, V, c["ilen"], c["olen"], c["ys"], g).to(dtype)
    return c


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


def make_prefix_case(B, T, V, H, seed, nsteps, snum=0, kind="randn", min_ratio=0.6, L=None):
    """log-probs (B,T,V) for prefix scoring.  kind as in make_logits; the `trained` kinds are built around a seeded
    label sequence per utterance (returned as "labels": the walk tests extend hypotheses along it)."""
    g = torch.Generator().manual_seed(seed)
    labels = None
    if kind == "randn":
        x = torch.randn(B, T, V, generator=g)
    xlens = torch.randint(max(int(min_ratio * T), 2), T + 1, (B,), generator=g)
    xlens[0] = T
    if kind != "randn":
        L = L or max(2, int(xlens.min()) // 4)
        labels = torch.randint(1, V - 1, (B, L), generator=g)
        x = make_logits(kind, B, T, V, xlens, torch.full((B,), L), labels, g)
    lp = torch.log_softmax(x, dim=-1)
    return {"lp": lp, "xlens": xlens, "H": H, "eos": V - 1, "nsteps": nsteps, "snum": snum, "seed": seed,
            "labels": labels}


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
        # BPE-sized vocabulary, joint decoding with ctc_weight 0.3 (BASELINE config 3's weights) and a constructed
        # near-tie (tests/_toy.make_search_input twin): pairs of hypotheses a few 1e-6 apart
        "v5000_near_tie": dict(T=66, V=5000, L=5, seed=45, beam=10, weights={"decoder": 0.7, "ctc": 0.3, "length_bonus": 0.1},
                               pre_beam_score_key="full", pre_beam_ratio=1.5, maxlenratio=0.0, twin=3e-5, peak=15.0, tboost=11.0),
        # the same vocabulary with every token scored by CTC (no pre-beam: recognize_beam_batch's CUDA path)
        "v5000_full_vocab": dict(T=44, V=5000, L=4, seed=46, beam=8, weights={"decoder": 0.5, "ctc": 0.5},
                                 pre_beam_score_key=None, pre_beam_ratio=1.5, maxlenratio=0.0, twin=2e-5, peak=15.0, tboost=11.0),
    }
