"""Parity of the CUDA CTC prefix scorer (through the C ABI) with CTCPrefixScoreTH.

Bar (BASELINE.json north_star): prefix scores within 1e-4 absolute of the reference.  Entries
the reference sets to its log(0) sentinel (-1e10, e.g. blank, unscored tokens, infeasible
prefixes) must be sentinel-like (< -1e9) here too; an fp32 number of that size has a spacing
of 1024, so they are compared by class, not by value.
"""
import numpy as np
import pytest
import torch

from tests import _cases
from tests._util import golden

pytestmark = pytest.mark.gpu
ATOL = 1e-4


@pytest.fixture(autouse=True, params=["mma", "fp32"])
def score_kernel(request):
    """Every test runs with the tensor-core score kernel (the default) and with the fp32 score kernels."""
    from espnet_b200 import _lib

    lib = _lib.load()
    lib.b200_ctc_debug_prefix_fp32_score(1 if request.param == "fp32" else 0)
    yield request.param
    lib.b200_ctc_debug_prefix_fp32_score(0)


def _close(got, ref, atol=ATOL):
    got, ref = np.asarray(got, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    real = np.abs(ref) < 1e9
    assert (got[~real] < -1e9).all()
    if real.any():
        # absolute bar, plus one fp32 spacing of the value itself for the large eos entries
        tol = atol + 2.4e-7 * np.abs(ref[real])
        assert (np.abs(got[real] - ref[real]) <= tol).all(), np.abs(got[real] - ref[real]).max()


@pytest.mark.parametrize("name", ["full", "partial", "single", "wide"])
def test_walk_matches_golden_reference(name):
    from espnet_b200.ctc_prefix_score import CTCPrefixScoreB200

    g = golden(f"prefix_{name}.npz")
    lp, xlens, H, eos, snum = torch.from_numpy(g["lp"]), g["xlens"], int(g["H"]), int(g["eos"]), int(g["snum"])
    B, T, V = lp.shape
    x = lp.clone().cuda()
    impl = CTCPrefixScoreB200(x, torch.from_numpy(xlens), 0, eos)
    # the constructor pads the caller's tensor in place, as the reference does
    for b in range(B):
        if xlens[b] < T:
            assert float(x[b, xlens[b]:, 1:].max()) <= -1e9 and float(x[b, xlens[b]:, 0].abs().max()) == 0.0
    y = [[eos] for _ in range(B * H)]
    state = None
    step = 0
    while f"scores_{step}" in g.files:
        ids = torch.from_numpy(g[f"ids_{step}"]).cuda() if snum > 0 else None
        scores, st = impl(y, state, ids)
        _close(scores.cpu().numpy(), g[f"scores_{step}"])
        best = torch.from_numpy(g[f"best_{step}"])
        state = impl.index_select_state(st, best.cuda())
        _close(state[0].cpu().numpy(), g[f"r_{step}"])
        _close(state[1].cpu().numpy(), g[f"s_{step}"])
        hyp_idx = (best // V + torch.arange(B).view(-1, 1) * H).view(-1)
        tok = (best % V).view(-1)
        y = [y[int(hyp_idx[k])] + [int(tok[k])] for k in range(B * H)]
        step += 1
    assert step >= 4


def _oracle_walk(c, nsteps, snum=0, windows=None, seed=0):
    """Runs the CUDA scorer and the C oracle side by side on a seeded hypothesis walk."""
    from espnet_b200.ctc_prefix_score import CTCPrefixScoreB200
    from oracle import ctc_oracle as O

    lp, xlens, H, eos = c["lp"], c["xlens"], c["H"], c["eos"]
    B, T, V = lp.shape
    g = torch.Generator().manual_seed(seed)
    impl = CTCPrefixScoreB200(lp.clone().cuda(), xlens, 0, eos)
    x_np = lp.numpy().copy()
    r_prev = np.repeat(O.prefix_init(x_np, xlens.numpy()), H, axis=2)
    s_prev, last = None, np.full(B * H, eos, dtype=np.int64)
    y = [[eos] for _ in range(B * H)]
    state = None
    for step in range(nsteps):
        ids = None
        if snum:
            ids = torch.stack([torch.randperm(V - 1, generator=g)[:snum] + 1 for _ in range(B * H)])
        scores, st = impl(y, state, None if ids is None else ids.cuda())
        ref, psi, _ = O.prefix_score(x_np, xlens.numpy(), H, 0, eos, r_prev, s_prev, last, step,
                                     scoring_ids=None if ids is None else ids.numpy())
        _close(scores.cpu().numpy(), ref)
        hyp = torch.randint(0, H, (B, H), generator=g)
        if ids is None:
            tok = torch.randint(1, V, (B, H), generator=g)
        else:
            pick = torch.randint(0, snum, (B, H), generator=g)
            tok = ids.view(B, H, snum)[torch.arange(B).view(-1, 1), hyp, pick]
        best = hyp * V + tok
        state = impl.index_select_state(st, best.cuda())
        parent = (hyp + torch.arange(B).view(-1, 1) * H).view(-1).numpy()
        token = tok.view(-1).numpy()
        r_prev, s_prev = O.prefix_select(x_np, H, 0, r_prev, psi, last, step, parent, token)
        _close(state[0].cpu().numpy(), r_prev)
        _close(state[1].cpu().numpy(), s_prev)
        last = token
        y = [y[int(parent[k])] + [int(token[k])] for k in range(B * H)]


def test_c3_shaped_walk_vs_oracle():
    """BASELINE config 3 shape (beam=10, T=500, V=5000) at B=3 utterances to bound oracle time."""
    _oracle_walk(_cases.make_prefix_case(3, 500, 5000, 10, seed=1237, nsteps=4), nsteps=4, seed=1)


def test_pre_beam_subset_vs_oracle():
    """scoring_ids = 15 candidates per hypothesis (1.5 x beam, BatchBeamSearch pre-beam)."""
    _oracle_walk(_cases.make_prefix_case(2, 200, 500, 10, seed=1238, nsteps=5), nsteps=5, snum=15, seed=2)


@pytest.mark.parametrize("H", [1, 5, 13, 20])
def test_hypothesis_counts(H):
    _oracle_walk(_cases.make_prefix_case(2, 60, 70, H, seed=40 + H, nsteps=3), nsteps=3, seed=H)


def test_repeated_last_label_and_eos():
    """Extending a prefix by its own last label uses phi = r_prev[:,b] only; eos gets r_sum[xlen-1]."""
    from espnet_b200.ctc_prefix_score import CTCPrefixScoreB200
    from oracle import ctc_oracle as O

    c = _cases.make_prefix_case(2, 40, 9, 2, seed=51, nsteps=2)
    lp, xlens, H, eos = c["lp"], c["xlens"], c["H"], c["eos"]
    B, T, V = lp.shape
    impl = CTCPrefixScoreB200(lp.clone().cuda(), xlens, 0, eos)
    x_np = lp.numpy().copy()
    r0 = np.repeat(O.prefix_init(x_np, xlens.numpy()), H, axis=2)
    y = [[eos]] * (B * H)
    _, st = impl(y, None)
    _, psi, _ = O.prefix_score(x_np, xlens.numpy(), H, 0, eos, r0, None, np.full(B * H, eos), 0)
    tok = torch.tensor([[3, 3], [5, 3]])
    best = torch.tensor([[0, 1], [0, 1]]) * V + tok
    state = impl.index_select_state(st, best.cuda())
    parent = np.array([0, 1, 2, 3])
    r1, s1 = O.prefix_select(x_np, H, 0, r0, psi, np.full(B * H, eos), 0, parent, tok.view(-1).numpy())
    y2 = [[eos, int(t)] for t in tok.view(-1)]
    scores, _ = impl(y2, state)
    ref, _, _ = O.prefix_score(x_np, xlens.numpy(), H, 0, eos, r1, s1, tok.view(-1).numpy(), 1)
    _close(scores.cpu().numpy(), ref)
    got = scores.cpu().numpy()
    assert (got[:, 0] < -1e9).all()                      # blank is never scored
    assert np.isfinite(got[:, eos]).all() and (got[:, eos] > -1e9).all()


def test_attention_window_margin():
    """margin > 0: the frame window comes from the attention weights (ctc_prefix_score.py:145-154)."""
    from espnet_b200.ctc_prefix_score import CTCPrefixScoreB200
    from oracle import ctc_oracle as O

    c = _cases.make_prefix_case(1, 80, 20, 3, seed=61, nsteps=2)
    lp, xlens, H, eos = c["lp"], c["xlens"], c["H"], c["eos"]
    B, T, V = lp.shape
    impl = CTCPrefixScoreB200(lp.clone().cuda(), xlens, 0, eos, margin=5)
    att = torch.zeros(B * H, T)
    att[0, 20] = 1.0
    att[1, 30] = 1.0
    att[2, 26] = 1.0
    scores, st = impl([[eos]] * (B * H), None, None, att.cuda())
    # reference arithmetic: f_min=20, f_max=30 -> start=min(1, max(15,0,1))=1, end=35
    assert st[2] == 20 and st[3] == 30
    x_np = lp.numpy().copy()
    r0 = np.repeat(O.prefix_init(x_np, xlens.numpy()), H, axis=2)
    ref, _, _ = O.prefix_score(x_np, xlens.numpy(), H, 0, eos, r0, None, np.full(B * H, eos), 0, start=1, end=35)
    _close(scores.cpu().numpy(), ref)


def test_scorer_adapter_batch_and_single_paths_agree():
    """CTCPrefixScorer: batch_score_partial (BatchBeamSearch) and score_partial (BeamSearch)
    give the same scores for the same prefix, as test_batch_beam_search_equal requires."""
    from espnet_b200.ctc import CTCEspnet1
    from espnet_b200.scorers import CTCPrefixScorer

    torch.manual_seed(3)
    V, D, T = 12, 8, 30
    ctc = CTCEspnet1(V, D, 0.0, ctc_type="b200").cuda()
    x = torch.randn(T, D, device="cuda")
    eos = V - 1
    a = CTCPrefixScorer(ctc, eos)
    b = CTCPrefixScorer(ctc, eos)
    st_a = a.batch_init_state(x)
    st_b = b.init_state(x)
    ids = torch.tensor([[1, 4, 7, eos]], device="cuda")
    y = torch.tensor([[eos]], device="cuda")
    sa, state_a = a.batch_score_partial(y, ids, [st_a], x)
    sb, state_b = b.score_partial(y[0], ids[0], st_b, x)
    np.testing.assert_allclose(sa[0, ids[0]].cpu().numpy(), sb.cpu().numpy(), atol=1e-5)
    # second step through select_state on both paths
    ha = a.select_state(state_a, 0, 4)
    hb = b.select_state(state_b, 1)
    y2 = torch.tensor([[eos, 4]], device="cuda")
    ids2 = torch.tensor([[2, 4, 9]], device="cuda")
    sa2, _ = a.batch_score_partial(y2, ids2, [ha], x)
    sb2, _ = b.score_partial(y2[0], ids2[0], hb, x)
    np.testing.assert_allclose(sa2[0, ids2[0]].cpu().numpy(), sb2.cpu().numpy(), atol=1e-5)


@pytest.mark.parametrize("T,V,H,snum,margin,kind", [(203, 517, 10, 15, 0, "randn"), (64, 40, 3, 39, 0, "randn"),
                                                    (333, 1000, 20, 60, 20, "randn"), (31, 33, 1, 1, 0, "randn"),
                                                    (203, 517, 10, 15, 0, "scale30"), (150, 300, 5, 40, 15, "trained")])
def test_candidate_path_from_column_copy_is_bit_identical(T, V, H, snum, margin, kind):
    """scoring_ids with and without the (B, V, T) column copy of x: the same bits, step after step -- ragged lengths,
    T not a multiple of the copy's 32-frame tiles, an attention window, candidates that repeat the last label and
    blank / eos / -1 entries (which the fill kernel has already answered), saturated posteriors (pairs whose scaled sum
    underflows and is redone in log space: by the fix kernel on one side, by the scoring warp itself on the other)."""
    from espnet_b200.ctc_prefix_score import CTCPrefixScoreB200

    c = _cases.make_prefix_case(3, T, V, H, seed=4242 + T, nsteps=4, kind=kind)
    lp, xlens, eos = c["lp"], c["xlens"], c["eos"]
    B = lp.size(0)
    g = torch.Generator().manual_seed(T)
    a = CTCPrefixScoreB200(lp.clone().cuda(), xlens, 0, eos, margin=margin, columns=True)
    b = CTCPrefixScoreB200(lp.clone().cuda(), xlens, 0, eos, margin=margin, columns=False)
    y = torch.full((B * H, 1), eos, dtype=torch.int64, device="cuda")
    sa = sb = None
    fixes = 0
    for step in range(4):
        ids = torch.stack([torch.randperm(V - 1, generator=g)[:snum] + 1 for _ in range(B * H)])
        ids[:, 0] = y[:, -1].cpu()                       # the repeated last label is always a candidate
        if snum > 3:
            ids[0, 1], ids[1, 2], ids[2, 3] = 0, eos, -1
        att_w = None
        if margin:
            centre = min(T - 1, 8 + 12 * step)
            att_w = torch.zeros(B * H, T)
            att_w[:, centre] = 1.0
            att_w = att_w.cuda()
        ga, sta = a(y, sa, ids.cuda(), att_w)
        gb, stb = b(y, sb, ids.cuda(), att_w)
        assert a._xc is not None and b._xc is None
        # (the queue counter of the four-kernel path: csrc/ctc_prefix.cu carve_prefix)
        fixes += int(b._work.view(torch.int32)[2 * B * T * ((H + 3) & ~3) + 3 * B * H])
        assert torch.equal(ga, gb)
        assert torch.equal(sta[1], stb[1])              # log_psi
        hyp = torch.randint(0, H, (B, H), generator=g)
        pick = torch.randint(0, snum, (B, H), generator=g)
        tok = ids.clamp(min=1).view(B, H, snum)[torch.arange(B).view(-1, 1), hyp, pick]
        tok = torch.where(tok == eos, torch.ones_like(tok), tok)
        best = (hyp * V + tok).cuda()
        sa = a.index_select_state(sta, best)
        sb = b.index_select_state(stb, best)
        assert torch.equal(sa[0], sb[0]) and torch.equal(sa[1], sb[1])
        parent = (hyp + torch.arange(B).view(-1, 1) * H).view(-1).cuda()
        y = torch.cat([y[parent], tok.view(-1, 1).cuda()], dim=1)
    if kind == "scale30":
        assert fixes > 0   # the log-space redo was exercised


def test_column_copy_is_the_transpose():
    from espnet_b200 import _lib

    lib = _lib.load()
    B, T, V = 2, 77, 101
    x = torch.randn(B, T, V, device="cuda")
    tc = lib.b200_ctc_prefix_column_stride(T)
    assert tc == 96
    xc = torch.full((B, V, tc), float("nan"), device="cuda")
    _lib.check(lib.b200_ctc_prefix_columns(x.data_ptr(), B, T, V, xc.data_ptr(), torch.cuda.current_stream().cuda_stream), "columns")
    assert torch.equal(xc[:, :, :T], x.transpose(1, 2))
    assert torch.isnan(xc[:, :, T:]).all()              # the pad is never written (and never read)


@pytest.mark.parametrize("margin", [0, 3])
def test_hypotheses_without_mass_are_not_queued_for_the_exact_kernel(margin):
    """A prefix that is longer than its utterance has frames (or whose window holds none of its mass) scores logzero
    for every token; such hypotheses used to send all V tokens to the exact (log-space) kernel, which returned the same
    logzero -- 724 us per position late in a windowed search.  On unit-scale inputs hardly anything else underflows:
    the queue never holds as much as one row."""
    from espnet_b200.ctc_prefix_score import CTCPrefixScoreB200

    B, T, V, H = 2, 12, 64, 4
    c = _cases.make_prefix_case(B, T, V, H, seed=77, nsteps=11)
    if margin:
        # (a window that STRADDLES the end of a shorter utterance still queues that utterance's tokens for a few
        # positions -- correct, only slower; the windowed case is run without padding)
        c["xlens"][:] = T
    impl = CTCPrefixScoreB200(c["lp"].clone().cuda(), c["xlens"], 0, c["eos"], margin=margin)
    g = torch.Generator().manual_seed(3)
    y = torch.full((B * H, 1), c["eos"], dtype=torch.int64, device="cuda")
    state = None
    empty_rows = 0
    for step in range(11):
        att_w = None
        if margin:
            att_w = torch.zeros(B * H, T)
            att_w[:, min(T - 1, step)] = 1.0
            att_w = att_w.cuda()
        scores, st = impl(y, state, None, att_w)
        queued = int(impl._work.view(torch.int32)[2 * B * T * ((H + 3) & ~3) + 3 * B * H])
        assert queued < V - 2, (step, queued)   # never a whole row (a few pairs of the padded frames do underflow)
        empty_rows += int((st[1][:, 1:V - 1] < -1e9).all(1).sum())
        hyp = torch.randint(0, H, (B, H), generator=g)
        tok = torch.full((B, H), 5)                         # "5 5 5 ...": L labels need 2L - 1 frames
        state = impl.index_select_state(st, (hyp * V + tok).cuda())
        parent = (hyp + torch.arange(B).view(-1, 1) * H).view(-1).cuda()
        y = torch.cat([y[parent], tok.view(-1, 1).cuda()], dim=1)
    assert empty_rows > 0   # the walk did reach hypotheses without mass
