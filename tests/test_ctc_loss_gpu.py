"""Parity of the CUDA CTC loss path (through the C ABI) with the reference.

Bars (BASELINE.json north_star): loss and gradients within 1e-5 relative in fp32 and 1e-2 for
bf16 logits.  "Relative" for the gradient is max|g - g_ref| / max|g_ref| per batch.  The
reference value is the reference's own CTC.loss_fn evaluated in float64 (recorded in
tests/golden, or torch.nn.CTCLoss in float64 at sizes without a fixture); the reference's
float32 path is itself only ~3e-4 away from that (measured below), so it is compared at the
looser of 1e-5 and its own float32-vs-float64 gap.
"""
import numpy as np
import pytest
import torch

from tests import _cases
from tests._util import golden

pytestmark = pytest.mark.gpu

TOL_F32 = 1e-5
TOL_BF16 = 1e-2


@pytest.fixture(autouse=True, params=["lin", "log"])
def sweep_domain(request):
    """Every test of this file runs twice: with the log-domain lattice sweeps (the default) and with the float64
    linear-domain sweeps (utterances they cannot certify are redone in the log domain)."""
    from espnet_b200 import _lib

    lib = _lib.load()
    was = lib.b200_ctc_debug_force_log_sweep(1 if request.param == "log" else 0)
    # (long label sequences, L > 255, have no linear-domain sweep: their second run takes the one-CTA sweep kernel
    #  instead of the thread-block-cluster one)
    lib.b200_ctc_debug_single_cta_sweep(1 if request.param == "log" else 0)
    yield request.param
    lib.b200_ctc_debug_force_log_sweep(was)
    lib.b200_ctc_debug_single_cta_sweep(0)


def _run(x_btv, ilen, olen, ys, grad_scale=None, dtype=None, strided=True, split=False):
    """Returns (nll (B,) float64 numpy, grad (B,T,V) float64 numpy, loss_sum).

    split=False: the fused b200_ctc_loss_fwd_bwd call (one pass over the logits when the rows are
    16-byte aligned); split=True: b200_ctc_loss_forward, then b200_ctc_loss_backward on the same
    workspace (the two-read path)."""
    from espnet_b200 import ctc_loss as CL

    B, T, V = x_btv.shape
    xd = x_btv.cuda() if dtype is None else x_btv.to(dtype).cuda()
    xt = xd.transpose(0, 1) if strided else xd.transpose(0, 1).contiguous()
    L = ys.shape[1]
    off = torch.arange(B) * L
    if split:
        nll, grad = CL.ctc_loss_split(xt, ys.cuda().reshape(-1), off, ilen, olen, L,
                                      grad_scale=(1.0 / B if grad_scale is None else grad_scale))
        torch.cuda.synchronize()
        return nll.double().cpu().numpy(), grad.transpose(0, 1).double().cpu().numpy(), float(nll.sum().item())
    nll, ls, grad = CL.ctc_loss_fwd_bwd(xt, ys.cuda().reshape(-1), off, ilen, olen, L,
                                        grad_scale=(1.0 / B if grad_scale is None else grad_scale))
    torch.cuda.synchronize()
    return nll.double().cpu().numpy(), grad.transpose(0, 1).double().cpu().numpy(), float(ls.item())


def _ref64(x_btv, ilen, olen, ys, dt=torch.float64, device="cuda"):
    """The reference formula (espnet2/asr/ctc.py:46-52) in float64."""
    B = x_btv.shape[0]
    xr = x_btv.to(device).to(dt).transpose(0, 1).detach().requires_grad_()
    tg = torch.cat([ys[i, : olen[i]] for i in range(B)]).to(device)
    nll = torch.nn.functional.ctc_loss(xr.log_softmax(2), tg, ilen, olen, reduction="none")
    (nll.sum() / B).backward()
    return nll.detach().double().cpu().numpy(), xr.grad.transpose(0, 1).double().cpu().numpy()


def _normrel(a, b):
    ok = np.isfinite(b)
    return np.abs(a[ok] - b[ok]).max() / max(np.abs(b[ok]).max(), 1e-30)


@pytest.mark.parametrize("split", [False, True])
@pytest.mark.parametrize("name", ["c1", "tiny", "v64", "repeats"])
def test_fp32_matches_golden_reference(name, split):
    g = golden(f"loss_{name}.npz")
    x, ilen, olen, ys = (torch.from_numpy(g[k]) for k in ("x", "ilen", "olen", "ys"))
    B = x.shape[0]
    nll, grad, ls = _run(x, ilen, olen, ys, split=split)
    ref_nll = g["nll_over_B_f64"] * B
    np.testing.assert_allclose(nll, ref_nll, rtol=TOL_F32)
    np.testing.assert_allclose(ls / B, g["loss_f64"], rtol=TOL_F32)
    assert _normrel(grad, g["grad_f64"]) <= TOL_F32
    # against the reference's float32 path: bounded by that path's own error
    gap = _normrel(g["grad_f32"].astype(np.float64), g["grad_f64"])
    assert _normrel(grad, g["grad_f32"].astype(np.float64)) <= max(TOL_F32, 2 * gap)


@pytest.mark.parametrize("split,V4", [(False, False), (True, False), (False, True)])
def test_edge_cases_match_reference(split, V4):
    """Empty target, single-frame utterance, infeasible utterances: inf loss and NaN gradient
    exactly where torch.nn.CTCLoss(zero_infinity=False) puts them.  V4 pads the vocabulary with
    three -inf columns (V=8: 16-byte rows) so that the one-pass kernels see the same problem."""
    g = golden("loss_edge.npz")
    x, ilen, olen, ys = (torch.from_numpy(g[k]) for k in ("x", "ilen", "olen", "ys"))
    B = x.shape[0]
    if V4:
        x = torch.cat([x, torch.full(x.shape[:2] + (3,), -float("inf"))], dim=2)
    nll, grad, _ = _run(x, ilen, olen, ys, split=split)
    if V4:
        assert (grad[..., 5:][~np.isnan(grad[..., 5:])] == 0).all()
        grad = grad[..., :5]
    ref_nll = g["nll_over_B_f64"] * B
    fin = np.isfinite(ref_nll)
    assert (np.isposinf(nll) == ~fin).all() and (~fin).sum() >= 1 and fin.sum() >= 2
    np.testing.assert_allclose(nll[fin], ref_nll[fin], rtol=TOL_F32)
    ref = g["grad_f64"]
    assert (np.isnan(grad) == np.isnan(ref)).all()
    assert _normrel(grad, ref) <= TOL_F32


@pytest.mark.parametrize("dtype,tol", [(torch.bfloat16, TOL_BF16), (torch.float16, 2e-3)])
def test_half_precision_logits(dtype, tol):
    c = _cases.make_loss_case(6, 120, 256, 30, seed=77, dtype=dtype)
    x = c["x"]
    nll, grad, _ = _run(x, c["ilen"], c["olen"], c["ys"])
    nll2, grad2, _ = _run(x, c["ilen"], c["olen"], c["ys"], split=True)
    np.testing.assert_allclose(nll2, nll, rtol=1e-6)
    assert _normrel(grad2, grad) <= tol
    rn, rg = _ref64(x.float(), c["ilen"], c["olen"], c["ys"])
    np.testing.assert_allclose(nll, rn, rtol=TOL_F32)  # fp32 accumulation: loss stays tight
    assert _normrel(grad, rg) <= tol


@pytest.mark.parametrize("B,T,V,L", [(5, 33, 7, 6), (2, 64, 1000, 20), (3, 40, 129, 12), (2, 300, 48, 170)])
def test_odd_shapes_vs_oracle(B, T, V, L):
    """Unaligned vocabularies (scalar path), L+1 > 160 (multi-batch gather), vs the C oracle."""
    from oracle import ctc_oracle as O

    c = _cases.make_loss_case(B, T, V, L, seed=100 + V)
    x = c["x"]
    nll, grad, _ = _run(x, c["ilen"], c["olen"], c["ys"])
    onll, ograd = O.ctc_loss(np.ascontiguousarray(x.numpy().transpose(1, 0, 2)), c["ys"].numpy().reshape(-1),
                             np.arange(B) * c["ys"].shape[1], c["ilen"].numpy(), c["olen"].numpy(),
                             grad_scale=1.0 / B)
    np.testing.assert_allclose(nll, onll, rtol=TOL_F32)
    assert _normrel(grad, ograd.transpose(1, 0, 2)) <= TOL_F32


def test_contiguous_and_strided_views_agree_bitwise():
    c = _cases.make_loss_case(4, 60, 64, 10, seed=5)
    a = _run(c["x"], c["ilen"], c["olen"], c["ys"], strided=True)
    b = _run(c["x"], c["ilen"], c["olen"], c["ys"], strided=False)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()


def test_deterministic_bitwise():
    c = _cases.make_loss_case(8, 100, 200, 25, seed=8, dtype=torch.bfloat16)
    a = _run(c["x"], c["ilen"], c["olen"], c["ys"])
    b = _run(c["x"], c["ilen"], c["olen"], c["ys"])
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()


@pytest.mark.parametrize("reduce", [True, False])
def test_module_forward_backward_like_reference(reduce):
    """espnet2 CTC.forward with ctc_type="b200" vs the same module with "builtin" (shared
    weights), loss and d loss / d hs_pad, through autograd (test/espnet2/asr/test_ctc.py)."""
    from espnet_b200.ctc import CTC

    torch.manual_seed(0)
    B, T, D, V = 4, 30, 16, 11
    a = CTC(V, D, ctc_type="b200", reduce=reduce).cuda()
    r = CTC(V, D, ctc_type="builtin", reduce=reduce).cuda().double()
    r.load_state_dict({k: v.double() for k, v in a.state_dict().items()})
    hs = torch.randn(B, T, D, device="cuda")
    hlens = torch.tensor([30, 22, 25, 30], device="cuda")
    ys_pad = torch.randint(1, V, (B, 6), device="cuda")
    ys_lens = torch.tensor([6, 3, 5, 1], device="cuda")
    h1 = hs.clone().requires_grad_()
    h2 = hs.double().requires_grad_()
    l1 = a(h1, hlens, ys_pad, ys_lens)
    l2 = r(h2, hlens, ys_pad, ys_lens)
    w = torch.linspace(0.5, 1.5, B, device="cuda")
    (l1.sum() if reduce else (l1 * w).sum()).backward()
    (l2.sum() if reduce else (l2 * w.double()).sum()).backward()
    np.testing.assert_allclose(l1.detach().double().cpu().numpy(), l2.detach().cpu().numpy(), rtol=1e-5)
    assert _normrel(h1.grad.double().cpu().numpy(), h2.grad.cpu().numpy()) <= 2e-5
    ga = a.ctc_lo.weight.grad.double().cpu().numpy()
    assert _normrel(ga, r.ctc_lo.weight.grad.cpu().numpy()) <= 2e-5


class _Spy(object):
    """Counts the C-ABI entry points a block of code calls."""

    def __init__(self, lib):
        self._lib, self.calls = lib, []

    def __getattr__(self, name):
        fn = getattr(self._lib, name)
        if name.startswith("b200_ctc_loss") or name.startswith("b200_ctc_grad"):
            def wrapped(*a, **k):
                self.calls.append(name)
                return fn(*a, **k)
            return wrapped
        return fn


def test_module_call_runs_the_one_pass_kernels(monkeypatch):
    """CTC(ctc_type="b200").forward + backward = ONE b200_ctc_loss_fwd_bwd (logits read once) and one
    b200_ctc_grad_rescale that finds nothing to do -- not the forward / backward pair that reads the logits twice."""
    from espnet_b200 import _lib
    from espnet_b200.ctc import CTC

    spy = _Spy(_lib.load())
    monkeypatch.setattr(_lib, "_lib", spy)
    torch.manual_seed(0)
    ctc = CTC(16, 8, ctc_type="b200").cuda()
    hs = torch.randn(3, 20, 8, device="cuda", requires_grad=True)
    loss = ctc(hs, torch.tensor([20, 15, 18]), torch.randint(1, 16, (3, 4), device="cuda"), torch.tensor([4, 2, 3]))
    loss.backward()
    torch.cuda.synchronize()
    assert [c for c in spy.calls if c != "b200_ctc_loss_workspace_bytes"] == ["b200_ctc_loss_fwd_bwd", "b200_ctc_grad_rescale"]
    spy.calls.clear()
    with torch.no_grad():      # evaluation: no gradient is produced
        ctc(hs, torch.tensor([20, 15, 18]), torch.randint(1, 16, (3, 4), device="cuda"), torch.tensor([4, 2, 3]))
    assert "b200_ctc_loss_fwd_bwd" not in spy.calls and "b200_ctc_loss_forward" in spy.calls


@pytest.mark.parametrize("hint", [1.0, 0.3])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_upstream_gradient_other_than_the_hint(hint, dtype):
    """loss_fn + backward with upstream gradients that differ from grad_scale_hint (ctc_weight, accum_grad, an AMP
    loss scale: espnet2/train/trainer.py:394-412): the rescale kernel must reproduce autograd's result."""
    from espnet_b200.ctc import CTC

    c = _cases.make_loss_case(5, 70, 96, 12, seed=21, dtype=dtype)
    x, ilen, olen, ys = c["x"], c["ilen"], c["olen"], c["ys"]
    B = x.shape[0]
    tg = torch.cat([ys[i, : olen[i]] for i in range(B)])
    a = CTC(96, 8, ctc_type="b200", grad_scale_hint=hint)
    r = CTC(96, 8, ctc_type="builtin")
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    for scale in (1.0, 0.3, 1024.0 * 0.3 / 4):
        xa = x.cuda().transpose(0, 1).detach().requires_grad_()
        xr = x.cuda().double().transpose(0, 1).detach().requires_grad_()
        la = a.loss_fn(xa, tg.cuda(), ilen, olen)          # lengths on the host: no synchronisation
        lr = r.loss_fn(xr, tg.cuda(), ilen, olen)
        (la * scale).backward()
        (lr * scale).backward()
        np.testing.assert_allclose(la.item(), lr.item(), rtol=1e-5)
        assert _normrel(xa.grad.double().cpu().numpy(), xr.grad.cpu().numpy()) <= tol


def test_reduce_false_per_utterance_upstream_gradient():
    from espnet_b200.ctc import CTC

    c = _cases.make_loss_case(4, 50, 64, 8, seed=22)
    x, ilen, olen, ys = c["x"], c["ilen"], c["olen"], c["ys"]
    tg = torch.cat([ys[i, : olen[i]] for i in range(4)])
    a, r = CTC(64, 8, ctc_type="b200", reduce=False), CTC(64, 8, ctc_type="builtin", reduce=False)
    xa = x.cuda().transpose(0, 1).detach().requires_grad_()
    xr = x.cuda().double().transpose(0, 1).detach().requires_grad_()
    w = torch.tensor([0.5, 2.0, 0.0, 1.0], device="cuda")
    (a.loss_fn(xa, tg.cuda(), ilen.cuda(), olen.cuda()) * w).sum().backward()
    (r.loss_fn(xr, tg.cuda(), ilen.cuda(), olen.cuda()) * w.double()).sum().backward()
    assert _normrel(xa.grad.double().cpu().numpy(), xr.grad.cpu().numpy()) <= 1e-5
    assert (xa.grad[:, 2] == 0).all()


def test_invalid_targets_are_loud():
    """A label outside [0,V), a blank inside the targets, olen > lmax: torch.nn.CTCLoss raises on the CPU
    (espnet2/asr/ctc.py:49); here such utterances get nll = NaN and NaN gradient rows, the others are untouched,
    and strict mode raises."""
    from espnet_b200 import ctc_loss as CL

    c = _cases.make_loss_case(4, 40, 32, 6, seed=23)
    x, ilen, olen, ys = c["x"], c["ilen"], c["olen"], c["ys"].clone()
    good = _run(x, ilen, olen, ys)
    ys[1, 0] = 32        # out of range
    ys[2, 1] = 0         # blank as a label
    nll, grad, _ = _run(x, ilen, olen, ys)
    assert np.isnan(nll[[1, 2]]).all() and np.isfinite(nll[[0, 3]]).all()
    assert (nll[[0, 3]] == good[0][[0, 3]]).all() and (grad[[0, 3]] == good[1][[0, 3]]).all()
    for b in (1, 2):
        assert np.isnan(grad[b, : int(ilen[b])]).all() and (grad[b, int(ilen[b]):] == 0).all()
    xt = x.cuda().transpose(0, 1).requires_grad_()
    off = torch.arange(4) * ys.shape[1]
    with pytest.raises(ValueError, match="invalid CTC targets"):
        CL.ctc_nll(xt, ys.cuda().reshape(-1), off, ilen, olen, ys.shape[1], strict=True)
    # olen beyond the lmax the workspace was sized for: clamped and flagged, nothing is written out of bounds
    ys = c["ys"]
    too_long = olen.clone()
    too_long[0] = 9
    nll = CL.ctc_nll(xt, ys.cuda().reshape(-1), off, ilen, too_long, 6)
    assert torch.isnan(nll[0]) and torch.isfinite(nll[1:]).all()
    with pytest.raises(ValueError):
        CL.ctc_nll(xt, ys.cuda().reshape(-1), off, ilen, too_long, 6, strict=True)
    CL.ctc_nll(xt, ys.cuda().reshape(-1), off, ilen, olen, 6, strict=True)   # valid input passes


def test_espnet1_module_gradients_match_builtin():
    from espnet_b200.ctc import CTCEspnet1

    torch.manual_seed(2)
    B, T, D, V = 3, 25, 8, 9
    a = CTCEspnet1(V, D, 0.0, ctc_type="b200").cuda()
    r = CTCEspnet1(V, D, 0.0, ctc_type="builtin").cuda().double()
    r.load_state_dict({k: v.double() for k, v in a.state_dict().items()})
    hs = torch.randn(B, T, D, device="cuda")
    ys_pad = torch.tensor([[1, 2, 3, -1], [4, 4, -1, -1], [5, 6, 7, 8]], device="cuda")
    h1, h2 = hs.clone().requires_grad_(), hs.double().requires_grad_()
    a(h1, [25, 20, 18], ys_pad).backward()
    r(h2, [25, 20, 18], ys_pad).backward()
    assert _normrel(h1.grad.double().cpu().numpy(), h2.grad.cpu().numpy()) <= 2e-5
    assert _normrel(a.ctc_lo.weight.grad.double().cpu().numpy(), r.ctc_lo.weight.grad.cpu().numpy()) <= 2e-5


def test_forced_align_builtin_and_b200_agree_with_reference_fixture():
    from espnet_b200.ctc import CTCEspnet1

    g = golden("viterbi.npz")
    for k in range(3):
        h, y, want = torch.from_numpy(g[f"h_{k}"]), torch.from_numpy(g[f"y_{k}"]), g[f"align_{k}"].tolist()
        V = h.shape[1]
        for kind, dev in (("b200", "cuda"), ("builtin", "cuda"), ("builtin", "cpu")):
            ctc = CTCEspnet1(V, V, 0.0, ctc_type=kind).to(dev)
            with torch.no_grad():
                ctc.ctc_lo.weight.copy_(torch.eye(V))
                ctc.ctc_lo.bias.zero_()
                assert [int(t) for t in ctc.forced_align(h.to(dev).unsqueeze(0), y)] == want


def test_espnet1_module_matches_builtin():
    from espnet_b200.ctc import CTCEspnet1

    torch.manual_seed(1)
    B, T, D, V = 3, 25, 8, 9
    a = CTCEspnet1(V, D, 0.0, ctc_type="b200").cuda()
    r = CTCEspnet1(V, D, 0.0, ctc_type="builtin").cuda()
    r.load_state_dict(a.state_dict())
    hs = torch.randn(B, T, D, device="cuda")
    ys_pad = torch.tensor([[1, 2, 3, -1], [4, 4, -1, -1], [5, 6, 7, 8]], device="cuda")
    la = a(hs, [25, 20, 18], ys_pad)
    lr = r(hs, [25, 20, 18], ys_pad)
    np.testing.assert_allclose(la.item(), lr.item(), rtol=1e-5)


def test_c2_full_size_bf16_properties_and_reference():
    """BASELINE config 2 (B=64, T=500, V=5000, L<=150, bf16): loss vs torch float64 on the
    GPU, gradient vs the same at the bf16 bar, plus size-independent properties."""
    c = _cases.make_loss_case(64, 500, 5000, 150, seed=1236, dtype=torch.bfloat16)
    x, ilen, olen, ys = c["x"], c["ilen"], c["olen"], c["ys"]
    nll, grad, ls = _run(x, ilen, olen, ys)
    rn, rg = _ref64(x.float(), ilen, olen, ys)
    np.testing.assert_allclose(nll, rn, rtol=TOL_F32)
    np.testing.assert_allclose(ls, rn.sum(), rtol=TOL_F32)
    assert _normrel(grad, rg) <= TOL_BF16
    # frames past ilen get exactly zero; inside, every row of the gradient sums to ~0
    for b in (0, 7, 63):
        assert (grad[b, int(ilen[b]):] == 0).all()
    rows = np.abs(grad.sum(-1))
    assert rows.max() <= 64 * 2e-3 / 64  # bf16 rounding of 5000 terms of size <= 1/B


def test_c2_full_size_fp32_gradient():
    c = _cases.make_loss_case(64, 500, 5000, 150, seed=1236)
    x, ilen, olen, ys = c["x"], c["ilen"], c["olen"], c["ys"]
    nll, grad, _ = _run(x, ilen, olen, ys)
    rn, rg = _ref64(x, ilen, olen, ys)
    np.testing.assert_allclose(nll, rn, rtol=TOL_F32)
    assert _normrel(grad, rg) <= 2e-5  # measured 9e-6..1.6e-5 at this size; torch fp32 is 3e-3


def test_c5_long_form_reduced_batch():
    """BASELINE config 5 shape (T=4000, V=5000, L<=1000) at B=2 to bound test time."""
    c = _cases.make_loss_case(2, 4000, 5000, 1000, seed=1239, dtype=torch.bfloat16)
    x, ilen, olen, ys = c["x"], c["ilen"], c["olen"], c["ys"]
    nll, grad, _ = _run(x, ilen, olen, ys)
    rn, rg = _ref64(x.float(), ilen, olen, ys)
    np.testing.assert_allclose(nll, rn, rtol=TOL_F32)
    assert _normrel(grad, rg) <= TOL_BF16


def test_upload_valid_frames_gives_the_same_result_as_a_full_copy():
    """The kernels never read frames past ilen: a device buffer that only received the valid frames
    (padding left as NaN) gives bitwise the results of a full host->device copy."""
    from espnet_b200 import ctc_loss as CL

    c = _cases.make_loss_case(6, 90, 256, 20, seed=321, dtype=torch.bfloat16)
    x, ilen, olen, ys = c["x"], c["ilen"], c["olen"], c["ys"]
    B, T, V = x.shape
    host = x.pin_memory()
    dev = torch.full_like(x, float("nan"), device="cuda")
    nbytes = CL.upload_valid_frames(host, dev, ilen)
    assert nbytes == int(ilen.clamp(max=T).sum()) * V * 2
    for b in range(B):
        n = int(ilen[b])
        assert torch.equal(dev[b, :n].cpu(), x[b, :n])
        assert torch.isnan(dev[b, n:].float()).all()
    L = ys.shape[1]
    off = torch.arange(B) * L
    a = CL.ctc_loss_fwd_bwd(dev.transpose(0, 1), ys.cuda().reshape(-1), off, ilen, olen, L, grad_scale=1.0 / B)
    r = CL.ctc_loss_fwd_bwd(x.cuda().transpose(0, 1), ys.cuda().reshape(-1), off, ilen, olen, L, grad_scale=1.0 / B)
    torch.cuda.synchronize()
    assert torch.equal(a[0], r[0]) and torch.equal(a[1], r[1])
    assert torch.equal(a[2].view(torch.int16), r[2].view(torch.int16))
    with pytest.raises(RuntimeError):
        CL.upload_valid_frames(host, dev.float(), ilen)


def test_repeated_calls_are_bitwise_identical_at_c2_size():
    """The kernels of a call are chained by programmatic dependent launch and per-utterance completion
    flags (prep runs under the stream kernel, the patch kernel under the sweeps); a missed dependency would
    show up as run-to-run differences.  25 back-to-back calls on one workspace, ragged C2-shaped batch."""
    from espnet_b200 import ctc_loss as CL

    c = _cases.make_loss_case(64, 500, 5000, 150, seed=4321, dtype=torch.bfloat16)
    B, L = 64, c["ys"].shape[1]
    xt = c["x"].cuda().transpose(0, 1)
    plan = CL._Plan(xt, c["ys"].cuda().view(-1), torch.arange(B) * L, c["ilen"], c["olen"], L, 0)
    ref = None
    for it in range(25):
        grad = torch.full((B, 500, 5000), float("nan"), dtype=torch.bfloat16, device="cuda").transpose(0, 1)
        nll, ls, g = CL.ctc_loss_fwd_bwd(xt, None, None, None, None, L, grad_scale=1.0 / B, grad_out=grad, plan=plan)
        if it % 3 == 0:
            torch.cuda.synchronize()
        cur = (nll.clone(), ls.clone(), g.contiguous().view(torch.int16).clone())
        if ref is None:
            ref = cur
        else:
            assert torch.equal(cur[0], ref[0]) and torch.equal(cur[1], ref[1]), f"nll differs at call {it}"
            assert torch.equal(cur[2], ref[2]), f"gradient differs at call {it}"
    assert not torch.isnan(ref[2].view(torch.bfloat16).float()).any()


def test_many_utterances_more_ctas_than_sms():
    """B = 400 utterances: the sweep grid (800 CTAs), the prep grid running under the stream kernel and the
    ticketed patch kernel all need several waves on 148 SMs; completion flags and counters must still line up."""
    c = _cases.make_loss_case(400, 48, 64, 10, seed=77)
    x, ilen, olen, ys = c["x"], c["ilen"], c["olen"], c["ys"]
    for _ in range(2):
        nll, grad, ls = _run(x, ilen, olen, ys)
        rn, rg = _ref64(x, ilen, olen, ys)
        np.testing.assert_allclose(nll, rn, rtol=TOL_F32)
        np.testing.assert_allclose(ls, rn.sum(), rtol=TOL_F32)
        assert _normrel(grad, rg) <= TOL_F32


@pytest.mark.parametrize("L", [260, 330])
def test_many_long_utterances_more_clusters_than_sms(L):
    """B = 48 utterances with L > 255: 48 x 2 clusters of 4 CTAs (384 CTAs) do not fit on 148 SMs at once, so the
    cluster sweeps run in waves while the patch kernel (L = 260: warp per row; L = 330: CTA per row) is staged
    behind them and follows the completion flags."""
    c = _cases.make_loss_case(48, 2 * L + 30, 32, L, seed=78)
    x, ilen, olen, ys = c["x"], c["ilen"], c["olen"], c["ys"]
    for _ in range(2):
        nll, grad, ls = _run(x, ilen, olen, ys)
        rn, rg = _ref64(x, ilen, olen, ys)
        np.testing.assert_allclose(nll, rn, rtol=TOL_F32)
        np.testing.assert_allclose(ls, rn.sum(), rtol=TOL_F32)
        # (tight lattices, ilen down to 0.6 T against up to 260 / 330 labels: bar as in the hand-over test below)
        _, fg = _ref64(x, ilen, olen, ys, dt=torch.float32)
        e_g, e_ref = _normrel(grad, rg), _normrel(fg, rg)
        print(f"L={L}: grad error {e_g:.2e}, torch float32 {e_ref:.2e}")
        assert e_g <= max(TOL_F32, e_ref), (e_g, e_ref)


@pytest.mark.parametrize("split", [False, True])
@pytest.mark.parametrize("Lmax", [255, 256, 257, 320, 321, 512, 513, 769, 1023])
def test_long_label_sequences_at_the_kernel_boundaries(Lmax, split, sweep_domain):
    """Label lengths around the hand-over points of the long-lattice kernels (staged sweep <= 255 < cluster sweep: 256
    label positions per CTA of the cluster; register patch kernel <= 319 < CTA-per-row patch kernel), with, in ONE batch:
    an utterance at Lmax (its last blank and last label state in different CTAs when Lmax is a multiple of 256), a
    short one (one active CTA in its cluster), an empty target, an infeasible one (inf loss, NaN gradient rows),
    long runs of a repeated label (skip transitions forbidden, duplicates merged across threads of different warps) and
    an utterance shorter than one 8-frame chunk."""
    g = torch.Generator().manual_seed(5000 + Lmax)
    B, V = 6, 48
    T = 2 * Lmax + 60
    x = torch.randn(B, T, V, generator=g) * 2.0
    ys = torch.randint(1, V, (B, Lmax), generator=g)
    ys[4, : Lmax // 2] = 7                      # a long run of one label: needs 2 frames per repeat
    olen = torch.tensor([Lmax, 3, 0, Lmax, Lmax // 2 + 5, 2])
    ilen = torch.tensor([T, T - 17, 40, Lmax - 1, T, 5])   # utterance 3 is infeasible (fewer frames than labels)
    nll, grad, _ = _run(x, ilen, olen, ys, split=split)
    rn, rg = _ref64(x, ilen, olen, ys)
    fin = np.isfinite(rn)
    assert list(fin) == [True, True, True, False, True, True]
    assert (np.isposinf(nll) == ~fin).all()
    np.testing.assert_allclose(nll[fin], rn[fin], rtol=TOL_F32)
    assert (np.isnan(grad) == np.isnan(rg)).all()
    # The tight lattices of this batch (T barely above what the labels need, a forced run of repeats) keep the states
    # that matter hundreds of log2 units below the frame maximum the sweeps re-centre on: float32 resolves them to
    # ~1e-5, the reference's own float32 evaluation (no re-centring: |alpha| ~ 4 T) to less.  Bar, as in
    # test_parity_distributions_gpu._check_loss: 1e-5 or the error of torch's float32 CTC on the same input,
    # whichever is larger (2e-5 at least for the float64 linear-domain sweeps, Lmax <= 255, 20-bit mantissas).
    _, fg = _ref64(x, ilen, olen, ys, dt=torch.float32)
    e_g, e_ref = _normrel(grad, rg), _normrel(fg, rg)
    print(f"Lmax={Lmax} split={split} {sweep_domain}: grad error {e_g:.2e}, torch float32 {e_ref:.2e}")
    assert e_g <= max(2e-5 if sweep_domain == "lin" and Lmax <= 255 else TOL_F32, e_ref), (e_g, e_ref)
    for b in range(B):
        assert (grad[b, int(ilen[b]):] == 0).all()
