"""CPU emulation of the fp64 scaled-linear lattice sweep (developer tool): numerics and the scaling / certificate
policy are settled here before the CUDA kernel is written.

    python tools/sim_fp64_sweep.py [kind ...]

For each input distribution of tests/_cases.KINDS: alpha / beta in float64 LINEAR domain with one global power-of-two
scale per 8-frame chunk (chosen from level measurements that arrive `lag` chunks late, as the skewed warps deliver
them), lattice stored as float32 log2 of the scaled values, occupancies as the patch kernel computes them (fp32),
gradient compared with torch.nn.CTCLoss in float64.  Reports the error, the smallest flush margin, and the
certificate  -1074 - G_alpha(t) - G_beta(t) + max(m_alpha, m_beta) - log2 Z  that bounds what flushed states can
have contributed.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from tests import _cases  # noqa: E402

CHF = 8
LOG2E = np.float32(1.4426950408889634)


def exp2_f64(lp32):
    """2^lp for float32 lp (<= 0, may be hugely negative) as float64: fractional part through an fp32 exp2
    (ex2.approx: ~2^-22 relative), integer part through the exponent field."""
    lp = lp32.astype(np.float64)
    e = np.floor(lp)
    f = (lp - e).astype(np.float32)
    m = np.exp2(f).astype(np.float32).astype(np.float64)
    out = np.where(e < -1100, 0.0, np.ldexp(m, np.maximum(e, -1100).astype(np.int64)))
    return out


def sweep(lp2, lab, lag):
    """Per-WARP block scaling: positions are cut into blocks of 32 (one warp each); block w keeps its own cumulative
    power-of-two scale Gw (true = stored * 2^-(Gff + Gw)); Gff is the per-frame feed-forward scale (the largest
    needed probability of a frame becomes ~1).  A block measures its level (own states + the boundary values that
    will flow in during the chunk) at the start of a chunk and applies the correction `lag` chunks later.
    Returns lattice (T, 2n) float32 log2 of the stored values with the block scales REMOVED (log2 true + Gff), the
    cumulative Gff (T,), log2 Z, the smallest flush margin, and the per-frame max over blocks of Gw."""
    T, K = lp2.shape
    L = K - 1
    n = L + 1
    nw = (n + 31) // 32
    skip = np.zeros(n)
    skip[1:L] = (lab[1:] != lab[:-1]).astype(np.float64)
    ff = -np.floor(lp2.max(1)).astype(np.int64)
    ff[0] = 0
    p = np.ldexp(exp2_f64(lp2), np.minimum(ff, 1100)[:, None].astype(np.int64))   # feed-forward folded into p
    Gff = np.cumsum(ff)
    ab = np.zeros(n)
    al = np.zeros(n)
    ab[0] = p[0, 0]
    if L > 0:
        al[0] = p[0, 1]
    lat = np.full((T, 2 * n), -3.0e38, dtype=np.float64)
    Gw = np.zeros(nw, dtype=np.int64)
    Gw_t = np.zeros((T, nw), dtype=np.int64)
    pend = [[] for _ in range(nw)]           # corrections decided but not yet applied (lag)
    hist = [[] for _ in range(nw)]
    min_margin = 1e9

    def store(t, w):
        lo, hi = 32 * w, min(32 * w + 32, n)
        with np.errstate(divide="ignore"):
            lat[t, 2 * lo:2 * hi:2] = np.where(ab[lo:hi] > 0, np.log2(np.ldexp(np.floor(np.ldexp(np.frexp(ab[lo:hi])[0], 21)), -21)) + np.frexp(ab[lo:hi])[1] - Gw[w], -3.0e38)
            lat[t, 2 * lo + 1:2 * hi:2] = np.where(al[lo:hi] > 0, np.log2(np.ldexp(np.floor(np.ldexp(np.frexp(al[lo:hi])[0], 21)), -21)) + np.frexp(al[lo:hi])[1] - Gw[w], -3.0e38)
        Gw_t[t, w] = Gw[w]

    for w in range(nw):
        store(0, w)
    nchunk = (T + CHF - 1) // CHF
    # boundary values (label state of the last position of block w at frame t) in block w's scale, with that scale
    bval = np.zeros((nw, T))
    bG = np.zeros((nw, T), dtype=np.int64)
    for w in range(nw):
        bval[w, 0] = al[min(32 * w + 31, n - 1)] if 32 * w + 31 < n else 0.0
    for c in range(nchunk):
        t0, t1 = max(c * CHF, 1), min(c * CHF + CHF, T)
        for w in range(nw):
            lo, hi = 32 * w, min(32 * w + 32, n)
            # measure (integer exponents): own states and the incoming boundary values of this chunk in my scale
            def expo(v):
                v = v[v > 0]
                return np.frexp(v)[1].astype(np.int64) - 1 if v.size else np.zeros(0, dtype=np.int64)
            lev = expo(np.concatenate((ab[lo:hi], al[lo:hi])))
            if w > 0 and t1 > t0:
                bv = bval[w - 1, t0 - 1:t1 - 1]
                e_in = (np.frexp(bv[bv > 0])[1].astype(np.int64) - 1) + (Gw[w] - bG[w - 1, t0 - 1:t1 - 1][bv > 0])
                lev = np.concatenate((lev, e_in))
            g = -int(lev.max()) if lev.size else 0
            Gw[w] += g
            if g:
                ab[lo:hi] = np.ldexp(ab[lo:hi], g)
                al[lo:hi] = np.ldexp(al[lo:hi], g)
            for t in range(t0, t1):
                nb_in = 0.0
                if w > 0:
                    nb_in = np.ldexp(bval[w - 1, t - 1], int(np.clip(Gw[w] - bG[w - 1, t - 1], -2000, 2000)))
                nbr = np.concatenate(([nb_in], al[lo:hi - 1]))
                nab = p[t, 0] * (ab[lo:hi] + nbr)
                pl = np.zeros(hi - lo)
                m = min(hi, L) - lo
                if m > 0:
                    pl[:m] = p[t, 1 + lo:1 + lo + m]
                nal = pl * (al[lo:hi] + ab[lo:hi] + skip[lo:hi] * nbr)
                ab[lo:hi], al[lo:hi] = nab, nal
                assert np.isfinite(nab).all() and np.isfinite(nal).all(), "overflow"
                bval[w, t] = al[hi - 1] if hi - lo == 32 else 0.0
                bG[w, t] = Gw[w]
                store(t, w)
                mxs = max(ab[lo:hi].max(), al[lo:hi].max())
                if mxs > 0:
                    min_margin = min(min_margin, np.log2(mxs) + 1074)
    wl, wl1 = L // 32, max(L - 1, 0) // 32
    with np.errstate(divide="ignore"):
        l1 = np.log2(ab[L]) - Gw[wl] if ab[L] > 0 else -np.inf
        l2_ = (np.log2(al[L - 1]) - Gw[wl1]) if (L > 0 and al[L - 1] > 0) else -np.inf
    m_ = max(l1, l2_)
    log2z = (m_ + np.log2(2.0 ** (l1 - m_) + 2.0 ** (l2_ - m_)) - Gff[-1]) if np.isfinite(m_) else -np.inf
    return lat, Gff, log2z, min_margin, Gw_t


def run_case(kind, B=4, T=500, V=200, L=150, seed=7, lag=6):
    c = _cases.make_loss_case_kind(kind, B, T, V, L, seed=seed)
    x, ilen, olen, ys = c["x"], c["ilen"], c["olen"], c["ys"]
    xr = x.double().transpose(0, 1).detach().requires_grad_()
    tg = torch.cat([ys[i, : olen[i]] for i in range(B)])
    nll = torch.nn.functional.ctc_loss(xr.log_softmax(2), tg, ilen, olen, reduction="none")
    (nll.sum() / B).backward()
    gref = xr.grad.transpose(0, 1).numpy()
    worst = dict(kind=kind, grad_rel=0.0, nll_rel=0.0, min_margin=1e9, cert_max=-1e9)
    for b in range(B):
        Tb, Lb = int(ilen[b]), int(olen[b])
        lab = ys[b, :Lb].numpy()
        xb = x[b, :Tb].numpy().astype(np.float32)
        # the stream kernel: row log2-sum-exp in fp32, lpg = x*log2e - lse2
        mx = xb.max(1, keepdims=True)
        ml = mx * LOG2E
        ssum = np.exp2((xb * LOG2E - ml).astype(np.float32)).astype(np.float32).sum(1, keepdims=True, dtype=np.float32)
        l2 = (ml + np.log2(ssum)).astype(np.float32)
        cols = np.concatenate(([0], lab))
        lp2 = np.maximum((xb[:, cols] * LOG2E - l2).astype(np.float32), np.float32(-1e30))
        la, Ga, za, m1, Gwa = sweep(lp2, lab, lag)
        lpr = lp2[::-1][:, np.concatenate(([0], np.arange(Lb, 0, -1)))]     # time- and label-reversed
        lb_, Gb, zb, m2, Gwb = sweep(lpr, lab[::-1], lag)
        lb_ = lb_[::-1]
        Gb = (Gb[-1] - np.concatenate(([0], Gb[:-1])))[::-1] * 0 + Gb[::-1]
        Gwb = Gwb[::-1]
        nll_b = -za * np.log(2.0)
        # occupancies as the patch kernel: un-normalised log2 occupancy per state, normalised per row in fp32
        n = Lb + 1
        S = 2 * Lb + 1
        a_bl, a_lb = la[:, 0::2], la[:, 1::2]
        # beta is stored in reversed state order: blank k <-> blank (L-k), label k <-> label (L-1-k)
        b_bl = lb_[:, 0::2][:, ::-1]
        b_lb = np.concatenate((lb_[:, 1::2][:, :Lb][:, ::-1], np.full((Tb, 1), -3e38, np.float32)), axis=1)
        ub = (a_bl + b_bl - lp2[:, :1])
        ul = np.full((Tb, n), -3e38, dtype=np.float64)
        ul[:, :Lb] = (a_lb[:, :Lb] + b_lb[:, :Lb] - lp2[:, 1:])
        m = np.maximum(ub.max(1), ul.max(1))[:, None]
        eb = np.exp2(np.maximum(ub - m, -200).astype(np.float32)).astype(np.float32) * (ub > -1e38)
        el = np.exp2(np.maximum(ul - m, -200).astype(np.float32)).astype(np.float32) * (ul > -1e38)
        tot = eb.sum(1) + el.sum(1)
        gam_b = eb.sum(1) / tot
        gam_l = el / tot[:, None]
        sm = np.exp2((xb * LOG2E - l2).astype(np.float32)).astype(np.float64)
        g = sm.copy()
        g[:, 0] -= gam_b
        for k in range(Lb):
            g[:, lab[k]] -= gam_l[:, k]
        g /= B
        ref = gref[b, :Tb]
        worst["grad_rel"] = max(worst["grad_rel"], np.abs(g - ref).max() / max(np.abs(gref).max(), 1.0 / B))
        worst["nll_rel"] = max(worst["nll_rel"], abs(nll_b - float(nll[b])) / max(abs(float(nll[b])), 1e-6))
        worst["min_margin"] = min(worst["min_margin"], m1, m2)
        # certificate (log2 units)
        ma = np.maximum((a_bl - lp2[:, :1]).max(1), (a_lb[:, :Lb] - lp2[:, 1:]).max(1) if Lb else -3e38)
        mb = np.maximum((b_bl - lp2[:, :1]).max(1), (b_lb[:, :Lb] - lp2[:, 1:]).max(1) if Lb else -3e38)
        cert = -1074.0 - Ga - Gb - np.minimum(Gwa.min(1), Gwb.min(1)) + np.maximum(ma, mb) - za
        worst["cert_max"] = max(worst["cert_max"], float(cert.max()))
        worst["zdiff"] = abs(za - zb)
    return worst


if __name__ == "__main__":
    kinds = sys.argv[1:] or _cases.KINDS
    for k in kinds:
        for lag in (0,):
            print(run_case(k, lag=lag), "lag", lag, flush=True)
