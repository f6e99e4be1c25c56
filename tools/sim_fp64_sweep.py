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
    """lp2: (T, L+1) float32 log2-probs of blank (col 0) and the labels; lab: (L,) labels.
    Returns lattice (T, 2L+2) float32 log2 of the scaled values [blank_i, label_i interleaved], G (T,) cumulative
    integer scale (true = stored * 2^-G), log2 Z."""
    T, K = lp2.shape
    L = K - 1
    n = L + 1
    skip = np.zeros(n)
    skip[1:L] = (lab[1:] != lab[:-1]).astype(np.float64)
    p = exp2_f64(lp2)                       # (T, K)
    ab = np.zeros(n)
    al = np.zeros(n)
    ab[0] = p[0, 0]
    if L > 0:
        al[0] = p[0, 1]
    lat = np.full((T, 2 * n), -3.0e38, dtype=np.float32)
    G = np.zeros(T, dtype=np.int64)
    g_hist = []                              # applied per chunk
    lvl_hist = []                            # measured level (max exponent) at the start of each chunk
    Gc = 0
    min_margin = 1e9

    def store(t):
        with np.errstate(divide="ignore"):
            lat[t, 0::2] = np.where(ab > 0, np.log2(ab), -3.0e38).astype(np.float32)
            lat[t, 1::2] = np.where(al > 0, np.log2(al), -3.0e38).astype(np.float32)

    store(0)
    for t in range(1, T):
        c, f = divmod(t, CHF)
        g = 0
        if f == 0:
            # level at the start of chunk c, published by the state warps
            mx = max(ab.max(), al.max())
            lvl_hist.append(np.floor(np.log2(mx)) if mx > 0 else -2000)
            # the helper sees the level of chunk c - lag only
            cm = len(lvl_hist) - 1 - lag
            if cm >= 0:
                applied = sum(g_hist[cm:])                       # shifts applied since that measurement
                drift = 0.0
                if cm >= 1:
                    drift = (lvl_hist[cm] - lvl_hist[cm - 1]) - g_hist[cm - 1]      # true change per chunk
                est_now = lvl_hist[cm] + applied + drift * lag
                g = int(np.clip(-np.rint(est_now + drift), -900, 900))   # centre at the end of this chunk
            g_hist.append(g)
            Gc += g
        G[t] = Gc
        nbr = np.concatenate(([0.0], al[:-1]))
        sc = np.ldexp(1.0, g)
        pb = p[t, 0] * sc
        pl = np.zeros(n)
        pl[:L] = p[t, 1:] * sc
        nab = pb * (ab + nbr)
        nal = pl * (al + ab + skip * nbr)
        ab, al = nab, nal
        store(t)
        mx = max(ab.max(), al.max())
        if mx > 0:
            min_margin = min(min_margin, np.log2(mx) + 1074)
        assert np.isfinite(mx), "overflow"
    z = ab[L] + (al[L - 1] if L > 0 else 0.0)
    with np.errstate(divide="ignore"):
        log2z = np.log2(z) - Gc if z > 0 else -np.inf
    return lat, G, log2z, min_margin


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
        la, Ga, za, m1 = sweep(lp2, lab, lag)
        lpr = lp2[::-1][:, np.concatenate(([0], np.arange(Lb, 0, -1)))]     # time- and label-reversed
        lb_, Gb, zb, m2 = sweep(lpr, lab[::-1], lag)
        lb_ = lb_[::-1]
        Gb = Gb[::-1]
        nll_b = -za * np.log(2.0)
        # occupancies as the patch kernel: un-normalised log2 occupancy per state, normalised per row in fp32
        n = Lb + 1
        S = 2 * Lb + 1
        a_bl, a_lb = la[:, 0::2], la[:, 1::2]
        # beta is stored in reversed state order: blank k <-> blank (L-k), label k <-> label (L-1-k)
        b_bl = lb_[:, 0::2][:, ::-1]
        b_lb = np.concatenate((lb_[:, 1::2][:, :Lb][:, ::-1], np.full((Tb, 1), -3e38, np.float32)), axis=1)
        ub = (a_bl + b_bl - lp2[:, :1]).astype(np.float32)
        ul = np.full((Tb, n), -3e38, dtype=np.float32)
        ul[:, :Lb] = (a_lb[:, :Lb] + b_lb[:, :Lb] - lp2[:, 1:]).astype(np.float32)
        m = np.maximum(ub.max(1), ul.max(1))[:, None]
        eb = np.exp2(np.maximum(ub - m, -200)).astype(np.float32) * (ub > -1e38)
        el = np.exp2(np.maximum(ul - m, -200)).astype(np.float32) * (ul > -1e38)
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
        cert = -1074.0 + Ga + Gb + np.maximum(ma, mb) - za
        worst["cert_max"] = max(worst["cert_max"], float(cert.max()))
        worst["zdiff"] = abs(za - zb)
    return worst


if __name__ == "__main__":
    kinds = sys.argv[1:] or _cases.KINDS
    for k in kinds:
        for lag in (1, 6):
            print(run_case(k, lag=lag), "lag", lag, flush=True)
