"""A seeded hypothesis walk that drives a prefix scorer under test and the REFERENCE's CTCPrefixScoreTH
(oracle/_ref, staged by oracle/stage_ref.py) side by side, comparing scores and selected states at every step."""
import pytest
import numpy as np
import torch


def walk_vs_reference_th(c, nsteps, make_ours, snum=0, margin=0, seed=0, follow=True, device="cpu", close=None):
    """make_ours(lp, xlens, eos, margin) -> object with
         step(y, state, ids, att) -> (scores numpy (n_bh,V), state)
         select(state, best (B,H)) -> (next state, r numpy (T,2,n_bh), s numpy (n_bh,))
    close(got, ref, step, what="scores", mag=None) asserts; mag: magnitude of the operands a score is the difference
    of (score = log_psi - s_prev: its float32 resolution is that of log_psi, not of the difference)."""
    from oracle import stage_ref

    if not stage_ref.activate():
        pytest.skip("oracle/_ref is not staged (python oracle/stage_ref.py in the build container)")
    from espnet.nets.ctc_prefix_score import CTCPrefixScoreTH

    lp, xlens, H, eos, labels = c["lp"], c["xlens"], c["H"], c["eos"], c.get("labels")
    B, T, V = lp.shape
    g = torch.Generator().manual_seed(seed)
    ours = make_ours(lp, xlens, eos, margin)
    ref = CTCPrefixScoreTH(lp.clone().to(device), xlens, 0, eos, margin=margin)
    y = [[eos] for _ in range(B * H)]
    so = sr = None
    for step in range(nsteps):
        ids = None
        if snum:
            ids = torch.stack([torch.randperm(V - 2, generator=g)[:snum] + 1 for _ in range(B * H)])
            if labels is not None and follow and step < labels.shape[1]:
                for u in range(B):   # the first hypothesis of every utterance can follow its labels
                    k, lab = u * H, int(labels[u, step])
                    if lab not in ids[k].tolist():
                        ids[k, 0] = lab
        att = None
        if margin:
            att = torch.zeros(B * H, T)
            centre = torch.randint(min(5 + 4 * step, T - 2), min(12 + 6 * step, T - 1), (B * H,), generator=g)
            att[torch.arange(B * H), centre] = 1.0
        s1, st1 = ours.step(y, so, ids, att)
        s2, st2 = ref(y, sr, None if ids is None else ids.to(device), None if att is None else att.to(device))
        lpsi = st2[1].detach().cpu().numpy()
        live = np.abs(lpsi) < 1e9
        close(s1, s2.cpu().numpy(), step, "scores", float(np.abs(lpsi[live]).max()) if live.any() else 0.0)
        hyp = torch.randint(0, H, (B, H), generator=g)
        if ids is None:
            tok = torch.randint(1, V - 1, (B, H), generator=g)
        else:
            pick = torch.randint(0, snum, (B, H), generator=g)
            tok = ids.view(B, H, snum)[torch.arange(B).view(-1, 1), hyp, pick]
        if labels is not None and follow and step < labels.shape[1]:
            hyp[:, 0] = 0
            tok[:, 0] = labels[:, step]
        best = hyp * V + tok
        so, r1, sv1 = ours.select(st1, best)
        sr = ref.index_select_state(st2, best.to(device))
        close(r1, sr[0].cpu().numpy(), step, "r")
        close(sv1, sr[1][:, 0].cpu().numpy(), step, "s")
        parent = (hyp + torch.arange(B).view(-1, 1) * H).view(-1)
        y = [y[int(parent[k])] + [int(tok.view(-1)[k])] for k in range(B * H)]
