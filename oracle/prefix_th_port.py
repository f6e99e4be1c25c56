"""Eager-PyTorch restatement of CTCPrefixScoreTH -- BASELINE INFRASTRUCTURE, NOT PRODUCT CODE.

The reference is Python and does not exist on the GPU box, so bench.py times this restatement
of espnet/nets/ctc_prefix_score.py:69-221 (same tensor shapes, the same per-frame
stack + logsumexp loop at :157-162, the same per-hypothesis Python loops at :136-142 and
:182-183) as the "reference on GPU" and "reference on CPU" step rates (kind "port").
It is also cross-checked against the C oracle in tests/test_oracle_vs_golden.py.
"""
import torch

LOGZERO = -10000000000.0


class PrefixScoreEager(object):
    def __init__(self, x, xlens, blank, eos):
        self.blank, self.eos = blank, eos
        self.B, self.T, self.O = x.shape
        self.device, self.dtype = x.device, x.dtype
        for i, l in enumerate(xlens):  # :46-51
            if l < self.T:
                x[i, l:, :] = LOGZERO
                x[i, l:, blank] = 0
        xn = x.transpose(0, 1)
        xb = xn[:, :, blank].unsqueeze(2).expand(-1, -1, self.O)
        self.x = torch.stack([xn, xb])  # (2, T, B, O)
        self.end_frames = torch.as_tensor(xlens) - 1
        self.idx_b = torch.arange(self.B, device=self.device)

    def step(self, last_ids, output_length, state):
        """scoring_ids=None path (what recognize_beam_batch uses on CUDA, rnn/decoders.py:758-764)."""
        n_bh = len(last_ids)
        n_hyps = n_bh // self.B
        T, O = self.T, self.O
        if state is None:  # :84-95
            r_prev = torch.full((T, 2, self.B, n_hyps), LOGZERO, dtype=self.dtype, device=self.device)
            r_prev[:, 1] = torch.cumsum(self.x[0, :, :, self.blank], 0).unsqueeze(2)
            r_prev = r_prev.view(-1, 2, n_bh)
            s_prev = 0.0
        else:
            r_prev, s_prev = state
        x_ = self.x.unsqueeze(3).repeat(1, 1, 1, n_hyps, 1).view(2, -1, n_bh, O)  # :120
        r = torch.full((T, 2, n_bh, O), LOGZERO, dtype=self.dtype, device=self.device)
        if output_length == 0:
            r[0, 0] = x_[0, 0]
        r_sum = torch.logsumexp(r_prev, 1)
        log_phi = r_sum.unsqueeze(2).repeat(1, 1, O)
        for idx in range(n_bh):  # :141-142
            log_phi[:, idx, last_ids[idx]] = r_prev[:, 1, idx]
        start, end = max(output_length, 1), T
        for t in range(start, end):  # :157-162
            rp = r[t - 1]
            rr = torch.stack([rp[0], log_phi[t - 1], rp[0], rp[1]]).view(2, 2, n_bh, O)
            r[t] = torch.logsumexp(rr, 1) + x_[:, t]
        log_phi_x = torch.cat((log_phi[0].unsqueeze(0), log_phi[:-1]), dim=0) + x_[0]
        log_psi = torch.logsumexp(torch.cat((log_phi_x[start:end], r[start - 1, 0].unsqueeze(0)), dim=0), dim=0)
        for si in range(n_bh):  # :182-183
            log_psi[si, self.eos] = r_sum[self.end_frames[si // n_hyps], si]
        log_psi[:, self.blank] = LOGZERO
        return (log_psi - s_prev), (r, log_psi)

    def select(self, state, best_ids):
        """:190-221 without scoring_idmap."""
        r, s = state
        n_bh = len(s)
        n_hyps = n_bh // self.B
        vidx = (best_ids + (self.idx_b * (n_hyps * self.O)).view(-1, 1)).view(-1)
        s_new = torch.index_select(s.view(-1), 0, vidx)
        s_new = s_new.view(-1, 1).repeat(1, self.O).view(n_bh, self.O)
        r_new = torch.index_select(r.view(-1, 2, n_bh * self.O), 2, vidx).view(-1, 2, n_bh)
        return r_new, s_new
