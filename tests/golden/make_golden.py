"""Generates the golden fixtures under tests/golden/ from the REFERENCE itself.

Run in the build container only (the reference tree does not exist on the GPU box):

    PYTHONPATH=/root/reference:tests/golden/stubs python tests/golden/make_golden.py

What is recorded (all from unmodified reference code, CPU, fixed seeds):
  loss_*.npz     espnet2/asr/ctc.py CTC.loss_fn (ctc_type="builtin") in float64 and float32:
                 loss, per-utterance nll, d loss / d logits.
  prefix_*.npz   espnet/nets/ctc_prefix_score.py CTCPrefixScoreTH: scores and selected states
                 for a seeded hypothesis walk (full vocabulary and scoring_ids variants),
                 plus the numpy CTCPrefixScore for the first hypothesis.
  viterbi.npz    espnet/nets/pytorch_backend/ctc.py CTC.forced_align.
  nbest.npz      espnet/nets/batch_beam_search.py BatchBeamSearch n-best with CTCPrefixScorer
                 (separate script: tests/golden/make_golden_nbest.py, also run by --nbest here).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, "stubs"))

from tests import _cases  # noqa: E402


def gen_loss():
    from espnet2.asr.ctc import CTC

    for name, case in _cases.loss_cases().items():
        x, ilen, olen, ys = case["x"], case["ilen"], case["olen"], case["ys"]
        B = x.shape[0]
        out = {"x": x.numpy(), "ilen": ilen.numpy(), "olen": olen.numpy(), "ys": ys.numpy()}
        for dt, tag in ((torch.float64, "f64"), (torch.float32, "f32")):
            for reduce in (True, False):
                ctc = CTC(x.shape[2], 4, ctc_type="builtin", reduce=reduce)
                xt = x.to(dt).transpose(0, 1).detach().requires_grad_()
                tg = torch.cat([ys[i, : olen[i]] for i in range(B)])
                loss = ctc.loss_fn(xt, tg, ilen, olen)
                if reduce:
                    loss.backward()
                    out[f"loss_{tag}"] = loss.detach().numpy()
                    out[f"grad_{tag}"] = xt.grad.transpose(0, 1).contiguous().numpy()
                else:
                    out[f"nll_over_B_{tag}"] = loss.detach().numpy()
        np.savez_compressed(os.path.join(HERE, f"loss_{name}.npz"), **out)
        print("loss", name, out["loss_f64"])


def gen_prefix():
    from espnet.nets.ctc_prefix_score import CTCPrefixScore, CTCPrefixScoreTH

    for name, case in _cases.prefix_cases().items():
        lp, xlens, H, eos, nsteps, snum = (case[k] for k in ("lp", "xlens", "H", "eos", "nsteps", "snum"))
        B, T, V = lp.shape
        g = torch.Generator().manual_seed(case["seed"] + 7)
        th = CTCPrefixScoreTH(lp.clone(), xlens, 0, eos)
        out = {"lp": lp.numpy(), "xlens": xlens.numpy(), "H": H, "eos": eos, "snum": snum}
        y = [[eos] for _ in range(B * H)]
        state = None
        for step in range(nsteps):
            ids = None
            if snum > 0:
                ids = torch.stack([torch.randperm(V, generator=g)[:snum] for _ in range(B * H)])
                out[f"ids_{step}"] = ids.numpy()
            scores, st = th(y, state, ids)
            out[f"scores_{step}"] = scores.numpy()
            # choose survivors: seeded (hyp, token) pairs per utterance, tokens never blank
            best = torch.empty(B, H, dtype=torch.long)
            for b in range(B):
                for h in range(H):
                    hyp = int(torch.randint(0, H, (1,), generator=g))
                    if snum > 0:
                        row = ids[b * H + hyp]
                        tok = int(row[int(torch.randint(0, snum, (1,), generator=g))])
                        if tok == 0:  # never extend with blank; stay inside the scored set
                            tok = int(row[row != 0][0])
                    else:
                        tok = int(torch.randint(1, V, (1,), generator=g))
                    if tok == 0:
                        tok = 1
                    best[b, h] = hyp * V + tok
            out[f"best_{step}"] = best.numpy()
            state = th.index_select_state(st, best)
            out[f"r_{step}"] = state[0].numpy()
            out[f"s_{step}"] = state[1][:, 0].numpy()
            hyp_idx = (best // V + torch.arange(B).view(-1, 1) * H).view(-1)
            tok = (best % V).view(-1)
            y = [y[int(hyp_idx[k])] + [int(tok[k])] for k in range(B * H)]
        # single-hypothesis numpy version on utterance 0, first step
        xn = lp[0, : int(xlens[0])].numpy()
        impl = CTCPrefixScore(xn, 0, eos, np)
        r0 = impl.initial_state()
        cs = np.arange(V)
        psi, _ = impl([eos], cs, r0)
        out["np_first_step_psi"] = psi
        np.savez_compressed(os.path.join(HERE, f"prefix_{name}.npz"), **out)
        print("prefix", name, out["scores_0"].shape)


def gen_viterbi():
    from espnet.nets.pytorch_backend.ctc import CTC

    out = {}
    for k, (T, V, L, seed) in enumerate(_cases.viterbi_cases()):
        g = torch.Generator().manual_seed(seed)
        ctc = CTC(V, V, 0.0, ctc_type="builtin")
        with torch.no_grad():
            ctc.ctc_lo.weight.copy_(torch.eye(V))
            ctc.ctc_lo.bias.zero_()
        h = torch.randn(T, V, generator=g)
        y = torch.randint(1, V, (L,), generator=g)
        if L >= 2:
            y[1] = y[0]
        with torch.no_grad():
            ali = ctc.forced_align(h.unsqueeze(0), y)
        out[f"h_{k}"] = h.numpy()
        out[f"y_{k}"] = y.numpy()
        out[f"align_{k}"] = np.asarray(ali, dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "viterbi.npz"), **out)
    print("viterbi ok")


if __name__ == "__main__":
    gen_loss()
    gen_prefix()
    gen_viterbi()
    if "--nbest" in sys.argv:
        from tests.golden import make_golden_nbest

        make_golden_nbest.main()
