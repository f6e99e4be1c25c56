"""Test helpers for the multi-utterance search loop: the REFERENCE's own attention decoder
(oracle/_ref: espnet/nets/pytorch_backend/rnn/decoders.py, attentions.py) with seeded random weights, run
 (a) through its own ``recognize_beam_batch`` -- the oracle -- and
 (b) as the ``DecoderStep`` callback of ``espnet_b200.recog_batch.recognize_beam_batch``, which then differs from (a)
     only in the glue and the CTC prefix scorer under test."""
from argparse import Namespace

import torch

from tests import _cases


def make_case(batch=4, T=40, odim=50, eprojs=16, seed=0, kind="trained", L=8):
    """Encoder states h (batch,T,eprojs), lengths, and CTC log-probs lpz (batch,T,odim) with an alignment in them."""
    c = _cases.make_prefix_case(batch, T, odim, 1, seed=seed, nsteps=1, kind=kind, L=L)
    g = torch.Generator().manual_seed(seed + 77)
    h = torch.randn(batch, T, eprojs, generator=g)
    return {"h": h, "hlens": c["xlens"].clone(), "lpz": c["lp"].clone(), "odim": odim, "eos": odim - 1}


def build_reference_decoder(odim, eprojs=16, dunits=24, att_dim=12, seed=0):
    """The reference Decoder (1-layer LSTM, additive attention), weights drawn from `seed`, eval mode."""
    from espnet.nets.pytorch_backend.rnn.attentions import AttAdd
    from espnet.nets.pytorch_backend.rnn.decoders import Decoder

    torch.manual_seed(seed)
    att = torch.nn.ModuleList([AttAdd(eprojs, dunits, att_dim)])
    dec = Decoder(eprojs, odim, "lstm", 1, dunits, odim - 1, odim - 1, att, verbose=0,
                  char_list=[str(i) for i in range(odim)], labeldist=None, lsm_weight=0.0, sampling_probability=0.0,
                  dropout=0.0, context_residual=False, replace_sos=False, num_encs=1)
    for p in dec.parameters():   # larger weights than the default init: the attention scores then discriminate
        p.data.mul_(3.0)
    return dec.eval()


def recog_args(beam, ctc_weight=0.3, nbest=3, margin=0, maxlenratio=0.0, minlenratio=0.0, penalty=0.0):
    return Namespace(beam_size=beam, penalty=penalty, ctc_weight=ctc_weight, maxlenratio=maxlenratio,
                     minlenratio=minlenratio, nbest=nbest, lm_weight=0.0, ctc_window_margin=margin, tgt_lang=None)


def reference_nbest(dec, case, args, device):
    """decoders.py:631-974 as it is: the oracle."""
    dec = dec.to(device)
    with torch.no_grad():
        out = dec.recognize_beam_batch(case["h"].to(device), case["hlens"].clone(), case["lpz"].to(device), args,
                                       [str(i) for i in range(case["odim"])], rnnlm=None, normalize_score=True)
    return [[{"yseq": [int(t) for t in h["yseq"]], "score": float(h["score"])} for h in hyps] for hyps in out]


class ReferenceDecoderStep(object):
    """The reference decoder's modules behind the DecoderStep interface (decoders.py:779-812, :853-893)."""

    def __init__(self, dec, h, hlens, beam):
        from espnet.nets.pytorch_backend.nets_utils import mask_by_length

        self.dec = dec
        h = mask_by_length(h, hlens, 0.0)
        batch = h.size(0)
        self.n_bb = batch * beam
        self.exp_hlens = hlens.repeat(beam).view(beam, batch).transpose(0, 1).contiguous().view(-1).tolist()
        self.exp_h = h.unsqueeze(1).repeat(1, beam, 1, 1).contiguous().view(self.n_bb, h.size(1), h.size(2))
        dec.att[0].reset()

    def start(self, n_bb):
        z = lambda: [torch.zeros(n_bb, self.dec.dunits, device=self.exp_h.device) for _ in range(self.dec.dlayers)]
        return {"z_prev": z(), "c_prev": z(), "z_list": z(), "c_list": z(), "a_prev": None}

    def step(self, vy, st):
        d = self.dec
        with torch.no_grad():
            ey = d.dropout_emb(d.embed(vy))
            att_c, att_w = d.att[0](self.exp_h, self.exp_hlens, d.dropout_dec[0](st["z_prev"][0]), st["a_prev"])
            ey = torch.cat((ey, att_c), dim=1)
            z_list, c_list = d.rnn_forward(ey, st["z_list"], st["c_list"], st["z_prev"], st["c_prev"])
            logits = d.output(d.dropout_dec[-1](z_list[-1]))
            logp = torch.log_softmax(logits, dim=1)
        return logp, att_w, {"z_list": z_list, "c_list": c_list, "att_w": att_w}

    def select(self, st, vidx):
        n = self.n_bb
        return {"a_prev": torch.index_select(st["att_w"].view(n, *st["att_w"].shape[1:]), 0, vidx),
                "z_prev": [torch.index_select(z.view(n, -1), 0, vidx) for z in st["z_list"]],
                "c_prev": [torch.index_select(c.view(n, -1), 0, vidx) for c in st["c_list"]],
                "z_list": st["z_list"], "c_list": st["c_list"]}
