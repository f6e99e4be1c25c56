"""CPU partial scorer for the beam-search tests, built on the C oracle (TEST INFRASTRUCTURE: the product's
scorer is espnet_b200.scorers.CTCPrefixScorer on the GPU; this one lets the search glue run without one)."""
import numpy as np
import torch

from oracle import ctc_oracle as O


class OracleCTCScorer(object):
    """batch_init_state / batch_score_partial / select_state of espnet/nets/scorers/ctc.py:88-127, one utterance."""

    def __init__(self, eos, per_hypothesis_only=True):
        self.eos = eos

    def batch_init_state(self, x):
        self.lp = np.ascontiguousarray(torch.log_softmax(x, dim=1).unsqueeze(0).numpy(), dtype=np.float32)
        self.xlens = np.array([x.shape[0]], dtype=np.int64)
        self.r0 = O.prefix_init(self.lp, self.xlens)
        return None

    def batch_score_partial(self, y, ids, state, x):
        n = y.shape[0]
        out_len = y.shape[1] - 1
        last = y[:, -1].numpy().astype(np.int64)
        if state[0] is None:
            r_prev, s_prev = np.repeat(self.r0, n, axis=2), None
        else:
            step, parent, token = state[0][0], [s[1] for s in state], [s[2] for s in state]
            r_prev, s_prev = O.prefix_select(self.lp, step["n"], 0, step["r_prev"], step["psi"], step["last"],
                                             step["out_len"], np.array(parent), np.array(token))
        sid = None if ids is None else ids.numpy()
        scores, psi, _ = O.prefix_score(self.lp, self.xlens, n, 0, self.eos, r_prev, s_prev, last, out_len,
                                        scoring_ids=sid)
        step = {"n": n, "r_prev": r_prev, "psi": psi, "last": last, "out_len": out_len}
        return torch.from_numpy(scores), step

    def select_state(self, state, i, new_id=None):
        return None if state is None else (state, int(i), int(new_id))
