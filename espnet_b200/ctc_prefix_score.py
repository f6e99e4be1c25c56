"""CTCPrefixScoreB200: the batched CTC prefix scorer on the B200 kernels.

Same surface as ``CTCPrefixScoreTH`` (espnet/nets/ctc_prefix_score.py:12-221):

    impl = CTCPrefixScoreB200(x, xlens, blank, eos, margin=0)
    scores, state = impl(y, state, scoring_ids=None, att_w=None)     # scores (n_bh, O)
    state = impl.index_select_state(state, best_ids)                 # after beam pruning

The difference is inside the state: the reference returns the forward variables r for every
(hypothesis, candidate) pair, (T, 2, n_bh, snum) -- 6.4 GB per step at B=32, beam=10, T=500,
O=5000 -- and ``index_select_state`` gathers the few survivors out of it.  Here ``__call__``
only computes the scores (one pass over x per step) and ``index_select_state`` recomputes
r (T, 2) for the survivors, which is what the next ``__call__`` needs.  The state tuples keep
the reference's positions: ``(r, log_psi, f_min, f_max, scoring_idmap)`` out of ``__call__``
(with ``r`` a :class:`PendingR`), ``(r_prev, s_prev, f_min, f_max)`` out of
``index_select_state`` (with ``s_prev`` of shape (n_bh,) instead of the reference's (n_bh, O)
copies of the same number).
"""
import torch

from . import _lib


class PendingR(object):
    """What is needed to recompute r for any (hypothesis, token) of one scoring step."""

    __slots__ = ("r_prev", "last_ids", "output_length", "start", "end", "n_bh")

    def __init__(self, r_prev, last_ids, output_length, start, end, n_bh):
        self.r_prev = r_prev
        self.last_ids = last_ids
        self.output_length = output_length
        self.start = start
        self.end = end
        self.n_bh = n_bh


class CTCPrefixScoreB200(object):
    """Batch processing of CTC prefix scores (Watanabe et al. 2017, Algorithm 2; Seki et al. 2019)."""

    def __init__(self, x, xlens, blank, eos, margin=0, columns=True):
        """Construct the scorer.

        :param torch.Tensor x: label posteriors (B, T, O), log-probabilities, CUDA.  Frames past
            xlens are overwritten in place exactly as the reference does
            (ctc_prefix_score.py:46-51) when x is float32 and contiguous.
        :param torch.Tensor xlens: input lengths (B,)
        :param int blank: blank label id
        :param int eos: end-of-sequence id
        :param int margin: margin parameter for windowing (0 means no windowing)
        :param bool columns: keep a (B, O, T) column copy of x for calls with scoring_ids (built at the
            first such call, once per scorer: B*T*O floats): the candidates' frames are then contiguous
            instead of O floats apart.  Same results bit for bit; False saves the memory.
        """
        if not x.is_cuda:
            raise RuntimeError("CTCPrefixScoreB200 runs on CUDA tensors only (there is no CPU fallback)")
        self.logzero = _lib.LOGZERO
        self.blank = int(blank)
        self.eos = int(eos)
        self.batch = x.size(0)
        self.input_length = x.size(1)
        self.odim = x.size(2)
        self.dtype = x.dtype
        self.device = x.device
        self._dev_index = x.device.index if x.device.index is not None else torch.cuda.current_device()
        self._lib = _lib.load()
        self._score_fn = self._lib.b200_ctc_prefix_score_columns
        self._select_fn = self._lib.b200_ctc_prefix_select
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.to(torch.float32).contiguous()
        self.x = x
        self.xlens = torch.as_tensor(xlens).to(device=self.device, dtype=torch.int64).contiguous()
        self.end_frames = torch.as_tensor(xlens) - 1
        B, T, O = self.batch, self.input_length, self.odim
        self.r0 = torch.empty((T, 2, B), dtype=torch.float32, device=self.device)
        self.xmax = torch.empty((B, O), dtype=torch.float32, device=self.device)
        rc = self._run(self._lib.b200_ctc_prefix_init, self.x.data_ptr(), self.xlens.data_ptr(), B, T, O, self.blank,
                       self.r0.data_ptr(), self.xmax.data_ptr(), self._stream())
        _lib.check(rc, "b200_ctc_prefix_init")
        self._x_ptr, self._xmax_ptr, self._xlens_ptr = self.x.data_ptr(), self.xmax.data_ptr(), self.xlens.data_ptr()
        self.margin = margin
        if margin > 0:
            self.frame_ids = torch.arange(T, dtype=torch.float32, device=self.device)
        self.idx_b = torch.arange(self.batch, device=self.device)
        self._work = None
        self.scoring_num = 0
        self._columns = bool(columns)
        self._xc = None

    def _stream(self):
        return torch._C._cuda_getCurrentRawStream(self._dev_index)

    def _run(self, fn, *args):
        """One library call on this scorer's device (the device guard only where the current device is another)."""
        if torch.cuda.current_device() == self._dev_index:
            return fn(*args)
        with torch.cuda.device(self.device):
            return fn(*args)

    def _workspace(self, n_bh):
        need = self._lib.b200_ctc_prefix_work_floats(self.input_length, n_bh, self.odim)
        if self._work is None or self._work.numel() < need:
            self._work = torch.empty(need, dtype=torch.float32, device=self.device)
        return self._work

    def _column_copy(self):
        if self._xc is None:
            B, T, O = self.batch, self.input_length, self.odim
            tc = self._lib.b200_ctc_prefix_column_stride(T)
            self._xc = torch.empty((B, O, tc), dtype=torch.float32, device=self.device)
            rc = self._run(self._lib.b200_ctc_prefix_columns, self.x.data_ptr(), B, T, O, self._xc.data_ptr(), self._stream())
            _lib.check(rc, "b200_ctc_prefix_columns")
        return self._xc

    def __call__(self, y, state, scoring_ids=None, att_w=None, last_ids=None):
        """Compute CTC prefix scores for next labels.

        :param list y: prefix label sequences (n_bh lists, or a (n_bh, len) tensor)
        :param tuple state: previous CTC state (from index_select_state) or None
        :param torch.Tensor scoring_ids: candidates per hypothesis (n_bh, snum) or None for all O
        :param torch.Tensor att_w: attention weights (n_bh, T) to decide the CTC window
        :param torch.Tensor last_ids: (not in the reference) the last label of every prefix as a contiguous int64
            CUDA tensor (n_bh,), for a search loop that has it already: saves the strided copy out of y
        :return: (ctc_local_scores (n_bh, O), new_state)
        """
        if last_ids is not None and torch.is_tensor(y):
            output_length = y.size(1) - 1
            if last_ids.device != self.device or last_ids.dtype != torch.int64 or not last_ids.is_contiguous():
                last_ids = last_ids.to(device=self.device, dtype=torch.int64).contiguous()
        elif torch.is_tensor(y):
            output_length = y.size(1) - 1
            last_ids = y[:, -1]
            if last_ids.device != self.device or last_ids.dtype != torch.int64:
                last_ids = last_ids.to(device=self.device, dtype=torch.int64)
            last_ids = last_ids.contiguous()
        else:
            output_length = len(y[0]) - 1  # ignore sos
            last_ids = torch.tensor([int(yi[-1]) for yi in y], dtype=torch.int64, device=self.device)
        n_bh = last_ids.numel()
        n_hyps = n_bh // self.batch  # each utterance has the same number of hypotheses
        T, O = self.input_length, self.odim
        self.scoring_num = scoring_ids.size(-1) if scoring_ids is not None else 0
        if state is None:
            r_prev = self.r0.unsqueeze(3).expand(T, 2, self.batch, n_hyps).reshape(T, 2, n_bh).contiguous()
            s_prev = None
            f_min_prev, f_max_prev = 0, 1
        else:
            r_prev, s_prev, f_min_prev, f_max_prev = state
            if not r_prev.is_contiguous():
                r_prev = r_prev.contiguous()
            if torch.is_tensor(s_prev):
                if s_prev.dim() == 2:  # the reference's (n_bh, O) layout
                    s_prev = s_prev[:, 0]
                if s_prev.dtype != torch.float32 or not s_prev.is_contiguous():
                    s_prev = s_prev.to(torch.float32).contiguous()
            else:
                s_prev = None if s_prev == 0.0 else torch.full((n_bh,), float(s_prev), device=self.device)

        # (the reference also builds scoring_idmap (n_bh, O) here, the inverse of scoring_ids, to find a candidate's
        # slice of its r tensor later (ctc_prefix_score.py:100-115,208-214); r is recomputed per survivor here, so the
        # map -- a 12.8 MB fill + scatter per step at C3 -- has no reader: the state keeps None in its place)
        scoring_idmap = None
        if self.scoring_num > 0 and (scoring_ids.device != self.device or scoring_ids.dtype != torch.int64
                                     or not scoring_ids.is_contiguous()):
            scoring_ids = scoring_ids.to(device=self.device, dtype=torch.int64).contiguous()

        # frame window from the attention weights (ctc_prefix_score.py:145-154)
        if att_w is not None and self.margin > 0:
            f_arg = torch.matmul(att_w.to(torch.float32), self.frame_ids)
            lo, hi = torch.stack(torch.aminmax(f_arg)).tolist()   # ONE device->host copy (the reference: two)
            f_min = max(int(lo), f_min_prev)
            f_max = max(int(hi), f_max_prev)
            start = min(f_max_prev, max(f_min - self.margin, output_length, 1))
            end = min(f_max + self.margin, T)
        else:
            f_min = f_max = 0
            start = max(output_length, 1)
            end = T

        both = torch.empty((2, n_bh, O), dtype=torch.float32, device=self.device)   # scores, log_psi: one allocation
        scores, log_psi = both[0], both[1]
        work = self._workspace(n_bh)
        xc = self._column_copy() if (self.scoring_num > 0 and self._columns) else None
        p0 = both.data_ptr()
        rc = self._run(
            self._score_fn,
            self._x_ptr, None if xc is None else xc.data_ptr(), self._xmax_ptr, self._xlens_ptr,
            self.batch, T, O, n_hyps,
            self.blank, self.eos, r_prev.data_ptr(), None if s_prev is None else s_prev.data_ptr(),
            last_ids.data_ptr(), output_length,
            None if scoring_ids is None else scoring_ids.data_ptr(), self.scoring_num, start, end,
            p0, p0 + 4 * n_bh * O, work.data_ptr(), self._stream())
        if rc:
            _lib.check(rc, "b200_ctc_prefix_score")
        pending = PendingR(r_prev, last_ids, output_length, start, end, n_bh)
        if self.dtype != torch.float32:
            scores = scores.to(self.dtype)
        return scores, (pending, log_psi, f_min, f_max, scoring_idmap)

    def select(self, state, parent, token):
        """r (T,2,n_new) and s (n_new,) of the hypotheses parent[k] extended by token[k]."""
        pending, log_psi, f_min, f_max, _ = state
        if parent.device != self.device or parent.dtype != torch.int64 or not parent.is_contiguous():
            parent = parent.to(device=self.device, dtype=torch.int64).contiguous()
        if token is not None:
            token = token.to(device=self.device, dtype=torch.int64).contiguous()
        n_new = parent.numel()
        T, O = self.input_length, self.odim
        n_hyps = pending.n_bh // self.batch
        r_new = torch.empty((T, 2, n_new), dtype=torch.float32, device=self.device)
        s_new = torch.empty((n_new,), dtype=torch.float32, device=self.device)
        rc = self._run(
            self._select_fn,
            self._x_ptr, self.batch, T, O, n_hyps, self.blank, pending.r_prev.data_ptr(),
            log_psi.data_ptr(), pending.last_ids.data_ptr(), pending.output_length, pending.start,
            pending.end, pending.n_bh, parent.data_ptr(), None if token is None else token.data_ptr(), n_new, None,
            r_new.data_ptr(), s_new.data_ptr(), self._stream())
        if rc:
            _lib.check(rc, "b200_ctc_prefix_select")
        return r_new, s_new, f_min, f_max

    def index_select_state(self, state, best_ids):
        """Select CTC states according to best ids (ctc_prefix_score.py:190-221).

        :param state: CTC state returned by __call__
        :param best_ids: index numbers selected by beam pruning (B, W), in hyp * O + token space
        :return: (r_new (T,2,B*W), s_new (B*W,), f_min, f_max)
        """
        # parent = best_ids // O + utterance * n_hyps and token = best_ids % O are derived inside the kernel
        return self.select(state, best_ids, None)

