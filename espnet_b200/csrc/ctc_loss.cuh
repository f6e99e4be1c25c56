// ctc_loss.cuh -- parameter block and workspace layout shared by the CTC loss kernels.
//
// Replaces espnet2/asr/ctc.py:46-52 (CTC.loss_fn: log_softmax + torch.nn.CTCLoss + /B)
// and its autograd backward.  Three kernels, all in log2 units:
//
// Fused forward+backward (b200_ctc_loss_fwd_bwd), all in log2 units:
//   stream    (HBM)     ctc_loss_stream.inl: the one pass over the logits.  Rows are staged in a
//                       shared-memory ring by 1-D bulk (TMA) copies; per row: log-sum-exp, the
//                       blank + label columns gathered into lpg[b][t][0..L], and w * softmax
//                       written in place and bulk-stored to the gradient.
//   sweep     (latency) ctc_loss_sweep.inl: one CTA per (utterance, direction), the alpha (or, on
//                       the reversed sequence, beta) recursion over lpg, one thread per label
//                       position, warps skewed against each other instead of barriers,
//                       re-centred so that values stay O(10) in fp32.
//   patch     (L2)      ctc_loss_occ.inl<PATCH>: occupancies from alpha+beta, normalised per row,
//                       written over the blank / label columns of the gradient.
// Separate forward / backward calls (autograd with an upstream gradient known only later):
//   rowstats -> sweep -> occ (forward), grad (backward, re-reads the logits) in ctc_loss_rows.inl.
#pragma once
#include "common.cuh"

namespace b200 {

// marks a merged duplicate label in LossWs::occ (softmax - occupancy lies in [-1, 1])
constexpr float kCorrSkip = 2.0f;
// LossWs::dupcode[k]: (index of the next position with the same label) + 1 in the low bits, 0 if none;
// kDupLater set if an earlier position carries the same label.  0 for a label that occurs once.
constexpr int kDupLater = 0x40000000;
constexpr int kDupNextMask = 0x3fffffff;

struct LossWs {
  float* lse2;      // [B][T]      row log2-sum-exp (log2 units)
  float* lpg;       // [B][T][Kp]  log2-prob of blank (k=0) and label k-1 (k>=1)
  float* lat;       // [2][B][T][Sp] alpha (dir 0) and beta in reversed-state order (dir 1)
  float* occ;       // [B][T][Kp]  softmax - occupancy of blank (k=0) / label k-1, duplicates merged into
                    //             the first occurrence; 2.0 marks a later duplicate (nothing to write)
  int* labels;      // [B][Kp]     labels as int32
  int* dup_next;    // [B][Kp]     next position with the same label, or -1
  int* is_first;    // [B][Kp]     1 if no earlier position has the same label
  int* dupcode;     // [B][Kp]     both of the above packed (see kDupLater)
  float* logz_beta; // [B]         log2 p(y|x) from the beta sweep (consistency check)
  float* nll_ws;    // [B]         copy of nll (nats) kept for the backward pass
  int* bad;         // [64]        [0] set when a label is outside [0,V) or equals blank, or olen > Lmax;
                    //             [1], [2] task counters of the rowstats / gradient kernels, [4] utterances
                    //             whose nll is final, [5] range tickets of the patch kernel; zeroed at the
                    //             start of every call
  int* badutt;      // [B]         1 if utterance b has an invalid label or olen[b] > Lmax: its nll is NaN and
                    //             its gradient rows are NaN (a corrupt target batch must not train silently)
  int* order;       // [B]         utterances by decreasing ilen (rank -> b): long lattices start first
  int* done;        // [B]         finished sweeps of utterance b (2 = its alpha and beta lattices are complete)
  // ---- linear-domain (fp64) sweeps, ctc_loss_lin.inl
  double* pd;       // [B][T][Kp]  2^(lpg + gff): the probabilities of blank / labels as doubles, scaled per frame
                    //             so that the largest of them is in [1, 2) (feed-forward scale)
  int* gff;         // [B][T]      that per-frame scale exponent (>= 0)
  int* gwarp;       // [2][B][ceil(T/8)][8]  block (warp) scale exponent of the sweep during chunk c:
                    //             true value = stored * 2^-(gcum + gwarp)
  int* gtot;        // [B]         sum over the frames of gff
  int* redo;        // [B]         1: the linear sweep could not certify utterance b (range), the log-domain
                    //             kernels redo it
  int64_t Kp, Sp;
  size_t bytes;
};

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Carves the workspace; base may be NULL to only compute the size.
inline LossWs carve_loss_ws(void* base, int64_t T, int64_t B, int64_t V, int64_t Lmax) {
  (void)V;
  LossWs w;
  // (one spare slot after the L+1 columns: the stream kernel leaves the frame's feed-forward exponent there)
  w.Kp = (int64_t)align_up((size_t)(Lmax + 2), 8);
  w.Sp = 2 * w.Kp;
  size_t off = 0;
  char* p = (char*)base;
  auto take = [&](size_t n) { size_t o = off; off = align_up(off + n, 256); return p ? p + o : (char*)0; };
  w.lse2 = (float*)take(sizeof(float) * B * T);
  w.lpg = (float*)take(sizeof(float) * B * T * w.Kp);
  w.lat = (float*)take(sizeof(float) * 2 * B * T * w.Sp);
  w.occ = (float*)take(sizeof(float) * B * T * w.Kp);
  w.labels = (int*)take(sizeof(int) * B * w.Kp);
  w.dup_next = (int*)take(sizeof(int) * B * w.Kp);
  w.is_first = (int*)take(sizeof(int) * B * w.Kp);
  w.dupcode = (int*)take(sizeof(int) * B * w.Kp);
  w.logz_beta = (float*)take(sizeof(float) * B);
  w.nll_ws = (float*)take(sizeof(float) * B);
  w.bad = (int*)take(sizeof(int) * 64);
  w.badutt = (int*)take(sizeof(int) * B);
  w.order = (int*)take(sizeof(int) * B);
  w.done = (int*)take(sizeof(int) * B);
  w.pd = (double*)take(sizeof(double) * B * T * w.Kp);
  w.gff = (int*)take(sizeof(int) * B * T);
  w.gwarp = (int*)take(sizeof(int) * 2 * B * ((T + 7) / 8) * 8);
  w.gtot = (int*)take(sizeof(int) * B);
  w.redo = (int*)take(sizeof(int) * B);
  w.bytes = off;
  return w;
}

struct LossParams {
  const void* logits;
  void* grad;
  int64_t T, B, V;
  int64_t stride_t, stride_b;    // logits element strides
  int64_t gstride_t, gstride_b;  // gradient element strides
  const int64_t* targets;
  const int64_t* tgt_offsets;
  const int64_t* ilen;
  const int64_t* olen;
  int64_t Lmax;
  int blank;
  const float* grad_weight;
  float grad_scale;
  float* nll;
  LossWs ws;
  int b0, b1;     // utterance range handled by this launch
  float* loss_sum; // loss_scale * sum_b nll[b], written by the sweep CTA that finishes last (may be NULL)
  float loss_scale;
  int prep_after_stream;  // fused call: the prep kernel is a programmatic dependent of the stream kernel
  int lin;        // the stream kernel also writes pd / gff for the linear-domain sweeps
  int redo_mode;  // log-domain sweep / patch kernels: only the utterances flagged in ws.redo
};

// label count of utterance b as every kernel uses it: olen clamped to the Lmax the workspace was sized for
// (the prep kernel flags the utterance when the clamp changes anything)
__device__ __forceinline__ int utt_labels(const LossParams& p, int b) {
  const int64_t L = p.olen[b];
  return (int)(L < 0 ? 0 : (L > p.Lmax ? p.Lmax : L));
}

}  // namespace b200
