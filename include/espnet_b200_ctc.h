/*
 * espnet_b200_ctc.h -- C ABI of libespnet_b200_ctc.so
 *
 * The drop-in boundary for ESPnet's CTC hot path on B200 (sm_100a).
 * Plain pointers, sizes and a cudaStream_t (passed as void*); no torch types,
 * no C++ exceptions.  Every entry point returns 0 on success and a non-zero
 * b200_status_t otherwise; b200_last_error() gives the thread-local message.
 *
 * All data pointers are DEVICE pointers unless the name ends in `_host`.
 * All buffers are owned by the caller; the library never allocates device
 * memory.  Calls are asynchronous and ordered on `stream`; they are safe to
 * call concurrently from several host threads (torch.nn.DataParallel calls
 * forward from one thread per GPU) because the library keeps no global
 * mutable state beyond a per-device cache of kernel attributes and the
 * thread-local error string.
 *
 * What each entry point replaces in the reference (kan-bayashi/espnet v0.9.5,
 * paths relative to the reference root):
 *
 *   b200_ctc_loss_forward / _backward / _fwd_bwd
 *       espnet2/asr/ctc.py:46-52  (CTC.loss_fn: log_softmax(2) + torch.nn.CTCLoss
 *       (reduction=sum|none) + "/ th_pred.size(1)") and its autograd backward;
 *       espnet/nets/pytorch_backend/ctc.py:52-65 (same, espnet1).
 *   b200_ctc_prefix_init
 *       espnet/nets/ctc_prefix_score.py:46-51 (in-place padding of x past xlens)
 *       and :84-92 (initial r_prev = cumsum of blank log-probs).
 *   b200_ctc_prefix_score
 *       espnet/nets/ctc_prefix_score.py:69-188 (CTCPrefixScoreTH.__call__), the
 *       score half: log_psi - s_prev for every (hypothesis, candidate).
 *   b200_ctc_prefix_columns / b200_ctc_prefix_score_columns
 *       the same call with scoring_ids (ctc_prefix_score.py:100-115, :159-163: the
 *       candidates' columns gathered with index_select; the pre-beam of
 *       espnet/nets/batch_beam_search.py:228-238) from a column copy of x.
 *   b200_ctc_prefix_select
 *       espnet/nets/ctc_prefix_score.py:190-221 (index_select_state) and
 *       espnet/nets/scorers/ctc.py:41-64 (select_state): the forward variables
 *       r (T,2) of the surviving (hypothesis, token) pairs only.
 *   b200_ctc_greedy
 *       espnet2/asr/ctc.py:103-111 (CTC.argmax over the vocabulary).
 *   b200_ctc_viterbi
 *       espnet/nets/pytorch_backend/ctc.py:153-216 (CTC.forced_align).
 *   b200_beam_prune
 *       espnet/nets/pytorch_backend/rnn/decoders.py:812-852 (score accumulation and
 *       per-utterance top-k of Decoder.recognize_beam_batch).
 *   b200_peer_sum_exchange
 *       espnet2/torch_utils/recursive_op.py:24,50 (the per-step all-reduces of
 *       espnet2/train/trainer.py:386-397).
 */
#ifndef ESPNET_B200_CTC_H_
#define ESPNET_B200_CTC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_CTC_ABI_VERSION 2

typedef enum {
  B200_OK = 0,
  B200_ERR_INVALID_ARGUMENT = 1,
  B200_ERR_MISALIGNED = 2,
  B200_ERR_UNSUPPORTED_SHAPE = 3,
  B200_ERR_WORKSPACE_TOO_SMALL = 4,
  B200_ERR_CUDA = 5
} b200_status_t;

typedef enum {
  B200_F32 = 0,
  B200_BF16 = 1,
  B200_F16 = 2
} b200_dtype_t;

/* The value CTCPrefixScoreTH uses for log(0) (ctc_prefix_score.py:34). */
#define B200_CTC_LOGZERO (-10000000000.0f)

/* ABI version of the loaded library (== B200_CTC_ABI_VERSION it was built with). */
int b200_ctc_abi_version(void);

/* Thread-local message of the last failing call on this thread ("" if none). */
const char* b200_last_error(void);

/* ------------------------------------------------------------------------
 * CTC loss, fused log_softmax + alpha/beta lattice + gradient w.r.t. logits
 * ------------------------------------------------------------------------
 * logits      (T,B,V) view with element strides (stride_t, stride_b, 1); the
 *             (B,T,V)-contiguous buffer ESPnet's ctc_lo produces, handed over
 *             as its .transpose(0,1) view (espnet2/asr/ctc.py:82), is
 *             stride_t = V, stride_b = T*V.
 * targets     int64 label ids; utterance b's labels are
 *             targets[tgt_offsets[b] .. tgt_offsets[b]+olen[b]).
 * ilen, olen  int64 (B,) frame and label counts.
 * Lmax        host-side upper bound on olen (sizes the lattice workspace).
 * nll         fp32 (B,) out: -log p(y_b | x_b) in nats; +inf if no alignment
 *             exists (torch.nn.CTCLoss, zero_infinity=False).
 * loss_sum    fp32 (1,) out, may be NULL: loss_scale * sum_b nll[b], summed in a fixed order by the
 *             lattice kernel that finishes last (CTC.loss_fn's "/ B" is loss_scale = 1/B: the
 *             reduced loss needs no further kernel).
 * workspace   b200_ctc_loss_workspace_bytes(...) bytes, 256-byte aligned; it
 *             carries the lattice from _forward to _backward and must stay
 *             untouched in between.
 */
size_t b200_ctc_loss_workspace_bytes(int64_t T, int64_t B, int64_t V, int64_t Lmax);

int b200_ctc_loss_forward(const void* logits, int dtype,
                          int64_t T, int64_t B, int64_t V,
                          int64_t stride_t, int64_t stride_b,
                          const int64_t* targets, const int64_t* tgt_offsets,
                          const int64_t* ilen, const int64_t* olen,
                          int64_t Lmax, int blank,
                          float* nll,
                          float* loss_sum /* may be NULL: loss_scale * sum_b nll[b] */, float loss_scale,
                          void* workspace, size_t workspace_bytes, void* stream);

/* grad_logits[t,b,v] = grad_scale * grad_weight[b] *
 *      (softmax(logits[t,b,:])[v] - occupancy[t,b,v])        for t <  ilen[b]
 *                    = 0                                     for t >= ilen[b]
 * grad_weight may be NULL (== 1 for every utterance).  Utterances whose nll is
 * +inf get NaN gradients for t < ilen[b], as torch does.
 * grad_logits has the dtype of logits and element strides (gstride_t, gstride_b, 1).
 * The logits, targets and lengths must be the ones given to _forward. */
int b200_ctc_loss_backward(const void* logits, int dtype,
                           int64_t T, int64_t B, int64_t V,
                           int64_t stride_t, int64_t stride_b,
                           const int64_t* targets, const int64_t* tgt_offsets,
                           const int64_t* ilen, const int64_t* olen,
                           int64_t Lmax, int blank,
                           const float* grad_weight, float grad_scale,
                           void* grad_logits, int64_t gstride_t, int64_t gstride_b,
                           void* workspace, size_t workspace_bytes, void* stream);

/* Forward and backward in one call (training with a known upstream gradient, and
 * the benchmark's "fwd+bwd" unit of work).  When the vocabulary rows are 16-byte
 * aligned and fit the shared-memory ring, the logits are read ONCE: the pass that
 * computes the row statistics also writes grad_scale*grad_weight*softmax to
 * grad_logits, and the lattice kernels then overwrite only the blank and label
 * columns.  Other shapes run _forward followed by _backward. */
int b200_ctc_loss_fwd_bwd(const void* logits, int dtype,
                          int64_t T, int64_t B, int64_t V,
                          int64_t stride_t, int64_t stride_b,
                          const int64_t* targets, const int64_t* tgt_offsets,
                          const int64_t* ilen, const int64_t* olen,
                          int64_t Lmax, int blank,
                          const float* grad_weight, float grad_scale,
                          float* nll, float* loss_sum, float loss_scale,
                          void* grad_logits, int64_t gstride_t, int64_t gstride_b,
                          void* workspace, size_t workspace_bytes, void* stream);

/* Measurement aid: ONLY the first stage of b200_ctc_loss_fwd_bwd (the
 * one pass over the logits: row statistics, label gather into the workspace and
 * grad_scale*grad_weight*softmax written to grad_logits).  bench.py times the
 * dominant HBM kernel with it; the result is not a gradient.  Returns
 * B200_ERR_UNSUPPORTED_SHAPE for shapes that _fwd_bwd runs on the three-pass path. */
int b200_ctc_loss_stream_pass(const void* logits, int dtype,
                              int64_t T, int64_t B, int64_t V,
                              int64_t stride_t, int64_t stride_b,
                              const int64_t* targets, const int64_t* tgt_offsets,
                              const int64_t* ilen, const int64_t* olen,
                              int64_t Lmax, int blank,
                              const float* grad_weight, float grad_scale,
                              void* grad_logits, int64_t gstride_t, int64_t gstride_b,
                              void* workspace, size_t workspace_bytes, void* stream);

/* Host-side helper for the label arguments of the calls above: lays [targets | tgt_offsets | ilen | olen]
 * out in the caller's PINNED host buffer (n_targets + 3B int64) and starts ONE asynchronous copy to `dev`
 * (same size) on `stream`.  tgt_offsets_host == NULL means "targets are concatenated" (what CTC.loss_fn
 * receives, espnet2/asr/ctc.py:85): the offsets are the exclusive prefix sum of olen.  The four device
 * arrays are then dev, dev + n_targets, dev + n_targets + B, dev + n_targets + 2B.  The pinned buffer may be
 * reused once the copy has completed (the caller orders that with an event). */
int b200_ctc_stage_labels(const int64_t* targets_host, int64_t n_targets,
                          const int64_t* tgt_offsets_host,
                          const int64_t* ilen_host, const int64_t* olen_host, int64_t B,
                          int64_t* pinned_host, int64_t* dev, void* stream);

/* The same with the pinned staging buffers owned by the library (a ring of 16 per device, grown on demand
 * with cudaHostAlloc -- page-locked HOST memory, never device memory -- each slot guarded by an event
 * recorded after its transfer).  The caller passes pageable host arrays and the device destination only.
 * The words travel by a small kernel on `stream` that reads the page-locked buffer over PCIe (no copy-engine
 * hand-over between the kernels of a step).  dev: 16-byte aligned, room for n_targets + 3*B words rounded
 * up to an even count. */
int b200_ctc_stage_labels_pinned(const int64_t* targets_host, int64_t n_targets,
                                 const int64_t* tgt_offsets_host,
                                 const int64_t* ilen_host, const int64_t* olen_host, int64_t B,
                                 int64_t* dev, void* stream);

/* Selects the lattice sweeps of _fwd_bwd.  on != 0 (the default): log-domain sweeps.  on == 0: float64
 * linear-domain sweeps (Lmax <= 255), whose patch kernel certifies every row it decodes and hands the
 * utterances it cannot certify (dynamic range) to the log-domain kernels; same results within the parity
 * bars, currently slower, kept selectable so that the parity tests exercise both.
 * Process-wide; returns the previous setting. */
int b200_ctc_debug_force_log_sweep(int on);

/* Score accumulation and per-utterance top-k of one output position of the joint CTC/attention search loop.
 * Replaces espnet/nets/pytorch_backend/rnn/decoders.py:812-852 (Decoder.recognize_beam_batch):
 *   local = att_weight * att_logp;  [ctc_scores != NULL: local[:, 0] = -1e10; local += ctc_weight * ctc_scores;]
 *   [first_position: local of hypotheses 1.. of every utterance = -1e10]
 *   eos_vscores = local[:, eos] + vscores;  acc = vscores + local with the eos column at -1e10;
 *   best_scores, best_ids = topk(acc per utterance over beam * V, k)   (sorted; ties in index order)
 *   tokens = best_ids % V;  parent_rows = best_ids / V + utterance * beam.
 * att_logp / ctc_scores: (batch*beam, V) fp32 contiguous device memory; vscores, eos_vscores: (batch, beam) fp32;
 * best_scores (batch, k) fp32; best_ids / tokens / parent_rows (batch, k) int64.  1 <= k <= 128.  float32 operation
 * order as in the eager formula (no FMA contraction): the values are the reference's bit for bit. */
int b200_beam_prune(const float* att_logp, const float* ctc_scores, const float* vscores, int64_t batch,
                    int64_t beam, int64_t V, int64_t k, float att_weight, float ctc_weight, int eos,
                    int first_position, float* best_scores, int64_t* best_ids, float* eos_vscores,
                    int64_t* tokens, int64_t* parent_rows, void* stream);

/* The pre-beam of the search loops (rnn/decoders.py:816-822, batch_beam_search.py:228-238):
 *   local = att_weight * att_logp; local[:, blank] = logzero; ids = topk(local, k)[1]
 * for every row of att_logp (n_rows, V) fp32, in one kernel.  ids int64 (n_rows, k), sorted by
 * value (descending), equal values in index order.  k <= 128, V <= 51200. */
int b200_pre_beam_candidates(const float* att_logp, int64_t n_rows, int64_t V, int64_t k,
                             float att_weight, int blank, int64_t* ids, void* stream);

/* Developer switch: on != 0 runs the lattice sweeps of utterances with 256 <= Lmax + 1 <= 1024 in one CTA per
 * (utterance, direction) instead of a 4-CTA thread-block cluster (A/B timing; the tests run both). */
void b200_ctc_debug_single_cta_sweep(int on);

/* Developer switch: on != 0 makes b200_ctc_prefix_score use its fp32 score kernels instead of the
 * tensor-core one (A/B timing; the tests run both). */
void b200_ctc_debug_prefix_fp32_score(int on);

/* Number of utterances of the last _forward / _fwd_bwd call on `workspace` whose targets were
 * invalid: a label outside [0,V), a label equal to blank, or olen[b] > Lmax.  torch.nn.CTCLoss
 * raises for such input on the CPU (espnet2/asr/ctc.py:49 calls it); here the kernels clamp the
 * label / the length so that nothing is read or written out of bounds, give the utterance
 * nll = NaN and NaN gradient rows, and this call reports the count.  Synchronises `stream`
 * (one B-int device->host copy): meant for a debug / strict mode, not for every step.
 * n_redone_host (optional): utterances of a _fwd_bwd call whose lattice the float64 linear-domain sweeps could not
 * certify (dynamic range) and which the log-domain sweeps therefore recomputed. */
int b200_ctc_loss_status(const void* workspace, int64_t T, int64_t B, int64_t V, int64_t Lmax,
                         int32_t* n_invalid_host, int32_t* n_redone_host /* may be NULL */, void* stream);

/* grad_logits[t,b,:] *= gout[b] / expected (gout_n == B) or gout[0] / expected (gout_n == 1), in
 * place, for the autograd backward of a one-pass _fwd_bwd call whose upstream gradient was not
 * known at forward time: the forward writes expected * d nll_b / d logits, the backward calls this
 * with the real upstream gradient gout (a DEVICE pointer: no host synchronisation).  Rows whose
 * ratio is exactly 1 are not touched, so the common case (the trainer's scale equals the hint
 * given at forward time) costs one launch of a kernel that reads B floats.  ilen (may be NULL)
 * lets it skip the zero rows past the end of each utterance. */
int b200_ctc_grad_rescale(void* grad_logits, int dtype, int64_t T, int64_t B, int64_t V,
                          int64_t gstride_t, int64_t gstride_b,
                          const float* gout, int64_t gout_n, float expected,
                          const int64_t* ilen, void* stream);

/* ------------------------------------------------------------------------
 * CTC prefix scoring (CTCPrefixScoreTH)
 * ------------------------------------------------------------------------
 * x        fp32 (B,T,V) contiguous log-probabilities (CTC.log_softmax output).
 * xlens    int64 (B,).
 * n_hyps   hypotheses per utterance; hypothesis h belongs to utterance h / n_hyps.
 * r        fp32 (T,2,n_bh) forward variables: [:,0,:] non-blank, [:,1,:] blank.
 */

/* Pads x past xlens in place (x[b,t>=xlens[b],:] = logzero, blank column = 0),
 * fills r0 (T,2,B): r0[:,0,:] = logzero, r0[:,1,b] = cumsum_t x[b,t,blank],
 * and xmax (B,V): max_t x[b,t,v] (the per-token scale of b200_ctc_prefix_score). */
int b200_ctc_prefix_init(float* x, const int64_t* xlens,
                         int64_t B, int64_t T, int64_t V, int blank,
                         float* r0, float* xmax, void* stream);

/* One step of prefix scoring.
 * r_prev      (T,2,n_bh)   state of the n_bh prefixes
 * s_prev      (n_bh,) or NULL (== 0): log_psi of each prefix
 * last_ids    int64 (n_bh,) last label of each prefix
 * output_len  len(prefix) - 1 (without <sos>), identical for all prefixes
 * scoring_ids int64 (n_bh,snum) candidates per prefix, or NULL (snum = 0): all V
 * start, end  frame window (ctc_prefix_score.py:145-154)
 * scores      fp32 (n_bh,V) out: log_psi - s_prev; logzero - s_prev for tokens
 *             not in scoring_ids and for blank; eos gets r_sum[xlens-1].
 * log_psi     fp32 (n_bh,V) out (may alias nothing; needed by _select)
 * work        fp32 scratch of b200_ctc_prefix_work_floats(T, n_bh, V) floats
 *
 * The scores are computed as scaled linear sums (one pass over x for all hypotheses of an
 * utterance); (hypothesis, token) pairs whose sum falls below 2^-100 of its scale are redone
 * exactly in log space by a second kernel, so the result never degrades to the sentinel where
 * the reference returns a finite log-sum-exp.  Any n_hyps >= 1 is accepted (more than 16
 * hypotheses per utterance are scored in groups of 16). */
size_t b200_ctc_prefix_work_floats(int64_t T, int64_t n_bh, int64_t V);

int b200_ctc_prefix_score(const float* x, const float* xmax, const int64_t* xlens,
                          int64_t B, int64_t T, int64_t V, int64_t n_hyps,
                          int blank, int eos,
                          const float* r_prev, const float* s_prev,
                          const int64_t* last_ids, int64_t output_len,
                          const int64_t* scoring_ids, int64_t snum,
                          int64_t start, int64_t end,
                          float* scores, float* log_psi, float* work, void* stream);

/* The candidate (pre-beam) path from a column copy of x.  With scoring_ids the step reads, per
 * (hypothesis, candidate), one word per frame from rows V floats apart (ctc_prefix_score.py:159-163
 * gathers the same elements with index_select): one DRAM sector per word.
 * b200_ctc_prefix_columns writes x_cols (B, V, Tc), Tc = b200_ctc_prefix_column_stride(T), once per
 * batch of utterances AFTER b200_ctc_prefix_init (it must see the padded x);
 * b200_ctc_prefix_score_columns is b200_ctc_prefix_score with that copy: same arguments, same results
 * bit for bit, the candidates' sums read contiguous columns.  x_cols == NULL or snum == 0: exactly
 * b200_ctc_prefix_score. */
int64_t b200_ctc_prefix_column_stride(int64_t T);
int b200_ctc_prefix_columns(const float* x, int64_t B, int64_t T, int64_t V, float* x_cols, void* stream);
int b200_ctc_prefix_score_columns(const float* x, const float* x_cols, const float* xmax,
                                  const int64_t* xlens, int64_t B, int64_t T, int64_t V,
                                  int64_t n_hyps, int blank, int eos,
                                  const float* r_prev, const float* s_prev,
                                  const int64_t* last_ids, int64_t output_len,
                                  const int64_t* scoring_ids, int64_t snum,
                                  int64_t start, int64_t end,
                                  float* scores, float* log_psi, float* work, void* stream);

/* Forward variables of the survivors: for new hypothesis k (0..n_new) with
 * parent hypothesis parent[k] (index into the n_bh prefixes of the step that
 * produced `work`) and appended token token[k]:
 *   r_new[t,:,k] for t in [0,T)   and   s_new[k] = log_psi[parent[k], token[k]].
 * Uses the phi rows that b200_ctc_prefix_score left in `work`.
 * token == NULL: parent[] holds the ids selected by beam pruning in
 * (local hypothesis * V + token) space, n_new / B per utterance, as
 * index_select_state receives them (ctc_prefix_score.py:203-207); the kernel
 * derives parent and token itself. */
int b200_ctc_prefix_select(const float* x, int64_t B, int64_t T, int64_t V,
                           int64_t n_hyps, int blank,
                           const float* r_prev, const float* log_psi,
                           const int64_t* last_ids, int64_t output_len,
                           int64_t start, int64_t end, int64_t n_bh,
                           const int64_t* parent, const int64_t* token, int64_t n_new,
                           const float* work,
                           float* r_new, float* s_new, void* stream);

/* ------------------------------------------------------------------------
 * Greedy path and forced alignment
 * ------------------------------------------------------------------------ */

/* ids[b,t] = argmax_v logits[b,t,v] (first maximum; a NaN counts as the maximum,
 * as torch.argmax), int64 (B,T) contiguous, for a (B,T,V) view with element
 * strides (stride_b, stride_t, 1). */
int b200_ctc_greedy(const void* logits, int dtype, int64_t B, int64_t T, int64_t V,
                    int64_t stride_b, int64_t stride_t, int64_t* ids, void* stream);

/* Viterbi alignment of one utterance: lpz fp32 (T,V) log-probs, y int64 (L,),
 * align int64 (T,) out: the label (or blank) emitted at each frame.
 * Scores are float64 and candidates are compared in the reference's order
 * (first maximum kept), so the alignment is the reference's bit for bit.
 * work: T * (2L+1) int16 back-pointers (b200_ctc_viterbi_work_bytes). */
size_t b200_ctc_viterbi_work_bytes(int64_t T, int64_t L);
int b200_ctc_viterbi(const float* lpz, int64_t T, int64_t V,
                     const int64_t* y, int64_t L, int blank,
                     int64_t* align, void* work, size_t work_bytes, void* stream);

/* ------------------------------------------------------------------------
 * Per-step loss exchange between the GPUs of one box
 * ------------------------------------------------------------------------
 * Replaces, for the (sum of nll, number of utterances) pair, the all-reduces of
 * espnet2/torch_utils/recursive_op.py:24,50 that espnet2/train/trainer.py:386-393 issues every iteration.
 * ring_ptrs_host[r] (HOST array of `world` device addresses): rank r's slot ring of
 * b200_peer_exchange_ring_floats(world) floats in peer-accessible (symmetric) memory, zeroed before the first
 * call.  local_sum * sum_scale: this rank's sum of nll (local_sum: device fp32, e.g. the loss_sum output of
 * b200_ctc_loss_fwd_bwd or the module's mean loss with sum_scale = B); local_count: this rank's number of
 * utterances.  pair: fp32 (4,) device buffer out: [0] total sum, [1] total count (summed in rank order: bitwise
 * identical on every rank), [2] global mean = [0]/[1], [3] local_count * world / [1] (what trainer.py:392-397's
 * "loss /= weight; loss *= world_size" multiplies the local mean loss by before backward).  seq: 1, 2, 3, ... the same on every rank for
 * the same step.  One kernel launch: peer
 * stores of the pair + sequence number into every ring, then a bounded wait (~2 s) on the local ring; if a
 * peer never arrives the totals are NaN and *status (device int32, may be NULL) is set to 1. */
size_t b200_peer_exchange_ring_floats(int64_t world);
int b200_peer_sum_exchange(const uint64_t* ring_ptrs_host, int64_t world, int64_t rank, int64_t seq,
                           const float* local_sum, float sum_scale, float local_count,
                           float* pair, int32_t* status, void* stream);

/* b200_ctc_viterbi without a caller-allocated work buffer.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  The back-pointer table lives in a
 * file-scope device array of B200_CTC_VITERBI_STATIC_BYTES (4 MiB) that every GPU loading the library
 * holds until exit; the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 * Limit: T * (2L+1) * 2 bytes <= B200_CTC_VITERBI_STATIC_BYTES (e.g. T=2000 with L<=523), else
 * B200_ERR_UNSUPPORTED_SHAPE.  All calls on a device share the array: calls that overlap on the GPU
 * (different streams) race; nothing in it needs clearing between calls.  Inputs, outputs and results are
 * those of b200_ctc_viterbi. */
#define B200_CTC_VITERBI_STATIC_BYTES (4u << 20)
size_t b200_ctc_viterbi_static_max_bytes(void);
int b200_ctc_viterbi_static(const float* lpz, int64_t T, int64_t V,
                            const int64_t* y, int64_t L, int blank,
                            int64_t* align, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ESPNET_B200_CTC_H_ */
