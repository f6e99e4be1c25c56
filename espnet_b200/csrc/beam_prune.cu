// beam_prune.cu -- score accumulation and per-utterance top-k of the joint CTC/attention search loop.
//
// One call replaces, per output position of Decoder.recognize_beam_batch (rnn/decoders.py:812-846):
//     local  = att_weight * log_softmax(decoder)           (:812)
//     local[:, 0] = logzero; local += ctc_weight * ctc     (:816, :829-833)   [when a CTC scorer is attached]
//     local[:, 1:, :] = logzero at position 0              (:836-838)
//     eos_vscores = local[:, :, eos] + vscores             (:841)
//     vscores = vscores + local, with the eos column at logzero; topk(beam) per utterance   (:842-846)
// i.e. ~10 elementwise launches over (batch*beam, V) and torch.topk over rows of beam*V floats (a multi-pass radix
// select of its own: ~76 us for 32 x 50 000 on a B200) -- here ONE kernel, one CTA per utterance: the values are
// recomputed from the two score matrices on every pass (they stay in L2) with the reference's float32 operation
// order (mul, mul, add, add: no FMA contraction, so the values and therefore the selected set are those of the
// eager formula), the k-th largest value is found by a 4 x 8-bit radix select on order-preserving integer keys,
// ties at the threshold are taken in index order, and the k winners leave sorted (value descending, index ascending).
#include "common.cuh"

namespace b200 {

constexpr int kPruneThreads = 1024;
constexpr int kPruneMaxK = 128;
#define B200_PRUNE_LZ (-10000000000.0f)

struct PruneParams {
  const float* att;   // (batch*beam, V) log-probs of the attention decoder
  const float* ctc;   // (batch*beam, V) CTC prefix scores or null
  const float* vs;    // (batch, beam) accumulated scores
  int batch, beam, V, k, eos, first;
  float att_w, ctc_w;
  float* best_scores;   // (batch, k)
  int64_t* best_ids;    // (batch, k)  h * V + token
  float* eos_v;         // (batch, beam)
  int64_t* tok;         // (batch, k)  best_ids % V                      (:847-849)
  int64_t* vidx;        // (batch, k)  best_ids / V + u * beam: row of the parent hypothesis (:850-852)
};

// local score of (hypothesis h of the utterance, token c), the reference's operations in the reference's order
__device__ __forceinline__ float prune_local(const PruneParams& p, int row, int h, int c) {
  float v = __fmul_rn(p.att_w, p.att[(int64_t)row * p.V + c]);
  if (p.ctc) {
    if (c == 0) v = B200_PRUNE_LZ;
    v = __fadd_rn(v, __fmul_rn(p.ctc_w, p.ctc[(int64_t)row * p.V + c]));
  }
  if (p.first && h >= 1) v = B200_PRUNE_LZ;
  return v;
}
__device__ __forceinline__ float prune_value(const PruneParams& p, int u, int h, int c, float vsh) {
  if (c == p.eos) return B200_PRUNE_LZ;
  return __fadd_rn(vsh, prune_local(p, u * p.beam + h, h, c));
}
// order-preserving key: larger float <-> larger unsigned (NaN sorts above +inf, as torch.topk treats it)
__device__ __forceinline__ uint32_t prune_key(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__global__ void __launch_bounds__(kPruneThreads) beam_prune_kernel(PruneParams p) {
  __shared__ int s_hist[256];
  __shared__ uint32_t s_prefix;   // the bits of the k-th largest key decided so far
  __shared__ int s_need;          // how many of the elements that match the prefix are still wanted
  __shared__ int s_scan[kPruneThreads / 32];
  __shared__ int s_cnt;
  __shared__ float s_val[kPruneMaxK];
  __shared__ int s_idx[kPruneMaxK];
  const int u = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = p.beam * p.V;
  if (tid == 0) { s_prefix = 0; s_need = p.k; s_cnt = 0; }
  // eos scores of the utterance's hypotheses (:841)
  for (int h = tid; h < p.beam; h += kPruneThreads)
    p.eos_v[u * p.beam + h] = __fadd_rn(prune_local(p, u * p.beam + h, h, p.eos), p.vs[u * p.beam + h]);

  // ---- radix select: after pass d the top 8*(d+1) bits of the k-th largest key are known
  for (int d = 0; d < 4; ++d) {
    const int shift = 24 - 8 * d;
    for (int i = tid; i < 256; i += kPruneThreads) s_hist[i] = 0;
    __syncthreads();
    const uint32_t prefix = s_prefix;
    const uint32_t mask = d == 0 ? 0u : (0xffffffffu << (shift + 8));
    for (int h = 0; h < p.beam; ++h) {
      const float vsh = p.vs[u * p.beam + h];
      for (int c = tid; c < p.V; c += kPruneThreads) {
        const uint32_t key = prune_key(prune_value(p, u, h, c, vsh));
        if ((key & mask) == prefix) atomicAdd(&s_hist[(key >> shift) & 255u], 1);
      }
    }
    __syncthreads();
    if (warp == 0) {
      // the bucket that holds the s_need-th largest of the matching elements: suffix sums from bucket 255 down
      int need = s_need;
      int c8[8];
      int tot = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { c8[j] = s_hist[255 - (lane * 8 + j)]; tot += c8[j]; }   // lane 0 holds the largest buckets
      int incl = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      int before = incl - tot;   // elements in larger buckets than this lane's
      int found = -1, left = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (found < 0 && before < need && before + c8[j] >= need) { found = 255 - (lane * 8 + j); left = need - before; }
        before += c8[j];
      }
      const unsigned who = __ballot_sync(0xffffffffu, found >= 0);
      if (who) {
        const int src = __ffs(who) - 1;
        found = __shfl_sync(0xffffffffu, found, src);
        left = __shfl_sync(0xffffffffu, left, src);
        if (lane == 0) { s_prefix = prefix | ((uint32_t)found << shift); s_need = left; }
      }
    }
    __syncthreads();
  }
  const uint32_t thr = s_prefix;   // key of the k-th largest value
  const int ties_wanted = s_need;  // how many elements EQUAL to it belong to the top k (>= 1)

  // ---- collect: everything above the threshold, and the first `ties_wanted` ties in index order.  Every thread owns a
  //      contiguous range of indices so that "in index order" is a block-wide exclusive scan of the tie counts.
  const int per = (n + kPruneThreads - 1) / kPruneThreads;
  const int j0 = min(tid * per, n), j1 = min(j0 + per, n);
  int my_ties = 0;
  {
    int h = j0 / p.V, c = j0 - h * p.V;
    float vsh = h < p.beam ? p.vs[u * p.beam + h] : 0.f;
    for (int j = j0; j < j1; ++j) {
      const float v = prune_value(p, u, h, c, vsh);
      const uint32_t key = prune_key(v);
      if (key > thr) {
        const int slot = atomicAdd(&s_cnt, 1);
        if (slot < kPruneMaxK) { s_val[slot] = v; s_idx[slot] = j; }
      } else if (key == thr) {
        ++my_ties;
      }
      if (++c == p.V) { c = 0; ++h; if (h < p.beam) vsh = p.vs[u * p.beam + h]; }
    }
  }
  // exclusive scan of my_ties over the block
  int incl = my_ties;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_scan[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = s_scan[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    s_scan[lane] = w;   // inclusive over warps
  }
  __syncthreads();
  int rank = incl - my_ties + (warp ? s_scan[warp - 1] : 0);   // ties before this thread's range
  const int above = s_cnt;                                     // (all `key > thr` elements are in by the barriers above)
  if (my_ties && rank < ties_wanted) {
    int h = j0 / p.V, c = j0 - h * p.V;
    float vsh = p.vs[u * p.beam + h];
    for (int j = j0; j < j1 && rank < ties_wanted; ++j) {
      const float v = prune_value(p, u, h, c, vsh);
      if (prune_key(v) == thr) {
        const int slot = above + rank;
        if (slot < kPruneMaxK) { s_val[slot] = v; s_idx[slot] = j; }
        ++rank;
      }
      if (++c == p.V) { c = 0; ++h; if (h < p.beam) vsh = p.vs[u * p.beam + h]; }
    }
  }
  __syncthreads();
  // ---- the k winners, sorted: value descending, index ascending (rank by counting)
  const int k = min(p.k, kPruneMaxK);
  if (tid < k) {
    const float v = s_val[tid];
    const int j = s_idx[tid];
    const uint32_t key = prune_key(v);
    int r = 0;
    for (int o = 0; o < k; ++o) {
      const uint32_t ko = prune_key(s_val[o]);
      r += (ko > key || (ko == key && s_idx[o] < j)) ? 1 : 0;
    }
    p.best_scores[(int64_t)u * p.k + r] = v;
    p.best_ids[(int64_t)u * p.k + r] = j;
    const int h = j / p.V;
    p.tok[(int64_t)u * p.k + r] = j - h * p.V;
    p.vidx[(int64_t)u * p.k + r] = h + u * p.beam;
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_beam_prune(const float* att_logp, const float* ctc_scores, const float* vscores, int64_t batch, int64_t beam,
                               int64_t V, int64_t k, float att_weight, float ctc_weight, int eos, int first_position,
                               float* best_scores, int64_t* best_ids, float* eos_vscores, int64_t* tokens, int64_t* parent_rows,
                               void* stream) {
  B200_CHECK_ARG(att_logp && vscores && best_scores && best_ids && eos_vscores && tokens && parent_rows, B200_ERR_INVALID_ARGUMENT,
                 "null pointer argument");
  B200_CHECK_ARG(batch >= 1 && beam >= 1 && V >= 2 && beam * V < (1ll << 31), B200_ERR_INVALID_ARGUMENT,
                 "need batch>=1, beam>=1, V>=2, beam*V<2^31, got batch=%lld beam=%lld V=%lld", (long long)batch, (long long)beam,
                 (long long)V);
  B200_CHECK_ARG(k >= 1 && k <= kPruneMaxK && k <= beam * V, B200_ERR_UNSUPPORTED_SHAPE, "k=%lld must lie in [1, min(%d, beam*V)]",
                 (long long)k, kPruneMaxK);
  B200_CHECK_ARG(eos >= 0 && eos < V, B200_ERR_INVALID_ARGUMENT, "eos=%d outside [0,%lld)", eos, (long long)V);
  PruneParams p;
  p.att = att_logp; p.ctc = ctc_scores; p.vs = vscores;
  p.batch = (int)batch; p.beam = (int)beam; p.V = (int)V; p.k = (int)k; p.eos = eos; p.first = first_position ? 1 : 0;
  p.att_w = att_weight; p.ctc_w = ctc_weight;
  p.best_scores = best_scores; p.best_ids = best_ids; p.eos_v = eos_vscores; p.tok = tokens; p.vidx = parent_rows;
  beam_prune_kernel<<<(unsigned)batch, kPruneThreads, 0, (cudaStream_t)stream>>>(p);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}
