// beam_prune.cu -- score accumulation and per-utterance top-k of the joint CTC/attention search loop.
//
// One call replaces, per output position of Decoder.recognize_beam_batch (rnn/decoders.py:812-846):
//     local  = att_weight * log_softmax(decoder)           (:812)
//     local[:, 0] = logzero; local += ctc_weight * ctc     (:816, :829-833)   [when a CTC scorer is attached]
//     local[:, 1:, :] = logzero at position 0              (:836-838)
//     eos_vscores = local[:, :, eos] + vscores             (:841)
//     vscores = vscores + local, with the eos column at logzero; topk(beam) per utterance   (:842-846)
// i.e. ~10 elementwise launches over (batch*beam, V) and torch.topk over rows of beam*V floats (a multi-pass radix
// select of its own: ~76 us for 32 x 50 000 on a B200) -- here ONE kernel, one 4-CTA cluster per utterance: the values are
// recomputed from the two score matrices on every pass (they stay in L2) with the reference's float32 operation
// order (mul, mul, add, add: no FMA contraction, so the values and therefore the selected set are those of the
// eager formula), the k-th largest value is found by a 4 x 8-bit radix select on order-preserving integer keys,
// ties at the threshold are taken in index order, and the k winners leave sorted (value descending, index ascending).
#include <cooperative_groups.h>

#include "common.cuh"

namespace b200 {
namespace cg = cooperative_groups;

constexpr int kPruneThreads = 1024;
constexpr int kPruneCtas = 4;   // CTAs (one cluster) per utterance
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

// A cluster of kPruneCtas CTAs per utterance: CTA r owns the r-th contiguous quarter of the utterance's beam * V elements
// (32 utterances alone would use 32 of 148 SMs, and every pass is latency per warp, not traffic), the digit histograms
// are merged through distributed shared memory, every CTA takes the same decision from the merged histogram, and the
// winners are collected in CTA 0's shared memory.
__global__ void __cluster_dims__(kPruneCtas, 1, 1) __launch_bounds__(kPruneThreads) beam_prune_kernel(PruneParams p) {
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ int s_hist[256];     // this CTA's histogram of the current digit
  __shared__ int s_tot[256];      // the cluster's
  __shared__ uint32_t s_prefix;   // the bits of the k-th largest key decided so far
  __shared__ int s_need;          // how many of the elements that match the prefix are still wanted
  __shared__ int s_scan[kPruneThreads / 32];
  __shared__ int s_ties;          // ties at the threshold inside this CTA's range
  __shared__ int s_cnt;           // (CTA 0's is the cluster's) elements above the threshold collected so far
  __shared__ float s_val[kPruneMaxK];
  __shared__ int s_idx[kPruneMaxK];
  const int r = (int)cluster.block_rank();
  const int u = blockIdx.x / kPruneCtas, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = p.beam * p.V;
  if (tid == 0) { s_prefix = 0; s_need = p.k; s_cnt = 0; }
  // eos scores of the utterance's hypotheses (:841)
  if (r == 0)
    for (int h = tid; h < p.beam; h += kPruneThreads)
      p.eos_v[u * p.beam + h] = __fadd_rn(prune_local(p, u * p.beam + h, h, p.eos), p.vs[u * p.beam + h]);

  // Every warp owns a contiguous range of the CTA's elements and walks it 32 elements at a time, four such rows per
  // trip with their eight loads issued before the first is used (the two score matrices sit in L2).
  constexpr int kWarps = kPruneThreads / 32;
  const int per_cta = (((n + kPruneCtas - 1) / kPruneCtas) + 31) & ~31;
  const int c0 = min(r * per_cta, n), c1 = min(c0 + per_cta, n);
  const int per_warp = (((c1 - c0 + kWarps - 1) / kWarps) + 31) & ~31;
  const int w0 = min(c0 + warp * per_warp, c1), w1 = min(w0 + per_warp, c1);
  auto walk = [&](auto&& visit) {  // visit(valid, j, value): called by all 32 lanes together, j ascending
    // (hypothesis, token) of the lane's element: one division per warp range, increments after that
    int hh = 0, cc = 0;
    if (w0 < w1) {
      const int j0 = min(w0 + lane, w1 - 1);
      hh = j0 / p.V;
      cc = j0 - hh * p.V;
    }
    for (int base = w0; base < w1; base += 128) {
      float a[4], c4[4], vs[4];
      int jj[4];
      bool ok[4], eosc[4], zero[4], dead[4];
#pragma unroll
      for (int x = 0; x < 4; ++x) {
        const int j = base + 32 * x + lane;
        jj[x] = j;
        ok[x] = j < w1;
        // (hh, cc) belong to min(j, the last valid element this lane reached): an invalid lane re-reads a valid address
        const int h = hh, c = cc;
        if (j + 32 < w1) {
          cc += 32;
          while (cc >= p.V) { cc -= p.V; ++hh; }
        }
        const int64_t o = (int64_t)(u * p.beam + h) * p.V + c;
        a[x] = p.att[o];
        c4[x] = p.ctc ? p.ctc[o] : 0.f;
        vs[x] = p.vs[u * p.beam + h];
        eosc[x] = c == p.eos;
        zero[x] = c == 0;
        dead[x] = p.first && h >= 1;
      }
#pragma unroll
      for (int x = 0; x < 4; ++x) {
        // the reference's operations in the reference's order (prune_local / prune_value)
        float v = __fmul_rn(p.att_w, a[x]);
        if (p.ctc) {
          if (zero[x]) v = B200_PRUNE_LZ;
          v = __fadd_rn(v, __fmul_rn(p.ctc_w, c4[x]));
        }
        if (dead[x]) v = B200_PRUNE_LZ;
        v = eosc[x] ? B200_PRUNE_LZ : __fadd_rn(vs[x], v);
        visit(ok[x], jj[x], v);
      }
    }
  };

  // ---- radix select: after pass d the top 8*(d+1) bits of the k-th largest key are known
  for (int d = 0; d < 4; ++d) {
    const int shift = 24 - 8 * d;
    for (int i = tid; i < 256; i += kPruneThreads) s_hist[i] = 0;
    __syncthreads();
    const uint32_t prefix = s_prefix;
    const uint32_t mask = d == 0 ? 0u : (0xffffffffu << (shift + 8));
    walk([&](bool ok, int, float v) {
      const uint32_t key = prune_key(v);
      const bool act = ok && (key & mask) == prefix;
      // scores of one utterance crowd into a few buckets: one shared-memory atomic per distinct bucket of the warp
      const uint32_t bucket = act ? ((key >> shift) & 255u) : 256u;
      const unsigned peers = __match_any_sync(0xffffffffu, bucket);
      if (act && lane == __ffs(peers) - 1) atomicAdd(&s_hist[bucket], __popc(peers));
    });
    cluster.sync();  // the histograms of all CTAs are complete
    if (tid < 256) {
      int t = 0;
#pragma unroll
      for (int rr = 0; rr < kPruneCtas; ++rr) t += *cluster.map_shared_rank(&s_hist[tid], rr);
      s_tot[tid] = t;
    }
    __syncthreads();
    if (warp == 0) {
      // the bucket that holds the s_need-th largest of the matching elements: suffix sums from bucket 255 down
      int need = s_need;
      int c8[8];
      int tot = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { c8[j] = s_tot[255 - (lane * 8 + j)]; tot += c8[j]; }   // lane 0 holds the largest buckets
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
    cluster.sync();  // every CTA has read this histogram (it is zeroed next); s_prefix / s_need are set
  }
  const uint32_t thr = s_prefix;   // key of the k-th largest value
  const int ties_wanted = s_need;  // how many elements EQUAL to it belong to the top k (>= 1)

  // ---- collect into CTA 0: everything above the threshold, and the first `ties_wanted` ties in index order -- the
  //      ranges of the CTAs and of their warps are contiguous and ascending, so "in index order" is the ties of the
  //      CTAs before this one, plus those of the warps before this one, plus the position inside the warp's own walk.
  float* val0 = cluster.map_shared_rank(s_val, 0);
  int* idx0 = cluster.map_shared_rank(s_idx, 0);
  int* cnt0 = cluster.map_shared_rank(&s_cnt, 0);
  int my_ties = 0;  // (warp-uniform)
  walk([&](bool ok, int j, float v) {
    const uint32_t key = prune_key(v);
    if (ok && key > thr) {
      const int slot = atomicAdd(cnt0, 1);
      if (slot < kPruneMaxK) { val0[slot] = v; idx0[slot] = j; }
    }
    my_ties += __popc(__ballot_sync(0xffffffffu, ok && key == thr));
  });
  if (lane == 0) s_scan[warp] = my_ties;
  __syncthreads();
  if (warp == 0) {
    int w = s_scan[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    s_scan[lane] = w;   // inclusive over warps
    if (lane == 31) s_ties = w;
  }
  cluster.sync();  // all elements above the threshold are in CTA 0; the CTAs' tie counts are visible
  int rank = warp ? s_scan[warp - 1] : 0;   // ties in the ranges before this warp's
  for (int rr = 0; rr < r; ++rr) rank += *cluster.map_shared_rank(&s_ties, rr);
  const int above = *cnt0;
  if (my_ties && rank < ties_wanted) {
    walk([&](bool ok, int j, float v) {
      const bool is_tie = ok && prune_key(v) == thr;
      const unsigned tie = __ballot_sync(0xffffffffu, is_tie);
      const int mine = rank + __popc(tie & ((1u << lane) - 1u));
      if (is_tie && mine < ties_wanted) {
        const int slot = above + mine;
        if (slot < kPruneMaxK) { val0[slot] = v; idx0[slot] = j; }
      }
      rank += __popc(tie);
    });
  }
  cluster.sync();  // CTA 0 holds the k winners; the other CTAs are done (nobody reads their shared memory any more)
  if (r != 0) return;
  // ---- the k winners, sorted: value descending, index ascending (rank by counting)
  const int k = min(p.k, kPruneMaxK);
  if (tid < k) {
    const float v = s_val[tid];
    const int j = s_idx[tid];
    const uint32_t key = prune_key(v);
    int rk = 0;
    for (int o = 0; o < k; ++o) {
      const uint32_t ko = prune_key(s_val[o]);
      rk += (ko > key || (ko == key && s_idx[o] < j)) ? 1 : 0;
    }
    p.best_scores[(int64_t)u * p.k + rk] = v;
    p.best_ids[(int64_t)u * p.k + rk] = j;
    const int h = j / p.V;
    p.tok[(int64_t)u * p.k + rk] = j - h * p.V;
    p.vidx[(int64_t)u * p.k + rk] = h + u * p.beam;
  }
}

// ------------------------------------------------------------------------------------------------ pre-beam candidates
// decoders.py:816-822 / batch_beam_search.py:228-238: the k best tokens of every hypothesis by the attention score,
//     local = att_weight * logp; local[:, blank] = logzero; topk(local, k)[1]
// (three eager launches, ~100 us of host time per position).  One CTA per hypothesis: the row's order-preserving keys
// go to shared memory once, the 4 x 8-bit radix select and the collection run there; ties in index order, ids sorted
// by value (descending) then index.
constexpr int kCandThreads = 256;
__global__ void __launch_bounds__(kCandThreads) pre_beam_topk_kernel(const float* __restrict__ logp, int V, int k, float w, int blank,
                                                                     int64_t* __restrict__ ids) {
  extern __shared__ uint32_t s_key[];  // [V]
  __shared__ int s_hist[256];
  __shared__ uint32_t s_prefix;
  __shared__ int s_need, s_cnt;
  __shared__ int s_scan[kCandThreads / 32];
  __shared__ uint32_t s_wkey[kPruneMaxK];
  __shared__ int s_widx[kPruneMaxK];
  const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int kWarps = kCandThreads / 32;
  const float* x = logp + (int64_t)row * V;
  for (int c = tid; c < V; c += kCandThreads) {
    float v = __fmul_rn(w, x[c]);
    if (c == blank) v = B200_PRUNE_LZ;
    s_key[c] = prune_key(v);
  }
  if (tid == 0) { s_prefix = 0; s_need = k; s_cnt = 0; }
  // contiguous, ascending ranges per warp (so that "in index order" is a scan over the warps)
  const int per_warp = (((V + kWarps - 1) / kWarps) + 31) & ~31;
  const int w0 = min(warp * per_warp, V), w1 = min(w0 + per_warp, V);
  for (int d = 0; d < 4; ++d) {
    const int shift = 24 - 8 * d;
    s_hist[tid] = 0;  // (256 threads, 256 buckets)
    __syncthreads();
    const uint32_t prefix = s_prefix;
    const uint32_t mask = d == 0 ? 0u : (0xffffffffu << (shift + 8));
    for (int base = w0; base < w1; base += 32) {
      const int c = base + lane;
      const uint32_t key = c < w1 ? s_key[c] : 0u;
      const bool act = c < w1 && (key & mask) == prefix;
      const uint32_t bucket = act ? ((key >> shift) & 255u) : 256u;
      const unsigned peers = __match_any_sync(0xffffffffu, bucket);
      if (act && lane == __ffs(peers) - 1) atomicAdd(&s_hist[bucket], __popc(peers));
    }
    __syncthreads();
    if (warp == 0) {
      const int need = s_need;
      int c8[8];
      int tot = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { c8[j] = s_hist[255 - (lane * 8 + j)]; tot += c8[j]; }
      int incl = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      int before = incl - tot;
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
  const uint32_t thr = s_prefix;
  const int ties_wanted = s_need;
  int my_ties = 0;
  for (int base = w0; base < w1; base += 32) {
    const int c = base + lane;
    const uint32_t key = c < w1 ? s_key[c] : 0u;
    if (c < w1 && key > thr) {
      const int slot = atomicAdd(&s_cnt, 1);
      if (slot < kPruneMaxK) { s_wkey[slot] = key; s_widx[slot] = c; }
    }
    my_ties += __popc(__ballot_sync(0xffffffffu, c < w1 && key == thr));
  }
  if (lane == 0) s_scan[warp] = my_ties;
  __syncthreads();
  int rank = 0;
  for (int q = 0; q < warp; ++q) rank += s_scan[q];
  const int above = s_cnt;
  if (my_ties && rank < ties_wanted) {
    for (int base = w0; base < w1; base += 32) {
      const int c = base + lane;
      const bool is_tie = c < w1 && s_key[c] == thr;
      const unsigned tie = __ballot_sync(0xffffffffu, is_tie);
      const int mine = rank + __popc(tie & ((1u << lane) - 1u));
      if (is_tie && mine < ties_wanted) {
        const int slot = above + mine;
        if (slot < kPruneMaxK) { s_wkey[slot] = thr; s_widx[slot] = c; }
      }
      rank += __popc(tie);
    }
  }
  __syncthreads();
  if (tid < k) {
    const uint32_t key = s_wkey[tid];
    const int c = s_widx[tid];
    int r = 0;
    for (int o = 0; o < k; ++o) r += (s_wkey[o] > key || (s_wkey[o] == key && s_widx[o] < c)) ? 1 : 0;
    ids[(int64_t)row * k + r] = c;
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
  beam_prune_kernel<<<(unsigned)batch * kPruneCtas, kPruneThreads, 0, (cudaStream_t)stream>>>(p);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

extern "C" int b200_pre_beam_candidates(const float* att_logp, int64_t n_rows, int64_t V, int64_t k, float att_weight, int blank,
                                        int64_t* ids, void* stream) {
  B200_CHECK_ARG(att_logp && ids, B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  B200_CHECK_ARG(n_rows >= 1 && V >= 1 && n_rows < (1ll << 31), B200_ERR_INVALID_ARGUMENT, "need n_rows>=1, V>=1, got n_rows=%lld V=%lld",
                 (long long)n_rows, (long long)V);
  B200_CHECK_ARG(k >= 1 && k <= kPruneMaxK && k <= V, B200_ERR_UNSUPPORTED_SHAPE, "k=%lld must lie in [1, min(%d, V)]", (long long)k,
                 kPruneMaxK);
  const size_t smem = sizeof(uint32_t) * (size_t)V;
  B200_CHECK_ARG(smem <= 200u * 1024u, B200_ERR_UNSUPPORTED_SHAPE, "V=%lld: a row of keys does not fit shared memory", (long long)V);
  if (smem > 40u * 1024u) B200_CUDA(set_max_smem_cached((const void*)pre_beam_topk_kernel, (int)smem));
  pre_beam_topk_kernel<<<(unsigned)n_rows, kCandThreads, smem, (cudaStream_t)stream>>>(att_logp, (int)V, (int)k, att_weight, blank, ids);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}
