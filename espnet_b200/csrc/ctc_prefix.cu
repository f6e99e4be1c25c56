// ctc_prefix.cu -- CTC prefix scoring (CTCPrefixScoreTH) for sm_100a.
//
// Replaces espnet/nets/ctc_prefix_score.py:23-221.  The reference materialises, for every
// (hypothesis, candidate) pair, the forward variables r (T,2) -- 6.4 GB per step at
// B=32, beam=10, T=500, V=5000 -- although the score itself does not depend on them:
//
//   log_psi(h,c) = LSE( r[start-1,n],  { phi_h(c)[t-1] + x[t,c] : t in [start,end) } )   (:165-180)
//   phi_h(c)[t]  = r_prev[t,b,h]                      if c == last label of h
//                = LSE(r_prev[t,n,h], r_prev[t,b,h])  otherwise                          (:133-142)
//
// so scoring all V candidates of all hypotheses of an utterance is one pass over the
// utterance's log-probs x[u,:,:] (read ONCE for the whole beam), and the forward variables
// are recomputed afterwards for the few (hypothesis, token) pairs that survive pruning.
//
//   prefix_init    pads x past xlens in place (:46-51), r0 = cumsum of blank (:84-92),
//                  xmax[u,c] = max_t x[u,t,c] (per-token scale of the score kernel)
//   prefix_phi     per hypothesis: phi rows, their maxima, scaled linear copies E = exp(phi - max)
//   prefix_score   per (utterance, token tile): sum_t E[h][t-1] * exp(x[t,c] - xmax[c]) for all
//                  hypotheses h of the utterance at once (fp32, terms in [0,1])
//   prefix_select  log-space r recursion (:157-162) for the survivors only
#include "common.cuh"

namespace b200 {

#define B200_LZ B200_CTC_LOGZERO

__device__ __forceinline__ float lse2_nat(float a, float b) {
  const float m = fmaxf(a, b);
  const float d = fminf(a, b) - m;  // <= 0; exp underflows to 0 for logzero operands
  return m + B200_LN2 * lg2f(1.0f + ex2f(d * B200_LOG2E));
}

// ------------------------------------------------------------------ init
__global__ void prefix_init_kernel(float* __restrict__ x, const int64_t* __restrict__ xlens, int T, int V,
                                   int B, int blank, float* __restrict__ r0, float* __restrict__ xmax) {
  const int u = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= V) return;
  const int xl = (int)xlens[u];
  float* col = x + (int64_t)u * T * V + c;
  float m = -3.0e38f, cum = 0.f;
  const bool is_blank = (c == blank);
  for (int t = 0; t < T; ++t) {
    float v;
    if (t >= xl) {
      v = is_blank ? 0.0f : B200_LZ;
      col[(int64_t)t * V] = v;
    } else {
      v = col[(int64_t)t * V];
    }
    m = fmaxf(m, v);
    if (is_blank) {
      cum += v;
      r0[((int64_t)t * 2 + 0) * B + u] = B200_LZ;
      r0[((int64_t)t * 2 + 1) * B + u] = cum;
    }
  }
  xmax[(int64_t)u * V + c] = m;
}

// ------------------------------------------------------------------ work layout
struct PrefixWork {
  float* EA;    // [B][T][NHP]  exp(r_sum[t]   - MA[h])   (c != last)
  float* EB;    // [B][T][NHP]  exp(r_prev_b[t] - MB[h])  (c == last)
  float* MA;    // [n_bh]
  float* MB;    // [n_bh]
  float* REND;  // [n_bh]  r_sum[xlen-1]
  int* fix_count;  // [1]        pairs queued for the exact (log-space) kernel, zeroed by the phi kernel
  int* fix_list;   // [n_bh * V] hypothesis * V + token
  int NHP;
};

__host__ __device__ inline int nh_pad(int n_hyps) { return (n_hyps + 3) & ~3; }

__host__ __device__ inline PrefixWork carve_prefix(float* work, int64_t B, int64_t T, int64_t n_hyps) {
  PrefixWork w;
  w.NHP = nh_pad((int)n_hyps);
  const int64_t n = B * T * w.NHP;
  w.EA = work;
  w.EB = work + n;
  w.MA = work + 2 * n;
  w.MB = w.MA + B * n_hyps;
  w.REND = w.MB + B * n_hyps;
  w.fix_count = reinterpret_cast<int*>(w.REND + B * n_hyps);
  w.fix_list = w.fix_count + 1;
  return w;
}

// ------------------------------------------------------------------ phi
// One CTA per hypothesis.  E rows are written for t in [start-1, end-1) (the only ones the
// score kernel reads); everything is relative to the maximum over that window.
__global__ void prefix_phi_kernel(const float* __restrict__ r_prev, const int64_t* __restrict__ xlens, int T,
                                  int n_hyps, int n_bh, int start, int end, PrefixWork w) {
  const int h = blockIdx.x;
  const int u = h / n_hyps, j = h - u * n_hyps;
  if (h == 0 && threadIdx.x == 0) *w.fix_count = 0;
  __shared__ float red[2][32];
  float ma = -3.0e38f, mb = -3.0e38f;
  for (int t = start - 1 + threadIdx.x; t < end - 1; t += blockDim.x) {
    const float rn = r_prev[((int64_t)t * 2 + 0) * n_bh + h];
    const float rb = r_prev[((int64_t)t * 2 + 1) * n_bh + h];
    ma = fmaxf(ma, lse2_nat(rn, rb));
    mb = fmaxf(mb, rb);
  }
  ma = warp_max(ma);
  mb = warp_max(mb);
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = ma; red[1][threadIdx.x >> 5] = mb; }
  __syncthreads();
  const int nw = blockDim.x >> 5;
  ma = warp_max((int)(threadIdx.x & 31) < nw ? red[0][threadIdx.x & 31] : -3.0e38f);
  mb = warp_max((int)(threadIdx.x & 31) < nw ? red[1][threadIdx.x & 31] : -3.0e38f);
  for (int t = start - 1 + threadIdx.x; t < end - 1; t += blockDim.x) {
    const float rn = r_prev[((int64_t)t * 2 + 0) * n_bh + h];
    const float rb = r_prev[((int64_t)t * 2 + 1) * n_bh + h];
    const int64_t o = ((int64_t)u * T + t) * w.NHP + j;
    w.EA[o] = ex2f((lse2_nat(rn, rb) - ma) * B200_LOG2E);
    w.EB[o] = ex2f((rb - mb) * B200_LOG2E);
  }
  if (threadIdx.x == 0) {
    w.MA[h] = ma;
    w.MB[h] = mb;
    const int te = (int)xlens[u] - 1;
    w.REND[h] = lse2_nat(r_prev[((int64_t)te * 2 + 0) * n_bh + h], r_prev[((int64_t)te * 2 + 1) * n_bh + h]);
  }
}

}  // namespace b200

#include "ctc_prefix_score.inl"
#include "ctc_prefix_mma.inl"
#include "ctc_prefix_host.inl"
