// ctc_prefix_score.inl -- the score and select kernels of CTC prefix scoring.
namespace b200 {

constexpr int kScoreThreads = 128;  // tokens per CTA
constexpr int kScoreTB = 64;        // frames of E staged in shared memory per block
constexpr int kScoreUnroll = 8;     // frames whose x loads are in flight per thread

struct ScoreParams {
  const float* x;       // (B,T,V) padded log-probs
  const float* xmax;    // (B,V)
  int B, T, V, n_hyps, n_bh, blank, eos;
  const float* r_prev;  // (T,2,n_bh)
  const float* s_prev;  // (n_bh) or null
  const int64_t* last_ids;
  int output_len, start, end;
  const int64_t* scoring_ids;
  int snum;
  float* scores;   // (n_bh,V)
  float* log_psi;  // (n_bh,V)
  PrefixWork w;
};

__device__ __forceinline__ float finish_psi(float acc, float scale_log, float init_term) {
  // log( acc * exp(scale_log) + exp(init_term) ), with acc == 0 meaning log(0)
  float psi = acc > 0.f ? scale_log + B200_LN2 * lg2f(acc) : B200_LZ;
  psi = fmaxf(psi, B200_LZ);
  if (init_term > B200_LZ) psi = lse2_nat(psi, init_term);
  return psi;
}

// All candidates: one thread per token, all hypotheses of the utterance at once.
template <int NH>
__global__ void __launch_bounds__(kScoreThreads) prefix_score_full_kernel(ScoreParams p) {
  __shared__ __align__(16) float sE[kScoreTB * NH];
  const int u = blockIdx.y;
  const int c = blockIdx.x * kScoreThreads + threadIdx.x;
  const bool live = c < p.V;
  const int cc = live ? c : p.V - 1;
  const int NHP = p.w.NHP;
  const float* xu = p.x + (int64_t)u * p.T * p.V + cc;
  const float mx = p.xmax[(int64_t)u * p.V + cc];
  const float mxl = mx * B200_LOG2E;
  float acc[NH];
#pragma unroll
  for (int j = 0; j < NH; ++j) acc[j] = 0.f;

  for (int tb = p.start; tb < p.end; tb += kScoreTB) {
    const int nt = min(kScoreTB, p.end - tb);
    __syncthreads();
    // E rows for frames tb-1 .. tb+nt-2 (term t uses phi[t-1])
    for (int i = threadIdx.x; i < nt * NH; i += kScoreThreads) {
      const int f = i / NH, j = i - f * NH;
      sE[i] = j < NHP ? p.w.EA[((int64_t)u * p.T + (tb - 1 + f)) * NHP + j] : 0.f;
    }
    __syncthreads();
    for (int f0 = 0; f0 < nt; f0 += kScoreUnroll) {
      float xv[kScoreUnroll];
#pragma unroll
      for (int k = 0; k < kScoreUnroll; ++k) {
        const int f = f0 + k;
        xv[k] = f < nt ? ld_nc_f32(xu + (int64_t)(tb + f) * p.V) : B200_LZ;
      }
#pragma unroll
      for (int k = 0; k < kScoreUnroll; ++k) {
        const int f = f0 + k;
        if (f < nt) {
          const float xh = ex2f(fmaf(xv[k], B200_LOG2E, -mxl));
          const float4* e4 = reinterpret_cast<const float4*>(sE + f * NH);
#pragma unroll
          for (int j4 = 0; j4 < NH / 4; ++j4) {
            const float4 e = e4[j4];
            acc[4 * j4 + 0] = fmaf(e.x, xh, acc[4 * j4 + 0]);
            acc[4 * j4 + 1] = fmaf(e.y, xh, acc[4 * j4 + 1]);
            acc[4 * j4 + 2] = fmaf(e.z, xh, acc[4 * j4 + 2]);
            acc[4 * j4 + 3] = fmaf(e.w, xh, acc[4 * j4 + 3]);
          }
        }
      }
    }
  }
  if (!live) return;
  const float x0 = (p.output_len == 0) ? p.x[(int64_t)u * p.T * p.V + c] : B200_LZ;  // r[0,n] = x[0,c]
#pragma unroll
  for (int j = 0; j < NH; ++j) {
    if (j < p.n_hyps) {
      const int h = u * p.n_hyps + j;
      float psi;
      if (c == p.blank) {
        psi = B200_LZ;
      } else if (c == p.eos) {
        psi = p.w.REND[h];
      } else if ((int64_t)c == p.last_ids[h]) {
        // the hypothesis' own last label: phi is r_prev[:,b] only (rare: one token per hypothesis)
        float a = 0.f;
        for (int t = p.start; t < p.end; ++t)
          a = fmaf(p.w.EB[((int64_t)u * p.T + t - 1) * NHP + j],
                   ex2f(fmaf(xu[(int64_t)t * p.V], B200_LOG2E, -mxl)), a);
        psi = finish_psi(a, p.w.MB[h] + mx, x0);
      } else {
        psi = finish_psi(acc[j], p.w.MA[h] + mx, x0);
      }
      const float sp = p.s_prev ? p.s_prev[h] : 0.f;
      p.log_psi[(int64_t)h * p.V + c] = psi;
      p.scores[(int64_t)h * p.V + c] = psi - sp;
    }
  }
}

// Pre-beam subset (scoring_ids): rows are first filled with logzero, then one thread per
// (hypothesis, candidate) pair.
__global__ void prefix_fill_kernel(ScoreParams p) {
  const int h = blockIdx.y;
  const int u = h / p.n_hyps;
  const float sp = p.s_prev ? p.s_prev[h] : 0.f;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < p.V; c += gridDim.x * blockDim.x) {
    float psi = B200_LZ;
    if (c == p.eos && c != p.blank) psi = p.w.REND[h];
    p.log_psi[(int64_t)h * p.V + c] = psi;
    p.scores[(int64_t)h * p.V + c] = psi - sp;
  }
  (void)u;
}

__global__ void prefix_score_partial_kernel(ScoreParams p) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= p.n_bh * p.snum) return;
  const int h = idx / p.snum;
  const int u = h / p.n_hyps, j = h - u * p.n_hyps;
  const int c = (int)p.scoring_ids[idx];
  if (c < 0 || c >= p.V || c == p.blank || c == p.eos) return;  // filled already
  const int NHP = p.w.NHP;
  const bool is_last = ((int64_t)c == p.last_ids[h]);
  const float* E = (is_last ? p.w.EB : p.w.EA) + (int64_t)u * p.T * NHP + j;
  const float* xu = p.x + (int64_t)u * p.T * p.V + c;
  const float mx = p.xmax[(int64_t)u * p.V + c];
  const float mxl = mx * B200_LOG2E;
  float a = 0.f;
  for (int t0 = p.start; t0 < p.end; t0 += kScoreUnroll) {
    float xv[kScoreUnroll], ev[kScoreUnroll];
#pragma unroll
    for (int k = 0; k < kScoreUnroll; ++k) {
      const int t = t0 + k;
      const bool ok = t < p.end;
      xv[k] = ok ? ld_nc_f32(xu + (int64_t)t * p.V) : B200_LZ;
      ev[k] = ok ? E[(int64_t)(t - 1) * NHP] : 0.f;
    }
#pragma unroll
    for (int k = 0; k < kScoreUnroll; ++k) a = fmaf(ev[k], ex2f(fmaf(xv[k], B200_LOG2E, -mxl)), a);
  }
  const float x0 = (p.output_len == 0) ? xu[0] : B200_LZ;
  const float psi = finish_psi(a, (is_last ? p.w.MB[h] : p.w.MA[h]) + mx, x0);
  const float sp = p.s_prev ? p.s_prev[h] : 0.f;
  p.log_psi[(int64_t)h * p.V + c] = psi;
  p.scores[(int64_t)h * p.V + c] = psi - sp;
}

// Forward variables of the survivors, ctc_prefix_score.py:123-131,157-162 in log space.
struct SelectParams {
  const float* x;
  int B, T, V, n_hyps, blank, n_bh, n_new;
  const float* r_prev;
  const float* log_psi;
  const int64_t* last_ids;
  int output_len, start, end;
  const int64_t* parent;
  const int64_t* token;
  float* r_new;  // (T,2,n_new)
  float* s_new;  // (n_new)
};

constexpr int kSelUnroll = 8;

__global__ void prefix_select_kernel(SelectParams p) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= p.n_new) return;
  const int h = (int)p.parent[k], c = (int)p.token[k];
  const int u = h / p.n_hyps;
  const bool is_last = ((int64_t)c == p.last_ids[h]);
  const float* xu = p.x + (int64_t)u * p.T * p.V;
  float pn = B200_LZ, pb = B200_LZ;
  // frames before the window keep log(0), except r[0,n] = x[0,c] for the first label (:130-131)
  for (int t = 0; t < min(p.start, p.T); ++t) {
    pn = (t == 0 && p.output_len == 0) ? xu[c] : B200_LZ;
    pb = B200_LZ;
    p.r_new[((int64_t)t * 2 + 0) * p.n_new + k] = pn;
    p.r_new[((int64_t)t * 2 + 1) * p.n_new + k] = pb;
  }
  for (int t0 = p.start; t0 < p.end; t0 += kSelUnroll) {
    float xc[kSelUnroll], xb[kSelUnroll], rn[kSelUnroll], rb[kSelUnroll];
#pragma unroll
    for (int i = 0; i < kSelUnroll; ++i) {
      const int t = min(t0 + i, p.end - 1);
      xc[i] = ld_nc_f32(xu + (int64_t)t * p.V + c);
      xb[i] = ld_nc_f32(xu + (int64_t)t * p.V + p.blank);
      rn[i] = p.r_prev[((int64_t)(t - 1) * 2 + 0) * p.n_bh + h];
      rb[i] = p.r_prev[((int64_t)(t - 1) * 2 + 1) * p.n_bh + h];
    }
#pragma unroll
    for (int i = 0; i < kSelUnroll; ++i) {
      const int t = t0 + i;
      if (t < p.end) {
        const float phi = is_last ? rb[i] : lse2_nat(rn[i], rb[i]);
        const float nn = lse2_nat(pn, phi) + xc[i];
        const float nb = lse2_nat(pn, pb) + xb[i];
        pn = nn; pb = nb;
        p.r_new[((int64_t)t * 2 + 0) * p.n_new + k] = pn;
        p.r_new[((int64_t)t * 2 + 1) * p.n_new + k] = pb;
      }
    }
  }
  for (int t = max(p.end, p.start); t < p.T; ++t) {
    p.r_new[((int64_t)t * 2 + 0) * p.n_new + k] = B200_LZ;
    p.r_new[((int64_t)t * 2 + 1) * p.n_new + k] = B200_LZ;
  }
  p.s_new[k] = p.log_psi[(int64_t)h * p.V + c];
}

}  // namespace b200
