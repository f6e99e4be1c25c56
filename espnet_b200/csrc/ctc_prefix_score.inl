// ctc_prefix_score.inl -- the score and select kernels of CTC prefix scoring.
namespace b200 {

constexpr int kScoreThreads = 128;  // tokens per CTA
constexpr int kScoreTB = 128;       // frames of E staged in shared memory per block
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
  int hyp_off;     // the full-vocabulary kernels handle hypotheses [hyp_off, hyp_off + NH) of every utterance
  const float* xc; // (B,V,Tc) column copy of x (b200_ctc_prefix_columns) or null
  int Tc;
  const int64_t* xlens;  // (B)
};

// The scaled linear sum loses every term below 2^-126 of the largest representable one (flush to zero).
// With acc >= 2^-100 the at most T lost terms change log_psi by less than T * 2^-26 (< 1e-4 up to T = 6700);
// below that the pair is queued for prefix_score_fix_kernel, which redoes it in log space.
constexpr float kScoreAccFloor = 7.888609052210118e-31f;  // 2^-100

__device__ __forceinline__ float finish_psi(float acc, float scale_log, float init_term) {
  // log( acc * exp(scale_log) + exp(init_term) ), with acc == 0 meaning log(0)
  float psi = acc > 0.f ? scale_log + B200_LN2 * lg2f(acc) : B200_LZ;
  psi = fmaxf(psi, B200_LZ);
  if (init_term > B200_LZ) psi = lse2_nat(psi, init_term);
  return psi;
}
// queue (hypothesis h, token c) for the exact kernel
// (m: the maximum of the hypothesis' phi over the window.  At logzero -- a window that does not overlap the frames the
// prefix has reached -- the exact sum is logzero as well, which is what the scaled sum already returned: nothing to redo.
// Without this test such a hypothesis queued all V of its tokens: 724 us per position in the windowed search loop.)
__device__ __forceinline__ void score_flag(const ScoreParams& p, float acc, int h, int c, float m) {
  // (likewise a window that lies in the utterance's padding: every non-blank log-prob there is logzero)
  if (acc < kScoreAccFloor && p.start < p.end && m > 0.5f * B200_LZ && p.start < (int)p.xlens[h / p.n_hyps]) {
    const int k = atomicAdd(p.w.fix_count, 1);
    p.w.fix_list[k] = h * p.V + c;
  }
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
      const int f = i / NH, j = i - f * NH + p.hyp_off;
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
    if (j + p.hyp_off < p.n_hyps) {
      const int h = u * p.n_hyps + j + p.hyp_off;
      float psi;
      if (c == p.blank) {
        psi = B200_LZ;
      } else if (c == p.eos) {
        psi = p.w.REND[h];
      } else {
        // (the hypothesis' own last label is overwritten by prefix_score_last_kernel)
        psi = finish_psi(acc[j], p.w.MA[h] + mx, x0);
        if ((int64_t)c != p.last_ids[h]) score_flag(p, acc[j], h, c, p.w.MA[h]);
      }
      const float sp = p.s_prev ? p.s_prev[h] : 0.f;
      p.log_psi[(int64_t)h * p.V + c] = psi;
      p.scores[(int64_t)h * p.V + c] = psi - sp;
    }
  }
}

// The one candidate per hypothesis that repeats its last label: phi is r_prev[:,b] only
// (ctc_prefix_score.py:136-142).  Inside the per-token loop of the full-vocabulary kernels this was a
// 500-iteration serial loop in one thread of almost every CTA, which the whole CTA waited for; here one warp per
// hypothesis sums over the frames in parallel and writes the entry.
__device__ __forceinline__ void prefix_score_last_warp(const ScoreParams& p, int h, int lane) {
  const int64_t c64 = p.last_ids[h];
  if (c64 < 0 || c64 >= p.V || c64 == p.blank || c64 == p.eos) return;
  const int c = (int)c64;
  const int u = h / p.n_hyps, j = h - u * p.n_hyps;
  const int NHP = p.w.NHP;
  const float* xu = p.x + (int64_t)u * p.T * p.V + c;
  const float mx = p.xmax[(int64_t)u * p.V + c];
  const float mxl = mx * B200_LOG2E;
  float a = 0.f;
  for (int t = p.start + lane; t < p.end; t += 32)
    a = fmaf(p.w.EB[((int64_t)u * p.T + t - 1) * NHP + j], ex2f(fmaf(ld_nc_f32(xu + (int64_t)t * p.V), B200_LOG2E, -mxl)), a);
  a = warp_sum(a);
  if (lane == 0) {
    const float x0 = (p.output_len == 0) ? xu[0] : B200_LZ;
    const float psi = finish_psi(a, p.w.MB[h] + mx, x0);
    score_flag(p, a, h, c, p.w.MB[h]);
    const float sp = p.s_prev ? p.s_prev[h] : 0.f;
    p.log_psi[(int64_t)h * p.V + c] = psi;
    p.scores[(int64_t)h * p.V + c] = psi - sp;
  }
}
// (after the one-token kernel, which writes the generic value to that entry first)
__global__ void __launch_bounds__(128) prefix_score_last_kernel(ScoreParams p) {
  const int h = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (h < p.n_bh) prefix_score_last_warp(p, h, threadIdx.x & 31);
}

// The same for vocabularies with V % 4 == 0: FOUR consecutive tokens per thread.  The one-token kernel above
// is bound by the shared-memory pipe (NH/4 16-byte reads of E per token and frame: 7.5 M warp-wide LDS.128
// per step at C3); here every E value feeds four accumulators.  x is read ONCE per step for the whole beam,
// 16 bytes per thread and frame, through a thread-private ring in shared memory filled with cp.async
// (kScore4Ring batches of 8 frames; a thread only ever reads what it copied itself, so cp.async.wait_group is the
// only synchronisation): with the loads in registers (two sets of 8) ncu showed 40 % of the stall samples on
// the first use of a batch and 28 % of the DRAM throughput -- the bytes in flight per SM, not the arithmetic,
// bound the kernel.
constexpr int kScore4Threads = 64;
constexpr int kScore4Batch = 8;  // frames per cp.async group
constexpr int kScore4Ring = 4;   // groups in the ring: kScore4Ring - 1 are in flight while one is consumed

__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <int NH>
__global__ void __launch_bounds__(kScore4Threads) prefix_score_full4_kernel(ScoreParams p) {
  __shared__ __align__(16) float sE[kScoreTB * NH];
  __shared__ __align__(16) uint4 sX[kScore4Ring * kScore4Batch * kScore4Threads];
  if (blockIdx.y == (unsigned)p.B) {
    // one more row of CTAs: the repeated-last-label entries, a warp per hypothesis (the token CTAs skip them)
    if (p.hyp_off != 0) return;  // (done by the launch of the first hypothesis group)
    const int nw = gridDim.x * (kScore4Threads / 32);
    for (int h = blockIdx.x * (kScore4Threads / 32) + (threadIdx.x >> 5); h < p.n_bh; h += nw)
      prefix_score_last_warp(p, h, threadIdx.x & 31);
    return;
  }
  const int u = blockIdx.y;
  const int c0 = 4 * (blockIdx.x * kScore4Threads + threadIdx.x);
  const bool live = c0 < p.V;  // V % 4 == 0: all four tokens or none
  const int cc = live ? c0 : p.V - 4;
  const int NHP = p.w.NHP;
  const float* xu = p.x + (int64_t)u * p.T * p.V + cc;
  const float4 mx4 = *reinterpret_cast<const float4*>(p.xmax + (int64_t)u * p.V + cc);
  const float mx[4] = {mx4.x, mx4.y, mx4.z, mx4.w};
  float mxl[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) mxl[q] = mx[q] * B200_LOG2E;
  float acc[4][NH];
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int j = 0; j < NH; ++j) acc[q][j] = 0.f;

  // batch m = frames start + 8 m .. start + 8 m + 7 (clamped), in ring slot m % kScore4Ring
  const int nbatch = (max(p.end - p.start, 0) + kScore4Batch - 1) / kScore4Batch;
  uint4* myx = sX + threadIdx.x;
  auto issue = [&](int m) {
    if (m < nbatch) {
      uint4* dst = myx + (m % kScore4Ring) * (kScore4Batch * kScore4Threads);
#pragma unroll
      for (int k = 0; k < kScore4Batch; ++k) {
        const int t = min(p.start + m * kScore4Batch + k, p.end - 1);
        cp_async_16(dst + k * kScore4Threads, xu + (int64_t)t * p.V);
      }
    }
    cp_async_commit();  // (an empty group keeps the count uniform)
  };
  for (int m = 0; m < kScore4Ring - 1; ++m) issue(m);
  int m = 0;
  static_assert(kScoreTB % kScore4Batch == 0, "batches do not straddle the blocks of E");
  for (int tb = p.start; tb < p.end; tb += kScoreTB) {
    const int nt = min(kScoreTB, p.end - tb);
    __syncthreads();
    for (int i = threadIdx.x; i < nt * NH; i += kScore4Threads) {
      const int f = i / NH, j = i - f * NH + p.hyp_off;
      sE[i] = j < NHP ? p.w.EA[((int64_t)u * p.T + (tb - 1 + f)) * NHP + j] : 0.f;
    }
    __syncthreads();
    for (int f0 = 0; f0 < nt; f0 += kScore4Batch, ++m) {
      issue(m + kScore4Ring - 1);  // into the slot of batch m - 1, which this thread has consumed
      cp_async_wait<kScore4Ring - 1>();  // batch m has landed
      const uint4* src = myx + (m % kScore4Ring) * (kScore4Batch * kScore4Threads);
#pragma unroll
      for (int k = 0; k < kScore4Batch; ++k) {
        const int f = f0 + k;
        if (f < nt) {
          const uint4 xv = src[k * kScore4Threads];
          float xh[4];
          xh[0] = ex2f(fmaf(__uint_as_float(xv.x), B200_LOG2E, -mxl[0]));
          xh[1] = ex2f(fmaf(__uint_as_float(xv.y), B200_LOG2E, -mxl[1]));
          xh[2] = ex2f(fmaf(__uint_as_float(xv.z), B200_LOG2E, -mxl[2]));
          xh[3] = ex2f(fmaf(__uint_as_float(xv.w), B200_LOG2E, -mxl[3]));
          const float4* e4 = reinterpret_cast<const float4*>(sE + f * NH);
#pragma unroll
          for (int j4 = 0; j4 < NH / 4; ++j4) {
            const float4 e = e4[j4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              acc[q][4 * j4 + 0] = fmaf(e.x, xh[q], acc[q][4 * j4 + 0]);
              acc[q][4 * j4 + 1] = fmaf(e.y, xh[q], acc[q][4 * j4 + 1]);
              acc[q][4 * j4 + 2] = fmaf(e.z, xh[q], acc[q][4 * j4 + 2]);
              acc[q][4 * j4 + 3] = fmaf(e.w, xh[q], acc[q][4 * j4 + 3]);
            }
          }
        }
      }
    }
  }
  cp_async_wait<0>();
  if (!live) return;
  float x0[4] = {B200_LZ, B200_LZ, B200_LZ, B200_LZ};  // r[0,n] = x[0,c]
  if (p.output_len == 0) {
    const float4 v = *reinterpret_cast<const float4*>(p.x + (int64_t)u * p.T * p.V + c0);
    x0[0] = v.x; x0[1] = v.y; x0[2] = v.z; x0[3] = v.w;
  }
#pragma unroll
  for (int j = 0; j < NH; ++j) {
    if (j + p.hyp_off < p.n_hyps) {
      const int h = u * p.n_hyps + j + p.hyp_off;
      const float sp = p.s_prev ? p.s_prev[h] : 0.f;
      const int64_t last = p.last_ids[h];
      float psi[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int c = c0 + q;
        if (c == p.blank) {
          psi[q] = B200_LZ;
        } else if (c == p.eos) {
          psi[q] = p.w.REND[h];
        } else {
          psi[q] = finish_psi(acc[q][j], p.w.MA[h] + mx[q], x0[q]);
          if ((int64_t)c != last) score_flag(p, acc[q][j], h, c, p.w.MA[h]);
        }
      }
      if (last >= c0 && last < c0 + 4 && last != p.blank && last != p.eos) {
        // the repeated-last-label entry belongs to the extra row of CTAs
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (c0 + q != (int)last) {
            p.log_psi[(int64_t)h * p.V + c0 + q] = psi[q];
            p.scores[(int64_t)h * p.V + c0 + q] = psi[q] - sp;
          }
      } else {
        *reinterpret_cast<float4*>(p.log_psi + (int64_t)h * p.V + c0) = make_float4(psi[0], psi[1], psi[2], psi[3]);
        *reinterpret_cast<float4*>(p.scores + (int64_t)h * p.V + c0) =
            make_float4(psi[0] - sp, psi[1] - sp, psi[2] - sp, psi[3] - sp);
      }
    }
  }
}

// The same computation with the log-probs arriving through a shared-memory RING of bulk (TMA) copies, as in the loss
// stream kernel: a CTA owns a tile of kRingTile consecutive tokens of one utterance (1 KB per frame); one producer
// warp issues one 1-D bulk copy per frame into the next free slot (kRingFrames frames per slot, completion counted
// on the slot's `full` mbarrier, slots handed back through `empty`), the consumer warps (thread = four tokens)
// take the slots in order.  Bytes in flight per SM are set by the ring, not by what a thread can keep in flight with
// cp.async (prefix_score_full4_kernel).  Measured at C3: 107 us against 114 us -- with 512-token tiles (330 CTAs, three
// per SM: two or three tiles per SM) it was SLOWER than cp.async (134 us); 256-token tiles give 660 CTAs, 4-5 per SM.
// The kernel is issue-bound (about 60 issue slots per thread and frame for 16 bytes of x), not HBM-bound.
constexpr int kRingTile = 256;     // tokens per CTA
constexpr int kRingFrames = 4;     // frames per slot
constexpr int kRingSlots = 6;
constexpr int kRingConsumers = 64;
constexpr int kRingRowBytes = kRingTile * 4;

template <int NH>
static size_t score_ring_smem() {
  return 256 + sizeof(float) * kScoreTB * NH + (size_t)kRingSlots * kRingFrames * kRingRowBytes;
}

// (Packed fp32x2 FMAs -- fma.rn.f32x2, SASS FFMA2 -- were tried for the 4 x NH accumulators: 232 vs 164 us per C3 step;
// 120 registers instead of 96 and the pack / unpack moves cost more than the halved FMA count saves.)
template <int NH>
__global__ void __launch_bounds__(kRingConsumers + 32) prefix_score_ring_kernel(ScoreParams p) {
  extern __shared__ __align__(128) unsigned char rg_smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(rg_smem);            // [kRingSlots]
  uint64_t* empty = full + kRingSlots;                               // [kRingSlots]
  float* sE = reinterpret_cast<float*>(rg_smem + 256);               // [kScoreTB][NH]
  unsigned char* ring = rg_smem + 256 + sizeof(float) * kScoreTB * NH;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (blockIdx.y == (unsigned)p.B) {
    // one more row of CTAs: the repeated-last-label entries, a warp per hypothesis (the token CTAs skip them)
    if (p.hyp_off != 0 || warp >= kRingConsumers / 32) return;
    const int nwp = gridDim.x * (kRingConsumers / 32);
    for (int h = blockIdx.x * (kRingConsumers / 32) + warp; h < p.n_bh; h += nwp) prefix_score_last_warp(p, h, lane);
    return;
  }
  const int u = blockIdx.y;
  const int tile0 = blockIdx.x * kRingTile;
  const int ntok = min(kRingTile, p.V - tile0);       // V % 4 == 0
  const uint32_t row_bytes = (uint32_t)ntok * 4u;
  const int nfr = max(p.end - p.start, 0);
  const int nstage = (nfr + kRingFrames - 1) / kRingFrames;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kRingSlots; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kRingConsumers / 32); }
    mbar_fence_init();
    fence_proxy_async();
  }
  __syncthreads();

  if (warp == kRingConsumers / 32) {
    // ------------------------------------------------------------------ producer
    if (lane == 0) {
      const float* xu = p.x + ((int64_t)u * p.T + p.start) * p.V + tile0;
      int slot = 0;
      uint32_t par = 0;  // parity of the `empty` completion that frees `slot` for its next use
      for (int s = 0; s < nstage; ++s) {
        if (s >= kRingSlots) mbar_wait(&empty[slot], par);
        const int n = min(kRingFrames, nfr - s * kRingFrames);
        mbar_arrive_expect_tx(&full[slot], (uint32_t)n * row_bytes);
        for (int k = 0; k < n; ++k)
          bulk_g2s(ring + ((size_t)slot * kRingFrames + k) * kRingRowBytes, xu + (int64_t)(s * kRingFrames + k) * p.V, row_bytes,
                   &full[slot]);
        if (++slot == kRingSlots) { slot = 0; if (s >= kRingSlots) par ^= 1u; }
      }
    }
    return;
  }

  // -------------------------------------------------------------------- consumers
  const int tid = threadIdx.x;  // 0 .. 127
  const int c0 = tile0 + 4 * tid;
  const bool live = 4 * tid < ntok;
  const int cc = live ? c0 : tile0;
  const int NHP = p.w.NHP;
  const float4 mx4 = *reinterpret_cast<const float4*>(p.xmax + (int64_t)u * p.V + cc);
  const float mx[4] = {mx4.x, mx4.y, mx4.z, mx4.w};
  float mxl[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) mxl[q] = mx[q] * B200_LOG2E;
  float acc[4][NH];
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int j = 0; j < NH; ++j) acc[q][j] = 0.f;
  const uint32_t my_off = live ? 16u * (uint32_t)tid : 0u;

  int slot = 0, st = 0;
  uint32_t par = 0;
  static_assert(kScoreTB % kRingFrames == 0, "slots do not straddle the blocks of E");
  for (int tb = p.start; tb < p.end; tb += kScoreTB) {
    const int nt = min(kScoreTB, p.end - tb);
    asm volatile("bar.sync 1, %0;" ::"n"(kRingConsumers) : "memory");
    // E rows for frames tb-1 .. tb+nt-2 (term t uses phi[t-1])
    for (int i = tid; i < nt * NH; i += kRingConsumers) {
      const int f = i / NH, j = i - f * NH + p.hyp_off;
      sE[i] = j < NHP ? p.w.EA[((int64_t)u * p.T + (tb - 1 + f)) * NHP + j] : 0.f;
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kRingConsumers) : "memory");
    for (int f0 = 0; f0 < nt; f0 += kRingFrames, ++st) {
      mbar_wait(&full[slot], par);
      const unsigned char* base = ring + (size_t)slot * kRingFrames * kRingRowBytes + my_off;
#pragma unroll
      for (int k = 0; k < kRingFrames; ++k) {
        const int f = f0 + k;
        if (f < nt) {
          const uint4 xv = *reinterpret_cast<const uint4*>(base + (size_t)k * kRingRowBytes);
          float xh[4];
          xh[0] = ex2f(fmaf(__uint_as_float(xv.x), B200_LOG2E, -mxl[0]));
          xh[1] = ex2f(fmaf(__uint_as_float(xv.y), B200_LOG2E, -mxl[1]));
          xh[2] = ex2f(fmaf(__uint_as_float(xv.z), B200_LOG2E, -mxl[2]));
          xh[3] = ex2f(fmaf(__uint_as_float(xv.w), B200_LOG2E, -mxl[3]));
          const float4* e4 = reinterpret_cast<const float4*>(sE + f * NH);
#pragma unroll
          for (int j4 = 0; j4 < NH / 4; ++j4) {
            const float4 e = e4[j4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              acc[q][4 * j4 + 0] = fmaf(e.x, xh[q], acc[q][4 * j4 + 0]);
              acc[q][4 * j4 + 1] = fmaf(e.y, xh[q], acc[q][4 * j4 + 1]);
              acc[q][4 * j4 + 2] = fmaf(e.z, xh[q], acc[q][4 * j4 + 2]);
              acc[q][4 * j4 + 3] = fmaf(e.w, xh[q], acc[q][4 * j4 + 3]);
            }
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
      if (++slot == kRingSlots) { slot = 0; par ^= 1u; }
    }
  }
  if (!live) return;
  float x0[4] = {B200_LZ, B200_LZ, B200_LZ, B200_LZ};  // r[0,n] = x[0,c]
  if (p.output_len == 0) {
    const float4 v = *reinterpret_cast<const float4*>(p.x + (int64_t)u * p.T * p.V + c0);
    x0[0] = v.x; x0[1] = v.y; x0[2] = v.z; x0[3] = v.w;
  }
#pragma unroll
  for (int j = 0; j < NH; ++j) {
    if (j + p.hyp_off < p.n_hyps) {
      const int h = u * p.n_hyps + j + p.hyp_off;
      const float sp = p.s_prev ? p.s_prev[h] : 0.f;
      const int64_t last = p.last_ids[h];
      float psi[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int c = c0 + q;
        if (c == p.blank) {
          psi[q] = B200_LZ;
        } else if (c == p.eos) {
          psi[q] = p.w.REND[h];
        } else {
          psi[q] = finish_psi(acc[q][j], p.w.MA[h] + mx[q], x0[q]);
          if ((int64_t)c != last) score_flag(p, acc[q][j], h, c, p.w.MA[h]);
        }
      }
      if (last >= c0 && last < c0 + 4 && last != p.blank && last != p.eos) {
        // the repeated-last-label entry belongs to the extra row of CTAs
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (c0 + q != (int)last) {
            p.log_psi[(int64_t)h * p.V + c0 + q] = psi[q];
            p.scores[(int64_t)h * p.V + c0 + q] = psi[q] - sp;
          }
      } else {
        *reinterpret_cast<float4*>(p.log_psi + (int64_t)h * p.V + c0) = make_float4(psi[0], psi[1], psi[2], psi[3]);
        *reinterpret_cast<float4*>(p.scores + (int64_t)h * p.V + c0) =
            make_float4(psi[0] - sp, psi[1] - sp, psi[2] - sp, psi[3] - sp);
      }
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

// One WARP per (hypothesis, candidate) pair: the lanes split the frames (each x load is its own 32-byte sector, V
// floats apart, so the loads of a warp are 32 independent sectors in flight) and the partial sums meet in one
// shuffle reduction.  (One THREAD per pair walking all frames -- 500 dependent strided loads, 8 in flight -- took
// 71 us at C3 with 15 candidates per hypothesis; this is the path BatchBeamSearch takes, batch_beam_search.py:228-238.)
__global__ void __launch_bounds__(256) prefix_score_partial_kernel(ScoreParams p) {
  const int lane = threadIdx.x & 31;
  const int idx = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (idx >= p.n_bh * p.snum) return;
  const int h = idx / p.snum;
  const int u = h / p.n_hyps, j = h - u * p.n_hyps;
  const int64_t c64 = p.scoring_ids[idx];
  if (c64 < 0 || c64 >= p.V || c64 == p.blank || c64 == p.eos) return;  // filled already
  const int c = (int)c64;
  const int NHP = p.w.NHP;
  const bool is_last = (c64 == p.last_ids[h]);
  const float* E = (is_last ? p.w.EB : p.w.EA) + (int64_t)u * p.T * NHP + j;
  const float* xu = p.x + (int64_t)u * p.T * p.V + c;
  const float mx = p.xmax[(int64_t)u * p.V + c];
  const float mxl = mx * B200_LOG2E;
  float a0 = 0.f, a1 = 0.f;
  int t = p.start + lane;
  for (; t + 32 < p.end; t += 64) {  // two independent loads per lane and trip
    const float x0 = ld_nc_f32(xu + (int64_t)t * p.V), x1 = ld_nc_f32(xu + (int64_t)(t + 32) * p.V);
    const float e0 = E[(int64_t)(t - 1) * NHP], e1 = E[(int64_t)(t + 31) * NHP];
    a0 = fmaf(e0, ex2f(fmaf(x0, B200_LOG2E, -mxl)), a0);
    a1 = fmaf(e1, ex2f(fmaf(x1, B200_LOG2E, -mxl)), a1);
  }
  if (t < p.end) a0 = fmaf(E[(int64_t)(t - 1) * NHP], ex2f(fmaf(ld_nc_f32(xu + (int64_t)t * p.V), B200_LOG2E, -mxl)), a0);
  const float a = warp_sum(a0 + a1);
  if (lane == 0) {
    const float x0 = (p.output_len == 0) ? xu[0] : B200_LZ;
    const float psi = finish_psi(a, (is_last ? p.w.MB[h] : p.w.MA[h]) + mx, x0);
    score_flag(p, a, h, c, is_last ? p.w.MB[h] : p.w.MA[h]);
    const float sp = p.s_prev ? p.s_prev[h] : 0.f;
    p.log_psi[(int64_t)h * p.V + c] = psi;
    p.scores[(int64_t)h * p.V + c] = psi - sp;
  }
}

// The candidate (pre-beam) step from the column copy of x (b200_ctc_prefix_columns: (B,V,Tc), a token's frames
// contiguous), ONE kernel.  The kernels above gather one 4-byte word per frame from rows V floats apart: every word
// costs a DRAM sector of its own (x is larger than L2 at C3: 34 us for 4800 pairs), and the step is four launches (phi,
// fill, partial, fix) that the host cannot issue as fast as the GPU runs them.  Here a CTA owns a hypothesis: it computes
// the hypothesis' phi and E columns into shared memory (what prefix_phi_kernel writes to the workspace), fills its score
// rows (prefix_fill_kernel), and each warp takes candidates -- 128-byte coalesced reads of the token's column, the same
// accumulation order as prefix_score_partial_kernel, and a sum that underflowed is redone in log space by the same
// warp on the spot (prefix_score_fix_kernel's arithmetic), so the results are the four-kernel path's bit for bit.
constexpr int kColsThreads = 512;
__global__ void __launch_bounds__(kColsThreads) prefix_score_cols_kernel(ScoreParams p, const int64_t* __restrict__ xlens) {
  extern __shared__ float s_e[];  // [4][n]: phi_a, phi_b (log), EA, EB of this hypothesis for frames start-1 .. end-2
  __shared__ float s_red[2][kColsThreads / 32];
  __shared__ float s_rend;
  const int h = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int u = h / p.n_hyps;
  const int n = p.end > p.start ? p.end - p.start : 0;
  float* s_pa = s_e;
  float* s_pb = s_e + n;
  float* s_ea = s_e + 2 * n;
  float* s_eb = s_e + 3 * n;
  float ma = -3.0e38f, mb = -3.0e38f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int t = p.start - 1 + i;
    const float rn = p.r_prev[((int64_t)t * 2 + 0) * p.n_bh + h];
    const float rb = p.r_prev[((int64_t)t * 2 + 1) * p.n_bh + h];
    const float pa = lse2_nat(rn, rb);
    s_pa[i] = pa;
    s_pb[i] = rb;
    ma = fmaxf(ma, pa);
    mb = fmaxf(mb, rb);
  }
  ma = warp_max(ma);
  mb = warp_max(mb);
  if (lane == 0) { s_red[0][warp] = ma; s_red[1][warp] = mb; }
  if (threadIdx.x == 0) {
    const int te = (int)xlens[u] - 1;
    s_rend = lse2_nat(p.r_prev[((int64_t)te * 2 + 0) * p.n_bh + h], p.r_prev[((int64_t)te * 2 + 1) * p.n_bh + h]);
  }
  __syncthreads();
  ma = warp_max(lane < nw ? s_red[0][lane] : -3.0e38f);
  mb = warp_max(lane < nw ? s_red[1][lane] : -3.0e38f);
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    s_ea[i] = ex2f((s_pa[i] - ma) * B200_LOG2E);
    s_eb[i] = ex2f((s_pb[i] - mb) * B200_LOG2E);
  }
  const float sp = p.s_prev ? p.s_prev[h] : 0.f;
  {
    // rows of the hypothesis: logzero everywhere, r_sum[xlen-1] at eos (ctc_prefix_score.py:182-186)
    float* lrow = p.log_psi + (int64_t)h * p.V;
    float* srow = p.scores + (int64_t)h * p.V;
    const float rend = s_rend;
    for (int c = threadIdx.x; c < p.V; c += blockDim.x) {
      const float psi = (c == p.eos && c != p.blank) ? rend : B200_LZ;
      lrow[c] = psi;
      srow[c] = psi - sp;
    }
  }
  __syncthreads();  // E complete; the fill precedes the candidates' stores
  const int64_t last = p.last_ids[h];
  for (int q = warp; q < p.snum; q += nw) {
    const int64_t c64 = p.scoring_ids[(int64_t)h * p.snum + q];
    if (c64 < 0 || c64 >= p.V || c64 == p.blank || c64 == p.eos) continue;  // filled already
    const int c = (int)c64;
    const bool is_last = (c64 == last);
    const float* col = p.xc + ((int64_t)u * p.V + c) * p.Tc;
    const float* e = (is_last ? s_eb : s_ea) - p.start;  // e[t] = E[t-1]
    const float mx = p.xmax[(int64_t)u * p.V + c];
    const float mxl = mx * B200_LOG2E;
    float a0 = 0.f, a1 = 0.f;
    int t = p.start + lane;
#pragma unroll 4
    for (; t + 32 < p.end; t += 64) {
      const float x0 = ld_nc_f32(col + t), x1 = ld_nc_f32(col + t + 32);
      a0 = fmaf(e[t], ex2f(fmaf(x0, B200_LOG2E, -mxl)), a0);
      a1 = fmaf(e[t + 32], ex2f(fmaf(x1, B200_LOG2E, -mxl)), a1);
    }
    if (t < p.end) a0 = fmaf(e[t], ex2f(fmaf(ld_nc_f32(col + t), B200_LOG2E, -mxl)), a0);
    const float a = __shfl_sync(0xffffffffu, warp_sum(a0 + a1), 0);  // (one value for the whole warp: the branch below is collective)
    const float x0 = (p.output_len == 0) ? col[0] : B200_LZ;
    float psi;
    if (a < kScoreAccFloor && n > 0 && (is_last ? mb : ma) > 0.5f * B200_LZ && p.start < (int)xlens[u]) {
      // the scaled sum underflowed: logsumexp over { phi[t-1] + x[t,c] } as written (ctc_prefix_score.py:165-180)
      const float* ph = (is_last ? s_pb : s_pa) - p.start;
      float m = -3.0e38f, sacc = 0.f;
      for (int tt = p.start + lane; tt < p.end; tt += 32) {
        const float v = ph[tt] + ld_nc_f32(col + tt);
        const float mn = fmaxf(m, v);
        sacc = sacc * ex2f((m - mn) * B200_LOG2E) + ex2f((v - mn) * B200_LOG2E);
        m = mn;
      }
      const float mw = warp_max(m);
      sacc = warp_sum(sacc * ex2f((m - mw) * B200_LOG2E));
      psi = sacc > 0.f ? mw + B200_LN2 * lg2f(sacc) : B200_LZ;
      psi = fmaxf(psi, B200_LZ);
      if (x0 > B200_LZ) psi = lse2_nat(psi, x0);
    } else {
      psi = finish_psi(a, (is_last ? mb : ma) + mx, x0);
    }
    if (lane == 0) {
      p.log_psi[(int64_t)h * p.V + c] = psi;
      p.scores[(int64_t)h * p.V + c] = psi - sp;
    }
  }
}

// x (B,T,V) -> xc (B,V,Tc): 32 x 32 tiles through shared memory, reads coalesced along V, writes along T.
__global__ void __launch_bounds__(256) prefix_columns_kernel(const float* __restrict__ x, int T, int V, int Tc,
                                                             float* __restrict__ xc) {
  __shared__ float tile[32][33];
  const int u = blockIdx.z, v0 = blockIdx.x * 32, t0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const float* xu = x + (int64_t)u * T * V;
  float* cu = xc + (int64_t)u * V * Tc;
  for (int i = ty; i < 32; i += 8)
    if (t0 + i < T && v0 + tx < V) tile[i][tx] = xu[(int64_t)(t0 + i) * V + v0 + tx];
  __syncthreads();
  for (int i = ty; i < 32; i += 8)
    if (v0 + i < V && t0 + tx < T) cu[(int64_t)(v0 + i) * Tc + t0 + tx] = tile[tx][i];
}

// Exact log-space redo of the queued (hypothesis, token) pairs (ctc_prefix_score.py:165-180 as written:
// logsumexp over { r[start-1,n], phi[t-1] + x[t,c] }), one warp per pair.  The queue is empty for ordinary
// posteriors; it fills when the joint spread of phi and x over the window exceeds what the scaled linear sum
// holds (log-probs tens of nats apart between frames: saturated posteriors, windows that exclude the frames
// that carry the mass).
__global__ void __launch_bounds__(256) prefix_score_fix_kernel(ScoreParams p) {
  const int lane = threadIdx.x & 31;
  const int nw = gridDim.x * (blockDim.x >> 5);
  const int n = min(*p.w.fix_count, p.n_bh * p.V);
  for (int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); q < n; q += nw) {
    const int code = p.w.fix_list[q];
    const int h = code / p.V, c = code - h * p.V;
    const int u = h / p.n_hyps;
    const bool is_last = ((int64_t)c == p.last_ids[h]);
    const float* xu = p.x + (int64_t)u * p.T * p.V + c;
    float m = -3.0e38f, s = 0.f;  // running maximum and sum of exp(v - m), per lane
    for (int t = p.start + lane; t < p.end; t += 32) {
      const float rn = p.r_prev[((int64_t)(t - 1) * 2 + 0) * p.n_bh + h];
      const float rb = p.r_prev[((int64_t)(t - 1) * 2 + 1) * p.n_bh + h];
      const float v = (is_last ? rb : lse2_nat(rn, rb)) + ld_nc_f32(xu + (int64_t)t * p.V);
      const float mn = fmaxf(m, v);
      s = s * ex2f((m - mn) * B200_LOG2E) + ex2f((v - mn) * B200_LOG2E);
      m = mn;
    }
    const float mw = warp_max(m);
    s = warp_sum(s * ex2f((m - mw) * B200_LOG2E));
    if (lane == 0) {
      float psi = s > 0.f ? mw + B200_LN2 * lg2f(s) : B200_LZ;
      psi = fmaxf(psi, B200_LZ);
      const float x0 = (p.output_len == 0) ? xu[0] : B200_LZ;
      if (x0 > B200_LZ) psi = lse2_nat(psi, x0);
      const float sp = p.s_prev ? p.s_prev[h] : 0.f;
      p.log_psi[(int64_t)h * p.V + c] = psi;
      p.scores[(int64_t)h * p.V + c] = psi - sp;
    }
  }
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
  const int64_t* token;  // null: parent[] holds beam-pruning ids in (hypothesis * V + token) space, `beam` per utterance
  int beam;
  float* r_new;  // (T,2,n_new)
  float* s_new;  // (n_new)
};

// parent hypothesis and appended token of survivor k (ctc_prefix_score.py:203-207 when only best_ids is given)
__device__ __forceinline__ void select_pair(const SelectParams& p, int k, int& h, int& c) {
  if (p.token) {
    h = (int)p.parent[k];
    c = (int)p.token[k];
  } else {
    const int64_t best = p.parent[k];
    h = (int)(best / p.V) + (k / p.beam) * p.n_hyps;
    c = (int)(best % p.V);
  }
}

// ONE WARP per survivor: lane i loads the inputs of frame t0 + i of a batch of 32 frames (and computes phi for it),
// the next batch is loaded before the current one is consumed, and the recursion itself runs in every lane; lane f
// keeps the values of frame f and stores them after the batch.
// The recursion is one latency chain per survivor, so its length is what the kernel costs (measured at C3: 157
// cycles per frame with the inputs fetched by shuffles right in front of their use and the two states computed one
// after the other).  Here the batch's inputs go through shared memory (one broadcast 16-byte load per frame, issued a
// frame ahead), each state's chain is  sub -> mul -> ex2 -> add -> lg2 -> fma  (the maximum and the frame's log-prob
// are folded into the fma's addend, off the chain), and the two states' ex2 / lg2 are issued alternately.
__device__ __forceinline__ float sel_ex2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sel_lg2(float x) {
  float y;
  asm volatile("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// one frame: (pn, pb) <- (lse(pn, phi) + xc, lse(pn, pb) + xb), natural-log units
__device__ __forceinline__ void select_frame(float& pn, float& pb, const float4 in) {
  const float dn = pn - in.x, db = pn - pb;
  const float basen = fmaxf(pn, in.x) + in.y, baseb = fmaxf(pn, pb) + in.z;
  const float en = sel_ex2(-fabsf(dn) * B200_LOG2E);
  const float eb = sel_ex2(-fabsf(db) * B200_LOG2E);
  const float ln = sel_lg2(1.0f + en);
  const float lb = sel_lg2(1.0f + eb);
  pn = fmaf(ln, B200_LN2, basen);
  pb = fmaf(lb, B200_LN2, baseb);
}

__global__ void __launch_bounds__(128) prefix_select_warp_kernel(SelectParams p) {
  __shared__ __align__(16) float4 s_in[4][32];  // per warp: {phi, x[c], x[blank], -} of the 32 frames of a batch
  const int lane = threadIdx.x & 31;
  const int k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (k >= p.n_new) return;
  float4* sw = s_in[threadIdx.x >> 5];
  int h, c;
  select_pair(p, k, h, c);
  const int u = h / p.n_hyps;
  const bool is_last = ((int64_t)c == p.last_ids[h]);
  const float* xu = p.x + (int64_t)u * p.T * p.V;
  const int pre = min(p.start, p.T);
  // frames before the window keep log(0), except r[0,n] = x[0,c] for the first label (:130-131)
  for (int t = lane; t < pre; t += 32) {
    p.r_new[((int64_t)t * 2 + 0) * p.n_new + k] = (t == 0 && p.output_len == 0) ? xu[c] : B200_LZ;
    p.r_new[((int64_t)t * 2 + 1) * p.n_new + k] = B200_LZ;
  }
  float pn = (pre == 1 && p.output_len == 0) ? xu[c] : B200_LZ, pb = B200_LZ;
  struct In { float xc, xb, rn, rb; };
  auto load = [&](int t0) {
    In q;
    const int t = max(min(t0 + lane, p.end - 1), p.start);
    q.xc = ld_nc_f32(xu + (int64_t)t * p.V + c);
    q.xb = ld_nc_f32(xu + (int64_t)t * p.V + p.blank);
    q.rn = p.r_prev[((int64_t)(t - 1) * 2 + 0) * p.n_bh + h];
    q.rb = p.r_prev[((int64_t)(t - 1) * 2 + 1) * p.n_bh + h];
    return q;
  };
  if (p.start < p.end) {
    In cur = load(p.start);
    for (int t0 = p.start; t0 < p.end; t0 += 32) {
      const In nxt = load(t0 + 32);
      const float phi_l = is_last ? cur.rb : lse2_nat(cur.rn, cur.rb);
      __syncwarp();  // every lane is done with the previous batch
      sw[lane] = make_float4(phi_l, cur.xc, cur.xb, 0.f);
      __syncwarp();
      float my_n = B200_LZ, my_b = B200_LZ;
      const int nf = min(32, p.end - t0);
      float4 in = sw[0];
      if (nf == 32) {
#pragma unroll
        for (int f = 0; f < 32; ++f) {
          const float4 nin = sw[(f + 1) & 31];  // (next frame's inputs: off the chain)
          select_frame(pn, pb, in);
          if (lane == f) { my_n = pn; my_b = pb; }
          in = nin;
        }
      } else {
        for (int f = 0; f < nf; ++f) {
          const float4 nin = sw[(f + 1) & 31];
          select_frame(pn, pb, in);
          if (lane == f) { my_n = pn; my_b = pb; }
          in = nin;
        }
      }
      if (lane < nf) {
        p.r_new[((int64_t)(t0 + lane) * 2 + 0) * p.n_new + k] = my_n;
        p.r_new[((int64_t)(t0 + lane) * 2 + 1) * p.n_new + k] = my_b;
      }
      cur = nxt;
    }
  }
  for (int t = max(p.end, p.start) + lane; t < p.T; t += 32) {
    p.r_new[((int64_t)t * 2 + 0) * p.n_new + k] = B200_LZ;
    p.r_new[((int64_t)t * 2 + 1) * p.n_new + k] = B200_LZ;
  }
  if (lane == 0) p.s_new[k] = p.log_psi[(int64_t)h * p.V + c];
}


}  // namespace b200
