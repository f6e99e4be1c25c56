// ctc_loss_sweep_cluster.inl -- the lattice sweeps for long label sequences (256 <= L + 1 <= 1024).
//
// ctc_sweep_kernel<4,1024> runs all 2L+1 states of an utterance in ONE CTA: 32 skewed warps on one SM, eight per
// scheduler, ~45 issue slots per warp and frame -- the recursion is issue-bound there (396 ns per frame at C5:
// T=4000, L<=1000, and B=16 utterances x 2 directions occupy 32 of the 148 SMs for 1.6 of the 2.5 ms of a step).
// Here the label dimension of one (utterance, direction) is split over a thread-block CLUSTER: kClusterCtas CTAs of
// 8 warps (256 label positions each) on as many SMs, two warps per scheduler, so every warp is again one latency
// chain.  The recursion is the one of ctc_sweep_kernel (same frame(), same integer re-centring); what changes is
// where the hand-off records live:
//   * the boundary value that the LAST warp of CTA r publishes per frame goes straight into the shared memory of
//     CTA r+1 (mapa + st.shared::cluster: distributed shared memory), where warp 0 polls it like a local record;
//   * the level record every warp publishes every second chunk is written to ALL active CTAs of the cluster, so
//     every warp still derives the same integer corrections from the same 32 records;
//   * the last blank / label state are written to CTA 0, which finishes log p(y|x) after a cluster barrier.
// Records stay self-validating {value, tag} 8-byte stores: no fence on the hand-off path.
namespace b200 {

constexpr int kClusterCtas = 4;
constexpr int kClusterWarps = 8;                              // state warps per CTA
constexpr int kClusterThreads = kClusterWarps * 32;
constexpr int kClusterMaxWarps = kClusterCtas * kClusterWarps;  // 32: L + 1 <= 1024
constexpr int kClusterRing = 512;                               // boundary records per warp (frames)
// Chunks of slack in the level feedback (see `delay`) and level records every kP-th chunk: 8 and 4 were 2 % faster at
// C5 (1.05 vs 1.07 ms) but lengthen the horizon over which the level is PREDICTED ((delay + kP) x 8 frames), and on
// tight lattices (T barely above what the labels need) the states then sit further from zero: gradient error 4.5e-5
// against 1.1e-5 relative (the one-CTA kernel, 4-frame chunks: 0.7e-5; torch's float32 CTC on that input: 1e-3).
constexpr int kClusterSlack = 0;

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// shared::cluster address of `local` (an address in this CTA's shared window) inside CTA `rank` of the cluster
__device__ __forceinline__ uint32_t cluster_map(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ void stc_tagged(uint32_t addr, float v, int tag) {
  // (relaxed, cluster scope: one 8-byte store that is not ordered against the stores before it -- a volatile store
  //  would make every hand-off wait for the previous one's round trip to the other SM)
  asm volatile("{\n.reg .b64 q;\nmov.b64 q, {%1,%2};\nst.relaxed.cluster.shared::cluster.b64 [%0], q;\n}" ::"r"(addr), "r"(__float_as_int(v)), "r"(tag) : "memory");
}
// generic address of a shared::cluster address (so that the stores below need no per-store window arithmetic)
__device__ __forceinline__ uint64_t cluster_generic(uint32_t cluster_addr) {
  uint64_t g;
  asm volatile("cvta.shared::cluster.u64 %0, %1;" : "=l"(g) : "l"((uint64_t)cluster_addr));
  return g;
}
// {v, tag} as ONE predicated 8-byte store to a generic address (own or a neighbour's shared memory): no branch, so
// the lane that publishes does not put a divergent region on the warp's chain (measured: 215-240 cycles per frame
// with `if (lane31) { remote ? ... : ... }`, against ~115 for the recursion itself)
__device__ __forceinline__ void stg_tagged_if(bool on, uint64_t gaddr, float v, int tag) {
  asm volatile("{\n.reg .pred p;\n.reg .b64 q;\nsetp.ne.s32 p, %3, 0;\nmov.b64 q, {%1,%2};\n@p st.relaxed.cluster.b64 [%0], q;\n}"
               ::"l"(gaddr), "r"(__float_as_int(v)), "r"(tag), "r"((int)on) : "memory");
}
__device__ __forceinline__ void stc_f32(uint32_t addr, float v) {
  asm volatile("st.volatile.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(kClusterThreads) ctc_sweep_cluster_kernel(LossParams p) {
  // chunks of 8 frames (the one-CTA kernel: 4): the per-chunk work -- boundary poll, level record, control arithmetic,
  // the loads of chunk c + 2 -- is paid half as often; the hand-off ring holds (32 + 3) chunks of skew
  constexpr int CHF = 8, kR = kClusterRing, kWR = 128, NWT = kClusterMaxWarps;
  static_assert(kR >= CHF * (kClusterMaxWarps + kClusterSlack + 3) && kWR >= 2 * (kClusterMaxWarps + kClusterSlack) + 4, "rings");
  extern __shared__ __align__(16) unsigned char sw_smem[];
  Tagged* s_bnd = reinterpret_cast<Tagged*>(sw_smem);          // [1 + kClusterWarps][kR]: row 0 arrives from CTA rank-1
  Tagged* s_wmx = s_bnd + (size_t)(1 + kClusterWarps) * kR;    // [kWR][NWT] level of global warp g at the start of chunk c
  int* s_hist = reinterpret_cast<int*>(s_wmx + (size_t)kWR * NWT);  // [kWR] total removed before chunk c
  __shared__ float s_fin[2];                                   // (CTA 0's copy is the one that is read)
  __shared__ int s_csum;

  pdl_wait();
  pdl_launch_dependents();
  const uint32_t rank = cluster_rank();
  const int b = p.ws.order[p.b0 + blockIdx.x];
  const int d = blockIdx.y;
  const int Tb = min((int)p.ilen[b], (int)p.T);
  const int L = utt_labels(p, b);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = (int)rank * kClusterThreads + threadIdx.x;   // logical label position / state pair of this thread
  const int gw = (int)rank * kClusterWarps + warp;           // global warp index
  const int Kp = (int)p.ws.Kp;
  const int64_t Sp = p.ws.Sp;
  const int nwa = L / 32 + 1;                                // warps that own at least one state
  const int nca = (nwa + kClusterWarps - 1) / kClusterWarps;  // CTAs of the cluster that do

  for (int k = threadIdx.x; k < (1 + kClusterWarps) * kR + kWR * NWT; k += blockDim.x) s_bnd[k].tag = -1;
  // nobody writes into a neighbour before every CTA of the cluster has cleared its records
  cluster_sync_all();

  if (Tb <= 0) {
    if (i == 0) {
      if (d == 0) sweep_finish_nll(p, b, (L == 0) ? 0.0f : __int_as_float(0x7f800000));
      else p.ws.logz_beta[b] = (L == 0) ? 0.0f : B200_NEG;
    }
    pdl_wait();
    return;  // (uniform over the cluster: nothing remote is touched after the barrier above)
  }

  float ab = B200_NEG, al = B200_NEG;
  int csum = 0;  // exact sum of everything removed so far
  if (gw < nwa) {
    const int* lab = p.ws.labels + (int64_t)b * Kp;
    const bool has_label = i < L;
    const bool active = i <= L;
    const int pi = d ? (L - 1 - i) : i;  // physical label position of logical position i
    bool skip_ok = false;
    if (has_label && i >= 1) skip_ok = lab[pi] != lab[d ? pi + 1 : pi - 1];
    const float skipmask = skip_ok ? 0.0f : B200_NEG;
    const int kcol = has_label ? 1 + pi : 0;
    const float* lp0 = p.ws.lpg + ((int64_t)b * p.T + (d ? (Tb - 1) : 0)) * Kp;
    const int lstep = d ? -Kp : Kp;
    float* outp = p.ws.lat + (((int64_t)d * p.B + b) * p.T + (d ? (Tb - 1) : 0)) * Sp + 2 * i;
    const int64_t ostep = d ? -Sp : Sp;
    const bool lane0 = lane == 0, lane31 = lane == 31;
    const bool poll = gw > 0;
    const Tagged* bsrc = s_bnd + (size_t)warp * kR;  // row w: what the warp before this one published
    // where this warp publishes: the next local row, or row 0 of the next CTA
    const bool remote = warp == kClusterWarps - 1 && gw + 1 < nwa;
    const uint64_t bdst = remote ? cluster_generic(cluster_map(smem_u32(s_bnd), rank + 1))
                                 : reinterpret_cast<uint64_t>(s_bnd + (size_t)(warp + 1) * kR);
    // lane r < nca publishes this warp's level records into CTA r of the cluster
    const uint64_t wmx_dst = cluster_generic(cluster_map(smem_u32(s_wmx), (uint32_t)(lane < nca ? lane : 0))) + (uint64_t)gw * 8u;
    // The level record of chunk c - delay must be there when chunk c starts.  Successive warps run one chunk apart, so
    // the last warp is nwa - 1 chunks (plus three hops between SMs) behind the first.
    const int delay = nwa + kClusterSlack;
#ifdef B200_SWEEP_CLOCKS
    // developer instrumentation: SM clock at four points of chunks 64..127 for eight warps of the longest utterance
    const int clk_slot = gw == 0 ? 0 : gw == 1 ? 1 : gw == 7 ? 2 : gw == 8 ? 3 : gw == 9 ? 4 : gw == 16 ? 5 : gw == 24 ? 6 : gw == 31 ? 7 : -1;
#define CL_STAMP(k) do { if (blockIdx.x == 0 && lane == 0 && clk_slot >= 0 && c >= 64 && c < 128) g_sweep_clk[((d * 8 + clk_slot) * 4 + (k)) * 64 + (c - 64)] = clock64(); } while (0)
#else
#define CL_STAMP(k) do { } while (0)
#endif
    float dstep = 0.f;
    float xprev = 0.f; int sprev = 0;
    bool have_prev = false;

    auto frame = [&](float lb_t, float ll_t, float bn, float o, int fr) {
      // (the frame of ctc_sweep_staged_kernel: the skip mask folded into the own-state side, two predicated
      //  subtractions instead of a select in front of one, the label state's ex2 / lg2 issued first)
      const float sh = __shfl_up_sync(0xffffffffu, al, 1);
      const float hi = fmaxf(al, ab);
      const float s1 = 1.0f + ex2f(-fabsf(al - ab));
      const float hi2 = hi - skipmask;
      const float lbo = lb_t - o, llo = ll_t - o;
      const float dl = sub_pick(lane0, bn, sh, hi2);
      const float db = sub_pick(lane0, bn, sh, ab);
      const float t = ex2f(-fabsf(dl));
      const float u = ex2f(-fabsf(db));
      const float nbr = lane0 ? bn : sh;
      const float sum = (dl <= 0.f) ? (s1 + t) : fmaf(s1, t, 1.0f);
      const float mxl = llo + (skip_ok ? fmaxf(hi, nbr) : hi);
      const float mxb = lbo + fmaxf(ab, nbr);
      al = mxl + lg2f(sum);
      ab = mxb + lg2f(1.0f + u);
      stg_tagged_if(lane31, bdst + (uint64_t)((fr & (kR - 1)) * 8), al, fr);
      if (active) *reinterpret_cast<float2*>(outp) = make_float2(ab, al);
      outp += ostep;
    };
    auto chunk_begin = [&](int c, int n) -> float {
      constexpr int kP = 2;
      if (c % kP) return 0.f;
      float corr = 0.f;
      const float wm = warp_max_redux(fmaxf(ab, al));
      // the level of this warp goes to every active CTA of the cluster (lanes 0 .. nca-1, one CTA each)
      stg_tagged_if(lane < nca, wmx_dst + (uint64_t)((c & (kWR - 1)) * NWT) * 8u, wm, c);
      if (lane31) s_hist[c & (kWR - 1)] = csum;
      __syncwarp();
      const int delay_p = (delay + kP - 1) / kP * kP;
      const int cm = c - delay_p;
      if (cm >= kP) {
        Tagged e;
        e.v = B200_NEG; e.tag = cm;
        if (lane < nwa) {
          const Tagged* src = &s_wmx[(cm & (kWR - 1)) * NWT + lane];
          do { e = lds_tagged(src); } while (e.tag != cm);
        }
        __syncwarp();
        const float xm = warp_max_redux(e.v);
        const int sm = s_hist[cm & (kWR - 1)];
        float delta = dstep;
        if (have_prev) delta = (xm - xprev + (float)(sm - sprev)) * (1.0f / (kP * CHF));
        xprev = xm; sprev = sm; have_prev = true;
        const float xhat = xm + (float)(sm - csum) + delta * (float)(CHF * delay_p);
        const float horizon = (float)(n == CHF ? kP * CHF : n);
        dstep = clamp_int_f(delta);
        corr = clamp_int_f(clamp_int_f(xhat + delta * horizon) - dstep * horizon);
      }
      return corr;
    };
    auto slow_chunk = [&](int c, int n) {
      const int f0 = c * CHF;
      float corr = 0.f;
      if (c == 0) {
        if (Tb > 1) dstep = clamp_int_f(floorf(ld_relaxed_f32(lp0 + lstep)));
        if (lane31) s_hist[0] = 0;
      } else {
        corr = chunk_begin(c, n);
      }
      float sb[CHF], sl[CHF];
#pragma unroll
      for (int f = 0; f < CHF; ++f) {
        const float* rowp = lp0 + (int64_t)(f0 + min(f, n - 1)) * lstep;
        sb[f] = ld_relaxed_f32(rowp);
        sl[f] = has_label ? ld_relaxed_f32(rowp + kcol) : B200_NEG;
      }
#pragma unroll
      for (int f = 0; f < CHF; ++f) {
        if (f < n) {
          const int fr = f0 + f;
          const float lb_t = sb[f], ll_t = sl[f];
          if (fr == 0) {
            if (i == 0) { ab = lb_t; al = ll_t; }
            stg_tagged_if(lane31, bdst, al, 0);
            if (active) *reinterpret_cast<float2*>(outp) = make_float2(ab, al);
            outp += ostep;
          } else {
            Tagged e;
            e.v = B200_NEG; e.tag = fr - 1;
            if (poll) {
              do { e = lds_tagged(&bsrc[(fr - 1) & (kR - 1)]); } while (e.tag != fr - 1);
            }
            frame(lb_t, ll_t, e.v, dstep + (f == 0 ? corr : 0.f), fr);
          }
        }
      }
      csum += (n - (c == 0 ? 1 : 0)) * (int)dstep + (int)corr;
    };
    const int nfull = Tb / CHF;
    auto load_chunk = [&](int c, float(&xb)[CHF], float(&xl)[CHF]) {
      const float* rowp = lp0 + (int64_t)(c * CHF) * lstep;
#pragma unroll
      for (int f = 0; f < CHF; ++f) {
        xb[f] = ld_relaxed_f32(rowp);
        xl[f] = has_label ? ld_relaxed_f32(rowp + kcol) : B200_NEG;
        rowp += lstep;
      }
    };
    auto fast_chunk = [&](int c, float(&xb)[CHF], float(&xl)[CHF], float(&nb)[CHF], float(&nl)[CHF]) {
      const int f0 = c * CHF;
      CL_STAMP(0);
      float first = xb[0] + xl[0];
#pragma unroll
      for (int f = 1; f < CHF; ++f) first += xb[f] + xl[f];
      asm volatile("" ::"f"(first));
      if (c + 2 < nfull) load_chunk(c + 2, nb, nl);
      const float corr = chunk_begin(c, CHF);
      CL_STAMP(1);
      float bn[CHF];
#pragma unroll
      for (int f = 0; f < CHF; ++f) bn[f] = B200_NEG;
      if (poll) {
        bool ok;
        do {
          ok = true;
#pragma unroll
          for (int f = 0; f < CHF; ++f) {
            const Tagged e = lds_tagged(&bsrc[(f0 + f - 1) & (kR - 1)]);
            bn[f] = e.v;
            ok = ok && (e.tag == f0 + f - 1);
          }
        } while (!ok);
      }
      CL_STAMP(2);
#pragma unroll
      for (int f = 0; f < CHF; ++f) frame(xb[f], xl[f], bn[f], dstep + (f == 0 ? corr : 0.f), f0 + f);
      CL_STAMP(3);
      csum += CHF * (int)dstep + (int)corr;
    };

    float rb0[CHF], rl0[CHF], rb1[CHF], rl1[CHF], rb2[CHF], rl2[CHF];
    if (1 < nfull) load_chunk(1, rb0, rl0);
    if (2 < nfull) load_chunk(2, rb1, rl1);
    slow_chunk(0, min(CHF, Tb));
    int c = 1;
    while (c < nfull) {
      fast_chunk(c, rb0, rl0, rb2, rl2);
      if (++c >= nfull) break;
      fast_chunk(c, rb1, rl1, rb0, rl0);
      if (++c >= nfull) break;
      fast_chunk(c, rb2, rl2, rb1, rl1);
      ++c;
    }
    if (nfull >= 1 && Tb > nfull * CHF) slow_chunk(nfull, Tb - nfull * CHF);

    // the last blank and the last label state go to CTA 0
    const uint32_t fin0 = cluster_map(smem_u32(s_fin), 0);
    if (i == L) stc_f32(fin0, ab);
    if (L == 0 && i == 0) stc_f32(fin0 + 4, B200_NEG);
    if (L > 0 && i == L - 1) stc_f32(fin0 + 4, al);
    if (gw == 0 && lane == 0) s_csum = csum;  // (CTA 0, read below; every warp holds the same sum)
  }
  // every state warp of every CTA has stored its last frame (and the values above) before anybody passes this
  __threadfence();
  cluster_sync_all();
  if (i == 0) {
    const float z = lse2(*reinterpret_cast<volatile float*>(&s_fin[0]), *reinterpret_cast<volatile float*>(&s_fin[1]));
    const bool feasible = z > -1.0e29f;
    const float logz2 = z + (float)(*reinterpret_cast<volatile int*>(&s_csum));
    if (d == 0) sweep_finish_nll(p, b, feasible ? -B200_LN2 * logz2 : __int_as_float(0x7f800000));
    else p.ws.logz_beta[b] = feasible ? logz2 : B200_NEG;
    __threadfence();
    atomicAdd(p.ws.done + b, 1);
  }
}

static size_t sweep_cluster_smem() {
  return sizeof(Tagged) * (size_t)(1 + kClusterWarps) * kClusterRing + sizeof(Tagged) * 128 * kClusterMaxWarps + sizeof(int) * 128;
}

static int launch_sweep_cluster(const LossParams& p, cudaStream_t st, bool pdl) {
  size_t smem = sweep_cluster_smem();
  const int nb = p.b1 - p.b0;
  // one latency chain per SM whenever the grid fits that way (see launch_sweep_t)
  if (2 * nb * kClusterCtas <= num_sms() && smem < 116 * 1024) smem = 116 * 1024;
  auto kern = ctc_sweep_cluster_kernel;
  B200_CUDA(set_max_smem_cached((const void*)kern, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)nb, 2, kClusterCtas);
  cfg.blockDim = dim3(kClusterThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 1;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = kClusterCtas;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 2 : 1;
  B200_CUDA(cudaLaunchKernelEx(&cfg, kern, p));
  return B200_OK;
}

}  // namespace b200
