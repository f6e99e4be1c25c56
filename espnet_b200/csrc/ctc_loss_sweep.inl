// ctc_loss_sweep.inl -- the alpha / beta lattice recursion (included by ctc_loss.cu).
//
// One CTA per utterance runs BOTH sweeps: thread i owns the logical states 2i (blank) and
// 2i+1 (label i) of the alpha recursion and, independently, of the beta recursion, which is
// the same recursion on the time- and label-reversed utterance (its states are stored in that
// reversed order).  The two dependency chains interleave in the same warps, so the second
// sweep rides in the issue slots the first one leaves empty.
//
// Per frame and chain the critical path is: shfl.up of the neighbour's label state (one shared
// memory word per warp boundary) -> ex2 -> lg2 -> add -> store of the warp boundary ->
// bar.sync.  Everything else is off that path:
//  * the label-gathered log-prob rows lpg[b][t][0..L] are staged in shared memory kSweepRing
//    chunks of CHF frames ahead by 1-D bulk copies (cp.async.bulk -> UBLKCP) signalling an
//    mbarrier, so the loop never waits on global memory;
//  * every frame subtracts an integer drift estimate, and once per chunk the measured maximum
//    over all states (REDUX.MAX) is removed, so values stay O(10..50) in log2 units and fp32
//    keeps ~4e-6 absolute resolution for any T; the integers are summed exactly and only enter
//    log p(y|x) at the end (the occupancy kernel normalises per frame and never needs them).
namespace b200 {

constexpr int kSweepRing = 4;

struct SweepChain {
  float ab, al;        // blank / label state of this thread
  float dstep;         // integer drift removed every frame
  float level_prev, corr_prev;
  int csum;            // exact sum of everything removed so far
  float skipmask;      // 0 if the s-2 transition is allowed, -1e30 otherwise
  float* outp;         // where this thread's float2 of the current frame goes
  int64_t ostep;       // +-Sp
  int kcol;            // column of this thread's label in an lpg row
};

// log2(2^hi * (1 + e0) + 2^c) where e0 = 2^(lo - hi) was computed off the critical path
__device__ __forceinline__ float lse3_pre(float hi, float s1, float c) {
  const float t = ex2f(-fabsf(c - hi));
  const float sum = (c <= hi) ? (s1 + t) : fmaf(s1, t, 1.0f);
  return fmaxf(hi, c) + lg2f(sum);
}
__device__ __forceinline__ float lse2_abs(float a, float b) {
  return fmaxf(a, b) + lg2f(1.0f + ex2f(-fabsf(a - b)));
}

template <int CHF, int MAXT>
__global__ void __launch_bounds__(MAXT) ctc_sweep_kernel(LossParams p) {
  extern __shared__ __align__(128) unsigned char sw_smem[];
  const int b = p.b0 + blockIdx.x;
  const int Tb = (int)p.ilen[b];
  const int L = (int)p.olen[b];
  const int i = threadIdx.x;
  const int lane = i & 31, warp = i >> 5;
  const int nw = blockDim.x >> 5;
  const int64_t Kp = p.ws.Kp, Sp = p.ws.Sp;

  __shared__ float s_bnd[2][2][36];   // [dir][frame parity][warp]; slot 32 holds log(0) for warp 0
  __shared__ float s_wmax[2][2][32];  // [dir][chunk parity][warp]
  __shared__ float s_fin[2][2];
  uint64_t* full = reinterpret_cast<uint64_t*>(sw_smem);  // [2][kSweepRing]
  float* ring = reinterpret_cast<float*>(sw_smem + 128);  // [2][kSweepRing][CHF*Kp]
  const int chunk_floats = CHF * (int)Kp;

  if (Tb <= 0) {
    if (i == 0) {
      p.nll[b] = p.ws.nll_ws[b] = (L == 0) ? 0.0f : __int_as_float(0x7f800000);
      p.ws.logz_beta[b] = (L == 0) ? 0.0f : B200_NEG;
    }
    return;
  }
  for (int k = i; k < 2 * 2 * 36; k += blockDim.x) (&s_bnd[0][0][0])[k] = B200_NEG;
  for (int k = i; k < 2 * 2 * 32; k += blockDim.x) (&s_wmax[0][0][0])[k] = B200_NEG;
  if (i == 0) {
    for (int s = 0; s < 2 * kSweepRing; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  __syncthreads();

  const float* lpg = p.ws.lpg + (int64_t)b * p.T * Kp;
  const int nchunks = (Tb + CHF - 1) / CHF;
  auto issue_chunk = [&](int c) {  // thread 0 only: chunk c of both directions
    const int f0 = c * CHF;
    const int n = min(CHF, Tb - f0);
    const int slot = c % kSweepRing;
    const uint32_t bytes = (uint32_t)(n * Kp * sizeof(float));
#pragma unroll
    for (int d = 0; d < 2; ++d) {
      const int phys0 = d ? (Tb - f0 - n) : f0;
      uint64_t* bar = &full[d * kSweepRing + slot];
      mbar_arrive_expect_tx(bar, bytes);
      bulk_g2s(ring + ((size_t)d * kSweepRing + slot) * chunk_floats, lpg + (int64_t)phys0 * Kp, bytes, bar);
    }
  };
  if (i == 0)
    for (int c = 0; c < kSweepRing - 1 && c < nchunks; ++c) issue_chunk(c);

  const int* lab = p.ws.labels + (int64_t)b * Kp;
  const bool has_label = i < L;
  const bool active = i <= L;
  SweepChain ch[2];
#pragma unroll
  for (int d = 0; d < 2; ++d) {
    const int pi = d ? (L - 1 - i) : i;  // physical label position of logical position i
    bool skip_ok = false;
    if (has_label && i >= 1) skip_ok = lab[pi] != lab[d ? pi + 1 : pi - 1];
    ch[d].skipmask = skip_ok ? 0.0f : B200_NEG;
    ch[d].kcol = has_label ? 1 + pi : 0;
    ch[d].ab = B200_NEG; ch[d].al = B200_NEG;
    ch[d].dstep = 0.f; ch[d].level_prev = 0.f; ch[d].corr_prev = 0.f; ch[d].csum = 0;
    ch[d].outp = p.ws.lat + (((int64_t)d * p.B + b) * p.T + (d ? (Tb - 1) : 0)) * Sp + 2 * i;
    ch[d].ostep = d ? -Sp : Sp;
  }
  const int bsrc = warp ? warp - 1 : 32;
  const bool lane0 = lane == 0, lane31 = lane == 31;

  for (int c = 0; c < nchunks; ++c) {
    const int slot = c % kSweepRing;
    const int f0 = c * CHF;
    const int n = min(CHF, Tb - f0);
    // fill the slot freed by the previous chunk (its last frame ended with a barrier)
    if (i == 0 && c + kSweepRing - 1 < nchunks) issue_chunk(c + kSweepRing - 1);
    float vb[2][CHF], vl[2][CHF];
#pragma unroll
    for (int d = 0; d < 2; ++d) {
      mbar_wait(&full[d * kSweepRing + slot], (uint32_t)((c / kSweepRing) & 1));
      const float* buf = ring + ((size_t)d * kSweepRing + slot) * chunk_floats;
#pragma unroll
      for (int f = 0; f < CHF; ++f) {
        if (f < n) {
          const float* rowp = buf + (size_t)(d ? (n - 1 - f) : f) * Kp;
          vb[d][f] = rowp[0];
          vl[d][f] = has_label ? rowp[ch[d].kcol] : B200_NEG;  // a missing label state stays at log(0)
        } else {
          vb[d][f] = 0.f; vl[d][f] = 0.f;
        }
      }
    }
    if (c == 0) {
#pragma unroll
      for (int d = 0; d < 2; ++d) {
        // frame 0: only the first blank and the first label are reachable
        if (i == 0) { ch[d].ab = vb[d][0]; ch[d].al = vl[d][0]; }
        if (active) *reinterpret_cast<float2*>(ch[d].outp) = make_float2(ch[d].ab, ch[d].al);
        ch[d].outp += ch[d].ostep;
        if (lane31) s_bnd[d][0][warp] = ch[d].al;
        if (n > 1) ch[d].dstep = floorf(vb[d][1]);
      }
    } else {
#pragma unroll
      for (int d = 0; d < 2; ++d) {
        const float wm = warp_max_redux(fmaxf(ch[d].ab, ch[d].al));
        if (lane0) s_wmax[d][c & 1][warp] = wm;
      }
    }
    if (nw > 1) __syncthreads(); else __syncwarp();
    float corr[2] = {0.f, 0.f};
    if (c > 0) {
#pragma unroll
      for (int d = 0; d < 2; ++d) {
        const float level = warp_max_redux(s_wmax[d][c & 1][lane]);
        corr[d] = floorf(level);
        // drift per frame seen over the last chunk, relative to the dstep that was applied
        ch[d].dstep = floorf(ch[d].dstep + (level - ch[d].level_prev + ch[d].corr_prev) * (1.0f / CHF) + 0.5f);
        ch[d].level_prev = level;
        ch[d].corr_prev = corr[d];
      }
    }
#pragma unroll
    for (int f = 0; f < CHF; ++f) {
      if (f < n && !(c == 0 && f == 0)) {
        const int par = (CHF & 1) ? ((f0 + f) & 1) : (f & 1);
        // neighbour's label state of the previous frame: lane-1 by shuffle, across the warp
        // boundary through one shared-memory word (loaded by every lane, in parallel with the shuffle)
        float nbr[2];
#pragma unroll
        for (int d = 0; d < 2; ++d) {
          const float bn = s_bnd[d][par ^ 1][bsrc];
          const float sh = __shfl_up_sync(0xffffffffu, ch[d].al, 1);
          nbr[d] = lane0 ? bn : sh;
        }
#pragma unroll
        for (int d = 0; d < 2; ++d) {
          SweepChain& s = ch[d];
          const float o = s.dstep + ((f == 0) ? corr[d] : 0.f);
          const float hi = fmaxf(s.al, s.ab);
          const float s1 = 1.0f + ex2f(-fabsf(s.al - s.ab));
          const float lb = vb[d][f] - o, ll = vl[d][f] - o;
          const float x = lse2_abs(s.ab, nbr[d]);
          const float y = lse3_pre(hi, s1, nbr[d] + s.skipmask);
          s.ab = lb + x;
          s.al = ll + y;
          if (lane31) s_bnd[d][par][warp] = s.al;
          if (active) *reinterpret_cast<float2*>(s.outp) = make_float2(s.ab, s.al);
          s.outp += s.ostep;
        }
        if (nw > 1) __syncthreads(); else __syncwarp();
      }
    }
    // exact bookkeeping of what this chunk removed (frame 0 of the utterance removes nothing)
#pragma unroll
    for (int d = 0; d < 2; ++d) ch[d].csum += (n - (c == 0 ? 1 : 0)) * (int)ch[d].dstep + (int)corr[d];
  }

  // log2 p(y|x) = lse2(state 2L, state 2L-1) + sum of offsets, from both sweeps
#pragma unroll
  for (int d = 0; d < 2; ++d) {
    if (i == L) s_fin[d][0] = ch[d].ab;
    if (L == 0 && i == 0) s_fin[d][1] = B200_NEG;
    if (L > 0 && i == L - 1) s_fin[d][1] = ch[d].al;
  }
  __syncthreads();
  if (i == 0) {
#pragma unroll
    for (int d = 0; d < 2; ++d) {
      const float z = lse2(s_fin[d][0], s_fin[d][1]);
      const bool feasible = z > -1.0e29f;
      const float logz2 = z + (float)ch[d].csum;
      if (d == 0) p.nll[b] = p.ws.nll_ws[b] = feasible ? -B200_LN2 * logz2 : __int_as_float(0x7f800000);
      else p.ws.logz_beta[b] = feasible ? logz2 : B200_NEG;
    }
  }
}

struct SweepCfg {
  int threads, chf, smem;
};

inline SweepCfg plan_sweep(const LossParams& p) {
  SweepCfg c;
  c.threads = (int)align_up((size_t)(p.Lmax + 1), 32);
  int chf = (int)(8192 / (p.ws.Kp * sizeof(float)));
  c.chf = chf >= 8 ? 8 : (chf >= 4 ? 4 : (chf >= 2 ? 2 : 1));
  c.smem = 128 + 2 * kSweepRing * c.chf * (int)p.ws.Kp * (int)sizeof(float);
  return c;
}

template <int CHF, int MAXT>
static int launch_sweep_t(const LossParams& p, const SweepCfg& c, cudaStream_t st) {
  auto kern = ctc_sweep_kernel<CHF, MAXT>;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, c.smem));
  kern<<<(unsigned)(p.b1 - p.b0), c.threads, c.smem, st>>>(p);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

static int launch_sweep(const LossParams& p, cudaStream_t st) {
  const SweepCfg c = plan_sweep(p);
  if (c.threads <= 256) {
    switch (c.chf) {
      case 8: return launch_sweep_t<8, 256>(p, c, st);
      case 4: return launch_sweep_t<4, 256>(p, c, st);
      case 2: return launch_sweep_t<2, 256>(p, c, st);
      default: return launch_sweep_t<1, 256>(p, c, st);
    }
  }
  switch (c.chf) {
    case 8: return launch_sweep_t<8, 1024>(p, c, st);
    case 4: return launch_sweep_t<4, 1024>(p, c, st);
    case 2: return launch_sweep_t<2, 1024>(p, c, st);
    default: return launch_sweep_t<1, 1024>(p, c, st);
  }
}

}  // namespace b200
