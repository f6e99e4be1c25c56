// ctc_loss_sweep.inl -- the alpha / beta lattice recursion (included by ctc_loss.cu).
//
// One CTA per (utterance, direction).  Thread i owns the logical states 2i (blank) and 2i+1
// (label i); the beta sweep is the same recursion on the time- and label-reversed utterance
// (its states are stored in that reversed order).  The recursion only flows from low to high
// states, so the warps of a CTA run SKEWED by one chunk of CHF frames instead of in lock step:
// warp w starts chunk c when warp w-1 has published the label state of its last lane for the
// frames of that chunk (a ring of floats in shared memory plus one progress word per warp).
// There is no barrier and no branch inside a full chunk: the CHF boundary values and the CHF
// log-prob pairs of the chunk are in registers before its first frame, and the per-frame
// critical path is shfl.up -> ex2 -> lg2 -> add inside a warp.
//
//  * the label-gathered log-prob rows lpg[b][t][0..L] are read straight into registers three
//    chunks ahead (each lane needs only its own column and the blank column);
//  * every frame subtracts an integer drift estimate and every chunk an integer correction,
//    both derived from the maximum over all states that every warp publishes per chunk (read
//    `delay` chunks later, when even the most skewed warp has published it), so values stay
//    O(10) in log2 units and fp32 keeps ~4e-6 absolute resolution for any T; all warps derive
//    the same integers from the same published data, their exact sum only enters log p(y|x)
//    at the end (the occupancy kernels normalise per frame and never need it).
namespace b200 {

// A value and the index of the frame (or chunk) it belongs to, written and read as ONE 8-byte
// shared-memory access: a record validates itself, so the hand-off needs no fence (a fence would
// also wait for the lattice stores to global memory that are in flight).
struct __align__(8) Tagged {
  float v;
  int tag;
};
__device__ __forceinline__ Tagged lds_tagged(const Tagged* p) {
  Tagged r;
  asm volatile("ld.volatile.shared.v2.b32 {%0,%1}, [%2];" : "=f"(r.v), "=r"(r.tag) : "r"(smem_u32(p)));
  return r;
}
__device__ __forceinline__ void sts_tagged(Tagged* p, float v, int tag) {
  asm volatile("st.volatile.shared.v2.b32 [%0], {%1,%2};" ::"r"(smem_u32(p)), "f"(v), "r"(tag) : "memory");
}

// log2(2^hi * s1 + 2^c) where s1 = 1 + 2^(lo - hi) was computed off the critical path
__device__ __forceinline__ float lse3_pre(float hi, float s1, float c) {
  const float t = ex2f(-fabsf(c - hi));
  const float sum = (c <= hi) ? (s1 + t) : fmaf(s1, t, 1.0f);
  return fmaxf(hi, c) + lg2f(sum);
}
__device__ __forceinline__ float lse2_abs(float a, float b) {
  return fmaxf(a, b) + lg2f(1.0f + ex2f(-fabsf(a - b)));
}
__device__ __forceinline__ float clamp_int_f(float x) {  // integer-valued, safely inside int32/fp32-exact range
  return fminf(fmaxf(rintf(x), -4194304.0f), 4194304.0f);
}

// Final value of utterance b; the CTA that delivers the last one also sums them (fixed order:
// the sum is deterministic).
// One call per utterance and pass (by its alpha CTA) once the utterance's nll is final for that pass; the caller that
// delivers the last one sums them in a fixed order.  (The redo pass of the log-domain kernels counts separately.)
__device__ __forceinline__ void sweep_count_nll(const LossParams& p) {
  if (p.loss_sum == nullptr) return;
  __threadfence();
  const int done = atomicAdd(p.ws.bad + (p.redo_mode ? 6 : 4), 1);
  if (done == (p.b1 - p.b0) - 1) {
    __threadfence();
    float s = 0.f;
    for (int o = p.b0; o < p.b1; ++o) s += ld_cg_f32(p.nll + o);
    *p.loss_sum = s * p.loss_scale;
  }
}
__device__ __forceinline__ void sweep_finish_nll(const LossParams& p, int b, float nll) {
  if (p.ws.badutt[b]) nll = __int_as_float(0x7fc00000);  // invalid labels / olen > Lmax: loud, not silent
  p.nll[b] = nll;
  p.ws.nll_ws[b] = nll;
  sweep_count_nll(p);
}

template <int MAXT> struct SweepRings {
  // hand-off ring in frames: > (maximum skew = delay <= MAXT/32 chunks) + 2 chunks; level ring in
  // chunks: > 2 * delay + 2.  Powers of two so that the ring indices are masks.
  static constexpr int kCHF = MAXT <= 256 ? 8 : 4;
  static constexpr int kR = MAXT <= 256 ? 128 : 256;
  static constexpr int kWR = MAXT <= 256 ? 32 : 128;
  static_assert(kR >= kCHF * (MAXT / 32 + 3) && kWR >= 2 * (MAXT / 32) + 4, "ring sizes");
};

template <int CHF, int MAXT>
__global__ void __launch_bounds__(MAXT) ctc_sweep_kernel(LossParams p) {
  extern __shared__ __align__(16) unsigned char sw_smem[];
  constexpr int kR = SweepRings<MAXT>::kR, kWR = SweepRings<MAXT>::kWR;
  static_assert(CHF == SweepRings<MAXT>::kCHF, "chunk size");
  // order / labels come from the prep kernel, which runs under the stream kernel of the same call
  pdl_wait();
  // the occupancy kernel may be staged from here on: it starts after EVERY sweep CTA has passed the wait above,
  // i.e. with prep complete (its counters zeroed), and follows the per-utterance completion flags
  pdl_launch_dependents();
  const int b = p.ws.order[p.b0 + blockIdx.x];
  const int d = blockIdx.y;
  if (p.redo_mode && !p.ws.redo[b]) {  // only what the linear sweeps could not certify
    if (threadIdx.x == 0 && d == 0) sweep_count_nll(p);
    return;
  }
  const int Tb = min((int)p.ilen[b], (int)p.T);
  const int L = utt_labels(p, b);
  const int i = threadIdx.x;
  const int lane = i & 31, warp = i >> 5;
  const int nw = blockDim.x >> 5;
  const int Kp = (int)p.ws.Kp;
  const int64_t Sp = p.ws.Sp;
  const int nwa = L / 32 + 1;  // warps that own at least one state (threads 0..L)
  Tagged* s_bnd = reinterpret_cast<Tagged*>(sw_smem);  // [nw][kR] label state of lane 31 of warp w at frame f
  Tagged* s_wmx = s_bnd + (size_t)nw * kR;             // [kWR][nw] maximum over the states of warp w at the start of chunk c
  int* s_hist = reinterpret_cast<int*>(s_wmx + (size_t)kWR * nw);  // [kWR] total removed before chunk c
  __shared__ float s_fin[2];

  if (Tb <= 0) {
    if (i == 0) {
      if (d == 0) sweep_finish_nll(p, b, (L == 0) ? 0.0f : __int_as_float(0x7f800000));
      else p.ws.logz_beta[b] = (L == 0) ? 0.0f : B200_NEG;
    }
    pdl_wait();
    return;
  }
  for (int k = i; k < nw * kR + kWR * nw; k += blockDim.x) s_bnd[k].tag = -1;
  __syncthreads();
  if (warp >= nwa) return;

  const int* lab = p.ws.labels + (int64_t)b * Kp;
  const bool has_label = i < L;
  const bool active = i <= L;
  const int pi = d ? (L - 1 - i) : i;  // physical label position of logical position i
  bool skip_ok = false;
  if (has_label && i >= 1) skip_ok = lab[pi] != lab[d ? pi + 1 : pi - 1];
  const float skipmask = skip_ok ? 0.0f : B200_NEG;
  const int kcol = has_label ? 1 + pi : 0;
  // logical frame f is physical frame f (alpha) or Tb-1-f (beta): rows of lpg and of the lattice
  const float* lp0 = p.ws.lpg + ((int64_t)b * p.T + (d ? (Tb - 1) : 0)) * Kp;
  const int lstep = d ? -Kp : Kp;
  float* outp = p.ws.lat + (((int64_t)d * p.B + b) * p.T + (d ? (Tb - 1) : 0)) * Sp + 2 * i;
  const int64_t ostep = d ? -Sp : Sp;
  const bool lane0 = lane == 0, lane31 = lane == 31;
  const bool poll = warp > 0;
  const Tagged* bsrc = s_bnd + (size_t)(poll ? warp - 1 : 0) * kR;
  Tagged* bdst = s_bnd + (size_t)warp * kR;
  // the most skewed warp publishes its level of chunk m when warp 0 is around chunk m + nwa - 1
  const int delay = nwa;

  float ab = B200_NEG, al = B200_NEG;
  float dstep = 0.f;
  int csum = 0;                      // exact sum of everything removed so far
  float xprev = 0.f; int sprev = 0;  // previous consumed level measurement and its csum
  bool have_prev = false;

  // one frame of the recursion; bn = label state of the last lane of the previous warp at the previous frame
  auto frame = [&](float lb_t, float ll_t, float bn, float o, int fr) {
    const float sh = __shfl_up_sync(0xffffffffu, al, 1);
    const float nbr = lane0 ? bn : sh;
    const float hi = fmaxf(al, ab);
    const float s1 = 1.0f + ex2f(-fabsf(al - ab));
    const float x = lse2_abs(ab, nbr);
    const float y = lse3_pre(hi, s1, nbr + skipmask);
    ab = (lb_t - o) + x;
    al = (ll_t - o) + y;
    if (lane31) sts_tagged(&bdst[fr & (kR - 1)], al, fr);
    if (active) *reinterpret_cast<float2*>(outp) = make_float2(ab, al);
    outp += ostep;
  };
  // start of chunk c >= 1 (n frames): publish this warp's level, consume everybody's of chunk
  // c - delay, derive the integer removals of this chunk (dstep per frame, corr once)
  auto chunk_begin = [&](int c, int n) -> float {
    // levels are published and consumed every kP-th chunk only (this is per-chunk overhead on the
    // critical path of every warp); dstep holds in between
    constexpr int kP = 2;
    if (c % kP) return 0.f;
    float corr = 0.f;
    const float wm = warp_max_redux(fmaxf(ab, al));
    if (lane31) {
      sts_tagged(&s_wmx[(c & (kWR - 1)) * nw + warp], wm, c);
      s_hist[c & (kWR - 1)] = csum;
    }
    __syncwarp();
    const int delay_p = (delay + kP - 1) / kP * kP;
    const int cm = c - delay_p;
    if (cm >= kP) {
      Tagged e;
      e.v = B200_NEG; e.tag = cm;
      if (lane < nwa) {
        const Tagged* src = &s_wmx[(cm & (kWR - 1)) * nw + lane];
        do { e = lds_tagged(src); } while (e.tag != cm);
      }
      __syncwarp();
      const float xm = warp_max_redux(e.v);
      const int sm = s_hist[cm & (kWR - 1)];
      // uncentred level y = x + S; drift per frame from two consecutive measurements
      float delta = dstep;
      if (have_prev) delta = (xm - xprev + (float)(sm - sprev)) * (1.0f / (kP * CHF));
      xprev = xm; sprev = sm; have_prev = true;
      const float xhat = xm + (float)(sm - csum) + delta * (float)(CHF * delay_p);  // level now
      // remove, over the next kP chunks, what brings the level back to zero at their end
      const float horizon = (float)(n == CHF ? kP * CHF : n);
      dstep = clamp_int_f(delta);
      corr = clamp_int_f(clamp_int_f(xhat + delta * horizon) - dstep * horizon);
    }
    return corr;
  };
  auto load_lp = [&](const float* q) -> float { return ld_relaxed_f32(q); };
  // generic (slow) chunk: the first chunk and a trailing partial chunk; reads lpg directly
  auto slow_chunk = [&](int c, int n) {
    const int f0 = c * CHF;
    float corr = 0.f;
    if (c == 0) {
      if (Tb > 1) dstep = clamp_int_f(floorf(load_lp(lp0 + lstep)));
      if (lane31) s_hist[0] = 0;
    } else {
      corr = chunk_begin(c, n);
    }
    // all loads of the chunk first (one memory latency, not one per frame)
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
          // frame 0: only the first blank and the first label are reachable
          if (i == 0) { ab = lb_t; al = ll_t; }
          if (lane31) sts_tagged(&bdst[0], al, 0);
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
  const int nfull = Tb / CHF;  // chunks 1 .. nfull-1 are full and take the fast path
  // registers of one full chunk, loaded two chunks ahead of their use
  auto load_chunk = [&](int c, float(&xb)[CHF], float(&xl)[CHF]) {
    const float* rowp = lp0 + (int64_t)(c * CHF) * lstep;
#pragma unroll
    for (int f = 0; f < CHF; ++f) {
      xb[f] = ld_relaxed_f32(rowp);
      xl[f] = has_label ? ld_relaxed_f32(rowp + kcol) : B200_NEG;  // a missing label state stays at log(0)
      rowp += lstep;
    }
  };
  // full chunk c >= 1: straight-line code, no branch inside the frames
  // nb / nl: the register set that is free during this chunk; it receives chunk c + 2 right after
  // the first use of xb / xl, so that a whole chunk of time lies between the issue of any load and
  // the next wait on a load scoreboard (ptxas may give several load batches the same scoreboard)
  auto fast_chunk = [&](int c, float(&xb)[CHF], float(&xl)[CHF], float(&nb)[CHF], float(&nl)[CHF]) {
    const int f0 = c * CHF;
    // first use of this chunk's registers (waits for their loads), then the prefetch
    float first = xb[0] + xl[0];
#pragma unroll
    for (int f = 1; f < CHF; ++f) first += xb[f] + xl[f];
    asm volatile("" ::"f"(first));
    if (c + 2 < nfull) load_chunk(c + 2, nb, nl);
    const float corr = chunk_begin(c, CHF);
    float bn[CHF];
#pragma unroll
    for (int f = 0; f < CHF; ++f) bn[f] = B200_NEG;
    if (poll) {
      // boundary values of frames f0-1 .. f0+CHF-2 from the previous warp
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
#pragma unroll
    for (int f = 0; f < CHF; ++f) frame(xb[f], xl[f], bn[f], dstep + (f == 0 ? corr : 0.f), f0 + f);
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

  // log2 p(y|x) = lse2(state 2L, state 2L-1) + sum of offsets
  if (i == L) s_fin[0] = ab;
  if (L == 0 && i == 0) s_fin[1] = B200_NEG;
  if (L > 0 && i == L - 1) s_fin[1] = al;
  asm volatile("bar.sync 1, %0;" ::"r"(nwa * 32) : "memory");
  if (i == 0) {
    const float z = lse2(s_fin[0], s_fin[1]);
    const bool feasible = z > -1.0e29f;
    const float logz2 = z + (float)csum;
    if (d == 0) sweep_finish_nll(p, b, feasible ? -B200_LN2 * logz2 : __int_as_float(0x7f800000));
    else p.ws.logz_beta[b] = feasible ? logz2 : B200_NEG;
    // this direction's lattice of utterance b is complete (every state warp stored its last frame before the
    // barrier above): the occupancy kernel may read it once both directions have signalled
    __threadfence();
    atomicAdd(p.ws.done + b, 1);
  }
}

// ------------------------------------------------------------------------------------------------
// The same recursion for L <= 255 with everything that is not the recursion itself moved to a HELPER
// WARP (the last warp of the CTA), because a state warp is one latency chain and every wait it executes
// is added to the chain:
//  * the log-prob rows of every full chunk are STAGED IN SHARED MEMORY: helper lane 0 brings the CHF rows
//    of a chunk (contiguous in lpg) in with one 1-D bulk (TMA) copy, kLpAhead chunks ahead, into a ring of
//    kLpSlots chunks, and the helper alone waits on the slots' mbarriers;
//  * the helper consumes the level records of the state warps and derives the integer removals
//    (dstep per frame, corr once) of every chunk; it publishes them as self-validating control records
//    {dstep, c}, {corr, c}, and only after the rows of chunk c + 1 have landed;
//  * a state warp reads the control records of chunk c + 1 and the log-prob columns of chunk c + 1 (LDS into
//    a second register set) at the START of chunk c, so neither latency is on its chain: per chunk it
//    executes two tag compares, 18 LDS issues, one REDUX + STS every second chunk and the boundary poll.
// Slot reuse needs no "empty" barrier: the helper stages chunk s only after it has consumed the level
// records of chunk >= s - kLpSlots of every warp (a warp publishes its level of chunk m after the last
// read of the rows of chunk m has long returned), see the inequality at the stage() call.
constexpr int kLpSlots = 16;   // chunks of rows [CHF frames][Kp columns] as the bulk copies deliver them
constexpr int kLpAhead = 4;    // bulk copies are issued this many chunks ahead of their first reader
constexpr int kCtlRing = 32;   // control records, in chunks

// (pick ? a : b) - c as two predicated subtractions: the one on `b` does not wait for `a` and vice versa
__device__ __forceinline__ float sub_pick(bool pick, float a, float b, float c) {
  float r;
  asm("{\n.reg .pred p;\nsetp.ne.s32 p, %4, 0;\n@p sub.f32 %0, %1, %3;\n@!p sub.f32 %0, %2, %3;\n}"
      : "=f"(r) : "f"(a), "f"(b), "f"(c), "r"((int)pick));
  return r;
}
__device__ __forceinline__ float lds_f32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ float4 lds_v4(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}

#ifdef B200_SWEEP_CLOCKS
// developer instrumentation: SM clock at four points of every full chunk, per direction and warp, of the
// first (longest) utterance; read back with b200_debug_sweep_clocks
__device__ long long g_sweep_clk[2 * 8 * 4 * 64];
#define SWEEP_STAMP(k) do { if (blockIdx.x == 0 && lane == 0 && c < 64) g_sweep_clk[((d * 8 + warp) * 4 + (k)) * 64 + c] = clock64(); } while (0)
#else
#define SWEEP_STAMP(k) do { } while (0)
#endif

__global__ void __launch_bounds__(288) ctc_sweep_staged_kernel(LossParams p) {
  constexpr int CHF = 8, MAXT = 256, kP = 2;
  extern __shared__ __align__(16) unsigned char sw_smem[];
  constexpr int kR = SweepRings<MAXT>::kR, kWR = SweepRings<MAXT>::kWR;
  static_assert(CHF == SweepRings<MAXT>::kCHF, "chunk size");
  const int d = blockIdx.y;
  const int i = threadIdx.x;
  const int lane = i & 31, warp = i >> 5;
  const int nw = (blockDim.x >> 5) - 1;  // state warps launched; warp nw is the helper
  const int Kp = (int)p.ws.Kp;
  const int64_t Sp = p.ws.Sp;
  Tagged* s_bnd = reinterpret_cast<Tagged*>(sw_smem);  // [nw][kR] label state of lane 31 of warp w at frame f
  Tagged* s_wmx = s_bnd + (size_t)nw * kR;             // [kWR][nw] maximum over the states of warp w at the start of chunk c
  Tagged* s_ctl = s_wmx + (size_t)kWR * nw;            // [kCtlRing][2] {dstep, c}, {corr, c}
  int* s_hist = reinterpret_cast<int*>(s_ctl + 2 * kCtlRing);  // [kWR] total removed before chunk c (helper only)
  const int slot_floats = CHF * Kp;
  float* s_lp = reinterpret_cast<float*>(s_hist + kWR);  // [kLpSlots][CHF][Kp] rows of lpg, chunk c >= 1 in slot (c - 1) % kLpSlots
  float* s_blk = s_lp + (size_t)kLpSlots * slot_floats;  // [kLpSlots][CHF] blank column of the chunk, frames in logical order
  uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_blk + kLpSlots * CHF);
  __shared__ float s_fin[2];
  __shared__ int s_csum;
  __shared__ float s_neg;  // log(0), read as the label column of threads without a label

  // shared-memory set-up overlaps the tail of the previous kernels; everything after the wait may use the
  // lpg rows of the stream kernel and the results of the prep kernel (utterance order, labels), which runs
  // under the stream kernel
  for (int k = i; k < nw * kR + kWR * nw + 2 * kCtlRing; k += blockDim.x) s_bnd[k].tag = -1;
  if (i == 0) {
    s_neg = B200_NEG;
    for (int k = 0; k < kLpSlots; ++k) mbar_init(&s_bar[k], 1);
    mbar_fence_init();
    fence_proxy_async();
  }
  pdl_wait();
  if (threadIdx.x == 0) TL_MIN(2);
  // the occupancy kernel may be staged from here on: it starts after EVERY sweep CTA has passed the wait above,
  // i.e. with prep complete (its counters zeroed), and follows the per-utterance completion flags
  pdl_launch_dependents();
  __syncthreads();
  const int b = p.ws.order[p.b0 + blockIdx.x];
  if (p.redo_mode && !p.ws.redo[b]) {  // only what the linear sweeps could not certify
    if (i == 0 && d == 0) sweep_count_nll(p);
    return;
  }
  const int Tb = min((int)p.ilen[b], (int)p.T);
  const int L = utt_labels(p, b);
  const int nwa = L / 32 + 1;  // warps that own at least one state (threads 0..L)
  if (Tb <= 0) {
    if (i == 0) {
      if (d == 0) sweep_finish_nll(p, b, (L == 0) ? 0.0f : __int_as_float(0x7f800000));
      else p.ws.logz_beta[b] = (L == 0) ? 0.0f : B200_NEG;
    }
    return;
  }
  const int nfull = Tb / CHF;  // chunks 1 .. nfull-1 are full and take the fast path
  // logical frame f is physical frame f (alpha) or Tb-1-f (beta): rows of lpg and of the lattice
  const float* lp0 = p.ws.lpg + ((int64_t)b * p.T + (d ? (Tb - 1) : 0)) * Kp;
  const int lstep = d ? -Kp : Kp;
  // the most skewed warp publishes its level of chunk m when warp 0 is around chunk m + nwa - 1; one more
  // pair of chunks lets the control record of chunk c be there when chunk c - 1 starts
  const int delay_p = (nwa + kP - 1) / kP * kP;
  if (warp >= nwa && warp != nw) return;

  if (warp == nw) {
    // ------------------------------------------------------------------ helper warp
    const float* lpg_b = p.ws.lpg + (int64_t)b * p.T * Kp;
    // rows of chunk c (logical frames c*CHF ..): physical rows c*CHF.. (alpha) or Tb-1-c*CHF-(CHF-1).. (beta)
    auto stage = [&](int c) {
      const int slot = (c - 1) % kLpSlots;
      const int prow = d ? (Tb - 1 - c * CHF - (CHF - 1)) : c * CHF;
      mbar_arrive_expect_tx(&s_bar[slot], (uint32_t)(slot_floats * 4));
      bulk_g2s(s_lp + (size_t)slot * slot_floats, lpg_b + (int64_t)prow * Kp, (uint32_t)(slot_floats * 4), &s_bar[slot]);
    };
    // wait for the rows of chunk c; the blank column (which every state thread needs) is packed so that a
    // thread reads the 8 values of a chunk with two 16-byte loads
    auto landed = [&](int c) {
      const int slot = (c - 1) % kLpSlots;
      mbar_wait(&s_bar[slot], (uint32_t)(((c - 1) / kLpSlots) & 1));
      if (lane < CHF) s_blk[slot * CHF + lane] = s_lp[(size_t)slot * slot_floats + (d ? (CHF - 1 - lane) : lane) * Kp];
      __syncwarp();
    };
    auto publish = [&](int c, float dstep, float corr) {
      if (lane == 0) {
        sts_tagged(&s_ctl[(c & (kCtlRing - 1)) * 2 + 1], corr, c);
        sts_tagged(&s_ctl[(c & (kCtlRing - 1)) * 2], dstep, c);
      }
    };
    if (lane == 0)
      for (int c = 1; c <= kLpAhead + 1 && c < nfull; ++c) stage(c);
    const int clast = nfull >= 1 ? (Tb > nfull * CHF ? nfull : nfull - 1) : 0;
    float dstep = 0.f;
    if (Tb > 1) dstep = clamp_int_f(floorf(ld_relaxed_f32(lp0 + lstep)));
    int csum = (min(CHF, Tb) - 1) * (int)dstep;  // exact sum of everything removed so far
    float xprev = 0.f; int sprev = 0;            // previous consumed level measurement and its csum
    bool have_prev = false;
    if (lane == 0) s_hist[0] = 0;
    if (1 < nfull) landed(1);
    publish(0, dstep, 0.f);
    for (int c = 1; c <= clast; ++c) {
      const int n = c < nfull ? CHF : Tb - nfull * CHF;
      float corr = 0.f;
      SWEEP_STAMP(0);
      if (c % kP == 0) {
        if (lane == 0) s_hist[c & (kWR - 1)] = csum;
        __syncwarp();
        const int cm = c - delay_p;
        if (cm >= kP) {
          Tagged e;
          e.v = B200_NEG; e.tag = cm;
          if (lane < nwa) {
            const Tagged* src = &s_wmx[(cm & (kWR - 1)) * nw + lane];
            do { e = lds_tagged(src); } while (e.tag != cm);
          }
          __syncwarp();
          const float xm = warp_max_redux(e.v);
          const int sm = s_hist[cm & (kWR - 1)];
          // uncentred level y = x + S; drift per frame from two consecutive measurements
          float delta = dstep;
          if (have_prev) delta = (xm - xprev + (float)(sm - sprev)) * (1.0f / (kP * CHF));
          xprev = xm; sprev = sm; have_prev = true;
          const float xhat = xm + (float)(sm - csum) + delta * (float)(CHF * delay_p);  // level now
          // remove, over the next kP chunks, what brings the level back to zero at their end
          const float horizon = (float)(n == CHF ? kP * CHF : n);
          dstep = clamp_int_f(delta);
          corr = clamp_int_f(clamp_int_f(xhat + delta * horizon) - dstep * horizon);
        }
      }
      SWEEP_STAMP(1);
      if (c + 1 < nfull) landed(c + 1);
      SWEEP_STAMP(2);
      publish(c, dstep, corr);
      csum += n * (int)dstep + (int)corr;
      // chunk s = c + 1 + kLpAhead replaces chunk s - kLpSlots: every warp has published its level of the last
      // even chunk <= c - delay_p, i.e. is done with the rows of all chunks up to c - 1 - delay_p, and
      // c + 1 + kLpAhead - kLpSlots <= c - 1 - delay_p  <=>  delay_p <= kLpSlots - kLpAhead - 2
      static_assert(MAXT / 32 <= kLpSlots - kLpAhead - 2, "lp ring");
      if (lane == 0 && c + 1 + kLpAhead < nfull) stage(c + 1 + kLpAhead);
      SWEEP_STAMP(3);
    }
    if (lane == 0) s_csum = csum;
    asm volatile("bar.sync 1, %0;" ::"r"((nwa + 1) * 32) : "memory");
    return;
  }

  // -------------------------------------------------------------------- state warps
  const int* lab = p.ws.labels + (int64_t)b * Kp;
  const bool has_label = i < L;
  const bool active = i <= L;
  const int pi = d ? (L - 1 - i) : i;  // physical label position of logical position i
  bool skip_ok = false;
  if (has_label && i >= 1) skip_ok = lab[pi] != lab[d ? pi + 1 : pi - 1];
  const float skipmask = skip_ok ? 0.0f : B200_NEG;
  const int kcol = has_label ? 1 + pi : 0;
  // shared-memory addresses of this thread's label column in the staged rows: first logical frame of slot 0,
  // step per frame, step per slot; a thread without a label reads log(0) from one fixed word instead
  const uint32_t a_lab0 = has_label ? smem_u32(s_lp) + (uint32_t)(((d ? (CHF - 1) * Kp : 0) + kcol) * 4) : smem_u32(&s_neg);
  const uint32_t a_fstep = has_label ? (uint32_t)(lstep * 4) : 0u;
  const uint32_t a_sstep = has_label ? (uint32_t)(slot_floats * 4) : 0u;
  const uint32_t a_blk0 = smem_u32(s_blk);
  float* outp = p.ws.lat + (((int64_t)d * p.B + b) * p.T + (d ? (Tb - 1) : 0)) * Sp + 2 * i;
  const int64_t ostep = d ? -Sp : Sp;
  const bool lane0 = lane == 0, lane31 = lane == 31;
  const bool poll = warp > 0;
  const Tagged* bsrc = s_bnd + (size_t)(poll ? warp - 1 : 0) * kR;
  Tagged* bdst = s_bnd + (size_t)warp * kR;

  float ab = B200_NEG, al = B200_NEG;

  // one frame of the recursion; bn = label state of the last lane of the previous warp at the previous frame
  // (the label state is the critical chain shfl -> sub -> ex2 -> add -> lg2 -> add -> add -> shfl: the skip
  // mask is folded into the own-state side, hi - skipmask, which is known before the shuffle returns, and
  // the label state's ex2 / lg2 are issued before the blank state's)
  auto frame = [&](float lb_t, float ll_t, float bn, float o, int fr) {
    const float sh = __shfl_up_sync(0xffffffffu, al, 1);
    const float hi = fmaxf(al, ab);
    const float s1 = 1.0f + ex2f(-fabsf(al - ab));
    const float hi2 = hi - skipmask;  // +1e30 when the skip transition is not allowed: its term vanishes
    const float lbo = lb_t - o, llo = ll_t - o;
    // lane 0 takes the boundary value: two predicated subtractions instead of a select in front of one
    const float dl = sub_pick(lane0, bn, sh, hi2);  // neighbour - hi2
    const float db = sub_pick(lane0, bn, sh, ab);   // neighbour - ab
    const float t = ex2f(-fabsf(dl));
    const float u = ex2f(-fabsf(db));
    const float nbr = lane0 ? bn : sh;
    const float sum = (dl <= 0.f) ? (s1 + t) : fmaf(s1, t, 1.0f);
    const float mxl = llo + (skip_ok ? fmaxf(hi, nbr) : hi);
    const float mxb = lbo + fmaxf(ab, nbr);
    al = mxl + lg2f(sum);
    ab = mxb + lg2f(1.0f + u);
    if (lane31) sts_tagged(&bdst[(fr + 1) & (kR - 1)], al, fr);
    if (active) *reinterpret_cast<float2*>(outp) = make_float2(ab, al);
    outp += ostep;
  };
  // every kP-th chunk each warp publishes the maximum over its states at the start of the chunk
  // (the reduction is issued at the start of the chunk, its result stored after the first frame: a warp issues
  // in order, and the store would otherwise wait for the reduction in front of the frames)
  auto level_of_warp = [&](int c) -> float { return (c % kP) ? 0.f : warp_max_redux(fmaxf(ab, al)); };
  auto publish_level = [&](int c, float wm) {
    if (c % kP == 0 && lane31) sts_tagged(&s_wmx[(c & (kWR - 1)) * nw + warp], wm, c);
  };
  auto ctl_dstep = [&](int c) { return lds_tagged(&s_ctl[(c & (kCtlRing - 1)) * 2]); };
  auto ctl_corr = [&](int c) { return lds_tagged(&s_ctl[(c & (kCtlRing - 1)) * 2 + 1]); };
  // generic (slow) chunk: the first chunk and a trailing partial chunk; reads lpg directly
  auto slow_chunk = [&](int c, int n) {
    const int f0 = c * CHF;
    float corr = 0.f, dstep = 0.f;
    if (c == 0) {
      if (Tb > 1) dstep = clamp_int_f(floorf(ld_relaxed_f32(lp0 + lstep)));
    } else {
      Tagged e;
      do { e = ctl_dstep(c); } while (e.tag != c);
      dstep = e.v;
      do { e = ctl_corr(c); } while (e.tag != c);
      corr = e.v;
    }
    // all loads of the chunk first (one memory latency, not one per frame)
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
          // frame 0: only the first blank and the first label are reachable
          if (i == 0) { ab = lb_t; al = ll_t; }
          if (lane31) sts_tagged(&bdst[1], al, 0);
          if (active) *reinterpret_cast<float2*>(outp) = make_float2(ab, al);
          outp += ostep;
        } else {
          Tagged e;
          e.v = B200_NEG; e.tag = fr - 1;
          if (poll) {
            do { e = lds_tagged(&bsrc[fr & (kR - 1)]); } while (e.tag != fr - 1);
          }
          frame(lb_t, ll_t, e.v, dstep + (f == 0 ? corr : 0.f), fr);
        }
      }
    }
  };
  // columns of the staged rows of full chunk c (they have landed: the control record of chunk c - 1 is valid)
  auto load_chunk = [&](int c, float(&xb)[CHF], float(&xl)[CHF]) {
    const uint32_t slot = (uint32_t)(c - 1) % kLpSlots;
    const float4 b0 = lds_v4(a_blk0 + slot * (CHF * 4)), b1 = lds_v4(a_blk0 + slot * (CHF * 4) + 16);
    uint32_t a = a_lab0 + slot * a_sstep;
#pragma unroll
    for (int f = 0; f < CHF; ++f) {
      xl[f] = lds_f32(a);
      a += a_fstep;
    }
    xb[0] = b0.x; xb[1] = b0.y; xb[2] = b0.z; xb[3] = b0.w; xb[4] = b1.x; xb[5] = b1.y; xb[6] = b1.z; xb[7] = b1.w;
  };
  // full chunk c >= 1: straight-line code, no branch inside the frames.  xb / xl: this chunk's log-probs;
  // nb / nl receive those of chunk c + 1; cd / cc: control records of chunk c as read one chunk ago
  auto fast_chunk = [&](int c, float(&xb)[CHF], float(&xl)[CHF], float(&nb)[CHF], float(&nl)[CHF], Tagged& cd, Tagged& cc) {
    const int f0 = c * CHF;
    SWEEP_STAMP(0);
    if (cd.tag != c) {  // (the helper writes corr before dstep, dstep is read before corr)
      do { cd = ctl_dstep(c); } while (cd.tag != c);
      cc = ctl_corr(c);
    }
    const float dstep = cd.v, corr = cc.v;
    // boundary values of frames f0-1 .. f0+CHF-2 from the previous warp, read before the other loads of this
    // chunk so that their latency is hidden when the records are already there.  One lane publishes them in
    // frame order and the LAST one is read first: if it is there, the reads after it see all of them.
    // (the record of frame fr lives at ring index fr + 1: the CHF records of a chunk are contiguous)
    const Tagged* brow = &bsrc[f0 & (kR - 1)];
    Tagged br[CHF];
    if (poll) {
      br[CHF - 1] = lds_tagged(brow + CHF - 1);
#pragma unroll
      for (int f = 0; f < CHF - 1; ++f) br[f] = lds_tagged(brow + f);
    }
    if (c + 1 < nfull) load_chunk(c + 1, nb, nl);
    cd = ctl_dstep(c + 1);
    cc = ctl_corr(c + 1);
    SWEEP_STAMP(1);
    const float wm = level_of_warp(c);
    SWEEP_STAMP(2);
    float bn[CHF];
#pragma unroll
    for (int f = 0; f < CHF; ++f) bn[f] = B200_NEG;
    if (poll) {
      if (br[CHF - 1].tag != f0 + CHF - 2) {
        do { br[CHF - 1] = lds_tagged(brow + CHF - 1); } while (br[CHF - 1].tag != f0 + CHF - 2);
#pragma unroll
        for (int f = 0; f < CHF - 1; ++f) br[f] = lds_tagged(brow + f);
      }
#pragma unroll
      for (int f = 0; f < CHF; ++f) bn[f] = br[f].v;
    }
    SWEEP_STAMP(3);
    frame(xb[0], xl[0], bn[0], dstep + corr, f0);
    publish_level(c, wm);
#pragma unroll
    for (int f = 1; f < CHF; ++f) frame(xb[f], xl[f], bn[f], dstep, f0 + f);
  };

  slow_chunk(0, min(CHF, Tb));
  if (nfull > 1) {
    float rb0[CHF], rl0[CHF], rb1[CHF], rl1[CHF];
    Tagged cd, cc;
    do { cd = ctl_dstep(0); } while (cd.tag != 0);  // the rows of chunk 1 have landed
    load_chunk(1, rb0, rl0);
    cd = ctl_dstep(1);
    cc = ctl_corr(1);
    int c = 1;
    while (c < nfull) {
      fast_chunk(c, rb0, rl0, rb1, rl1, cd, cc);
      if (++c >= nfull) break;
      fast_chunk(c, rb1, rl1, rb0, rl0, cd, cc);
      ++c;
    }
  }
  if (nfull >= 1 && Tb > nfull * CHF) slow_chunk(nfull, Tb - nfull * CHF);

  // log2 p(y|x) = lse2(state 2L, state 2L-1) + sum of offsets
  if (i == L) s_fin[0] = ab;
  if (L == 0 && i == 0) s_fin[1] = B200_NEG;
  if (L > 0 && i == L - 1) s_fin[1] = al;
  asm volatile("bar.sync 1, %0;" ::"r"((nwa + 1) * 32) : "memory");
  if (i == 0) {
    const float z = lse2(s_fin[0], s_fin[1]);
    const bool feasible = z > -1.0e29f;
    const float logz2 = z + (float)s_csum;
    if (d == 0) sweep_finish_nll(p, b, feasible ? -B200_LN2 * logz2 : __int_as_float(0x7f800000));
    else p.ws.logz_beta[b] = feasible ? logz2 : B200_NEG;
    // this direction's lattice of utterance b is complete (every state warp stored its last frame before the
    // barrier above): the occupancy kernel may read it once both directions have signalled
    __threadfence();
    atomicAdd(p.ws.done + b, 1);
    TL_MAX(3);
    TL_SET(64 + 2 * b + d);
  }
}

template <int MAXT>
static size_t sweep_smem_bytes(int nw) {
  const size_t kR = SweepRings<MAXT>::kR, kWR = SweepRings<MAXT>::kWR;
  return sizeof(Tagged) * nw * kR + sizeof(Tagged) * kWR * nw + sizeof(int) * kWR;
}

template <int CHF, int MAXT>
static int launch_sweep_t(const LossParams& p, int threads, cudaStream_t st, bool pdl) {
  size_t smem = sweep_smem_bytes<MAXT>(threads / 32);
  // A sweep CTA is one latency chain: give it an SM of its own whenever the grid fits that way.
  // (Staged under the previous kernel by PDL, CTAs would otherwise pile up on the SMs that free
  // up first.)  Asking for more than half of the shared memory caps residency at one CTA per SM.
  if (2 * (p.b1 - p.b0) <= num_sms() && smem < 116 * 1024) smem = 116 * 1024;
  auto kern = ctc_sweep_kernel<CHF, MAXT>;
  if (smem > 48 * 1024) B200_CUDA(set_max_smem_cached((const void*)kern, (int)smem));
  B200_CUDA(launch_kernel(kern, dim3((unsigned)(p.b1 - p.b0), 2), dim3(threads), smem, st, pdl, p));
  return B200_OK;
}

constexpr size_t kSweepPadBytes = 114 * 1024;

static int launch_sweep_staged(const LossParams& p, int threads, cudaStream_t st, bool pdl) {
  size_t smem = sweep_smem_bytes<256>(threads / 32) + sizeof(Tagged) * 2 * kCtlRing +
                (size_t)kLpSlots * 8 * (p.ws.Kp + 1) * sizeof(float) + kLpSlots * sizeof(uint64_t);
  threads += 32;  // the helper warp
  // One latency chain per SM whenever the grid fits that way (see launch_sweep_t): 114 KB is more than half of an SM's
  // shared memory (two sweep CTAs cannot share an SM) and leaves room for ONE CTA of the patch kernel (111.9 KB + the
  // 1 KB the system reserves per CTA: 233344 of 233472 bytes) next to a running sweep CTA, so the patch kernel works
  // through the utterances that have finished on ALL SMs, not only on the 20 without a sweep CTA.  (Measured at C2:
  // 189.2 against 190.7 us per step with a 116 KB pad -- the patch warps that share an SM also take issue slots from
  // the sweep's latency chain, so most of what they gain the sweep loses.)
  if (2 * (p.b1 - p.b0) <= num_sms() && smem < kSweepPadBytes) smem = kSweepPadBytes;
  auto kern = ctc_sweep_staged_kernel;
  if (smem > 48 * 1024) B200_CUDA(set_max_smem_cached((const void*)kern, (int)smem));
  B200_CUDA(launch_kernel(kern, dim3((unsigned)(p.b1 - p.b0), 2), dim3(threads), smem, st, pdl, p));
  return B200_OK;
}

#ifdef B200_SWEEP_CLOCKS
}  // namespace b200
extern "C" int b200_debug_sweep_clocks(long long* host) {
  return (int)cudaMemcpyFromSymbol(host, b200::g_sweep_clk, sizeof(b200::g_sweep_clk));
}
namespace b200 {
#endif

}  // namespace b200
#include "ctc_loss_sweep_cluster.inl"
namespace b200 {
static int g_single_cta_sweep = 0;  // developer switch: long label sequences in one CTA (b200_ctc_debug_single_cta_sweep)

// pdl: staged under the stream kernel of the same call
static int launch_sweep(const LossParams& p, cudaStream_t st, bool pdl = false) {
  const int threads = (int)align_up((size_t)(p.Lmax + 1), 32);
  if (threads <= 256) return launch_sweep_staged(p, threads, st, pdl);
  if (!g_single_cta_sweep) return launch_sweep_cluster(p, st, pdl);
  return launch_sweep_t<4, 1024>(p, threads, st, pdl);
}

}  // namespace b200
