// ctc_loss_lin.inl -- the alpha / beta lattice recursion in the LINEAR domain, in float64 (included by ctc_loss.cu).
//
// The log-domain sweep (ctc_loss_sweep.inl) pays shfl -> ex2 -> lg2 -> adds per frame: 85..100 cycles of dependent
// latency (tools/latbench.cu).  In the linear domain a frame is
//     al' = pl * (al + ab + skip * nbr),   ab' = pb * (ab + nbr)
// and with q = pl * (al + ab) computed while the shuffle is in flight the chain is shfl -> DFMA: 38..40 cycles
// (B200: DFMA 8.4 cycles, two 32-bit shuffles 30).  What the log domain gave for free is range; here it comes from
//   * float64 (2^-1074 .. 2^1023),
//   * a FEED-FORWARD scale per frame, applied by the stream kernel: the probabilities arrive as
//     pd[t][k] = p_t(k) * 2^gff(t) with the largest needed probability of the frame in [1, 2),
//   * a FEEDBACK scale per warp ("block") and chunk of 8 frames: at the start of a chunk a warp brings the largest
//     of its own states and of the boundary values that will flow in from the previous warp during the chunk to
//     ~1 by an exact power of two.  Warps only ever talk to their predecessor, so the scale travels with the
//     boundary record and there is no global controller:  true value = stored * 2^-(gcum(t) + gwarp(chunk, warp)).
// Nothing is converted here: the lattice keeps the HIGH WORD of every state (sign, 11-bit exponent, 20 mantissa
// bits: 2^-21 relative, the resolution the fp32 log2 lattice had), and the occupancy kernel works from exponents
// and mantissas directly.  States below 2^-1074 of their block's scale flush to zero; the occupancy kernel bounds
// what they could have contributed (a certificate per row) and hands utterances it cannot certify to the
// log-domain kernels (LossWs::redo), so the result never depends on the range being enough.
namespace b200 {

constexpr int kLinCHF = 8;      // frames per chunk
constexpr int kLinSlots = 10;   // chunks of pd rows in shared memory
constexpr int kLinAhead = 3;    // bulk copies run this many chunks ahead of warp 0
constexpr int kLinR = 128;      // boundary ring, frames

// label state of lane 31 of a warp after frame `tag`, in that warp's block scale g
struct __align__(16) BRec {
  int lo, hi, tag, g;
};
__device__ __forceinline__ BRec lds_brec(const BRec* p) {
  BRec r;
  asm volatile("ld.volatile.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r.lo), "=r"(r.hi), "=r"(r.tag), "=r"(r.g) : "r"(smem_u32(p)));
  return r;
}
__device__ __forceinline__ void sts_brec(BRec* p, double v, int tag, int g) {
  asm volatile("st.volatile.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(smem_u32(p)), "r"(__double2loint(v)), "r"(__double2hiint(v)),
               "r"(tag), "r"(g)
               : "memory");
}
__device__ __forceinline__ int lds_volatile_s32(const int* p) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)));
  return v;
}
__device__ __forceinline__ void sts_volatile_s32(int* p, int v) {
  asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ double lds_f64(uint32_t a) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
  return v;
}
// exponent field of a non-negative double given its high word (0 for zero and subnormals)
__device__ __forceinline__ int expo_field(int hi) { return (hi >> 20) & 0x7ff; }
// v * 2^d for any integer d, by exponent arithmetic: zero / subnormal in, or a result below the normal range, gives 0
__device__ __forceinline__ double scale_pow2(int lo, int hi, int d) {
  const int e = expo_field(hi);
  int ne = e + d;
  if (e == 0 || ne <= 0) return 0.0;
  if (ne > 2046) ne = 2046;
  return __hiloint2double((hi & 0x800fffff) | (ne << 20), lo);
}
__device__ __forceinline__ double pow2_double(int g) {  // |g| <= 1022
  return __hiloint2double((1023 + g) << 20, 0);
}

struct LinSmem {
  BRec* bnd;       // [nw][kLinR]
  int* ready;      // [32] chunk c staged (tag)
  int* prog;       // [8]  last chunk whose rows state warp w has consumed
  double* pd;      // [kLinSlots][CHF][Kp]
  uint64_t* bar;   // [kLinSlots]
};

__device__ __forceinline__ LinSmem carve_lin_smem(unsigned char* base, int nw, int Kp) {
  LinSmem s;
  s.bnd = reinterpret_cast<BRec*>(base);
  s.ready = reinterpret_cast<int*>(s.bnd + (size_t)nw * kLinR);
  s.prog = s.ready + 32;
  s.pd = reinterpret_cast<double*>(s.prog + 8 + 8);  // (keeps 16-byte alignment)
  s.bar = reinterpret_cast<uint64_t*>(s.pd + (size_t)kLinSlots * kLinCHF * Kp);
  return s;
}
static size_t lin_smem_bytes(int nw, int64_t Kp) {
  return sizeof(BRec) * (size_t)nw * kLinR + sizeof(int) * 48 + sizeof(double) * (size_t)kLinSlots * kLinCHF * (size_t)Kp +
         sizeof(uint64_t) * kLinSlots;
}

__global__ void __launch_bounds__(288) ctc_sweep_lin_kernel(LossParams p) {
  constexpr int CHF = kLinCHF;
  extern __shared__ __align__(16) unsigned char lin_smem[];
  const int d = blockIdx.y;
  const int i = threadIdx.x;
  const int lane = i & 31, warp = i >> 5;
  const int nw = (blockDim.x >> 5) - 1;  // state warps launched; warp nw is the helper
  const int Kp = (int)p.ws.Kp;
  const int64_t Sp = p.ws.Sp;
  const LinSmem sm = carve_lin_smem(lin_smem, nw, Kp);
  __shared__ double s_finv[2];
  __shared__ int s_fing[2];
  __shared__ int s_gtot;
  __shared__ double s_zero;

  // shared-memory set-up overlaps the tail of the previous kernels; after the wait the rows of the stream kernel
  // (pd, gff) and the results of the prep kernel (utterance order, labels) may be used
  for (int k = i; k < nw * kLinR; k += blockDim.x) sm.bnd[k].tag = -1;
  if (i < 32) sm.ready[i] = -1;
  if (i < 8) sm.prog[i] = -1;
  if (i == 0) {
    s_zero = 0.0;
    for (int k = 0; k < kLinSlots; ++k) mbar_init(&sm.bar[k], 1);
    mbar_fence_init();
    fence_proxy_async();
  }
  pdl_wait();
  pdl_launch_dependents();
  __syncthreads();
  const int b = p.ws.order[p.b0 + blockIdx.x];
  const int Tb = min((int)p.ilen[b], (int)p.T);
  const int L = utt_labels(p, b);
  const int nwa = L / 32 + 1;  // warps that own at least one state (threads 0..L)
  if (Tb <= 0) {
    if (i == 0) {
      if (d == 0) sweep_finish_nll(p, b, (L == 0) ? 0.0f : __int_as_float(0x7f800000));
      else p.ws.logz_beta[b] = (L == 0) ? 0.0f : B200_NEG;
      __threadfence();
      atomicAdd(p.ws.done + b, 1);
    }
    return;
  }
  const int nchunk = (Tb + CHF - 1) / CHF;
  if (warp >= nwa && warp != nw) return;
  const int chunk_floats = CHF * Kp;  // doubles per slot
  // frames of chunk c, and the slot row that holds logical frame 8c + j (beta: the rows arrive in physical order)
  auto frames_in = [&](int c) { return min(CHF, Tb - c * CHF); };

  if (warp == nw) {
    // ------------------------------------------------------------------ helper warp: staging + gcum
    const double* pd_b = p.ws.pd + (int64_t)b * p.T * Kp;
    const int* gff_b = p.ws.gff + (int64_t)b * p.T;
    int* gcum_b = p.ws.gcum + ((int64_t)d * p.B + b) * p.T;
    auto stage = [&](int c) {
      const int n = frames_in(c);
      const int slot = c % kLinSlots;
      const int prow = d ? (Tb - c * CHF - n) : c * CHF;  // first physical row of the chunk
      const uint32_t bytes = (uint32_t)(n * Kp * 8);
      mbar_arrive_expect_tx(&sm.bar[slot], bytes);
      bulk_g2s(sm.pd + (size_t)slot * chunk_floats, pd_b + (int64_t)prow * Kp, bytes, &sm.bar[slot]);
    };
    if (lane == 0)
      for (int c = 0; c < kLinAhead && c < nchunk; ++c) stage(c);
    int run = 0;  // cumulative feed-forward exponent before this chunk
    for (int c = 0; c < nchunk; ++c) {
      const int n = frames_in(c);
      mbar_wait(&sm.bar[c % kLinSlots], (uint32_t)((c / kLinSlots) & 1));
      if (lane == 0) sts_volatile_s32(&sm.ready[c & 31], c);
      // cumulative scale up to and including each frame of the chunk, in sweep order
      {
        const int f = c * CHF + lane;          // logical frame of this lane
        const int t = d ? (Tb - 1 - f) : f;    // physical frame
        int g = (lane < n) ? gff_b[t] : 0;
#pragma unroll
        for (int o = 1; o < CHF; o <<= 1) {
          const int u = __shfl_up_sync(0xffffffffu, g, o);
          if (lane >= o) g += u;
        }
        if (lane < n) gcum_b[t] = run + g;
        run += __shfl_sync(0xffffffffu, g, n - 1);
      }
      // the slot of chunk c + kLinAhead is free once every state warp has consumed chunk c + kLinAhead - kLinSlots
      const int nx = c + kLinAhead;
      if (nx < nchunk) {
        const int need = nx - kLinSlots;
        if (need >= 0) {
          for (;;) {
            const int pr = (lane < nwa) ? lds_volatile_s32(&sm.prog[lane]) : 0x7fffffff;
            if (__all_sync(0xffffffffu, pr >= need)) break;
            __nanosleep(64);
          }
        }
        if (lane == 0) stage(nx);
      }
    }
    if (lane == 0) {
      s_gtot = run;
      if (d == 0) p.ws.gtot[b] = run;
    }
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
  const int kcol = has_label ? 1 + pi : 0;
  const bool lane0 = lane == 0, lane31 = lane == 31;
  const bool poll = warp > 0;
  const BRec* bsrc = sm.bnd + (size_t)(poll ? warp - 1 : 0) * kLinR;
  BRec* bdst = sm.bnd + (size_t)warp * kLinR;
  int* outp = reinterpret_cast<int*>(p.ws.lat) + (((int64_t)d * p.B + b) * p.T + (d ? (Tb - 1) : 0)) * Sp + 2 * i;
  const int64_t ostep = d ? -Sp : Sp;
  int* gw_out = p.ws.gwarp + (((int64_t)d * p.B + b) * ((p.T + 7) / 8)) * 8 + warp;
  const uint32_t a_pd = smem_u32(sm.pd);
  const uint32_t a_zero = smem_u32(&s_zero);

  double ab = 0.0, al = 0.0;
  int Gw = 0;

  // probabilities of chunk c for this thread: blank column and own label column, frames in logical order
  auto load_chunk = [&](int c, double(&xb)[CHF], double(&xl)[CHF]) {
    while (lds_volatile_s32(&sm.ready[c & 31]) != c) { }
    const int n = frames_in(c);
    const uint32_t base = a_pd + (uint32_t)((c % kLinSlots) * chunk_floats * 8);
#pragma unroll
    for (int f = 0; f < CHF; ++f) {
      const int j = min(f, n - 1);
      const uint32_t row = base + (uint32_t)((d ? (n - 1 - j) : j) * Kp * 8);
      xb[f] = lds_f64(row);
      xl[f] = lds_f64(has_label ? row + (uint32_t)(kcol * 8) : a_zero);
    }
  };

  auto run_chunk = [&](int c, const double(&xb)[CHF], const double(&xl)[CHF]) {
    const int n = frames_in(c);
    const int f0 = c * CHF;
    const int j0 = (c == 0) ? 1 : 0;
    if (lane0) sts_volatile_s32(&sm.prog[warp], c);  // (the rows of chunk c are in registers)
    // boundary records of frames f0 + j - 1 for j in [j0, n): the last one is published last
    BRec br[CHF];
#pragma unroll
    for (int f = 0; f < CHF; ++f) { br[f].lo = 0; br[f].hi = 0; br[f].tag = 0; br[f].g = 0; }
    if (poll && n > j0) {
      const int last = f0 + n - 2;
      BRec r;
      do { r = lds_brec(&bsrc[last & (kLinR - 1)]); } while (r.tag != last);
#pragma unroll
      for (int f = 0; f < CHF; ++f) {
        const int fr = f0 + f - 1;
        if (f >= j0 && f < n) br[f] = lds_brec(&bsrc[fr & (kLinR - 1)]);
      }
    }
    // ---- feedback scale of this block for the chunk: own states and what will flow in, to ~1
    if (c > 0) {
      const int hown = max(__double2hiint(ab), __double2hiint(al));  // non-negative doubles order like their high words
      int lev = expo_field(__reduce_max_sync(0xffffffffu, hown));
      if (poll) {
#pragma unroll
        for (int f = 0; f < CHF; ++f) {
          const int e = expo_field(br[f].hi);
          if (f < n && e != 0) lev = max(lev, e + (Gw - br[f].g));
        }
      }
      int g = 0;
      if (lev != 0) g = max(min(1023 - lev, 1000), -1000);
      if (g != 0) {
        const double sc = pow2_double(g);
        ab *= sc;
        al *= sc;
        Gw += g;
      }
    }
    if (lane0) gw_out[(int64_t)c * 8] = Gw;
    double bn[CHF];
#pragma unroll
    for (int f = 0; f < CHF; ++f) bn[f] = poll ? scale_pow2(br[f].lo, br[f].hi, Gw - br[f].g) : 0.0;

    if (c == 0) {
      // frame 0: only the first blank and the first label are reachable
      if (i == 0) { ab = xb[0]; al = xl[0]; }
      if (lane31) sts_brec(&bdst[0], al, 0, Gw);
      if (active) *reinterpret_cast<int2*>(outp) = make_int2(__double2hiint(ab), __double2hiint(al));
      outp += ostep;
    }
#pragma unroll
    for (int f = 0; f < CHF; ++f) {
      if (f >= j0 && f < n) {
        const int fr = f0 + f;
        const int slo = __shfl_up_sync(0xffffffffu, __double2loint(al), 1);
        const int shi = __shfl_up_sync(0xffffffffu, __double2hiint(al), 1);
        const double q = xl[f] * (al + ab);           // (in flight while the shuffle returns)
        const double plk = skip_ok ? xl[f] : 0.0;
        const double nbr = lane0 ? bn[f] : __hiloint2double(shi, slo);
        const double nal = fma(plk, nbr, q);
        const double nab = xb[f] * (ab + nbr);
        al = nal;
        ab = nab;
        if (lane31) sts_brec(&bdst[fr & (kLinR - 1)], al, fr, Gw);
        if (active) *reinterpret_cast<int2*>(outp) = make_int2(__double2hiint(ab), __double2hiint(al));
        outp += ostep;
      }
    }
  };

  {
    double xb0[CHF], xl0[CHF], xb1[CHF], xl1[CHF];
    load_chunk(0, xb0, xl0);
    int c = 0;
    while (c < nchunk) {
      if (c + 1 < nchunk) load_chunk(c + 1, xb1, xl1);
      run_chunk(c, xb0, xl0);
      if (++c >= nchunk) break;
      if (c + 1 < nchunk) load_chunk(c + 1, xb0, xl0);
      run_chunk(c, xb1, xl1);
      ++c;
    }
  }

  // p(y|x) = state 2L + state 2L-1 (each in its own block scale), times 2^-(sum of the feed-forward exponents)
  if (i == L) { s_finv[0] = ab; s_fing[0] = Gw; }
  if (L == 0 && i == 0) { s_finv[1] = 0.0; s_fing[1] = 0; }
  if (L > 0 && i == L - 1) { s_finv[1] = al; s_fing[1] = Gw; }
  asm volatile("bar.sync 1, %0;" ::"r"((nwa + 1) * 32) : "memory");
  if (i == 0) {
    const double va = s_finv[0], vb = s_finv[1];
    const double la = va > 0.0 ? log2(va) - (double)s_fing[0] : -1.0e300;
    const double lb = vb > 0.0 ? log2(vb) - (double)s_fing[1] : -1.0e300;
    const double m = fmax(la, lb);
    const bool feasible = m > -1.0e299;
    const double logz2 = feasible ? m + log2(exp2(la - m) + exp2(lb - m)) - (double)s_gtot : 0.0;
    if (d == 0) {
      // a zero here is either an utterance without any alignment or a sweep that ran out of range: the log-domain
      // kernels decide (they cost nothing when the flag stays clear)
      if (!feasible) p.ws.redo[b] = 1;
      sweep_finish_nll(p, b, feasible ? (float)(-0.6931471805599453 * logz2) : __int_as_float(0x7f800000));
    } else {
      p.ws.logz_beta[b] = feasible ? (float)logz2 : B200_NEG;
    }
    __threadfence();
    atomicAdd(p.ws.done + b, 1);
  }
}

static int launch_sweep_lin(const LossParams& p, cudaStream_t st, bool pdl) {
  const int threads = (int)align_up((size_t)(p.Lmax + 1), 32) + 32;  // + the helper warp
  size_t smem = lin_smem_bytes(threads / 32 - 1, p.ws.Kp);
  if (2 * (p.b1 - p.b0) <= num_sms() && smem < 116 * 1024) smem = 116 * 1024;  // one latency chain per SM
  auto kern = ctc_sweep_lin_kernel;
  if (smem > 48 * 1024) B200_CUDA(set_max_smem_cached((const void*)kern, (int)smem));
  B200_CUDA(launch_kernel(kern, dim3((unsigned)(p.b1 - p.b0), 2), dim3(threads), smem, st, pdl, p));
  return B200_OK;
}

// the linear sweep serves label sequences of up to 255 (8 state warps) whose staged rows fit shared memory
static bool lin_sweep_ok(const LossParams& p) {
  return p.Lmax + 1 <= 256 && lin_smem_bytes((int)(align_up((size_t)(p.Lmax + 1), 32) / 32), p.ws.Kp) <= (size_t)max_smem_optin() - 2048;
}

}  // namespace b200
