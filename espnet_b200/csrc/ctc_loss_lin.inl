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

constexpr int kLinCHF = 8;        // frames per chunk
constexpr int kLinMaxSlots = 32;  // chunks of pd rows in shared memory (as many as fit, see lin_slots)
constexpr int kLinReady = 64;     // ring of "chunk staged" tags
// The boundary ring (frames) must hold everything a warp can publish before its successor has read it.  A warp can
// lead its successor by at most the row ring (the helper does not stage chunk s before EVERY warp has consumed chunk
// s - nslots), so the ring is sized from nslots: 8 * (nslots + 2) frames rounded up to a power of two (lin_plan).
// (With a fixed 128-frame ring and 20 row slots warp 0 -- which polls nobody and gains ~450 cycles per chunk on the
// others -- overwrote records its successor had not read yet, and the successor waited forever.)

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
  const int ne = min(e + d, 2046);
  const bool zero = (e == 0) | (ne <= 0);  // (selects, no branches: this sits in front of every chunk)
  return __hiloint2double(zero ? 0 : ((hi & 0x800fffff) | (ne << 20)), zero ? 0 : lo);
}
__device__ __forceinline__ double pow2_double(int g) {  // |g| <= 1022
  return __hiloint2double((1023 + g) << 20, 0);
}

struct LinSmem {
  BRec* bnd;       // [nw][ring]
  int* ready;      // [kLinReady] chunk c staged (tag)
  int* prog;       // [8]  last chunk whose rows state warp w has consumed
  uint64_t* bar;   // [kLinMaxSlots]
  double* pd;      // [nslots][CHF][Kp]
};

__device__ __forceinline__ LinSmem carve_lin_smem(unsigned char* base, int nw, int ring) {
  LinSmem s;
  s.bnd = reinterpret_cast<BRec*>(base);
  s.ready = reinterpret_cast<int*>(s.bnd + (size_t)nw * ring);
  s.prog = s.ready + kLinReady;
  s.bar = reinterpret_cast<uint64_t*>(s.prog + 8);
  s.pd = reinterpret_cast<double*>(s.bar + kLinMaxSlots);
  return s;
}
static size_t lin_smem_fixed(int nw, int ring) {
  return sizeof(BRec) * (size_t)nw * ring + sizeof(int) * (kLinReady + 8) + sizeof(uint64_t) * kLinMaxSlots;
}
struct LinPlan {
  int nslots, ring;
  size_t smem;
};
// As many chunk slots as fit: the rows come from DRAM (~2 us) and a chunk lasts 8 frames x ~20 ns, so the copies must
// run a dozen chunks ahead of the first warp; the most skewed warp is nw chunks behind it.  The boundary ring follows
// from the slot count (see above): 256 frames serve up to 30 slots, 128 frames up to 14.
static LinPlan lin_plan(int nw, int64_t Kp) {
  const size_t slot = sizeof(double) * kLinCHF * (size_t)Kp;
  LinPlan best{0, 128, 0};
  const int rings[2] = {256, 128};
  for (int ring : rings) {
    const size_t fixed = lin_smem_fixed(nw, ring);
    const size_t total = (size_t)max_smem_optin() - 4096;
    if (fixed + slot > total) continue;
    int64_t n = (int64_t)((total - fixed) / slot);
    const int64_t cap = ring / kLinCHF - 2;
    if (n > cap) n = cap;
    if (n > kLinMaxSlots) n = kLinMaxSlots;
    if (n > best.nslots) best = LinPlan{(int)n, ring, fixed + slot * (size_t)n};
  }
  return best;
}

__global__ void __launch_bounds__(288) ctc_sweep_lin_kernel(LossParams p, int nslots, int ring) {
  constexpr int CHF = kLinCHF;
  extern __shared__ __align__(16) unsigned char lin_smem[];
  const int d = blockIdx.y;
  const int i = threadIdx.x;
  const int lane = i & 31, warp = i >> 5;
  const int nw = (blockDim.x >> 5) - 1;  // state warps launched; warp nw is the helper
  const int Kp = (int)p.ws.Kp;
  const int64_t Sp = p.ws.Sp;
  const LinSmem sm = carve_lin_smem(lin_smem, nw, ring);
  const int rmask = ring - 1;
  __shared__ double s_finv[2];
  __shared__ int s_fing[2];
  __shared__ int s_gtot;
  __shared__ double s_zero;
  __shared__ double s_bn[8][kLinCHF];  // boundary values of the chunk in the block's scale, per state warp

  // shared-memory set-up overlaps the tail of the previous kernels; after the wait the rows of the stream kernel
  // (pd, gff) and the results of the prep kernel (utterance order, labels) may be used
  for (int k = i; k < nw * ring; k += blockDim.x) sm.bnd[k].tag = -1;
  if (i < kLinReady) sm.ready[i] = -1;
  if (i < 8) sm.prog[i] = -1;
  if (i == 0) {
    s_zero = 0.0;
    for (int k = 0; k < nslots; ++k) mbar_init(&sm.bar[k], 1);
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
  auto frames_in = [&](int c) { return min(CHF, Tb - c * CHF); };

  if (warp == nw) {
    // ------------------------------------------------------------------ helper warp: staging
    const double* pd_b = p.ws.pd + (int64_t)b * p.T * Kp;
    // (chunks are staged and awaited in order: the slot indices are counted, not computed -- an integer division per
    // chunk is ~40 dependent instructions)
    int st_slot = 0, next_stage = 0;
    auto stage = [&](int c) {  // lane 0
      const int n = frames_in(c);
      const int prow = d ? (Tb - c * CHF - n) : c * CHF;  // first physical row of the chunk
      const uint32_t bytes = (uint32_t)(n * Kp * 8);
      mbar_arrive_expect_tx(&sm.bar[st_slot], bytes);
      bulk_g2s(sm.pd + (size_t)st_slot * chunk_floats, pd_b + (int64_t)prow * Kp, bytes, &sm.bar[st_slot]);
    };
    // stage every chunk whose slot is free: chunk s replaces chunk s - nslots, which every state warp must have
    // consumed.  Never sleeps (__nanosleep costs about a microsecond, three chunks of the sweep): the copies run as far
    // ahead as the ring allows, so this only spins when the ring is full, and then there is nothing else to do.
    auto stage_free = [&](bool must) {
      for (;;) {
        if (next_stage >= nchunk) return;
        const int need = next_stage - nslots;
        if (need >= 0) {
          const int pr = (lane < nwa) ? lds_volatile_s32(&sm.prog[lane]) : 0x7fffffff;
          if (!__all_sync(0xffffffffu, pr >= need)) {
            if (must) continue;
            return;
          }
        }
        if (lane == 0) stage(next_stage);
        st_slot = (st_slot + 1 == nslots) ? 0 : st_slot + 1;
        ++next_stage;
        must = false;
      }
    };
    stage_free(false);
    int w_slot = 0;
    uint32_t w_par = 0;
    for (int c = 0; c < nchunk; ++c) {
      if (next_stage <= c) stage_free(true);  // (cannot happen with nslots > nw + 2, kept for safety)
      mbar_wait(&sm.bar[w_slot], w_par);
      if (++w_slot == nslots) { w_slot = 0; w_par ^= 1u; }
      if (lane == 0) sts_volatile_s32(&sm.ready[c & (kLinReady - 1)], c);
      stage_free(false);
    }
    // sum of the feed-forward exponents of the utterance (enters log p(y|x) and the occupancy kernel's reference level)
    int gsum = 0;
    {
      const int* gff_b = p.ws.gff + (int64_t)b * p.T;
      for (int t = lane; t < Tb; t += 32) gsum += gff_b[t];
      gsum = __reduce_add_sync(0xffffffffu, gsum);
    }
    if (lane == 0) {
      s_gtot = gsum;
      if (d == 0) p.ws.gtot[b] = gsum;
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
  const BRec* bsrc = sm.bnd + (size_t)(poll ? warp - 1 : 0) * ring;
  BRec* bdst = sm.bnd + (size_t)warp * ring;
  int* outp = reinterpret_cast<int*>(p.ws.lat) + (((int64_t)d * p.B + b) * p.T + (d ? (Tb - 1) : 0)) * Sp + 2 * i;
  const int64_t ostep = d ? -Sp : Sp;
  int* gw_out = p.ws.gwarp + (((int64_t)d * p.B + b) * ((p.T + 7) / 8)) * 8 + warp;
  const uint32_t a_pd = smem_u32(sm.pd);
  const uint32_t a_zero = smem_u32(&s_zero);

  double ab = 0.0, al = 0.0;
  int Gw = 0;
  int ld_slot = 0;  // slot of the next chunk to load (chunks are loaded in order)
  const uint32_t row_bytes = (uint32_t)(Kp * 8), slot_bytes = (uint32_t)(chunk_floats * 8);
  const uint32_t row_step = d ? (uint32_t)(-(int)row_bytes) : row_bytes;
  const uint32_t col_off = (uint32_t)(kcol * 8);

  // probabilities of chunk c for this thread: blank column and own label column, frames in logical order
  auto load_chunk = [&](int c, double(&xb)[CHF], double(&xl)[CHF]) {
    while (lds_volatile_s32(&sm.ready[c & (kLinReady - 1)]) != c) { }
    const int n = frames_in(c);
    // logical frame j of the chunk sits in slot row j (alpha) or n - 1 - j (beta: the rows arrive in physical order)
    uint32_t row = a_pd + (uint32_t)ld_slot * slot_bytes + (d ? (uint32_t)(n - 1) * row_bytes : 0u);
    ld_slot = (ld_slot + 1 == nslots) ? 0 : ld_slot + 1;
#pragma unroll
    for (int f = 0; f < CHF; ++f) {
      xb[f] = lds_f64(row);
      xl[f] = lds_f64(has_label ? row + col_off : a_zero);
      if (f < n - 1) row += row_step;
    }
  };

  // FULL: a chunk of 8 frames that is not the first one -- straight-line code, no branch inside the frames
  auto run_chunk = [&](auto full_tag, int c, const double(&xb)[CHF], const double(&xl)[CHF]) {
    constexpr bool FULL = decltype(full_tag)::value;
    const int n = FULL ? CHF : frames_in(c);
    const int f0 = c * CHF;
    const int j0 = (!FULL && c == 0) ? 1 : 0;
    if (lane0) sts_volatile_s32(&sm.prog[warp], c);  // (the rows of chunk c are in registers)
    // Boundary records of frames f0 + j - 1 for j in [j0, n): LANE j reads record j (one load for the warp; the last
    // one is published last), contributes it to the level of the block, converts it to the block's new scale, and
    // leaves it in shared memory for lane 0.  Doing this per lane instead of eight times in every lane keeps the
    // front of a chunk short.
    BRec r;
    r.lo = 0; r.hi = 0; r.tag = 0; r.g = 0;
    const bool mine = poll && lane >= j0 && lane < n;  // this lane holds a record
    if (poll && n > j0) {
      const BRec* src = &bsrc[(f0 + lane - 1) & rmask];
      const int want = f0 + lane - 1;
      for (;;) {
        if (mine) r = lds_brec(src);
        if (__all_sync(0xffffffffu, !mine || r.tag == want)) break;
      }
    }
    // ---- feedback scale of this block for the chunk: own states and what will flow in, to ~1
    if (FULL || c > 0) {
      const int er = expo_field(r.hi);
      const int cand = (mine && er != 0) ? er + (Gw - r.g) : 0;
      const int hown = max(__double2hiint(ab), __double2hiint(al));  // non-negative doubles order like their high words
      const int lev = __reduce_max_sync(0xffffffffu, max(expo_field(hown), cand));
      // Upward the step is capped (the rest follows in the next chunk).  DOWNWARD it is not: when a value arrives
      // from the previous block that is thousands of binades above this block's own states (a block whose junk has
      // decayed for many chunks, rescaled to 1 every time, and which the alignment now enters), the block must take
      // the incoming scale at once -- its own states are then below 2^-1022 of it and flush to zero, legitimately.
      // (A cap of -1000 here saturated the incoming value instead: garbage that the certificate could not see.)
      const int g = (lev != 0) ? min(1023 - lev, 1000) : 0;
      const double sc = (g >= -1000) ? pow2_double(g) : 0.0;
      ab *= sc;
      al *= sc;
      Gw += g;
    }
    if (lane0) gw_out[(int64_t)c * 8] = Gw;
    double bn[CHF];
    if (poll) {
      if (lane < CHF) s_bn[warp][lane] = mine ? scale_pow2(r.lo, r.hi, Gw - r.g) : 0.0;
      __syncwarp();
#pragma unroll
      for (int f = 0; f < CHF; ++f) bn[f] = s_bn[warp][f];
      __syncwarp();
    } else {
#pragma unroll
      for (int f = 0; f < CHF; ++f) bn[f] = 0.0;
    }

    if (!FULL && c == 0) {
      // frame 0: only the first blank and the first label are reachable
      if (i == 0) { ab = xb[0]; al = xl[0]; }
      if (lane31) sts_brec(&bdst[0], al, 0, Gw);
      if (active) *reinterpret_cast<int2*>(outp) = make_int2(__double2hiint(ab), __double2hiint(al));
      outp += ostep;
    }
#pragma unroll
    for (int f = 0; f < CHF; ++f) {
      if (FULL || (f >= j0 && f < n)) {
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
        if (lane31) sts_brec(&bdst[fr & rmask], al, fr, Gw);
        if (active) *reinterpret_cast<int2*>(outp) = make_int2(__double2hiint(ab), __double2hiint(al));
        outp += ostep;
      }
    }
  };
  auto run = [&](int c, const double(&xb)[CHF], const double(&xl)[CHF]) {
    if (c > 0 && frames_in(c) == CHF) run_chunk(std::true_type{}, c, xb, xl);
    else run_chunk(std::false_type{}, c, xb, xl);
  };

  {
    double xb0[CHF], xl0[CHF], xb1[CHF], xl1[CHF];
    load_chunk(0, xb0, xl0);
    int c = 0;
    while (c < nchunk) {
      if (c + 1 < nchunk) load_chunk(c + 1, xb1, xl1);
      run(c, xb0, xl0);
      if (++c >= nchunk) break;
      if (c + 1 < nchunk) load_chunk(c + 1, xb0, xl0);
      run(c, xb1, xl1);
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
  const int nw = (int)(align_up((size_t)(p.Lmax + 1), 32) / 32);
  const int threads = (nw + 1) * 32;  // + the helper warp
  const LinPlan pl = lin_plan(nw, p.ws.Kp);
  auto kern = ctc_sweep_lin_kernel;
  B200_CUDA(set_max_smem_cached((const void*)kern, (int)pl.smem));
  B200_CUDA(launch_kernel(kern, dim3((unsigned)(p.b1 - p.b0), 2), dim3(threads), pl.smem, st, pdl, p, pl.nslots, pl.ring));
  return B200_OK;
}

// the linear sweep serves label sequences of up to 255 (8 state warps) whose staged rows leave room for a deep ring
static bool lin_sweep_ok(const LossParams& p) {
  if (p.Lmax + 1 > 256) return false;
  const int nw = (int)(align_up((size_t)(p.Lmax + 1), 32) / 32);
  return lin_plan(nw, p.ws.Kp).nslots >= nw + 4;
}

}  // namespace b200
