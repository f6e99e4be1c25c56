// ctc_loss_occ.inl -- occupancy ("gamma") kernel of the CTC loss (included by ctc_loss.cu).
//
// gamma_t(s) ~ 2^(alpha_t(s) + beta_t(s) - lp_t(l'_s)) is normalised over the row (sum_s gamma
// = 1, so neither the re-centring offsets of the sweeps nor log p(y|x) are needed); states of
// the same label are merged in label order (deterministic, no atomics).
//
// A row of work = frame t of utterance b = three contiguous pieces of the workspace: alpha[b][t]
// (2L+1 floats), beta[b][t] (2L+1, reversed state order) and lpg[b][t] (L+1).  Every WARP runs its
// own pipeline: lane 0 brings the three pieces of the next rows into the warp's shared-memory
// stages with 1-D bulk (TMA) copies that complete on the stage's mbarrier, all lanes reduce the
// row that has landed (lane j owns label positions j, j+32, ...).  Nothing of a row waits in
// registers, so the depth of the pipeline is set by shared memory (2-4 rows per warp, 16 warps per
// SM at L <= 150) and is independent of L: the same code serves L = 1 and L = 1023.
// The valid rows, in (utterance, frame) order, are cut into equal contiguous ranges, one per warp of the
// (persistent) grid: a warp streams through its range with the copies `stages` rows ahead of the reduction.
//
//   PATCH = false  writes corr[b][t][0..L] = softmax - occupancy of the blank and label columns
//                  for the separate backward pass (ctc_grad_kernel patches it into the softmax);
//   PATCH = true   the stream kernel has already written w * softmax to every column of the
//                  gradient: overwrite the blank and the (first-occurrence) label columns with
//                  w * (softmax - occupancy), and whole rows with NaN for infeasible utterances.
namespace b200 {

__device__ __forceinline__ uint32_t raw_bits(float v) { return __float_as_uint(v); }
__device__ __forceinline__ uint32_t raw_bits(__nv_bfloat16 v) { return __bfloat16_as_ushort(v); }
__device__ __forceinline__ uint32_t raw_bits(__half v) { return __half_as_ushort(v); }

constexpr int kOccMaxStages = 4;  // rows in flight per warp

// floats of one stage (alpha row, beta row, lpg row) and of one warp's share of shared memory:
// stages, un-normalised occupancies [2][Kp], label columns [Kp], 16-bit duplicate codes [Kp]
// (+16: the block scale exponents of the alpha and of the beta sweep for the row's chunks, linear lattice only)
__host__ __device__ inline int64_t occ_stage_floats(int64_t Kp, bool lin) { return 5 * Kp + (lin ? 16 : 0); }
__host__ __device__ inline int64_t occ_warp_floats(int64_t Kp, int stages, bool lin) {
  return stages * occ_stage_floats(Kp, lin) + 3 * Kp + Kp / 2;
}
// 16-bit form of LossWs::dupcode (next position + 1 <= 1023)
constexpr int kDup16Later = 0x8000;
constexpr int kDup16NextMask = 0x7fff;

// NJ > 0: L + 1 <= 32 * NJ, the row's values stay in registers between the passes (unrolled), and with
//         four stages TWO rows are reduced together (two independent dependency chains per warp: the
//         kernel is bound by the latency of one row's ~500 dependent instructions, not by issue slots);
// NJ = 0: any L, every pass re-reads shared memory, one row at a time.
// LIN: the lattice holds the HIGH WORDS of float64 states of the linear-domain sweeps (ctc_loss_lin.inl):
//      true value = word-as-double * 2^-(gcum(t) + gwarp(chunk, block)); otherwise fp32 log2 values.
template <typename T, bool PATCH, int NJ, bool LIN>
__global__ void __launch_bounds__(256) ctc_occ_kernel(LossParams p, int stages) {
  extern __shared__ __align__(128) unsigned char occ_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int64_t Kp = p.ws.Kp, Sp = p.ws.Sp;
  const int T_ = (int)p.T;
  const int64_t stagef = occ_stage_floats(Kp, LIN);
  float* wbase = reinterpret_cast<float*>(occ_smem) + warp * occ_warp_floats(Kp, stages, LIN);
  float* scratch = wbase + stages * stagef;               // [2][Kp] un-normalised occupancies of the label states
  int* s_lab = reinterpret_cast<int*>(scratch + 2 * Kp);  // [Kp] label columns of the unit's utterance
  unsigned short* s_dup = reinterpret_cast<unsigned short*>(s_lab + Kp);  // [Kp] its duplicate codes
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<float*>(occ_smem) + nwarps * occ_warp_floats(Kp, stages, LIN)) +
                   warp * kOccMaxStages;
  if (lane == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
    fence_proxy_async();
  }
  __syncwarp();
  if (lane == 0) TL_MIN(6);
  bool tl_first = true;
  (void)tl_first;
  uint32_t phase = 0;  // bit s: parity of the next completion of stage s
  T* grad = reinterpret_cast<T*>(p.grad);
  const bool pairs = NJ > 0 && stages == kOccMaxStages;

  // ---- this warp's share: a contiguous range [r0, r1) of the valid rows, in (utterance, frame) order.
  // Lengths are inputs of the call, so this runs before the wait on the sweeps.
  // Utterances are taken SHORTEST FIRST (logical index i -> utterance order[nb - 1 - i]): their sweeps finish
  // first.  The ranges are handed out through a ticket counter, and a warp waits for the completion flags of
  // the utterances of its range only: CTAs of this grid that become resident while long sweeps are still
  // running (on the SMs that host no sweep CTA, and on those whose sweep CTA has finished; a CTA of this
  // kernel and a sweep CTA do not fit on one SM together) work through the finished utterances meanwhile.
  const int nb = p.b1 - p.b0;
  auto phys = [&](int i) { return p.ws.order[p.b0 + nb - 1 - i]; };
  // (redo pass: only the utterances the linear sweeps could not certify; their log-domain sweeps signal 3 and 4)
  auto frames_of = [&](int i) {
    if (i >= nb) return 0;
    const int ub = phys(i);
    if (p.redo_mode && !p.ws.redo[ub]) return 0;
    return max(min((int)p.ilen[ub], T_), 0);
  };
  const int done_target = p.redo_mode ? 4 : 2;
  auto wait_done = [&](int ub) {  // lane 0 only
    while (ld_acquire_s32(p.ws.done + ub) < done_target) __nanosleep(256);
  };
  int tot = 0;
  for (int base = 0; base < nb; base += 32) tot += frames_of(base + lane);
  tot = __reduce_add_sync(0xffffffffu, tot);
  const int total_warps = gridDim.x * nwarps;
  // three ranges per warp of the grid (measured at C2: 190.4 / 186.7 / 185.7 / 187.0 / 188.3 us per step for
  // 1 / 2 / 3 / 4 / 6): early CTAs take more of them, but every range refills the warp's copy pipeline
  constexpr int split = 3;
  const int chunk = (tot + total_warps * split - 1) / (total_warps * split);
  for (;;) {
  int ticket = 0;
  if (lane == 0) ticket = atomicAdd(p.ws.bad + (p.redo_mode ? 7 : 5), 1);
  ticket = __shfl_sync(0xffffffffu, ticket, 0);
  const int r0 = min(ticket * chunk, tot);
  const int n = min(r0 + chunk, tot) - r0;
  if (n <= 0) break;
  // producer cursor (lane 0 issues copies `stages` rows ahead) and consumer cursor, both warp-uniform;
  // pi / bi: logical (shortest-first) indices of the utterances pb / b
  int pi = 0, pb = 0, pt = 0, pTb = 0, pL = 0;
  {
    int run = 0;
    for (int base = 0; base < nb; base += 32) {
      const int v = frames_of(base + lane);
      int incl = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
      }
      const int excl = run + incl - v;
      const unsigned hit = __ballot_sync(0xffffffffu, v > 0 && r0 >= excl && r0 < excl + v);
      if (hit) {
        const int src = __ffs(hit) - 1;
        pi = base + src;
        pb = phys(pi);
        pt = r0 - __shfl_sync(0xffffffffu, excl, src);
        pTb = __shfl_sync(0xffffffffu, v, src);
        break;
      }
      run += __shfl_sync(0xffffffffu, incl, 31);
    }
    pL = utt_labels(p, pb);
  }
  int bi = pi, b = pb, t = pt, Tb = pTb, L = pL;  // consumer
  float wgt = 1.f;
  bool poison = false;
  bool lin_skip = false;        // LIN: nothing to do for this utterance here
  int lin_k0 = 0, lin_cert = -0x40000000;  // LIN: scale reference of the utterance; largest certificate term so far
  // next utterance with frames (there is one whenever this is called: the range has rows left)
  auto next_utt = [&](int& ui, int& ub, int& uTb, int& uL) {
    do {
      ++ui;
      ub = phys(ui);
      uTb = (p.redo_mode && !p.ws.redo[ub]) ? 0 : max(min((int)p.ilen[ub], T_), 0);
    } while (uTb == 0);
    uL = utt_labels(p, ub);
  };
  // the producer runs ahead of the consumer: it is the one that waits for the lattices of an utterance
  if (lane == 0) wait_done(pb);
  __syncwarp();

  // utterance-level state of the consumer: weight, label columns and duplicate codes
  auto enter = [&]() {
    if (PATCH) {
      wgt = p.grad_scale * (p.grad_weight ? p.grad_weight[b] : 1.0f);
      poison = !isfinite(p.ws.nll_ws[b]) || isnan(wgt);
      if (LIN && p.ws.redo[b]) poison = false;  // (the redo pass owns this utterance)
    }
    if (LIN) {
      // log2 p(y|x) and the sum of the feed-forward exponents: sum_s alpha_t(s) beta_t(s) / p_t(s) = p(y|x) in every
      // frame, so every occupancy term is known to lie below this level and the relevant ones just below it
      lin_skip = p.ws.redo[b] != 0 || !isfinite(p.ws.nll_ws[b]);
      lin_k0 = p.ws.gtot[b] + (int)floorf(-p.ws.nll_ws[b] * B200_LOG2E) + 2046;
      lin_cert = -0x40000000;
    }
    const int* lab = p.ws.labels + (int64_t)b * Kp;
    const int* dup = p.ws.dupcode + (int64_t)b * Kp;
    for (int k = lane; k < L; k += 32) {
      const int dc = dup[k];
      s_lab[k] = lab[k];
      s_dup[k] = (unsigned short)((dc & kDupNextMask) | ((dc & kDupLater) ? kDup16Later : 0));
    }
    __syncwarp();
  };
  // LIN: the rows of utterance b decoded so far are certified only if no flushed state could have mattered; otherwise
  // the utterance goes to the log-domain kernels (any warp may raise the flag; the redo pass rewrites all of its rows)
  auto lin_flush = [&]() {
    if (LIN) {
      const int c = __reduce_max_sync(0xffffffffu, lin_cert);
      if (lane == 0 && c > -48 && !lin_skip) p.ws.redo[b] = 1;
      lin_cert = -0x40000000;
    }
  };
  // copies of the producer's current row into stage s, then the cursor moves on
  int issued = 0;
  auto issue = [&](int s) {
    if (lane == 0) {
      const uint32_t nA = (uint32_t)(((2 * pL + 1) * 4 + 15) & ~15);  // bytes of a lattice row that are used
      const uint32_t nL = LIN ? (uint32_t)(Kp * 4) : (uint32_t)(((pL + 1) * 4 + 15) & ~15);  // (LIN: gff in the last slot)
      const int64_t row = (int64_t)pb * T_ + pt;
      float* st = wbase + s * stagef;
      mbar_arrive_expect_tx(&bars[s], 2 * nA + nL + (LIN ? 64u : 0u));
      bulk_g2s(st, p.ws.lat + row * Sp, nA, &bars[s]);
      bulk_g2s(st + Sp, p.ws.lat + ((int64_t)p.B * T_ + row) * Sp, nA, &bars[s]);
      bulk_g2s(st + 2 * Sp, p.ws.lpg + row * Kp, nL, &bars[s]);
      if (LIN) {
        const int64_t nch = (T_ + 7) / 8;
        bulk_g2s(st + 5 * Kp, p.ws.gwarp + ((int64_t)pb * nch + pt / 8) * 8, 32, &bars[s]);
        bulk_g2s(st + 5 * Kp + 8, p.ws.gwarp + (((int64_t)p.B + pb) * nch + (pTb - 1 - pt) / 8) * 8, 32, &bars[s]);
      }
    }
    ++issued;
    if (++pt == pTb && issued < n) {
      next_utt(pi, pb, pTb, pL);
      pt = 0;
      if (lane == 0) wait_done(pb);
    }
  };
  for (int i = 0; i < min(stages, n); ++i) issue(i);
  enter();

  {
    // un-normalised log2 occupancy of the blank state k (k <= L) and of the label state k (k < L)
    // LIN: log2 of (alpha word * beta word) relative to the utterance's reference level kt, from exponent fields and
    // 20-bit mantissas; -3e38 when either state is zero.  cert: the certificate term of a state with a zero word
    // (what a flushed state could have carried, relative to p(y|x)); structural zeros (states that frame t cannot
    // reach from either end) are not flushes.
    auto lin_dec = [&](int wa, int wb, int ga, int gb, int kt, float lp, bool reach_a, bool reach_b, int& cert) -> float {
      const int ea = wa >> 20, eb = wb >> 20;
      if (ea == 0 || eb == 0) {
        if ((ea == 0 && reach_a) || (eb == 0 && reach_b)) {
          // (a word with exponent field 0 is a zero OR a subnormal: anything below 2^-1022, i.e. field 1)
          const int c = (ea ? ea : 1) + (eb ? eb : 1) - ga - gb - kt + 1 - (int)floorf(lp);
          cert = max(cert, c);
        }
        return -3.0e38f;
      }
      const float ma = __int_as_float(0x3f800000 | ((wa & 0xfffff) << 3));
      const float mb = __int_as_float(0x3f800000 | ((wb & 0xfffff) << 3));
      return (float)(ea + eb - ga - gb - kt) + lg2f(ma * mb);
    };
    int lin_t = 0, lin_Tb = 0;  // frame / length of the row being decoded (reachability)
    auto logocc = [&](const float* st, int k, float& ub, float& ul) {
      if (LIN) {
        const int* si = reinterpret_cast<const int*>(st);
        const int kt = lin_k0 + (int)st[2 * Sp + Kp - 1];  // + gff(t)
        const int2 a = *reinterpret_cast<const int2*>(si + 2 * k);
        const int ga = si[5 * Kp + (k >> 5)];
        const int fa = 2 * lin_t + 1, fb = 2 * (lin_Tb - 1 - lin_t) + 1;  // highest reachable state index from either end
        ub = lin_dec(a.x, si[Sp + 2 * (L - k)], ga, si[5 * Kp + 8 + ((L - k) >> 5)], kt, st[2 * Sp], 2 * k <= fa,
                     2 * (L - k) <= fb, lin_cert) - st[2 * Sp];
        ul = -3.0e38f;
        if (k < L)
          ul = lin_dec(a.y, si[Sp + 2 * (L - k) - 1], ga, si[5 * Kp + 8 + ((L - 1 - k) >> 5)], kt, st[2 * Sp + 1 + k],
                       2 * k + 1 <= fa, 2 * (L - 1 - k) + 1 <= fb, lin_cert) - st[2 * Sp + 1 + k];
        return;
      }
      const float2 a = *reinterpret_cast<const float2*>(st + 2 * k);
      ub = a.x + st[Sp + 2 * (L - k)] - st[2 * Sp];
      ul = -3.0e38f;
      if (k < L) ul = a.y + st[Sp + 2 * (L - k) - 1] - st[2 * Sp + 1 + k];
    };
    // blank column of row t, then the label columns through emit()
    auto emit_blank = [&](int t, const float* st, float w_blank, float inv) {
      if (lane == 0) {
        const float g0 = ex2f(st[2 * Sp]) - w_blank * inv;
        if (PATCH) grad[(int64_t)b * p.gstride_b + (int64_t)t * p.gstride_t + p.blank] = ElemTraits<T>::from_float(wgt * g0);
        else p.ws.occ[((int64_t)b * T_ + t) * Kp] = g0;
      }
    };
    auto emit = [&](int t, const float* st, const float* scr, float inv, int k) {
      const int dc = s_dup[k];
      float* out = p.ws.occ + ((int64_t)b * T_ + t) * Kp;
      if (dc & kDup16Later) {  // merged into the first occurrence of its label
        if (!PATCH) out[1 + k] = kCorrSkip;
        return;
      }
      float o = scr[k];
      int nx = (dc & kDup16NextMask) - 1;
      while (nx >= 0) {  // later occurrences of the same label, in label order
        o += scr[nx];
        nx = (s_dup[nx] & kDup16NextMask) - 1;
      }
      const float pk = ex2f(st[2 * Sp + 1 + k]);
      if (PATCH) {
        // The column already holds w * softmax from the stream kernel.  Where the occupancy is too
        // small to change that value in the output type (most (frame, label) pairs: the occupancy
        // is concentrated around the alignment), nothing is written: every 2-byte store here costs
        // a 32-byte sector fill and write-back in DRAM.
        const T with = ElemTraits<T>::from_float(wgt * (pk - o * inv));
        const T without = ElemTraits<T>::from_float(wgt * pk);
        if (raw_bits(with) != raw_bits(without) || p.redo_mode)  // (redo pass: the first pass may have patched it)
          grad[(int64_t)b * p.gstride_b + (int64_t)t * p.gstride_t + s_lab[k]] = with;
      } else {
        out[1 + k] = pk - o * inv;
      }
    };

    // RR rows (frames t, t+1, ...) in stages st[0..RR), everything unrolled so that the chains interleave
    auto process_regs = [&](auto rr_tag, int t, const float* const* st) {
      constexpr int RR = decltype(rr_tag)::value;
      constexpr int N = NJ > 0 ? NJ : 1;
      float ub[RR][N], ul[RR][N], m[RR], w_all[RR], w_blank[RR], inv[RR];
#pragma unroll
      for (int x = 0; x < RR; ++x) {
        m[x] = -3.0e38f;
#pragma unroll
        for (int j = 0; j < N; ++j) {
          const int k = lane + 32 * j;
          ub[x][j] = -3.0e38f; ul[x][j] = -3.0e38f;
          lin_t = t + x;
          if (k <= L) logocc(st[x], k, ub[x][j], ul[x][j]);
          m[x] = fmaxf(m[x], fmaxf(ub[x][j], ul[x][j]));
        }
      }
#pragma unroll
      for (int x = 0; x < RR; ++x) m[x] = warp_max_redux(m[x]);  // (REDUX.MAX on order-preserving ints: one instruction instead of five dependent shuffles)
#pragma unroll
      for (int x = 0; x < RR; ++x) {
        w_all[x] = 0.f; w_blank[x] = 0.f;
#pragma unroll
        for (int j = 0; j < N; ++j) {
          const int k = lane + 32 * j;
          ub[x][j] = ex2f(ub[x][j] - m[x]);  // -3e38 - m underflows to 0
          ul[x][j] = ex2f(ul[x][j] - m[x]);
          w_blank[x] += ub[x][j];
          w_all[x] += ul[x][j];
          if (k <= L) scratch[x * Kp + k] = ul[x][j];
        }
      }
#pragma unroll
      for (int x = 0; x < RR; ++x) {
        w_all[x] = warp_sum(w_all[x] + w_blank[x]);
        w_blank[x] = warp_sum(w_blank[x]);
        inv[x] = 1.0f / w_all[x];
      }
      __syncwarp();  // scratch complete
#pragma unroll
      for (int x = 0; x < RR; ++x) {
        emit_blank(t + x, st[x], w_blank[x], inv[x]);
#pragma unroll
        for (int j = 0; j < N; ++j)
          if (lane + 32 * j < L) emit(t + x, st[x], scratch + x * Kp, inv[x], lane + 32 * j);
      }
    };
    auto process_loop = [&](int t, const float* st) {
      float m = -3.0e38f, w_all = 0.f, w_blank = 0.f;
      lin_t = t;
      for (int k = lane; k <= L; k += 32) {
        float ub, ul;
        logocc(st, k, ub, ul);
        m = fmaxf(m, fmaxf(ub, ul));
      }
      m = warp_max_redux(m);
      for (int k = lane; k <= L; k += 32) {
        float ub, ul;
        logocc(st, k, ub, ul);
        ub = ex2f(ub - m);
        ul = ex2f(ul - m);
        w_blank += ub;
        w_all += ul;
        scratch[k] = ul;
      }
      w_all = warp_sum(w_all + w_blank);
      w_blank = warp_sum(w_blank);
      const float inv = 1.0f / w_all;
      __syncwarp();  // scratch complete
      emit_blank(t, st, w_blank, inv);
      for (int k = lane; k < L; k += 32) emit(t, st, scratch, inv, k);
    };

    for (int r = 0; r < n;) {
      const int s = r % stages;
      const int cnt = (pairs && r + 1 < n && t + 1 < Tb) ? 2 : 1;  // two frames of the same utterance
      const float* st[2] = {wbase + s * stagef, wbase + ((s + 1) % stages) * stagef};
      for (int i = 0; i < cnt; ++i) {
        const int si = (s + i) % stages;
        mbar_wait(&bars[si], (phase >> si) & 1u);
        phase ^= 1u << si;
      }
      if (tl_first && lane == 0) { TL_MIN(4); tl_first = false; }
      if (PATCH && poison) {
        // infeasible utterance: every gradient element of its frames is NaN (as torch)
        const T nanv = ElemTraits<T>::from_float(__int_as_float(0x7fc00000));
        for (int x = 0; x < cnt; ++x) {
          T* dst = grad + (int64_t)b * p.gstride_b + (int64_t)(t + x) * p.gstride_t;
          for (int v = lane; v < (int)p.V; v += 32) dst[v] = nanv;
        }
      } else if (LIN && lin_skip) {
      } else if constexpr (NJ > 0) {
        lin_Tb = Tb;
        if (cnt == 2) process_regs(std::integral_constant<int, 2>{}, t, st);
        else process_regs(std::integral_constant<int, 1>{}, t, st);
      } else {
        lin_Tb = Tb;
        process_loop(t, st[0]);
      }
      __syncwarp();  // every lane is done with the stages (and with scratch) before they are refilled
      for (int i = 0; i < cnt; ++i)
        if (issued < n) issue((s + i) % stages);
      r += cnt;
      t += cnt;
      if (t == Tb && r < n) {
        lin_flush();
        next_utt(bi, b, Tb, L);
        t = 0;
        enter();
      }
    }
    lin_flush();
  }
  }  // next ticket
  if (lane == 0) TL_MAX(5);
  // this grid must not be seen as finished before the sweeps it ran under
  pdl_wait();
}

struct OccPlan {
  int warps, stages;
  size_t smem;
};

// Largest pipeline that fits: 8 warps x up to 4 stages with two CTAs per SM if the rows are short,
// fewer warps / stages and one CTA per SM for long label sequences (a row is 20 * Kp bytes).
static OccPlan plan_occ(int64_t Kp, bool lin) {
  const size_t budgets[2] = {112u * 1024u, (size_t)max_smem_optin() - 1024u};
  for (size_t budget : budgets) {
    for (int warps = 8; warps >= 1; warps >>= 1) {
      for (int stages = kOccMaxStages; stages >= 2; --stages) {
        const size_t smem = (size_t)warps * ((size_t)occ_warp_floats(Kp, stages, lin) * 4 + kOccMaxStages * 8);
        if (smem <= budget) return OccPlan{warps, stages, smem};
      }
    }
  }
  return OccPlan{0, 0, 0};
}

template <typename T, bool PATCH, int NJ, bool LIN>
static int launch_occ_nj(const LossParams& p, cudaStream_t st) {
  const OccPlan pl = plan_occ(p.ws.Kp, LIN);
  B200_CHECK_ARG(pl.warps > 0, B200_ERR_UNSUPPORTED_SHAPE, "Lmax=%lld: a lattice row does not fit in shared memory",
                 (long long)p.Lmax);
  B200_CUDA(set_max_smem_cached((const void*)ctc_occ_kernel<T, PATCH, NJ, LIN>, (int)pl.smem));
  const int64_t units = (int64_t)(p.b1 - p.b0) * p.T;  // rows, at least one per warp
  int per_sm = 1;
  B200_CUDA(max_blocks_per_sm_cached((const void*)ctc_occ_kernel<T, PATCH, NJ, LIN>, pl.warps * 32, pl.smem, &per_sm));
  if (per_sm < 1) per_sm = 1;
  int64_t g = (units + pl.warps - 1) / pl.warps;
  const int64_t cap = (int64_t)num_sms() * per_sm;  // persistent: every CTA resident
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  B200_CUDA(launch_kernel(ctc_occ_kernel<T, PATCH, NJ, LIN>, dim3((unsigned)g), dim3(pl.warps * 32), pl.smem, st, PATCH, p,
                          pl.stages));
  return B200_OK;
}

}  // namespace b200
#include "ctc_loss_occ_wide.inl"
namespace b200 {

template <typename T, bool PATCH, bool LIN>
static int launch_occ_t(const LossParams& p, cudaStream_t st) {
  if (p.Lmax + 1 <= 32 * 5) return launch_occ_nj<T, PATCH, 5, LIN>(p, st);
  if (LIN || p.Lmax + 1 <= 32 * 10) return launch_occ_nj<T, PATCH, 10, LIN>(p, st);  // (LIN: Lmax <= 255)
  // long rows: a CTA per row (developer switch b200_ctc_debug_single_cta_sweep: the warp-per-row kernel)
  if (!LIN && p.Lmax + 1 <= kWideThreads * kWideIter && !g_single_cta_sweep) return launch_occ_wide<T, PATCH>(p, st);
  return launch_occ_nj<T, PATCH, 0, LIN>(p, st);
}

// occupancies into ws.occ (separate forward / backward calls)
static int launch_occ(const LossParams& p, cudaStream_t st) { return launch_occ_t<float, false, false>(p, st); }

// blank / label columns of the gradient patched in place (fused forward + backward); lin: the lattice of the
// linear-domain sweeps
static int launch_patch(const LossParams& p, int dtype, cudaStream_t st, bool lin = false) {
  if (lin) {
    switch (dtype) {
      case B200_F32: return launch_occ_t<float, true, true>(p, st);
      case B200_BF16: return launch_occ_t<__nv_bfloat16, true, true>(p, st);
      case B200_F16: return launch_occ_t<__half, true, true>(p, st);
    }
  }
  switch (dtype) {
    case B200_F32: return launch_occ_t<float, true, false>(p, st);
    case B200_BF16: return launch_occ_t<__nv_bfloat16, true, false>(p, st);
    case B200_F16: return launch_occ_t<__half, true, false>(p, st);
  }
  set_error("unknown dtype %d", dtype);
  return B200_ERR_INVALID_ARGUMENT;
}

}  // namespace b200
