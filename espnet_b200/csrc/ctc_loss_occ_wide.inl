// ctc_loss_occ_wide.inl -- the occupancy / patch kernel for long label sequences (L + 1 > 320).
//
// ctc_occ_kernel gives every WARP its own rows; at L = 1000 a row is 20 KB of shared memory per stage, so only four
// warps fit on an SM and each of them walks 63 strided iterations per pass: 644 us at C5 for 1.3 GB of traffic.
// Here a CTA of 8 warps works on ONE row at a time (each thread owns label positions tid, tid + 256, ...: at most four,
// kept in registers between the passes), a ninth warp brings the rows in with 1-D bulk (TMA) copies three stages ahead,
// and three CTAs share an SM.  Per row: two named barriers (row maximum, row sums); the partial results and the
// un-normalised occupancies are double-buffered by row parity, and a stage is released by one arrival per warp, so
// there is no third barrier.  Work distribution (ticketed ranges of valid rows, shortest utterances first, the producer
// waits for the completion flags of the sweeps) is the one of ctc_occ_kernel.
namespace b200 {

constexpr int kWideWarps = 8;
constexpr int kWideThreads = kWideWarps * 32;
constexpr int kWideStages = 3;
constexpr int kWideIter = 4;  // label positions per thread: L + 1 <= 1024

__host__ __device__ inline size_t occ_wide_smem(int64_t Kp) {
  // stages (alpha, beta, lpg rows), un-normalised occupancies [2][Kp], label columns [Kp], duplicate codes [Kp] u16,
  // partial maxima / sums [2][3][8], barriers
  return (size_t)kWideStages * 5 * Kp * 4 + 2 * Kp * 4 + Kp * 4 + Kp * 2 + 2 * 3 * kWideWarps * 4 + 2 * kWideStages * 8 + 64;
}

template <typename T, bool PATCH>
__global__ void __launch_bounds__(kWideThreads + 32) ctc_occ_wide_kernel(LossParams p) {
  extern __shared__ __align__(128) unsigned char ow_smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t Kp = p.ws.Kp, Sp = p.ws.Sp;
  const int T_ = (int)p.T;
  const int64_t stagef = 5 * Kp;
  float* stages = reinterpret_cast<float*>(ow_smem);
  float* scratch = stages + kWideStages * stagef;                       // [2][Kp]
  int* s_lab = reinterpret_cast<int*>(scratch + 2 * Kp);                // [Kp]
  unsigned short* s_dup = reinterpret_cast<unsigned short*>(s_lab + Kp);  // [Kp]
  float* s_red = reinterpret_cast<float*>(s_dup + Kp + (Kp & 1));       // [2][3][kWideWarps]
  uint64_t* full = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(s_red + 2 * 3 * kWideWarps) + 15) & ~(uintptr_t)15);
  uint64_t* empty = full + kWideStages;
  __shared__ int s_ticket;
  if (tid == 0) {
    for (int s = 0; s < kWideStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kWideWarps); }
    mbar_fence_init();
    fence_proxy_async();
  }
  __syncthreads();
  T* grad = reinterpret_cast<T*>(p.grad);
  const bool producer = warp == kWideWarps;

  const int nb = p.b1 - p.b0;
  auto phys = [&](int i) { return p.ws.order[p.b0 + nb - 1 - i]; };  // shortest utterances first
  auto frames_of = [&](int i) { return i < nb ? max(min((int)p.ilen[phys(i)], T_), 0) : 0; };
  int tot = 0;
  for (int base = 0; base < nb; base += 32) tot += frames_of(base + lane);
  tot = __reduce_add_sync(0xffffffffu, tot);
  constexpr int split = 3;
  const int chunk = (tot + gridDim.x * split - 1) / (gridDim.x * split);
  uint32_t full_par = 0, empty_par = 0;  // bit s: parity of the next completion of full[s] / empty[s] this thread waits for
  int seq = 0;                           // rows this CTA has pushed through the ring so far (stage = seq % kWideStages)

  for (;;) {
    __syncthreads();  // (everybody is done with the previous ticket: s_ticket, label tables)
    if (tid == 0) s_ticket = atomicAdd(p.ws.bad + 5, 1);
    __syncthreads();
    const int r0 = min(s_ticket * chunk, tot);
    const int n = min(r0 + chunk, tot) - r0;
    if (n <= 0) break;
    // cursor at row r0: logical utterance index, utterance, frame, its length
    int ci = 0, cb = 0, ct = 0, cTb = 0;
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
          ci = base + src;
          cb = phys(ci);
          ct = r0 - __shfl_sync(0xffffffffu, excl, src);
          cTb = __shfl_sync(0xffffffffu, v, src);
          break;
        }
        run += __shfl_sync(0xffffffffu, incl, 31);
      }
    }
    auto next_utt = [&](int& ui, int& ub, int& uTb) {
      do {
        ++ui;
        ub = phys(ui);
        uTb = max(min((int)p.ilen[ub], T_), 0);
      } while (uTb == 0);
    };

    if (producer) {
      // ---------------------------------------------------------------- producer warp (lane 0 issues)
      int L = utt_labels(p, cb);
      if (lane == 0) while (ld_acquire_s32(p.ws.done + cb) < 2) __nanosleep(256);
      __syncwarp();
      for (int r = 0; r < n; ++r, ++seq) {
        const int s = seq % kWideStages;
        if (seq >= kWideStages) {
          mbar_wait(&empty[s], (empty_par >> s) & 1u);
          empty_par ^= 1u << s;
        }
        if (lane == 0) {
          const uint32_t nA = (uint32_t)(((2 * L + 1) * 4 + 15) & ~15);
          const uint32_t nL = (uint32_t)(((L + 1) * 4 + 15) & ~15);
          const int64_t row = (int64_t)cb * T_ + ct;
          float* st = stages + s * stagef;
          mbar_arrive_expect_tx(&full[s], 2 * nA + nL);
          bulk_g2s(st, p.ws.lat + row * Sp, nA, &full[s]);
          bulk_g2s(st + Sp, p.ws.lat + ((int64_t)p.B * T_ + row) * Sp, nA, &full[s]);
          bulk_g2s(st + 2 * Sp, p.ws.lpg + row * Kp, nL, &full[s]);
        }
        if (++ct == cTb && r + 1 < n) {
          next_utt(ci, cb, cTb);
          ct = 0;
          L = utt_labels(p, cb);
          if (lane == 0) while (ld_acquire_s32(p.ws.done + cb) < 2) __nanosleep(256);
          __syncwarp();
        }
      }
      continue;
    }

    // ------------------------------------------------------------------ consumers: 8 warps, one row at a time
    int L = utt_labels(p, cb);
    float wgt = 1.f;
    bool poison = false;
    auto enter = [&]() {  // label columns and duplicate codes of utterance cb (between two CTA-wide barriers)
      // The utterance's nll (written by its alpha sweep before it signals) decides whether its rows are poisoned: the
      // consumers wait for the completion flags themselves -- the producer warp's wait orders only ITS copies, and a
      // consumer can get here before the sweeps of the next utterance have finished (a stale nll from an earlier
      // call made this decision at random).
      if (tid == 0) while (ld_acquire_s32(p.ws.done + cb) < 2) __nanosleep(64);
      asm volatile("bar.sync 1, %0;" ::"n"(kWideThreads) : "memory");
      if (PATCH) {
        wgt = p.grad_scale * (p.grad_weight ? p.grad_weight[cb] : 1.0f);
        poison = !isfinite(ld_cg_f32(p.ws.nll_ws + cb)) || isnan(wgt);
      }
      const int* lab = p.ws.labels + (int64_t)cb * Kp;
      const int* dup = p.ws.dupcode + (int64_t)cb * Kp;
      for (int k = tid; k < L; k += kWideThreads) {
        const int dc = dup[k];
        s_lab[k] = lab[k];
        s_dup[k] = (unsigned short)((dc & kDupNextMask) | ((dc & kDupLater) ? kDup16Later : 0));
      }
      asm volatile("bar.sync 1, %0;" ::"n"(kWideThreads) : "memory");
    };
    enter();
    for (int r = 0; r < n; ++r, ++seq) {
      const int s = seq % kWideStages;
      const float* st = stages + s * stagef;
      float* scr = scratch + (seq & 1) * Kp;
      float* red = s_red + (seq & 1) * 3 * kWideWarps;
      mbar_wait(&full[s], (full_par >> s) & 1u);
      full_par ^= 1u << s;
      if (PATCH && poison) {
        // infeasible utterance: every gradient element of its frames is NaN (as torch)
        const T nanv = ElemTraits<T>::from_float(__int_as_float(0x7fc00000));
        T* dst = grad + (int64_t)cb * p.gstride_b + (int64_t)ct * p.gstride_t;
        for (int v = tid; v < (int)p.V; v += kWideThreads) dst[v] = nanv;
      } else {
        // un-normalised log2 occupancy of the blank state k (k <= L) and of the label state k (k < L)
        float ub[kWideIter], ul[kWideIter];
        float m = -3.0e38f;
#pragma unroll
        for (int j = 0; j < kWideIter; ++j) {
          const int k = tid + kWideThreads * j;
          ub[j] = -3.0e38f; ul[j] = -3.0e38f;
          if (k <= L) {
            const float2 a = *reinterpret_cast<const float2*>(st + 2 * k);
            ub[j] = a.x + st[Sp + 2 * (L - k)] - st[2 * Sp];
            if (k < L) ul[j] = a.y + st[Sp + 2 * (L - k) - 1] - st[2 * Sp + 1 + k];
          }
          m = fmaxf(m, fmaxf(ub[j], ul[j]));
        }
        m = warp_max_redux(m);
        if (lane == 0) red[warp] = m;
        asm volatile("bar.sync 1, %0;" ::"n"(kWideThreads) : "memory");
        m = red[lane & (kWideWarps - 1)];
        m = warp_max_redux(m);
        float w_all = 0.f, w_blank = 0.f;
#pragma unroll
        for (int j = 0; j < kWideIter; ++j) {
          const int k = tid + kWideThreads * j;
          ub[j] = ex2f(ub[j] - m);  // -3e38 - m underflows to 0
          ul[j] = ex2f(ul[j] - m);
          w_blank += ub[j];
          w_all += ul[j];
          if (k <= L) scr[k] = ul[j];
        }
        w_all = warp_sum(w_all + w_blank);
        w_blank = warp_sum(w_blank);
        if (lane == 0) { red[kWideWarps + warp] = w_all; red[2 * kWideWarps + warp] = w_blank; }
        asm volatile("bar.sync 1, %0;" ::"n"(kWideThreads) : "memory");
        w_all = warp_sum(lane < kWideWarps ? red[kWideWarps + lane] : 0.f);
        w_blank = warp_sum(lane < kWideWarps ? red[2 * kWideWarps + lane] : 0.f);
        const float inv = 1.0f / w_all;
        float* out = p.ws.occ + ((int64_t)cb * T_ + ct) * Kp;
        if (tid == 0) {
          const float g0 = ex2f(st[2 * Sp]) - w_blank * inv;
          if (PATCH) grad[(int64_t)cb * p.gstride_b + (int64_t)ct * p.gstride_t + p.blank] = ElemTraits<T>::from_float(wgt * g0);
          else out[0] = g0;
        }
#pragma unroll
        for (int j = 0; j < kWideIter; ++j) {
          const int k = tid + kWideThreads * j;
          if (k >= L) continue;
          const int dc = s_dup[k];
          if (dc & kDup16Later) {  // merged into the first occurrence of its label
            if (!PATCH) out[1 + k] = kCorrSkip;
            continue;
          }
          float o = scr[k];
          int nx = (dc & kDup16NextMask) - 1;
          while (nx >= 0) {  // later occurrences of the same label, in label order
            o += scr[nx];
            nx = (s_dup[nx] & kDup16NextMask) - 1;
          }
          const float pk = ex2f(st[2 * Sp + 1 + k]);
          if (PATCH) {
            const T with = ElemTraits<T>::from_float(wgt * (pk - o * inv));
            const T without = ElemTraits<T>::from_float(wgt * pk);
            if (raw_bits(with) != raw_bits(without))
              grad[(int64_t)cb * p.gstride_b + (int64_t)ct * p.gstride_t + s_lab[k]] = with;
          } else {
            out[1 + k] = pk - o * inv;
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);  // this warp is done with the stage
      if (++ct == cTb && r + 1 < n) {
        next_utt(ci, cb, cTb);
        ct = 0;
        L = utt_labels(p, cb);
        enter();
      }
    }
  }
  // this grid must not be seen as finished before the sweeps it ran under
  pdl_wait();
}

template <typename T, bool PATCH>
static int launch_occ_wide(const LossParams& p, cudaStream_t st) {
  const size_t smem = occ_wide_smem(p.ws.Kp);
  B200_CHECK_ARG(smem <= (size_t)max_smem_optin(), B200_ERR_UNSUPPORTED_SHAPE, "Lmax=%lld: a lattice row does not fit in shared memory",
                 (long long)p.Lmax);
  B200_CUDA(set_max_smem_cached((const void*)ctc_occ_wide_kernel<T, PATCH>, (int)smem));
  int per_sm = 1;
  B200_CUDA(max_blocks_per_sm_cached((const void*)ctc_occ_wide_kernel<T, PATCH>, kWideThreads + 32, smem, &per_sm));
  if (per_sm < 1) per_sm = 1;
  const int64_t units = (int64_t)(p.b1 - p.b0) * p.T;
  int64_t g = (int64_t)num_sms() * per_sm;  // persistent: every CTA resident
  if (g > units) g = units;
  if (g < 1) g = 1;
  B200_CUDA(launch_kernel(ctc_occ_wide_kernel<T, PATCH>, dim3((unsigned)g), dim3(kWideThreads + 32), smem, st, PATCH, p));
  return B200_OK;
}

}  // namespace b200
