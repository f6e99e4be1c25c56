// ctc_loss_occ.inl -- occupancy ("gamma") kernel of the CTC loss (included by ctc_loss.cu).
//
// One warp per (utterance, frame) row, rows dealt round-robin to all warps of the grid, every
// value of a row held in registers (lane j owns the label positions j, j+32, ...).
// gamma_t(s) ~ 2^(alpha_t(s) + beta_t(s) - lp_t(l'_s)) is normalised over the row (sum_s gamma
// = 1, so neither the re-centring offsets of the sweeps nor log p(y|x) are needed); states of
// the same label are merged in label order (deterministic, no atomics).
//
//   PATCH = false  writes corr[b][t][0..L] = softmax - occupancy of the blank and label columns
//                  for the separate backward pass (ctc_grad_kernel patches it into the softmax);
//   PATCH = true   the stream kernel has already written w * softmax to every column of the
//                  gradient: overwrite the blank and the (first-occurrence) label columns with
//                  w * (softmax - occupancy), and whole rows with NaN for infeasible utterances.
namespace b200 {

constexpr int kOccWarps = 8;

__device__ __forceinline__ uint32_t raw_bits(float v) { return __float_as_uint(v); }
__device__ __forceinline__ uint32_t raw_bits(__nv_bfloat16 v) { return __bfloat16_as_ushort(v); }
__device__ __forceinline__ uint32_t raw_bits(__half v) { return __half_as_ushort(v); }
constexpr int kOccFrames = 4;  // consecutive frames of one utterance per task

template <int NJ>
struct OccRaw {
  float a0[NJ], b0[NJ], a1[NJ], b1[NJ], ll[NJ];
  float lp0;
};

// Work unit = kOccFrames consecutive frames of one utterance, pulled by each WARP from an atomic
// ticket counter: everything that depends only on the utterance (label columns, duplicate codes,
// weight) is loaded once per unit, the lattice rows of the unit are double-buffered in registers.
template <typename T, int NJ, bool PATCH>
__global__ void __launch_bounds__(kOccWarps * 32, NJ <= 5 ? 2 : 1) ctc_occ_kernel(LossParams p) {
  __shared__ float s_occ[kOccWarps][NJ * 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t Kp = p.ws.Kp, Sp = p.ws.Sp;
  const int T_ = (int)p.T;
  const int nblk = (T_ + kOccFrames - 1) / kOccFrames;
  const int ntask = (p.b1 - p.b0) * nblk;
  float* occw = s_occ[warp];
  T* grad = reinterpret_cast<T*>(p.grad);
  int* counter = p.ws.bad + 3;
  pdl_wait();  // staged under the sweeps: their lattices (and the streamed gradient) are complete from here

  // issue every global load of row (b, t) (nothing waits on them here)
  auto issue = [&](int b, int t, int L, OccRaw<NJ>& w) {
    const int64_t row = (int64_t)b * T_ + t;
    const float* A = p.ws.lat + row * Sp;
    const float* Bt = p.ws.lat + ((int64_t)p.B * T_ + row) * Sp;
    const float* lp = p.ws.lpg + row * Kp;
    w.lp0 = ld_nc_f32(lp);
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int k = lane + 32 * j;
      const int kb = min(k, L);              // clamped: a valid blank state
      const int kl = min(k, max(L - 1, 0));  // clamped: a valid label state (if L > 0)
      w.a0[j] = ld_nc_f32(A + 2 * kb);
      w.b0[j] = ld_nc_f32(Bt + 2 * (L - kb));
      w.a1[j] = 0.f; w.b1[j] = 0.f; w.ll[j] = 0.f;
      if (L > 0) {
        w.a1[j] = ld_nc_f32(A + 2 * kl + 1);
        w.b1[j] = ld_nc_f32(Bt + 2 * (L - kl) - 1);
        w.ll[j] = ld_nc_f32(lp + 1 + kl);
      }
    }
  };

  int q = fetch_ticket(counter, lane);
  while (q < ntask) {
    const int b = p.b0 + q / nblk;
    const int t0 = (q % nblk) * kOccFrames;
    q = fetch_ticket(counter, lane);  // next ticket: its latency hides behind this unit
    const int Tb = min((int)p.ilen[b], T_);
    if (t0 >= Tb) continue;
    const int t1 = min(t0 + kOccFrames, Tb);
    const int L = (int)p.olen[b];
    const int* lab = p.ws.labels + (int64_t)b * Kp;
    const int* dup = p.ws.dupcode + (int64_t)b * Kp;
    float wgt = 1.f;
    bool poison = false;
    if (PATCH) {
      wgt = p.grad_scale * (p.grad_weight ? p.grad_weight[b] : 1.0f);
      poison = isinf(p.ws.nll_ws[b]) || isnan(wgt);
    }
    int col[NJ], dc[NJ];
    int anydup = 0;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int k = lane + 32 * j;
      col[j] = 0; dc[j] = 0;
      if (k < L) { col[j] = lab[k]; dc[j] = dup[k]; }
      anydup |= dc[j];
    }
    const bool merge = __any_sync(0xffffffffu, anydup != 0);

    auto process = [&](int t, const OccRaw<NJ>& w) {
      const int64_t row = (int64_t)b * T_ + t;
      if (PATCH && poison) {
        // infeasible utterance: every gradient element of its frames is NaN (as torch)
        T* dst = grad + (int64_t)b * p.gstride_b + (int64_t)t * p.gstride_t;
        const T nanv = ElemTraits<T>::from_float(__int_as_float(0x7fc00000));
        for (int v = lane; v < (int)p.V; v += 32) dst[v] = nanv;
        return;
      }
      const float lp0 = w.lp0;
      float ub[NJ], ul[NJ];
      float m = -3.0e38f;
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int k = lane + 32 * j;
        ub[j] = (k <= L) ? (w.a0[j] + w.b0[j] - lp0) : -3.0e38f;
        ul[j] = (k < L) ? (w.a1[j] + w.b1[j] - w.ll[j]) : -3.0e38f;
        m = fmaxf(m, fmaxf(ub[j], ul[j]));
      }
      m = warp_max(m);
      float w_all = 0.f, w_blank = 0.f;
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        ub[j] = ex2f(ub[j] - m);  // -3e38 - m underflows to 0
        ul[j] = ex2f(ul[j] - m);
        w_blank += ub[j];
        w_all += ul[j];
      }
      w_all = warp_sum(w_all + w_blank);
      w_blank = warp_sum(w_blank);
      const float inv = 1.0f / w_all;
      if (merge) {
        // merge the occupancies of repeated labels into their first occurrence, in label order
        __syncwarp();
#pragma unroll
        for (int j = 0; j < NJ; ++j) occw[lane + 32 * j] = ul[j];
        __syncwarp();
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          if (dc[j] > 0 && !(dc[j] & kDupLater)) {
            float o = ul[j];
            int nx = (dc[j] & kDupNextMask) - 1;
            while (nx >= 0) {
              o += occw[nx];
              nx = (dup[nx] & kDupNextMask) - 1;
            }
            ul[j] = o;
          }
        }
      }
      if (PATCH && (p.debug & 16)) {
        if (w_blank * inv == 123.f) p.ws.occ[row] = inv;  // developer variant: no scattered stores
      } else if (PATCH) {
        T* dst = grad + (int64_t)b * p.gstride_b + (int64_t)t * p.gstride_t;
        if (lane == 0) dst[p.blank] = ElemTraits<T>::from_float(wgt * (ex2f(lp0) - w_blank * inv));
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          const int k = lane + 32 * j;
          if (k < L && !(dc[j] & kDupLater)) {
            // The column already holds w * softmax from the stream kernel.  Where the occupancy is
            // too small to change that value in the output type (most (frame, label) pairs: the
            // occupancy is concentrated around the alignment), nothing is written: every 2-byte
            // store here costs a 32-byte sector fill and write-back in DRAM.
            const float pk = ex2f(w.ll[j]);
            const T with = ElemTraits<T>::from_float(wgt * (pk - ul[j] * inv));
            const T without = ElemTraits<T>::from_float(wgt * pk);
            if (raw_bits(with) != raw_bits(without)) dst[col[j]] = with;
          }
        }
      } else {
        float* out = p.ws.occ + row * Kp;
        if (lane == 0) out[0] = ex2f(lp0) - w_blank * inv;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          const int k = lane + 32 * j;
          if (k < L) out[1 + k] = (dc[j] & kDupLater) ? kCorrSkip : ex2f(w.ll[j]) - ul[j] * inv;
        }
      }
    };

    // two register sets: the loads of the next row are in flight while this row is reduced and stored
    OccRaw<NJ> ra, rb;
    issue(b, t0, L, ra);
    for (int t = t0; t < t1; t += 2) {
      if (t + 1 < t1) issue(b, t + 1, L, rb);
      process(t, ra);
      if (t + 1 >= t1) break;
      if (t + 2 < t1) issue(b, t + 2, L, ra);
      process(t + 1, rb);
    }
  }
}

template <typename T, int NJ, bool PATCH>
static int launch_occ_nj(const LossParams& p, cudaStream_t st) {
  const int64_t nrows = ((int64_t)(p.b1 - p.b0) * ((p.T + kOccFrames - 1) / kOccFrames));  // tasks
  int per_sm = 1;
  B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ctc_occ_kernel<T, NJ, PATCH>, kOccWarps * 32, 0));
  if (per_sm < 1) per_sm = 1;
  int64_t g = (nrows + kOccWarps - 1) / kOccWarps;
  const int64_t cap = (int64_t)num_sms() * per_sm;  // persistent: every CTA resident
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  B200_CUDA(launch_kernel(ctc_occ_kernel<T, NJ, PATCH>, dim3((unsigned)g), dim3(kOccWarps * 32), 0, st, PATCH, p));
  return B200_OK;
}

template <typename T, bool PATCH>
static int launch_occ_t(const LossParams& p, cudaStream_t st) {
  const int nj = (int)((p.Lmax + 1 + 31) / 32);
  if (nj <= 5) return launch_occ_nj<T, 5, PATCH>(p, st);
  if (nj <= 10) return launch_occ_nj<T, 10, PATCH>(p, st);
  if (nj <= 20) return launch_occ_nj<T, 20, PATCH>(p, st);
  return launch_occ_nj<T, 32, PATCH>(p, st);
}

// occupancies into ws.occ (separate forward / backward calls)
static int launch_occ(const LossParams& p, cudaStream_t st) { return launch_occ_t<float, false>(p, st); }

// blank / label columns of the gradient patched in place (fused forward + backward)
static int launch_patch(const LossParams& p, int dtype, cudaStream_t st) {
  switch (dtype) {
    case B200_F32: return launch_occ_t<float, true>(p, st);
    case B200_BF16: return launch_occ_t<__nv_bfloat16, true>(p, st);
    case B200_F16: return launch_occ_t<__half, true>(p, st);
  }
  set_error("unknown dtype %d", dtype);
  return B200_ERR_INVALID_ARGUMENT;
}

}  // namespace b200
