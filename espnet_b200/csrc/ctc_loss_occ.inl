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

template <int NJ>
struct OccRaw {
  float a0[NJ], b0[NJ], a1[NJ], b1[NJ], ll[NJ];
  float lp0;
  int b, t, L;
};

template <typename T, int NJ, bool PATCH>
__global__ void __launch_bounds__(kOccWarps * 32) ctc_occ_kernel(LossParams p) {
  __shared__ float s_occ[kOccWarps][NJ * 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t Kp = p.ws.Kp, Sp = p.ws.Sp;
  const int T_ = (int)p.T;
  const int64_t nrows = (int64_t)(p.b1 - p.b0) * T_;
  const int64_t stride = (int64_t)gridDim.x * kOccWarps;
  float* occw = s_occ[warp];
  T* grad = reinterpret_cast<T*>(p.grad);

  // next row >= r of this warp's list that lies inside its utterance, or nrows
  auto next_valid = [&](int64_t r) {
    while (r < nrows) {
      const int b = p.b0 + (int)(r / T_);
      if ((int)(r % T_) < min((int)p.ilen[b], T_)) break;
      r += stride;
    }
    return r;
  };
  // issue every global load of row r (nothing waits on them here)
  auto issue = [&](int64_t r, OccRaw<NJ>& w) {
    w.b = p.b0 + (int)(r / T_);
    w.t = (int)(r % T_);
    w.L = (int)p.olen[w.b];
    const int64_t row = (int64_t)w.b * T_ + w.t;
    const float* A = p.ws.lat + row * Sp;
    const float* Bt = p.ws.lat + ((int64_t)p.B * T_ + row) * Sp;
    const float* lp = p.ws.lpg + row * Kp;
    w.lp0 = ld_nc_f32(lp);
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int k = lane + 32 * j;
      const int kb = min(k, w.L);                // clamped: a valid blank state
      const int kl = min(k, max(w.L - 1, 0));    // clamped: a valid label state (if L > 0)
      w.a0[j] = ld_nc_f32(A + 2 * kb);
      w.b0[j] = ld_nc_f32(Bt + 2 * (w.L - kb));
      w.a1[j] = 0.f; w.b1[j] = 0.f; w.ll[j] = 0.f;
      if (w.L > 0) {
        w.a1[j] = ld_nc_f32(A + 2 * kl + 1);
        w.b1[j] = ld_nc_f32(Bt + 2 * (w.L - kl) - 1);
        w.ll[j] = ld_nc_f32(lp + 1 + kl);
      }
    }
  };
  auto process = [&](const OccRaw<NJ>& w) {
    const int b = w.b, t = w.t, L = w.L;
    const int64_t row = (int64_t)b * T_ + t;
    const int* lab = p.ws.labels + (int64_t)b * Kp;
    const int* dup = p.ws.dupcode + (int64_t)b * Kp;
    float wgt = 1.f;
    if (PATCH) {
      wgt = p.grad_scale * (p.grad_weight ? p.grad_weight[b] : 1.0f);
      if (isinf(p.ws.nll_ws[b]) || isnan(wgt)) {
        // infeasible utterance: every gradient element of its frames is NaN (as torch)
        T* dst = grad + (int64_t)b * p.gstride_b + (int64_t)t * p.gstride_t;
        const T nanv = ElemTraits<T>::from_float(__int_as_float(0x7fc00000));
        for (int v = lane; v < (int)p.V; v += 32) dst[v] = nanv;
        return;
      }
    }
    const float lp0 = w.lp0;
    float ub[NJ], ul[NJ];
    int col[NJ], dc[NJ];
    float m = -3.0e38f;
    int anydup = 0;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int k = lane + 32 * j;
      col[j] = 0; dc[j] = 0;
      if (k < L) { col[j] = lab[k]; dc[j] = dup[k]; }
      ub[j] = (k <= L) ? (w.a0[j] + w.b0[j] - lp0) : -3.0e38f;
      ul[j] = (k < L) ? (w.a1[j] + w.b1[j] - w.ll[j]) : -3.0e38f;
      m = fmaxf(m, fmaxf(ub[j], ul[j]));
      anydup |= dc[j];
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
    if (__any_sync(0xffffffffu, anydup != 0)) {
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
    if (PATCH) {
      T* dst = grad + (int64_t)b * p.gstride_b + (int64_t)t * p.gstride_t;
      if (lane == 0) dst[p.blank] = ElemTraits<T>::from_float(wgt * (ex2f(lp0) - w_blank * inv));
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int k = lane + 32 * j;
        if (k < L && !(dc[j] & kDupLater))
          dst[col[j]] = ElemTraits<T>::from_float(wgt * (ex2f(w.ll[j]) - ul[j] * inv));
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
  int64_t r = next_valid((int64_t)blockIdx.x * kOccWarps + warp);
  if (r < nrows) issue(r, ra);
  while (r < nrows) {
    int64_t r2 = next_valid(r + stride);
    if (r2 < nrows) issue(r2, rb);
    process(ra);
    if (r2 >= nrows) break;
    r = next_valid(r2 + stride);
    if (r < nrows) issue(r, ra);
    process(rb);
  }
}

template <typename T, int NJ, bool PATCH>
static int launch_occ_nj(const LossParams& p, cudaStream_t st) {
  const int64_t nrows = (int64_t)(p.b1 - p.b0) * p.T;
  int per_sm = 1;
  B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ctc_occ_kernel<T, NJ, PATCH>, kOccWarps * 32, 0));
  if (per_sm < 1) per_sm = 1;
  int64_t g = (nrows + kOccWarps - 1) / kOccWarps;
  const int64_t cap = (int64_t)num_sms() * per_sm;  // persistent: every CTA resident
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  ctc_occ_kernel<T, NJ, PATCH><<<(unsigned)g, kOccWarps * 32, 0, st>>>(p);
  B200_CUDA(cudaGetLastError());
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
