// ctc_loss_gamma.inl -- occupancy kernel of the CTC loss (included by ctc_loss.cu).
//
// One warp per (utterance, frame) row, tasks pulled from an atomic counter as in
// ctc_loss_rows.inl.  gamma_t(s) ~ 2^(alpha_t(s) + beta_t(s) - lp_t(l'_s)) is normalised over
// the row (sum_s gamma = 1 exactly in exact arithmetic, so neither the re-centring offsets
// of the sweeps nor log p(y|x) are needed), states of the same label are merged in label
// order (deterministic, no atomics) and the result corr[b][t][0..L] = softmax - occupancy of
// the blank and label columns is what the gradient kernel patches into the streamed softmax.
namespace b200 {

constexpr int kGammaU = 10;  // lattice states per lane loaded in one batch (covers 2L+1 <= 320)

__global__ void __launch_bounds__(kRowsMaxWarps * 32, 1) ctc_gamma_kernel(LossParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t Kp = p.ws.Kp, Sp = p.ws.Sp;
  const size_t per_warp = 2 * Kp * sizeof(int) + Sp * sizeof(float);
  int* s_nxt = reinterpret_cast<int*>(smem + warp * per_warp);
  int* s_fst = s_nxt + Kp;
  float* gam = reinterpret_cast<float*>(s_fst + Kp);

  const int nblk = (int)((p.T + kFramesPerTask - 1) / kFramesPerTask);
  const int ntask = (p.b1 - p.b0) * nblk;
  int* counter = p.ws.bad + 3;
  int cached_b = -1;

  int q = fetch_ticket(counter, lane);
  while (q < ntask) {
    const int q_next = fetch_ticket(counter, lane);
    const int b = p.b0 + q / nblk;
    const int t0 = (q % nblk) * kFramesPerTask;
    const int Tb = (int)p.ilen[b];
    const int t1 = min(min(t0 + kFramesPerTask, (int)p.T), Tb);
    const int L = (int)p.olen[b];
    q = q_next;
    if (t0 >= Tb) continue;
    if (b != cached_b) {
      __syncwarp();
      for (int k = lane; k < L; k += 32) {
        s_nxt[k] = p.ws.dup_next[(int64_t)b * Kp + k];
        s_fst[k] = p.ws.is_first[(int64_t)b * Kp + k];
      }
      cached_b = b;
      __syncwarp();
    }
    const int S = 2 * L + 1;
    for (int t = t0; t < t1; ++t) {
      const int64_t row = (int64_t)b * p.T + t;
      const float* A = p.ws.lat + row * Sp;
      const float* Bt = p.ws.lat + ((int64_t)p.B * p.T + row) * Sp;
      const float* lp = p.ws.lpg + row * Kp;
      float m = -3.0e38f;
      for (int base = 0; base < S; base += 32 * kGammaU) {
        float a[kGammaU], bb[kGammaU], l[kGammaU];
#pragma unroll
        for (int i = 0; i < kGammaU; ++i) {
          const int s = base + lane + 32 * i;
          const int sc = s < S ? s : 0;
          a[i] = ld_nc_f32(A + sc);
          bb[i] = ld_nc_f32(Bt + (S - 1 - sc));
          l[i] = ld_nc_f32(lp + ((sc & 1) ? 1 + (sc >> 1) : 0));
        }
#pragma unroll
        for (int i = 0; i < kGammaU; ++i) {
          const int s = base + lane + 32 * i;
          if (s < S) {
            const float u = a[i] + bb[i] - l[i];
            gam[s] = u;
            m = fmaxf(m, u);
          }
        }
      }
      m = warp_max(m);
      float w_all = 0.f, w_blank = 0.f;
      for (int s = lane; s < S; s += 32) {
        const float w = ex2f(gam[s] - m);
        gam[s] = w;
        w_all += w;
        if (!(s & 1)) w_blank += w;
      }
      w_all = warp_sum(w_all);
      w_blank = warp_sum(w_blank);
      const float inv = 1.0f / w_all;
      __syncwarp();
      // corr[k] = softmax - occupancy of column k (blank, labels); kCorrSkip marks a later
      // duplicate of a label (its occupancy was merged into the first occurrence)
      float* out = p.ws.occ + row * Kp;
      if (lane == 0) out[0] = ex2f(ld_nc_f32(lp)) - w_blank * inv;
      for (int k = lane; k < L; k += 32) {
        float o = kCorrSkip;
        if (s_fst[k]) {
          o = gam[2 * k + 1];
          for (int j = s_nxt[k]; j >= 0; j = s_nxt[j]) o += gam[2 * j + 1];
          o = ex2f(ld_nc_f32(lp + 1 + k)) - o * inv;
        }
        out[1 + k] = o;
      }
      __syncwarp();
    }
  }
}

}  // namespace b200
