// ctc_loss_rows.inl -- the two streaming kernels of the CTC loss (included by ctc_loss.cu).
//
// Both walk the (utterance, frame) rows of the logits with one warp per row.  A row
// (V elements, contiguous) is staged in shared memory by ONE 1-D bulk copy
// (cp.async.bulk, SASS UBLKCP) that signals an mbarrier; each warp keeps a ring of
// kStages rows so that (kStages-1) rows are in flight while one is being processed.
// The gradient kernel rewrites the staged row in place and returns it with a bulk
// shared->global copy, so the logits are read once and the gradient written once.
namespace b200 {

template <typename T>
__device__ __forceinline__ T neg_inf_elem();
template <> __device__ __forceinline__ float neg_inf_elem<float>() { return __int_as_float(0xff800000); }
template <> __device__ __forceinline__ __nv_bfloat16 neg_inf_elem<__nv_bfloat16>() {
  return __ushort_as_bfloat16((unsigned short)0xff80);
}
template <> __device__ __forceinline__ __half neg_inf_elem<__half>() {
  return __ushort_as_half((unsigned short)0xfc00);
}

struct RowIter {
  int64_t r, R, NW;
  int T, b0;
  const int64_t* ilen;
  int b, t;
  // advance to the first row >= r (stepping by NW) whose frame is inside the utterance
  __device__ __forceinline__ void settle() {
    while (r < R) {
      b = b0 + (int)(r / T);
      t = (int)(r - (int64_t)(b - b0) * T);
      if (t < (int)ilen[b]) return;
      r += NW;
    }
  }
  __device__ __forceinline__ bool valid() const { return r < R; }
  __device__ __forceinline__ void next() { r += NW; settle(); }
};

template <typename T, bool BULK, bool GRAD>
__global__ void __launch_bounds__(256, 1) ctc_rows_kernel(LossParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = blockDim.x >> 5;
  const int NS = p.stages;
  const int row_bytes = p.row_bytes;
  const int64_t Kp = p.ws.Kp, Sp = p.ws.Sp;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem) + warp * NS;
  const int bar_bytes = (nwarps * NS * 8 + 127) & ~127;
  const int scratch_bytes = GRAD ? (int)((Sp + Kp) * sizeof(float)) : 0;
  const int per_warp = NS * row_bytes + scratch_bytes;
  unsigned char* wbase = smem + bar_bytes + (size_t)warp * per_warp;
  float* gam = reinterpret_cast<float*>(wbase + NS * row_bytes);
  float* cv = gam + Sp;

  if (BULK) {
    if (lane == 0) {
      for (int s = 0; s < NS; ++s) mbar_init(&bars[s], 1);
      mbar_fence_init();
    }
    __syncwarp();
  }

  const T* logits = reinterpret_cast<const T*>(p.logits);
  T* grad = reinterpret_cast<T*>(p.grad);
  const uint32_t copy_bytes = (uint32_t)(p.V * sizeof(T));
  const int vpad = row_bytes / (int)sizeof(T);
  const int nvec = row_bytes / 16;
  constexpr int PER = ElemTraits<T>::kPerVec;

  RowIter it_load, it_proc;
  it_load.R = (int64_t)(p.b1 - p.b0) * p.T;
  it_load.NW = (int64_t)gridDim.x * nwarps;
  it_load.T = (int)p.T;
  it_load.b0 = p.b0;
  it_load.ilen = p.ilen;
  it_load.r = (int64_t)blockIdx.x * nwarps + warp;
  it_load.b = 0; it_load.t = 0;
  it_load.settle();
  it_proc = it_load;

  auto issue = [&](const RowIter& it, int stage) {
    const T* src = logits + (int64_t)it.b * p.stride_b + (int64_t)it.t * p.stride_t;
    unsigned char* dst = wbase + (size_t)stage * row_bytes;
    if (BULK) {
      if (lane == 0) {
        mbar_arrive_expect_tx(&bars[stage], copy_bytes);
        bulk_g2s(dst, src, copy_bytes, &bars[stage]);
      }
    } else {
      T* d = reinterpret_cast<T*>(dst);
      for (int v = lane; v < vpad; v += 32) d[v] = v < p.V ? src[v] : neg_inf_elem<T>();
      __syncwarp();
    }
  };

  // prologue: NS-1 rows in flight
  for (int s = 0; s < NS - 1; ++s) {
    if (it_load.valid()) { issue(it_load, s); it_load.next(); }
  }

  for (int i = 0; it_proc.valid(); ++i, it_proc.next()) {
    const int stage = i % NS;
    const int b = it_proc.b, t = it_proc.t;
    const int64_t row = (int64_t)b * p.T + t;
    const int L = (int)p.olen[b];
    const int* lab = p.ws.labels + (int64_t)b * Kp;
    T* buf = reinterpret_cast<T*>(wbase + (size_t)stage * row_bytes);
    const uint4* buf4 = reinterpret_cast<const uint4*>(buf);

    float inv = 0.f, wb = 0.f;
    if (GRAD) {
      // occupancies gamma_t(s) ~ 2^(alpha + beta - lp), normalised over the row
      const int S = 2 * L + 1;
      const float* A = p.ws.lat + row * Sp;
      const float* Bt = p.ws.lat + ((int64_t)p.B * p.T + row) * Sp;
      const float* lp = p.ws.lpg + row * Kp;
      float m = -3.0e38f;
      for (int s = lane; s < S; s += 32) {
        const float u = A[s] + Bt[S - 1 - s] - ((s & 1) ? lp[1 + (s >> 1)] : lp[0]);
        gam[s] = u;
        m = fmaxf(m, u);
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
      wb = warp_sum(w_blank);
      inv = 1.0f / w_all;
      __syncwarp();
    }

    if (BULK) mbar_wait(&bars[stage], (uint32_t)((i / NS) & 1));

    if (!GRAD) {
      float m = -3.0e38f;
      for (int j = lane; j < nvec; j += 32) {
        float f[PER];
        unpack16<T>(buf4[j], f);
#pragma unroll
        for (int e = 0; e < PER; ++e) m = fmaxf(m, f[e]);
      }
      m = warp_max(m);
      const float ml = m * B200_LOG2E;
      float s = 0.f;
      for (int j = lane; j < nvec; j += 32) {
        float f[PER];
        unpack16<T>(buf4[j], f);
#pragma unroll
        for (int e = 0; e < PER; ++e) s += ex2f(fmaf(f[e], B200_LOG2E, -ml));
      }
      s = warp_sum(s);
      const float l2 = ml + lg2f(s);
      if (lane == 0) p.ws.lse2[row] = l2;
      float* lpo = p.ws.lpg + row * Kp;
      for (int k = lane; k <= L; k += 32) {
        const int v = k ? lab[k - 1] : p.blank;
        lpo[k] = fmaf(ElemTraits<T>::to_float(buf[v]), B200_LOG2E, -l2);
      }
      __syncwarp();
    } else {
      float wgt = p.grad_scale * (p.grad_weight ? p.grad_weight[b] : 1.0f);
      if (isinf(p.ws.nll_ws[b])) wgt = __int_as_float(0x7fc00000);
      const float l2 = p.ws.lse2[row];
      const int* nxt = p.ws.dup_next + (int64_t)b * Kp;
      const int* fst = p.ws.is_first + (int64_t)b * Kp;
      // values of the blank and label columns, computed from the untouched row
      for (int k = lane; k < L; k += 32) {
        if (fst[k]) {
          float occ = gam[2 * k + 1];
          for (int j = nxt[k]; j >= 0; j = nxt[j]) occ += gam[2 * j + 1];
          const float xl = ElemTraits<T>::to_float(buf[lab[k]]);
          cv[k] = wgt * (ex2f(fmaf(xl, B200_LOG2E, -l2)) - occ * inv);
        }
      }
      float cvb = 0.f;
      if (lane == 0)
        cvb = wgt * (ex2f(fmaf(ElemTraits<T>::to_float(buf[p.blank]), B200_LOG2E, -l2)) - wb * inv);
      __syncwarp();
      // softmax of the whole row, in place
      uint4* out4 = reinterpret_cast<uint4*>(buf);
      for (int j = lane; j < nvec; j += 32) {
        float f[PER];
        unpack16<T>(buf4[j], f);
#pragma unroll
        for (int e = 0; e < PER; ++e) f[e] = wgt * ex2f(fmaf(f[e], B200_LOG2E, -l2));
        out4[j] = pack16<T>(f);
      }
      __syncwarp();
      for (int k = lane; k < L; k += 32)
        if (fst[k]) buf[lab[k]] = ElemTraits<T>::from_float(cv[k]);
      if (lane == 0) buf[p.blank] = ElemTraits<T>::from_float(cvb);
      T* dst = grad + (int64_t)b * p.gstride_b + (int64_t)t * p.gstride_t;
      if (BULK) {
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          bulk_s2g(dst, buf, copy_bytes);
          bulk_commit();
          bulk_wait_read<1>();  // the previous row's store has left its stage
        }
      } else {
        __syncwarp();
        for (int v = lane; v < p.V; v += 32) dst[v] = buf[v];
      }
      __syncwarp();
    }

    // refill the stage processed one iteration ago
    if (it_load.valid()) { issue(it_load, (i + NS - 1) % NS); it_load.next(); }
  }
  if (GRAD && BULK && lane == 0) bulk_wait_all<0>();

  if (GRAD) {
    // frames past the end of each utterance get a zero gradient
    const int64_t R = (int64_t)(p.b1 - p.b0) * p.T;
    const int64_t NW = (int64_t)gridDim.x * nwarps;
    for (int64_t r = (int64_t)blockIdx.x * nwarps + warp; r < R; r += NW) {
      const int b = p.b0 + (int)(r / p.T);
      const int t = (int)(r - (int64_t)(b - p.b0) * p.T);
      if (t < (int)p.ilen[b]) continue;
      T* dst = grad + (int64_t)b * p.gstride_b + (int64_t)t * p.gstride_t;
      if (BULK) {
        uint4* d4 = reinterpret_cast<uint4*>(dst);
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (int j = lane; j < nvec; j += 32) st_cs_v4(d4 + j, z);
      } else {
        for (int v = lane; v < p.V; v += 32) dst[v] = ElemTraits<T>::from_float(0.f);
      }
    }
  }
}

}  // namespace b200
