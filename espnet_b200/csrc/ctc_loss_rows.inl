// ctc_loss_rows.inl -- the two streaming kernels of the CTC loss (included by ctc_loss.cu).
//
// Work unit ("task") = kFramesPerTask consecutive frames of one utterance.  Every WARP pulls
// its own tasks from an atomic counter (ragged utterance lengths balance, no CTA-wide
// barrier).  A warp caches its utterance's labels in its slice of shared memory and owns one
// (utterance, frame) row at a time.  A row is read in chunks of kChunkVec 16-byte vectors per
// lane with coalesced ld.global.nc loads (5 KB in flight per warp).
//
//   rowstats: max (packed bf16/fp16 max for half types) and sum-exp in registers, online
//             across chunks; each chunk is also stashed in the warp's shared-memory buffer so
//             that the blank/label columns are gathered from shared memory, not from HBM.
//   grad:     softmax * weight (the weight is folded into the exponent), packed to the output
//             dtype into the warp's shared-memory buffer, blank/label columns patched there
//             with the pre-computed (softmax - occupancy), and the chunk leaves with ONE bulk
//             shared->global copy (cp.async.bulk, SASS UBLKCP) per 5 KB.
namespace b200 {

constexpr int kRowsMaxWarps = 24;
constexpr int kGradMaxWarps = 16;
constexpr int kFramesPerTask = 8;
constexpr int kChunkVec = 10;              // 16-byte vectors per lane per chunk
constexpr int kChunkBytes = kChunkVec * 32 * 16;
constexpr int kGather = 5;                 // label slots per lane held in registers (L+1 <= 160)

template <typename T> __device__ __forceinline__ uint32_t neg_inf_word();
template <> __device__ __forceinline__ uint32_t neg_inf_word<float>() { return 0xff800000u; }
template <> __device__ __forceinline__ uint32_t neg_inf_word<__nv_bfloat16>() { return 0xff80ff80u; }
template <> __device__ __forceinline__ uint32_t neg_inf_word<__half>() { return 0xfc00fc00u; }

// max over the elements of a chunk held in registers
template <typename T>
__device__ __forceinline__ float chunk_max(const uint4 (&q)[kChunkVec]);
template <>
__device__ __forceinline__ float chunk_max<float>(const uint4 (&q)[kChunkVec]) {
  float m0 = -3.0e38f, m1 = -3.0e38f;
#pragma unroll
  for (int i = 0; i < kChunkVec; ++i) {
    m0 = fmaxf(m0, fmaxf(__uint_as_float(q[i].x), __uint_as_float(q[i].y)));
    m1 = fmaxf(m1, fmaxf(__uint_as_float(q[i].z), __uint_as_float(q[i].w)));
  }
  return fmaxf(m0, m1);
}
template <>
__device__ __forceinline__ float chunk_max<__nv_bfloat16>(const uint4 (&q)[kChunkVec]) {
  __nv_bfloat162 m0 = *reinterpret_cast<const __nv_bfloat162*>(&q[0].x);
  __nv_bfloat162 m1 = *reinterpret_cast<const __nv_bfloat162*>(&q[0].y);
#pragma unroll
  for (int i = 0; i < kChunkVec; ++i) {
    m0 = __hmax2(m0, *reinterpret_cast<const __nv_bfloat162*>(&q[i].x));
    m1 = __hmax2(m1, *reinterpret_cast<const __nv_bfloat162*>(&q[i].y));
    m0 = __hmax2(m0, *reinterpret_cast<const __nv_bfloat162*>(&q[i].z));
    m1 = __hmax2(m1, *reinterpret_cast<const __nv_bfloat162*>(&q[i].w));
  }
  m0 = __hmax2(m0, m1);
  return fmaxf(__low2float(m0), __high2float(m0));
}
template <>
__device__ __forceinline__ float chunk_max<__half>(const uint4 (&q)[kChunkVec]) {
  __half2 m0 = *reinterpret_cast<const __half2*>(&q[0].x);
  __half2 m1 = *reinterpret_cast<const __half2*>(&q[0].y);
#pragma unroll
  for (int i = 0; i < kChunkVec; ++i) {
    m0 = __hmax2(m0, *reinterpret_cast<const __half2*>(&q[i].x));
    m1 = __hmax2(m1, *reinterpret_cast<const __half2*>(&q[i].y));
    m0 = __hmax2(m0, *reinterpret_cast<const __half2*>(&q[i].z));
    m1 = __hmax2(m1, *reinterpret_cast<const __half2*>(&q[i].w));
  }
  m0 = __hmax2(m0, m1);
  return fmaxf(__low2float(m0), __high2float(m0));
}

__device__ __forceinline__ int fetch_ticket(int* counter, int lane) {
  int t = 0;
  if (lane == 0) t = atomicAdd(counter, 1);
  return __shfl_sync(0xffffffffu, t, 0);
}

struct RowTask {
  int b, t0, t1, Tb, L;
};

// Pulls the next task; returns false when the queue is exhausted.
__device__ __forceinline__ bool next_task(const LossParams& p, int* counter, int lane, int nblk, int ntask,
                                          RowTask* k) {
  const int q = fetch_ticket(counter, lane);
  if (q >= ntask) return false;
  k->b = p.b0 + q / nblk;
  k->t0 = (q % nblk) * kFramesPerTask;
  k->t1 = min(k->t0 + kFramesPerTask, (int)p.T);
  k->Tb = (int)p.ilen[k->b];
  k->L = utt_labels(p, k->b);
  return true;
}

// ---------------------------------------------------------------------------- rowstats
template <typename T, bool VEC>
__global__ void __launch_bounds__(kRowsMaxWarps * 32, 1) ctc_rowstats_kernel(LossParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t Kp = p.ws.Kp;
  const size_t per_warp = Kp * sizeof(int) + (VEC ? kChunkBytes : 0);
  int* s_lab = reinterpret_cast<int*>(smem + warp * per_warp);
  uint4* buf4 = reinterpret_cast<uint4*>(s_lab + Kp);
  const T* bufT = reinterpret_cast<const T*>(buf4);

  const T* logits = reinterpret_cast<const T*>(p.logits);
  const int V = (int)p.V;
  const int nvec = (int)((int64_t)V * sizeof(T) / 16);
  constexpr int PER = ElemTraits<T>::kPerVec;
  constexpr int CHUNK_ELEMS = kChunkVec * 32 * PER;
  const int nblk = (int)((p.T + kFramesPerTask - 1) / kFramesPerTask);
  const int ntask = (p.b1 - p.b0) * nblk;
  int* counter = p.ws.bad + 1;
  int cached_b = -1;
  RowTask k;

  while (next_task(p, counter, lane, nblk, ntask, &k)) {
    if (k.t0 >= k.Tb) continue;
    const int L = k.L;
    if (k.b != cached_b) {
      __syncwarp();
      const int* lab = p.ws.labels + (int64_t)k.b * Kp;
      for (int i = lane; i < L; i += 32) s_lab[i] = lab[i];
      cached_b = k.b;
      __syncwarp();
    }
    int col[kGather];
#pragma unroll
    for (int i = 0; i < kGather; ++i) {
      const int kk = lane + 32 * i;
      col[i] = (kk >= 1 && kk <= L) ? s_lab[kk - 1] : p.blank;
    }
    const int t_end = min(k.t1, k.Tb);
    for (int t = k.t0; t < t_end; ++t) {
      const int64_t row = (int64_t)k.b * p.T + t;
      const T* src = logits + (int64_t)k.b * p.stride_b + (int64_t)t * p.stride_t;
      float* lpo = p.ws.lpg + row * Kp;
      float xg[kGather];
#pragma unroll
      for (int i = 0; i < kGather; ++i) xg[i] = 0.f;
      float m_run = -3.0e38f, s_run = 0.f;
      if (VEC) {
        const uint4* src4 = reinterpret_cast<const uint4*>(src);
        for (int c0 = 0; c0 < nvec; c0 += 32 * kChunkVec) {
          uint4 q[kChunkVec];
          const uint32_t ni = neg_inf_word<T>();
#pragma unroll
          for (int i = 0; i < kChunkVec; ++i) {
            const int j = c0 + lane + 32 * i;
            q[i] = j < nvec ? ld_nc_v4(src4 + j) : make_uint4(ni, ni, ni, ni);
          }
          // stash the chunk: the blank/label columns are gathered from shared memory
#pragma unroll
          for (int i = 0; i < kChunkVec; ++i) buf4[lane + 32 * i] = q[i];
          const float m = fmaxf(warp_max(chunk_max<T>(q)), m_run);
          const float ml = m * B200_LOG2E;
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int i = 0; i < kChunkVec; ++i) {
            float f[PER];
            unpack16<T>(q[i], f);
#pragma unroll
            for (int e = 0; e < PER; e += 2) {
              s0 += ex2f(fmaf(f[e], B200_LOG2E, -ml));
              s1 += ex2f(fmaf(f[e + 1], B200_LOG2E, -ml));
            }
          }
          const float s = warp_sum(s0 + s1);
          s_run = s_run * ex2f((m_run - m) * B200_LOG2E) + s;
          m_run = m;
          __syncwarp();
          const int lo = c0 * PER, hi = lo + CHUNK_ELEMS;
#pragma unroll
          for (int i = 0; i < kGather; ++i)
            if (col[i] >= lo && col[i] < hi) xg[i] = ElemTraits<T>::to_float(bufT[col[i] - lo]);
          for (int kk = lane + 32 * kGather; kk <= L; kk += 32) {  // L+1 > 160: spill to lpg directly
            const int v = s_lab[kk - 1];
            if (v >= lo && v < hi) lpo[kk] = ElemTraits<T>::to_float(bufT[v - lo]) * B200_LOG2E;
          }
          __syncwarp();
        }
      } else {
        float m = -3.0e38f;
        for (int v = lane; v < V; v += 32) m = fmaxf(m, ElemTraits<T>::to_float(src[v]));
        m = warp_max(m);
        const float ml = m * B200_LOG2E;
        float s = 0.f;
        for (int v = lane; v < V; v += 32) s += ex2f(fmaf(ElemTraits<T>::to_float(src[v]), B200_LOG2E, -ml));
        s_run = warp_sum(s);
        m_run = m;
#pragma unroll
        for (int i = 0; i < kGather; ++i) xg[i] = ElemTraits<T>::to_float(src[col[i]]);
        for (int kk = lane + 32 * kGather; kk <= L; kk += 32)
          lpo[kk] = ElemTraits<T>::to_float(src[s_lab[kk - 1]]) * B200_LOG2E;
      }
      const float l2 = m_run * B200_LOG2E + lg2f(s_run);
      if (lane == 0) p.ws.lse2[row] = l2;
#pragma unroll
      for (int i = 0; i < kGather; ++i) {
        const int kk = lane + 32 * i;
        if (kk <= L) lpo[kk] = fmaf(xg[i], B200_LOG2E, -l2);
      }
      for (int kk = lane + 32 * kGather; kk <= L; kk += 32) lpo[kk] -= l2;
    }
  }
}

// ---------------------------------------------------------------------------- gradient
template <typename T> __device__ __forceinline__ uint32_t nan_word();
template <> __device__ __forceinline__ uint32_t nan_word<float>() { return 0x7fc00000u; }
template <> __device__ __forceinline__ uint32_t nan_word<__nv_bfloat16>() { return 0x7fc07fc0u; }
template <> __device__ __forceinline__ uint32_t nan_word<__half>() { return 0x7e007e00u; }
template <typename T> __device__ __forceinline__ uint32_t sign_word();
template <> __device__ __forceinline__ uint32_t sign_word<float>() { return 0x80000000u; }
template <> __device__ __forceinline__ uint32_t sign_word<__nv_bfloat16>() { return 0x80008000u; }
template <> __device__ __forceinline__ uint32_t sign_word<__half>() { return 0x80008000u; }

template <typename T, bool VEC>
__global__ void __launch_bounds__(kGradMaxWarps * 32, 1) ctc_grad_kernel(LossParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t Kp = p.ws.Kp;
  const size_t per_warp = Kp * sizeof(int) + (VEC ? 2 * kChunkBytes : 0);
  int* s_lab = reinterpret_cast<int*>(smem + warp * per_warp);
  unsigned char* obuf = reinterpret_cast<unsigned char*>(s_lab + Kp);

  const T* logits = reinterpret_cast<const T*>(p.logits);
  T* grad = reinterpret_cast<T*>(p.grad);
  const int V = (int)p.V;
  const int nvec = (int)((int64_t)V * sizeof(T) / 16);
  constexpr int PER = ElemTraits<T>::kPerVec;
  constexpr int CHUNK_ELEMS = kChunkVec * 32 * PER;
  const int nblk = (int)((p.T + kFramesPerTask - 1) / kFramesPerTask);
  const int ntask = (p.b1 - p.b0) * nblk;
  int* counter = p.ws.bad + 2;
  int cached_b = -1;
  uint32_t nbuf = 0;  // chunks stored so far by this warp (selects the output buffer)
  RowTask k;

  while (next_task(p, counter, lane, nblk, ntask, &k)) {
    const int L = k.L;
    if (k.b != cached_b && k.t0 < k.Tb) {
      __syncwarp();
      const int* lab = p.ws.labels + (int64_t)k.b * Kp;
      for (int i = lane; i < L; i += 32) s_lab[i] = lab[i];
      cached_b = k.b;
      __syncwarp();
    }
    int col[kGather];
#pragma unroll
    for (int i = 0; i < kGather; ++i) {
      const int kk = lane + 32 * i;
      col[i] = (k.t0 < k.Tb && kk >= 1 && kk <= L) ? s_lab[kk - 1] : p.blank;
    }
    const float wgt = p.grad_scale * (p.grad_weight ? p.grad_weight[k.b] : 1.0f);
    const bool poison = !isfinite(p.ws.nll_ws[k.b]) || isnan(wgt);  // infeasible utterance: NaN rows (torch)
    const bool neg = wgt < 0.f;
    const float lw = lg2f(fabsf(wgt));  // weight folded into the exponent; -inf for wgt == 0

    for (int t = k.t0; t < k.t1; ++t) {
      const int64_t row = (int64_t)k.b * p.T + t;
      const T* src = logits + (int64_t)k.b * p.stride_b + (int64_t)t * p.stride_t;
      T* dst = grad + (int64_t)k.b * p.gstride_b + (int64_t)t * p.gstride_t;
      if (t >= k.Tb) {
        // frames past the end of the utterance: zero gradient
        if (VEC) {
          uint4* d4 = reinterpret_cast<uint4*>(dst);
          const uint4 z = make_uint4(0, 0, 0, 0);
          for (int j = lane; j < nvec; j += 32) st_cs_v4(d4 + j, z);
        } else {
          for (int v = lane; v < V; v += 32) dst[v] = ElemTraits<T>::from_float(0.f);
        }
        continue;
      }
      const float* corr = p.ws.occ + row * Kp;
      const float l2w = ld_nc_f32(p.ws.lse2 + row) - lw;
      float cr[kGather];
#pragma unroll
      for (int i = 0; i < kGather; ++i) {
        const int kk = lane + 32 * i;
        cr[i] = ld_nc_f32(corr + (kk <= L ? kk : 0));
      }
      if (VEC) {
        const uint4* src4 = reinterpret_cast<const uint4*>(src);
        for (int c0 = 0; c0 < nvec; c0 += 32 * kChunkVec) {
          unsigned char* buf = obuf + (nbuf & 1) * kChunkBytes;
          ++nbuf;
          uint4 q[kChunkVec];
#pragma unroll
          for (int i = 0; i < kChunkVec; ++i) {
            const int j = c0 + lane + 32 * i;
            q[i] = j < nvec ? ld_nc_v4(src4 + j) : make_uint4(0, 0, 0, 0);
          }
          // the bulk store that last used this buffer (two chunks ago) must have read it
          if (lane == 0) bulk_wait_read<1>();
          __syncwarp();
          uint4* b4 = reinterpret_cast<uint4*>(buf);
#pragma unroll
          for (int i = 0; i < kChunkVec; ++i) {
            float f[PER];
            unpack16<T>(q[i], f);
#pragma unroll
            for (int e = 0; e < PER; ++e) f[e] = ex2f(fmaf(f[e], B200_LOG2E, -l2w));
            uint4 o = pack16<T>(f);
            if (neg) { const uint32_t sw = sign_word<T>(); o.x ^= sw; o.y ^= sw; o.z ^= sw; o.w ^= sw; }
            if (poison) { const uint32_t nw = nan_word<T>(); o = make_uint4(nw, nw, nw, nw); }
            b4[lane + 32 * i] = o;
          }
          __syncwarp();
          // blank and label columns: weight * (softmax - occupancy), duplicates already merged
          const int lo = c0 * PER, hi = lo + CHUNK_ELEMS;
          T* bT = reinterpret_cast<T*>(buf);
          if (!poison) {
#pragma unroll
            for (int i = 0; i < kGather; ++i) {
              const int kk = lane + 32 * i;
              if (kk <= L && cr[i] < 1.5f && col[i] >= lo && col[i] < hi)
                bT[col[i] - lo] = ElemTraits<T>::from_float(wgt * cr[i]);
            }
            for (int kk = lane + 32 * kGather; kk <= L; kk += 32) {
              const int v = s_lab[kk - 1];
              const float c = corr[kk];
              if (c < 1.5f && v >= lo && v < hi) bT[v - lo] = ElemTraits<T>::from_float(wgt * c);
            }
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            const int bytes = min(kChunkBytes, (nvec - c0) * 16);
            bulk_s2g(reinterpret_cast<unsigned char*>(dst) + (size_t)c0 * 16, buf, (uint32_t)bytes);
            bulk_commit();
          }
        }
      } else {
        const float w = poison ? __int_as_float(0x7fc00000) : wgt;
        const float l2 = p.ws.lse2[row];
        for (int v = lane; v < V; v += 32)
          dst[v] = ElemTraits<T>::from_float(w * ex2f(fmaf(ElemTraits<T>::to_float(src[v]), B200_LOG2E, -l2)));
        __syncwarp();
        for (int kk = lane; kk <= L; kk += 32) {
          const float c = corr[kk];
          const int v = kk ? s_lab[kk - 1] : p.blank;
          if (c < 1.5f) dst[v] = ElemTraits<T>::from_float(w * c);
        }
        __syncwarp();
      }
    }
  }
  if (VEC && lane == 0) bulk_wait_all<0>();
}

}  // namespace b200
