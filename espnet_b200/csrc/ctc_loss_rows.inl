// ctc_loss_rows.inl -- the two streaming kernels of the CTC loss (included by ctc_loss.cu).
//
// Work unit ("task") = kFramesPerTask consecutive frames of one utterance.  Every WARP
// pulls its own tasks from an atomic counter (ragged utterance lengths balance, and no
// CTA-wide barrier is ever needed); the next ticket is fetched while the current task
// runs.  A warp caches its utterance's labels in its own slice of shared memory and owns
// one (utterance, frame) row at a time.  A row is read with coalesced 128-bit
// ld.global.nc loads, kChunkVec vectors per lane in flight, reduced in registers (online
// log-sum-exp across chunks) and -- in the gradient kernel -- written back with 128-bit
// streaming stores.  Up to 24 warps per SM keep >100 KB of loads in flight per SM.
namespace b200 {

constexpr int kRowsMaxWarps = 24;
constexpr int kFramesPerTask = 8;
constexpr int kChunkVec = 10;  // 16-byte vectors per lane per chunk (10*32*16 B = 5 KB per warp)
constexpr int kGather = 5;     // label columns per lane fetched in one batch (covers L+1 <= 160)

template <typename T> __device__ __forceinline__ uint32_t neg_inf_word();
template <> __device__ __forceinline__ uint32_t neg_inf_word<float>() { return 0xff800000u; }
template <> __device__ __forceinline__ uint32_t neg_inf_word<__nv_bfloat16>() { return 0xff80ff80u; }
template <> __device__ __forceinline__ uint32_t neg_inf_word<__half>() { return 0xfc00fc00u; }

// Row log2-sum-exp of x*log2(e): returns l2 such that log2 softmax(x)[v] = x[v]*log2e - l2.
template <typename T, bool VEC>
__device__ __forceinline__ float row_lse2(const T* __restrict__ src, int V, int nvec, int lane, int dbg = 0) {
  constexpr int PER = ElemTraits<T>::kPerVec;
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
      float m = -3.0e38f;
#pragma unroll
      for (int i = 0; i < kChunkVec; ++i) {
        float f[PER];
        unpack16<T>(q[i], f);
#pragma unroll
        for (int e = 0; e < PER; ++e) m = fmaxf(m, f[e]);
      }
      m = fmaxf(warp_max(m), m_run);
      const float ml = m * B200_LOG2E;
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < kChunkVec; ++i) {
        float f[PER];
        unpack16<T>(q[i], f);
#pragma unroll
        for (int e = 0; e < PER; ++e) s += (dbg & 1) ? fmaf(f[e], B200_LOG2E, -ml) : ex2f(fmaf(f[e], B200_LOG2E, -ml));
      }
      s = warp_sum(s);
      s_run = s_run * ex2f((m_run - m) * B200_LOG2E) + s;
      m_run = m;
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
  }
  return m_run * B200_LOG2E + lg2f(s_run);
}

__device__ __forceinline__ int fetch_ticket(int* counter, int lane) {
  int t = 0;
  if (lane == 0) t = atomicAdd(counter, 1);
  return __shfl_sync(0xffffffffu, t, 0);
}

template <typename T, bool VEC, bool GRAD>
__global__ void __launch_bounds__(kRowsMaxWarps * 32, 1) ctc_rows_kernel(LossParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t Kp = p.ws.Kp;
  int* s_lab = reinterpret_cast<int*>(smem) + (size_t)warp * Kp;  // per-warp label cache

  const T* logits = reinterpret_cast<const T*>(p.logits);
  T* grad = reinterpret_cast<T*>(p.grad);
  const int V = (int)p.V;
  const int nvec = (int)((int64_t)V * sizeof(T) / 16);
  constexpr int PER = ElemTraits<T>::kPerVec;
  const int nblk = (int)((p.T + kFramesPerTask - 1) / kFramesPerTask);
  const int ntask = (p.b1 - p.b0) * nblk;
  int* counter = p.ws.bad + (GRAD ? 2 : 1);
  int cached_b = -1;

  int q = fetch_ticket(counter, lane);
  while (q < ntask) {
    const int q_next = fetch_ticket(counter, lane);  // overlaps with this task
    const int b = p.b0 + q / nblk;
    const int t0 = (q % nblk) * kFramesPerTask;
    const int t1 = min(t0 + kFramesPerTask, (int)p.T);
    const int Tb = (int)p.ilen[b];
    const int L = (int)p.olen[b];
    q = q_next;
    if (!GRAD && t0 >= Tb) continue;
    if (b != cached_b && t0 < Tb) {
      __syncwarp();
      const int* lab = p.ws.labels + (int64_t)b * Kp;
      for (int k = lane; k < L; k += 32) s_lab[k] = lab[k];
      cached_b = b;
      __syncwarp();
    }

    float wgt = 0.f;
    if (GRAD) {
      wgt = p.grad_scale * (p.grad_weight ? p.grad_weight[b] : 1.0f);
      if (isinf(p.ws.nll_ws[b])) wgt = __int_as_float(0x7fc00000);
    }

    for (int t = t0; t < t1; ++t) {
      const int64_t row = (int64_t)b * p.T + t;
      const T* src = logits + (int64_t)b * p.stride_b + (int64_t)t * p.stride_t;
      // column of slot k: blank for k == 0, label k-1 otherwise (clamped when k > L)
      int col[kGather];
#pragma unroll
      for (int i = 0; i < kGather; ++i) {
        const int k = lane + 32 * i;
        col[i] = (k >= 1 && k <= L) ? s_lab[k - 1] : p.blank;
      }
      if (!GRAD) {
        if (t >= Tb) break;
        // gather loads are issued first (raw, unconverted) so they overlap with the row stream
        uint32_t raw[kGather];
#pragma unroll
        for (int i = 0; i < kGather; ++i) raw[i] = (p.debug & 2) ? 0u : ld_nc_raw(src + col[i]);
        const float l2 = row_lse2<T, VEC>(src, V, nvec, lane, p.debug);
        if (lane == 0) p.ws.lse2[row] = l2;
        float* lpo = p.ws.lpg + row * Kp;
#pragma unroll
        for (int i = 0; i < kGather; ++i) {
          const int k = lane + 32 * i;
          if (k <= L && !(p.debug & 4)) lpo[k] = fmaf(raw_to_float<T>(raw[i]), B200_LOG2E, -l2);
        }
        for (int k = lane + 32 * kGather; k <= L; k += 32)
          lpo[k] = fmaf(ElemTraits<T>::to_float(src[s_lab[k - 1]]), B200_LOG2E, -l2);
      } else {
        T* dst = grad + (int64_t)b * p.gstride_b + (int64_t)t * p.gstride_t;
        if (t >= Tb) {
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
        // prologue loads (one batch): row log-sum-exp, label logits, occupancies
        const float* occ = p.ws.occ + row * Kp;
        const float l2 = ld_nc_f32(p.ws.lse2 + row);
        uint32_t raw[kGather];
        float oc[kGather];
#pragma unroll
        for (int i = 0; i < kGather; ++i) {
          const int k = lane + 32 * i;
          raw[i] = ld_nc_raw(src + col[i]);
          oc[i] = ld_nc_f32(occ + (k <= L ? k : 0));
        }
        // softmax of the whole row, streamed
        if (VEC) {
          const uint4* src4 = reinterpret_cast<const uint4*>(src);
          uint4* d4 = reinterpret_cast<uint4*>(dst);
          for (int c0 = 0; c0 < nvec; c0 += 32 * kChunkVec) {
            uint4 qv[kChunkVec];
#pragma unroll
            for (int i = 0; i < kChunkVec; ++i) {
              const int j = c0 + lane + 32 * i;
              if (j < nvec) qv[i] = ld_nc_v4(src4 + j);
            }
#pragma unroll
            for (int i = 0; i < kChunkVec; ++i) {
              const int j = c0 + lane + 32 * i;
              if (j < nvec) {
                float f[PER];
                unpack16<T>(qv[i], f);
#pragma unroll
                for (int e = 0; e < PER; ++e) f[e] = wgt * ex2f(fmaf(f[e], B200_LOG2E, -l2));
                st_cs_v4(d4 + j, pack16<T>(f));
              }
            }
          }
        } else {
          for (int v = lane; v < V; v += 32)
            dst[v] = ElemTraits<T>::from_float(
                wgt * ex2f(fmaf(ElemTraits<T>::to_float(src[v]), B200_LOG2E, -l2)));
        }
        __syncwarp();
        // blank and label columns: softmax - occupancy (occ < 0 marks a merged duplicate)
#pragma unroll
        for (int i = 0; i < kGather; ++i) {
          const int k = lane + 32 * i;
          if (k <= L && oc[i] >= 0.f)
            dst[col[i]] = ElemTraits<T>::from_float(
                wgt * (ex2f(fmaf(raw_to_float<T>(raw[i]), B200_LOG2E, -l2)) - oc[i]));
        }
        for (int k = lane + 32 * kGather; k <= L; k += 32) {
          const float o = occ[k];
          const int v = s_lab[k - 1];
          if (o >= 0.f)
            dst[v] = ElemTraits<T>::from_float(
                wgt * (ex2f(fmaf(ElemTraits<T>::to_float(src[v]), B200_LOG2E, -l2)) - o));
        }
        __syncwarp();
      }
    }
  }
}

}  // namespace b200
