// ctc_loss_rows.inl -- the two streaming kernels of the CTC loss (included by ctc_loss.cu).
//
// Work unit ("task") = kFramesPerTask consecutive frames of one utterance; CTAs pull
// tasks from an atomic counter so that ragged utterance lengths balance.  Per task the
// utterance's labels are cached in shared memory; inside a task one warp owns one
// (utterance, frame) row at a time.  A row is read with coalesced 128-bit
// ld.global.nc loads, kChunkVec vectors per lane in flight, reduced in registers
// (online log-sum-exp across chunks), and -- in the gradient kernel -- written back with
// 128-bit streaming stores.  16 warps per SM keep >100 KB of loads in flight per SM.
namespace b200 {

constexpr int kRowsWarps = 16;
constexpr int kFramesPerTask = 32;
constexpr int kChunkVec = 20;  // 16-byte vectors per lane per chunk (20*32*16 B = 10 KB per warp)

template <typename T> __device__ __forceinline__ uint32_t neg_inf_word();
template <> __device__ __forceinline__ uint32_t neg_inf_word<float>() { return 0xff800000u; }
template <> __device__ __forceinline__ uint32_t neg_inf_word<__nv_bfloat16>() { return 0xff80ff80u; }
template <> __device__ __forceinline__ uint32_t neg_inf_word<__half>() { return 0xfc00fc00u; }

struct RowTask {
  int b, t0, t1;  // frames [t0, t1) of utterance b
};

// Row log2-sum-exp of x*log2(e): returns l2 such that log2 softmax(x)[v] = x[v]*log2e - l2.
template <typename T, bool VEC>
__device__ __forceinline__ float row_lse2(const T* __restrict__ src, int V, int nvec, int lane) {
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
        for (int e = 0; e < PER; ++e) s += ex2f(fmaf(f[e], B200_LOG2E, -ml));
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

template <typename T, bool VEC, bool GRAD>
__global__ void __launch_bounds__(kRowsWarps * 32, 1) ctc_rows_kernel(LossParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t Kp = p.ws.Kp, Sp = p.ws.Sp;
  int* s_lab = reinterpret_cast<int*>(smem);              // [Kp]
  int* s_nxt = s_lab + Kp;                                // [Kp] (GRAD)
  int* s_fst = s_nxt + Kp;                                // [Kp] (GRAD)
  float* gam = reinterpret_cast<float*>(s_fst + Kp) + (size_t)warp * Sp;  // [Sp] per warp (GRAD)
  __shared__ int s_task;

  const T* logits = reinterpret_cast<const T*>(p.logits);
  T* grad = reinterpret_cast<T*>(p.grad);
  const int V = (int)p.V;
  const int nvec = (int)((int64_t)V * sizeof(T) / 16);
  constexpr int PER = ElemTraits<T>::kPerVec;
  const int nblk = (int)((p.T + kFramesPerTask - 1) / kFramesPerTask);
  const int ntask = (p.b1 - p.b0) * nblk;
  int* counter = p.ws.bad + (GRAD ? 2 : 1);
  int cached_b = -1;

  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) s_task = atomicAdd(counter, 1);
    __syncthreads();
    const int q = s_task;
    if (q >= ntask) break;
    const int b = p.b0 + q / nblk;
    const int t0 = (q % nblk) * kFramesPerTask;
    const int t1 = min(t0 + kFramesPerTask, (int)p.T);
    const int Tb = (int)p.ilen[b];
    const int L = (int)p.olen[b];
    if (!GRAD && t0 >= Tb) continue;
    if (b != cached_b && t0 < Tb) {
      const int* lab = p.ws.labels + (int64_t)b * Kp;
      for (int k = threadIdx.x; k < L; k += blockDim.x) {
        s_lab[k] = lab[k];
        if (GRAD) {
          s_nxt[k] = p.ws.dup_next[(int64_t)b * Kp + k];
          s_fst[k] = p.ws.is_first[(int64_t)b * Kp + k];
        }
      }
      cached_b = b;
      __syncthreads();
    }

    float wgt = 0.f;
    if (GRAD) {
      wgt = p.grad_scale * (p.grad_weight ? p.grad_weight[b] : 1.0f);
      if (isinf(p.ws.nll_ws[b])) wgt = __int_as_float(0x7fc00000);
    }

    for (int t = t0 + warp; t < t1; t += kRowsWarps) {
      const int64_t row = (int64_t)b * p.T + t;
      const T* src = logits + (int64_t)b * p.stride_b + (int64_t)t * p.stride_t;
      if (!GRAD) {
        if (t >= Tb) break;
        // gather loads first: they overlap with the row stream
        float xg[4];
        const int ng = (L + 1 + 31) / 32;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int k = lane + 32 * i;
          xg[i] = (k <= L) ? ElemTraits<T>::to_float(src[k ? s_lab[k - 1] : p.blank]) : 0.f;
        }
        const float l2 = row_lse2<T, VEC>(src, V, nvec, lane);
        if (lane == 0) p.ws.lse2[row] = l2;
        float* lpo = p.ws.lpg + row * Kp;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int k = lane + 32 * i;
          if (k <= L) lpo[k] = fmaf(xg[i], B200_LOG2E, -l2);
        }
        for (int i = 4; i < ng; ++i) {
          const int k = lane + 32 * i;
          if (k <= L) lpo[k] = fmaf(ElemTraits<T>::to_float(src[s_lab[k - 1]]), B200_LOG2E, -l2);
        }
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
        // occupancies gamma_t(s) ~ 2^(alpha + beta - lp), normalised over the row
        const int S = 2 * L + 1;
        const float* A = p.ws.lat + row * Sp;
        const float* Bt = p.ws.lat + ((int64_t)p.B * p.T + row) * Sp;
        const float* lp = p.ws.lpg + row * Kp;
        const float l2 = p.ws.lse2[row];
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
        w_blank = warp_sum(w_blank);
        const float inv = 1.0f / w_all;
        __syncwarp();

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
        // blank and label columns: softmax - occupancy (duplicates merged in label order)
        for (int k = lane; k < L; k += 32) {
          if (s_fst[k]) {
            float occ = gam[2 * k + 1];
            for (int j = s_nxt[k]; j >= 0; j = s_nxt[j]) occ += gam[2 * j + 1];
            const int v = s_lab[k];
            const float xl = ElemTraits<T>::to_float(src[v]);
            dst[v] = ElemTraits<T>::from_float(wgt * (ex2f(fmaf(xl, B200_LOG2E, -l2)) - occ * inv));
          }
        }
        if (lane == 0) {
          const float xb = ElemTraits<T>::to_float(src[p.blank]);
          dst[p.blank] = ElemTraits<T>::from_float(wgt * (ex2f(fmaf(xb, B200_LOG2E, -l2)) - w_blank * inv));
        }
        __syncwarp();
      }
    }
  }
}

}  // namespace b200
