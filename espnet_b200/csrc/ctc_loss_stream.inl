// ctc_loss_stream.inl -- the ONE pass over the logits of the fused forward+backward call
// (included by ctc_loss.cu).
//
// The gradient of the CTC loss w.r.t. the logits is  w * (softmax - occupancy), and the
// occupancy is non-zero only in the blank column and the <= L label columns of a row.  So the
// row pass that computes the log-sum-exp can already write  w * softmax  to EVERY column of the
// gradient; the lattice kernels then only patch L+1 columns per row (ctc_occ_kernel<PATCH>).
// HBM traffic of the whole call: logits read once + gradient written once (+ the lattice).
//
// One persistent CTA per SM: a ring of row slots in shared memory, one producer warp and
// kStreamWarps consumer warps.
//   producer  walks the CTA's rows (row r of the (B,T) grid belongs to CTA r % grid), issues
//             one 1-D bulk copy (cp.async.bulk, SASS UBLKCP, L2 evict-first) per valid row into
//             the next free slot, completion counted on the slot's `full` mbarrier; rows past
//             the end of their utterance are zero-filled directly.
//   consumer  groups of G warps (G = 1, 2, 4 by row size) take the filled slots round-robin:
//             row max; exponentials relative to it, summed in fp32 and written back IN PLACE
//             (fp32 rows as fp32, 16-bit rows as fp16 scaled by 2^14, so each element costs ONE
//             ex2); the blank and label columns gathered into lpg[b][t][0..L] (log2 units, what
//             the sweeps read); then the slot is rescaled to w * softmax in the output type and
//             leaves with one bulk shared->global copy; when that copy has read the slot, the
//             group arrives on the slot's `empty` mbarrier and the producer refills it.
// Bytes in flight per SM are set by the ring (~200 KB), not by registers or occupancy.
#include <type_traits>
namespace b200 {

constexpr int kStreamWarps = 16;
constexpr int kStreamGather = 3;  // label slots per thread held in registers
constexpr int kStreamMaxSlots = 32;
constexpr int kStreamSmemBudget = 200 * 1024;

__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                              uint64_t* bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
          "r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
      : "memory");
}

// running maximum of packed elements
template <typename T> __device__ __forceinline__ void vmax_acc(const uint4& q, uint32_t& m0, uint32_t& m1);
template <> __device__ __forceinline__ void vmax_acc<float>(const uint4& q, uint32_t& m0, uint32_t& m1) {
  m0 = __float_as_uint(fmaxf(__uint_as_float(m0), fmaxf(__uint_as_float(q.x), __uint_as_float(q.y))));
  m1 = __float_as_uint(fmaxf(__uint_as_float(m1), fmaxf(__uint_as_float(q.z), __uint_as_float(q.w))));
}
template <> __device__ __forceinline__ void vmax_acc<__nv_bfloat16>(const uint4& q, uint32_t& m0, uint32_t& m1) {
  __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&m0), b = *reinterpret_cast<__nv_bfloat162*>(&m1);
  a = __hmax2(a, *reinterpret_cast<const __nv_bfloat162*>(&q.x));
  b = __hmax2(b, *reinterpret_cast<const __nv_bfloat162*>(&q.y));
  a = __hmax2(a, *reinterpret_cast<const __nv_bfloat162*>(&q.z));
  b = __hmax2(b, *reinterpret_cast<const __nv_bfloat162*>(&q.w));
  m0 = *reinterpret_cast<uint32_t*>(&a); m1 = *reinterpret_cast<uint32_t*>(&b);
}
template <> __device__ __forceinline__ void vmax_acc<__half>(const uint4& q, uint32_t& m0, uint32_t& m1) {
  __half2 a = *reinterpret_cast<__half2*>(&m0), b = *reinterpret_cast<__half2*>(&m1);
  a = __hmax2(a, *reinterpret_cast<const __half2*>(&q.x));
  b = __hmax2(b, *reinterpret_cast<const __half2*>(&q.y));
  a = __hmax2(a, *reinterpret_cast<const __half2*>(&q.z));
  b = __hmax2(b, *reinterpret_cast<const __half2*>(&q.w));
  m0 = *reinterpret_cast<uint32_t*>(&a); m1 = *reinterpret_cast<uint32_t*>(&b);
}
template <typename T> __device__ __forceinline__ float vmax_final(uint32_t m0, uint32_t m1);
template <> __device__ __forceinline__ float vmax_final<float>(uint32_t m0, uint32_t m1) {
  return fmaxf(__uint_as_float(m0), __uint_as_float(m1));
}
template <> __device__ __forceinline__ float vmax_final<__nv_bfloat16>(uint32_t m0, uint32_t m1) {
  return fmaxf(fmaxf(__uint_as_float(m0 << 16), __uint_as_float(m0 & 0xffff0000u)),
               fmaxf(__uint_as_float(m1 << 16), __uint_as_float(m1 & 0xffff0000u)));
}
template <> __device__ __forceinline__ float vmax_final<__half>(uint32_t m0, uint32_t m1) {
  const __half2 a = *reinterpret_cast<__half2*>(&m0), b = *reinterpret_cast<__half2*>(&m1);
  return fmaxf(fmaxf(__low2float(a), __high2float(a)), fmaxf(__low2float(b), __high2float(b)));
}

struct StreamRow {
  int b, t;
  bool in_range, valid;
};
// lane `lane` looks at entry (base + lane) of this CTA's row list
__device__ __forceinline__ StreamRow stream_row(const LossParams& p, int64_t base, int lane, int64_t nrows) {
  StreamRow r;
  const int64_t g = (int64_t)blockIdx.x + (base + lane) * (int64_t)gridDim.x;
  r.in_range = g < nrows;
  r.b = 0; r.t = 0; r.valid = false;
  if (r.in_range) {
    r.b = p.b0 + (int)(g / p.T);
    r.t = (int)(g % p.T);
    r.valid = r.t < (int)p.ilen[r.b];
  }
  return r;
}

// In-place intermediate of a row: e = 2^(x*log2e - max) as fp32 for fp32 rows, as fp16 * 2^14 for
// 16-bit rows (normal fp16 down to e = 2^-28, denormal to 2^-38).
template <typename T> struct StreamMid;
template <> struct StreamMid<float> {
  static constexpr float kBias = 0.f;
  static constexpr float kUnbias = 1.f;
  __device__ static __forceinline__ uint4 pack(const float* e) { return pack16<float>(e); }
  __device__ static __forceinline__ void unpack(const uint4& q, float* e) { unpack16<float>(q, e); }
};
template <> struct StreamMid<__nv_bfloat16> {
  static constexpr float kBias = 14.f;
  static constexpr float kUnbias = 1.f / 16384.f;
  __device__ static __forceinline__ uint4 pack(const float* e) { return pack16<__half>(e); }
  __device__ static __forceinline__ void unpack(const uint4& q, float* e) { unpack16<__half>(q, e); }
};
template <> struct StreamMid<__half> : StreamMid<__nv_bfloat16> {};

// 2^x for x <= ~1 as a double: fraction through ex2.approx (fp32, 2^-22 relative), integer part through the exponent
// field; below 2^-1021 the result is 0 (what the linear sweeps treat as "impossible")
__device__ __forceinline__ double exp2_scaled(float x) {
  const float e = floorf(x);
  const float m = ex2f(x - e);  // [1, 2)
  if (!(e > -1021.0f)) return 0.0;
  const double d = (double)m;
  return __hiloint2double(__double2hiint(d) + ((int)e << 20), __double2loint(d));
}

__device__ __forceinline__ void group_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

template <typename T, int G>
__global__ void __launch_bounds__((kStreamWarps + 1) * 32, 1) ctc_stream_kernel(LossParams p, int nslots, int slot_bytes, int park_floats) {
  extern __shared__ __align__(128) unsigned char st_smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(st_smem);  // [kStreamMaxSlots]
  uint64_t* empty = full + kStreamMaxSlots;               // [kStreamMaxSlots]
  unsigned char* slots = st_smem + 2 * kStreamMaxSlots * sizeof(uint64_t);
  __shared__ float s_mx[kStreamWarps], s_sm[kStreamWarps], s_gx[kStreamWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int PER = ElemTraits<T>::kPerVec;
  constexpr int GT = 32 * G;               // threads per row group
  constexpr int NGRP = kStreamWarps / G;   // row groups per CTA
  const int V = (int)p.V;
  const int nvec = (int)((int64_t)V * sizeof(T) / 16);
  const uint32_t row_bytes = (uint32_t)nvec * 16u;
  const int64_t nrows = (int64_t)(p.b1 - p.b0) * p.T;
  const int64_t Kp = p.ws.Kp;
  const T* logits = reinterpret_cast<const T*>(p.logits);
  T* grad = reinterpret_cast<T*>(p.grad);

  if (threadIdx.x == 0) {
    for (int s = 0; s < nslots; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_fence_init();
  }
  __syncthreads();
  pdl_wait();  // (launched as a programmatic dependent of whatever precedes it on the stream: labels, logits)
  if (threadIdx.x == 0) TL_MIN(0);
  // This kernel reads the caller's int64 labels itself (clamped exactly as the prep kernel clamps them), so
  // it does not wait for the prep kernel: prep is launched AFTER it as a programmatic dependent and runs
  // on the free thread slots of the SMs while the rows stream.
  pdl_launch_dependents();

  if (warp == kStreamWarps) {
    // ------------------------------------------------------------------ producer
    const uint64_t pol = l2_evict_first_policy();
    int k = 0;  // ordinal of the next valid row of this CTA
    for (int64_t base = 0;; base += 32) {
      const StreamRow r = stream_row(p, base, lane, nrows);
      const unsigned im = __ballot_sync(0xffffffffu, r.in_range);
      const unsigned vm = __ballot_sync(0xffffffffu, r.valid);
      if (im == 0) break;
      unsigned m = vm;
      while (m) {
        const int l = __ffs(m) - 1;
        m &= m - 1;
        const int bb = __shfl_sync(0xffffffffu, r.b, l), tt = __shfl_sync(0xffffffffu, r.t, l);
        if (lane == 0) {
          const int slot = k % nslots, round = k / nslots;
          if (round > 0) mbar_wait(&empty[slot], (uint32_t)((round - 1) & 1));
          mbar_arrive_expect_tx(&full[slot], row_bytes);
          bulk_g2s_hint(slots + (size_t)slot * slot_bytes,
                        logits + (int64_t)bb * p.stride_b + (int64_t)tt * p.stride_t, row_bytes, &full[slot], pol);
        }
        ++k;
      }
      __syncwarp();
      // rows past the end of their utterance: zero gradient
      unsigned zm = im & ~vm;
      while (zm) {
        const int l = __ffs(zm) - 1;
        zm &= zm - 1;
        const int bb = __shfl_sync(0xffffffffu, r.b, l), tt = __shfl_sync(0xffffffffu, r.t, l);
        uint4* d4 = reinterpret_cast<uint4*>(grad + (int64_t)bb * p.gstride_b + (int64_t)tt * p.gstride_t);
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (int j = lane; j < nvec; j += 32) st_cs_v4(d4 + j, z);
      }
    }
    return;
  }

  // -------------------------------------------------------------------- consumers
  const int grp = warp / G, wg = warp % G;
  const int gtid = wg * 32 + lane;
  const int nfull = nvec / GT;
  const bool tail = gtid + GT * nfull < nvec;
  const int bar_id = 1 + grp;
  int k = 0;
  for (int64_t base = 0;; base += 32) {
    const StreamRow r = stream_row(p, base, lane, nrows);
    const unsigned im = __ballot_sync(0xffffffffu, r.in_range);
    unsigned m = __ballot_sync(0xffffffffu, r.valid);
    if (im == 0) break;
    for (; m; m &= m - 1, ++k) {
      if (k % NGRP != grp) continue;
      const int l = __ffs(m) - 1;
      const int b = __shfl_sync(0xffffffffu, r.b, l), t = __shfl_sync(0xffffffffu, r.t, l);
      const int slot = k % nslots, round = k / nslots;
      unsigned char* sp = slots + (size_t)slot * slot_bytes;
      uint4* s4 = reinterpret_cast<uint4*>(sp);
      const T* sT = reinterpret_cast<const T*>(sp);
      const int64_t row = (int64_t)b * p.T + t;
      const int L = utt_labels(p, b);
      const int64_t* tgt = p.targets + p.tgt_offsets[b];
      auto label_at = [&](int k) -> int {
        const int64_t v = tgt[k];
        return (int)(v < 0 ? 0 : (v >= p.V ? p.V - 1 : v));
      };
      const float wgt = p.grad_scale * (p.grad_weight ? p.grad_weight[b] : 1.0f);
      // columns this thread gathers (global loads issued before the row is waited for)
      int col[kStreamGather];
#pragma unroll
      for (int g = 0; g < kStreamGather; ++g) {
        const int kk = gtid + GT * g;
        col[g] = (kk >= 1 && kk <= L) ? label_at(kk - 1) : p.blank;
      }

      mbar_wait(&full[slot], (uint32_t)(round & 1));

      // raw values of the blank / label columns (the slot is overwritten below)
      float* lpo = p.ws.lpg + row * Kp;
      float xg[kStreamGather];
#pragma unroll
      for (int g = 0; g < kStreamGather; ++g) xg[g] = ElemTraits<T>::to_float(sT[col[g]]);
      // (label positions beyond the register slots: raw value parked in this row group's strip of shared memory
      // behind the ring -- in global memory it was 0.4 GB of extra traffic per C5 step; lpg itself only ever
      // receives FINAL values, because the sweeps running under this kernel treat any non-NaN word of lpg as valid)
      float* park = reinterpret_cast<float*>(slots + (size_t)nslots * slot_bytes) + (size_t)grp * park_floats -
                    GT * kStreamGather;
      // largest raw value among the blank / label columns (for the feed-forward scale of the linear sweeps)
      float gx = -3.0e38f;
#pragma unroll
      for (int g = 0; g < kStreamGather; ++g)
        if (gtid + GT * g <= L) gx = fmaxf(gx, xg[g]);
      // (eight label loads in flight per batch: one load at a time cost an L2 round trip per label position --
      //  13 of them per row and thread at L = 1000, where this kernel ran at 3.0 TB/s against 5.4 at L = 150)
      for (int kk0 = gtid + GT * kStreamGather; kk0 <= L; kk0 += GT * 8) {
        int cc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int kk = kk0 + GT * j;
          cc[j] = kk <= L ? label_at(kk - 1) : p.blank;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int kk = kk0 + GT * j;
          if (kk <= L) {
            const float v = ElemTraits<T>::to_float(sT[cc[j]]);
            gx = fmaxf(gx, v);
            park[kk] = v * B200_LOG2E;
          }
        }
      }
      // Every thread of the group owns the vectors gtid + GT*u; the first nfull of them exist for
      // every thread (no bounds checks in the unrolled loops), at most one more for some.
      auto drive = [&](auto pass) {
        int u = 0;
        for (; u + 4 <= nfull; u += 4) pass(gtid + GT * u, std::integral_constant<int, 4>{});
        for (; u < nfull; ++u) pass(gtid + GT * u, std::integral_constant<int, 1>{});
        if (tail) pass(gtid + GT * nfull, std::integral_constant<int, 1>{});
      };
      // pass 1: row maximum (packed for the 16-bit types)
      uint32_t m0 = neg_inf_word<T>(), m1 = m0;
      drive([&](int j0, auto n_) {
        constexpr int N = decltype(n_)::value;
        uint4 q[N];
#pragma unroll
        for (int u = 0; u < N; ++u) q[u] = s4[j0 + GT * u];
#pragma unroll
        for (int u = 0; u < N; ++u) vmax_acc<T>(q[u], m0, m1);
      });
      float mx = warp_max(vmax_final<T>(m0, m1));
      if (G > 1) {
        if (lane == 0) s_mx[warp] = mx;
        group_sync(bar_id, GT);
#pragma unroll
        for (int g = 0; g < G; ++g) mx = fmaxf(mx, s_mx[grp * G + g]);
      } else {
        __syncwarp();
      }
      const float mlog = mx * B200_LOG2E;
      const float ml = mlog - StreamMid<T>::kBias;
      // pass 2: exponentials, summed in fp32 and kept in place
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
      drive([&](int j0, auto n_) {
        constexpr int N = decltype(n_)::value;
        uint4 q[N];
#pragma unroll
        for (int u = 0; u < N; ++u) q[u] = s4[j0 + GT * u];
#pragma unroll
        for (int u = 0; u < N; ++u) {
          float f[PER];
          unpack16<T>(q[u], f);
#pragma unroll
          for (int e = 0; e < PER; e += 4) {
            f[e] = ex2f(fmaf(f[e], B200_LOG2E, -ml));
            f[e + 1] = ex2f(fmaf(f[e + 1], B200_LOG2E, -ml));
            f[e + 2] = ex2f(fmaf(f[e + 2], B200_LOG2E, -ml));
            f[e + 3] = ex2f(fmaf(f[e + 3], B200_LOG2E, -ml));
            s0 += f[e]; s1 += f[e + 1]; s2 += f[e + 2]; s3 += f[e + 3];
          }
          s4[j0 + GT * u] = StreamMid<T>::pack(f);
        }
      });
      float ssum = warp_sum((s0 + s1) + (s2 + s3));
      gx = warp_max(gx);
      if (G > 1) {
        if (lane == 0) { s_sm[warp] = ssum; s_gx[warp] = gx; }
        group_sync(bar_id, GT);
        ssum = 0.f;
#pragma unroll
        for (int g = 0; g < G; ++g) { ssum += s_sm[grp * G + g]; gx = fmaxf(gx, s_gx[grp * G + g]); }
      } else {
        __syncwarp();
      }
      // row log2-sum-exp.  The bias is taken out of the sum (an exact power of two) BEFORE the logarithm: adding
      // -14 and +14.0000x in fp32 would quantise the result to 1e-6, which a converged model's per-frame
      // log-probability (~ -1e-6) does not survive over 500 frames
      const float l2 = mlog + lg2f(ssum * StreamMid<T>::kUnbias);
      // blank and label columns -> lpg (log2-probabilities)
      // feed-forward scale of the linear sweeps: the largest needed probability of the frame lands in [1, 2)
      const float maxlp = fmaf(gx, B200_LOG2E, -l2);
      const float gf = (p.lin && maxlp > -1.0e29f) ? -floorf(maxlp) : 0.f;
      double* pdo = p.ws.pd + row * Kp;
#pragma unroll
      for (int g = 0; g < kStreamGather; ++g) {
        const int kk = gtid + GT * g;
        if (kk <= L) {
          const float lp = fmaxf(fmaf(xg[g], B200_LOG2E, -l2), B200_NEG);
          st_relaxed_f32(lpo + kk, lp);
          if (p.lin) pdo[kk] = exp2_scaled(lp + gf);
        }
      }
      for (int kk0 = gtid + GT * kStreamGather; kk0 <= L; kk0 += GT * 8) {
        float pv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int kk = kk0 + GT * j;
          pv[j] = kk <= L ? park[kk] : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int kk = kk0 + GT * j;
          if (kk <= L) {
            const float lp = fmaxf(pv[j] - l2, B200_NEG);
            st_relaxed_f32(lpo + kk, lp);
            if (p.lin) pdo[kk] = exp2_scaled(lp + gf);
          }
        }
      }
      if (p.lin && gtid == 0) {
        p.ws.gff[row] = (int)gf;
        st_relaxed_f32(lpo + Kp - 1, gf);
      }
      if (gtid == 0) p.ws.lse2[row] = l2;
      // pass 3: e * (w / sum) in the output type, in place
      const float scale = wgt / ssum;
      drive([&](int j0, auto n_) {
        constexpr int N = decltype(n_)::value;
        uint4 q[N];
#pragma unroll
        for (int u = 0; u < N; ++u) q[u] = s4[j0 + GT * u];
#pragma unroll
        for (int u = 0; u < N; ++u) {
          float f[PER];
          StreamMid<T>::unpack(q[u], f);
#pragma unroll
          for (int e = 0; e < PER; ++e) f[e] *= scale;
          s4[j0 + GT * u] = pack16<T>(f);
        }
      });
      fence_proxy_async();
      if (G > 1) group_sync(bar_id, GT); else __syncwarp();
      if (gtid == 0) {
        bulk_s2g(grad + (int64_t)b * p.gstride_b + (int64_t)t * p.gstride_t, sp, row_bytes);
        bulk_commit();
        bulk_wait_read<0>();
        mbar_arrive(&empty[slot]);
      }
    }
  }
  if (gtid == 0) bulk_wait_all<0>();
  if (gtid == 0) TL_MAX(1);
}

struct StreamCfg {
  int nslots, slot_bytes, smem, group, park_floats;
};

// The one-pass kernel needs whole rows resident in shared memory and 16-byte aligned rows.
static bool plan_stream(const LossParams& p, int es, StreamCfg* c) {
  const int64_t rb = p.V * es;
  if (rb % 16 != 0) return false;
  c->slot_bytes = (int)align_up((size_t)rb, 128);
  int64_t n = kStreamSmemBudget / c->slot_bytes;
  if (n > kStreamMaxSlots) n = kStreamMaxSlots;
  if (n < 4) return false;
  c->nslots = (int)n;
  c->smem = 2 * kStreamMaxSlots * (int)sizeof(uint64_t) + c->nslots * c->slot_bytes;
  // warps per row: leave at least a third of the ring to loads in flight
  c->group = 1;
  while (c->group < 4 && kStreamWarps / c->group > (c->nslots * 2) / 3) c->group *= 2;
  // a strip per row group for the raw values of the label positions that do not fit the register slots
  const int gt = 32 * c->group, ngrp = kStreamWarps / c->group;
  c->park_floats = (int)align_up((size_t)((p.Lmax + 1 > gt * kStreamGather) ? (p.Lmax + 1 - gt * kStreamGather) : 0), 4);
  const int64_t park_bytes = (int64_t)ngrp * c->park_floats * 4;
  while (c->nslots > 4 && (int64_t)c->nslots * c->slot_bytes + park_bytes > kStreamSmemBudget) --c->nslots;
  if ((int64_t)c->nslots * c->slot_bytes + park_bytes > kStreamSmemBudget) return false;
  c->smem = 2 * kStreamMaxSlots * (int)sizeof(uint64_t) + c->nslots * c->slot_bytes + (int)park_bytes;
  if (kStreamWarps / c->group > c->nslots - 1) return false;
  return true;
}

template <typename T>
static int launch_stream_t(const LossParams& p, const StreamCfg& c, cudaStream_t st) {
  const int64_t nrows = (int64_t)(p.b1 - p.b0) * p.T;
  int64_t g = num_sms();
  if (g > nrows) g = nrows;
  auto go = [&](auto kern) -> int {
    B200_CUDA(set_max_smem_cached((const void*)kern, c.smem));
    B200_CUDA(launch_kernel(kern, dim3((unsigned)g), dim3((kStreamWarps + 1) * 32), (size_t)c.smem, st,
                            true, p, c.nslots, c.slot_bytes, c.park_floats));
    return B200_OK;
  };
  if (c.group == 1) return go(ctc_stream_kernel<T, 1>);
  if (c.group == 2) return go(ctc_stream_kernel<T, 2>);
  return go(ctc_stream_kernel<T, 4>);
}

static int launch_stream(const LossParams& p, int dtype, const StreamCfg& c, cudaStream_t st) {
  switch (dtype) {
    case B200_F32: return launch_stream_t<float>(p, c, st);
    case B200_BF16: return launch_stream_t<__nv_bfloat16>(p, c, st);
    case B200_F16: return launch_stream_t<__half>(p, c, st);
  }
  set_error("unknown dtype %d", dtype);
  return B200_ERR_INVALID_ARGUMENT;
}

}  // namespace b200
