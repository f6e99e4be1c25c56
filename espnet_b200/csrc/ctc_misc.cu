// ctc_misc.cu -- greedy path (row argmax) and Viterbi forced alignment.
//
//   b200_ctc_greedy   espnet2/asr/ctc.py:103-111 / pytorch_backend/ctc.py:144-151 (CTC.argmax):
//                     ids[b,t] = argmax_v logits[b,t,v], first maximum, NaN counts as the maximum
//                     (torch.argmax).  One warp per row, 128-bit loads: the row is read once (HBM).
//   b200_ctc_viterbi  espnet/nets/pytorch_backend/ctc.py:153-216 (CTC.forced_align): max instead of
//                     log-sum-exp on the same lattice, float64 scores as the reference's numpy
//                     arrays, candidates compared in the reference's order [s, s-1, s-2] with the
//                     first maximum kept, including its wrap-around read of state -1 for s = 0.
#include "common.cuh"

#include <mutex>

namespace b200 {

// ------------------------------------------------------------------------------ greedy
// a is better than b: larger value, or equal value at a smaller index; NaN beats every number
__device__ __forceinline__ bool arg_better(float va, int ia, float vb, int ib) {
  const bool na = va != va, nb = vb != vb;
  if (na || nb) return na && (!nb || ia < ib);
  return va > vb || (va == vb && ia < ib);
}

constexpr int kGreedyWarps = 8;

template <typename T, bool VEC>
__global__ void __launch_bounds__(kGreedyWarps * 32) ctc_greedy_kernel(const T* __restrict__ logits, int64_t B, int64_t Tn,
                                                                      int V, int64_t stride_b, int64_t stride_t,
                                                                      int64_t* __restrict__ ids) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int PER = ElemTraits<T>::kPerVec;
  const int64_t nrows = B * Tn;
  const int nvec = (int)((int64_t)V * sizeof(T) / 16);
  for (int64_t r = (int64_t)blockIdx.x * kGreedyWarps + warp; r < nrows; r += (int64_t)gridDim.x * kGreedyWarps) {
    const T* src = logits + (r / Tn) * stride_b + (r % Tn) * stride_t;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    if (VEC) {
      const uint4* s4 = reinterpret_cast<const uint4*>(src);
      for (int j0 = lane; j0 < nvec; j0 += 128) {
        uint4 q[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = j0 + 32 * u;
          q[u] = j < nvec ? ld_nc_v4(s4 + j) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = j0 + 32 * u;
          if (j < nvec) {
            float f[PER];
            unpack16<T>(q[u], f);
#pragma unroll
            for (int e = 0; e < PER; ++e)
              if (arg_better(f[e], j * PER + e, best, bi)) { best = f[e]; bi = j * PER + e; }
          }
        }
      }
    } else {
      for (int v = lane; v < V; v += 32) {
        const float x = ElemTraits<T>::to_float(src[v]);
        if (arg_better(x, v, best, bi)) { best = x; bi = v; }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (arg_better(ov, oi, best, bi)) { best = ov; bi = oi; }
    }
    if (lane == 0) ids[r] = bi;
  }
}

template <typename T>
static int launch_greedy(const void* logits, int64_t B, int64_t Tn, int64_t V, int64_t stride_b, int64_t stride_t,
                         int64_t* ids, cudaStream_t st) {
  const int es = (int)sizeof(T);
  const bool vec = (reinterpret_cast<uintptr_t>(logits) & 15) == 0 && (V * es) % 16 == 0 && (stride_b * es) % 16 == 0 &&
                   (stride_t * es) % 16 == 0;
  int64_t g = (B * Tn + kGreedyWarps - 1) / kGreedyWarps;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  if (vec)
    ctc_greedy_kernel<T, true><<<(unsigned)g, kGreedyWarps * 32, 0, st>>>((const T*)logits, B, Tn, (int)V, stride_b,
                                                                         stride_t, ids);
  else
    ctc_greedy_kernel<T, false><<<(unsigned)g, kGreedyWarps * 32, 0, st>>>((const T*)logits, B, Tn, (int)V, stride_b,
                                                                          stride_t, ids);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// ------------------------------------------------------------------------------ Viterbi
constexpr double kVitLogZero = -100000000000.0;  // the reference's "log of zero"

__device__ __forceinline__ int vit_ext(const int64_t* y, int blank, int s) { return (s & 1) ? (int)y[s >> 1] : blank; }

// One CTA per utterance; scores of two consecutive frames ping-pong in shared memory.
__global__ void __launch_bounds__(1024) ctc_viterbi_kernel(const float* __restrict__ lpz, int T, int V,
                                                           const int64_t* __restrict__ y, int L, int blank,
                                                           int64_t* __restrict__ align, int16_t* __restrict__ path) {
  extern __shared__ double vit_delta[];  // [2][S]
  const int S = 2 * L + 1;
  double* cur = vit_delta;
  double* nxt = vit_delta + S;
  for (int s = threadIdx.x; s < S; s += blockDim.x) {
    double v = kVitLogZero;
    if (s < 2) v = (double)lpz[vit_ext(y, blank, s)];
    cur[s] = v;
  }
  __syncthreads();
  for (int t = 1; t < T; ++t) {
    const float* row = lpz + (int64_t)t * V;
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
      const int e = vit_ext(y, blank, s);
      const int ncand = (e == blank || s < 2 || e == vit_ext(y, blank, s - 2)) ? 2 : 3;
      int best_s = s;
      double best = cur[s];
      for (int j = 1; j < ncand; ++j) {
        const int ps = s - j;
        const double v = cur[ps < 0 ? ps + S : ps];  // the reference reads logdelta[t-1, -1] for s = 0
        if (v > best) { best = v; best_s = ps; }
      }
      nxt[s] = best + (double)row[e];
      path[(int64_t)t * S + s] = (int16_t)best_s;
    }
    __syncthreads();
    double* tmp = cur; cur = nxt; nxt = tmp;
  }
  if (threadIdx.x == 0) {
    // back-trace (T dependent reads of what this CTA wrote)
    const double a = cur[S - 1];
    const double b = cur[(S - 2) < 0 ? S - 2 + S : S - 2];
    int st = (a >= b) ? S - 1 : S - 2;
    for (int t = T - 1; t >= 0; --t) {
      const int idx = st < 0 ? st + S : st;
      align[t] = vit_ext(y, blank, idx);
      if (t > 0) st = path[(int64_t)t * S + idx];
    }
  }
}

// ------------------------------------------------------------------------------ gradient rescale
// grad[t,b,:] *= gout[b] / expected for t < ilen[b]  (autograd backward of the one-pass call: the forward
// already wrote expected * d nll_b / d logits; the upstream gradient is only known now).  Rows whose ratio is
// exactly 1 -- the whole batch when the trainer's scale matches the hint -- are not touched: a CTA whose rows
// all skip exits after reading B floats.
template <typename T>
__global__ void __launch_bounds__(256) ctc_grad_rescale_kernel(T* __restrict__ grad, int64_t Tn, int64_t B, int V,
                                                               int64_t gstride_t, int64_t gstride_b,
                                                               const float* __restrict__ gout, int gout_n, float expected,
                                                               const int64_t* __restrict__ ilen) {
  constexpr int PER = ElemTraits<T>::kPerVec;
  const int nvec = V > 0 ? (int)((int64_t)V * sizeof(T) / 16) : 0;  // V < 0: rows are not 16-byte aligned
  V = V > 0 ? V : -V;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const float inv = 1.0f / expected;
  pdl_launch_dependents();
  pdl_wait();  // (a programmatic dependent of the grid before it: the upstream gradient, the patch kernel's stores)
  {
    // the usual case first: every ratio is exactly 1 and there is nothing to do
    bool same = true;
    for (int b = lane; b < (gout_n == 1 ? 1 : (int)B); b += 32) same = same && (gout[b] * inv == 1.0f);
    if (__all_sync(0xffffffffu, same)) return;
  }
  for (int64_t r = (int64_t)blockIdx.x * nwarp + warp; r < B * Tn; r += (int64_t)gridDim.x * nwarp) {
    const int64_t b = r / Tn, t = r - b * Tn;
    const float ratio = gout[gout_n == 1 ? 0 : b] * inv;
    if (ratio == 1.0f) continue;
    if (ilen && t >= ilen[b]) continue;  // zeros stay zeros
    T* row = grad + b * gstride_b + t * gstride_t;
    uint4* r4 = reinterpret_cast<uint4*>(row);
    for (int j = lane; j < nvec; j += 32) {
      float f[PER];
      unpack16<T>(r4[j], f);
#pragma unroll
      for (int e = 0; e < PER; ++e) f[e] *= ratio;
      r4[j] = pack16<T>(f);
    }
    for (int v = nvec * PER + lane; v < V; v += 32) row[v] = ElemTraits<T>::from_float(ElemTraits<T>::to_float(row[v]) * ratio);
  }
}

template <typename T>
static int launch_rescale(void* grad, int64_t Tn, int64_t B, int64_t V, int64_t gstride_t, int64_t gstride_b,
                          const float* gout, int gout_n, float expected, const int64_t* ilen, cudaStream_t st) {
  const int es = (int)sizeof(T);
  const bool vec = (reinterpret_cast<uintptr_t>(grad) & 15) == 0 && (gstride_b * es) % 16 == 0 && (gstride_t * es) % 16 == 0;
  int64_t g = (B * Tn + 7) / 8;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (g > cap) g = cap;
  // (unaligned rows: V is passed as 0 vectors by giving the scalar loop the whole row)
  B200_CUDA(launch_kernel(ctc_grad_rescale_kernel<T>, dim3((unsigned)g), dim3(256), 0, st, true, (T*)grad, Tn, B,
                          vec ? (int)V : -(int)V, gstride_t, gstride_b, gout, gout_n, expected, ilen));
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_ctc_grad_rescale(void* grad, int dtype, int64_t T, int64_t B, int64_t V, int64_t gstride_t,
                                     int64_t gstride_b, const float* gout, int64_t gout_n, float expected,
                                     const int64_t* ilen, void* stream) {
  B200_CHECK_ARG(grad && gout, B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  B200_CHECK_ARG(T >= 1 && B >= 1 && V >= 1 && V < (1ll << 30), B200_ERR_INVALID_ARGUMENT,
                 "need T>=1, B>=1, 1<=V<2^30, got T=%lld B=%lld V=%lld", (long long)T, (long long)B, (long long)V);
  B200_CHECK_ARG(gout_n == 1 || gout_n == B, B200_ERR_INVALID_ARGUMENT, "gout_n=%lld must be 1 or B=%lld",
                 (long long)gout_n, (long long)B);
  B200_CHECK_ARG(expected != 0.0f && expected == expected, B200_ERR_INVALID_ARGUMENT, "expected scale must be a non-zero number");
  B200_CHECK_ARG(gstride_t >= 0 && gstride_b >= 0, B200_ERR_INVALID_ARGUMENT, "negative stride");
  cudaStream_t st = (cudaStream_t)stream;
  switch (dtype) {
    case B200_F32: return launch_rescale<float>(grad, T, B, V, gstride_t, gstride_b, gout, (int)gout_n, expected, ilen, st);
    case B200_BF16: return launch_rescale<__nv_bfloat16>(grad, T, B, V, gstride_t, gstride_b, gout, (int)gout_n, expected, ilen, st);
    case B200_F16: return launch_rescale<__half>(grad, T, B, V, gstride_t, gstride_b, gout, (int)gout_n, expected, ilen, st);
  }
  set_error("dtype must be B200_F32, B200_BF16 or B200_F16, got %d", dtype);
  return B200_ERR_INVALID_ARGUMENT;
}

extern "C" int b200_ctc_greedy(const void* logits, int dtype, int64_t B, int64_t T, int64_t V, int64_t stride_b,
                               int64_t stride_t, int64_t* ids, void* stream) {
  B200_CHECK_ARG(logits && ids, B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  B200_CHECK_ARG(B >= 1 && T >= 1 && V >= 1 && V < (1ll << 30), B200_ERR_INVALID_ARGUMENT,
                 "need B>=1, T>=1, 1<=V<2^30, got B=%lld T=%lld V=%lld", (long long)B, (long long)T, (long long)V);
  B200_CHECK_ARG(stride_b >= 0 && stride_t >= 0, B200_ERR_INVALID_ARGUMENT, "negative stride");
  cudaStream_t st = (cudaStream_t)stream;
  switch (dtype) {
    case B200_F32: return launch_greedy<float>(logits, B, T, V, stride_b, stride_t, ids, st);
    case B200_BF16: return launch_greedy<__nv_bfloat16>(logits, B, T, V, stride_b, stride_t, ids, st);
    case B200_F16: return launch_greedy<__half>(logits, B, T, V, stride_b, stride_t, ids, st);
  }
  set_error("dtype must be B200_F32, B200_BF16 or B200_F16, got %d", dtype);
  return B200_ERR_INVALID_ARGUMENT;
}

extern "C" size_t b200_ctc_viterbi_work_bytes(int64_t T, int64_t L) {
  if (T < 1 || L < 0) return 0;
  return (size_t)T * (size_t)(2 * L + 1) * sizeof(int16_t);
}

extern "C" int b200_ctc_viterbi(const float* lpz, int64_t T, int64_t V, const int64_t* y, int64_t L, int blank,
                                int64_t* align, void* work, size_t work_bytes, void* stream) {
  B200_CHECK_ARG(lpz && align && work && (y || L == 0), B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  B200_CHECK_ARG(T >= 1 && V >= 1 && L >= 0 && blank >= 0 && blank < V, B200_ERR_INVALID_ARGUMENT,
                 "need T>=1, V>=1, L>=0, 0<=blank<V, got T=%lld V=%lld L=%lld blank=%d", (long long)T, (long long)V,
                 (long long)L, blank);
  B200_CHECK_ARG(2 * L + 1 <= 32767, B200_ERR_UNSUPPORTED_SHAPE,
                 "L=%lld: back-pointers are int16 as in the reference (2L+1 <= 32767)", (long long)L);
  B200_CHECK_ARG(T < (1ll << 31), B200_ERR_UNSUPPORTED_SHAPE, "T too large");
  B200_CHECK_ARG(work_bytes >= b200_ctc_viterbi_work_bytes(T, L), B200_ERR_WORKSPACE_TOO_SMALL,
                 "work has %zu bytes, need %zu", work_bytes, b200_ctc_viterbi_work_bytes(T, L));
  const int S = (int)(2 * L + 1);
  const size_t smem = 2 * (size_t)S * sizeof(double);
  B200_CHECK_ARG(smem <= (size_t)max_smem_optin(), B200_ERR_UNSUPPORTED_SHAPE,
                 "L=%lld needs %zu bytes of shared memory", (long long)L, smem);
  if (smem > 48 * 1024)
    B200_CUDA(cudaFuncSetAttribute(ctc_viterbi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int threads = (S + 31) / 32 * 32;
  if (threads > 1024) threads = 1024;
  ctc_viterbi_kernel<<<1, threads, smem, (cudaStream_t)stream>>>(lpz, (int)T, (int)V, y, (int)L, blank, align,
                                                                 (int16_t*)work);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// ------------------------------------------------------------------------------ Viterbi without caller scratch
// b200_ctc_viterbi_static: the same alignment with the back-pointer table in a file-scope device array instead of a
// caller-allocated work buffer, for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that
// forbid cudaMalloc).  No counterpart in the reference.
//   * the array costs B200_CTC_VITERBI_STATIC_BYTES of device memory on EVERY GPU that loads the library, until exit;
//   * in host code `g_vit_scratch` is a host-side stand-in, not a device address: the address a kernel can use comes
//     from cudaGetSymbolAddress, fetched once per device with that device current and kept here;
//   * every call on a device shares the array: calls that overlap on the GPU (two streams) race.  Nothing in it has
//     to start at zero -- the kernel writes every back-pointer of frames 1..T-1 before the back-trace reads them --
//     so what the previous call left behind is simply overwritten.
static __device__ __align__(256) unsigned char g_vit_scratch[B200_CTC_VITERBI_STATIC_BYTES];

namespace {
std::mutex g_vit_mu;
void* g_vit_ptr[64] = {nullptr};
}  // namespace

static int vit_static_ptr(void** out) {
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  B200_CHECK_ARG(dev >= 0 && dev < 64, B200_ERR_INVALID_ARGUMENT, "device index %d outside [0,64)", dev);
  std::lock_guard<std::mutex> lk(g_vit_mu);
  if (g_vit_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    B200_CUDA(cudaGetSymbolAddress((void**)&ptr, g_vit_scratch));  // (the current device's instance of the array)
    B200_CHECK_ARG(ptr != nullptr, B200_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
    g_vit_ptr[dev] = ptr;
  }
  *out = g_vit_ptr[dev];
  return B200_OK;
}

extern "C" size_t b200_ctc_viterbi_static_max_bytes(void) { return (size_t)B200_CTC_VITERBI_STATIC_BYTES; }

extern "C" int b200_ctc_viterbi_static(const float* lpz, int64_t T, int64_t V, const int64_t* y, int64_t L, int blank,
                                       int64_t* align, void* stream) {
  B200_CHECK_ARG(T >= 1 && L >= 0, B200_ERR_INVALID_ARGUMENT, "need T>=1, L>=0, got T=%lld L=%lld", (long long)T,
                 (long long)L);
  const size_t need = b200_ctc_viterbi_work_bytes(T, L);
  B200_CHECK_ARG(need <= (size_t)B200_CTC_VITERBI_STATIC_BYTES, B200_ERR_UNSUPPORTED_SHAPE,
                 "T=%lld, L=%lld need %zu bytes of back-pointers, the static scratch holds %zu: use b200_ctc_viterbi "
                 "with a caller-allocated work buffer",
                 (long long)T, (long long)L, need, (size_t)B200_CTC_VITERBI_STATIC_BYTES);
  void* work = nullptr;
  const int rc = vit_static_ptr(&work);
  if (rc) return rc;
  return b200_ctc_viterbi(lpz, T, V, y, L, blank, align, work, (size_t)B200_CTC_VITERBI_STATIC_BYTES, stream);
}
