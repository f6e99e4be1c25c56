// common.cuh -- shared device helpers and host-side error plumbing for
// libespnet_b200_ctc.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/espnet_b200_ctc.h"

namespace b200 {

// ---------------------------------------------------------------- errors
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
int num_sms();        // SM count of the current device (cached per device)
int max_smem_optin(); // max dynamic shared memory per block (opt-in)
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) / the occupancy query, remembered per (kernel, device, size):
// both cost microseconds of host time, and a training step launches the same kernels with the same sizes forever
cudaError_t set_max_smem_cached(const void* kern, int bytes);
cudaError_t max_blocks_per_sm_cached(const void* kern, int threads, size_t smem, int* out);

#define B200_CHECK_ARG(cond, code, ...)            \
  do {                                             \
    if (!(cond)) {                                 \
      ::b200::set_error(__VA_ARGS__);              \
      return (code);                               \
    }                                              \
  } while (0)

#define B200_CUDA(expr)                                         \
  do {                                                          \
    cudaError_t e__ = (expr);                                   \
    if (e__ != cudaSuccess) return ::b200::cuda_fail(e__, #expr); \
  } while (0)

// Kernel launch with (optionally) programmatic stream serialization: the kernel may be staged
// while its predecessor in the stream is still running and synchronises with it on the device
// (griddepcontrol.wait / flags in global memory).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                 bool pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------- math
// Everything on the lattice is kept in log2 units so that the recursion uses
// the MUFU ex2/lg2 units directly.
#define B200_LOG2E 1.4426950408889634f
#define B200_LN2 0.6931471805599453f
// Finite stand-in for log(0) inside the lattice kernels (keeps a - a == 0).
#define B200_NEG (-1.0e30f)

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2f(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// log2(2^a + 2^b): one ex2 + one lg2.
__device__ __forceinline__ float lse2(float a, float b) {
  float m = fmaxf(a, b);
  float d = fminf(a, b) - m;  // <= 0
  return m + lg2f(1.0f + ex2f(d));
}
// log2(2^a + 2^b + 2^c): two ex2 + one lg2.
__device__ __forceinline__ float lse3(float a, float b, float c) {
  float hi = fmaxf(a, b), lo = fminf(a, b);
  float m = fmaxf(hi, c);
  float mid = fminf(hi, c);
  return m + lg2f(1.0f + ex2f(lo - m) + ex2f(mid - m));
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// Order-preserving float <-> int map so that REDUX.MAX.S32 reduces floats.
__device__ __forceinline__ int float_ord(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : (i ^ 0x7fffffff);
}
__device__ __forceinline__ float ord_float(int i) {
  return __int_as_float(i >= 0 ? i : (i ^ 0x7fffffff));
}
__device__ __forceinline__ float warp_max_redux(float v) {
  return ord_float(__reduce_max_sync(0xffffffffu, float_ord(v)));
}

// ---------------------------------------------------------------- element types
template <typename T> struct ElemTraits;
template <> struct ElemTraits<float> {
  static constexpr int kDtype = B200_F32;
  static constexpr int kPerVec = 4;  // elements per 16-byte vector
  __device__ static __forceinline__ float to_float(float v) { return v; }
  __device__ static __forceinline__ float from_float(float v) { return v; }
};
template <> struct ElemTraits<__nv_bfloat16> {
  static constexpr int kDtype = B200_BF16;
  static constexpr int kPerVec = 8;
  __device__ static __forceinline__ float to_float(__nv_bfloat16 v) { return __bfloat162float(v); }
  __device__ static __forceinline__ __nv_bfloat16 from_float(float v) { return __float2bfloat16_rn(v); }
};
template <> struct ElemTraits<__half> {
  static constexpr int kDtype = B200_F16;
  static constexpr int kPerVec = 8;
  __device__ static __forceinline__ float to_float(__half v) { return __half2float(v); }
  __device__ static __forceinline__ __half from_float(float v) { return __float2half_rn(v); }
};

// Unpack one 16-byte vector into floats / pack floats back.
template <typename T>
__device__ __forceinline__ void unpack16(const uint4& q, float* f);
template <>
__device__ __forceinline__ void unpack16<float>(const uint4& q, float* f) {
  f[0] = __uint_as_float(q.x); f[1] = __uint_as_float(q.y);
  f[2] = __uint_as_float(q.z); f[3] = __uint_as_float(q.w);
}
template <>
__device__ __forceinline__ void unpack16<__nv_bfloat16>(const uint4& q, float* f) {
  // bf16 -> fp32 is a 16-bit shift
  f[0] = __uint_as_float(q.x << 16); f[1] = __uint_as_float(q.x & 0xffff0000u);
  f[2] = __uint_as_float(q.y << 16); f[3] = __uint_as_float(q.y & 0xffff0000u);
  f[4] = __uint_as_float(q.z << 16); f[5] = __uint_as_float(q.z & 0xffff0000u);
  f[6] = __uint_as_float(q.w << 16); f[7] = __uint_as_float(q.w & 0xffff0000u);
}
template <>
__device__ __forceinline__ void unpack16<__half>(const uint4& q, float* f) {
  const __half2* h = reinterpret_cast<const __half2*>(&q);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __half22float2(h[i]);
    f[2 * i] = t.x; f[2 * i + 1] = t.y;
  }
}
template <typename T>
__device__ __forceinline__ uint4 pack16(const float* f);
template <>
__device__ __forceinline__ uint4 pack16<float>(const float* f) {
  return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                    __float_as_uint(f[3]));
}
template <>
__device__ __forceinline__ uint4 pack16<__nv_bfloat16>(const float* f) {
  uint4 q;
  __nv_bfloat162 a = __floats2bfloat162_rn(f[0], f[1]);
  __nv_bfloat162 b = __floats2bfloat162_rn(f[2], f[3]);
  __nv_bfloat162 c = __floats2bfloat162_rn(f[4], f[5]);
  __nv_bfloat162 d = __floats2bfloat162_rn(f[6], f[7]);
  q.x = *reinterpret_cast<uint32_t*>(&a); q.y = *reinterpret_cast<uint32_t*>(&b);
  q.z = *reinterpret_cast<uint32_t*>(&c); q.w = *reinterpret_cast<uint32_t*>(&d);
  return q;
}
template <>
__device__ __forceinline__ uint4 pack16<__half>(const float* f) {
  uint4 q;
  __half2 a = __floats2half2_rn(f[0], f[1]);
  __half2 b = __floats2half2_rn(f[2], f[3]);
  __half2 c = __floats2half2_rn(f[4], f[5]);
  __half2 d = __floats2half2_rn(f[6], f[7]);
  q.x = *reinterpret_cast<uint32_t*>(&a); q.y = *reinterpret_cast<uint32_t*>(&b);
  q.z = *reinterpret_cast<uint32_t*>(&c); q.w = *reinterpret_cast<uint32_t*>(&d);
  return q;
}

// ---------------------------------------------------------------- mbarrier + 1-D bulk (TMA) copies
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
// global -> shared bulk copy (SASS: UBLKCP), completion counted on `bar`.
// dst, src and bytes must be multiples of 16.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// shared -> global bulk copy, tracked by the bulk async-group of the issuing thread.
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// make generic-proxy smem writes visible to the async proxy (before bulk_s2g)
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// programmatic dependent launch (PDL): no-ops when the kernel was launched without the attribute
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// L2-coherent loads for data another kernel of the same call is still producing
__device__ __forceinline__ float ld_cg_f32(const float* p) {
  float r;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ int ld_acquire_s32(const int* p) {
  int r;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ float ld_relaxed_f32(const float* p) {
  float r;
  asm volatile("ld.relaxed.gpu.global.f32 %0, [%1];" : "=f"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_relaxed_f32(float* p, float v) {
  asm volatile("st.relaxed.gpu.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// raw (unconverted) element loads: issue several, convert later, so that the loads overlap
__device__ __forceinline__ uint32_t ld_nc_raw(const float* p) {
  uint32_t r;
  asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ld_nc_raw(const __nv_bfloat16* p) {
  unsigned short r;
  asm volatile("ld.global.nc.u16 %0, [%1];" : "=h"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ld_nc_raw(const __half* p) {
  unsigned short r;
  asm volatile("ld.global.nc.u16 %0, [%1];" : "=h"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ float ld_nc_f32(const float* p) {
  float r;
  asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}
template <typename T> __device__ __forceinline__ float raw_to_float(uint32_t r);
template <> __device__ __forceinline__ float raw_to_float<float>(uint32_t r) { return __uint_as_float(r); }
template <> __device__ __forceinline__ float raw_to_float<__nv_bfloat16>(uint32_t r) {
  return __uint_as_float(r << 16);
}
template <> __device__ __forceinline__ float raw_to_float<__half>(uint32_t r) {
  return __half2float(__ushort_as_half((unsigned short)r));
}
__device__ __forceinline__ void st_cs_v4(void* p, const uint4& v) {
  asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

}  // namespace b200
