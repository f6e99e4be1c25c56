// Developer microbenchmark: how many bytes in flight per SM does a row-streaming read need on B200?
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>
__device__ __forceinline__ uint4 ldnc(const uint4* p) {
  uint4 r; asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x),"=r"(r.y),"=r"(r.z),"=r"(r.w) : "l"(p)); return r; }
template <int U, int WORK, bool DB>
__global__ void __launch_bounds__(1024,1) k(const uint4* __restrict__ src, int rows, int nvec, unsigned* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  unsigned acc = 0;
  const int NW = gridDim.x * nw;
  const int chunks = (nvec + 32 * U - 1) / (32 * U);
  const long total = (long)rows * chunks;
  long it = (long)(blockIdx.x * nw + warp);
  uint4 cur[U], nxt[U];
  auto load = [&](uint4* q, long i) {
    const int r = (int)(i / chunks), c = (int)(i % chunks);
    const uint4* p = src + (long)r * nvec + c * 32 * U;
#pragma unroll
    for (int j = 0; j < U; ++j) { int v = c * 32 * U + lane + 32 * j; q[j] = v < nvec ? ldnc(p + lane + 32 * j) : make_uint4(0,0,0,0); }
  };
  if (it < total) load(cur, it);
  for (; it < total; it += NW) {
    if (DB && it + NW < total) load(nxt, it + NW);
#pragma unroll
    for (int j = 0; j < U; ++j) {
      unsigned x = cur[j].x ^ cur[j].y ^ cur[j].z ^ cur[j].w;
#pragma unroll
      for (int w = 0; w < WORK; ++w) x = x * 1664525u + 1013904223u;   // dependent ALU work per vector
      acc += x;
    }
    if (DB) {
#pragma unroll
      for (int j = 0; j < U; ++j) cur[j] = nxt[j];
    } else if (it + NW < total) load(cur, it + NW);
  }
  if (acc == 0x12345678u) out[0] = acc;
}
__device__ __forceinline__ float ex2f(float x){float y; asm("ex2.approx.ftz.f32 %0, %1;":"=f"(y):"f"(x)); return y;}
// the real rowstats arithmetic: bf16 unpack, max pass, exp-sum pass, warp reductions per chunk
template <int U, int NACC>
__global__ void __launch_bounds__(768,1) kreal(const uint4* __restrict__ src, int rows, int nvec, float* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int NW = gridDim.x * nw;
  float tot = 0.f;
  for (int r = blockIdx.x * nw + warp; r < rows; r += NW) {
    float m_run = -3e38f, s_run = 0.f;
    for (int c0 = 0; c0 < nvec; c0 += 32 * U) {
      uint4 q[U];
#pragma unroll
      for (int j = 0; j < U; ++j) { int v = c0 + lane + 32 * j; q[j] = v < nvec ? ldnc(src + (long)r * nvec + v) : make_uint4(0xff80ff80u,0xff80ff80u,0xff80ff80u,0xff80ff80u); }
      float mm[NACC];
#pragma unroll
      for (int a = 0; a < NACC; ++a) mm[a] = -3e38f;
#pragma unroll
      for (int j = 0; j < U; ++j) {
        const unsigned w[4] = {q[j].x, q[j].y, q[j].z, q[j].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          mm[(2*e) % NACC] = fmaxf(mm[(2*e) % NACC], __uint_as_float(w[e] << 16));
          mm[(2*e+1) % NACC] = fmaxf(mm[(2*e+1) % NACC], __uint_as_float(w[e] & 0xffff0000u));
        }
      }
      float m = mm[0];
#pragma unroll
      for (int a = 1; a < NACC; ++a) m = fmaxf(m, mm[a]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
      m = fmaxf(m, m_run);
      const float ml = m * 1.4426950408889634f;
      float ss[NACC];
#pragma unroll
      for (int a = 0; a < NACC; ++a) ss[a] = 0.f;
#pragma unroll
      for (int j = 0; j < U; ++j) {
        const unsigned w[4] = {q[j].x, q[j].y, q[j].z, q[j].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          ss[(2*e) % NACC] += ex2f(fmaf(__uint_as_float(w[e] << 16), 1.4426950408889634f, -ml));
          ss[(2*e+1) % NACC] += ex2f(fmaf(__uint_as_float(w[e] & 0xffff0000u), 1.4426950408889634f, -ml));
        }
      }
      float s = ss[0];
#pragma unroll
      for (int a = 1; a < NACC; ++a) s += ss[a];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      s_run = s_run * ex2f((m_run - m) * 1.4426950408889634f) + s;
      m_run = m;
    }
    tot += m_run + s_run;
  }
  if (tot == 1.2345f) out[0] = tot;
}
template <int U, int NACC>
void runreal(const uint4* d, int rows, int nvec, float* out, int warps, const char* name) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) kreal<U, NACC><<<148, warps * 32>>>(d, rows, nvec, out);
  cudaEventRecord(a);
  for (int i = 0; i < 10; ++i) kreal<U, NACC><<<148, warps * 32>>>(d, rows, nvec, out);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); ms /= 10;
  printf("%-28s warps=%2d  %.1f us  %.0f GB/s  (err %d)\n", name, warps, ms * 1e3, (double)rows * nvec * 16 / ms / 1e6, (int)cudaGetLastError());
}
template <int U, int WORK, bool DB>
void run(const uint4* d, int rows, int nvec, unsigned* out, int warps, const char* name) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) k<U, WORK, DB><<<148, warps * 32>>>(d, rows, nvec, out);
  cudaEventRecord(a);
  for (int i = 0; i < 10; ++i) k<U, WORK, DB><<<148, warps * 32>>>(d, rows, nvec, out);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); ms /= 10;
  printf("%-28s warps=%2d  %.1f us  %.0f GB/s  (err %d)\n", name, warps, ms * 1e3, (double)rows * nvec * 16 / ms / 1e6, (int)cudaGetLastError());
}
int main() {
  const int rows = 32000, nvec = 625;  // 320 MB of 10 KB rows
  uint4* d; unsigned* out; cudaMalloc(&d, (size_t)rows * nvec * 16); cudaMalloc(&out, 4);
  cudaMemset(d, 1, (size_t)rows * nvec * 16);
  for (int w : {8, 16, 24}) {
    runreal<10, 1>(d, rows, nvec, (float*)out, w, "real U10 acc1");
    runreal<10, 4>(d, rows, nvec, (float*)out, w, "real U10 acc4");
    runreal<5, 4>(d, rows, nvec, (float*)out, w, "real U5 acc4");
    runreal<20, 4>(d, rows, nvec, (float*)out, w, "real U20 acc4");
  }
  for (int w : {16}) {
    run<10, 0, false>(d, rows, nvec, out, w, "U10 work0 single");
    run<10, 0, true>(d, rows, nvec, out, w, "U10 work0 double");
    run<10, 16, false>(d, rows, nvec, out, w, "U10 work16 single");
    run<10, 16, true>(d, rows, nvec, out, w, "U10 work16 double");
    run<20, 16, false>(d, rows, nvec, out, w, "U20 work16 single");
    run<5, 16, true>(d, rows, nvec, out, w, "U5 work16 double");
  }
  return 0;
}
