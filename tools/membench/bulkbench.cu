// Developer microbenchmark: throughput of 1-D bulk (TMA) copies of W-token row pieces at a row stride of V floats,
// through a ring of S slots of R rows per CTA -- the access pattern of the prefix score kernels (x is (B,T,V) fp32,
// a CTA owns W tokens of one utterance and walks the T frames).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulkbench bulkbench.cu && ./bulkbench
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mb_init(uint64_t* b, uint32_t n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void mb_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mb_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory"); }
__device__ __forceinline__ void mb_wait(uint64_t* b, uint32_t par) {
  asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(s32(b)), "r"(par) : "memory");
}
__device__ __forceinline__ void bulk(void* dst, const void* src, uint32_t bytes, uint64_t* b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(dst)), "l"(src), "r"(bytes), "r"(s32(b)) : "memory");
}

// grid (tiles, B); block: 1 producer warp + NC consumer warps; READ: the consumers read every word of the slot
template <bool READ>
__global__ void k(const float* x, int T, int V, int W, int R, int S, int NC, float* out) {
  extern __shared__ __align__(128) unsigned char sm[];
  uint64_t* full = (uint64_t*)sm;
  uint64_t* empty = full + 16;
  float* ring = (float*)(sm + 256);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile0 = blockIdx.x * W;
  const int ntok = min(W, V - tile0);
  const uint32_t rb = ntok * 4;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mb_init(&full[s], 1); mb_init(&empty[s], NC); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  const int nst = (T + R - 1) / R;
  if (warp == NC) {
    if (lane == 0) {
      const float* xu = x + (int64_t)blockIdx.y * T * V + tile0;
      for (int s = 0; s < nst; ++s) {
        const int slot = s % S;
        if (s >= S) mb_wait(&empty[slot], ((s / S) - 1) & 1);
        const int n = min(R, T - s * R);
        mb_expect(&full[slot], n * rb);
        for (int r = 0; r < n; ++r) bulk(ring + ((size_t)slot * R + r) * W, xu + (int64_t)(s * R + r) * V, rb, &full[slot]);
      }
    }
    return;
  }
  float acc = 0.f;
  for (int s = 0; s < nst; ++s) {
    const int slot = s % S;
    mb_wait(&full[slot], (s / S) & 1);
    if (READ) {
      const float4* p = (const float4*)(ring + (size_t)slot * R * W);
      for (int i = warp * 32 + lane; i < R * W / 4; i += NC * 32) { const float4 v = p[i]; acc += v.x + v.y + v.z + v.w; }
    }
    __syncwarp();
    if (lane == 0) mb_arrive(&empty[slot]);
  }
  if (acc == 1.2345f) out[0] = acc;
}

int main(int argc, char** argv) {
  const int B = 32, T = 500, V = 5000;
  float* x; float* out;
  cudaMalloc(&x, (size_t)B * T * V * 4);
  cudaMalloc(&out, 4);
  cudaMemset(x, 0, (size_t)B * T * V * 4);
  float* flush; cudaMalloc(&flush, 256u << 20);
  struct Cfg { int W, R, S, NC; } cfgs[] = {
      {256, 4, 6, 2}, {256, 8, 4, 4}, {256, 8, 8, 4}, {256, 16, 4, 4}, {512, 8, 4, 4}, {512, 8, 6, 4}, {512, 16, 4, 4},
      {1024, 4, 4, 4}, {1024, 8, 4, 4}, {1024, 8, 6, 4}, {1280, 8, 4, 4}, {2500, 4, 4, 4}, {2500, 4, 5, 4}, {5000, 2, 4, 4}, {5000, 2, 5, 4},
      {128, 8, 4, 2}, {128, 16, 4, 2}, {128, 16, 8, 2}, {64, 16, 8, 1}, {64, 32, 8, 1}};
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int read = 0; read < 2; ++read)
    for (auto c : cfgs) {
      const size_t smem = 256 + (size_t)c.S * c.R * c.W * 4;
      if (smem > 227 * 1024) continue;
      auto kern = read ? k<true> : k<false>;
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      int per_sm = 0;
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, (c.NC + 1) * 32, smem);
      dim3 grid((V + c.W - 1) / c.W, B);
      float best = 1e9f;
      for (int rep = 0; rep < 5; ++rep) {
        cudaMemsetAsync(flush, rep, 256u << 20);
        cudaEventRecord(e0);
        kern<<<grid, (c.NC + 1) * 32, smem>>>(x, T, V, c.W, c.R, c.S, c.NC, out);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
      }
      const cudaError_t err = cudaGetLastError();
      printf("read=%d W=%4d R=%2d S=%d NC=%d smem=%6zu CTAs=%4d per_sm=%2d waves=%.2f : %7.1f us  %.2f TB/s %s\n", read, c.W, c.R, c.S, c.NC,
             smem, grid.x * grid.y, per_sm, grid.x * grid.y / (148.0 * per_sm), best * 1e3, (double)B * T * V * 4 / (best * 1e-3) / 1e12,
             err ? cudaGetErrorString(err) : "");
    }
  return 0;
}
