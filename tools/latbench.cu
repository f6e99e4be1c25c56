// latbench.cu -- dependent-chain latencies that bound the lattice sweep (developer tool, sm_100a).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/latbench tools/latbench.cu && tools/latbench
// One CTA, `warps` warps; every chain is a loop of N dependent iterations timed with clock64 by lane 0 of
// warp 0.  Prints cycles per iteration.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float lg2f(float x) { float y; asm volatile("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

constexpr int N = 4096;

template <int MODE>
__global__ void chain(float* out, long long* cyc, float seed, float lp) {
  float a = seed + threadIdx.x * 1e-3f, b = seed * 0.5f;
  float acc = 0.f;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 8
  for (int k = 0; k < N; ++k) {
    if (MODE == 0) {  // FADD only
      a = a + lp;
    } else if (MODE == 1) {  // SHFL.UP + FADD
      a = __shfl_up_sync(0xffffffffu, a, 1) + lp;
    } else if (MODE == 2) {  // EX2 + FADD
      a = ex2f(a) + lp;
    } else if (MODE == 3) {  // LG2 + FADD
      a = lg2f(a) + 1.5f;
    } else if (MODE == 4) {  // the log-domain frame: shfl -> lse3_pre and lse2 -> add
      const float sh = __shfl_up_sync(0xffffffffu, a, 1);
      const float hi = fmaxf(a, b);
      const float s1 = 1.0f + ex2f(-fabsf(a - b));
      const float x = fmaxf(b, sh) + lg2f(1.0f + ex2f(-fabsf(b - sh)));
      const float t = ex2f(-fabsf(sh - hi));
      const float sum = (sh <= hi) ? (s1 + t) : fmaf(s1, t, 1.0f);
      const float y = fmaxf(hi, sh) + lg2f(sum);
      b = lp + x;
      a = lp + y;
    } else if (MODE == 5) {  // linear-domain frame: shfl -> fma -> add -> mul
      const float sh = __shfl_up_sync(0xffffffffu, a, 1);
      const float nb = b + sh;
      const float na = fmaf(sh, seed, a) + b;
      b = nb * lp;
      a = na * lp;
    } else if (MODE == 6) {  // log-domain frame without the shuffle on the chain (neighbour from 2 steps ago)
      const float sh = __shfl_up_sync(0xffffffffu, acc, 1);
      acc = a;
      const float hi = fmaxf(a, b);
      const float s1 = 1.0f + ex2f(-fabsf(a - b));
      const float x = fmaxf(b, sh) + lg2f(1.0f + ex2f(-fabsf(b - sh)));
      const float t = ex2f(-fabsf(sh - hi));
      const float sum = (sh <= hi) ? (s1 + t) : fmaf(s1, t, 1.0f);
      const float y = fmaxf(hi, sh) + lg2f(sum);
      b = lp + x;
      a = lp + y;
    } else if (MODE == 7) {  // FFMA only
      a = fmaf(a, seed, lp);
    } else if (MODE == 8) {  // SHFL.IDX (bfly) + FADD
      a = __shfl_xor_sync(0xffffffffu, a, 1) + lp;
    } else if (MODE == 9) {  // shared-memory round trip: STS -> LDS of the neighbour's word
      __shared__ float sm[1024 + 1];
      sm[threadIdx.x + 1] = a;
      __syncwarp();
      a = sm[threadIdx.x] + lp;
      __syncwarp();
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = a + b + acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void run(const char* name, int warps, float seed, float lp) {
  float* out; long long* cyc;
  cudaMalloc(&out, 4096 * sizeof(float));
  cudaMalloc(&cyc, 8 * sizeof(long long));
  for (int r = 0; r < 2; ++r) chain<MODE><<<1, warps * 32>>>(out, cyc, seed, lp);
  cudaDeviceSynchronize();
  long long h = 0;
  cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-44s warps=%d  %7.1f cycles/iter  (%s)\n", name, warps, (double)h / N, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  const int ws[3] = {1, 5, 8};
  for (int wi = 0; wi < 3; ++wi) {
    const int w = ws[wi];
    run<0>("FADD", w, 1.0f, 0.25f);
    run<7>("FFMA", w, 1.0f, 0.25f);
    run<1>("SHFL.UP + FADD", w, 1.0f, 0.25f);
    run<8>("SHFL.BFLY + FADD", w, 1.0f, 0.25f);
    run<2>("MUFU.EX2 + FADD", w, -1.0f, -2.0f);
    run<3>("MUFU.LG2 + FADD", w, 3.0f, 0.f);
    run<9>("STS -> syncwarp -> LDS + FADD", w, 1.0f, 0.25f);
    run<4>("log-domain frame (shfl on chain)", w, -3.0f, -0.5f);
    run<6>("log-domain frame (shfl off chain)", w, -3.0f, -0.5f);
    run<5>("linear-domain frame", w, 0.5f, 0.4f);
  }
  return 0;
}
