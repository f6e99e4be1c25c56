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
  double da = a, db = b;
  const double dseed = seed, dlp = lp;
  int ea = threadIdx.x;
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
    } else if (MODE == 10) {  // DFMA only
      da = fma(da, dseed, dlp);
    } else if (MODE == 11) {  // DADD only
      da = da + dlp;
    } else if (MODE == 12) {  // DMUL only
      da = da * dseed;
    } else if (MODE == 13) {  // 64-bit SHFL.UP (two 32-bit shuffles) + DADD
      const int lo = __shfl_up_sync(0xffffffffu, __double2loint(da), 1);
      const int hi = __shfl_up_sync(0xffffffffu, __double2hiint(da), 1);
      da = __hiloint2double(hi, lo) + dlp;
    } else if (MODE == 14) {  // fp64 scaled-linear frame: shfl64 -> DFMA (label state), blank state off the chain
      const int lo = __shfl_up_sync(0xffffffffu, __double2loint(da), 1);
      const int hi = __shfl_up_sync(0xffffffffu, __double2hiint(da), 1);
      const double sh = __hiloint2double(hi, lo);
      const double q = dlp * (da + db);          // p * (al + ab): known before the shuffle returns
      const double nb = dseed * (db + sh);       // blank state
      da = fma(dlp, sh, q);                      // label state: p * (al + ab + nbr)
      db = nb;
    } else if (MODE == 15) {  // the same + conversion of both states to log2 fp32 and a float2 store (issue load)
      const int lo = __shfl_up_sync(0xffffffffu, __double2loint(da), 1);
      const int hi = __shfl_up_sync(0xffffffffu, __double2hiint(da), 1);
      const double sh = __hiloint2double(hi, lo);
      const double q = dlp * (da + db);
      const double nb = dseed * (db + sh);
      da = fma(dlp, sh, q);
      db = nb;
      float l2[2];
      const double v[2] = {da, db};
#pragma unroll
      for (int z = 0; z < 2; ++z) {
        const int h = __double2hiint(v[z]), l = __double2loint(v[z]);
        const int e = (h >> 20) - 1023;
        const unsigned m = (((unsigned)h << 3) | ((unsigned)l >> 29)) & 0x7fffffu;
        l2[z] = (float)e + lg2f(__uint_as_float(m | 0x3f800000u));
      }
      reinterpret_cast<float2*>(out)[(k & 63) * 1024 + threadIdx.x] = make_float2(l2[0], l2[1]);
    } else if (MODE == 16) {  // fp32 value + own integer exponent (software extended range) frame
      const float sh = __shfl_up_sync(0xffffffffu, a, 1);
      const int she = __shfl_up_sync(0xffffffffu, ea, 1);
      const int d = max(min(she - ea, 126), -126);
      const float kf = __int_as_float((d + 127) << 23);
      const float na = fmaf(sh, kf, a + b) * lp;
      const float nb2 = fmaf(sh, kf, b) * seed;
      const int ex = (__float_as_int(na) >> 23) - 127;   // renormalise
      ea += ex;
      const float sc = __int_as_float((127 - ex) << 23);
      a = na * sc;
      b = nb2 * sc;
    } else if (MODE == 9) {  // shared-memory round trip: STS -> LDS of the neighbour's word
      __shared__ float sm[1024 + 1];
      sm[threadIdx.x + 1] = a;
      __syncwarp();
      a = sm[threadIdx.x] + lp;
      __syncwarp();
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = a + b + acc + (float)(da + db) + (float)ea;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void run(const char* name, int warps, float seed, float lp) {
  float* out; long long* cyc;
  cudaMalloc(&out, 64 * 1024 * 2 * sizeof(float) + 4096 * sizeof(float));
  cudaMalloc(&cyc, 8 * sizeof(long long));
  for (int r = 0; r < 2; ++r) chain<MODE><<<1, warps * 32>>>(out, cyc, seed, lp);
  cudaDeviceSynchronize();
  long long h = 0;
  cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-44s warps=%d  %7.1f cycles/iter  (%s)\n", name, warps, (double)h / N, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  const int ws[4] = {1, 5, 8, 16};
  for (int wi = 0; wi < 4; ++wi) {
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
    run<10>("DFMA", w, 0.999f, 0.25f);
    run<11>("DADD", w, 1.0f, 0.25f);
    run<12>("DMUL", w, 0.9999f, 0.25f);
    run<13>("SHFL.UP x2 (64 bit) + DADD", w, 1.0f, 0.25f);
    run<14>("fp64 scaled-linear frame", w, 0.45f, 0.4f);
    run<15>("fp64 frame + log2 conversion + store", w, 0.45f, 0.4f);
    run<16>("fp32 + integer exponent frame", w, 0.45f, 0.4f);
  }
  return 0;
}
