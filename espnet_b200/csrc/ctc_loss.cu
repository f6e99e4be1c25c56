// ctc_loss.cu -- fused log_softmax + CTC alpha/beta lattice + gradient for sm_100a.
// See ctc_loss.cuh for the kernel map and include/espnet_b200_ctc.h for the ABI.
#include "ctc_loss.cuh"

#include <mutex>

// Developer instrumentation (-DB200_TIMELINE): global-timer stamps of the kernels of one call -- even slots keep the
// minimum (first CTA / warp to get there), odd slots the maximum; slots 64 + 2*b + d: end of the sweep of utterance b,
// direction d.  Read back / reset with b200_debug_timeline (tools/dev_timeline.py).
#ifdef B200_TIMELINE
namespace b200 {
__device__ unsigned long long g_tl[1024];
__device__ __forceinline__ unsigned long long tl_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
}  // namespace b200
#define TL_MIN(i) atomicMin(&b200::g_tl[i], b200::tl_now())
#define TL_MAX(i) atomicMax(&b200::g_tl[i], b200::tl_now())
#define TL_SET(i) (b200::g_tl[i] = b200::tl_now())
extern "C" int b200_debug_timeline(unsigned long long* host, int reset) {
  if (reset) {
    unsigned long long init[1024];
    for (int i = 0; i < 1024; ++i) init[i] = (i < 64 && !(i & 1)) ? ~0ull : 0ull;
    return (int)cudaMemcpyToSymbol(b200::g_tl, init, sizeof(init));
  }
  return (int)cudaMemcpyFromSymbol(host, b200::g_tl, sizeof(unsigned long long) * 1024);
}
#else
#define TL_MIN(i) ((void)0)
#define TL_MAX(i) ((void)0)
#define TL_SET(i) ((void)0)
#endif

namespace b200 {

// ============================================================================
// prep: labels -> int32, duplicate chains (deterministic, atomics-free gradient)
// ============================================================================
__global__ void ctc_prep_kernel(LossParams p) {
  pdl_launch_dependents();  // the stream kernel may be staged (it waits for this grid before it reads labels)
  const int b = p.b0 + blockIdx.x;
  // olen beyond the Lmax the workspace was sized for is clamped (and flagged): nothing below may index past Kp
  const int64_t Lraw = p.olen[b];
  const int64_t L = Lraw < 0 ? 0 : (Lraw > p.Lmax ? p.Lmax : Lraw);
  const int64_t* y = p.targets + p.tgt_offsets[b];
  int* lab = p.ws.labels + (int64_t)b * p.ws.Kp;
  int* nxt = p.ws.dup_next + (int64_t)b * p.ws.Kp;
  int* fst = p.ws.is_first + (int64_t)b * p.ws.Kp;
  extern __shared__ int s_lab[];
  // the counters of the kernels that follow (task queues, finished-sweep count) start at zero; this grid
  // completes before any of them runs, so no separate memset node is needed in front of a call
  if (blockIdx.x == 0 && threadIdx.x < 8) p.ws.bad[threadIdx.x] = 0;
  __shared__ int s_bad;
  if (threadIdx.x == 0) { p.ws.done[b] = 0; p.ws.redo[b] = 0; s_bad = (Lraw != L) ? 1 : 0; }
  __syncthreads();
  for (int64_t i = threadIdx.x; i < L; i += blockDim.x) {
    int64_t v = y[i];
    if (v < 0 || v >= p.V || v == p.blank) {
      s_bad = 1;
      v = v < 0 ? 0 : (v >= p.V ? p.V - 1 : v);
    }
    s_lab[i] = (int)v;
    lab[i] = (int)v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    p.ws.badutt[b] = s_bad;
  }
  int* dc = p.ws.dupcode + (int64_t)b * p.ws.Kp;
  for (int64_t i = threadIdx.x; i < L; i += blockDim.x) {
    const int v = s_lab[i];
    // branch-free scan (every thread reads the same word: a broadcast)
    int first = 1, n = 0x7fffffff;
    for (int j = 0; j < (int)L; ++j) {
      const bool same = s_lab[j] == v;
      if (same && j < (int)i) first = 0;
      if (same && j > (int)i) n = min(n, j);
    }
    if (n == 0x7fffffff) n = -1;
    fst[i] = first;
    nxt[i] = n;
    dc[i] = (first ? 0 : kDupLater) | (n + 1);
  }
  // rank of this utterance by decreasing frame count (ties by index): order[rank] = b
  __shared__ int s_rank;
  if (threadIdx.x == 0) s_rank = 0;
  __syncthreads();
  {
    const int64_t mine = p.ilen[b];
    int cnt = 0;
    for (int o = p.b0 + threadIdx.x; o < p.b1; o += blockDim.x) {
      const int64_t other = p.ilen[o];
      cnt += (other > mine || (other == mine && o < b)) ? 1 : 0;
    }
    if (cnt) atomicAdd(&s_rank, cnt);
  }
  __syncthreads();
  if (threadIdx.x == 0) p.ws.order[p.b0 + s_rank] = b;
  // launched as a programmatic dependent of the stream kernel: this grid must not be seen as finished
  // before the kernel it ran under (the sweeps wait for THIS grid only)
  if (p.prep_after_stream) pdl_wait();
}

// deterministic sum of the per-utterance nll
__global__ void ctc_loss_sum_kernel(const float* nll, int B, float* out) {
  float s = 0.f;
  for (int b = threadIdx.x; b < B; b += 32) s += nll[b];
  s = warp_sum(s);
  if (threadIdx.x == 0) *out = s;
}

}  // namespace b200

#include "ctc_loss_sweep.inl"
#include "ctc_loss_lin.inl"
#include "ctc_loss_rows.inl"
#include "ctc_loss_occ.inl"
#include "ctc_loss_stream.inl"
#include "ctc_loss_host.inl"
