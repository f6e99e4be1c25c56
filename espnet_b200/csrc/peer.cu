// peer.cu -- per-step exchange of (sum of nll, number of utterances) between the GPUs of one box, over peer memory.
//
// Replaces, for this one message, the two NCCL all-reduces of espnet2/torch_utils/recursive_op.py:24,50
// (recursive_average, called every iteration by espnet2/train/trainer.py:386-393 before backward).  The message is
// 8 bytes; an NCCL all-reduce of it is pure launch + protocol latency.  Here every rank owns a small slot ring in
// SYMMETRIC memory (allocated and exchanged by the host, torch.distributed._symmetric_memory), and ONE kernel
// launch per step
//   1. stores this rank's pair and then, with release semantics at system scope, the step's sequence number into
//      its slot of EVERY peer's ring (NVLink peer stores),
//   2. waits until the slots of all ranks in ITS OWN ring carry the step's sequence number (acquire loads of local
//      memory: no traffic while waiting),
//   3. sums the pairs in rank order (every rank gets bitwise the same result) and writes the total.
// Slot of rank r for step s: ring[(s % kPeerRing)][r] = {sum, n, seq, pad}.  A rank stores step s only after it has
// read all slots of step s-1 (stream order), so when anyone overwrites the entry of step s with step s + kPeerRing
// every rank has long finished reading it (kPeerRing >= 2 suffices).
// The wait is bounded: if a peer never arrives (a rank died) the kernel gives up after ~2 s, writes NaN and sets the
// status word, so a broken job fails loudly instead of hanging the GPU.
#include "common.cuh"

namespace b200 {

constexpr int kPeerRing = 4;
constexpr int kPeerMaxWorld = 16;

struct PeerPtrs {
  float* ring[kPeerMaxWorld];  // ring[r]: device address of rank r's slot ring, as seen from this device
};

__device__ __forceinline__ void st_relaxed_sys_f32(float* p, float v) {
  asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void st_release_sys_s32(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys_s32(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(32) peer_sum_exchange_kernel(PeerPtrs pp, int world, int rank, int seq,
                                                              const float* local_sum, float sum_scale, float my_n, float* pair,
                                                              int* status) {
  const int lane = threadIdx.x;
  const int entry = ((seq % kPeerRing) * world) * 4;
  const float my_sum = *local_sum * sum_scale;
  if (lane < world) {
    float* dst = pp.ring[lane] + entry + rank * 4;
    st_relaxed_sys_f32(dst + 0, my_sum);
    st_relaxed_sys_f32(dst + 1, my_n);
    st_release_sys_s32(reinterpret_cast<int*>(dst + 2), seq);
  }
  float s = 0.f, n = 0.f;
  bool ok = true;
  if (lane < world) {
    const float* src = pp.ring[rank] + entry + lane * 4;
    const long long t0 = clock64();
    while (ld_acquire_sys_s32(reinterpret_cast<const int*>(src + 2)) != seq) {
      if (clock64() - t0 > 4000000000ll) { ok = false; break; }  // ~2 s at 2 GHz
    }
    s = src[0];
    n = src[1];
  }
  ok = __all_sync(0xffffffffu, ok);
  // fixed order: rank 0, 1, 2, ... (bitwise the same total on every rank)
  float ts = 0.f, tn = 0.f;
  for (int r = 0; r < world; ++r) {
    ts += __shfl_sync(0xffffffffu, s, r);
    tn += __shfl_sync(0xffffffffu, n, r);
  }
  if (lane == 0) {
    const float bad = __int_as_float(0x7fc00000);
    pair[0] = ok ? ts : bad;
    pair[1] = ok ? tn : bad;
    pair[2] = ok ? ts / tn : bad;                  // global mean loss
    pair[3] = ok ? my_n * (float)world / tn : bad;  // "loss /= weight; loss *= world_size" relative to the local mean
    if (!ok && status) *status = 1;
  }
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_peer_exchange_ring_floats(int64_t world) {
  if (world < 1 || world > kPeerMaxWorld) return 0;
  return (size_t)(kPeerRing * world * 4);
}

extern "C" int b200_peer_sum_exchange(const uint64_t* ring_ptrs_host, int64_t world, int64_t rank, int64_t seq,
                                      const float* local_sum, float sum_scale, float local_count, float* pair,
                                      int32_t* status, void* stream) {
  B200_CHECK_ARG(ring_ptrs_host && pair && local_sum, B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  B200_CHECK_ARG(world >= 1 && world <= kPeerMaxWorld, B200_ERR_UNSUPPORTED_SHAPE, "world=%lld: 1..%d ranks of one box",
                 (long long)world, kPeerMaxWorld);
  B200_CHECK_ARG(rank >= 0 && rank < world, B200_ERR_INVALID_ARGUMENT, "rank=%lld outside [0,%lld)", (long long)rank,
                 (long long)world);
  B200_CHECK_ARG(seq >= 1 && seq < (1ll << 31), B200_ERR_INVALID_ARGUMENT, "seq=%lld must be in [1, 2^31)", (long long)seq);
  PeerPtrs pp;
  for (int r = 0; r < kPeerMaxWorld; ++r) pp.ring[r] = r < world ? reinterpret_cast<float*>(ring_ptrs_host[r]) : nullptr;
  for (int r = 0; r < world; ++r)
    B200_CHECK_ARG(pp.ring[r] != nullptr && (ring_ptrs_host[r] & 15) == 0, B200_ERR_MISALIGNED, "ring pointer of rank %d", r);
  peer_sum_exchange_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pp, (int)world, (int)rank, (int)seq, local_sum, sum_scale,
                                                               local_count, pair, status);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}
