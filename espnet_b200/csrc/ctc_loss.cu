// ctc_loss.cu -- fused log_softmax + CTC alpha/beta lattice + gradient for sm_100a.
// See ctc_loss.cuh for the kernel map and include/espnet_b200_ctc.h for the ABI.
#include "ctc_loss.cuh"

namespace b200 {

// ============================================================================
// prep: labels -> int32, duplicate chains (deterministic, atomics-free gradient)
// ============================================================================
__global__ void ctc_prep_kernel(LossParams p) {
  const int b = p.b0 + blockIdx.x;
  const int64_t L = p.olen[b];
  const int64_t* y = p.targets + p.tgt_offsets[b];
  int* lab = p.ws.labels + (int64_t)b * p.ws.Kp;
  int* nxt = p.ws.dup_next + (int64_t)b * p.ws.Kp;
  int* fst = p.ws.is_first + (int64_t)b * p.ws.Kp;
  extern __shared__ int s_lab[];
  for (int64_t i = threadIdx.x; i < L; i += blockDim.x) {
    int64_t v = y[i];
    if (v < 0 || v >= p.V || v == p.blank) {
      atomicExch(p.ws.bad, 1);
      v = v < 0 ? 0 : (v >= p.V ? p.V - 1 : v);
    }
    s_lab[i] = (int)v;
    lab[i] = (int)v;
  }
  __syncthreads();
  for (int64_t i = threadIdx.x; i < L; i += blockDim.x) {
    const int v = s_lab[i];
    int first = 1;
    for (int64_t j = 0; j < i; ++j)
      if (s_lab[j] == v) { first = 0; break; }
    int n = -1;
    for (int64_t j = i + 1; j < L; ++j)
      if (s_lab[j] == v) { n = (int)j; break; }
    fst[i] = first;
    nxt[i] = n;
  }
}

// ============================================================================
// sweep: the alpha / beta recursion, one CTA per (utterance, direction)
// ============================================================================
// Thread i owns the logical states 2i (blank) and 2i+1 (label i).  dir 1 runs
// the same recursion on the time- and label-reversed utterance, which is the
// beta recursion; its states are stored in that reversed (logical) order.
// Every frame the values are re-centred by floor(max of the previous frame) so
// that they stay O(10): in fp32 that keeps ~1e-6 absolute resolution in log2
// units however long the utterance is.  The integer offsets are summed exactly.
constexpr int kSweepU = 4;  // frames of lpg prefetched per batch

__global__ void __launch_bounds__(1024) ctc_sweep_kernel(LossParams p) {
  const int b = p.b0 + blockIdx.x;
  const int dir = blockIdx.y;
  const int Tb = (int)p.ilen[b];
  const int L = (int)p.olen[b];
  const int i = threadIdx.x;
  const int lane = i & 31, warp = i >> 5;
  const int nw = blockDim.x >> 5;
  const int64_t Kp = p.ws.Kp, Sp = p.ws.Sp;

  __shared__ float s_bnd[2][32];
  __shared__ float s_wmax[2][32];
  __shared__ float s_fin[2];

  if (Tb <= 0) {
    if (i == 0) {
      if (dir == 0) p.nll[b] = p.ws.nll_ws[b] = (L == 0) ? 0.0f : __int_as_float(0x7f800000);
      else p.ws.logz_beta[b] = (L == 0) ? 0.0f : B200_NEG;
    }
    return;
  }
  for (int k = i; k < 64; k += blockDim.x) { (&s_bnd[0][0])[k] = B200_NEG; (&s_wmax[0][0])[k] = B200_NEG; }
  __syncthreads();

  const int* lab = p.ws.labels + (int64_t)b * Kp;
  const bool has_label = i < L;
  const bool active = i <= L;
  const int pi = dir ? (L - 1 - i) : i;  // physical label position of logical i
  int my_lab = -1, prev_lab = -1;
  if (has_label) {
    my_lab = lab[pi];
    if (i >= 1) prev_lab = lab[dir ? pi + 1 : pi - 1];
  }
  const bool skip_ok = has_label && i >= 1 && my_lab != prev_lab;

  const float* lpg = p.ws.lpg + (int64_t)b * p.T * Kp;
  float* out = p.ws.lat + (((int64_t)dir * p.B + b) * p.T) * Sp + 2 * i;
  const int kcol = has_label ? 1 + pi : 0;

  float cb[kSweepU], cl[kSweepU], qb[kSweepU], ql[kSweepU];
  auto load_batch = [&](int base, float* vb, float* vl) {
#pragma unroll
    for (int j = 0; j < kSweepU; ++j) {
      const int tt = base + j;
      if (tt < Tb) {
        const int t = dir ? (Tb - 1 - tt) : tt;
        const float* row = lpg + (int64_t)t * Kp;
        vb[j] = __ldg(row);
        vl[j] = __ldg(row + kcol);
      } else {
        vb[j] = 0.f; vl[j] = 0.f;
      }
    }
  };

  // frame 0
  float ab = B200_NEG, al = B200_NEG;
  {
    const int t0 = dir ? (Tb - 1) : 0;
    const float* row = lpg + (int64_t)t0 * Kp;
    if (i == 0) {
      ab = __ldg(row);
      if (has_label) al = __ldg(row + kcol);
    }
    if (active) *reinterpret_cast<float2*>(out + (int64_t)t0 * Sp) = make_float2(ab, al);
    float m = warp_max_redux(fmaxf(ab, al));
    if (lane == 31) s_bnd[0][warp] = al;
    if (lane == 0) s_wmax[0][warp] = m;
  }
  int csum = 0;
  load_batch(1, cb, cl);
  __syncthreads();

  for (int base = 1; base < Tb; base += kSweepU) {
    load_batch(base + kSweepU, qb, ql);
#pragma unroll
    for (int j = 0; j < kSweepU; ++j) {
      const int tt = base + j;
      if (tt < Tb) {
        const int rd = (tt - 1) & 1, wr = tt & 1;
        // neighbour: label state of the previous logical position
        float nb = __shfl_up_sync(0xffffffffu, al, 1);
        if (lane == 0) nb = warp ? s_bnd[rd][warp - 1] : B200_NEG;
        // re-centring offset: floor(max over all states of the previous frame)
        const float pm = warp_max_redux(s_wmax[rd][lane]);
        const float off = floorf(pm);
        csum += (int)off;
        const float lb = cb[j] - off;
        const float ll = cl[j] - off;
        const float nab = lb + lse2(ab, nb);
        const float nal = ll + lse3(al, ab, skip_ok ? nb : B200_NEG);
        ab = active ? nab : B200_NEG;
        al = has_label ? nal : B200_NEG;
        const int t = dir ? (Tb - 1 - tt) : tt;
        if (active) *reinterpret_cast<float2*>(out + (int64_t)t * Sp) = make_float2(ab, al);
        const float m = warp_max_redux(fmaxf(ab, al));
        if (lane == 31) s_bnd[wr][warp] = al;
        if (lane == 0) s_wmax[wr][warp] = m;
        if (nw > 1) __syncthreads(); else __syncwarp();
      }
    }
#pragma unroll
    for (int j = 0; j < kSweepU; ++j) { cb[j] = qb[j]; cl[j] = ql[j]; }
  }

  // log2 p(y|x) = lse2(state 2L, state 2L-1) + sum of offsets
  if (i == L) s_fin[0] = ab;
  if (L == 0 && i == 0) s_fin[1] = B200_NEG;
  if (L > 0 && i == L - 1) s_fin[1] = al;
  __syncthreads();
  if (i == 0) {
    const float z = lse2(s_fin[0], s_fin[1]);
    const bool feasible = z > -1.0e29f;
    const float logz2 = z + (float)csum;
    if (dir == 0) p.nll[b] = p.ws.nll_ws[b] = feasible ? -B200_LN2 * logz2 : __int_as_float(0x7f800000);
    else p.ws.logz_beta[b] = feasible ? logz2 : B200_NEG;
  }
}

// deterministic sum of the per-utterance nll
__global__ void ctc_loss_sum_kernel(const float* nll, int B, float* out) {
  float s = 0.f;
  for (int b = threadIdx.x; b < B; b += 32) s += nll[b];
  s = warp_sum(s);
  if (threadIdx.x == 0) *out = s;
}

}  // namespace b200

#include "ctc_loss_rows.inl"
#include "ctc_loss_host.inl"
