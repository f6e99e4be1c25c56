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
//
// The label-gathered log-prob rows lpg[b][t][0..L] are staged into shared memory
// kSweepRing chunks ahead by 1-D bulk copies (cp.async.bulk -> UBLKCP) that signal
// an mbarrier, so the sequential loop never waits on global memory.
//
// Every frame the values are re-centred by an integer offset (a per-frame drift estimate
// plus, once per staged chunk, the measured maximum over all states), so the reduction is
// off the critical path.  Values therefore stay O(10..50) in log2 units and keep ~4e-6
// absolute resolution in fp32 however long the utterance is; the integer offsets are
// summed exactly.
constexpr int kSweepRing = 4;
constexpr int kSweepMaxChf = 8;  // frames per staged chunk (upper bound)

__global__ void __launch_bounds__(1024) ctc_sweep_kernel(LossParams p, int chf) {
  extern __shared__ __align__(128) unsigned char sw_smem[];
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
  uint64_t* full = reinterpret_cast<uint64_t*>(sw_smem);
  float* ring = reinterpret_cast<float*>(sw_smem + 128);
  const int chunk_floats = chf * (int)Kp;

  if (Tb <= 0) {
    if (i == 0) {
      if (dir == 0) p.nll[b] = p.ws.nll_ws[b] = (L == 0) ? 0.0f : __int_as_float(0x7f800000);
      else p.ws.logz_beta[b] = (L == 0) ? 0.0f : B200_NEG;
    }
    return;
  }
  for (int k = i; k < 64; k += blockDim.x) { (&s_bnd[0][0])[k] = B200_NEG; (&s_wmax[0][0])[k] = B200_NEG; }
  if (i == 0) {
    for (int s = 0; s < kSweepRing; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  __syncthreads();

  const float* lpg = p.ws.lpg + (int64_t)b * p.T * Kp;
  const int nchunks = (Tb + chf - 1) / chf;
  auto issue_chunk = [&](int c) {  // thread 0 only
    const int f0 = c * chf;
    const int n = min(chf, Tb - f0);
    const int phys0 = dir ? (Tb - f0 - n) : f0;
    const int slot = c % kSweepRing;
    const uint32_t bytes = (uint32_t)(n * Kp * sizeof(float));
    mbar_arrive_expect_tx(&full[slot], bytes);
    bulk_g2s(ring + (size_t)slot * chunk_floats, lpg + (int64_t)phys0 * Kp, bytes, &full[slot]);
  };
  if (i == 0)
    for (int c = 0; c < kSweepRing - 1 && c < nchunks; ++c) issue_chunk(c);

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
  float* out = p.ws.lat + (((int64_t)dir * p.B + b) * p.T) * Sp + 2 * i;
  const int kcol = has_label ? 1 + pi : 0;

  float ab = B200_NEG, al = B200_NEG;
  // Re-centring: every frame subtracts the integer `dstep`; once per chunk the level
  // (max over all states) is measured and removed by a one-shot `corr`, and dstep is
  // re-estimated from the measured drift.  csum is the exact sum of what was subtracted.
  int csum = 0;
  float dstep = 0.f, level_prev = 0.f, corr_prev = 0.f;
  float* outp = out + (int64_t)(dir ? (Tb - 1) : 0) * Sp;
  const int64_t ostep = dir ? -Sp : Sp;
  const int bsrc = warp ? warp - 1 : 0;

  for (int c = 0; c < nchunks; ++c) {
    const int slot = c % kSweepRing;
    const int f0 = c * chf;
    const int n = min(chf, Tb - f0);
    // fill the slot freed by the previous chunk (its last frame ended with a barrier)
    if (i == 0 && c + kSweepRing - 1 < nchunks) issue_chunk(c + kSweepRing - 1);
    mbar_wait(&full[slot], (uint32_t)((c / kSweepRing) & 1));
    const float* buf = ring + (size_t)slot * chunk_floats;
    float vb[kSweepMaxChf], vl[kSweepMaxChf];
#pragma unroll
    for (int f = 0; f < kSweepMaxChf; ++f) {
      if (f < n) {
        const float* rowp = buf + (size_t)(dir ? (n - 1 - f) : f) * Kp;
        vb[f] = rowp[0];
        vl[f] = has_label ? rowp[kcol] : B200_NEG;  // a missing label state stays at log(0)
      } else {
        vb[f] = 0.f; vl[f] = 0.f;
      }
    }
    if (c == 0) {
      // frame 0: only the first blank and the first label are reachable
      if (i == 0) { ab = vb[0]; al = vl[0]; }
      if (active) *reinterpret_cast<float2*>(outp) = make_float2(ab, al);
      outp += ostep;
      if (lane == 31) s_bnd[0][warp] = al;
      if (n > 1) dstep = floorf(vb[1]);
    } else {
      // publish this warp's level; it is picked up after the next barrier
      const float wm = warp_max_redux(fmaxf(ab, al));
      if (lane == 0) s_wmax[c & 1][warp] = wm;
    }
    if (nw > 1) __syncthreads(); else __syncwarp();
    float corr = 0.f;
    if (c > 0) {
      const float level = warp_max_redux(s_wmax[c & 1][lane]);
      corr = floorf(level);
      // drift per frame seen over the last chunk, relative to the dstep that was applied
      dstep = floorf(dstep + (level - level_prev + corr_prev) / (float)chf + 0.5f);
      level_prev = level;
      corr_prev = corr;
    }
#pragma unroll
    for (int f = 0; f < kSweepMaxChf; ++f) {
      if (f < n && !(c == 0 && f == 0)) {
        const int tt = f0 + f;
        const int rd = (tt - 1) & 1, wr = tt & 1;
        const float o = dstep + ((f == 0 || (c == 0 && f == 1)) ? corr : 0.f);
        float nb = __shfl_up_sync(0xffffffffu, al, 1);
        const float bn = s_bnd[rd][bsrc];
        if (lane == 0) nb = warp ? bn : B200_NEG;
        const float x = lse2(ab, nb);
        const float y = lse3(al, ab, skip_ok ? nb : B200_NEG);
        ab = (vb[f] - o) + x;
        al = (vl[f] - o) + y;
        csum += (int)o;
        if (lane == 31) s_bnd[wr][warp] = al;
        if (active) *reinterpret_cast<float2*>(outp) = make_float2(ab, al);
        outp += ostep;
        if (nw > 1) __syncthreads(); else __syncwarp();
      }
    }
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
#include "ctc_loss_gamma.inl"
#include "ctc_loss_host.inl"
