// ctc_prefix_mma.inl -- the full-vocabulary score kernel of CTC prefix scoring on the tensor cores.
//
// For one utterance the scaled sums of ctc_prefix_score.py:144-148,
//     acc[h][c] = sum_t E[t-1][h] * 2^(x[t][c] * log2e - max_t x[.][c] * log2e),
// are a (hypotheses x frames) x (frames x tokens) product with at most 16 rows.  The fp32 kernels above spend one
// FFMA per (hypothesis, frame, token) -- ~100 issue slots per thread and frame for 16 bytes of x, 57 % issue
// utilisation at 3 TB/s (ncu, r02) -- so the product goes to mma.sync m16n8k8 here: rows = hypotheses, k = 8
// frames, n = 8 tokens, ~6.5 issue slots per element of x.  TF32 operands carry 11 significant bits, not enough
// for the 1e-4 bar on log_psi, so both operands are split into a TF32 head and a TF32 tail and three products are
// accumulated in fp32 (head*head + tail*head + head*tail; what is dropped is below 2^-21 relative per term).
//
// The rows of x arrive as in prefix_score_ring_kernel: one producer lane, 1-D bulk copies of a tile's rows into a
// ring of 8-frame slots (rows padded by 8 floats: the B fragments of a warp then read 32 distinct banks).
namespace b200 {

// tokens per CTA.  192 (three consumer warps) instead of 256: at C3 (V = 5000, 32 utterances) that is 864 CTAs of
// 32 KB, six on most SMs and five on the rest, against 660 CTAs of 40 KB that put five on some SMs and four on
// the others -- the SMs with the most CTAs set the kernel's time (C3 step: 150 -> 144.5 us; 128-token tiles: 146.4).
constexpr int kMmaTile = 192;
constexpr int kMmaWarpTok = 64;        // tokens per consumer warp (8 n-tiles)
constexpr int kMmaNT = kMmaWarpTok / 8;
constexpr int kMmaConsumers = (kMmaTile / kMmaWarpTok) * 32;
constexpr int kMmaRow = kMmaTile + 8;  // floats between two rows of a slot
constexpr int kMmaSlots = 4;
constexpr int kMmaTB = 64;             // frames of E staged per block
constexpr int kMmaES = 24;             // floats between two frames of the staged E (16 used)

static size_t score_mma_smem() {
  return 256 + sizeof(float) * kMmaTB * kMmaES + sizeof(float) * (size_t)kMmaSlots * 8 * kMmaRow;
}

__device__ __forceinline__ float mma_lds_f32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ float4 mma_lds_v4(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// v = head + tail with head = the 19 leading bits of v (exactly a TF32 value); the tail is exact in fp32
__device__ __forceinline__ void tf32_split(float v, uint32_t& head, uint32_t& tail) {
  head = __float_as_uint(v) & 0xffffe000u;
  tail = __float_as_uint(v - __uint_as_float(head));
}

template <int DBG>
__global__ void __launch_bounds__(kMmaConsumers + 32, 7) prefix_score_mma_kernel(ScoreParams p) {
  extern __shared__ __align__(128) unsigned char mm_smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(mm_smem);  // [kMmaSlots]
  uint64_t* empty = full + kMmaSlots;                     // [kMmaSlots]
  float* sE = reinterpret_cast<float*>(mm_smem + 256);    // [kMmaTB][kMmaES]
  float* ring = sE + kMmaTB * kMmaES;                 // [kMmaSlots][8][kMmaRow]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NCW = kMmaConsumers / 32;
  if (blockIdx.y == (unsigned)p.B) {
    // one more row of CTAs: the repeated-last-label entries, a warp per hypothesis (the token CTAs skip them)
    if (p.hyp_off != 0 || warp >= NCW) return;
    const int nwp = gridDim.x * NCW;
    for (int h = blockIdx.x * NCW + warp; h < p.n_bh; h += nwp) prefix_score_last_warp(p, h, lane);
    return;
  }
  const int u = blockIdx.y;
  const int tile0 = blockIdx.x * kMmaTile;
  const int ntok = min(kMmaTile, p.V - tile0);  // V % 4 == 0
  const uint32_t row_bytes = (uint32_t)ntok * 4u;
  const int nfr = max(p.end - p.start, 0);
  const int nstage = (nfr + 7) / 8;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kMmaSlots; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NCW); }
    mbar_fence_init();
    fence_proxy_async();
  }
  __syncthreads();

  if (warp == NCW) {
    // ------------------------------------------------------------------ producer
    if (lane == 0) {
      const float* xu = p.x + ((int64_t)u * p.T + p.start) * p.V + tile0;
      int slot = 0;
      uint32_t par = 0;  // parity of the `empty` completion that frees `slot` for its next use
      for (int s = 0; s < nstage; ++s) {
        if (s >= kMmaSlots) mbar_wait(&empty[slot], par);
        const int n = min(8, nfr - s * 8);
        mbar_arrive_expect_tx(&full[slot], (uint32_t)n * row_bytes);
        for (int k = 0; k < n; ++k)
          bulk_g2s(ring + ((size_t)slot * 8 + k) * kMmaRow, xu + (int64_t)(s * 8 + k) * p.V, row_bytes, &full[slot]);
        if (++slot == kMmaSlots) { slot = 0; if (s >= kMmaSlots) par ^= 1u; }
      }
    }
    return;
  }

  // -------------------------------------------------------------------- consumers
  // Token layout of a warp's 64 tokens over the eight n-tiles: column `gid` of n-tile (half, j) is token
  // wtok + 32*half + 4*gid + j, so that a thread's four tokens of a half are CONTIGUOUS: one 16-byte shared-memory
  // load per (row, half) instead of four 4-byte ones, and float4 stores in the epilogue.  (The MMA does not care which
  // token a column stands for.)  With rows 264 floats apart the eight lanes of a quarter-warp read eight distinct
  // 16-byte bank groups.
  const int tid = threadIdx.x;          // 0 .. kMmaConsumers-1
  const int gid = lane >> 2, tig = lane & 3;
  const int wtok = warp * kMmaWarpTok;  // first token of this warp inside the tile
  const int NHP = p.w.NHP;
  float mxl[kMmaNT];
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int c = tile0 + wtok + 32 * half + 4 * gid;
    float4 m = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < p.V) m = *reinterpret_cast<const float4*>(p.xmax + (int64_t)u * p.V + c);
    mxl[4 * half + 0] = m.x * B200_LOG2E; mxl[4 * half + 1] = m.y * B200_LOG2E;
    mxl[4 * half + 2] = m.z * B200_LOG2E; mxl[4 * half + 3] = m.w * B200_LOG2E;
  }
  float acc[kMmaNT][4];
#pragma unroll
  for (int nt = 0; nt < kMmaNT; ++nt)
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[nt][q] = 0.f;
  const uint32_t b_off = (uint32_t)((tig * kMmaRow + wtok + 4 * gid) * 4);
  const uint32_t ring_u32 = smem_u32(ring);

  int slot = 0;
  uint32_t par = 0;
  for (int tb = p.start; tb < p.end; tb += kMmaTB) {
    const int nt_blk = min(kMmaTB, p.end - tb);
    const float* sEb = sE;
    asm volatile("bar.sync 1, %0;" ::"n"(kMmaConsumers) : "memory");
    // E rows for frames tb-1 .. tb+nt_blk-2 (term t uses phi[t-1]); zero beyond the block and beyond the hypotheses.
    // (Double-buffered cp.async copies of 32-frame blocks, one block ahead, were SLOWER: 156.5 vs 151 us per C3 step.)
    for (int i = tid; i < kMmaTB * 16; i += kMmaConsumers) {
      const int f = i >> 4, j = (i & 15) + p.hyp_off;
      sE[f * kMmaES + (i & 15)] = (f < nt_blk && j < NHP) ? p.w.EA[((int64_t)u * p.T + (tb - 1 + f)) * NHP + j] : 0.f;
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kMmaConsumers) : "memory");
    for (int f0 = 0; f0 < nt_blk; f0 += 8) {
      // A fragments: hypotheses gid / gid + 8, frames f0 + tig / f0 + tig + 4
      uint32_t ah[4], al[4];
      {
        const float* e = sEb + (f0 + tig) * kMmaES + gid;
        tf32_split(e[0], ah[0], al[0]);
        tf32_split(e[8], ah[1], al[1]);
        tf32_split(e[4 * kMmaES], ah[2], al[2]);
        tf32_split(e[4 * kMmaES + 8], ah[3], al[3]);
      }
      const int nrow = nt_blk - f0;  // rows of the slot that hold frames (8 except in the last slot)
      mbar_wait(&full[slot], par);
      const uint32_t base = ring_u32 + (uint32_t)(slot * 8 * kMmaRow * 4) + b_off;
      // half by half, phase by phase (the asm statements keep this order): the loads, the exponentials, the splits,
      // and 12 MMAs that alternate between four accumulators
      const bool masked = nrow < 8;  // last slot: rows >= nrow were never written by this call (E is zero there,
                                     // but 0 * NaN is not)
      const bool keep0 = tig < nrow, keep1 = tig + 4 < nrow;
#pragma unroll
      for (int half = 0; half < (DBG == 1 ? 0 : 2); ++half) {
        const float4 r0 = mma_lds_v4(base + half * 128);                     // frame tig, tokens j = 0..3
        const float4 r1 = mma_lds_v4(base + half * 128 + 4 * kMmaRow * 4);   // frame tig + 4
        float xr[8] = {r0.x, r1.x, r0.y, r1.y, r0.z, r1.z, r0.w, r1.w};      // [2*j + {0,1}]
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          xr[2 * q] = ex2f(fmaf(xr[2 * q], B200_LOG2E, -mxl[4 * half + q]));
          xr[2 * q + 1] = ex2f(fmaf(xr[2 * q + 1], B200_LOG2E, -mxl[4 * half + q]));
        }
        if (masked) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (!keep0) xr[2 * q] = 0.f;
            if (!keep1) xr[2 * q + 1] = 0.f;
          }
        }
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) tf32_split(xr[q], hi[q], lo[q]);
#pragma unroll
        for (int q = 0; q < 4; ++q) mma_tf32(acc[4 * half + q], al, hi[2 * q], hi[2 * q + 1]);
#pragma unroll
        for (int q = 0; q < 4; ++q) mma_tf32(acc[4 * half + q], ah, lo[2 * q], lo[2 * q + 1]);
#pragma unroll
        for (int q = 0; q < 4; ++q) mma_tf32(acc[4 * half + q], ah, hi[2 * q], hi[2 * q + 1]);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
      if (++slot == kMmaSlots) { slot = 0; par ^= 1u; }
    }
  }

  if (DBG == 1) {  // (developer variant without the arithmetic: keep the exact kernel's queue empty)
#pragma unroll
    for (int nt = 0; nt < kMmaNT; ++nt)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[nt][q] = 1.f;
  }
  // ---- D fragments: hypotheses gid / gid + 8; column 2*tig + q of n-tile (half, j) is token
  //      wtok + 32*half + 4*(2*tig + q) + j: for fixed (half, q) the four j are contiguous
#pragma unroll
  for (int hrow = 0; hrow < 2; ++hrow) {
    const int jh = gid + 8 * hrow + p.hyp_off;
    if (jh >= p.n_hyps) continue;
    const int h = u * p.n_hyps + jh;
    const float sp = p.s_prev ? p.s_prev[h] : 0.f;
    const int64_t last = p.last_ids[h];
    const float ma = p.w.MA[h];
    const float rend = p.w.REND[h];
#pragma unroll
    for (int half = 0; half < 2; ++half) {
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int c0 = tile0 + wtok + 32 * half + 8 * tig + 4 * q;
        if (c0 >= p.V) continue;  // (V % 4 == 0: all four tokens are inside or outside)
        const float4 mx4 = *reinterpret_cast<const float4*>(p.xmax + (int64_t)u * p.V + c0);
        float4 x04 = make_float4(B200_LZ, B200_LZ, B200_LZ, B200_LZ);  // r[0,n] = x[0,c]
        if (p.output_len == 0) x04 = *reinterpret_cast<const float4*>(p.x + (int64_t)u * p.T * p.V + c0);
        const float mxs[4] = {mx4.x, mx4.y, mx4.z, mx4.w}, x0s[4] = {x04.x, x04.y, x04.z, x04.w};
        float psi[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c = c0 + j;
          const float a = acc[4 * half + j][2 * hrow + q];
          if (c == p.blank) {
            psi[j] = B200_LZ;
          } else if (c == p.eos) {
            psi[j] = rend;
          } else {
            psi[j] = finish_psi(a, ma + mxs[j], x0s[j]);
            if ((int64_t)c != last) score_flag(p, a, h, c, ma);
          }
        }
        float* lp = p.log_psi + (int64_t)h * p.V + c0;
        float* sc = p.scores + (int64_t)h * p.V + c0;
        if (last >= c0 && last < c0 + 4 && last != p.blank && last != p.eos) {
          // the repeated-last-label entry belongs to the extra row of CTAs
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (c0 + j != (int)last) {
              lp[j] = psi[j];
              sc[j] = psi[j] - sp;
            }
        } else {
          *reinterpret_cast<float4*>(lp) = make_float4(psi[0], psi[1], psi[2], psi[3]);
          *reinterpret_cast<float4*>(sc) = make_float4(psi[0] - sp, psi[1] - sp, psi[2] - sp, psi[3] - sp);
        }
      }
    }
  }
}

}  // namespace b200
