// ctc_prefix_host.inl -- extern "C" entry points of CTC prefix scoring.
using namespace b200;

// developer switch: the fp32 score kernels instead of the tensor-core one (A/B timing, parity of the fallback)
static int g_prefix_no_mma = 0;
extern "C" void b200_ctc_debug_prefix_fp32_score(int on) { g_prefix_no_mma = on; }

extern "C" int b200_ctc_prefix_init(float* x, const int64_t* xlens, int64_t B, int64_t T, int64_t V, int blank,
                                    float* r0, float* xmax, void* stream) {
  B200_CHECK_ARG(x && xlens && r0 && xmax, B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  B200_CHECK_ARG(B >= 1 && T >= 1 && V >= 2 && B < 65536, B200_ERR_INVALID_ARGUMENT,
                 "need 1<=B<65536, T>=1, V>=2, got B=%lld T=%lld V=%lld", (long long)B, (long long)T, (long long)V);
  B200_CHECK_ARG(blank >= 0 && blank < V, B200_ERR_INVALID_ARGUMENT, "blank=%d outside [0,%lld)", blank,
                 (long long)V);
  dim3 grid((unsigned)((V + 127) / 128), (unsigned)B);
  prefix_init_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(x, xlens, (int)T, (int)V, (int)B, blank, r0, xmax);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

extern "C" size_t b200_ctc_prefix_work_floats(int64_t T, int64_t n_bh, int64_t V) {
  if (T < 1 || n_bh < 1 || V < 1) return 0;
  // 2 * B*T*pad4(n_hyps) + 3*n_bh, bounded by pad4(n_hyps)*B <= 4*n_bh; then the queue of the exact kernel
  return (size_t)(2 * T * 4 * n_bh + 4 * n_bh + 4 + n_bh * V);
}

static int check_prefix(int64_t B, int64_t T, int64_t V, int64_t n_hyps, int blank, int64_t start, int64_t end) {
  B200_CHECK_ARG(B >= 1 && T >= 1 && V >= 2 && B < 65536, B200_ERR_INVALID_ARGUMENT,
                 "need 1<=B<65536, T>=1, V>=2, got B=%lld T=%lld V=%lld", (long long)B, (long long)T, (long long)V);
  B200_CHECK_ARG(n_hyps >= 1 && n_hyps < (1 << 20), B200_ERR_INVALID_ARGUMENT, "n_hyps=%lld", (long long)n_hyps);
  B200_CHECK_ARG(B * n_hyps * V < (1ll << 31), B200_ERR_UNSUPPORTED_SHAPE,
                 "n_bh * V = %lld does not fit 31 bits", (long long)(B * n_hyps * V));
  B200_CHECK_ARG(blank >= 0 && blank < V, B200_ERR_INVALID_ARGUMENT, "blank=%d outside [0,%lld)", blank,
                 (long long)V);
  B200_CHECK_ARG(start >= 1 && end <= T, B200_ERR_INVALID_ARGUMENT,
                 "frame window [%lld,%lld) must satisfy 1 <= start, end <= T=%lld", (long long)start,
                 (long long)end, (long long)T);
  return B200_OK;
}

// all candidates on the tensor cores (ctc_prefix_mma.inl): up to 16 hypotheses of every utterance per launch
static bool launch_full_mma(const ScoreParams& p, cudaStream_t st) {
  const bool vec4 = p.V % 4 == 0 && ((uintptr_t)p.x | (uintptr_t)p.xmax | (uintptr_t)p.scores | (uintptr_t)p.log_psi) % 16 == 0;
  if (!vec4 || g_prefix_no_mma == 1) return false;
  const size_t smem = score_mma_smem();
  dim3 grid((unsigned)((p.V + kMmaTile - 1) / kMmaTile), (unsigned)p.B + 1);  // row B: the repeated-last-label entries
#ifdef B200_PREFIX_DBG
  if (g_prefix_no_mma == 2) {  // developer build: the copies alone
    if (set_max_smem_cached((const void*)prefix_score_mma_kernel<1>, (int)smem) != cudaSuccess) return false;
    prefix_score_mma_kernel<1><<<grid, kMmaConsumers + 32, smem, st>>>(p);
    return true;
  }
#endif
  if (set_max_smem_cached((const void*)prefix_score_mma_kernel<0>, (int)smem) != cudaSuccess) return false;
  prefix_score_mma_kernel<0><<<grid, kMmaConsumers + 32, smem, st>>>(p);
  return true;
}

template <int NH>
static void launch_full(const ScoreParams& p, cudaStream_t st) {
  // four tokens per thread when the rows of x, xmax, scores and log_psi can be read / written 16 bytes at a time
  const bool vec4 = p.V % 4 == 0 && NH <= 16 && ((uintptr_t)p.x | (uintptr_t)p.xmax | (uintptr_t)p.scores | (uintptr_t)p.log_psi) % 16 == 0;
  if (vec4) {
    // bulk-copy ring (row B of the grid: the repeated-last-label entries)
    const size_t smem = score_ring_smem<NH>();
    if (set_max_smem_cached((const void*)prefix_score_ring_kernel<NH>, (int)smem) == cudaSuccess) {
      dim3 grid((unsigned)((p.V + kRingTile - 1) / kRingTile), (unsigned)p.B + 1);
      prefix_score_ring_kernel<NH><<<grid, kRingConsumers + 32, smem, st>>>(p);
      return;
    }
  }
  if (vec4) {
    // row B of the grid: the repeated-last-label entries
    dim3 grid((unsigned)((p.V / 4 + kScore4Threads - 1) / kScore4Threads), (unsigned)p.B + 1);
    prefix_score_full4_kernel<NH><<<grid, kScore4Threads, 0, st>>>(p);
    return;
  }
  dim3 grid((unsigned)((p.V + kScoreThreads - 1) / kScoreThreads), (unsigned)p.B);
  prefix_score_full_kernel<NH><<<grid, kScoreThreads, 0, st>>>(p);
  prefix_score_last_kernel<<<(unsigned)((p.n_bh + 3) / 4), 128, 0, st>>>(p);
}

extern "C" int64_t b200_ctc_prefix_column_stride(int64_t T) { return T < 1 ? 0 : (T + 31) & ~(int64_t)31; }

extern "C" int b200_ctc_prefix_columns(const float* x, int64_t B, int64_t T, int64_t V, float* x_cols, void* stream) {
  B200_CHECK_ARG(x && x_cols, B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  B200_CHECK_ARG(B >= 1 && T >= 1 && V >= 2 && B < 65536 && (T + 31) / 32 < 65536, B200_ERR_INVALID_ARGUMENT,
                 "need 1<=B<65536, 1<=T<2^21, V>=2, got B=%lld T=%lld V=%lld", (long long)B, (long long)T, (long long)V);
  dim3 grid((unsigned)((V + 31) / 32), (unsigned)((T + 31) / 32), (unsigned)B);
  prefix_columns_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, (int)T, (int)V, (int)b200_ctc_prefix_column_stride(T), x_cols);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

static int prefix_score_impl(const float* x, const float* x_cols, const float* xmax, const int64_t* xlens, int64_t B, int64_t T,
                             int64_t V, int64_t n_hyps, int blank, int eos, const float* r_prev,
                             const float* s_prev, const int64_t* last_ids, int64_t output_len,
                             const int64_t* scoring_ids, int64_t snum, int64_t start, int64_t end,
                             float* scores, float* log_psi, float* work, void* stream) {
  int rc = check_prefix(B, T, V, n_hyps, blank, start, end);
  if (rc) return rc;
  B200_CHECK_ARG(x && xmax && xlens && r_prev && last_ids && scores && log_psi && work,
                 B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  B200_CHECK_ARG(eos >= 0 && eos < V, B200_ERR_INVALID_ARGUMENT, "eos=%d outside [0,%lld)", eos, (long long)V);
  B200_CHECK_ARG(snum >= 0 && (snum == 0 || scoring_ids), B200_ERR_INVALID_ARGUMENT,
                 "snum=%lld needs scoring_ids", (long long)snum);
  cudaStream_t st = (cudaStream_t)stream;
  ScoreParams p;
  p.x = x; p.xmax = xmax;
  p.B = (int)B; p.T = (int)T; p.V = (int)V; p.n_hyps = (int)n_hyps; p.n_bh = (int)(B * n_hyps);
  p.blank = blank; p.eos = eos;
  p.r_prev = r_prev; p.s_prev = s_prev; p.last_ids = last_ids;
  p.output_len = (int)output_len; p.start = (int)start; p.end = (int)end;
  p.scoring_ids = scoring_ids; p.snum = (int)snum;
  p.scores = scores; p.log_psi = log_psi;
  p.w = carve_prefix(work, B, T, n_hyps);
  p.xc = x_cols; p.Tc = (int)b200_ctc_prefix_column_stride(T);
  p.xlens = xlens;
  if (snum > 0 && x_cols) {
    // candidates from the column copy: phi, fill, sums and the exact redo in one kernel, a CTA per hypothesis
    const size_t smem = sizeof(float) * 4 * (size_t)(end > start ? end - start : 0);
    if (smem <= 200u * 1024u &&
        (smem <= 40u * 1024u || set_max_smem_cached((const void*)prefix_score_cols_kernel, (int)smem) == cudaSuccess)) {
      prefix_score_cols_kernel<<<p.n_bh, kColsThreads, smem, st>>>(p, xlens);
      B200_CUDA(cudaGetLastError());
      return B200_OK;
    }
  }
  prefix_phi_kernel<<<p.n_bh, 128, 0, st>>>(r_prev, xlens, (int)T, (int)n_hyps, p.n_bh, p.start, p.end, p.w);
  B200_CUDA(cudaGetLastError());
  p.hyp_off = 0;
  if (snum == 0) {
    bool mma = true;
    for (; mma && p.hyp_off < (int)n_hyps; p.hyp_off += 16) mma = launch_full_mma(p, st);
    p.hyp_off = 0;
    if (mma) {
    } else if (n_hyps <= 4) launch_full<4>(p, st);
    else if (n_hyps <= 8) launch_full<8>(p, st);
    else if (n_hyps <= 12) launch_full<12>(p, st);
    else if (n_hyps <= 16) launch_full<16>(p, st);
    else {
      // wider beams (the LibriSpeech decode configs use beam_size 60): groups of 16 hypotheses, x is read once
      // per group
      for (; p.hyp_off < (int)n_hyps; p.hyp_off += 16) launch_full<16>(p, st);
      p.hyp_off = 0;
    }
  } else {
    dim3 grid((unsigned)((V + 1023) / 1024), (unsigned)p.n_bh);
    prefix_fill_kernel<<<grid, 256, 0, st>>>(p);
    B200_CUDA(cudaGetLastError());
    const int n = p.n_bh * p.snum;
    prefix_score_partial_kernel<<<(n + 7) / 8, 256, 0, st>>>(p);  // a warp per (hypothesis, candidate)
  }
  B200_CUDA(cudaGetLastError());
  // pairs whose scaled sum underflowed: exact log-space redo (an empty queue costs one launch)
  prefix_score_fix_kernel<<<2 * num_sms(), 256, 0, st>>>(p);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

extern "C" int b200_ctc_prefix_score(const float* x, const float* xmax, const int64_t* xlens, int64_t B, int64_t T,
                                     int64_t V, int64_t n_hyps, int blank, int eos, const float* r_prev,
                                     const float* s_prev, const int64_t* last_ids, int64_t output_len,
                                     const int64_t* scoring_ids, int64_t snum, int64_t start, int64_t end,
                                     float* scores, float* log_psi, float* work, void* stream) {
  return prefix_score_impl(x, nullptr, xmax, xlens, B, T, V, n_hyps, blank, eos, r_prev, s_prev, last_ids, output_len,
                           scoring_ids, snum, start, end, scores, log_psi, work, stream);
}

extern "C" int b200_ctc_prefix_score_columns(const float* x, const float* x_cols, const float* xmax, const int64_t* xlens,
                                             int64_t B, int64_t T, int64_t V, int64_t n_hyps, int blank, int eos,
                                             const float* r_prev, const float* s_prev, const int64_t* last_ids,
                                             int64_t output_len, const int64_t* scoring_ids, int64_t snum, int64_t start,
                                             int64_t end, float* scores, float* log_psi, float* work, void* stream) {
  return prefix_score_impl(x, x_cols, xmax, xlens, B, T, V, n_hyps, blank, eos, r_prev, s_prev, last_ids, output_len,
                           scoring_ids, snum, start, end, scores, log_psi, work, stream);
}

extern "C" int b200_ctc_prefix_select(const float* x, int64_t B, int64_t T, int64_t V, int64_t n_hyps, int blank,
                                      const float* r_prev, const float* log_psi, const int64_t* last_ids,
                                      int64_t output_len, int64_t start, int64_t end, int64_t n_bh,
                                      const int64_t* parent, const int64_t* token, int64_t n_new,
                                      const float* work, float* r_new, float* s_new, void* stream) {
  (void)work;
  B200_CHECK_ARG(x && r_prev && log_psi && last_ids && parent && r_new && s_new,
                 B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  B200_CHECK_ARG(token != nullptr || (n_new > 0 && n_new % B == 0), B200_ERR_INVALID_ARGUMENT,
                 "token is null (parent holds best_ids): n_new=%lld must be a multiple of B=%lld", (long long)n_new,
                 (long long)B);
  B200_CHECK_ARG(B >= 1 && T >= 1 && V >= 2 && n_hyps >= 1 && n_bh == B * n_hyps, B200_ERR_INVALID_ARGUMENT,
                 "inconsistent sizes B=%lld n_hyps=%lld n_bh=%lld", (long long)B, (long long)n_hyps, (long long)n_bh);
  B200_CHECK_ARG(start >= 1 && end <= T, B200_ERR_INVALID_ARGUMENT, "bad frame window [%lld,%lld)",
                 (long long)start, (long long)end);
  if (n_new <= 0) return B200_OK;
  SelectParams p;
  p.x = x; p.B = (int)B; p.T = (int)T; p.V = (int)V; p.n_hyps = (int)n_hyps; p.blank = blank;
  p.n_bh = (int)n_bh; p.n_new = (int)n_new;
  p.r_prev = r_prev; p.log_psi = log_psi; p.last_ids = last_ids;
  p.output_len = (int)output_len; p.start = (int)start; p.end = (int)end;
  p.parent = parent; p.token = token; p.r_new = r_new; p.s_new = s_new;
  p.beam = token ? 1 : (int)(n_new / B);
  prefix_select_warp_kernel<<<(unsigned)((n_new + 3) / 4), 128, 0, (cudaStream_t)stream>>>(p);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}
