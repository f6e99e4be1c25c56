// ctc_loss_host.inl -- argument checks, launch configuration and the extern "C" entry
// points of the CTC loss (included by ctc_loss.cu).
namespace b200 {

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static int elem_size(int dtype) { return dtype == B200_F32 ? 4 : 2; }

struct RowsCfg {
  int smem_bytes, grid, warps;
  bool vec;
};

// Shared memory of the streaming kernels is capped so that a lattice-sweep CTA of another
// utterance group can be resident on the same SM.
static const int kRowsSmemBudget = 176 * 1024;

static int plan_rows(const LossParams& p, bool grad_kernel, bool vec, RowsCfg* c) {
  c->vec = vec;
  const size_t per_warp = p.ws.Kp * sizeof(int) + (vec ? (grad_kernel ? 2 : 1) * kChunkBytes : 0);
  int warps = (int)(kRowsSmemBudget / per_warp);
  const int maxw = grad_kernel ? kGradMaxWarps : kRowsMaxWarps;
  if (warps > maxw) warps = maxw;
  if (warps < 1) {
    set_error("Lmax=%lld needs %zu bytes of shared memory per warp", (long long)p.Lmax, per_warp);
    return B200_ERR_UNSUPPORTED_SHAPE;
  }
  c->warps = warps;
  c->smem_bytes = (int)(per_warp * warps);
  const int nblk = (int)((p.T + kFramesPerTask - 1) / kFramesPerTask);
  int64_t g = ((int64_t)(p.b1 - p.b0) * nblk + warps - 1) / warps;
  const int sms = num_sms();
  if (g > sms) g = sms;
  if (g < 1) g = 1;
  c->grid = (int)g;
  return B200_OK;
}

template <typename T, bool VEC, bool GRAD>
static int launch_rows_t(const LossParams& p, const RowsCfg& c, cudaStream_t st) {
  if (GRAD) {
    auto kern = ctc_grad_kernel<T, VEC>;
    if (c.smem_bytes > 48 * 1024)
      B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, c.smem_bytes));
    kern<<<c.grid, c.warps * 32, c.smem_bytes, st>>>(p);
  } else {
    auto kern = ctc_rowstats_kernel<T, VEC>;
    if (c.smem_bytes > 48 * 1024)
      B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, c.smem_bytes));
    kern<<<c.grid, c.warps * 32, c.smem_bytes, st>>>(p);
  }
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

template <bool GRAD>
static int launch_rows(const LossParams& p, int dtype, const RowsCfg& c, cudaStream_t st) {
  switch (dtype) {
    case B200_F32:
      return c.vec ? launch_rows_t<float, true, GRAD>(p, c, st) : launch_rows_t<float, false, GRAD>(p, c, st);
    case B200_BF16:
      return c.vec ? launch_rows_t<__nv_bfloat16, true, GRAD>(p, c, st)
                   : launch_rows_t<__nv_bfloat16, false, GRAD>(p, c, st);
    case B200_F16:
      return c.vec ? launch_rows_t<__half, true, GRAD>(p, c, st) : launch_rows_t<__half, false, GRAD>(p, c, st);
  }
  set_error("unknown dtype %d", dtype);
  return B200_ERR_INVALID_ARGUMENT;
}

static int check_common(const void* logits, int dtype, int64_t T, int64_t B, int64_t V, int64_t stride_t,
                        int64_t stride_b, const int64_t* targets, const int64_t* tgt_offsets,
                        const int64_t* ilen, const int64_t* olen, int64_t Lmax, int blank, void* workspace,
                        size_t workspace_bytes) {
  B200_CHECK_ARG(dtype == B200_F32 || dtype == B200_BF16 || dtype == B200_F16, B200_ERR_INVALID_ARGUMENT,
                 "dtype must be B200_F32, B200_BF16 or B200_F16, got %d", dtype);
  B200_CHECK_ARG(T >= 1 && B >= 1 && V >= 2, B200_ERR_INVALID_ARGUMENT,
                 "need T>=1, B>=1, V>=2, got T=%lld B=%lld V=%lld", (long long)T, (long long)B, (long long)V);
  B200_CHECK_ARG(T < (1ll << 30) && B < (1ll << 30) && V < (1ll << 30), B200_ERR_INVALID_ARGUMENT,
                 "T, B, V must each be below 2^30");
  B200_CHECK_ARG(logits && ilen && olen && tgt_offsets && (targets || Lmax == 0), B200_ERR_INVALID_ARGUMENT,
                 "null pointer argument");
  B200_CHECK_ARG(stride_t >= 0 && stride_b >= 0, B200_ERR_INVALID_ARGUMENT, "negative logits stride");
  B200_CHECK_ARG(blank >= 0 && blank < V, B200_ERR_INVALID_ARGUMENT, "blank=%d outside [0,%lld)", blank,
                 (long long)V);
  B200_CHECK_ARG(Lmax >= 0, B200_ERR_INVALID_ARGUMENT, "Lmax=%lld is negative", (long long)Lmax);
  B200_CHECK_ARG(Lmax + 1 <= 1024, B200_ERR_UNSUPPORTED_SHAPE,
                 "Lmax=%lld: the lattice sweep supports at most 1023 labels per utterance", (long long)Lmax);
  B200_CHECK_ARG(workspace != nullptr, B200_ERR_INVALID_ARGUMENT, "workspace is null");
  B200_CHECK_ARG((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, B200_ERR_MISALIGNED,
                 "workspace must be 256-byte aligned");
  const size_t need = carve_loss_ws(nullptr, T, B, V, Lmax).bytes;
  B200_CHECK_ARG(workspace_bytes >= need, B200_ERR_WORKSPACE_TOO_SMALL, "workspace has %zu bytes, need %zu",
                 workspace_bytes, need);
  return B200_OK;
}

static bool rows_vec_ok(const void* base, int64_t V, int64_t st, int64_t sb, int es) {
  return aligned16(base) && (V * es) % 16 == 0 && (st * es) % 16 == 0 && (sb * es) % 16 == 0;
}

static LossParams make_params(const void* logits, int64_t T, int64_t B, int64_t V, int64_t stride_t,
                              int64_t stride_b, const int64_t* targets, const int64_t* tgt_offsets,
                              const int64_t* ilen, const int64_t* olen, int64_t Lmax, int blank, float* nll,
                              void* workspace) {
  LossParams p;
  memset(&p, 0, sizeof(p));
  p.logits = logits;
  p.T = T; p.B = B; p.V = V;
  p.stride_t = stride_t; p.stride_b = stride_b;
  p.targets = targets; p.tgt_offsets = tgt_offsets;
  p.ilen = ilen; p.olen = olen;
  p.Lmax = Lmax; p.blank = blank;
  p.nll = nll;
  p.ws = carve_loss_ws(workspace, T, B, V, Lmax);
  p.b0 = 0; p.b1 = (int)B;
  return p;
}

static int launch_prep(const LossParams& p, cudaStream_t st);
// Which lattice sweeps the fused call runs (b200_ctc_debug_force_log_sweep): 1 = log domain (the default: measured
// faster at C2, 50 + 42 us against 64 + 53 us for the float64 linear-domain sweeps and their patch kernel, which also
// cost the stream kernel 9 us for the float64 rows), 0 = linear domain with the log-domain redo pass.
static int g_force_log_sweep = 1;

static int run_forward(LossParams p, int dtype, float* loss_sum, float loss_scale, cudaStream_t st) {
  const int es = elem_size(dtype);
  RowsCfg cfg;
  int rc = plan_rows(p, false, rows_vec_ok(p.logits, p.V, p.stride_t, p.stride_b, es), &cfg);
  if (rc) return rc;
  B200_CUDA(cudaMemsetAsync(p.ws.bad, 0, 64 * sizeof(int), st));
  rc = launch_prep(p, st);
  if (rc) return rc;
  rc = launch_rows<false>(p, dtype, cfg, st);
  if (rc) return rc;
  p.loss_sum = loss_sum;  // written by the sweep CTA that finishes last
  p.loss_scale = loss_scale;
  rc = launch_sweep(p, st);
  if (rc) return rc;
  // occupancies for the backward pass (cheap, and keeps _backward a pure stream)
  return launch_occ(p, st);
}

static int launch_prep(const LossParams& p, cudaStream_t st) {
  int threads = (int)align_up((size_t)(p.Lmax > 0 ? p.Lmax : 1), 32);
  if (threads > 1024) threads = 1024;
  const size_t sm = sizeof(int) * (size_t)(p.Lmax > 0 ? p.Lmax : 1);
  B200_CUDA(launch_kernel(ctc_prep_kernel, dim3((unsigned)(p.b1 - p.b0)), dim3(threads), sm, st,
                          p.prep_after_stream != 0, p));
  return B200_OK;
}

// One pass over the logits (stream), the two lattice sweeps, and the column patch.
// Returns false in *done when the shape needs the general three-pass path.
static int run_fused(LossParams p, int dtype, float* loss_sum, float loss_scale, cudaStream_t st, bool* done,
                     bool stream_only = false) {
  *done = false;
  const int es = elem_size(dtype);
  StreamCfg sc;
  if (!rows_vec_ok(p.logits, p.V, p.stride_t, p.stride_b, es) ||
      !rows_vec_ok(p.grad, p.V, p.gstride_t, p.gstride_b, es) || !plan_stream(p, es, &sc))
    return B200_OK;
  p.loss_sum = loss_sum;  // (the prep kernel zeroes the workspace counters)
  p.loss_scale = loss_scale;
  // The stream kernel reads the int64 labels itself, so the 6 us label-prep kernel (int32 labels, duplicate
  // codes, length ranking, workspace counters: inputs of the sweeps and of the patch) is launched AFTER it
  // as a programmatic dependent and runs under it.  stream -> prep -> sweeps -> patch are all chained by
  // programmatic dependent launch; every grid ends with griddepcontrol.wait so that completion is transitive.
  p.prep_after_stream = 1;
  p.lin = (lin_sweep_ok(p) && !g_force_log_sweep) ? 1 : 0;
  int rc = launch_stream(p, dtype, sc, st);
  if (rc) return rc;
  if (stream_only) { *done = true; return B200_OK; }
  rc = launch_prep(p, st);
  if (rc) return rc;
  if (p.lin) {
    // linear-domain (float64) sweeps, then the patch kernel, which certifies every row it decodes; utterances it
    // cannot certify are redone by the log-domain sweeps and a second patch pass (both exit at once otherwise)
    rc = launch_sweep_lin(p, st, true);
    if (rc) return rc;
    rc = launch_patch(p, dtype, st, true);
    if (rc) return rc;
    p.redo_mode = 1;
  }
  rc = launch_sweep(p, st, true);
  if (rc) return rc;
  rc = launch_patch(p, dtype, st);
  if (rc) return rc;
  *done = true;
  return B200_OK;
}

static int run_backward(LossParams p, int dtype, cudaStream_t st) {
  const int es = elem_size(dtype);
  const bool vec = rows_vec_ok(p.logits, p.V, p.stride_t, p.stride_b, es) &&
                   rows_vec_ok(p.grad, p.V, p.gstride_t, p.gstride_b, es);
  RowsCfg cfg;
  int rc = plan_rows(p, true, vec, &cfg);
  if (rc) return rc;
  B200_CUDA(cudaMemsetAsync(p.ws.bad + 2, 0, sizeof(int), st));
  return launch_rows<true>(p, dtype, cfg, st);
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_ctc_loss_workspace_bytes(int64_t T, int64_t B, int64_t V, int64_t Lmax) {
  if (T < 1 || B < 1 || V < 1 || Lmax < 0) return 0;
  return carve_loss_ws(nullptr, T, B, V, Lmax).bytes;
}

extern "C" int b200_ctc_loss_forward(const void* logits, int dtype, int64_t T, int64_t B, int64_t V,
                                     int64_t stride_t, int64_t stride_b, const int64_t* targets,
                                     const int64_t* tgt_offsets, const int64_t* ilen, const int64_t* olen,
                                     int64_t Lmax, int blank, float* nll, float* loss_sum, float loss_scale,
                                     void* workspace, size_t workspace_bytes, void* stream) {
  int rc = check_common(logits, dtype, T, B, V, stride_t, stride_b, targets, tgt_offsets, ilen, olen, Lmax,
                        blank, workspace, workspace_bytes);
  if (rc) return rc;
  B200_CHECK_ARG(nll != nullptr, B200_ERR_INVALID_ARGUMENT, "nll is null");
  LossParams p = make_params(logits, T, B, V, stride_t, stride_b, targets, tgt_offsets, ilen, olen, Lmax,
                             blank, nll, workspace);
  return run_forward(p, dtype, loss_sum, loss_scale, (cudaStream_t)stream);
}

extern "C" int b200_ctc_loss_backward(const void* logits, int dtype, int64_t T, int64_t B, int64_t V,
                                      int64_t stride_t, int64_t stride_b, const int64_t* targets,
                                      const int64_t* tgt_offsets, const int64_t* ilen, const int64_t* olen,
                                      int64_t Lmax, int blank, const float* grad_weight, float grad_scale,
                                      void* grad_logits, int64_t gstride_t, int64_t gstride_b,
                                      void* workspace, size_t workspace_bytes, void* stream) {
  int rc = check_common(logits, dtype, T, B, V, stride_t, stride_b, targets, tgt_offsets, ilen, olen, Lmax,
                        blank, workspace, workspace_bytes);
  if (rc) return rc;
  B200_CHECK_ARG(grad_logits != nullptr, B200_ERR_INVALID_ARGUMENT, "grad_logits is null");
  B200_CHECK_ARG(gstride_t >= V || T == 1, B200_ERR_INVALID_ARGUMENT, "gradient rows overlap (gstride_t < V)");
  LossParams p = make_params(logits, T, B, V, stride_t, stride_b, targets, tgt_offsets, ilen, olen, Lmax,
                             blank, nullptr, workspace);
  p.grad = grad_logits;
  p.gstride_t = gstride_t; p.gstride_b = gstride_b;
  p.grad_weight = grad_weight; p.grad_scale = grad_scale;
  return run_backward(p, dtype, (cudaStream_t)stream);
}

extern "C" int b200_ctc_loss_fwd_bwd(const void* logits, int dtype, int64_t T, int64_t B, int64_t V,
                                     int64_t stride_t, int64_t stride_b, const int64_t* targets,
                                     const int64_t* tgt_offsets, const int64_t* ilen, const int64_t* olen,
                                     int64_t Lmax, int blank, const float* grad_weight, float grad_scale,
                                     float* nll, float* loss_sum, float loss_scale, void* grad_logits,
                                     int64_t gstride_t, int64_t gstride_b, void* workspace, size_t workspace_bytes,
                                     void* stream) {
  int rc = check_common(logits, dtype, T, B, V, stride_t, stride_b, targets, tgt_offsets, ilen, olen, Lmax,
                        blank, workspace, workspace_bytes);
  if (rc) return rc;
  B200_CHECK_ARG(nll != nullptr, B200_ERR_INVALID_ARGUMENT, "nll is null");
  B200_CHECK_ARG(grad_logits != nullptr, B200_ERR_INVALID_ARGUMENT, "grad_logits is null");
  B200_CHECK_ARG(gstride_t >= V || T == 1, B200_ERR_INVALID_ARGUMENT, "gradient rows overlap (gstride_t < V)");
  {
    LossParams p = make_params(logits, T, B, V, stride_t, stride_b, targets, tgt_offsets, ilen, olen, Lmax,
                               blank, nll, workspace);
    p.grad = grad_logits;
    p.gstride_t = gstride_t; p.gstride_b = gstride_b;
    p.grad_weight = grad_weight; p.grad_scale = grad_scale;
    bool done = false;
    rc = run_fused(p, dtype, loss_sum, loss_scale, (cudaStream_t)stream, &done);
    if (rc || done) return rc;
  }
  // general shapes: three passes (rowstats, lattice, gradient)
  rc = b200_ctc_loss_forward(logits, dtype, T, B, V, stride_t, stride_b, targets, tgt_offsets, ilen, olen,
                             Lmax, blank, nll, loss_sum, loss_scale, workspace, workspace_bytes, stream);
  if (rc) return rc;
  return b200_ctc_loss_backward(logits, dtype, T, B, V, stride_t, stride_b, targets, tgt_offsets, ilen, olen,
                                Lmax, blank, grad_weight, grad_scale, grad_logits, gstride_t, gstride_b,
                                workspace, workspace_bytes, stream);
}

extern "C" int b200_ctc_loss_stream_pass(const void* logits, int dtype, int64_t T, int64_t B, int64_t V,
                                         int64_t stride_t, int64_t stride_b, const int64_t* targets,
                                         const int64_t* tgt_offsets, const int64_t* ilen, const int64_t* olen,
                                         int64_t Lmax, int blank, const float* grad_weight, float grad_scale,
                                         void* grad_logits, int64_t gstride_t, int64_t gstride_b,
                                         void* workspace, size_t workspace_bytes, void* stream) {
  int rc = check_common(logits, dtype, T, B, V, stride_t, stride_b, targets, tgt_offsets, ilen, olen, Lmax,
                        blank, workspace, workspace_bytes);
  if (rc) return rc;
  B200_CHECK_ARG(grad_logits != nullptr, B200_ERR_INVALID_ARGUMENT, "grad_logits is null");
  LossParams p = make_params(logits, T, B, V, stride_t, stride_b, targets, tgt_offsets, ilen, olen, Lmax,
                             blank, nullptr, workspace);
  p.grad = grad_logits;
  p.gstride_t = gstride_t; p.gstride_b = gstride_b;
  p.grad_weight = grad_weight; p.grad_scale = grad_scale;
  bool done = false;
  rc = run_fused(p, dtype, nullptr, 1.0f, (cudaStream_t)stream, &done, true);
  if (rc) return rc;
  B200_CHECK_ARG(done, B200_ERR_UNSUPPORTED_SHAPE, "this shape runs on the three-pass path (no one-pass stream stage)");
  return B200_OK;
}

extern "C" int b200_ctc_debug_force_log_sweep(int on) {
  const int was = g_force_log_sweep;
  g_force_log_sweep = on ? 1 : 0;
  return was;
}

// the label side of a call laid out in one host buffer: targets | offsets (B) | ilen (B) | olen (B)
static int stage_labels_fill(const int64_t* targets_host, int64_t n_targets, const int64_t* tgt_offsets_host,
                             const int64_t* ilen_host, const int64_t* olen_host, int64_t B, int64_t* pinned_host) {
  B200_CHECK_ARG((targets_host || n_targets == 0) && ilen_host && olen_host && pinned_host,
                 B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  B200_CHECK_ARG(n_targets >= 0 && B >= 1, B200_ERR_INVALID_ARGUMENT, "bad sizes");
  int64_t* q = pinned_host;
  if (n_targets) memcpy(q, targets_host, sizeof(int64_t) * (size_t)n_targets);
  q += n_targets;
  if (tgt_offsets_host) {
    memcpy(q, tgt_offsets_host, sizeof(int64_t) * (size_t)B);
  } else {  // concatenated targets: exclusive prefix sum of the lengths
    int64_t run = 0;
    for (int64_t b = 0; b < B; ++b) { q[b] = run; run += olen_host[b] > 0 ? olen_host[b] : 0; }
  }
  memcpy(q + B, ilen_host, sizeof(int64_t) * (size_t)B);
  memcpy(q + 2 * B, olen_host, sizeof(int64_t) * (size_t)B);
  return B200_OK;
}

extern "C" int b200_ctc_stage_labels(const int64_t* targets_host, int64_t n_targets, const int64_t* tgt_offsets_host,
                                     const int64_t* ilen_host, const int64_t* olen_host, int64_t B,
                                     int64_t* pinned_host, int64_t* dev, void* stream) {
  B200_CHECK_ARG(dev != nullptr, B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  const int rc = stage_labels_fill(targets_host, n_targets, tgt_offsets_host, ilen_host, olen_host, B, pinned_host);
  if (rc) return rc;
  B200_CUDA(cudaMemcpyAsync(dev, pinned_host, sizeof(int64_t) * (size_t)(n_targets + 3 * B), cudaMemcpyHostToDevice,
                            (cudaStream_t)stream));
  return B200_OK;
}

// ---- the same with the pinned staging buffers owned by the library: a small ring per device, each slot guarded by
// an event recorded after its copy (a slot is rewritten only when that copy has completed)
namespace b200 {
// The staged labels travel by a KERNEL that reads the page-locked host buffer over PCIe (cudaHostAlloc memory is
// device-accessible under unified addressing) instead of a copy-engine transfer: between two kernels of one stream a
// cudaMemcpyAsync costs ~14 us of engine hand-over (measured at C2, tools/dev_gap.py), this kernel ~3 us, and as a
// programmatic dependent its launch hides under the tail of the grid before it.
__global__ void __launch_bounds__(1024) ctc_label_fetch_kernel(const int4* __restrict__ src, int4* __restrict__ dst, int n16) {
  pdl_launch_dependents();
  pdl_wait();  // earlier grids of the stream may still read the destination (a recycled allocation)
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += gridDim.x * blockDim.x) {
    int4 v;
    asm volatile("ld.volatile.global.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(src + i) : "memory");
    dst[i] = v;
  }
}
}  // namespace b200

namespace {
struct PinSlot {
  int64_t* buf = nullptr;
  size_t cap = 0;
  cudaEvent_t ev = nullptr;
  bool busy = false;
};
struct PinRing {
  PinSlot slot[16];
  int next = 0;
};
PinRing g_pin[64];
std::mutex g_pin_mu;
}  // namespace

extern "C" int b200_ctc_stage_labels_pinned(const int64_t* targets_host, int64_t n_targets, const int64_t* tgt_offsets_host,
                                            const int64_t* ilen_host, const int64_t* olen_host, int64_t B, int64_t* dev,
                                            void* stream) {
  B200_CHECK_ARG(n_targets >= 0 && B >= 1, B200_ERR_INVALID_ARGUMENT, "bad sizes");
  int d = 0;
  B200_CUDA(cudaGetDevice(&d));
  B200_CHECK_ARG(d >= 0 && d < 64, B200_ERR_INVALID_ARGUMENT, "device index %d outside [0,64)", d);
  B200_CHECK_ARG(dev != nullptr && (reinterpret_cast<uintptr_t>(dev) & 15) == 0, B200_ERR_INVALID_ARGUMENT,
                 "the device destination must be 16-byte aligned");
  const size_t n = (size_t)(n_targets + 3 * B);
  std::lock_guard<std::mutex> lk(g_pin_mu);
  PinRing& ring = g_pin[d];
  PinSlot& s = ring.slot[ring.next];
  ring.next = (ring.next + 1) % 16;
  if (s.busy) B200_CUDA(cudaEventSynchronize(s.ev));
  if (s.cap < n) {
    if (s.buf) B200_CUDA(cudaFreeHost(s.buf));
    s.buf = nullptr;
    s.cap = 0;
    const size_t cap = n < 16384 ? 16384 : n + n / 2 + 2;
    B200_CUDA(cudaHostAlloc((void**)&s.buf, cap * sizeof(int64_t), cudaHostAllocDefault));  // HOST memory
    s.cap = cap;
  }
  if (!s.ev) B200_CUDA(cudaEventCreateWithFlags(&s.ev, cudaEventDisableTiming));
  const int rc = stage_labels_fill(targets_host, n_targets, tgt_offsets_host, ilen_host, olen_host, B, s.buf);
  if (rc) return rc;
  const int n16 = (int)((n * sizeof(int64_t) + 15) / 16);  // (the pinned buffer holds >= n + 1 words; see the header for dev)
  B200_CUDA(launch_kernel(ctc_label_fetch_kernel, dim3((unsigned)((n16 + 1023) / 1024 > 8 ? 8 : (n16 + 1023) / 1024)), dim3(1024), 0,
                          (cudaStream_t)stream, true, reinterpret_cast<const int4*>(s.buf), reinterpret_cast<int4*>(dev), n16));
  B200_CUDA(cudaEventRecord(s.ev, (cudaStream_t)stream));
  s.busy = true;
  return B200_OK;
}

extern "C" void b200_ctc_debug_single_cta_sweep(int on) { g_single_cta_sweep = on ? 1 : 0; }

extern "C" int b200_ctc_loss_status(const void* workspace, int64_t T, int64_t B, int64_t V, int64_t Lmax,
                                    int32_t* n_invalid_host, int32_t* n_redone_host, void* stream) {
  B200_CHECK_ARG(workspace && n_invalid_host, B200_ERR_INVALID_ARGUMENT, "null pointer argument");
  B200_CHECK_ARG(T >= 1 && B >= 1 && V >= 1 && Lmax >= 0 && B < (1ll << 30), B200_ERR_INVALID_ARGUMENT, "bad sizes");
  const LossWs ws = carve_loss_ws(const_cast<void*>(workspace), T, B, V, Lmax);
  int* host = (int*)malloc(sizeof(int) * 2 * (size_t)B);
  B200_CHECK_ARG(host != nullptr, B200_ERR_INVALID_ARGUMENT, "out of host memory");
  cudaError_t e = cudaMemcpyAsync(host, ws.badutt, sizeof(int) * (size_t)B, cudaMemcpyDeviceToHost, (cudaStream_t)stream);
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(host + B, ws.redo, sizeof(int) * (size_t)B, cudaMemcpyDeviceToHost, (cudaStream_t)stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)stream);
  int n = 0, r = 0;
  if (e == cudaSuccess)
    for (int64_t b = 0; b < B; ++b) { n += host[b] ? 1 : 0; r += host[B + b] ? 1 : 0; }
  free(host);
  if (e != cudaSuccess) return cuda_fail(e, "b200_ctc_loss_status");
  *n_invalid_host = n;
  if (n_redone_host) *n_redone_host = r;
  return B200_OK;
}
