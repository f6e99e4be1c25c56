/*
 * oracle/ctc_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C CPU restatement of the two recursions on ESPnet's CTC hot path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
 * arm may build, load or call this file.  The product (espnet_b200/) never does.
 *
 * What it restates (paths relative to the reference root, kan-bayashi/espnet v0.9.5):
 *
 *  1. CTC loss + gradient w.r.t. pre-softmax logits: what
 *     espnet2/asr/ctc.py:46-52 (CTC.loss_fn) computes, i.e. log_softmax(2) followed
 *     by torch.nn.CTCLoss(reduction=sum|none) and its backward.  The lattice
 *     arithmetic itself is NOT in the reference tree: it lives in the third-party
 *     dependency PyTorch (pinned by the reference at TH_VERSION := 1.4.0,
 *     tools/Makefile:7; torch 2.11.0 in this image).  The recursion below is the
 *     published algorithm (Graves et al., ICML 2006, eq. 6-8, 10-11, 16) in log
 *     space with blank = 0 default, the same one torch.nn.CTCLoss documents.
 *  2. CTC prefix scoring: espnet/nets/ctc_prefix_score.py:69-188
 *     (CTCPrefixScoreTH.__call__), :190-221 (index_select_state), :23-67
 *     (constructor padding) and the single-hypothesis numpy version :224-310.
 *  3. Viterbi forced alignment: espnet/nets/pytorch_backend/ctc.py:153-216.
 *
 * Parity pinning: the reference holds no golden vectors for this path
 * (SURVEY.md section 4).  This oracle is pinned against outputs of the reference
 * itself (espnet2 CTC.loss_fn, CTCPrefixScoreTH, CTCPrefixScore, BatchBeamSearch)
 * generated in the build container by tests/golden/make_golden.py and committed
 * under tests/golden/ (see tests/test_oracle_vs_golden.py).
 *
 * Build: make -C oracle   (gcc -O2 -fopenmp -shared -fPIC)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_LOGZERO (-10000000000.0)

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* ------------------------------------------------------------------ */
/* log-sum-exp helpers, double and float flavours                      */
/* ------------------------------------------------------------------ */
static inline double lse2_d(double a, double b) {
  if (a == -INFINITY) return b;
  if (b == -INFINITY) return a;
  double m = a > b ? a : b;
  return m + log(exp(a - m) + exp(b - m));
}
static inline double lse3_d(double a, double b, double c) {
  double m = a > b ? a : b;
  if (c > m) m = c;
  if (m == -INFINITY) return -INFINITY;
  return m + log(exp(a - m) + exp(b - m) + exp(c - m));
}
static inline float lse2_f(float a, float b) {
  if (a == -INFINITY) return b;
  if (b == -INFINITY) return a;
  float m = a > b ? a : b;
  return m + logf(expf(a - m) + expf(b - m));
}
static inline float lse3_f(float a, float b, float c) {
  float m = a > b ? a : b;
  if (c > m) m = c;
  if (m == -INFINITY) return -INFINITY;
  return m + logf(expf(a - m) + expf(b - m) + expf(c - m));
}

/* ------------------------------------------------------------------ */
/* 1. CTC loss and gradient (double accumulation)                      */
/* ------------------------------------------------------------------ */
/*
 * logits : float (T,B,V) view with element strides (stride_t, stride_b, 1)
 * targets: int64, utterance b's labels at targets[tgt_offsets[b] + i], i < olen[b]
 * nll    : double (B,) out, -log p(y|x) in nats, +inf when infeasible
 * grad   : double (T,B,V) contiguous out or NULL:
 *            grad[t,b,v] = grad_scale * gw[b] * (softmax[t,b,v] - occupancy[t,b,v]), t < ilen[b]
 *            0 for t >= ilen[b]; NaN rows for infeasible utterances (torch behaviour)
 * Follows SURVEY.md 7.0 / espnet2/asr/ctc.py:46-52 + torch.nn.CTCLoss.
 */
int oracle_ctc_loss(const float* logits, int64_t T, int64_t B, int64_t V,
                    int64_t stride_t, int64_t stride_b,
                    const int64_t* targets, const int64_t* tgt_offsets,
                    const int64_t* ilen, const int64_t* olen, int blank,
                    const double* gw, double grad_scale,
                    double* nll, double* grad) {
  int status = 0;
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t b = 0; b < B; ++b) {
    const int64_t Tb = ilen[b], L = olen[b], S = 2 * L + 1;
    const int64_t* y = targets + tgt_offsets[b];
    if (Tb > T || Tb < 0 || L < 0) { status = 1; continue; }
    double* lp = (double*)malloc(sizeof(double) * (size_t)(Tb > 0 ? Tb : 1) * V);
    double* alpha = (double*)malloc(sizeof(double) * (size_t)(Tb > 0 ? Tb : 1) * S);
    double* beta = (double*)malloc(sizeof(double) * (size_t)(Tb > 0 ? Tb : 1) * S);
    /* log_softmax over the vocabulary, espnet2/asr/ctc.py:48 */
    for (int64_t t = 0; t < Tb; ++t) {
      const float* row = logits + t * stride_t + b * stride_b;
      double m = -INFINITY;
      for (int64_t v = 0; v < V; ++v) if (row[v] > m) m = row[v];
      double s = 0.0;
      for (int64_t v = 0; v < V; ++v) s += exp((double)row[v] - m);
      double lse = m + log(s);
      for (int64_t v = 0; v < V; ++v) lp[t * V + v] = (double)row[v] - lse;
    }
#define EXT(s_) (((s_) & 1) ? (int)y[(s_) >> 1] : blank)
    double loglik = -INFINITY;
    if (Tb > 0) {
      /* alpha: Graves eq. 6-7 */
      for (int64_t s = 0; s < S; ++s) alpha[s] = -INFINITY;
      alpha[0] = lp[blank];
      if (S > 1) alpha[1] = lp[EXT(1)];
      for (int64_t t = 1; t < Tb; ++t) {
        const double* ap = alpha + (t - 1) * S;
        double* ac = alpha + t * S;
        for (int64_t s = 0; s < S; ++s) {
          double a0 = ap[s];
          double a1 = s >= 1 ? ap[s - 1] : -INFINITY;
          double a2 = (s >= 2 && (s & 1) && EXT(s) != EXT(s - 2)) ? ap[s - 2] : -INFINITY;
          ac[s] = lse3_d(a0, a1, a2) + lp[t * V + EXT(s)];
        }
      }
      loglik = lse2_d(alpha[(Tb - 1) * S + S - 1],
                      S > 1 ? alpha[(Tb - 1) * S + S - 2] : -INFINITY);
    } else if (L == 0) {
      loglik = 0.0;
    }
    nll[b] = -loglik;
    if (grad) {
      const double w = grad_scale * (gw ? gw[b] : 1.0);
      for (int64_t t = Tb; t < T; ++t)
        for (int64_t v = 0; v < V; ++v) grad[(t * B + b) * V + v] = 0.0;
      if (Tb > 0 && loglik == -INFINITY) {
        for (int64_t t = 0; t < Tb; ++t)
          for (int64_t v = 0; v < V; ++v) grad[(t * B + b) * V + v] = NAN;
      } else if (Tb > 0) {
        /* beta: Graves eq. 10-11 (beta includes the emission at t, as in torch) */
        double* bl = beta + (Tb - 1) * S;
        for (int64_t s = 0; s < S; ++s) bl[s] = -INFINITY;
        bl[S - 1] = lp[(Tb - 1) * V + blank];
        if (S > 1) bl[S - 2] = lp[(Tb - 1) * V + EXT(S - 2)];
        for (int64_t t = Tb - 2; t >= 0; --t) {
          const double* bn = beta + (t + 1) * S;
          double* bc = beta + t * S;
          for (int64_t s = 0; s < S; ++s) {
            double b0 = bn[s];
            double b1 = s + 1 < S ? bn[s + 1] : -INFINITY;
            double b2 = (s + 2 < S && (s & 1) && EXT(s) != EXT(s + 2)) ? bn[s + 2] : -INFINITY;
            bc[s] = lse3_d(b0, b1, b2) + lp[t * V + EXT(s)];
          }
        }
        /* gradient: Graves eq. 16 with the log_softmax Jacobian folded in */
        double* acc = (double*)malloc(sizeof(double) * V);
        for (int64_t t = 0; t < Tb; ++t) {
          for (int64_t v = 0; v < V; ++v) acc[v] = -INFINITY;
          for (int64_t s = 0; s < S; ++s) {
            int c = EXT(s);
            acc[c] = lse2_d(acc[c], alpha[t * S + s] + beta[t * S + s]);
          }
          for (int64_t v = 0; v < V; ++v) {
            double l = lp[t * V + v];
            double occ = acc[v] == -INFINITY ? 0.0 : exp(acc[v] - loglik - l);
            grad[(t * B + b) * V + v] = w * (exp(l) - occ);
          }
        }
        free(acc);
      }
    }
#undef EXT
    free(lp); free(alpha); free(beta);
  }
  return status;
}

/* ------------------------------------------------------------------ */
/* 2. CTC prefix scoring                                                */
/* ------------------------------------------------------------------ */

/* Constructor side effects, espnet/nets/ctc_prefix_score.py:46-51, and the
 * initial state :84-92.  x (B,T,V) float is modified in place.
 * r0 (T,2,B) float: [:,0,:]=logzero, [:,1,b]=cumsum_t x[b,t,blank]. */
void oracle_prefix_init(float* x, const int64_t* xlens, int64_t B, int64_t T, int64_t V,
                        int blank, float* r0) {
  for (int64_t b = 0; b < B; ++b) {
    for (int64_t t = xlens[b]; t < T; ++t) {
      float* row = x + (b * T + t) * V;
      for (int64_t v = 0; v < V; ++v) row[v] = (float)ORACLE_LOGZERO;
      row[blank] = 0.0f;
    }
    float c = 0.0f;
    for (int64_t t = 0; t < T; ++t) {
      c += x[(b * T + t) * V + blank];
      r0[(t * 2 + 0) * B + b] = (float)ORACLE_LOGZERO;
      r0[(t * 2 + 1) * B + b] = c;
    }
  }
}

/*
 * One call of CTCPrefixScoreTH.__call__ (ctc_prefix_score.py:69-188), float
 * arithmetic as the reference's dtype=float32 path.
 *
 * x (B,T,V) padded log-probs; r_prev (T,2,n_bh); s_prev (n_bh,) or NULL (= 0);
 * last_ids (n_bh,); scoring_ids (n_bh,snum) or NULL with snum = 0 (all V).
 * Outputs: scores (n_bh,V) = log_psi - s_prev; log_psi (n_bh,V);
 *          r_full (T,2,n_bh,snum_eff) or NULL, snum_eff = snum ? snum : V.
 */
int oracle_prefix_score(const float* x, const int64_t* xlens,
                        int64_t B, int64_t T, int64_t V, int64_t n_hyps,
                        int blank, int eos,
                        const float* r_prev, const float* s_prev,
                        const int64_t* last_ids, int64_t output_len,
                        const int64_t* scoring_ids, int64_t snum,
                        int64_t start, int64_t end,
                        float* scores, float* log_psi, float* r_full) {
  const int64_t n_bh = B * n_hyps;
  const int64_t SN = snum > 0 ? snum : V;
  const float LZ = (float)ORACLE_LOGZERO;
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t h = 0; h < n_bh; ++h) {
    const int64_t u = h / n_hyps;
    float* r_sum = (float*)malloc(sizeof(float) * T);
    float* rn = (float*)malloc(sizeof(float) * T);
    float* rb = (float*)malloc(sizeof(float) * T);
    for (int64_t t = 0; t < T; ++t)  /* :133 r_sum = logsumexp(r_prev, 1) */
      r_sum[t] = lse2_f(r_prev[(t * 2 + 0) * n_bh + h], r_prev[(t * 2 + 1) * n_bh + h]);
    float* psi_row = log_psi + h * V;
    if (snum > 0) for (int64_t v = 0; v < V; ++v) psi_row[v] = LZ;  /* :168-170 */
    for (int64_t k = 0; k < SN; ++k) {
      const int64_t c = snum > 0 ? scoring_ids[h * snum + k] : k;
      const int is_last = (c == last_ids[h]);
      for (int64_t t = 0; t < T; ++t) { rn[t] = LZ; rb[t] = LZ; }  /* :123-129 */
      if (output_len == 0) rn[0] = x[(u * T + 0) * V + c];          /* :130-131 */
      /* :157-162 forward variables */
      for (int64_t t = start; t < end; ++t) {
        float phi = is_last ? r_prev[((t - 1) * 2 + 1) * n_bh + h] : r_sum[t - 1];
        float n_new = lse2_f(rn[t - 1], phi) + x[(u * T + t) * V + c];
        float b_new = lse2_f(rn[t - 1], rb[t - 1]) + x[(u * T + t) * V + blank];
        rn[t] = n_new; rb[t] = b_new;
      }
      /* :165-180 log_psi = logsumexp over {log_phi_x[start:end], r[start-1,0]}.
       * log_phi_x[t] = log_phi[max(t-1,0)] + x[t]  (the cat at :165) */
      float m = rn[start - 1];
      for (int64_t t = start; t < end; ++t) {
        int64_t tp = t - 1 > 0 ? t - 1 : 0;
        float phi = is_last ? r_prev[(tp * 2 + 1) * n_bh + h] : r_sum[tp];
        float v = phi + x[(u * T + t) * V + c];
        if (v > m) m = v;
      }
      double acc = exp((double)rn[start - 1] - m);
      for (int64_t t = start; t < end; ++t) {
        int64_t tp = t - 1 > 0 ? t - 1 : 0;
        float phi = is_last ? r_prev[(tp * 2 + 1) * n_bh + h] : r_sum[tp];
        acc += exp((double)(phi + x[(u * T + t) * V + c]) - m);
      }
      psi_row[c] = (float)(m + log(acc));
      if (r_full) {
        for (int64_t t = 0; t < T; ++t) {
          r_full[((t * 2 + 0) * n_bh + h) * SN + k] = rn[t];
          r_full[((t * 2 + 1) * n_bh + h) * SN + k] = rb[t];
        }
      }
    }
    psi_row[eos] = r_sum[xlens[u] - 1];  /* :182-183 */
    psi_row[blank] = LZ;                 /* :186 */
    const float sp = s_prev ? s_prev[h] : 0.0f;
    for (int64_t v = 0; v < V; ++v) scores[h * V + v] = psi_row[v] - sp;  /* :188 */
    free(r_sum); free(rn); free(rb);
  }
  return 0;
}

/*
 * index_select_state (ctc_prefix_score.py:190-221) without materialising r for
 * every candidate: recomputes r (T,2) for new hypothesis k from its parent
 * prefix parent[k] and the appended token token[k].
 * r_new (T,2,n_new), s_new (n_new,).
 */
int oracle_prefix_select(const float* x, int64_t B, int64_t T, int64_t V,
                         int64_t n_hyps, int blank,
                         const float* r_prev, const float* log_psi,
                         const int64_t* last_ids, int64_t output_len,
                         int64_t start, int64_t end, int64_t n_bh,
                         const int64_t* parent, const int64_t* token, int64_t n_new,
                         float* r_new, float* s_new) {
  const float LZ = (float)ORACLE_LOGZERO;
  (void)B;
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t k = 0; k < n_new; ++k) {
    const int64_t h = parent[k], c = token[k], u = h / n_hyps;
    const int is_last = (c == last_ids[h]);
    float pn = LZ, pb = LZ;
    for (int64_t t = 0; t < T; ++t) {
      float n_new_ = LZ, b_new = LZ;
      if (t == 0 && output_len == 0) n_new_ = x[(u * T + 0) * V + c];
      if (t >= start && t < end) {
        float rs = lse2_f(r_prev[((t - 1) * 2 + 0) * n_bh + h], r_prev[((t - 1) * 2 + 1) * n_bh + h]);
        float phi = is_last ? r_prev[((t - 1) * 2 + 1) * n_bh + h] : rs;
        n_new_ = lse2_f(pn, phi) + x[(u * T + t) * V + c];
        b_new = lse2_f(pn, pb) + x[(u * T + t) * V + blank];
      }
      r_new[(t * 2 + 0) * n_new + k] = n_new_;
      r_new[(t * 2 + 1) * n_new + k] = b_new;
      pn = n_new_; pb = b_new;
    }
    s_new[k] = log_psi[h * V + c];
  }
  return 0;
}

/* ------------------------------------------------------------------ */
/* 3. Viterbi forced alignment, espnet/nets/pytorch_backend/ctc.py:153-216 */
/* ------------------------------------------------------------------ */
/* lpz (T,V) float log-probs; y (L,) labels; align (T,) out labels per frame. */
int oracle_ctc_viterbi(const float* lpz, int64_t T, int64_t V, const int64_t* y, int64_t L,
                       int blank, int64_t* align) {
  const int64_t S = 2 * L + 1;
  double* delta = (double*)malloc(sizeof(double) * T * S);
  int32_t* path = (int32_t*)malloc(sizeof(int32_t) * T * S);
#define EXT(s_) (((s_) & 1) ? (int)y[(s_) >> 1] : blank)
  for (int64_t i = 0; i < T * S; ++i) { delta[i] = -100000000000.0; path[i] = -1; }
  delta[0] = lpz[EXT(0)];
  if (S > 1) delta[1] = lpz[EXT(1)];
  for (int64_t t = 1; t < T; ++t) {
    for (int64_t s = 0; s < S; ++s) {
      /* candidates in the reference's order [s, s-1, (s-2)]; numpy argmax keeps
       * the first maximum.  For s < 1 the reference indexes logdelta[t-1, -1]
       * (python wrap-around to the last state), restated here as is. */
      int ncand = (EXT(s) == blank || s < 2 || EXT(s) == EXT(s - 2)) ? 2 : 3;
      int64_t best_s = s;
      double best = delta[(t - 1) * S + s];
      for (int j = 1; j < ncand; ++j) {
        int64_t ps = s - j;
        int64_t idx = ps < 0 ? ps + S : ps;
        double v = delta[(t - 1) * S + idx];
        if (v > best) { best = v; best_s = ps; }
      }
      delta[t * S + s] = best + lpz[t * V + EXT(s)];
      path[t * S + s] = (int32_t)best_s;
    }
  }
  int64_t st;
  {
    double a = delta[(T - 1) * S + S - 1];
    double b = delta[(T - 1) * S + ((S - 2) < 0 ? S - 2 + S : S - 2)];
    st = (a >= b) ? S - 1 : S - 2;
  }
  for (int64_t t = T - 1; t >= 0; --t) {
    int64_t idx = st < 0 ? st + S : st;
    align[t] = EXT(idx);
    if (t > 0) st = path[t * S + idx];
  }
#undef EXT
  free(delta); free(path);
  return 0;
}
