/*
 * llama_oracle.c -- CPU oracle for the decoder forward path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this file's library.  The product (distributed_inference_server_b200/) never does.
 *
 * PARITY UNPINNED BY THE REFERENCE.  its-me-ojas/distributed-inference-server has no forward
 * pass to restate: crates/inference/src/worker.rs:1 and crates/inference/src/lib.rs:1-3 are
 * comments, and no tensor/ML crate appears in Cargo.toml:11-45 or Cargo.lock.  There are no
 * golden vectors for this path (the only tests are JSON/validator proptests at
 * crates/core/src/models.rs:447-473 and crates/core/src/validator.rs:340-433).  What this file
 * restates instead:
 *   - the public Llama decoder architecture (RMSNorm -> QKV -> RoPE(rotate-half) -> GQA causal
 *     attention -> O -> residual -> RMSNorm -> SwiGLU MLP -> residual; final RMSNorm; lm_head),
 *     which the reference names as its target model at .kiro/specs/.../requirements.md:5 and
 *     design.md:5;
 *   - the sampling controls the reference's request model carries: temperature and top_p only
 *     (crates/core/src/models.rs:64-70; ranges from crates/core/src/validator.rs:22-25), with
 *     temperature 0 meaning greedy.
 * It is pinned instead against an independent implementation, transformers 5.5.0
 * LlamaForCausalLM (CPU, fp32): see oracle/make_golden.py and tests/golden/.
 *
 * Two arithmetic modes:
 *   act_round = 0 : activations stay fp32 everywhere (the "fp32 reference" of north_star).
 *   act_round = 1 : activations are rounded to bf16 (RNE) at exactly the points where the GPU
 *                   path stores bf16 (GEMM inputs, q, cached k/v, attention output, SiLU-gate
 *                   output); accumulation stays fp32.  Used for token-exactness checks.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

typedef struct {
  int32_t vocab, hidden, layers, n_heads, n_kv_heads, head_dim, inter, max_ctx;
  float rope_theta, rms_eps;
} orc_config;

enum {
  ORC_EMBED = 0, ORC_ATTN_NORM = 1, ORC_WQ = 2, ORC_WK = 3, ORC_WV = 4, ORC_WO = 5,
  ORC_MLP_NORM = 6, ORC_WGATE = 7, ORC_WUP = 8, ORC_WDOWN = 9, ORC_FINAL_NORM = 10,
  ORC_LM_HEAD = 11, ORC_NKINDS = 12
};

typedef struct {
  float *attn_norm, *wq, *wk, *wv, *wo, *mlp_norm, *wgate, *wup, *wdown;
} orc_layer;

typedef struct {
  float *k, *v; /* [layers][max_ctx][n_kv_heads*head_dim] */
  int len;
} orc_seq;

typedef struct orc_model {
  orc_config c;
  int act_round;
  float *embed, *final_norm, *lm_head;
  orc_layer *L;
  float *rope_cos, *rope_sin; /* [max_ctx][head_dim/2] */
  orc_seq *seqs;
  int n_seqs, cap_seqs;
} orc_model;

/* ------------------------------------------------------------------ bf16 + RNG */

static inline float bf16_round(float x) {
  uint32_t u;
  memcpy(&u, &x, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) { u |= 0x00400000u; u &= 0xffff0000u; } /* NaN */
  else { u += 0x7fffu + ((u >> 16) & 1u); u &= 0xffff0000u; }
  memcpy(&x, &u, 4);
  return x;
}

ORC_API void orc_bf16_round_inplace(float *x, size_t n) {
  for (size_t i = 0; i < n; i++) x[i] = bf16_round(x[i]);
}

/* splitmix64 finaliser over (seed, key, idx): the synthetic-weight and sampling RNG. */
static inline uint64_t mix64(uint64_t seed, uint64_t key, uint64_t idx) {
  uint64_t z = seed * 0x9E3779B97F4A7C15ull + key * 0xD1B54A32D192ED03ull + idx + 0x632BE59BD9B4E019ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

static inline float u01(uint64_t h) { return (float)(h >> 40) * (1.0f / 16777216.0f); }

ORC_API float orc_uniform(uint64_t seed, uint64_t key, uint64_t idx) { return u01(mix64(seed, key, idx)); }

/* value = offset + (2u-1)*amp, optionally rounded to bf16.  (2u-1) is exact in fp32. */
ORC_API void orc_synth_tensor(float *dst, size_t n, uint64_t seed, int kind, int layer, float amp,
                              float offset, int round_bf16) {
  uint64_t key = (uint64_t)layer * 16u + (uint64_t)kind;
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < n; i++) {
    float u = u01(mix64(seed, key, i));
    float s = u * 2.0f - 1.0f;
    float v = s * amp;
    v = v + offset;
    dst[i] = round_bf16 ? bf16_round(v) : v;
  }
}

/* amplitude / offset convention for synthetic models (shared with the engine's own generator,
 * which is written independently in csrc/synth.cuh). std = amp/sqrt(3). */
ORC_API void orc_synth_params(const orc_config *c, int kind, float *amp, float *offset) {
  const float a = 0.034641016f; /* 0.02*sqrt(3) */
  *offset = 0.0f;
  switch (kind) {
    case ORC_ATTN_NORM: case ORC_MLP_NORM: case ORC_FINAL_NORM: *amp = 0.1f; *offset = 1.0f; break;
    case ORC_WO: case ORC_WDOWN: *amp = a / sqrtf(2.0f * (float)c->layers); break;
    default: *amp = a; break;
  }
}

/* ------------------------------------------------------------------ primitives */

ORC_API void orc_rmsnorm(const float *x, const float *g, float *y, int T, int H, float eps, int round_out) {
#pragma omp parallel for schedule(static)
  for (int t = 0; t < T; t++) {
    const float *xr = x + (size_t)t * H;
    float *yr = y + (size_t)t * H;
    double ss = 0.0;
    for (int i = 0; i < H; i++) ss += (double)xr[i] * (double)xr[i];
    float inv = 1.0f / sqrtf((float)(ss / (double)H) + eps);
    for (int i = 0; i < H; i++) {
      float v = xr[i] * inv * g[i];
      yr[i] = round_out ? bf16_round(v) : v;
    }
  }
}

/* Y[T,N] = X[T,K] * W[N,K]^T, fp32 accumulate. */
ORC_API void orc_linear(const float *X, const float *W, float *Y, int T, int N, int K) {
#pragma omp parallel for schedule(static)
  for (int n = 0; n < N; n++) {
    const float *w = W + (size_t)n * K;
    int t = 0;
    for (; t + 4 <= T; t += 4) {
      const float *x0 = X + (size_t)t * K, *x1 = x0 + K, *x2 = x1 + K, *x3 = x2 + K;
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma omp simd reduction(+ : a0, a1, a2, a3)
      for (int k = 0; k < K; k++) {
        float wk = w[k];
        a0 += wk * x0[k]; a1 += wk * x1[k]; a2 += wk * x2[k]; a3 += wk * x3[k];
      }
      Y[(size_t)(t + 0) * N + n] = a0; Y[(size_t)(t + 1) * N + n] = a1;
      Y[(size_t)(t + 2) * N + n] = a2; Y[(size_t)(t + 3) * N + n] = a3;
    }
    for (; t < T; t++) {
      const float *x0 = X + (size_t)t * K;
      float a0 = 0.f;
#pragma omp simd reduction(+ : a0)
      for (int k = 0; k < K; k++) a0 += w[k] * x0[k];
      Y[(size_t)t * N + n] = a0;
    }
  }
}

/* cos/sin tables computed in double, stored fp32: [max_ctx][d/2], angle = pos * theta^(-2i/d). */
ORC_API void orc_rope_table(float *cs, float *sn, int max_ctx, int d, float theta) {
  int half = d / 2;
  for (int p = 0; p < max_ctx; p++)
    for (int i = 0; i < half; i++) {
      double inv_freq = pow((double)theta, -2.0 * (double)i / (double)d);
      double a = (double)p * inv_freq;
      cs[(size_t)p * half + i] = (float)cos(a);
      sn[(size_t)p * half + i] = (float)sin(a);
    }
}

/* rotate-half RoPE on one head vector (HF convention): pairs (i, i+d/2). */
static inline void rope_head(float *v, const float *cs, const float *sn, int d) {
  int half = d / 2;
  for (int i = 0; i < half; i++) {
    float a = v[i], b = v[i + half];
    v[i] = a * cs[i] - b * sn[i];
    v[i + half] = b * cs[i] + a * sn[i];
  }
}

/* x: [T][n_heads_total*d] in place; positions[T]. */
ORC_API void orc_rope(float *x, const int *positions, int T, int n_heads, int d, const float *cs,
                      const float *sn) {
  int half = d / 2;
#pragma omp parallel for schedule(static)
  for (int t = 0; t < T; t++)
    for (int h = 0; h < n_heads; h++)
      rope_head(x + ((size_t)t * n_heads + h) * d, cs + (size_t)positions[t] * half,
                sn + (size_t)positions[t] * half, d);
}

/* One query token against a contiguous K/V history.
 * q: [n_heads][d]; K,V: [ctx][n_kv][d] (row stride kv_stride floats); out: [n_heads][d]. */
ORC_API void orc_attention_one(const float *q, const float *K, const float *V, float *out, int ctx,
                               int n_heads, int n_kv, int d, size_t kv_stride) {
  int group = n_heads / n_kv;
  float scale = 1.0f / sqrtf((float)d);
  float *s = (float *)malloc((size_t)ctx * sizeof(float));
  for (int h = 0; h < n_heads; h++) {
    int kvh = h / group;
    const float *qh = q + (size_t)h * d;
    float m = -INFINITY;
    for (int j = 0; j < ctx; j++) {
      const float *kj = K + (size_t)j * kv_stride + (size_t)kvh * d;
      float acc = 0.f;
      for (int i = 0; i < d; i++) acc += qh[i] * kj[i];
      acc *= scale;
      s[j] = acc;
      if (acc > m) m = acc;
    }
    double denom = 0.0;
    for (int j = 0; j < ctx; j++) { s[j] = expf(s[j] - m); denom += (double)s[j]; }
    float inv = (float)(1.0 / denom);
    float *oh = out + (size_t)h * d;
    for (int i = 0; i < d; i++) oh[i] = 0.f;
    for (int j = 0; j < ctx; j++) {
      const float *vj = V + (size_t)j * kv_stride + (size_t)kvh * d;
      float p = s[j];
      for (int i = 0; i < d; i++) oh[i] += p * vj[i];
    }
    for (int i = 0; i < d; i++) oh[i] *= inv;
  }
  free(s);
}

ORC_API void orc_silu_mul(const float *gate, const float *up, float *y, size_t n, int round_out) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < n; i++) {
    float g = gate[i];
    float v = (g / (1.0f + expf(-g))) * up[i];
    y[i] = round_out ? bf16_round(v) : v;
  }
}

/* ------------------------------------------------------------------ sampler
 * temperature == 0 -> argmax, lowest index wins ties.
 * else p = softmax(logits / temperature); nucleus = { i : p_i >= tau } where tau is the
 * probability of the last token of the shortest descending-sorted prefix whose mass >= top_p
 * (ties at tau are all kept); draw u = uniform(seed, key=0xC0FFEE, idx=step); walk the kept
 * tokens in ascending index order and return the first whose running mass exceeds u * kept_mass.
 * (models.rs:64-70 has temperature and top_p only -- no top_k.) */
typedef struct { float p; int i; } orc_pi;
static int cmp_pi(const void *a, const void *b) {
  const orc_pi *x = (const orc_pi *)a, *y = (const orc_pi *)b;
  if (x->p > y->p) return -1;
  if (x->p < y->p) return 1;
  return x->i - y->i;
}

ORC_API int orc_argmax(const float *logits, int V, float *margin) {
  int best = 0;
  float b = logits[0], second = -INFINITY;
  for (int i = 1; i < V; i++) {
    float x = logits[i];
    if (x > b) { second = b; b = x; best = i; }
    else if (x > second) second = x;
  }
  if (margin) *margin = b - second;
  return best;
}

ORC_API int orc_sample(const float *logits, int V, float temperature, float top_p, uint64_t seed,
                       uint64_t step, float *out_tau, float *out_kept_mass) {
  if (temperature <= 0.f) return orc_argmax(logits, V, NULL);
  float inv_t = 1.0f / temperature;
  float m = -INFINITY;
  for (int i = 0; i < V; i++) { float z = logits[i] * inv_t; if (z > m) m = z; }
  float *p = (float *)malloc((size_t)V * sizeof(float));
  double sum = 0.0;
  for (int i = 0; i < V; i++) { p[i] = expf(logits[i] * inv_t - m); sum += (double)p[i]; }
  float inv = (float)(1.0 / sum);
  for (int i = 0; i < V; i++) p[i] *= inv;
  float tau = 0.f;
  if (top_p < 1.0f) {
    orc_pi *s = (orc_pi *)malloc((size_t)V * sizeof(orc_pi));
    for (int i = 0; i < V; i++) { s[i].p = p[i]; s[i].i = i; }
    qsort(s, (size_t)V, sizeof(orc_pi), cmp_pi);
    double cum = 0.0;
    tau = s[V - 1].p;
    for (int i = 0; i < V; i++) { cum += (double)s[i].p; if (cum >= (double)top_p) { tau = s[i].p; break; } }
    free(s);
  }
  double kept = 0.0;
  for (int i = 0; i < V; i++) if (p[i] >= tau) kept += (double)p[i];
  float u = u01(mix64(seed, 0xC0FFEEull, step));
  double target = (double)u * kept, run = 0.0;
  int tok = -1, last_kept = 0;
  for (int i = 0; i < V; i++) {
    if (p[i] >= tau) { run += (double)p[i]; last_kept = i; if (run > target) { tok = i; break; } }
  }
  if (tok < 0) tok = last_kept;
  if (out_tau) *out_tau = tau;
  if (out_kept_mass) *out_kept_mass = (float)kept;
  free(p);
  return tok;
}

/* ------------------------------------------------------------------ model */

static size_t kind_elems(const orc_config *c, int kind) {
  size_t H = c->hidden, qd = (size_t)c->n_heads * c->head_dim, kd = (size_t)c->n_kv_heads * c->head_dim;
  switch (kind) {
    case ORC_EMBED: case ORC_LM_HEAD: return (size_t)c->vocab * H;
    case ORC_ATTN_NORM: case ORC_MLP_NORM: case ORC_FINAL_NORM: return H;
    case ORC_WQ: return qd * H;
    case ORC_WK: case ORC_WV: return kd * H;
    case ORC_WO: return H * qd;
    case ORC_WGATE: case ORC_WUP: return (size_t)c->inter * H;
    case ORC_WDOWN: return H * (size_t)c->inter;
  }
  return 0;
}

ORC_API size_t orc_tensor_elems(const orc_config *c, int kind) { return kind_elems(c, kind); }

static float *falloc(size_t n) {
  float *p = (float *)malloc(n * sizeof(float));
  if (!p) { fprintf(stderr, "oracle: out of memory (%zu floats)\n", n); abort(); }
  return p;
}

ORC_API orc_model *orc_create(const orc_config *cfg, int act_round) {
  orc_model *m = (orc_model *)calloc(1, sizeof(orc_model));
  m->c = *cfg;
  m->act_round = act_round;
  m->embed = falloc(kind_elems(cfg, ORC_EMBED));
  m->final_norm = falloc(kind_elems(cfg, ORC_FINAL_NORM));
  m->lm_head = falloc(kind_elems(cfg, ORC_LM_HEAD));
  m->L = (orc_layer *)calloc((size_t)cfg->layers, sizeof(orc_layer));
  for (int l = 0; l < cfg->layers; l++) {
    orc_layer *y = &m->L[l];
    y->attn_norm = falloc(kind_elems(cfg, ORC_ATTN_NORM));
    y->wq = falloc(kind_elems(cfg, ORC_WQ));
    y->wk = falloc(kind_elems(cfg, ORC_WK));
    y->wv = falloc(kind_elems(cfg, ORC_WV));
    y->wo = falloc(kind_elems(cfg, ORC_WO));
    y->mlp_norm = falloc(kind_elems(cfg, ORC_MLP_NORM));
    y->wgate = falloc(kind_elems(cfg, ORC_WGATE));
    y->wup = falloc(kind_elems(cfg, ORC_WUP));
    y->wdown = falloc(kind_elems(cfg, ORC_WDOWN));
  }
  size_t half = (size_t)cfg->head_dim / 2;
  m->rope_cos = falloc((size_t)cfg->max_ctx * half);
  m->rope_sin = falloc((size_t)cfg->max_ctx * half);
  orc_rope_table(m->rope_cos, m->rope_sin, cfg->max_ctx, cfg->head_dim, cfg->rope_theta);
  return m;
}

ORC_API float *orc_tensor(orc_model *m, int kind, int layer) {
  switch (kind) {
    case ORC_EMBED: return m->embed;
    case ORC_FINAL_NORM: return m->final_norm;
    case ORC_LM_HEAD: return m->lm_head;
  }
  if (layer < 0 || layer >= m->c.layers) return NULL;
  orc_layer *y = &m->L[layer];
  switch (kind) {
    case ORC_ATTN_NORM: return y->attn_norm;
    case ORC_WQ: return y->wq;
    case ORC_WK: return y->wk;
    case ORC_WV: return y->wv;
    case ORC_WO: return y->wo;
    case ORC_MLP_NORM: return y->mlp_norm;
    case ORC_WGATE: return y->wgate;
    case ORC_WUP: return y->wup;
    case ORC_WDOWN: return y->wdown;
  }
  return NULL;
}

ORC_API void orc_fill_synthetic(orc_model *m, uint64_t seed, int round_bf16) {
  for (int kind = 0; kind < ORC_NKINDS; kind++) {
    int per_layer = !(kind == ORC_EMBED || kind == ORC_FINAL_NORM || kind == ORC_LM_HEAD);
    int nl = per_layer ? m->c.layers : 1;
    for (int l = 0; l < nl; l++) {
      float amp, off;
      orc_synth_params(&m->c, kind, &amp, &off);
      orc_synth_tensor(orc_tensor(m, kind, l), kind_elems(&m->c, kind), seed, kind, per_layer ? l : 0,
                       amp, off, round_bf16);
    }
  }
}

ORC_API int orc_seq_new(orc_model *m) {
  for (int i = 0; i < m->n_seqs; i++)
    if (m->seqs[i].k == NULL) { /* reuse freed slot */
      size_t n = (size_t)m->c.layers * m->c.max_ctx * m->c.n_kv_heads * m->c.head_dim;
      m->seqs[i].k = falloc(n); m->seqs[i].v = falloc(n); m->seqs[i].len = 0;
      return i;
    }
  if (m->n_seqs == m->cap_seqs) {
    m->cap_seqs = m->cap_seqs ? m->cap_seqs * 2 : 8;
    m->seqs = (orc_seq *)realloc(m->seqs, (size_t)m->cap_seqs * sizeof(orc_seq));
  }
  size_t n = (size_t)m->c.layers * m->c.max_ctx * m->c.n_kv_heads * m->c.head_dim;
  orc_seq *s = &m->seqs[m->n_seqs];
  s->k = falloc(n); s->v = falloc(n); s->len = 0;
  return m->n_seqs++;
}

ORC_API void orc_seq_free(orc_model *m, int id) {
  if (id < 0 || id >= m->n_seqs) return;
  free(m->seqs[id].k); free(m->seqs[id].v);
  m->seqs[id].k = m->seqs[id].v = NULL; m->seqs[id].len = 0;
}

ORC_API int orc_seq_len(orc_model *m, int id) { return m->seqs[id].len; }

/* Fill a sequence's K/V history with synthetic values without running a prefill
 * (used only by the bounded CPU timing sample in bench.py). */
ORC_API void orc_seq_fill_synthetic(orc_model *m, int id, int len, uint64_t seed) {
  orc_seq *s = &m->seqs[id];
  size_t kvd = (size_t)m->c.n_kv_heads * m->c.head_dim;
  for (int l = 0; l < m->c.layers; l++) {
    size_t off = (size_t)l * m->c.max_ctx * kvd;
    orc_synth_tensor(s->k + off, (size_t)len * kvd, seed + (uint64_t)id, 14, l, 1.0f, 0.f, 1);
    orc_synth_tensor(s->v + off, (size_t)len * kvd, seed + (uint64_t)id, 15, l, 1.0f, 0.f, 1);
  }
  s->len = len;
}

ORC_API void orc_destroy(orc_model *m) {
  if (!m) return;
  for (int i = 0; i < m->n_seqs; i++) { free(m->seqs[i].k); free(m->seqs[i].v); }
  free(m->seqs);
  for (int l = 0; l < m->c.layers; l++) {
    orc_layer *y = &m->L[l];
    free(y->attn_norm); free(y->wq); free(y->wk); free(y->wv); free(y->wo);
    free(y->mlp_norm); free(y->wgate); free(y->wup); free(y->wdown);
  }
  free(m->L); free(m->embed); free(m->final_norm); free(m->lm_head);
  free(m->rope_cos); free(m->rope_sin);
  free(m);
}

/* Forward T tokens.  Token t belongs to sequence seq_ids[t] at position positions[t]; a
 * sequence's tokens must appear in increasing position order and extend its history
 * contiguously.  logit_rows[t] >= 0 asks for that token's logits in logits[logit_rows[t]*V].
 * hidden_out (optional): final-norm input residual [T][H] for debugging.
 * Returns 0, or -1 on a bad argument. */
ORC_API int orc_forward(orc_model *m, int T, const int *tokens, const int *seq_ids, const int *positions,
                        const int *logit_rows, float *logits, float *hidden_out) {
  const orc_config *c = &m->c;
  const int H = c->hidden, d = c->head_dim, nh = c->n_heads, nkv = c->n_kv_heads, I = c->inter;
  const size_t qd = (size_t)nh * d, kvd = (size_t)nkv * d;
  const int R = m->act_round;
  for (int t = 0; t < T; t++) {
    if (tokens[t] < 0 || tokens[t] >= c->vocab) return -1;
    if (seq_ids[t] < 0 || seq_ids[t] >= m->n_seqs || !m->seqs[seq_ids[t]].k) return -1;
    if (positions[t] < 0 || positions[t] >= c->max_ctx) return -1;
  }
  float *x = falloc((size_t)T * H), *n = falloc((size_t)T * H);
  float *q = falloc((size_t)T * qd), *k = falloc((size_t)T * kvd), *v = falloc((size_t)T * kvd);
  float *att = falloc((size_t)T * qd), *o = falloc((size_t)T * H);
  float *g = falloc((size_t)T * I), *u = falloc((size_t)T * I), *a = falloc((size_t)T * I);

  for (int t = 0; t < T; t++) memcpy(x + (size_t)t * H, m->embed + (size_t)tokens[t] * H, (size_t)H * sizeof(float));

  for (int l = 0; l < c->layers; l++) {
    orc_layer *y = &m->L[l];
    orc_rmsnorm(x, y->attn_norm, n, T, H, c->rms_eps, R);
    orc_linear(n, y->wq, q, T, (int)qd, H);
    orc_linear(n, y->wk, k, T, (int)kvd, H);
    orc_linear(n, y->wv, v, T, (int)kvd, H);
    orc_rope(q, positions, T, nh, d, m->rope_cos, m->rope_sin);
    orc_rope(k, positions, T, nkv, d, m->rope_cos, m->rope_sin);
    if (R) { orc_bf16_round_inplace(q, (size_t)T * qd); orc_bf16_round_inplace(k, (size_t)T * kvd);
             orc_bf16_round_inplace(v, (size_t)T * kvd); }
    /* write all K/V of this batch first, then attend causally by position */
    for (int t = 0; t < T; t++) {
      orc_seq *s = &m->seqs[seq_ids[t]];
      size_t off = ((size_t)l * c->max_ctx + positions[t]) * kvd;
      memcpy(s->k + off, k + (size_t)t * kvd, kvd * sizeof(float));
      memcpy(s->v + off, v + (size_t)t * kvd, kvd * sizeof(float));
    }
#pragma omp parallel for schedule(dynamic, 1)
    for (int t = 0; t < T; t++) {
      orc_seq *s = &m->seqs[seq_ids[t]];
      size_t off = (size_t)l * c->max_ctx * kvd;
      orc_attention_one(q + (size_t)t * qd, s->k + off, s->v + off, att + (size_t)t * qd, positions[t] + 1,
                        nh, nkv, d, kvd);
    }
    if (R) orc_bf16_round_inplace(att, (size_t)T * qd);
    orc_linear(att, y->wo, o, T, H, (int)qd);
    for (size_t i = 0; i < (size_t)T * H; i++) x[i] += o[i];
    orc_rmsnorm(x, y->mlp_norm, n, T, H, c->rms_eps, R);
    orc_linear(n, y->wgate, g, T, I, H);
    orc_linear(n, y->wup, u, T, I, H);
    orc_silu_mul(g, u, a, (size_t)T * I, R);
    orc_linear(a, y->wdown, o, T, H, I);
    for (size_t i = 0; i < (size_t)T * H; i++) x[i] += o[i];
  }
  for (int t = 0; t < T; t++) {
    orc_seq *s = &m->seqs[seq_ids[t]];
    if (positions[t] + 1 > s->len) s->len = positions[t] + 1;
  }
  if (hidden_out) memcpy(hidden_out, x, (size_t)T * H * sizeof(float));
  if (logit_rows && logits) {
    int nrows = 0;
    for (int t = 0; t < T; t++) if (logit_rows[t] >= 0) nrows++;
    if (nrows) {
      float *xs = falloc((size_t)nrows * H), *ns = falloc((size_t)nrows * H), *lg = falloc((size_t)nrows * c->vocab);
      int *rows = (int *)malloc((size_t)nrows * sizeof(int));
      int r = 0;
      for (int t = 0; t < T; t++) if (logit_rows[t] >= 0) { memcpy(xs + (size_t)r * H, x + (size_t)t * H, (size_t)H * sizeof(float)); rows[r++] = logit_rows[t]; }
      orc_rmsnorm(xs, m->final_norm, ns, nrows, H, c->rms_eps, R);
      orc_linear(ns, m->lm_head, lg, nrows, c->vocab, H);
      for (r = 0; r < nrows; r++) memcpy(logits + (size_t)rows[r] * c->vocab, lg + (size_t)r * c->vocab, (size_t)c->vocab * sizeof(float));
      free(xs); free(ns); free(lg); free(rows);
    }
  }
  free(x); free(n); free(q); free(k); free(v); free(att); free(o); free(g); free(u); free(a);
  return 0;
}

/* Greedy/sampled generation for a padded batch, the shape InferenceBatch describes
 * (design.md:241-248): prompts[b] has prompt_lens[b] tokens.  out_tokens: [B][max_new];
 * out_margins (optional): top1-top2 logit gap per generated token; out_logits (optional):
 * [B][n_logit_steps][V] logits of the first n_logit_steps generated positions. */
ORC_API int orc_generate(orc_model *m, int B, const int *prompts, const int *prompt_lens, int prompt_stride,
                         int max_new, float temperature, float top_p, const uint64_t *seeds, int eos_id,
                         int *out_tokens, int *out_lens, float *out_margins, float *out_logits,
                         int n_logit_steps) {
  const int V = m->c.vocab;
  int *sid = (int *)malloc((size_t)B * sizeof(int));
  int *cur = (int *)malloc((size_t)B * sizeof(int));
  int *done = (int *)calloc((size_t)B, sizeof(int));
  float *lg = falloc((size_t)B * V);
  int rc = 0;
  for (int b = 0; b < B; b++) { sid[b] = orc_seq_new(m); out_lens[b] = 0; }
  /* prefill, one sequence at a time */
  for (int b = 0; b < B && rc == 0; b++) {
    int n = prompt_lens[b];
    int *pos = (int *)malloc((size_t)n * sizeof(int)), *ids = (int *)malloc((size_t)n * sizeof(int));
    int *rows = (int *)malloc((size_t)n * sizeof(int));
    for (int t = 0; t < n; t++) { pos[t] = t; ids[t] = sid[b]; rows[t] = (t == n - 1) ? b : -1; }
    rc = orc_forward(m, n, prompts + (size_t)b * prompt_stride, ids, pos, rows, lg, NULL);
    free(pos); free(ids); free(rows);
  }
  for (int step = 0; step < max_new && rc == 0; step++) {
    int nact = 0;
    for (int b = 0; b < B; b++) {
      if (done[b]) continue;
      float margin = 0.f;
      int tok;
      if (temperature <= 0.f) tok = orc_argmax(lg + (size_t)b * V, V, &margin);
      else { orc_argmax(lg + (size_t)b * V, V, &margin); tok = orc_sample(lg + (size_t)b * V, V, temperature, top_p, seeds ? seeds[b] : 0, (uint64_t)step, NULL, NULL); }
      if (out_logits && step < n_logit_steps)
        memcpy(out_logits + ((size_t)b * n_logit_steps + step) * V, lg + (size_t)b * V, (size_t)V * sizeof(float));
      out_tokens[(size_t)b * max_new + step] = tok;
      if (out_margins) out_margins[(size_t)b * max_new + step] = margin;
      out_lens[b] = step + 1;
      cur[b] = tok;
      if (tok == eos_id) done[b] = 1; else nact++;
    }
    if (step == max_new - 1 || nact == 0) break;
    int *toks = (int *)malloc((size_t)nact * sizeof(int)), *ids = (int *)malloc((size_t)nact * sizeof(int));
    int *pos = (int *)malloc((size_t)nact * sizeof(int)), *rows = (int *)malloc((size_t)nact * sizeof(int));
    int j = 0;
    for (int b = 0; b < B; b++) if (!done[b]) { toks[j] = cur[b]; ids[j] = sid[b]; pos[j] = prompt_lens[b] + step; rows[j] = b; j++; }
    rc = orc_forward(m, nact, toks, ids, pos, rows, lg, NULL);
    free(toks); free(ids); free(pos); free(rows);
  }
  for (int b = 0; b < B; b++) orc_seq_free(m, sid[b]);
  free(sid); free(cur); free(done); free(lg);
  return rc;
}

/* bench.py sets the thread count explicitly: a launcher such as torch.distributed.run exports OMP_NUM_THREADS=1,
 * which would silently turn the CPU baseline into a one-core figure. */
ORC_API void orc_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

ORC_API int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
