// k_elementwise.cu -- the small HBM-bound kernels of the decoder step:
//   add_rmsnorm      embedding gather | residual add of split-K partials, then RMSNorm -> bf16
//   rope_kv_write    split-K reduce of the fused QKV projection, RoPE, q -> bf16, k/v -> KV pages
//   reduce_partials  split-K reduce alone (feeds the tensor-parallel all-reduce)
//   sample           argmax / temperature + nucleus sampler (warp-shuffle reductions)
//   synth_bf16       counter-hash synthetic weights
// None of these has a reference counterpart; they implement §3.3's decoder step.
#include "kernels.h"
#include "ptx.cuh"

namespace dis {

bool &pdl_enabled() {
  static thread_local bool on = false;
  return on;
}

// ------------------------------------------------------------------------------ block reductions
// kThreads = 0: use blockDim.x (a multiple of 32)
template <int kThreads>
__device__ __forceinline__ float block_sum(float v, float *sm) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int nw = kThreads ? kThreads / 32 : (int)(blockDim.x >> 5);
  __syncthreads();
  if (l == 0) sm[w] = v;
  __syncthreads();
  float r = (l < nw) ? sm[l] : 0.f;
  r = warp_sum(r);
  return r;  // every thread has the total
}
template <int kThreads>
__device__ __forceinline__ float block_max(float v, float *sm) {
  v = warp_max(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int nw = kThreads ? kThreads / 32 : (int)(blockDim.x >> 5);
  __syncthreads();
  if (l == 0) sm[w] = v;
  __syncthreads();
  float r = (l < nw) ? sm[l] : -INFINITY;
  r = warp_max(r);
  return r;
}

// ------------------------------------------------------------------------------ add_rmsnorm
// One CTA per row with H/4 threads (<= 1024): every thread owns one float4 column group per pass, so
// the S split-K slices of a row are fetched with S independent 16-byte loads per thread in flight.
// Measured and rejected (round 2): fetching the gain (a weight, cold in L2 once per step) BEFORE griddepcontrol.wait to
// take its latency off the chain wait -> row loads -> reduction -> gain -> store: graph-only step 3.505 / 3.500 ms
// against 3.492 / 3.491 (alternating runs on one box, profiles/r2_norm_early_gain_*.json) -- not faster.
__global__ void __launch_bounds__(1024)
add_rmsnorm_kernel(float *__restrict__ x, const uint32_t *__restrict__ tokens,
                   const uint16_t *__restrict__ embed, const float *__restrict__ partial, int n_partial,
                   long long partial_stride, const float *__restrict__ gain, uint16_t *__restrict__ xn,
                   const int *__restrict__ row_map, int H, float eps) {
  __shared__ float red[32];
  extern __shared__ float row[];  // H floats (only used when a thread owns more than one group)
  const int t = blockIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  const int src = row_map ? row_map[t] : t;
  const int step = blockDim.x * 4;
  const bool single = H <= step;
  float ss = 0.f;
  float4 keep = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = threadIdx.x * 4; i < H; i += step) {
    float4 v;
    if (tokens) {
      const uint2 e = *reinterpret_cast<const uint2 *>(embed + (size_t)tokens[src] * H + i);
      const float2 a = bf16x2_to_float2(e.x), b = bf16x2_to_float2(e.y);
      v = make_float4(a.x, a.y, b.x, b.y);
    } else {
      v = *reinterpret_cast<const float4 *>(x + (size_t)src * H + i);
      const float *pp = partial + (size_t)src * H + i;
#pragma unroll 4
      for (int s = 0; s < n_partial; s++) {
        const float4 p = *reinterpret_cast<const float4 *>(pp + (size_t)s * partial_stride);
        v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
      }
    }
    if (!row_map) *reinterpret_cast<float4 *>(x + (size_t)t * H + i) = v;
    if (single) keep = v; else *reinterpret_cast<float4 *>(row + i) = v;
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  ss = block_sum<0>(ss, red);
  const float inv = rsqrtf(ss / (float)H + eps);
  for (int i = threadIdx.x * 4; i < H; i += step) {
    const float4 v = single ? keep : *reinterpret_cast<const float4 *>(row + i);
    const float4 g = *reinterpret_cast<const float4 *>(gain + i);
    uint2 o;
    o.x = float2_to_bf16x2(v.x * inv * g.x, v.y * inv * g.y);
    o.y = float2_to_bf16x2(v.z * inv * g.z, v.w * inv * g.w);
    *reinterpret_cast<uint2 *>(xn + (size_t)t * H + i) = o;
  }
}

// rows from which the row kernels switch to 256-thread CTAs (measured: single-request TTFT 9.00 -> 8.73 ms at 512
// tokens, 5.64 -> 5.92 ms at 256 tokens; 135 -> 76 us per norm at 8192)
static int wide_rows_threshold() { return 512; }

cudaError_t launch_add_rmsnorm(float *x, const uint32_t *tokens, const uint16_t *embed, const float *partial,
                               int n_partial, long long partial_stride, const float *gain, uint16_t *xn,
                               const int *row_map, int T, int H, float eps, cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  if (H % 4 != 0 || H * 4 > 200 * 1024) { set_last_error("add_rmsnorm: H must be a multiple of 4 and <= 51200"); return cudaErrorInvalidValue; }
  int threads = (H / 4 + 31) / 32 * 32;
  threads = threads > 1024 ? 1024 : threads;
  // prefill-sized T is a streaming job: 256-thread CTAs put 8 rows per SM in different phases (load / reduce /
  // store), where two 1024-thread rows leave the memory pipe idle during every reduction (T = 8192: 135 -> see DESIGN)
  if (T >= wide_rows_threshold() && threads > 256) threads = 256;
  const size_t smem = H <= threads * 4 ? 0 : (size_t)H * sizeof(float);
  if (smem > 48 * 1024) DIS_CUDA_TRY(cudaFuncSetAttribute(add_rmsnorm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  return launch_k(add_rmsnorm_kernel, dim3(T), dim3(threads), smem, st, x, tokens, embed, partial, n_partial, partial_stride, gain,
                  xn, row_map, H, eps);
}

// ------------------------------------------------------------------------------ reduce_partials
__global__ void reduce_partials_kernel(const float *__restrict__ partial, int n_partial, long long stride,
                                       float *__restrict__ out, long long n4) {
  pdl_launch_dependents();
  pdl_wait();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 v = reinterpret_cast<const float4 *>(partial)[i];
    for (int s = 1; s < n_partial; s++) {
      const float4 p = reinterpret_cast<const float4 *>(partial + (size_t)s * stride)[i];
      v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
    }
    reinterpret_cast<float4 *>(out)[i] = v;
  }
}

cudaError_t launch_reduce_partials(const float *partial, int n_partial, long long stride, float *out, long long n,
                                   cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  if (n % 4 != 0) { set_last_error("reduce_partials: n must be a multiple of 4"); return cudaErrorInvalidValue; }
  const long long n4 = n / 4;
  const int blocks = (int)min((long long)kNumSMs * 8, (n4 + 255) / 256);
  return launch_k(reduce_partials_kernel, dim3(blocks), dim3(256), 0, st, partial, n_partial, stride, out, n4);
}

// ------------------------------------------------------------------------------ silu_mul_reduce
// act[t][i] = bf16(silu(sum_s p[s][t][i]) * sum_s p[s][t][I + i]) for a split-K gate-up projection whose
// shard is too small to fill the GPU without split-K (tensor parallel ranks).
__global__ void silu_mul_reduce_kernel(const float *__restrict__ partial, int n_partial, long long stride,
                                       uint16_t *__restrict__ act, int T, int I) {
  pdl_launch_dependents();
  pdl_wait();
  const long long n4 = (long long)T * I / 4;
  for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < n4; j += (long long)gridDim.x * blockDim.x) {
    const long long t = (j * 4) / I, i = (j * 4) - t * I;
    const float *g = partial + (size_t)t * 2 * I + i;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
#pragma unroll 4
    for (int s = 0; s < n_partial; s++) {
      const float4 pg = *reinterpret_cast<const float4 *>(g + (size_t)s * stride);
      const float4 pu = *reinterpret_cast<const float4 *>(g + (size_t)s * stride + I);
      a.x += pg.x; a.y += pg.y; a.z += pg.z; a.w += pg.w;
      b.x += pu.x; b.y += pu.y; b.z += pu.z; b.w += pu.w;
    }
    uint2 o;
    o.x = float2_to_bf16x2((a.x / (1.0f + expf(-a.x))) * b.x, (a.y / (1.0f + expf(-a.y))) * b.y);
    o.y = float2_to_bf16x2((a.z / (1.0f + expf(-a.z))) * b.z, (a.w / (1.0f + expf(-a.w))) * b.w);
    *reinterpret_cast<uint2 *>(act + (size_t)t * I + i) = o;
  }
}

cudaError_t launch_silu_mul_reduce(const float *partial, int n_partial, long long stride, uint16_t *act, int T, int I,
                                   cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  if (I % 4 != 0) { set_last_error("silu_mul_reduce: I must be a multiple of 4"); return cudaErrorInvalidValue; }
  const long long n4 = (long long)T * I / 4;
  const int blocks = (int)min((long long)kNumSMs * 4, (n4 + 255) / 256);
  return launch_k(silu_mul_reduce_kernel, dim3(blocks), dim3(256), 0, st, partial, n_partial, stride, act, T, I);
}

// ------------------------------------------------------------------------------ rope_kv_write
// One CTA per token.  A work unit is four adjacent rotate-half pairs (i..i+3, i+d/2..i+d/2+3) of a q
// or k head, or four adjacent v elements: two (one) float4 loads per split-K slice, all independent.
__device__ __forceinline__ float4 sum_slices(const float *p, int n_partial, long long stride) {
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
  for (int s = 0; s < n_partial; s++) {
    const float4 v = *reinterpret_cast<const float4 *>(p + (size_t)s * stride);
    a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
  }
  return a;
}

__global__ void __launch_bounds__(1024)
rope_kv_write_kernel(const float *__restrict__ qkv, int n_partial, long long partial_stride,
                     uint16_t *__restrict__ q, uint16_t *__restrict__ kv_pool,
                     const int *__restrict__ page_table, int max_pages, const int *__restrict__ pt_row,
                     const int *__restrict__ pos, const float *__restrict__ rope_cos,
                     const float *__restrict__ rope_sin, int n_heads, int n_kv, int d, int page_size) {
  const int t = blockIdx.x;
  const int half = d >> 1, hq = half >> 2;  // hq: units per head
  pdl_launch_dependents();
  pdl_wait();
  const int p = pos[t];
  const int nq = n_heads * d, nk = n_kv * d;
  const int ntot = nq + 2 * nk;
  const float *base = qkv + (size_t)t * ntot;
  const int page = page_table[(size_t)pt_row[t] * max_pages + p / page_size];
  const int slot = p % page_size;
  // [page][2][n_kv][page_size][d]
  uint16_t *kpage = kv_pool + (size_t)page * 2 * n_kv * page_size * d;
  uint16_t *vpage = kpage + (size_t)n_kv * page_size * d;
  const float *cs = rope_cos + (size_t)p * half, *sn = rope_sin + (size_t)p * half;

  const int q_units = n_heads * hq, k_units = n_kv * hq, v_units = nk >> 2;
  for (int u = threadIdx.x; u < q_units + k_units + v_units; u += blockDim.x) {
    if (u < q_units + k_units) {
      const bool is_q = u < q_units;
      const int uu = is_q ? u : u - q_units;
      const int h = uu / hq, i = (uu - h * hq) * 4;
      const int col = (is_q ? 0 : nq) + h * d + i;
      const float4 a = sum_slices(base + col, n_partial, partial_stride);
      const float4 b = sum_slices(base + col + half, n_partial, partial_stride);
      const float4 c = *reinterpret_cast<const float4 *>(cs + i), sv = *reinterpret_cast<const float4 *>(sn + i);
      uint2 lo, hi;
      lo.x = float2_to_bf16x2(a.x * c.x - b.x * sv.x, a.y * c.y - b.y * sv.y);
      lo.y = float2_to_bf16x2(a.z * c.z - b.z * sv.z, a.w * c.w - b.w * sv.w);
      hi.x = float2_to_bf16x2(b.x * c.x + a.x * sv.x, b.y * c.y + a.y * sv.y);
      hi.y = float2_to_bf16x2(b.z * c.z + a.z * sv.z, b.w * c.w + a.w * sv.w);
      uint16_t *dst = is_q ? q + (size_t)t * nq + h * d + i : kpage + ((size_t)h * page_size + slot) * d + i;
      *reinterpret_cast<uint2 *>(dst) = lo;
      *reinterpret_cast<uint2 *>(dst + half) = hi;
    } else {
      const int uu = (u - q_units - k_units) * 4;
      const int h = uu / d, i = uu - h * d;
      const float4 a = sum_slices(base + nq + nk + uu, n_partial, partial_stride);
      uint2 o;
      o.x = float2_to_bf16x2(a.x, a.y);
      o.y = float2_to_bf16x2(a.z, a.w);
      *reinterpret_cast<uint2 *>(vpage + ((size_t)h * page_size + slot) * d + i) = o;
    }
  }
}

cudaError_t launch_rope_kv_write(const float *qkv, int n_partial, long long partial_stride, uint16_t *q,
                                 uint16_t *kv_pool, const int *page_table, int max_pages, const int *pt_row,
                                 const int *pos, const float *rope_cos, const float *rope_sin, int T,
                                 int n_heads, int n_kv, int d, int page_size, cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  if (d % 8 != 0) { set_last_error("rope_kv_write: head_dim must be a multiple of 8"); return cudaErrorInvalidValue; }
  const int units = (n_heads + n_kv) * (d / 8) + n_kv * d / 4;
  int threads = (units + 31) / 32 * 32;
  threads = threads > 1024 ? 1024 : threads;
  if (T >= wide_rows_threshold() && threads > 256) threads = 256;  // streaming job: many small CTAs per SM hide each other's index chasing
  return launch_k(rope_kv_write_kernel, dim3(T), dim3(threads), 0, st, qkv, n_partial, partial_stride, q, kv_pool, page_table,
                  max_pages, pt_row, pos, rope_cos, rope_sin, n_heads, n_kv, d, page_size);
}

// ------------------------------------------------------------------------------ sampler
constexpr int kSampThreads = 1024;

struct ArgMax { float v; int i; };
__device__ __forceinline__ ArgMax argmax_better(ArgMax a, ArgMax b) {
  return (b.v > a.v || (b.v == a.v && b.i < a.i)) ? b : a;
}
__device__ __forceinline__ ArgMax warp_argmax(ArgMax a) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ArgMax b;
    b.v = __shfl_xor_sync(0xffffffffu, a.v, o);
    b.i = __shfl_xor_sync(0xffffffffu, a.i, o);
    a = argmax_better(a, b);
  }
  return a;
}

// kept-mass of { i : bits(p_i) >= thr } for one row; p_i = exp(z_i - m) * inv_z
__device__ __forceinline__ float mass_at_least(const float *row, int V, float inv_t, float m, float inv_z,
                                               uint32_t thr, float *red) {
  float s = 0.f;
  for (int i = threadIdx.x; i < V; i += kSampThreads) {
    const float p = expf(row[i] * inv_t - m) * inv_z;
    if (__float_as_uint(p) >= thr) s += p;
  }
  return block_sum<kSampThreads>(s, red);
}

__global__ void __launch_bounds__(kSampThreads)
sample_kernel(const float *__restrict__ logits, long long ld, int V, const float *__restrict__ temperature,
              const float *__restrict__ top_p, const unsigned long long *__restrict__ seeds,
              const unsigned long long *__restrict__ steps, uint32_t *__restrict__ out, int vocab_offset,
              float *__restrict__ out_val, float *__restrict__ out_sumexp, float *__restrict__ out_logprob) {
  __shared__ float red[32];
  __shared__ ArgMax red_am[32];
  __shared__ float scan[kSampThreads / 32];
  __shared__ int s_tok;
  const int t = blockIdx.x;
  pdl_wait();
  const float *row = logits + (size_t)t * ld;
  const float temp = temperature ? temperature[t] : 0.f;

  // ---- argmax (always: it is also the max for the softmax)
  ArgMax am{-INFINITY, 0x7fffffff};
  for (int i = threadIdx.x; i < V; i += kSampThreads) {
    const float v = row[i];
    if (v > am.v) { am.v = v; am.i = i; }  // increasing i: first max kept
  }
  am = warp_argmax(am);
  if ((threadIdx.x & 31) == 0) red_am[threadIdx.x >> 5] = am;
  __syncthreads();
  if (threadIdx.x < 32) {
    am = red_am[threadIdx.x];
    am = warp_argmax(am);
    if (threadIdx.x == 0) red_am[0] = am;
  }
  __syncthreads();
  am = red_am[0];
  if (temp <= 0.f) {
    // TokenEvent.logprob (crates/core/src/models.rs:276): log softmax(logits)[token].  The argmax element has
    // z - max = 0, so its log-probability is -log(sum exp(z - max)).  A vocab-parallel rank reports the sum over
    // its own slice relative to its own max (out_sumexp); the candidate merge combines the slices.
    float se = 0.f;
    if (out_sumexp || out_logprob) {
      for (int i = threadIdx.x; i < V; i += kSampThreads) se += expf(row[i] - am.v);
      se = block_sum<kSampThreads>(se, red);
    }
    if (threadIdx.x == 0) {
      out[t] = (uint32_t)(am.i + vocab_offset);
      if (out_val) out_val[t] = am.v;
      if (out_sumexp) out_sumexp[t] = se;
      if (out_logprob) out_logprob[t] = -logf(se);
    }
    return;
  }

  // ---- temperature softmax + nucleus threshold
  const float inv_t = 1.0f / temp;
  // max of z = logits*inv_t; inv_t > 0 so it is the argmax element's z
  float m = -INFINITY;
  for (int i = threadIdx.x; i < V; i += kSampThreads) m = fmaxf(m, row[i] * inv_t);
  m = block_max<kSampThreads>(m, red);
  float z = 0.f;
  for (int i = threadIdx.x; i < V; i += kSampThreads) z += expf(row[i] * inv_t - m);
  z = block_sum<kSampThreads>(z, red);
  const float inv_z = 1.0f / z;
  const float tp = top_p ? top_p[t] : 1.0f;
  uint32_t tau = 0u;
  float kept = 1.0f;
  if (tp < 1.0f) {
    // largest threshold whose kept mass still reaches top_p (bit patterns of positive floats are ordered)
    uint32_t lo = 0u, hi = 0x3f800001u;  // mass(lo) >= tp holds for lo = 0
    while (hi - lo > 1u) {
      const uint32_t mid = lo + ((hi - lo) >> 1);
      const float ms = mass_at_least(row, V, inv_t, m, inv_z, mid, red);
      if (ms >= tp) lo = mid; else hi = mid;
    }
    tau = lo;
  }
  kept = mass_at_least(row, V, inv_t, m, inv_z, tau, red);

  // ---- draw: first kept token, in index order, whose running mass exceeds u * kept
  const float u = hash_u01(mix64(seeds ? seeds[t] : 0ull, 0xC0FFEEull, steps ? steps[t] : 0ull));
  const float target = u * kept;
  const int chunk = (V + kSampThreads - 1) / kSampThreads;
  const int i0 = threadIdx.x * chunk, i1 = min(V, i0 + chunk);
  float local = 0.f;
  for (int i = i0; i < i1; i++) {
    const float p = expf(row[i] * inv_t - m) * inv_z;
    if (__float_as_uint(p) >= tau) local += p;
  }
  // exclusive scan over threads
  float incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float n = __shfl_up_sync(0xffffffffu, incl, o);
    if ((threadIdx.x & 31) >= o) incl += n;
  }
  if (threadIdx.x == 0) s_tok = -1;
  if ((threadIdx.x & 31) == 31) scan[threadIdx.x >> 5] = incl;
  __syncthreads();
  if (threadIdx.x < 32) {
    float w = scan[threadIdx.x];
    float wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float n = __shfl_up_sync(0xffffffffu, wi, o);
      if (threadIdx.x >= o) wi += n;
    }
    scan[threadIdx.x] = wi - w;  // exclusive warp offsets
  }
  __syncthreads();
  const float excl = scan[threadIdx.x >> 5] + incl - local;
  if (local > 0.f && excl <= target && target < excl + local) {
    float run = excl;
    int tok = -1, last = -1;
    for (int i = i0; i < i1; i++) {
      const float p = expf(row[i] * inv_t - m) * inv_z;
      if (__float_as_uint(p) >= tau) { run += p; last = i; if (run > target) { tok = i; break; } }
    }
    if (tok < 0) tok = last;
    atomicMax(&s_tok, tok);  // intervals are disjoint: at most one thread gets here
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int tok = s_tok;
    if (tok < 0) tok = am.i;  // rounding left no interval: fall back to the mode
    out[t] = (uint32_t)(tok + vocab_offset);
    if (out_val) out_val[t] = row[tok];
    // log-probability of the drawn token under softmax(logits / temperature), before the nucleus cut
    if (out_logprob) out_logprob[t] = row[tok] * inv_t - m - logf(z);
  }
}

cudaError_t launch_sample(const float *logits, long long ld, int T, int V, const float *temperature,
                          const float *top_p, const unsigned long long *seeds, const unsigned long long *steps,
                          uint32_t *out_tokens, int vocab_offset, float *out_val, float *out_sumexp, float *out_logprob,
                          cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  return launch_k(sample_kernel, dim3(T), dim3(kSampThreads), 0, st, logits, ld, V, temperature, top_p, seeds, steps, out_tokens,
                  vocab_offset, out_val, out_sumexp, out_logprob);
}

// Final step of a vocab-parallel greedy argmax: cand_val/cand_tok are [tp][T] (all-gathered).
__global__ void argmax_merge_kernel(const float *__restrict__ cand_val, const uint32_t *__restrict__ cand_tok,
                                    const float *__restrict__ cand_sum, int tp, int T, uint32_t *__restrict__ out,
                                    float *__restrict__ out_logprob) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  float bv = cand_val[t];
  uint32_t bt = cand_tok[t];
  for (int r = 1; r < tp; r++) {
    const float v = cand_val[(size_t)r * T + t];
    const uint32_t k = cand_tok[(size_t)r * T + t];
    if (v > bv || (v == bv && k < bt)) { bv = v; bt = k; }
  }
  out[t] = bt;
  if (out_logprob && cand_sum) {  // the winner's value is the global max: log p = -log sum_r s_r * exp(m_r - max)
    float se = 0.f;
    for (int r = 0; r < tp; r++) se += cand_sum[(size_t)r * T + t] * expf(cand_val[(size_t)r * T + t] - bv);
    out_logprob[t] = -logf(se);
  }
}

cudaError_t launch_argmax_merge(const float *cand_val, const uint32_t *cand_tok, const float *cand_sum, int tp, int T,
                                uint32_t *out, float *out_logprob, cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  argmax_merge_kernel<<<(T + 127) / 128, 128, 0, st>>>(cand_val, cand_tok, cand_sum, tp, T, out, out_logprob);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------ synthetic weights
__global__ void synth_bf16_kernel(uint16_t *__restrict__ dst, size_t n, unsigned long long seed,
                                  unsigned long long key, float amp, float offset, size_t idx0,
                                  size_t row_len, size_t src_ld, size_t src_col0) {
  // element j of dst (row-major [rows][row_len]) is element idx0 + (j / row_len) * src_ld + src_col0 + j % row_len
  // of the full tensor: lets a tensor-parallel shard be generated in place.
  for (size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x; j < n; j += (size_t)gridDim.x * blockDim.x) {
    const size_t r = j / row_len, c = j - r * row_len;
    const size_t i = idx0 + r * src_ld + src_col0 + c;
    const float u = hash_u01(mix64(seed, key, i));
    const float s = u * 2.0f - 1.0f;  // exact
    const float v = __fadd_rn(__fmul_rn(s, amp), offset);
    dst[j] = float_to_bf16_bits(v);
  }
}

cudaError_t launch_synth_bf16(uint16_t *dst, size_t n, unsigned long long seed, int kind, int layer, float amp,
                              float offset, size_t idx0, size_t row_len, size_t src_ld, size_t src_col0,
                              cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  const unsigned long long key = (unsigned long long)layer * 16ull + (unsigned long long)kind;
  const int blocks = (int)min((size_t)kNumSMs * 16, (n + 255) / 256);
  synth_bf16_kernel<<<blocks, 256, 0, st>>>(dst, n, seed, key, amp, offset, idx0, row_len, src_ld, src_col0);
  return cudaGetLastError();
}

__global__ void bf16_to_f32_kernel(const uint16_t *__restrict__ src, float *__restrict__ dst, size_t n) {
  for (size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x; j < n; j += (size_t)gridDim.x * blockDim.x)
    dst[j] = bf16_bits_to_float(src[j]);
}
cudaError_t launch_bf16_to_f32(const uint16_t *src, float *dst, size_t n, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  bf16_to_f32_kernel<<<(int)min((size_t)1024, (n + 255) / 256), 256, 0, st>>>(src, dst, n);
  return cudaGetLastError();
}

}  // namespace dis
