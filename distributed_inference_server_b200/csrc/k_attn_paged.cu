// k_attn_paged.cu -- attention of single query tokens over a paged bf16 KV cache.
//
// One CTA handles one (query token, kv head, split of the history) for up to 4 query heads of the
// GQA group, so K and V are read from HBM once per group.  64-token K/V tiles are staged into
// shared memory through a 3-4 stage cp.async ring (64-96 KB in flight per CTA, two CTAs per SM);
// scores use 16 elements per lane and a short shuffle reduction, the softmax is online across
// tiles, and P.V is accumulated with 8 output dims per thread.  HBM-bound: algorithmic
// bytes = 2 (K,V) * ctx * n_kv * d * 2 B per query token (+ q and out, negligible).
//
// KV page pool layout, per layer: [page][2 (K,V)][n_kv][page_size][d] bf16 -- each (page, kv head)
// is one contiguous page_size*d block.
//
// With n_splits > 1 every split writes (m, l, unnormalised o) to scratch and a combine pass merges
// them (flash-decoding); that keeps all 148 SMs busy at batch 1.
#include "kernels.h"

namespace dis {

constexpr int kAttnThreads = 128;
constexpr int kAttnTile = 64;  // tokens per pipeline stage

template <int D>
struct AttnCfg {
  static constexpr int kStageBytes = 2 * kAttnTile * D * 2;              // K tile + V tile, bf16
  static constexpr int kStages = D == 128 ? 3 : 4;                       // 96 KB / 64 KB -> 2+ CTAs per SM
  static constexpr int kSmemBytes = kStages * kStageBytes;
};

__device__ __forceinline__ void cp_async_16(uint32_t dst, const void *src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int D, int G>
__global__ void __launch_bounds__(kAttnThreads, 2)
attn_paged_kernel(const uint16_t *__restrict__ q, const uint16_t *__restrict__ kv_pool,
                  const int *__restrict__ page_table, int max_pages, const int *__restrict__ pt_row,
                  const int *__restrict__ ctx_len, uint16_t *__restrict__ out, float *__restrict__ scratch,
                  int n_splits, int T, int n_heads, int n_kv, int page_size, int groups_per_kv) {
  using Cfg = AttnCfg<D>;
  constexpr int LPT = D / 16;              // lanes per token in the QK phase (16 elements each)
  constexpr int TPW = 32 / LPT;            // tokens per warp per pass
  constexpr int TPI = TPW * (kAttnThreads / 32);
  constexpr int CPR = D / 8;               // 16-byte chunks per K/V row
  constexpr int RPAR = kAttnThreads / CPR;  // V rows processed in parallel
  constexpr int kChunks = 2 * kAttnTile * CPR;  // 16-byte chunks per stage

  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(16) float s_p[kAttnTile][G];
  __shared__ float s_scale[G], s_m[G], s_l[G];

  const int t = blockIdx.z;
  const int kvh = blockIdx.y / groups_per_kv;
  const int sub = blockIdx.y - kvh * groups_per_kv;
  const int split = blockIdx.x;
  const int heads_per_kv = n_heads / n_kv;
  const int h0 = kvh * heads_per_kv + sub * G;  // first query head of this CTA
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  const int ctx = ctx_len[t];
  int chunk = (ctx + n_splits - 1) / n_splits;
  chunk = (chunk + kAttnTile - 1) / kAttnTile * kAttnTile;
  const int begin = split * chunk;
  const int end = min(ctx, begin + chunk);
  const int n_tiles = end > begin ? (end - begin + kAttnTile - 1) / kAttnTile : 0;

  const int *pt = page_table + (size_t)pt_row[t] * max_pages;
  const size_t page_stride = (size_t)2 * n_kv * page_size * D;
  const uint16_t *kbase = kv_pool + (size_t)kvh * page_size * D;
  const size_t v_off = (size_t)n_kv * page_size * D;
  const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(smem);

  // stage <- tokens [begin + tile*64, +64): K rows then V rows, [row][D] bf16, 16 B per cp.async
  auto load_tile = [&](int tile, int stage) {
    const int tile0 = begin + tile * kAttnTile;
    const uint32_t sbase = smem_base + stage * Cfg::kStageBytes;
#pragma unroll
    for (int id = threadIdx.x; id < kChunks; id += kAttnThreads) {
      const int kv = id / (kAttnTile * CPR);
      const int row = (id / CPR) % kAttnTile;
      const int c = id % CPR;
      const int pos = tile0 + row;
      const bool valid = pos < end;
      const uint16_t *src = kbase;
      if (valid) src = kbase + (size_t)pt[pos / page_size] * page_stride + (kv ? v_off : 0) + (size_t)(pos % page_size) * D + c * 8;
      cp_async_16(sbase + (uint32_t)(kv * kAttnTile * D * 2 + row * D * 2 + c * 16), src, valid ? 16 : 0);
    }
  };

  // ---- q fragment: elements [j*8, j*8+8) and [D/2 + j*8, D/2 + j*8 + 8) of each head, fp32 registers
  const int j = lane % LPT, grp = lane / LPT;
  float qf[G][16];
#pragma unroll
  for (int g = 0; g < G; g++) {
    const uint16_t *qp = q + ((size_t)t * n_heads + h0 + g) * D;
    const uint4 a = *reinterpret_cast<const uint4 *>(qp + j * 8);
    const uint4 b = *reinterpret_cast<const uint4 *>(qp + D / 2 + j * 8);
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 8; i++) { const float2 f = bf16x2_to_float2(w[i]); qf[g][2 * i] = f.x; qf[g][2 * i + 1] = f.y; }
  }
  const float scale = rsqrtf((float)D);

  float acc[G][8];
#pragma unroll
  for (int g = 0; g < G; g++)
#pragma unroll
    for (int i = 0; i < 8; i++) acc[g][i] = 0.f;
  if (threadIdx.x < G) { s_m[threadIdx.x] = -INFINITY; s_l[threadIdx.x] = 0.f; }

  // ---- pipeline prologue
#pragma unroll
  for (int s = 0; s < Cfg::kStages - 1; s++) {
    if (s < n_tiles) load_tile(s, s);
    cp_async_commit();
  }
  const int vr = threadIdx.x / CPR, vc = threadIdx.x % CPR;

  for (int tile = 0; tile < n_tiles; tile++) {
    cp_async_wait<Cfg::kStages - 2>();
    __syncthreads();  // tile's bytes are visible to all; the stage consumed last iteration is free
    {
      const int nt = tile + Cfg::kStages - 1;
      if (nt < n_tiles) load_tile(nt, nt % Cfg::kStages);
      cp_async_commit();
    }
    const int stage = tile % Cfg::kStages;
    const uint8_t *ks = smem + stage * Cfg::kStageBytes;
    const uint8_t *vs = ks + kAttnTile * D * 2;
    const int tile_len = min(kAttnTile, end - (begin + tile * kAttnTile));
    // ---- scores: 16 elements per lane, LPT lanes per token
#pragma unroll
    for (int base = warp * TPW; base < kAttnTile; base += TPI) {
      const int tok = base + grp;
      const uint4 k0 = *reinterpret_cast<const uint4 *>(ks + tok * D * 2 + j * 16);
      const uint4 k1 = *reinterpret_cast<const uint4 *>(ks + tok * D * 2 + D + j * 16);
      const uint32_t w[8] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w};
      float kf[16];
#pragma unroll
      for (int i = 0; i < 8; i++) { const float2 f = bf16x2_to_float2(w[i]); kf[2 * i] = f.x; kf[2 * i + 1] = f.y; }
#pragma unroll
      for (int g = 0; g < G; g++) {
        float d = 0.f;
#pragma unroll
        for (int i = 0; i < 16; i++) d = fmaf(qf[g][i], kf[i], d);
#pragma unroll
        for (int o = LPT / 2; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        if (j == 0) s_p[tok][g] = tok < tile_len ? d * scale : -INFINITY;
      }
    }
    __syncthreads();
    // ---- online softmax bookkeeping: warp g owns head g
    if (warp < G) {
      const float a = s_p[lane][warp], b = s_p[lane + 32][warp];
      const float mx = warp_max(fmaxf(a, b));
      const float m_old = s_m[warp];
      const float m_new = fmaxf(m_old, mx);
      const float pa = expf(a - m_new), pb = expf(b - m_new);  // exp(-inf) = 0 for masked tokens
      s_p[lane][warp] = pa;
      s_p[lane + 32][warp] = pb;
      const float sum = warp_sum(pa + pb);
      if (lane == 0) {
        const float sc = expf(m_old - m_new);  // 0 on the first tile
        s_scale[warp] = sc;
        s_m[warp] = m_new;
        s_l[warp] = s_l[warp] * sc + sum;
      }
    }
    __syncthreads();
    // ---- P.V: 8 output dims per thread, RPAR rows in parallel
#pragma unroll
    for (int g = 0; g < G; g++) {
      const float sc = s_scale[g];
#pragma unroll
      for (int i = 0; i < 8; i++) acc[g][i] *= sc;
    }
#pragma unroll 4
    for (int tok = vr; tok < tile_len; tok += RPAR) {
      const uint4 v = *reinterpret_cast<const uint4 *>(vs + tok * D * 2 + vc * 16);
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
      float vf[8];
#pragma unroll
      for (int i = 0; i < 4; i++) { const float2 f = bf16x2_to_float2(w[i]); vf[2 * i] = f.x; vf[2 * i + 1] = f.y; }
      float p[G];
#pragma unroll
      for (int g = 0; g < G; g++) p[g] = s_p[tok][g];
#pragma unroll
      for (int g = 0; g < G; g++)
#pragma unroll
        for (int i = 0; i < 8; i++) acc[g][i] = fmaf(p[g], vf[i], acc[g][i]);
    }
  }
  cp_async_wait<0>();
  __syncthreads();

  // ---- reduce the RPAR row slots through the (now free) stage memory
  float *s_red = reinterpret_cast<float *>(smem);  // [RPAR][G][D]
#pragma unroll
  for (int g = 0; g < G; g++)
#pragma unroll
    for (int i = 0; i < 8; i++) s_red[(vr * G + g) * D + vc * 8 + i] = acc[g][i];
  __syncthreads();
  for (int e = threadIdx.x; e < G * D; e += kAttnThreads) {
    const int g = e / D, dd = e - g * D;
    float o = 0.f;
#pragma unroll
    for (int r = 0; r < RPAR; r++) o += s_red[(r * G + g) * D + dd];
    const int h = h0 + g;
    if (n_splits == 1) {
      out[((size_t)t * n_heads + h) * D + dd] = float_to_bf16_bits(o / s_l[g]);
    } else {
      // scratch: ml [T][n_heads][S][2] then o [T][n_heads][S][D]
      float *ml = scratch + (((size_t)t * n_heads + h) * n_splits + split) * 2;
      float *op = scratch + (size_t)T * n_heads * n_splits * 2 + (((size_t)t * n_heads + h) * n_splits + split) * D;
      op[dd] = o;
      if (dd == 0) { ml[0] = s_m[g]; ml[1] = s_l[g]; }
    }
  }
}

template <int D>
__global__ void attn_combine_kernel(const float *__restrict__ scratch, uint16_t *__restrict__ out, int n_splits,
                                    int T, int n_heads) {
  const int t = blockIdx.y, h = blockIdx.x, dd = threadIdx.x;
  const float *ml = scratch + ((size_t)t * n_heads + h) * n_splits * 2;
  const float *op = scratch + (size_t)T * n_heads * n_splits * 2 + ((size_t)t * n_heads + h) * n_splits * D;
  float M = -INFINITY;
  for (int s = 0; s < n_splits; s++) M = fmaxf(M, ml[2 * s]);
  float L = 0.f, o = 0.f;
  for (int s = 0; s < n_splits; s++) {
    const float w = expf(ml[2 * s] - M);
    L += ml[2 * s + 1] * w;
    o += op[(size_t)s * D + dd] * w;
  }
  out[((size_t)t * n_heads + h) * D + dd] = float_to_bf16_bits(o / L);
}

size_t attn_scratch_elems(int T, int n_heads, int d, int n_splits) {
  if (n_splits <= 1) return 0;
  return (size_t)T * n_heads * n_splits * (2 + (size_t)d);
}

template <int D, int G>
static cudaError_t attn_launch(const uint16_t *q, const uint16_t *kv_pool, const int *page_table, int max_pages,
                               const int *pt_row, const int *ctx, uint16_t *out, float *scratch, int n_splits,
                               int T, int n_heads, int n_kv, int page_size, int groups_per_kv, cudaStream_t st) {
  dim3 grid(n_splits, n_kv * groups_per_kv, T);
  static bool attr_set = false;
  if (!attr_set) {
    DIS_CUDA_TRY(cudaFuncSetAttribute(attn_paged_kernel<D, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnCfg<D>::kSmemBytes));
    attr_set = true;
  }
  attn_paged_kernel<D, G><<<grid, kAttnThreads, AttnCfg<D>::kSmemBytes, st>>>(q, kv_pool, page_table, max_pages, pt_row, ctx, out,
                                                         scratch, n_splits, T, n_heads, n_kv, page_size,
                                                         groups_per_kv);
  DIS_CUDA_TRY(cudaGetLastError());
  if (n_splits > 1) {
    attn_combine_kernel<D><<<dim3(n_heads, T), D, 0, st>>>(scratch, out, n_splits, T, n_heads);
    DIS_CUDA_TRY(cudaGetLastError());
  }
  return cudaSuccess;
}

cudaError_t launch_attn_paged(const uint16_t *q, const uint16_t *kv_pool, const int *page_table, int max_pages,
                              const int *pt_row, const int *ctx, uint16_t *out, float *scratch, int n_splits,
                              int T, int n_heads, int n_kv, int d, int page_size, cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  if (n_kv <= 0 || n_heads % n_kv != 0) { set_last_error("attn_paged: n_heads must be a multiple of n_kv"); return cudaErrorInvalidValue; }
  if (n_splits < 1 || (n_splits > 1 && !scratch)) { set_last_error("attn_paged: n_splits > 1 needs scratch"); return cudaErrorInvalidValue; }
  if (T > 65535) { set_last_error("attn_paged: T > 65535 tokens per launch"); return cudaErrorInvalidValue; }
  const int hpk = n_heads / n_kv;
  const int G = hpk >= 4 ? 4 : hpk;  // query heads per CTA
  if (hpk % G != 0 || (G != 1 && G != 2 && G != 4)) { set_last_error("attn_paged: GQA group must be 1, 2 or a multiple of 4"); return cudaErrorInvalidValue; }
  const int gpk = hpk / G;
#define DIS_ATTN_CASE(DD, GG)                                                                                  \
  if (d == DD && G == GG)                                                                                      \
    return attn_launch<DD, GG>(q, kv_pool, page_table, max_pages, pt_row, ctx, out, scratch, n_splits, T,      \
                               n_heads, n_kv, page_size, gpk, st);
  DIS_ATTN_CASE(128, 4) DIS_ATTN_CASE(128, 2) DIS_ATTN_CASE(128, 1)
  DIS_ATTN_CASE(64, 4) DIS_ATTN_CASE(64, 2) DIS_ATTN_CASE(64, 1)
#undef DIS_ATTN_CASE
  set_last_error("attn_paged: head_dim must be 64 or 128");
  return cudaErrorInvalidValue;
}

}  // namespace dis
