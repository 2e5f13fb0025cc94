// k_attn_paged.cu -- attention of single query tokens over a paged bf16 KV cache.
//
// One CTA handles one (query token, kv head, split of the history) for all query heads of the GQA
// group, so K and V are read from HBM once per group.  64-token K/V tiles are staged by TMA into
// 128B-swizzled shared memory through a 3-4 stage mbarrier ring (64-96 KB in flight per CTA, two
// CTAs per SM); Q K^T and P V run on mma.sync m16n8k16 (bf16 in, fp32 accumulate) with ldmatrix operand
// loads -- on CUDA cores the 4 FMA per KV byte of a 4:1 GQA group is issue-bound well below the HBM
// rate (measured: 56 us vs 5 us per layer at batch 32, ctx 512).  HBM-bound: algorithmic
// bytes = 2 (K,V) * ctx * n_kv * d * 2 B per query token (+ q and out, negligible).
//
// KV page pool layout, per layer: [page][2 (K,V)][n_kv][page_size][d] bf16 -- each (page, kv head)
// is one contiguous page_size*d block.
//
// With n_splits > 1 every split writes (m, l, unnormalised o) to scratch and a combine pass merges
// them (flash-decoding); that keeps all 148 SMs busy at batch 1.
#include <cstdlib>
#include <mutex>

#include "../../include/dis_engine.h"
#include "kernels.h"
#include "ptx.cuh"

namespace dis {

constexpr int kAttnThreads = 128;
constexpr int kAttnTile = 64;  // tokens per pipeline stage: 16 per warp

template <int D>
struct AttnCfg {
  static constexpr int kHalves = D / 64;                                  // 128-byte column blocks per row
  static constexpr int kHalfBytes = kAttnTile * 128;                      // one [64 tokens][64 dims] block
  static constexpr int kStageBytes = 2 * kHalves * kHalfBytes;            // K tile + V tile, bf16
  static constexpr int kStages = D == 128 ? 3 : 4;                        // 96 KB / 64 KB -> two CTAs per SM
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 64;    // + alignment slack + barriers
};

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
// D(16x8, f32) += A(16x16, bf16, row) * B(16x8, bf16, col)
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// One CTA = one (query token, kv head, split).  The GQA group's query heads (<= 8) are rows 0..G-1 of a
// 16-row MMA tile; rows G..15 are zero padding (the kernel is HBM-bound, the wasted MMA rows are free).
// K/V tiles of 64 tokens arrive by TMA (one elected thread, 128B-swizzled boxes of min(page,64) rows,
// mbarrier completion) into a 3-4 stage ring.  Each of the 4 warps owns 16 of a tile's 64 tokens and
// keeps its own online-softmax state (m, l, O[16 x D] in mma accumulator fragments); the four states
// are merged once at the end.  P is rounded to bf16 for the P.V MMA (the usual flash-attention choice).
// Rows past the context inside a box hold stale (finite) cache contents and get probability 0; the
// engine zero-fills the pool once so that they can never be NaN.
// Decode steps fuse the RoPE + KV-write stage into this kernel (rope.qkv != nullptr): the CTA sums the split-K
// slices of the fused QKV projection for its query heads and -- if the new position falls in its split -- for
// its kv head, rotates q and k, writes the new K/V row to the page pool for later steps and patches it into
// the shared-memory tile that TMA fetched.  Because nothing but that one row depends on this step, the K/V
// stream is started BEFORE griddepcontrol.wait: attention's HBM traffic overlaps the QKV projection's tail.
struct AttnRope {
  const float *qkv;          // [S][T][(nh + 2 nkv) * d] fp32 split-K slices, or nullptr: q / pages are ready
  int n_partial;
  long long stride;
  const int *pos;            // [T]
  const float *cos_t, *sin_t;  // [max_ctx][d/2]
  uint16_t *kv_base;         // this layer's page pool
};

__device__ __forceinline__ float4 attn_sum_slices(const float *p, int n, long long stride) {
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
  for (int s = 0; s < n; s++) {
    const float4 v = *reinterpret_cast<const float4 *>(p + (size_t)s * stride);
    a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
  }
  return a;
}

template <int D, int G>
__global__ void __launch_bounds__(kAttnThreads, 2)
attn_paged_kernel(const __grid_constant__ CUtensorMap kv_map, long long row0, const AttnRope rope, const uint16_t *__restrict__ q,
                  const int *__restrict__ page_table, int max_pages, const int *__restrict__ pt_row,
                  const int *__restrict__ ctx_len, uint16_t *__restrict__ out, float *__restrict__ scratch,
                  int n_splits, int T, int n_heads, int n_kv, int page_shift, int n_stages) {
  using Cfg = AttnCfg<D>;
  constexpr int KS = D / 16;  // k-steps of the QK^T MMA
  constexpr int NT = D / 8;   // n-tiles of the P.V MMA

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // 128B swizzle atoms are 1024-byte aligned
  uint8_t *smem = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t bar_base = smem_base + n_stages * Cfg::kStageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };

  const int t = blockIdx.z;
  const int kvh = blockIdx.y;
  const int split = blockIdx.x;
  const int h0 = kvh * G;  // first query head of this CTA (G = n_heads / n_kv)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, tq = lane & 3;  // mma fragment coordinates

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&kv_map);
    for (int s = 0; s < n_stages; s++) mbar_init(full_bar(s), 1);
    fence_barrier_init();
    fence_proxy_async();
  }
  pdl_launch_dependents();
  const bool fused = rope.qkv != nullptr;
  __shared__ __align__(16) uint16_t s_q[G][D];
  __shared__ __align__(16) uint16_t s_kvnew[2][D];
  if (!fused) pdl_wait();  // q and this step's K/V rows come from the previous kernels
  __syncthreads();

  const int page_size = 1 << page_shift;
  const int box_rows = min(page_size, kAttnTile);
  const int ctx = ctx_len[t];
  int chunk = (ctx + n_splits - 1) / n_splits;
  chunk = (chunk + kAttnTile - 1) / kAttnTile * kAttnTile;
  const int begin = split * chunk;
  const int end = min(ctx, begin + chunk);
  const int n_tiles = end > begin ? (end - begin + kAttnTile - 1) / kAttnTile : 0;
  const int *pt = page_table + (size_t)pt_row[t] * max_pages;

  // pool rows: [page][2 (K,V)][n_kv][page_size] of d elements each
  auto load_tile = [&](int tile, int stage) {  // one thread
    const int tile0 = begin + tile * kAttnTile;
    const uint32_t sbase = smem_base + stage * Cfg::kStageBytes;
    mbar_expect_tx(full_bar(stage), Cfg::kStageBytes);
    for (int r = 0; r < kAttnTile; r += box_rows) {
      const int pos = tile0 + r;
      const long long page = pt[min(pos >> page_shift, max_pages - 1)];
      const long long krow = row0 + ((page * 2 + 0) * n_kv + kvh) * page_size + (pos & (page_size - 1));
      const long long vrow = krow + (long long)n_kv * page_size;
#pragma unroll
      for (int h = 0; h < Cfg::kHalves; h++) {
        tma_load_2d(sbase + h * Cfg::kHalfBytes + r * 128, &kv_map, h * 64, (int)krow, full_bar(stage), kEvictFirst);
        tma_load_2d(sbase + (Cfg::kHalves + h) * Cfg::kHalfBytes + r * 128, &kv_map, h * 64, (int)vrow, full_bar(stage), kEvictFirst);
      }
    }
  };

  // One stage is kept back for the refill at the top of iteration 0.  Filling the whole ring up front (nothing consumes a
  // tile before the RoPE prologue is over) was measured in round 2: 3.491 / 3.495 against 3.531 / 3.479 ms per step at
  // batch 32, 4.269 against 4.253 at batch 64 -- no gain, not kept.
  if (threadIdx.x == 0) {
    for (int s = 0; s < n_stages - 1; s++)
      if (s < n_tiles) load_tile(s, s);
  }

  const int new_pos = ctx - 1;                               // the position this decode step appends
  const bool owner = fused && new_pos >= begin && new_pos < end;  // the split that holds it
  if (fused) {
    pdl_wait();  // the QKV projection's slices
    constexpr int HQ = D / 8;  // float4 rotate-half items per head
    const int nq = n_heads * D, nk = n_kv * D;
    const float *base = rope.qkv + (size_t)t * (nq + 2 * nk);
    const float *cs = rope.cos_t + (size_t)new_pos * (D / 2), *sn = rope.sin_t + (size_t)new_pos * (D / 2);
    for (int item = threadIdx.x; item < (G + 1) * HQ + D / 4; item += kAttnThreads) {
      if (item < (G + 1) * HQ) {
        const bool is_q = item < G * HQ;
        if (!is_q && !owner) continue;
        const int hh = is_q ? item / HQ : 0, i = (item % HQ) * 4;
        const int col = is_q ? (h0 + hh) * D + i : nq + kvh * D + i;
        const float4 a = attn_sum_slices(base + col, rope.n_partial, rope.stride);
        const float4 b = attn_sum_slices(base + col + D / 2, rope.n_partial, rope.stride);
        const float4 c = *reinterpret_cast<const float4 *>(cs + i), sv = *reinterpret_cast<const float4 *>(sn + i);
        uint2 lo, hi;
        lo.x = float2_to_bf16x2(a.x * c.x - b.x * sv.x, a.y * c.y - b.y * sv.y);
        lo.y = float2_to_bf16x2(a.z * c.z - b.z * sv.z, a.w * c.w - b.w * sv.w);
        hi.x = float2_to_bf16x2(b.x * c.x + a.x * sv.x, b.y * c.y + a.y * sv.y);
        hi.y = float2_to_bf16x2(b.z * c.z + a.z * sv.z, b.w * c.w + a.w * sv.w);
        uint16_t *dst = is_q ? s_q[hh] : s_kvnew[0];
        *reinterpret_cast<uint2 *>(dst + i) = lo;
        *reinterpret_cast<uint2 *>(dst + D / 2 + i) = hi;
      } else if (owner) {
        const int i = (item - (G + 1) * HQ) * 4;
        const float4 a = attn_sum_slices(base + nq + nk + kvh * D + i, rope.n_partial, rope.stride);
        uint2 o;
        o.x = float2_to_bf16x2(a.x, a.y);
        o.y = float2_to_bf16x2(a.z, a.w);
        *reinterpret_cast<uint2 *>(s_kvnew[1] + i) = o;
      }
    }
    __syncthreads();
    if (owner && threadIdx.x < 2 * (D / 8)) {  // the new row goes to the page pool for the steps to come
      const int kv = threadIdx.x / (D / 8), c = threadIdx.x % (D / 8);
      const long long page = pt[new_pos >> page_shift];
      uint16_t *dst = rope.kv_base + ((((size_t)page * 2 + kv) * n_kv + kvh) * page_size + (new_pos & (page_size - 1))) * D + c * 8;
      *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(s_kvnew[kv] + c * 8);
    }
  }

  // ---- Q as the A operand: row g = head g (g < G), everything else zero
  uint32_t qa[KS][2];
#pragma unroll
  for (int ks = 0; ks < KS; ks++) {
    qa[ks][0] = 0u; qa[ks][1] = 0u;
    if (g < G) {
      const uint16_t *qp = fused ? s_q[g] + ks * 16 + tq * 2 : q + ((size_t)t * n_heads + h0 + g) * D + ks * 16 + tq * 2;
      qa[ks][0] = *reinterpret_cast<const uint32_t *>(qp);
      qa[ks][1] = *reinterpret_cast<const uint32_t *>(qp + 8);
    }
  }
  const float scale_log2 = rsqrtf((float)D) * 1.4426950408889634f;  // scores are kept in the log2 domain

  float oacc[NT][4];
#pragma unroll
  for (int n = 0; n < NT; n++) { oacc[n][0] = 0.f; oacc[n][1] = 0.f; oacc[n][2] = 0.f; oacc[n][3] = 0.f; }
  float m_run = -INFINITY, l_run = 0.f;  // row g; l_run is this thread's partial sum (quad-reduced at the end)

  const int wtok = warp * 16;  // this warp's token slice inside a tile
  for (int tile = 0; tile < n_tiles; tile++) {
    __syncthreads();  // everyone is done with the stage consumed last iteration
    if (threadIdx.x == 0) {
      const int nt = tile + n_stages - 1;
      if (nt < n_tiles) load_tile(nt, nt % n_stages);
    }
    const int stage = tile % n_stages;
    mbar_wait(full_bar(stage), (uint32_t)(tile / n_stages) & 1u);
    const uint32_t ks_addr = smem_base + stage * Cfg::kStageBytes;
    const uint32_t vs_addr = ks_addr + Cfg::kHalves * Cfg::kHalfBytes;
    const int tile_len = min(kAttnTile, end - (begin + tile * kAttnTile));
    if (owner && tile == n_tiles - 1) {
      // the tile TMA fetched predates this step's K/V row: patch it in (same 128B swizzle as the TMA boxes)
      const int r = new_pos - (begin + tile * kAttnTile);
      if (threadIdx.x < 2 * (D / 8)) {
        const int kv = threadIdx.x / (D / 8), c = threadIdx.x % (D / 8);
        const uint32_t dst = (kv ? vs_addr : ks_addr) + (c >> 3) * Cfg::kHalfBytes + r * 128 + (((c & 7) ^ (r & 7)) << 4);
        const uint4 v = *reinterpret_cast<const uint4 *>(s_kvnew[kv] + c * 8);
        asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
      }
      __syncthreads();
    }
    if (wtok >= tile_len) continue;  // warp-uniform: nothing of this tile belongs to this warp

    // ---- S = Q K^T for this warp's 16 tokens (two n-tiles of 8)
    float sacc[2][4];
#pragma unroll
    for (int n = 0; n < 2; n++) { sacc[n][0] = 0.f; sacc[n][1] = 0.f; sacc[n][2] = 0.f; sacc[n][3] = 0.f; }
    {
      const int mi = lane >> 3, r = lane & 7;
      const int tok = wtok + (mi >> 1) * 8 + r;  // matrices 0,1 -> n-tile 0; 2,3 -> n-tile 1
#pragma unroll
      for (int ks = 0; ks < KS; ks++) {
        const int c = ks * 2 + (mi & 1);  // 16-byte chunk of the row
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4(ks_addr + (c >> 3) * Cfg::kHalfBytes + tok * 128 + (((c & 7) ^ (tok & 7)) << 4), b0, b1, b2, b3);
        mma_bf16_16816(sacc[0], qa[ks][0], 0u, qa[ks][1], 0u, b0, b1);
        mma_bf16_16816(sacc[1], qa[ks][0], 0u, qa[ks][1], 0u, b2, b3);
      }
    }
    // ---- online softmax on row g (tokens wtok + n*8 + tq*2 + {0,1})
    float mx = -INFINITY;
#pragma unroll
    for (int n = 0; n < 2; n++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int tok = wtok + n * 8 + tq * 2 + e;
        const float v = tok < tile_len ? sacc[n][e] * scale_log2 : -INFINITY;
        sacc[n][e] = v;
        mx = fmaxf(mx, v);
      }
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    const float m_new = fmaxf(m_run, mx);  // finite: token wtok is valid
    const float corr = exp2f(m_run - m_new);  // 0 on the first tile
    m_run = m_new;
    float psum = 0.f;
    uint32_t pa[2];
#pragma unroll
    for (int n = 0; n < 2; n++) {
      const float p0 = exp2f(sacc[n][0] - m_new), p1 = exp2f(sacc[n][1] - m_new);
      psum += p0 + p1;
      pa[n] = float2_to_bf16x2(p0, p1);
    }
    l_run = l_run * corr + psum;
#pragma unroll
    for (int n = 0; n < NT; n++) { oacc[n][0] *= corr; oacc[n][1] *= corr; }
    // ---- O += P V : A = P (rows g; rows g+8 are padding -> zero), B = V^T fragments via ldmatrix.trans
    {
      const int mi = lane >> 3, r = lane & 7;
      const int tok = wtok + (mi & 1) * 8 + r;  // matrices 0,2 -> tokens 0-7; 1,3 -> tokens 8-15
#pragma unroll
      for (int n = 0; n < NT; n += 2) {
        const int c = n + (mi >> 1);
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4_trans(vs_addr + (c >> 3) * Cfg::kHalfBytes + tok * 128 + (((c & 7) ^ (tok & 7)) << 4), b0, b1, b2, b3);
        mma_bf16_16816(oacc[n], pa[0], 0u, pa[1], 0u, b0, b1);
        mma_bf16_16816(oacc[n + 1], pa[0], 0u, pa[1], 0u, b2, b3);
      }
    }
  }
  __syncthreads();  // all TMA loads issued were waited for above: the stage memory is free

  // ---- merge the four warps' (m, l, O) through the stage memory
  float *s_o = reinterpret_cast<float *>(smem);        // [4 warps][G][D]
  float *s_m = s_o + 4 * G * D;                        // [4][G]
  float *s_l = s_m + 4 * G;                            // [4][G]
  l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
  l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
  if (g < G) {
#pragma unroll
    for (int n = 0; n < NT; n++)
      *reinterpret_cast<float2 *>(s_o + ((size_t)(warp * G + g) * D + n * 8 + tq * 2)) = make_float2(oacc[n][0], oacc[n][1]);
    if (tq == 0) { s_m[warp * G + g] = m_run; s_l[warp * G + g] = l_run; }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < G * D; e += kAttnThreads) {
    const int hh = e / D, dd = e - hh * D;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < 4; w++) M = fmaxf(M, s_m[w * G + hh]);
    float L = 0.f, o = 0.f;
#pragma unroll
    for (int w = 0; w < 4; w++) {
      const float mw = s_m[w * G + hh];
      const float wgt = mw == -INFINITY ? 0.f : exp2f(mw - M);
      L += s_l[w * G + hh] * wgt;
      o += s_o[(size_t)(w * G + hh) * D + dd] * wgt;
    }
    const int h = h0 + hh;
    if (n_splits == 1) {
      out[((size_t)t * n_heads + h) * D + dd] = float_to_bf16_bits(o / L);
    } else {
      // scratch: ml [T][n_heads][S][2] then o [T][n_heads][S][D]; m is in the log2 domain
      float *ml = scratch + (((size_t)t * n_heads + h) * n_splits + split) * 2;
      float *op = scratch + (size_t)T * n_heads * n_splits * 2 + (((size_t)t * n_heads + h) * n_splits + split) * D;
      op[dd] = o;
      if (dd == 0) { ml[0] = M; ml[1] = L; }
    }
  }
}

// ---------------------------------------------------------------------------------------- prefill
// Attention of a BLOCK of up to 16 consecutive query tokens of one sequence.
// One CTA = (query block, kv head); warp w owns query head h0 + w: a 16-row MMA tile whose rows are the
// block's tokens.  All warps walk the same 64-token K/V tiles (TMA ring shared by the GQA group, so K/V
// are read once per 16 tokens x G heads), causal mask by position, online softmax per row, P in bf16.
template <int D, int G>
__global__ void __launch_bounds__(G * 32)
attn_prefill_kernel(const __grid_constant__ CUtensorMap kv_map, long long row0, const uint16_t *__restrict__ q,
                    const int *__restrict__ page_table, int max_pages, const int *__restrict__ qb_tok0,
                    const int *__restrict__ qb_len, const int *__restrict__ qb_pos0, const int *__restrict__ qb_ptrow,
                    uint16_t *__restrict__ out, int n_heads, int n_kv, int page_shift) {
  using Cfg = AttnCfg<D>;
  constexpr int KS = D / 16, NT = D / 8;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + Cfg::kStages * Cfg::kStageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };

  const int qb = blockIdx.x, kvh = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, tq = lane & 3;
  const int h = kvh * G + warp;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&kv_map);
    for (int s = 0; s < Cfg::kStages; s++) mbar_init(full_bar(s), 1);
    fence_barrier_init();
    fence_proxy_async();
  }
  pdl_launch_dependents();
  pdl_wait();
  __syncthreads();

  const int tok0 = qb_tok0[qb], len = qb_len[qb], pos0 = qb_pos0[qb];
  const int *pt = page_table + (size_t)qb_ptrow[qb] * max_pages;
  const int page_size = 1 << page_shift;
  const int box_rows = min(page_size, kAttnTile);
  const int end = pos0 + len;  // keys [0, end)
  const int n_tiles = (end + kAttnTile - 1) / kAttnTile;

  auto load_tile = [&](int tile, int stage) {  // one thread
    const int tile0 = tile * kAttnTile;
    const uint32_t sbase = smem_base + stage * Cfg::kStageBytes;
    mbar_expect_tx(full_bar(stage), Cfg::kStageBytes);
    for (int r = 0; r < kAttnTile; r += box_rows) {
      const int pos = tile0 + r;
      const long long page = pt[min(pos >> page_shift, max_pages - 1)];
      const long long krow = row0 + ((page * 2 + 0) * n_kv + kvh) * page_size + (pos & (page_size - 1));
      const long long vrow = krow + (long long)n_kv * page_size;
#pragma unroll
      for (int hh = 0; hh < Cfg::kHalves; hh++) {
        tma_load_2d(sbase + hh * Cfg::kHalfBytes + r * 128, &kv_map, hh * 64, (int)krow, full_bar(stage), kEvictLast);
        tma_load_2d(sbase + (Cfg::kHalves + hh) * Cfg::kHalfBytes + r * 128, &kv_map, hh * 64, (int)vrow, full_bar(stage), kEvictLast);
      }
    }
  };
  if (threadIdx.x == 0)
    for (int s = 0; s < Cfg::kStages - 1; s++)
      if (s < n_tiles) load_tile(s, s);

  // ---- Q fragments: rows g and g+8 are tokens tok0+g, tok0+g+8 of head h
  uint32_t qa[KS][4];
  {
    const bool v0 = g < len, v1 = g + 8 < len;
    const uint16_t *q0 = q + ((size_t)(tok0 + g) * n_heads + h) * D + tq * 2;
    const uint16_t *q1 = q + ((size_t)(tok0 + g + 8) * n_heads + h) * D + tq * 2;
#pragma unroll
    for (int ks = 0; ks < KS; ks++) {
      qa[ks][0] = v0 ? *reinterpret_cast<const uint32_t *>(q0 + ks * 16) : 0u;
      qa[ks][1] = v1 ? *reinterpret_cast<const uint32_t *>(q1 + ks * 16) : 0u;
      qa[ks][2] = v0 ? *reinterpret_cast<const uint32_t *>(q0 + ks * 16 + 8) : 0u;
      qa[ks][3] = v1 ? *reinterpret_cast<const uint32_t *>(q1 + ks * 16 + 8) : 0u;
    }
  }
  const float scale_log2 = rsqrtf((float)D) * 1.4426950408889634f;
  float oacc[NT][4];
#pragma unroll
  for (int n = 0; n < NT; n++) { oacc[n][0] = 0.f; oacc[n][1] = 0.f; oacc[n][2] = 0.f; oacc[n][3] = 0.f; }
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  const int qp0 = pos0 + g, qp1 = pos0 + g + 8;  // query positions of this thread's two rows

  for (int tile = 0; tile < n_tiles; tile++) {
    __syncthreads();
    if (threadIdx.x == 0) {
      const int nt = tile + Cfg::kStages - 1;
      if (nt < n_tiles) load_tile(nt, nt % Cfg::kStages);
    }
    const int stage = tile % Cfg::kStages;
    mbar_wait(full_bar(stage), (uint32_t)(tile / Cfg::kStages) & 1u);
    const uint32_t ks_addr = smem_base + stage * Cfg::kStageBytes;
    const uint32_t vs_addr = ks_addr + Cfg::kHalves * Cfg::kHalfBytes;
    const int tile0 = tile * kAttnTile;

    // ---- S = Q K^T : 16 rows x 64 keys
    float sacc[8][4];
#pragma unroll
    for (int n = 0; n < 8; n++) { sacc[n][0] = 0.f; sacc[n][1] = 0.f; sacc[n][2] = 0.f; sacc[n][3] = 0.f; }
    {
      const int mi = lane >> 3, r = lane & 7;
#pragma unroll
      for (int ks = 0; ks < KS; ks++) {
        const int c = ks * 2 + (mi & 1);
#pragma unroll
        for (int np = 0; np < 4; np++) {  // n-tile pairs
          const int tok = np * 16 + (mi >> 1) * 8 + r;
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4(ks_addr + (c >> 3) * Cfg::kHalfBytes + tok * 128 + (((c & 7) ^ (tok & 7)) << 4), b0, b1, b2, b3);
          mma_bf16_16816(sacc[2 * np], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], b0, b1);
          mma_bf16_16816(sacc[2 * np + 1], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], b2, b3);
        }
      }
    }
    // ---- causal mask + online softmax (rows g and g+8)
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int n = 0; n < 8; n++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int kp = tile0 + n * 8 + tq * 2 + e;
        const float v0 = kp <= qp0 ? sacc[n][e] * scale_log2 : -INFINITY;
        const float v1 = kp <= qp1 ? sacc[n][2 + e] * scale_log2 : -INFINITY;
        sacc[n][e] = v0; sacc[n][2 + e] = v1;
        mx0 = fmaxf(mx0, v0); mx1 = fmaxf(mx1, v1);
      }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m_run[0], mx0), mn1 = fmaxf(m_run[1], mx1);
    // a row with nothing visible yet keeps m = -inf: use 0 as the reference so exp2(-inf - 0) = 0
    const float r0 = mn0 == -INFINITY ? 0.f : mn0, r1 = mn1 == -INFINITY ? 0.f : mn1;
    const float c0 = exp2f(m_run[0] - r0), c1 = exp2f(m_run[1] - r1);
    m_run[0] = mn0; m_run[1] = mn1;
    float ps0 = 0.f, ps1 = 0.f;
    uint32_t pa[4][4];
#pragma unroll
    for (int kk = 0; kk < 4; kk++) {
      const float a0 = exp2f(sacc[2 * kk][0] - r0), a1 = exp2f(sacc[2 * kk][1] - r0);
      const float b0 = exp2f(sacc[2 * kk][2] - r1), b1 = exp2f(sacc[2 * kk][3] - r1);
      const float c2 = exp2f(sacc[2 * kk + 1][0] - r0), c3 = exp2f(sacc[2 * kk + 1][1] - r0);
      const float d0 = exp2f(sacc[2 * kk + 1][2] - r1), d1 = exp2f(sacc[2 * kk + 1][3] - r1);
      ps0 += a0 + a1 + c2 + c3;
      ps1 += b0 + b1 + d0 + d1;
      pa[kk][0] = float2_to_bf16x2(a0, a1); pa[kk][1] = float2_to_bf16x2(b0, b1);
      pa[kk][2] = float2_to_bf16x2(c2, c3); pa[kk][3] = float2_to_bf16x2(d0, d1);
    }
    l_run[0] = l_run[0] * c0 + ps0;
    l_run[1] = l_run[1] * c1 + ps1;
#pragma unroll
    for (int n = 0; n < NT; n++) { oacc[n][0] *= c0; oacc[n][1] *= c0; oacc[n][2] *= c1; oacc[n][3] *= c1; }
    // ---- O += P V
    {
      const int mi = lane >> 3, r = lane & 7;
#pragma unroll
      for (int kk = 0; kk < 4; kk++) {
        const int tok = kk * 16 + (mi & 1) * 8 + r;
#pragma unroll
        for (int n = 0; n < NT; n += 2) {
          const int c = n + (mi >> 1);
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4_trans(vs_addr + (c >> 3) * Cfg::kHalfBytes + tok * 128 + (((c & 7) ^ (tok & 7)) << 4), b0, b1, b2, b3);
          mma_bf16_16816(oacc[n], pa[kk][0], pa[kk][1], pa[kk][2], pa[kk][3], b0, b1);
          mma_bf16_16816(oacc[n + 1], pa[kk][0], pa[kk][1], pa[kk][2], pa[kk][3], b2, b3);
        }
      }
    }
  }
  // ---- normalise and store rows g, g+8
  l_run[0] += __shfl_xor_sync(0xffffffffu, l_run[0], 1); l_run[0] += __shfl_xor_sync(0xffffffffu, l_run[0], 2);
  l_run[1] += __shfl_xor_sync(0xffffffffu, l_run[1], 1); l_run[1] += __shfl_xor_sync(0xffffffffu, l_run[1], 2);
  const float i0 = 1.0f / l_run[0], i1 = 1.0f / l_run[1];
  if (g < len) {
    uint16_t *o = out + ((size_t)(tok0 + g) * n_heads + h) * D + tq * 2;
#pragma unroll
    for (int n = 0; n < NT; n++) *reinterpret_cast<uint32_t *>(o + n * 8) = float2_to_bf16x2(oacc[n][0] * i0, oacc[n][1] * i0);
  }
  if (g + 8 < len) {
    uint16_t *o = out + ((size_t)(tok0 + g + 8) * n_heads + h) * D + tq * 2;
#pragma unroll
    for (int n = 0; n < NT; n++) *reinterpret_cast<uint32_t *>(o + n * 8) = float2_to_bf16x2(oacc[n][2] * i1, oacc[n][3] * i1);
  }
}

template <int D>
__global__ void attn_combine_kernel(const float *__restrict__ scratch, uint16_t *__restrict__ out, int n_splits,
                                    int T, int n_heads) {
  const int t = blockIdx.y, h = blockIdx.x, dd = threadIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  const float *ml = scratch + ((size_t)t * n_heads + h) * n_splits * 2;
  const float *op = scratch + (size_t)T * n_heads * n_splits * 2 + ((size_t)t * n_heads + h) * n_splits * D;
  float M = -INFINITY;
  for (int s = 0; s < n_splits; s++) M = fmaxf(M, ml[2 * s]);
  float L = 0.f, o = 0.f;
  for (int s = 0; s < n_splits; s++) {
    const float w = ml[2 * s] == -INFINITY ? 0.f : exp2f(ml[2 * s] - M);
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
static cudaError_t attn_launch(const TmaMap &kv_map, long long row0, const AttnRope &rope, const uint16_t *q, const int *page_table,
                               int max_pages, const int *pt_row, const int *ctx, uint16_t *out, float *scratch,
                               int n_splits, int T, int n_heads, int n_kv, int page_shift, cudaStream_t st) {
  dim3 grid(n_splits, n_kv, T);
  static DeviceOnce once;
  DIS_CUDA_TRY(once.run([&] { return cudaFuncSetAttribute(attn_paged_kernel<D, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnCfg<D>::kSmemBytes); }));
  // ring depth: fewer stages leave shared memory for the NEXT kernel's CTAs to become resident (and prefetch their
  // weights) while this one is still streaming; DIS_ATTN_STAGES overrides (A/B runs), 2 <= stages <= AttnCfg::kStages
  static int stages = 0;
  if (stages == 0) {
    const char *e = getenv("DIS_ATTN_STAGES");
    stages = e ? atoi(e) : AttnCfg<D>::kStages;
    stages = stages < 2 ? 2 : (stages > AttnCfg<D>::kStages ? AttnCfg<D>::kStages : stages);
  }
  const int smem = stages * AttnCfg<D>::kStageBytes + 1024 + 64;
  DIS_CUDA_TRY(launch_k(attn_paged_kernel<D, G>, grid, dim3(kAttnThreads), (size_t)smem, st,
                        *reinterpret_cast<const CUtensorMap *>(&kv_map), row0, rope, q, page_table, max_pages, pt_row, ctx, out,
                        scratch, n_splits, T, n_heads, n_kv, page_shift, stages));
  if (n_splits > 1)
    DIS_CUDA_TRY(launch_k(attn_combine_kernel<D>, dim3(n_heads, T), dim3(D), 0, st, (const float *)scratch, out, n_splits, T, n_heads));
  return cudaSuccess;
}

// The pool as a 2-D tensor of d-element rows, boxes of {64 dims, min(page_size, 64) rows}, 128B swizzle.
cudaError_t make_kv_map(TmaMap *out, const uint16_t *pool, long long total_rows, int d, int page_size) {
  return make_tma_map_2d(out, pool, (uint64_t)total_rows, (uint64_t)d, (uint32_t)(page_size < kAttnTile ? page_size : kAttnTile));
}

static cudaError_t attn_dispatch(const TmaMap &kv_map, long long row0, const AttnRope &rope, const uint16_t *q,
                                 const int *page_table, int max_pages, const int *pt_row, const int *ctx, uint16_t *out,
                                 float *scratch, int n_splits, int T, int n_heads, int n_kv, int d, int page_size,
                                 cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  if (n_kv <= 0 || n_heads % n_kv != 0) { set_last_error("attn_paged: n_heads must be a multiple of n_kv"); return cudaErrorInvalidValue; }
  if (n_splits < 1 || (n_splits > 1 && !scratch)) { set_last_error("attn_paged: n_splits > 1 needs scratch"); return cudaErrorInvalidValue; }
  if (T > 65535 || n_kv > 65535) { set_last_error("attn_paged: more than 65535 tokens or kv heads per launch"); return cudaErrorInvalidValue; }
  if (page_size < 8 || (page_size & (page_size - 1))) { set_last_error("attn_paged: page_size must be a power of two >= 8"); return cudaErrorInvalidValue; }
  int page_shift = 0;
  while ((1 << page_shift) < page_size) page_shift++;
  const int G = n_heads / n_kv;  // query heads per kv head, all handled by one CTA
#define DIS_ATTN_CASE(DD, GG)                                                                                        \
  if (d == DD && G == GG)                                                                                            \
    return attn_launch<DD, GG>(kv_map, row0, rope, q, page_table, max_pages, pt_row, ctx, out, scratch, n_splits, T, \
                               n_heads, n_kv, page_shift, st);
  DIS_ATTN_CASE(128, 8) DIS_ATTN_CASE(128, 4) DIS_ATTN_CASE(128, 2) DIS_ATTN_CASE(128, 1)
  DIS_ATTN_CASE(64, 8) DIS_ATTN_CASE(64, 4) DIS_ATTN_CASE(64, 2) DIS_ATTN_CASE(64, 1)
#undef DIS_ATTN_CASE
  set_last_error("attn_paged: head_dim must be 64 or 128 and the GQA group 1, 2, 4 or 8");
  return cudaErrorInvalidValue;
}

cudaError_t launch_attn_paged(const TmaMap &kv_map, long long row0, const uint16_t *q, const int *page_table,
                              int max_pages, const int *pt_row, const int *ctx, uint16_t *out, float *scratch,
                              int n_splits, int T, int n_heads, int n_kv, int d, int page_size, cudaStream_t st) {
  AttnRope none{};
  return attn_dispatch(kv_map, row0, none, q, page_table, max_pages, pt_row, ctx, out, scratch, n_splits, T, n_heads, n_kv, d,
                       page_size, st);
}

// Decode step: RoPE + KV write of the step's new token (position ctx[t] - 1 == pos[t]) fused in.
cudaError_t launch_attn_decode_fused(const TmaMap &kv_map, long long row0, uint16_t *kv_layer_base, const float *qkv,
                                     int n_partial, long long partial_stride, const int *pos, const float *rope_cos,
                                     const float *rope_sin, const int *page_table, int max_pages, const int *pt_row,
                                     const int *ctx, uint16_t *out, float *scratch, int n_splits, int T, int n_heads,
                                     int n_kv, int d, int page_size, cudaStream_t st) {
  AttnRope r{qkv, n_partial, partial_stride, pos, rope_cos, rope_sin, kv_layer_base};
  return attn_dispatch(kv_map, row0, r, nullptr, page_table, max_pages, pt_row, ctx, out, scratch, n_splits, T, n_heads, n_kv, d,
                       page_size, st);
}

template <int D, int G>
static cudaError_t prefill_launch(const TmaMap &kv_map, long long row0, const uint16_t *q, const int *page_table,
                                  int max_pages, const int *qb_tok0, const int *qb_len, const int *qb_pos0,
                                  const int *qb_ptrow, uint16_t *out, int n_blocks, int n_heads, int n_kv,
                                  int page_shift, cudaStream_t st) {
  static DeviceOnce once;
  DIS_CUDA_TRY(once.run([&] { return cudaFuncSetAttribute(attn_prefill_kernel<D, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnCfg<D>::kSmemBytes); }));
  return launch_k(attn_prefill_kernel<D, G>, dim3(n_blocks, n_kv), dim3(G * 32), AttnCfg<D>::kSmemBytes, st,
                  *reinterpret_cast<const CUtensorMap *>(&kv_map), row0, q, page_table, max_pages, qb_tok0, qb_len, qb_pos0,
                  qb_ptrow, out, n_heads, n_kv, page_shift);
}

// Query blocks: block b covers tokens [qb_tok0[b], +qb_len[b]) of the flattened batch (qb_len <= 16), which are
// consecutive positions qb_pos0[b].. of the sequence whose page-table row is qb_ptrow[b].
cudaError_t launch_attn_prefill(const TmaMap &kv_map, long long row0, const uint16_t *q, const int *page_table,
                                int max_pages, const int *qb_tok0, const int *qb_len, const int *qb_pos0,
                                const int *qb_ptrow, uint16_t *out, int n_blocks, int n_heads, int n_kv, int d,
                                int page_size, cudaStream_t st) {
  if (n_blocks <= 0) return cudaSuccess;
  if (n_kv <= 0 || n_heads % n_kv != 0 || n_kv > 65535) { set_last_error("attn_prefill: bad head counts"); return cudaErrorInvalidValue; }
  if (page_size < 8 || (page_size & (page_size - 1))) { set_last_error("attn_prefill: page_size must be a power of two >= 8"); return cudaErrorInvalidValue; }
  int page_shift = 0;
  while ((1 << page_shift) < page_size) page_shift++;
  const int G = n_heads / n_kv;
#define DIS_PF_CASE(DD, GG)                                                                                        \
  if (d == DD && G == GG)                                                                                          \
    return prefill_launch<DD, GG>(kv_map, row0, q, page_table, max_pages, qb_tok0, qb_len, qb_pos0, qb_ptrow, out, \
                                  n_blocks, n_heads, n_kv, page_shift, st);
  DIS_PF_CASE(128, 8) DIS_PF_CASE(128, 4) DIS_PF_CASE(128, 2) DIS_PF_CASE(128, 1)
  DIS_PF_CASE(64, 8) DIS_PF_CASE(64, 4) DIS_PF_CASE(64, 2) DIS_PF_CASE(64, 1)
#undef DIS_PF_CASE
  set_last_error("attn_prefill: head_dim must be 64 or 128 and the GQA group 1, 2, 4 or 8");
  return cudaErrorInvalidValue;
}

// ---------------------------------------------------------------------------------------------------
// dis_k_attn_paged_static: the split-KV attention above for hosts that must not allocate.
//
// Who needs it: a serving process that has already handed every free byte of HBM to its KV-page pool (this
// engine does exactly that with kv_mem_fraction) has nothing left for cudaMalloc to give, and a request
// loop that must stay free of cudaMalloc/cudaFree -- both synchronise the whole device and stall every
// other stream -- cannot afford the call even when memory is available.  Such a host still wants the
// split-KV path (it is what keeps all SMs busy at small batch), which needs room for the per-split
// (m, l, o) partials.  Here that room is a fixed array inside the library image instead of a buffer the
// caller provides.
namespace {
constexpr size_t kStaticScratchFloats = DIS_ATTN_STATIC_SCRATCH_FLOATS;
// Lives in every CUDA context that loads this library, from module load until the process exits.
static __device__ __align__(256) float g_attn_static_scratch[kStaticScratchFloats];

// In host code the array's name is only a handle for the runtime, not a device address: the address each
// device gave its own copy has to be asked for, once per device, while that device is current.
cudaError_t static_scratch_of_current_device(float **out) {
  static std::mutex mu;
  static float *ptrs[64] = {};
  int dev = -1;
  DIS_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { set_last_error("attn_paged_static: device ordinal out of range"); return cudaErrorInvalidDevice; }
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[dev]) {
    float *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_attn_static_scratch);
    if (e != cudaSuccess || !p) {
      set_last_error(std::string("attn_paged_static: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
      return e != cudaSuccess ? e : cudaErrorUnknown;
    }
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return cudaSuccess;
}
}  // namespace

}  // namespace dis

extern "C" DIS_API int dis_k_attn_paged_static(const uint16_t *q, const uint16_t *kv_pool, int64_t n_pages,
                                               const int32_t *page_table, int32_t max_pages, const int32_t *pt_row,
                                               const int32_t *ctx, uint16_t *out, int32_t n_splits, int32_t T,
                                               int32_t n_heads, int32_t n_kv, int32_t d, int32_t page_size, void *stream) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    dis::set_last_error("no CUDA device: there is no CPU fallback");
    return DIS_ERR_INFERENCE_FAILED;
  }
  if (T <= 0 || n_splits < 1) { dis::set_last_error("attn_paged_static: T and n_splits must be positive"); return DIS_ERR_INVALID_ARGUMENT; }
  const size_t need = dis::attn_scratch_elems(T, n_heads, d, n_splits);
  if (need > dis::kStaticScratchFloats) {
    dis::set_last_error("attn_paged_static: T*n_heads*n_splits*(2+d) = " + std::to_string(need) +
                        " floats of split-KV partials do not fit the built-in scratch of " +
                        std::to_string(dis::kStaticScratchFloats) + " floats (DIS_ATTN_STATIC_SCRATCH_FLOATS); use fewer "
                        "splits or tokens per call, or dis_k_attn_paged with your own buffer");
    return DIS_ERR_INVALID_ARGUMENT;
  }
  float *scratch = nullptr;
  if (dis::static_scratch_of_current_device(&scratch) != cudaSuccess) return DIS_ERR_INFERENCE_FAILED;
  dis::TmaMap map;  // encoded on the host stack; no device allocation
  if (dis::make_kv_map(&map, kv_pool, (long long)n_pages * 2 * n_kv * page_size, d, page_size) != cudaSuccess)
    return DIS_ERR_INFERENCE_FAILED;
  // Nothing in the scratch has to start at zero: every (m, l, o) slot the combine pass reads is written by a
  // split CTA of this same call first, so whatever the previous call left behind is never observed.
  const cudaError_t e = dis::launch_attn_paged(map, 0, q, page_table, max_pages, pt_row, ctx, out, n_splits > 1 ? scratch : nullptr,
                                               n_splits, T, n_heads, n_kv, d, page_size, static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? DIS_OK : DIS_ERR_INFERENCE_FAILED;
}
