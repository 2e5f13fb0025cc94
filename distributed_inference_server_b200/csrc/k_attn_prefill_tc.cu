// k_attn_prefill_tc.cu -- causal prefill attention on the 5th-generation tensor cores:
// tcgen05.mma with S = Q K^T and the per-tile P V product accumulated in TMEM, K/V tiles staged by TMA.
// No reference counterpart (the reference has no tensor code at all).
//
// One CTA = (query block, kv head).  A query block is up to QT = 128 / G consecutive tokens of one sequence
// (G = query heads per kv head: 32 tokens for Llama-3-8B's 4:1 GQA, 16 for 70B's 8:1); its G heads x QT tokens
// are the 128 rows (= 128 TMEM lanes) of every MMA, row = head * QT + token, so K and V are fetched once per
// 128 query rows instead of once per 64 as in the mma.sync kernel (k_attn_paged.cu), and the QK^T / PV
// products run at tcgen05 rate instead of through ldmatrix + mma.sync.
//
//   warp 0      TMA producer: 64-key K/V tiles (128B-swizzled boxes out of the page pool) into a 3-stage ring
//   warp 1      one thread issues the MMAs:   S_j   = Q   K_j^T   M128 N64  K16 x 8   (A, B K-major)       -> TMEM S[j&1]
//                                             O_j   = P_j V_j     M128 N128 K16 x 4   (B = V, MN-major)    -> TMEM O[j&1]
//   warps 2-5   one thread per row: S_j from TMEM, scale, causal mask, running max / sum (no shuffles: a thread
//               owns a whole row), P_j = exp2(s - m) as bf16 into shared memory (the A operand of the PV MMA),
//               then O_{j-1} from TMEM into the fp32 accumulator it keeps in registers: acc = acc * corr + O_{j-1}
// S, P and O are double-buffered, so the tensor pipe computes S_{j+1} while the softmax of tile j runs and
// P_j V_j while the row threads fold O_{j-1}.  Tensor-bound: 4 * rows * keys * 128 flop per (block, kv head).
#include <cuda.h>

#include <cstdlib>
#include <cstring>

#include "kernels.h"
#include "ptx.cuh"

namespace dis {

constexpr int kPfThreads = 192;
constexpr int kPfTile = 64;    // keys per stage
constexpr int kPfStages = 3;
constexpr int kPfD = 128;
constexpr int kPfHalf = kPfTile * 128;                  // one [64 keys][64 dims] block, 8 KB
constexpr int kPfStageBytes = 4 * kPfHalf;               // K halves + V halves, 32 KB
constexpr int kPfQBytes = 2 * 128 * 128;                 // Q: two [128 rows][64 dims] halves, 32 KB
constexpr int kPfPBytes = 128 * 128;                     // P: [128 rows][64 keys] bf16, 16 KB
constexpr int kPfSmemBytes = kPfQBytes + kPfStages * kPfStageBytes + 2 * kPfPBytes + 256 + 1024;
constexpr int kPfTmemCols = 512;                         // S0 | S1 (64 each) | O0 | O1 (128 each)

// shared-memory matrix descriptor, MN-major operand, 128B swizzle: element (mn, k) lives at
//   (mn % 64) * 2 + (mn / 64) * LBO + (k % 8) * 128 + (k / 8) * SBO   -- the [keys][dims] V tile as TMA stores it
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
         (1ull << 46) | (2ull << 61);
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int G>
__global__ void __launch_bounds__(kPfThreads, 1)
attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap kv_map, long long row0, const uint16_t *__restrict__ q,
                       const int *__restrict__ page_table, int max_pages, const int *__restrict__ qb_tok0,
                       const int *__restrict__ qb_len, const int *__restrict__ qb_pos0, const int *__restrict__ qb_ptrow,
                       uint16_t *__restrict__ out, int n_heads, int n_kv, int page_shift) {
  constexpr int QT = 128 / G;  // tokens per query block
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *smem = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t q_base = smem_base;
  const uint32_t kv_base = q_base + kPfQBytes;
  const uint32_t p_base = kv_base + kPfStages * kPfStageBytes;
  const uint32_t bar_base = p_base + 2 * kPfPBytes;
  auto kv_full = [&](int s) { return bar_base + 8u * s; };
  auto kv_empty = [&](int s) { return bar_base + 8u * (kPfStages + s); };
  auto s_full = [&](int b) { return bar_base + 8u * (2 * kPfStages + b); };
  auto p_full = [&](int b) { return bar_base + 8u * (2 * kPfStages + 2 + b); };
  auto o_full = [&](int b) { return bar_base + 8u * (2 * kPfStages + 4 + b); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kPfStages + 6);

  const int qb = blockIdx.x, kvh = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&kv_map);
    for (int s = 0; s < kPfStages; s++) { mbar_init(kv_full(s), 1); mbar_init(kv_empty(s), 1); }
    for (int b = 0; b < 2; b++) { mbar_init(s_full(b), 1); mbar_init(p_full(b), 128); mbar_init(o_full(b), 1); }
    fence_barrier_init();
    fence_proxy_async();
  }
  if (warp == 1) tmem_alloc<kPfTmemCols>(tmem_slot);
  pdl_launch_dependents();
  pdl_wait();  // q and this chunk's K/V rows come from the previous kernels
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int tok0 = qb_tok0[qb], len = qb_len[qb], pos0 = qb_pos0[qb];
  const int *pt = page_table + (size_t)qb_ptrow[qb] * max_pages;
  const int page_size = 1 << page_shift;
  const int box_rows = min(page_size, kPfTile);
  const int end = pos0 + len;  // keys [0, end)
  const int n_tiles = (end + kPfTile - 1) / kPfTile;

  // ---- Q tile: row = head * QT + token, K-major, 128B swizzle (rows of padding tokens are zero)
  if (warp >= 2) {
    const int row = (warp & 3) * 32 + lane;
    const int hl = row / QT, t = row % QT;
    const uint4 *src = reinterpret_cast<const uint4 *>(q + ((size_t)(tok0 + t) * n_heads + kvh * G + hl) * kPfD);
#pragma unroll
    for (int c = 0; c < 16; c++) {
      const uint4 v = t < len ? src[c] : make_uint4(0u, 0u, 0u, 0u);
      const uint32_t off = (uint32_t)(c >> 3) * (128 * 128) + (uint32_t)(row >> 3) * 1024 + (uint32_t)(row & 7) * 128 +
                           (uint32_t)(((c & 7) ^ (row & 7)) << 4);
      *reinterpret_cast<uint4 *>(smem + off) = v;
    }
    fence_proxy_async();
  }
  __syncthreads();

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      for (int tile = 0; tile < n_tiles; tile++) {
        const int stage = tile % kPfStages;
        if (tile >= kPfStages) mbar_wait(kv_empty(stage), (uint32_t)(tile / kPfStages - 1) & 1u);
        const uint32_t sbase = kv_base + stage * kPfStageBytes;
        mbar_expect_tx(kv_full(stage), kPfStageBytes);
        for (int r = 0; r < kPfTile; r += box_rows) {
          const int pos = tile * kPfTile + r;
          const long long page = pt[min(pos >> page_shift, max_pages - 1)];
          const long long krow = row0 + ((page * 2 + 0) * n_kv + kvh) * page_size + (pos & (page_size - 1));
          const long long vrow = krow + (long long)n_kv * page_size;
#pragma unroll
          for (int hh = 0; hh < 2; hh++) {
            tma_load_2d(sbase + hh * kPfHalf + r * 128, &kv_map, hh * 64, (int)krow, kv_full(stage), kEvictLast);
            tma_load_2d(sbase + (2 + hh) * kPfHalf + r * 128, &kv_map, hh * 64, (int)vrow, kv_full(stage), kEvictLast);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc<kPfTile>();              // M128 N64, A and B K-major
      constexpr uint32_t idesc_o = umma_idesc<kPfD>() | (1u << 16);    // M128 N128, B (= V) MN-major
      auto issue_s = [&](int j) {
        const int stage = j % kPfStages;
        mbar_wait(kv_full(stage), (uint32_t)(j / kPfStages) & 1u);
        tc_fence_after();
        const uint32_t kb = kv_base + stage * kPfStageBytes;
#pragma unroll
        for (int k = 0; k < kPfD / 16; k++) {
          const uint32_t a = q_base + (k >> 2) * (128 * 128) + (k & 3) * 32;
          const uint32_t b = kb + (k >> 2) * kPfHalf + (k & 3) * 32;
          umma_bf16(tmem_base + (uint32_t)(j & 1) * kPfTile, umma_desc_sw128(a), umma_desc_sw128(b), idesc_s, k > 0 ? 1u : 0u);
        }
        umma_commit(s_full(j & 1));
      };
      auto issue_o = [&](int j) {
        const int stage = j % kPfStages;
        mbar_wait(p_full(j & 1), (uint32_t)(j >> 1) & 1u);
        tc_fence_after();
        const uint32_t vb = kv_base + stage * kPfStageBytes + 2 * kPfHalf;
        const uint32_t pb = p_base + (j & 1) * kPfPBytes;
#pragma unroll
        for (int k = 0; k < kPfTile / 16; k++)
          umma_bf16(tmem_base + 2 * kPfTile + (uint32_t)(j & 1) * kPfD, umma_desc_sw128(pb + k * 32),
                    umma_desc_mn_sw128(vb + k * 2048, kPfHalf, 1024), idesc_o, k > 0 ? 1u : 0u);
        umma_commit(o_full(j & 1));
        umma_commit(kv_empty(stage));  // K_j was read by S_j, V_j by this product: the stage may be refilled
      };
      issue_s(0);
      for (int j = 0; j < n_tiles; j++) {
        if (j + 1 < n_tiles) issue_s(j + 1);
        issue_o(j);
      }
    }
  } else {
    // ===== softmax + O accumulation: one thread per row =====
    const int quarter = warp & 3;  // a warp may only touch TMEM lanes [32 * (warp % 4), +32)
    const int row = quarter * 32 + lane;
    const int hl = row / QT, t = row % QT;
    const int qp = pos0 + t;  // this row's query position
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const float scale_log2 = rsqrtf((float)kPfD) * 1.4426950408889634f;
    float acc[kPfD];
#pragma unroll
    for (int i = 0; i < kPfD; i++) acc[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f, corr_prev = 1.f;
    auto fold_o = [&](int j, float corr) {  // acc = acc * corr + O_j
      mbar_wait(o_full(j & 1), (uint32_t)(j >> 1) & 1u);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < kPfD; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32(lane_addr + 2 * kPfTile + (uint32_t)(j & 1) * kPfD + c, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i++) acc[c + i] = fmaf(acc[c + i], corr, __uint_as_float(v[i]));
      }
    };
    for (int j = 0; j < n_tiles; j++) {
      mbar_wait(s_full(j & 1), (uint32_t)(j >> 1) & 1u);
      tc_fence_after();
      uint32_t s0[32], s1[32];
      tmem_ld_32x32(lane_addr + (uint32_t)(j & 1) * kPfTile, s0);
      tmem_ld_32x32(lane_addr + (uint32_t)(j & 1) * kPfTile + 32, s1);
      tmem_ld_wait();
      const int tile0 = j * kPfTile;
      // One warp per scheduler runs this loop, so its length in instructions IS the kernel's time (ncu: 1060 warp
      // instructions per 64-key tile before this diet, tensor pipe 21 % busy): the causal mask is applied only on the
      // tiles that cross the diagonal, the scale is folded into the exponent's FMA, ex2.approx is issued directly.
      float mx = -INFINITY;
      if (tile0 + kPfTile - 1 > qp) {
#pragma unroll
        for (int i = 0; i < 32; i++) {
          if (tile0 + i > qp) s0[i] = 0xff800000u;        // -inf
          if (tile0 + 32 + i > qp) s1[i] = 0xff800000u;
        }
      }
#pragma unroll
      for (int i = 0; i < 32; i++) mx = fmaxf(mx, fmaxf(__uint_as_float(s0[i]), __uint_as_float(s1[i])));
      const float m_new = fmaxf(m_run, mx * scale_log2);   // scale > 0: the max commutes with it; finite from tile 0 on
      const float corr = ex2_approx(m_run - m_new);        // 0 at the first tile (m_run = -inf)
      const float neg_m = -m_new;
      float ps0 = 0.f, ps1 = 0.f;
      uint8_t *prow = smem + (p_base - smem_base) + (j & 1) * kPfPBytes + (row >> 3) * 1024 + (row & 7) * 128;
#pragma unroll
      for (int c = 0; c < 8; c++) {  // 8 keys = one 16-byte chunk of the swizzled row
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
          const int i = c * 8 + e * 2;
          const float p0 = ex2_approx(fmaf(__uint_as_float(i < 32 ? s0[i] : s1[i - 32]), scale_log2, neg_m));
          const float p1 = ex2_approx(fmaf(__uint_as_float(i + 1 < 32 ? s0[i + 1] : s1[i + 1 - 32]), scale_log2, neg_m));
          ps0 += p0; ps1 += p1;
          pk[e] = float2_to_bf16x2(p0, p1);
        }
        *reinterpret_cast<uint4 *>(prow + ((c ^ (row & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      }
      const float ps = ps0 + ps1;
      l_run = l_run * corr + ps;
      m_run = m_new;
      fence_proxy_async();   // the PV MMA reads P through the async proxy
      tc_fence_before();     // ... and this thread's TMEM reads of S_j are done before the issuer reuses the buffer
      mbar_arrive(p_full(j & 1));
      if (j > 0) fold_o(j - 1, corr_prev);
      corr_prev = corr;
    }
    fold_o(n_tiles - 1, corr_prev);
    if (t < len) {
      const float inv = 1.0f / l_run;
      uint4 *o = reinterpret_cast<uint4 *>(out + ((size_t)(tok0 + t) * n_heads + kvh * G + hl) * kPfD);
#pragma unroll
      for (int c = 0; c < 16; c++) {
        uint4 v;
        v.x = float2_to_bf16x2(acc[c * 8 + 0] * inv, acc[c * 8 + 1] * inv);
        v.y = float2_to_bf16x2(acc[c * 8 + 2] * inv, acc[c * 8 + 3] * inv);
        v.z = float2_to_bf16x2(acc[c * 8 + 4] * inv, acc[c * 8 + 5] * inv);
        v.w = float2_to_bf16x2(acc[c * 8 + 6] * inv, acc[c * 8 + 7] * inv);
        o[c] = v;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<kPfTmemCols>(tmem_base);
}

// Tokens per query block of the prefill attention for this head geometry: 128 / G on the tcgen05 kernel
// (head_dim 128, GQA group 4 or 8), 16 on the mma.sync kernel.  DIS_ATTN_PREFILL=mma forces the latter (A/B runs).
static bool prefill_tc_enabled() {
  static int on = -1;
  if (on < 0) {
    const char *e = getenv("DIS_ATTN_PREFILL");
    on = (e && !strcmp(e, "mma")) ? 0 : 1;
  }
  return on == 1;
}

bool attn_prefill_uses_tc(int n_heads, int n_kv, int d) {
  if (n_kv <= 0 || n_heads % n_kv != 0) return false;
  const int G = n_heads / n_kv;
  return prefill_tc_enabled() && d == kPfD && (G == 4 || G == 8);
}

int attn_prefill_block_tokens(int n_heads, int n_kv, int d) {
  return attn_prefill_uses_tc(n_heads, n_kv, d) ? 128 / (n_heads / n_kv) : 16;
}

template <int G>
static cudaError_t prefill_tc_launch(const TmaMap &kv_map, long long row0, const uint16_t *q, const int *page_table,
                                     int max_pages, const int *qb_tok0, const int *qb_len, const int *qb_pos0,
                                     const int *qb_ptrow, uint16_t *out, int n_blocks, int n_heads, int n_kv,
                                     int page_shift, cudaStream_t st) {
  static DeviceOnce once;
  DIS_CUDA_TRY(once.run([&] { return cudaFuncSetAttribute(attn_prefill_tc_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPfSmemBytes); }));
  return launch_k(attn_prefill_tc_kernel<G>, dim3(n_blocks, n_kv), dim3(kPfThreads), (size_t)kPfSmemBytes, st,
                  *reinterpret_cast<const CUtensorMap *>(&kv_map), row0, q, page_table, max_pages, qb_tok0, qb_len, qb_pos0,
                  qb_ptrow, out, n_heads, n_kv, page_shift);
}

// Same contract as launch_attn_prefill (k_attn_paged.cu) with query blocks of up to attn_prefill_block_tokens() tokens.
cudaError_t launch_attn_prefill_tc(const TmaMap &kv_map, long long row0, const uint16_t *q, const int *page_table,
                                   int max_pages, const int *qb_tok0, const int *qb_len, const int *qb_pos0,
                                   const int *qb_ptrow, uint16_t *out, int n_blocks, int n_heads, int n_kv, int d,
                                   int page_size, cudaStream_t st) {
  if (n_blocks <= 0) return cudaSuccess;
  if (n_kv <= 0 || n_heads % n_kv != 0 || n_kv > 65535 || d != kPfD) { set_last_error("attn_prefill_tc: bad head geometry"); return cudaErrorInvalidValue; }
  if (page_size < 8 || (page_size & (page_size - 1))) { set_last_error("attn_prefill_tc: page_size must be a power of two >= 8"); return cudaErrorInvalidValue; }
  int page_shift = 0;
  while ((1 << page_shift) < page_size) page_shift++;
  const int G = n_heads / n_kv;
  if (G == 4) return prefill_tc_launch<4>(kv_map, row0, q, page_table, max_pages, qb_tok0, qb_len, qb_pos0, qb_ptrow, out, n_blocks, n_heads, n_kv, page_shift, st);
  if (G == 8) return prefill_tc_launch<8>(kv_map, row0, q, page_table, max_pages, qb_tok0, qb_len, qb_pos0, qb_ptrow, out, n_blocks, n_heads, n_kv, page_shift, st);
  set_last_error("attn_prefill_tc: the GQA group must be 4 or 8");
  return cudaErrorInvalidValue;
}

}  // namespace dis
