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

// ---------------------------------------------------------------------------------------- two CTAs per SM
// ncu on the kernel above (profiles/r2_attn_prefill_tc_full.txt): tensor pipe 21 % busy, XU 22 %, one warp per scheduler
// walking ~1000 dependent instructions per 64-key tile -- it is bound by the latency of its own instruction stream, and
// 237 registers + 512 TMEM columns + 165 KB of shared memory allow one CTA per SM.  This variant is built to fit TWO:
//   * O stays in TMEM and accumulates there over the tiles (tcgen05.mma accumulate), so the 128-register accumulator and
//     the 128 FMAs + TMEM reads of the per-tile fold are gone.  The softmax reference m_ref is only raised when the running
//     maximum exceeds it by more than 8 (log2 domain): P = exp2(s - m_ref) <= 256 stays exact enough in bf16/fp32, the
//     reference cancels in the final division by l, and the rescale of O (tcgen05.ld, scale, tcgen05.st, warp-uniformly
//     decided) happens only on the few tiles where a row's maximum still grows;
//   * TMEM: S0 | S1 | O = 256 columns; shared memory: Q 32 KB + two K and two V slots of 16 KB + one P tile = 112 KB;
//     K and V have their own barriers, so K_{j+1} can be fetched as soon as S_{j-1} has read its slot, a whole tile
//     before it is needed.
constexpr int kPf2KSlot = 2 * kPfHalf;                   // one K (or V) tile: two [64 keys][64 dims] halves, 16 KB
constexpr int kPf2SmemBytes = kPfQBytes + 4 * kPf2KSlot + kPfPBytes + 256;
constexpr int kPf2TmemCols = 256;
constexpr float kPf2Lazy = 8.0f;

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]),
        "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]),
        "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

template <int G>
__global__ void __launch_bounds__(kPfThreads, 2)
attn_prefill_tc2_kernel(const __grid_constant__ CUtensorMap kv_map, long long row0, const uint16_t *__restrict__ q,
                        const int *__restrict__ page_table, int max_pages, const int *__restrict__ qb_tok0,
                        const int *__restrict__ qb_len, const int *__restrict__ qb_pos0, const int *__restrict__ qb_ptrow,
                        uint16_t *__restrict__ out, int n_heads, int n_kv, int page_shift) {
  constexpr int QT = 128 / G;
  extern __shared__ __align__(1024) uint8_t smem_al[];
  const uint32_t smem_base = smem_u32(smem_al);
  if (smem_base & 1023u) __trap();  // the 128B-swizzle atoms need 1024-byte alignment and there is no room for slack
  uint8_t *smem = smem_al;
  const uint32_t q_base = smem_base;
  const uint32_t k_base = q_base + kPfQBytes;            // K slots 0, 1
  const uint32_t v_base = k_base + 2 * kPf2KSlot;        // V slots 0, 1
  const uint32_t p_base = v_base + 2 * kPf2KSlot;
  const uint32_t bar_base = p_base + kPfPBytes;
  auto k_full = [&](int s) { return bar_base + 8u * s; };
  auto k_empty = [&](int s) { return bar_base + 8u * (2 + s); };
  auto v_full = [&](int s) { return bar_base + 8u * (4 + s); };
  auto v_empty = [&](int s) { return bar_base + 8u * (6 + s); };
  auto s_full = [&](int b) { return bar_base + 8u * (8 + b); };
  const uint32_t p_full = bar_base + 8u * 10, o_full = bar_base + 8u * 11;
  const uint32_t tmem_slot = bar_base + 8u * 12;

  // later query blocks of a sequence have more key tiles: launch them first, so that the longest CTAs do not form the tail
  const int qb = (int)gridDim.x - 1 - (int)blockIdx.x, kvh = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&kv_map);
    for (int s = 0; s < 2; s++) {
      mbar_init(k_full(s), 1); mbar_init(k_empty(s), 1); mbar_init(v_full(s), 1); mbar_init(v_empty(s), 1); mbar_init(s_full(s), 1);
    }
    mbar_init(p_full, 128); mbar_init(o_full, 1);
    fence_barrier_init();
    fence_proxy_async();
  }
  if (warp == 1) tmem_alloc<kPf2TmemCols>(tmem_slot);
  pdl_launch_dependents();
  pdl_wait();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int tok0 = qb_tok0[qb], len = qb_len[qb], pos0 = qb_pos0[qb];
  const int *pt = page_table + (size_t)qb_ptrow[qb] * max_pages;
  const int page_size = 1 << page_shift;
  const int box_rows = min(page_size, kPfTile);
  const int end = pos0 + len;
  const int n_tiles = (end + kPfTile - 1) / kPfTile;

  if (warp >= 2) {  // Q tile: row = head * QT + token, K-major, 128B swizzle (rows of padding tokens are zero)
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
    // ===== TMA producer: K_t as soon as S_{t-2} has read its slot, V_{t-1} as soon as P_{t-3} V_{t-3} has =====
    if (lane == 0) {
      auto load = [&](int tile, bool is_v) {
        const int slot = tile & 1;
        const uint32_t full = is_v ? v_full(slot) : k_full(slot), empty = is_v ? v_empty(slot) : k_empty(slot);
        if (tile >= 2) mbar_wait(empty, (uint32_t)(tile / 2 - 1) & 1u);
        const uint32_t sbase = (is_v ? v_base : k_base) + slot * kPf2KSlot;
        mbar_expect_tx(full, kPf2KSlot);
        for (int r = 0; r < kPfTile; r += box_rows) {
          const int pos = tile * kPfTile + r;
          const long long page = pt[min(pos >> page_shift, max_pages - 1)];
          const long long krow = row0 + ((page * 2 + (is_v ? 1 : 0)) * n_kv + kvh) * page_size + (pos & (page_size - 1));
#pragma unroll
          for (int hh = 0; hh < 2; hh++) tma_load_2d(sbase + hh * kPfHalf + r * 128, &kv_map, hh * 64, (int)krow, full, kEvictLast);
        }
      };
      for (int t = 0; t <= n_tiles; t++) {
        if (t < n_tiles) load(t, false);
        if (t >= 1) load(t - 1, true);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc<kPfTile>();
      constexpr uint32_t idesc_o = umma_idesc<kPfD>() | (1u << 16);
      auto issue_s = [&](int j) {
        const int slot = j & 1;
        mbar_wait(k_full(slot), (uint32_t)(j >> 1) & 1u);
        tc_fence_after();
        const uint32_t kb = k_base + slot * kPf2KSlot;
#pragma unroll
        for (int k = 0; k < kPfD / 16; k++) {
          const uint32_t a = q_base + (k >> 2) * (128 * 128) + (k & 3) * 32;
          const uint32_t b = kb + (k >> 2) * kPfHalf + (k & 3) * 32;
          umma_bf16(tmem_base + (uint32_t)slot * kPfTile, umma_desc_sw128(a), umma_desc_sw128(b), idesc_s, k > 0 ? 1u : 0u);
        }
        umma_commit(s_full(slot));
        umma_commit(k_empty(slot));
      };
      auto issue_o = [&](int j) {
        const int slot = j & 1;
        mbar_wait(p_full, (uint32_t)j & 1u);
        mbar_wait(v_full(slot), (uint32_t)(j >> 1) & 1u);
        tc_fence_after();
        const uint32_t vb = v_base + slot * kPf2KSlot;
#pragma unroll
        for (int k = 0; k < kPfTile / 16; k++)
          umma_bf16(tmem_base + 2 * kPfTile, umma_desc_sw128(p_base + k * 32), umma_desc_mn_sw128(vb + k * 2048, kPfHalf, 1024),
                    idesc_o, (j > 0 || k > 0) ? 1u : 0u);
        umma_commit(o_full);
        umma_commit(v_empty(slot));
      };
      issue_s(0);
      for (int j = 0; j < n_tiles; j++) {
        if (j + 1 < n_tiles) issue_s(j + 1);
        issue_o(j);
      }
    }
  } else {
    // ===== softmax: one thread per row; O lives in TMEM =====
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int hl = row / QT, t = row % QT;
    const int qp = pos0 + t;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const uint32_t o_addr = lane_addr + 2 * kPfTile;
    const float scale_log2 = rsqrtf((float)kPfD) * 1.4426950408889634f;
    float m_ref = -INFINITY, l_run = 0.f;
    for (int j = 0; j < n_tiles; j++) {
      mbar_wait(s_full(j & 1), (uint32_t)(j >> 1) & 1u);
      tc_fence_after();
      uint32_t s0[32], s1[32];
      tmem_ld_32x32(lane_addr + (uint32_t)(j & 1) * kPfTile, s0);
      tmem_ld_32x32(lane_addr + (uint32_t)(j & 1) * kPfTile + 32, s1);
      tmem_ld_wait();
      const int tile0 = j * kPfTile;
      if (tile0 + kPfTile - 1 > qp) {  // only the tiles that cross the diagonal need the causal mask
#pragma unroll
        for (int i = 0; i < 32; i++) {
          if (tile0 + i > qp) s0[i] = 0xff800000u;
          if (tile0 + 32 + i > qp) s1[i] = 0xff800000u;
        }
      }
      // four independent chains each for the maximum and the row sum: two warps per scheduler do not hide the latency of
      // one 32-deep dependent chain (ncu: "wait" and "long scoreboard" are the top two stall reasons)
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < 32; i++) mx4[i & 3] = fmaxf(mx4[i & 3], fmaxf(__uint_as_float(s0[i]), __uint_as_float(s1[i])));
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      const float m_new = fmaxf(m_ref, mx * scale_log2);
      if (j == 0) {
        m_ref = m_new;  // finite: key 0 is visible to every query; O is first written by P_0 V_0 (accumulate off)
      } else {
        // P_{j-1} V_{j-1} must have finished before P's buffer is rewritten and before O may be rescaled
        mbar_wait(o_full, (uint32_t)(j - 1) & 1u);
        tc_fence_after();
        const bool need = m_new > m_ref + kPf2Lazy;
        if (__any_sync(0xffffffffu, need)) {  // tcgen05.ld / st are warp-collective: decide per warp, scale per lane
          const float f = need ? ex2_approx(m_ref - m_new) : 1.0f;
          if (need) m_ref = m_new;
          l_run *= f;
#pragma unroll
          for (int c = 0; c < kPfD; c += 32) {
            uint32_t v[32];
            tmem_ld_32x32(o_addr + c, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; i++) v[i] = __float_as_uint(__uint_as_float(v[i]) * f);
            tmem_st_32x32(o_addr + c, v);
          }
          tmem_st_wait();
        }
      }
      const float neg_m = -m_ref;
      float ps4[4] = {0.f, 0.f, 0.f, 0.f};
      uint8_t *prow = smem + (p_base - smem_base) + (row >> 3) * 1024 + (row & 7) * 128;
#pragma unroll
      for (int c = 0; c < 8; c++) {
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
          const int i = c * 8 + e * 2;
          const float p0 = ex2_approx(fmaf(__uint_as_float(i < 32 ? s0[i] : s1[i - 32]), scale_log2, neg_m));
          const float p1 = ex2_approx(fmaf(__uint_as_float(i + 1 < 32 ? s0[i + 1] : s1[i + 1 - 32]), scale_log2, neg_m));
          ps4[e] += p0 + p1;
          pk[e] = float2_to_bf16x2(p0, p1);
        }
        *reinterpret_cast<uint4 *>(prow + ((c ^ (row & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      }
      l_run += (ps4[0] + ps4[1]) + (ps4[2] + ps4[3]);
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(p_full);
    }
    mbar_wait(o_full, (uint32_t)(n_tiles - 1) & 1u);
    tc_fence_after();
    const float inv = 1.0f / l_run;
    uint4 *o = reinterpret_cast<uint4 *>(out + ((size_t)(tok0 + t) * n_heads + kvh * G + hl) * kPfD);
#pragma unroll
    for (int c = 0; c < kPfD; c += 32) {
      uint32_t v[32];
      tmem_ld_32x32(o_addr + c, v);
      tmem_ld_wait();
      if (t < len) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
          uint4 w;
          w.x = float2_to_bf16x2(__uint_as_float(v[i * 8 + 0]) * inv, __uint_as_float(v[i * 8 + 1]) * inv);
          w.y = float2_to_bf16x2(__uint_as_float(v[i * 8 + 2]) * inv, __uint_as_float(v[i * 8 + 3]) * inv);
          w.z = float2_to_bf16x2(__uint_as_float(v[i * 8 + 4]) * inv, __uint_as_float(v[i * 8 + 5]) * inv);
          w.w = float2_to_bf16x2(__uint_as_float(v[i * 8 + 6]) * inv, __uint_as_float(v[i * 8 + 7]) * inv);
          o[c / 8 + i] = w;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<kPf2TmemCols>(tmem_base);
}

// Tokens per query block of the prefill attention for this head geometry: 128 / G on the tcgen05 kernel
// (head_dim 128, GQA group 4 or 8), 16 on the mma.sync kernel.  DIS_ATTN_PREFILL=mma forces the latter (A/B runs).
static int prefill_variant() {  // 0: mma.sync kernel, 1: tcgen05 with O folded in registers, 2: tcgen05, two CTAs per SM
  static int v = -1;
  if (v < 0) {
    const char *e = getenv("DIS_ATTN_PREFILL");
    v = !e ? 2 : (!strcmp(e, "mma") ? 0 : (!strcmp(e, "tc1") ? 1 : 2));
  }
  return v;
}
static bool prefill_tc_enabled() { return prefill_variant() != 0; }

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
  DIS_CUDA_TRY(once.run([&] {
    cudaError_t e = cudaFuncSetAttribute(attn_prefill_tc_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPfSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(attn_prefill_tc2_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPf2SmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(attn_prefill_tc2_kernel<G>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    return e;
  }));
  if (prefill_variant() == 2)
    return launch_k(attn_prefill_tc2_kernel<G>, dim3(n_blocks, n_kv), dim3(kPfThreads), (size_t)kPf2SmemBytes, st,
                    *reinterpret_cast<const CUtensorMap *>(&kv_map), row0, q, page_table, max_pages, qb_tok0, qb_len, qb_pos0,
                    qb_ptrow, out, n_heads, n_kv, page_shift);
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
