// k_gemm_tcgen05.cu -- the projection GEMM of the decoder, written for sm_100a:
// TMA (cp.async.bulk.tensor) stages bf16 tiles into 128B-swizzled shared memory, one elected thread
// issues tcgen05.mma with the fp32 accumulator in TMEM, four epilogue warps read it back with
// tcgen05.ld.  No reference counterpart (the reference has no tensor code at all).
//
// Orientation ("swap-AB"): the WEIGHT tile is the MMA's A operand (128 output features = 128 TMEM
// lanes) and the TOKEN tile is the B operand (UMMA N = 32/64/128 tokens = TMEM columns).  Decode
// batches of 32..128 tokens therefore fill a whole 128-row MMA instead of padding M to 128, and the
// weight matrix -- the only large operand -- is streamed exactly once per launch.
//
//   Y[t][n] = sum_k X[t][k] * W[n][k]         X: [T][K] bf16 (K-major B),  W: [N][K] bf16 (K-major A)
//
// Epilogues:
//   partial : out[s][t][n] fp32, s = split-K slice (consumers sum the slices: add_rmsnorm,
//             rope_kv_write, the sampler reads slice 0 of the lm_head logits)
//   silu    : two accumulators (gate rows, up rows of the fused [2I][K] weight) ->
//             act[t][n] = bf16(silu(gate) * up)
//
// Roofline: HBM-bound for T <= ~128 (AI ~ T flop/B against ~210 flop/B machine balance).
// Algorithmic bytes per launch = 2*N*K (weights) + 2*T*K (activations) + 4*S*T*N (partials) or 2*T*I.
#include <cuda.h>

#include <cstdlib>

#include "kernels.h"
#include "ptx.cuh"

namespace dis {

// (mbarrier / TMA / PDL wrappers: ptx.cuh)
// (tcgen05 / TMEM wrappers and the UMMA descriptors: ptx.cuh)
// ------------------------------------------------------------------------------------ tile config

// The ring is sized at launch: a stage holds (1 or 2) weight tiles of tile_rows x 64 bf16 and one token tile, so a
// short weight tile (tensor-parallel shard, balanced gate-up tile) buys MORE stages -- the k-loop of a weight-
// streaming GEMM advances one k-block per (memory latency / stages), whatever the stage holds.
constexpr int kMaxStages = 12;
constexpr int kMaxDynSmem = 227 * 1024;

template <int BN, bool kDual>
struct GemmCfg {
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kAccCols = (kDual ? 2 : 1) * BN;
  static constexpr int kTmemCols = kAccCols <= 32 ? 32 : kAccCols <= 64 ? 64 : kAccCols <= 128 ? 128 : kAccCols <= 256 ? 256 : 512;
  static int stage_bytes(int tile_rows) { return (kDual ? 2 : 1) * tile_rows * kBK * 2 + kBBytes; }
  // The MMA always reads a 128-row A tile: behind a shorter tile it reads (and ignores) whatever follows, so the
  // allocation is padded to keep the last stage's 16 KB read inside the CTA's shared memory.
  static int smem_bytes(int stages, int tile_rows) {
    return stages * stage_bytes(tile_rows) + 1024 /*align slack*/ + 256 /*barriers*/ + (tile_rows < kBM ? kBM * kBK * 2 : 0);
  }
};

template <int BN, bool kDual>
__global__ void __launch_bounds__(kGemmThreads)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap w_map, const __grid_constant__ CUtensorMap x_map,
                    float *__restrict__ out_partial, uint16_t *__restrict__ out_act, int T, int N, int K,
                    int w_row0, int up_row_off, int kb_per_split, int tile_rows, int n_stages, int early_tma) {
  using Cfg = GemmCfg<BN, kDual>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_bytes = (uint32_t)tile_rows * kBK * 2;  // a multiple of 1024 (tile_rows % 8 == 0): swizzle atoms stay aligned
  const uint32_t stage_bytes = (kDual ? 2u : 1u) * a_bytes + Cfg::kBBytes;
  const uint32_t bar_base = smem_base + (uint32_t)n_stages * stage_bytes;
  // barriers: full[n_stages], empty[n_stages], tmem_full, then the TMEM base address word
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (n_stages + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * n_stages);
  const uint32_t tmem_slot = tmem_full_bar + 8u;
  auto a_tile = [&](int s) { return smem_base + s * stage_bytes; };
  auto a2_tile = [&](int s) { return a_tile(s) + a_bytes; };
  auto b_tile = [&](int s) { return a_tile(s) + (kDual ? 2u : 1u) * a_bytes; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // tile_rows (<= 128, the weight map's box height) output features per CTA: a 128-row MMA runs on a
  // partly filled A tile; lanes >= tile_rows compute on stale shared memory and are never stored.
  const int n0 = blockIdx.x * tile_rows;
  const uint32_t stage_tx = stage_bytes;
  const int t0 = blockIdx.y * BN;
  const int num_kb = (K + kBK - 1) / kBK;
  const int kb0 = blockIdx.z * kb_per_split;
  const int kb1 = min(num_kb, kb0 + kb_per_split);

  const int pre = min(kb1 - kb0, n_stages);
  auto issue_weights = [&]() {
    for (int i = 0; i < pre; i++) {
      mbar_expect_tx(full_bar(i), stage_tx);
      tma_load_2d(a_tile(i), &w_map, (kb0 + i) * kBK, w_row0 + n0, full_bar(i), kEvictFirst);
      if (kDual) tma_load_2d(a2_tile(i), &w_map, (kb0 + i) * kBK, w_row0 + up_row_off + n0, full_bar(i), kEvictFirst);
    }
  };
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&w_map);
    tma_prefetch_desc(&x_map);
    for (int s = 0; s < n_stages; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
    fence_proxy_async();
    // Weights depend neither on the previous kernel nor on this CTA's TMEM allocation: the thread that has just
    // initialised the barriers can fill the ring with weight tiles at once, before the CTA-wide barrier below
    // (DIS_GEMM_EARLY_TMA=0 keeps the round-1 order for A/B runs)
    if (early_tma) issue_weights();
  }
  if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  pdl_launch_dependents();  // the next kernel may start its own prologue / weight prefetch
  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      // Weights do not depend on the previous kernel: fill the ring with weight tiles first, and only
      // then wait for the producer of the activations (programmatic dependent launch).
      if (!early_tma) issue_weights();
      pdl_wait();
      for (int i = 0; i < pre; i++) tma_load_2d(b_tile(i), &x_map, (kb0 + i) * kBK, t0, full_bar(i), kEvictLast);
      int s = pre == n_stages ? 0 : pre;
      uint32_t phase = pre == n_stages ? 1u : 0u;
      for (int kb = kb0 + pre; kb < kb1; kb++) {
        mbar_wait(empty_bar(s), phase ^ 1u);
        mbar_expect_tx(full_bar(s), stage_tx);
        tma_load_2d(a_tile(s), &w_map, kb * kBK, w_row0 + n0, full_bar(s), kEvictFirst);
        if (kDual) tma_load_2d(a2_tile(s), &w_map, kb * kBK, w_row0 + up_row_off + n0, full_bar(s), kEvictFirst);
        tma_load_2d(b_tile(s), &x_map, kb * kBK, t0, full_bar(s), kEvictLast);
        if (++s == n_stages) { s = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc<BN>();
      int s = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; kb++) {
        mbar_wait(full_bar(s), phase);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < kBK / 16; k++) {
          const uint32_t acc = (kb > kb0 || k > 0) ? 1u : 0u;
          const uint64_t bd = umma_desc_sw128(b_tile(s) + k * 32);
          umma_bf16(tmem_base, umma_desc_sw128(a_tile(s) + k * 32), bd, idesc, acc);
          if (kDual) umma_bf16(tmem_base + BN, umma_desc_sw128(a2_tile(s) + k * 32), bd, idesc, acc);
        }
        umma_commit(empty_bar(s));  // slot reusable once these MMAs have read it
        if (++s == n_stages) { s = 0; phase ^= 1u; }
      }
      umma_commit(tmem_full_bar);  // accumulator complete
    }
  } else {
    // ===== epilogue: TMEM -> registers -> global =====
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    pdl_wait();  // global writes only after every prerequisite grid has completed
    const int quarter = warp & 3;  // a warp may only touch TMEM lanes [32*(warp%4), +32)
    const int n = (quarter * 32 + lane < tile_rows) ? n0 + quarter * 32 + lane : N;  // N = masked
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
      uint32_t g[32];
      tmem_ld_32x32(lane_addr + c, g);
      if (kDual) {
        uint32_t u[32];
        tmem_ld_32x32(lane_addr + BN + c, u);
        tmem_ld_wait();
        if (n < N) {
#pragma unroll
          for (int jj = 0; jj < 32; jj++) {
            const int t = t0 + c + jj;
            if (t < T) {
              const float gv = __uint_as_float(g[jj]), uv = __uint_as_float(u[jj]);
              out_act[(size_t)t * N + n] = float_to_bf16_bits((gv / (1.0f + expf(-gv))) * uv);
            }
          }
        }
      } else {
        tmem_ld_wait();
        if (n < N) {
          float *o = out_partial + ((size_t)blockIdx.z * T + t0 + c) * N + n;
#pragma unroll
          for (int jj = 0; jj < 32; jj++)
            if (t0 + c + jj < T) o[(size_t)jj * N] = __uint_as_float(g[jj]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

// ------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

static_assert(sizeof(TmaMap) == sizeof(CUtensorMap), "TmaMap must mirror CUtensorMap");

cudaError_t make_tma_map_2d(TmaMap *out, const void *base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) { set_last_error("cuTensorMapEncodeTiled is not available from this driver"); return cudaErrorNotSupported; }
  if (cols % 8 != 0 || ((uintptr_t)base & 15u)) { set_last_error("TMA map: row length must be a multiple of 8 bf16 and base 16-byte aligned"); return cudaErrorInvalidValue; }
  if (box_rows == 0 || box_rows > 256) { set_last_error("TMA map: box rows must be 1..256"); return cudaErrorInvalidValue; }
  const cuuint64_t dims[2] = {cols, rows};
  const cuuint64_t strides[1] = {cols * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kBK, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(reinterpret_cast<CUtensorMap *>(out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base),
                        dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_last_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r)); return cudaErrorInvalidValue; }
  return cudaSuccess;
}

// UMMA N (token tile): 32 / 64 for decode batches (HBM-bound), 128 and 256 for prefill (the wider the token tile,
// the fewer bytes per flop cross the L2 -> SM fabric, which is what bounds the 128-wide tile)
int gemm_token_tile(int T) {
  // 128-token tiles up to T = 1023: twice the CTAs of 256-token tiles, two per SM (one CTA's epilogue overlaps the
  // other's k-loop) -- TTFT of a single 256 / 512-token request 7.1 -> 5.7 / 9.9 -> 9.0 ms; equal at 1024, and from
  // 2048 the 256-token tile (fewer bytes per flop through L2) wins: 33.6 vs 35.5 ms.
  return T <= 32 ? 32 : (T <= 64 ? 64 : (T < 1024 ? 128 : 256));
}

// Split-K factor: enough CTAs to keep every SM streaming (two CTAs per SM), each with >= 4 k-blocks.
int gemm_pick_split_k(int T, int N, int K) {
  const int bn = gemm_token_tile(T);
  const int tiles = ((N + kBM - 1) / kBM) * ((T + bn - 1) / bn);
  const int num_kb = (K + kBK - 1) / kBK;
  if (tiles >= kNumSMs) return 1;
  int s = (2 * kNumSMs) / tiles;
  s = s > 16 ? 16 : s;
  s = s > num_kb / 4 ? num_kb / 4 : s;
  if (s < 1) s = 1;
  const int per = (num_kb + s - 1) / s;
  return (num_kb + per - 1) / per;  // no empty slices
}

template <int BN, bool kDual>
static cudaError_t gemm_launch(const TmaMap &w_map, const TmaMap &x_map, float *out, uint16_t *act, int T, int N,
                               int K, int w_row0, int up_off, int split_k, int tile_rows, cudaStream_t st) {
  using Cfg = GemmCfg<BN, kDual>;
  auto kern = gemm_tcgen05_kernel<BN, kDual>;
  static DeviceOnce once;
  DIS_CUDA_TRY(once.run([&] { return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem); }));
  const int num_kb = (K + kBK - 1) / kBK;
  const int per = (num_kb + split_k - 1) / split_k;
  if ((num_kb + per - 1) / per != split_k) { set_last_error("gemm: split_k leaves an empty K slice; use gemm_pick_split_k"); return cudaErrorInvalidValue; }
  if (tile_rows < 8 || tile_rows > kBM || tile_rows % 8) { set_last_error("gemm: tile_rows must be a multiple of 8 in [8, 128]"); return cudaErrorInvalidValue; }
  dim3 grid((N + tile_rows - 1) / tile_rows, (T + BN - 1) / BN, split_k);
  // more CTAs than SMs: leave room for two per SM (one CTA's epilogue overlaps the other's stream, and under PDL
  // the next kernel's CTA can prefetch beside this one); otherwise one CTA per SM gets the whole shared memory
  const long long ctas = (long long)grid.x * grid.y * grid.z;
  // DIS_GEMM_SMEM_KB2 / DIS_GEMM_SMEM_KB1 (A/B runs): shared memory per CTA when two / one CTA per SM is intended; a
  // smaller ring leaves room for the NEXT kernel's CTAs to become resident and prefetch their weights meanwhile
  static int kb2 = 0, kb1 = 0;
  if (kb2 == 0) {
    const char *e2 = getenv("DIS_GEMM_SMEM_KB2"), *e1 = getenv("DIS_GEMM_SMEM_KB1");
    kb2 = e2 ? atoi(e2) : 110; kb1 = e1 ? atoi(e1) : 220;
    kb2 = kb2 < 48 ? 48 : (kb2 > 110 ? 110 : kb2); kb1 = kb1 < 64 ? 64 : (kb1 > 220 ? 220 : kb1);
  }
  const int budget = (ctas > kNumSMs && BN <= 128 && Cfg::kTmemCols <= 256) ? kb2 * 1024 : kb1 * 1024;
  int stages = (budget - Cfg::smem_bytes(0, tile_rows)) / Cfg::stage_bytes(tile_rows);
  stages = stages > kMaxStages ? kMaxStages : (stages < 2 ? 2 : stages);
  stages = stages > per ? (per < 2 ? 2 : per) : stages;  // never more stages than k-blocks
  const int smem = Cfg::smem_bytes(stages, tile_rows);
  if (smem > kMaxDynSmem) { set_last_error("gemm: tile does not fit shared memory"); return cudaErrorInvalidValue; }
  static int early = -1;
  if (early < 0) { const char *e = getenv("DIS_GEMM_EARLY_TMA"); early = (e && atoi(e) == 0) ? 0 : 1; }
  return launch_k(kern, grid, dim3(kGemmThreads), (size_t)smem, st, *reinterpret_cast<const CUtensorMap *>(&w_map),
                  *reinterpret_cast<const CUtensorMap *>(&x_map), out, act, T, N, K, w_row0, up_off, per, tile_rows, stages, early);
}

cudaError_t launch_gemm_partial(const TmaMap &w_map, const TmaMap &x_map, float *out, int T, int N, int K,
                                int w_row0, int split_k, cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  if (split_k < 1) { set_last_error("gemm: split_k must be >= 1"); return cudaErrorInvalidValue; }
  switch (gemm_token_tile(T)) {
    case 32: return gemm_launch<32, false>(w_map, x_map, out, nullptr, T, N, K, w_row0, 0, split_k, kBM, st);
    case 64: return gemm_launch<64, false>(w_map, x_map, out, nullptr, T, N, K, w_row0, 0, split_k, kBM, st);
    case 128: return gemm_launch<128, false>(w_map, x_map, out, nullptr, T, N, K, w_row0, 0, split_k, kBM, st);
    default: return gemm_launch<256, false>(w_map, x_map, out, nullptr, T, N, K, w_row0, 0, split_k, kBM, st);
  }
}

// Output features per CTA for the single-pass (no split-K) gate-up projection: as few CTAs as possible
// with at most one per SM, so that every SM streams the same share of the weight
// (128-row tiles put 14336 rows on 112 of 148 SMs: measured 76 % of the 148-SM rate).
int gemm_pick_tile_rows(int N) {
  int r = ((N + kNumSMs - 1) / kNumSMs + 7) / 8 * 8;
  return r > kBM ? kBM : (r < 8 ? 8 : r);
}

cudaError_t launch_gemm_silu(const TmaMap &w_map, const TmaMap &x_map, uint16_t *act, int T, int I, int K,
                             int w_row0, int tile_rows, cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  switch (gemm_token_tile(T)) {
    case 32: return gemm_launch<32, true>(w_map, x_map, nullptr, act, T, I, K, w_row0, I, 1, tile_rows, st);
    case 64: return gemm_launch<64, true>(w_map, x_map, nullptr, act, T, I, K, w_row0, I, 1, tile_rows, st);
    case 128: return gemm_launch<128, true>(w_map, x_map, nullptr, act, T, I, K, w_row0, I, 1, tile_rows, st);
    default: return gemm_launch<256, true>(w_map, x_map, nullptr, act, T, I, K, w_row0, I, 1, tile_rows, st);
  }
}

}  // namespace dis
