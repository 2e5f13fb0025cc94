// k_gemm_persist.cu -- the projection GEMM for PREFILL-sized token counts, sm_100a.
//
// Same orientation and operand staging as k_gemm_tcgen05.cu (weights = MMA A operand, 128 features = 128 TMEM
// lanes; tokens = B operand; TMA into 128B-swizzled shared memory; one thread issues tcgen05.mma), but built for
// the tensor-bound regime, where that kernel's one-tile-per-CTA schedule leaves the tensor pipe idle while four
// warps drain the accumulator (ncu launch list, T = 8192: 1.0 PFLOP/s on the K = 4096 projections against
// 1.3 PFLOP/s on the K = 14336 one, whose longer k-loop amortises the same epilogue):
//   * persistent: one CTA per SM walks tiles blockIdx.x, blockIdx.x + grid, ...; the TMA ring never drains
//     between tiles;
//   * the accumulator is double-buffered in TMEM (2 x 256 columns = all 512): the MMA warp starts tile i+1 in
//     one buffer while the epilogue warps read tile i from the other (acc_full / acc_empty mbarriers);
//   * tiles are 128 features x 256 tokens.  The gate-up projection runs as gemm_persist_pair_kernel below (64 gate
//     rows over their 64 up rows in ONE 128 x 256 MMA); the kDual = true instantiation of this kernel (2 x 128
//     features x 128 tokens, two N = 128 MMAs per k-step) is the measured alternative it replaced (1.17 against
//     1.28 PFLOP/s) and is no longer instantiated.
// No split-K: callers use it only when there are at least two tiles per SM (gemm_use_persist).
#include <cuda.h>

#include "kernels.h"
#include "ptx.cuh"

namespace dis {

template <int BN, bool kDual>
struct PersistCfg {
  static constexpr int kABytes = kBM * kBK * 2;                       // 16 KB
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = (kDual ? 2 : 1) * kABytes + kBBytes;  // 48 KB for <256,false> and <128,true>
  static constexpr int kStages = 4;
  static constexpr int kAccCols = (kDual ? 2 : 1) * BN;               // 256
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
  static_assert(kAccCols == 256, "two accumulator buffers must fill the 512 TMEM columns");
};

template <int BN, bool kDual>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_persist_kernel(const __grid_constant__ CUtensorMap w_map, const __grid_constant__ CUtensorMap x_map,
                    float *__restrict__ out_f32, uint16_t *__restrict__ out_act, int T, int N, int K, int w_row0,
                    int up_row_off, int tiles_n, int tiles_t, int t_fast) {
  using Cfg = PersistCfg<BN, kDual>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + Cfg::kStages * Cfg::kStageBytes;
  // barriers: full[kStages], empty[kStages], acc_full[2], acc_empty[2]; then the TMEM base word
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (Cfg::kStages + s); };
  auto acc_full = [&](int b) { return bar_base + 8u * (2 * Cfg::kStages + b); };
  auto acc_empty = [&](int b) { return bar_base + 8u * (2 * Cfg::kStages + 2 + b); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * Cfg::kStages + 4);
  auto a_tile = [&](int s) { return smem_base + s * Cfg::kStageBytes; };
  auto a2_tile = [&](int s) { return a_tile(s) + Cfg::kABytes; };
  auto b_tile = [&](int s) { return a_tile(s) + (kDual ? 2 : 1) * Cfg::kABytes; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = (K + kBK - 1) / kBK;
  const int n_tiles = tiles_n * tiles_t;
  // tile -> (weight tile, token tile).  The ~148 tiles in flight together should cover ALL tiles of the smaller
  // operand (it stays in L2 and is re-read every round) and only a few of the larger one (streamed from HBM
  // once): token tile fastest when the weight matrix is the larger operand (gate-up: 235 MB of weights against
  // 67 MB of activations at T = 8192 -- weight-fastest order would stream the weights once per token tile).
  auto tile_n0 = [&](int tile) { return (t_fast ? tile / tiles_t : tile % tiles_n) * kBM; };
  auto tile_t0 = [&](int tile) { return (t_fast ? tile % tiles_t : tile / tiles_n) * BN; };
  const uint64_t w_hint = t_fast ? kEvictFirst : kEvictLast, x_hint = t_fast ? kEvictLast : kEvictFirst;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&w_map);
    tma_prefetch_desc(&x_map);
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int b = 0; b < 2; b++) { mbar_init(acc_full(b), 1); mbar_init(acc_empty(b), 4); }  // 4 epilogue warps
    fence_barrier_init();
    fence_proxy_async();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  pdl_launch_dependents();

  if (warp == 0) {
    // ===== TMA producer: one flat stream of (tile, k-block) stages =====
    if (lane == 0) {
      pdl_wait();  // activations come from the previous kernel
      int s = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int n0 = tile_n0(tile), t0 = tile_t0(tile);
        for (int kb = 0; kb < num_kb; kb++) {
          mbar_wait(empty_bar(s), phase ^ 1u);  // passes on the first lap
          mbar_expect_tx(full_bar(s), Cfg::kStageBytes);
          tma_load_2d(a_tile(s), &w_map, kb * kBK, w_row0 + n0, full_bar(s), w_hint);
          if (kDual) tma_load_2d(a2_tile(s), &w_map, kb * kBK, w_row0 + up_row_off + n0, full_bar(s), w_hint);
          tma_load_2d(b_tile(s), &x_map, kb * kBK, t0, full_bar(s), x_hint);
          if (++s == Cfg::kStages) { s = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread): accumulator buffers alternate per tile =====
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc<BN>();
      int s = 0, buf = 0;
      uint32_t phase = 0, aphase[2] = {0u, 0u};
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        mbar_wait(acc_empty(buf), aphase[buf] ^ 1u);  // the epilogue has drained this buffer (passes on first use)
        tc_fence_after();
        const uint32_t acc_addr = tmem_base + (uint32_t)(buf * Cfg::kAccCols);
        for (int kb = 0; kb < num_kb; kb++) {
          mbar_wait(full_bar(s), phase);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < kBK / 16; k++) {
            const uint32_t acc = (kb > 0 || k > 0) ? 1u : 0u;
            const uint64_t bd = umma_desc_sw128(b_tile(s) + k * 32);
            umma_bf16(acc_addr, umma_desc_sw128(a_tile(s) + k * 32), bd, idesc, acc);
            if (kDual) umma_bf16(acc_addr + BN, umma_desc_sw128(a2_tile(s) + k * 32), bd, idesc, acc);
          }
          umma_commit(empty_bar(s));  // slot reusable once these MMAs have read it
          if (++s == Cfg::kStages) { s = 0; phase ^= 1u; }
        }
        umma_commit(acc_full(buf));  // this tile's accumulator is complete
        aphase[buf] ^= 1u;
        buf ^= 1;
      }
    }
  } else {
    // ===== epilogue warps: TMEM -> registers -> global, one tile behind the MMA warp =====
    const int quarter = warp & 3;  // a warp may only touch TMEM lanes [32*(warp%4), +32)
    int buf = 0;
    uint32_t aphase[2] = {0u, 0u};
    bool waited = false;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const int n = tile_n0(tile) + quarter * 32 + lane, t0 = tile_t0(tile);
      mbar_wait(acc_full(buf), aphase[buf]);
      tc_fence_after();
      if (!waited) { pdl_wait(); waited = true; }  // global writes only after every prerequisite grid has completed
      const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * Cfg::kAccCols);
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
            float *o = out_f32 + (size_t)(t0 + c) * N + n;
#pragma unroll
            for (int jj = 0; jj < 32; jj++)
              if (t0 + c + jj < T) o[(size_t)jj * N] = __uint_as_float(g[jj]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty(buf));
      aphase[buf] ^= 1u;
      buf ^= 1;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------ gate-up, paired tile
// The gate-up projection with the SiLU-gate epilogue as ONE 128 x 256 MMA per k-step: the A tile is 64 gate rows
// (TMEM lanes 0-63) over their 64 up rows (lanes 64-127), so the tile keeps the 256-token width -- a 128-wide
// token tile re-reads the A operand from shared memory twice as often and measured 1.17 PFLOP/s against 1.3-1.4
// for the 256-wide plain tiles -- and its accumulator still fits twice into TMEM.  gate and up of one feature sit
// in different lane quarters, i.e. different epilogue warps: the up warps (quarters 2, 3) pass their values to the
// gate warps (quarters 0, 1) for the even 32-column chunks and the gate warps theirs to the up warps for the odd
// ones, through shared memory, double-buffered, one named barrier (64 threads: an up warp + a gate warp) per pair
// of chunks -- all four warps do an equal share of the SiLU and of the stores.
constexpr int kPairChunkFloats = 32 * 32;                    // [token column][lane]
constexpr int kPairXchgBytes = 2 /*pairs*/ * 2 /*buffers*/ * 2 /*directions*/ * kPairChunkFloats * 4;  // 32 KB

__device__ __forceinline__ void named_bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void named_bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }

__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_persist_pair_kernel(const __grid_constant__ CUtensorMap w_map64, const __grid_constant__ CUtensorMap x_map,
                         uint16_t *__restrict__ out_act, int T, int I, int K, int w_row0, int tiles_n, int tiles_t,
                         int t_fast) {
  constexpr int BN = 256, kHalf = 64;
  using Cfg = PersistCfg<256, false>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t xchg_base = smem_base + Cfg::kStages * Cfg::kStageBytes;
  const uint32_t bar_base = xchg_base + kPairXchgBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (Cfg::kStages + s); };
  auto acc_full = [&](int b) { return bar_base + 8u * (2 * Cfg::kStages + b); };
  auto acc_empty = [&](int b) { return bar_base + 8u * (2 * Cfg::kStages + 2 + b); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * Cfg::kStages + 4);
  auto a_tile = [&](int s) { return smem_base + s * Cfg::kStageBytes; };
  auto b_tile = [&](int s) { return a_tile(s) + Cfg::kABytes; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = (K + kBK - 1) / kBK;
  const int n_tiles = tiles_n * tiles_t;
  auto tile_n0 = [&](int tile) { return (t_fast ? tile / tiles_t : tile % tiles_n) * kHalf; };
  auto tile_t0 = [&](int tile) { return (t_fast ? tile % tiles_t : tile / tiles_n) * BN; };
  const uint64_t w_hint = t_fast ? kEvictFirst : kEvictLast, x_hint = t_fast ? kEvictLast : kEvictFirst;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&w_map64);
    tma_prefetch_desc(&x_map);
    for (int s = 0; s < Cfg::kStages; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int b = 0; b < 2; b++) { mbar_init(acc_full(b), 1); mbar_init(acc_empty(b), 4); }
    fence_barrier_init();
    fence_proxy_async();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  pdl_launch_dependents();

  if (warp == 0) {
    if (lane == 0) {
      pdl_wait();
      int s = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int n0 = tile_n0(tile), t0 = tile_t0(tile);
        for (int kb = 0; kb < num_kb; kb++) {
          mbar_wait(empty_bar(s), phase ^ 1u);
          mbar_expect_tx(full_bar(s), Cfg::kStageBytes);
          tma_load_2d(a_tile(s), &w_map64, kb * kBK, w_row0 + n0, full_bar(s), w_hint);                      // gate rows -> lanes 0-63
          tma_load_2d(a_tile(s) + kHalf * kBK * 2, &w_map64, kb * kBK, w_row0 + I + n0, full_bar(s), w_hint);  // up rows -> lanes 64-127
          tma_load_2d(b_tile(s), &x_map, kb * kBK, t0, full_bar(s), x_hint);
          if (++s == Cfg::kStages) { s = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc<BN>();
      int s = 0, buf = 0;
      uint32_t phase = 0, aphase[2] = {0u, 0u};
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        mbar_wait(acc_empty(buf), aphase[buf] ^ 1u);
        tc_fence_after();
        const uint32_t acc_addr = tmem_base + (uint32_t)(buf * BN);
        for (int kb = 0; kb < num_kb; kb++) {
          mbar_wait(full_bar(s), phase);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < kBK / 16; k++)
            umma_bf16(acc_addr, umma_desc_sw128(a_tile(s) + k * 32), umma_desc_sw128(b_tile(s) + k * 32), idesc,
                      (kb > 0 || k > 0) ? 1u : 0u);
          umma_commit(empty_bar(s));
          if (++s == Cfg::kStages) { s = 0; phase ^= 1u; }
        }
        umma_commit(acc_full(buf));
        aphase[buf] ^= 1u;
        buf ^= 1;
      }
    }
  } else {
    const int quarter = warp & 3;
    const bool is_up = quarter >= 2;
    const int pair = quarter & 1;  // gate quarter q works with up quarter q + 2
    // exchange area of this pair: [2 buffers][2 directions][32 columns][32 lanes] floats
    float *xchg = reinterpret_cast<float *>(smem_raw + (xchg_base - smem_u32(smem_raw))) + (size_t)pair * 4 * kPairChunkFloats;
    const int bar_id = 1 + pair;  // named barrier of the pair (64 threads)
    int step = 0;                 // chunk pairs handled so far: selects the exchange buffer
    int buf = 0;
    uint32_t aphase[2] = {0u, 0u};
    bool waited = false;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const int n = tile_n0(tile) + pair * 32 + lane, t0 = tile_t0(tile);
      mbar_wait(acc_full(buf), aphase[buf]);
      tc_fence_after();
      if (!waited) { pdl_wait(); waited = true; }
      const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * BN);
      // Columns in pairs of 32-column chunks: the gate warp finishes the even chunk (it needs the up values of it),
      // the up warp the odd one (it needs the gate values): each warp sends one chunk, keeps one, stores one.
#pragma unroll 1
      for (int c = 0; c < BN; c += 64, step++) {
        float *give = xchg + ((step & 1) * 2 + (is_up ? 0 : 1)) * kPairChunkFloats;  // what this warp sends
        float *take = xchg + ((step & 1) * 2 + (is_up ? 1 : 0)) * kPairChunkFloats;  // what it receives
        const int c_send = c + (is_up ? 0 : 32), c_keep = c + (is_up ? 32 : 0);
        uint32_t sv[32], kv[32];
        tmem_ld_32x32(lane_addr + c_send, sv);
        tmem_ld_32x32(lane_addr + c_keep, kv);
        tmem_ld_wait();
#pragma unroll
        for (int jj = 0; jj < 32; jj++) give[jj * 32 + lane] = __uint_as_float(sv[jj]);
        named_bar_sync(bar_id, 64);  // both directions written; the buffers of step - 1 are free again after this
        if (n < I) {
#pragma unroll
          for (int jj = 0; jj < 32; jj++) {
            const int t = t0 + c_keep + jj;
            if (t < T) {
              const float other = take[jj * 32 + lane], mine = __uint_as_float(kv[jj]);
              const float gv = is_up ? other : mine, uv = is_up ? mine : other;
              out_act[(size_t)t * I + n] = float_to_bf16_bits((gv / (1.0f + expf(-gv))) * uv);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty(buf));
      aphase[buf] ^= 1u;
      buf ^= 1;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// Worth it from two tiles per SM (tiles: 128 features, or 64 gate + 64 up features, x 256 tokens).
bool gemm_use_persist(int T, int N, bool dual) {
  if (T < 512) return false;
  const int rows = dual ? 64 : kBM;  // output features per tile
  const long long tiles = (long long)((N + rows - 1) / rows) * ((T + 255) / 256);
  return tiles >= 2 * kNumSMs;
}

template <int BN, bool kDual>
static cudaError_t persist_launch(const TmaMap &w_map, const TmaMap &x_map, float *out, uint16_t *act, int T, int N,
                                  int K, int w_row0, int up_off, cudaStream_t st) {
  using Cfg = PersistCfg<BN, kDual>;
  auto kern = gemm_persist_kernel<BN, kDual>;
  static DeviceOnce once;
  DIS_CUDA_TRY(once.run([&] { return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes); }));
  static int n_sm = 0;
  if (!n_sm) {
    int dev = 0;
    DIS_CUDA_TRY(cudaGetDevice(&dev));
    DIS_CUDA_TRY(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  }
  const int tiles_n = (N + kBM - 1) / kBM, tiles_t = (T + BN - 1) / BN;
  const int n_tiles = tiles_n * tiles_t;
  const int grid = n_tiles < n_sm ? n_tiles : n_sm;
  const int t_fast = (long long)(kDual ? 2 : 1) * N > (long long)T ? 1 : 0;  // weight rows vs token rows (same K)
  return launch_k(kern, dim3(grid), dim3(kGemmThreads), (size_t)Cfg::kSmemBytes, st, *reinterpret_cast<const CUtensorMap *>(&w_map),
                  *reinterpret_cast<const CUtensorMap *>(&x_map), out, act, T, N, K, w_row0, up_off, tiles_n, tiles_t, t_fast);
}

// w_map: 128-row boxes; x_map: 256-row boxes.  out[t][n] fp32 (one slice).
cudaError_t launch_gemm_persist_partial(const TmaMap &w_map, const TmaMap &x_map256, float *out, int T, int N, int K,
                                        int w_row0, cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  return persist_launch<256, false>(w_map, x_map256, out, nullptr, T, N, K, w_row0, 0, st);
}

// w_map64: 64-row boxes over the fused [2I][K] weight (gate rows, then up rows); x_map256: 256-row boxes.
// act[t][n] = bf16(silu(gate) * up).
cudaError_t launch_gemm_persist_silu(const TmaMap &w_map64, const TmaMap &x_map256, uint16_t *act, int T, int I, int K,
                                     int w_row0, cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  using Cfg = PersistCfg<256, false>;
  constexpr int smem = Cfg::kSmemBytes + kPairXchgBytes;
  static DeviceOnce once;
  DIS_CUDA_TRY(once.run([&] { return cudaFuncSetAttribute(gemm_persist_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); }));
  static int n_sm = 0;
  if (!n_sm) {
    int dev = 0;
    DIS_CUDA_TRY(cudaGetDevice(&dev));
    DIS_CUDA_TRY(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  }
  const int tiles_n = (I + 63) / 64, tiles_t = (T + 255) / 256;
  const int n_tiles = tiles_n * tiles_t;
  const int grid = n_tiles < n_sm ? n_tiles : n_sm;
  const int t_fast = 2LL * I > (long long)T ? 1 : 0;
  return launch_k(gemm_persist_pair_kernel, dim3(grid), dim3(kGemmThreads), (size_t)smem, st,
                  *reinterpret_cast<const CUtensorMap *>(&w_map64), *reinterpret_cast<const CUtensorMap *>(&x_map256), act, T, I,
                  K, w_row0, tiles_n, tiles_t, t_fast);
}

}  // namespace dis
