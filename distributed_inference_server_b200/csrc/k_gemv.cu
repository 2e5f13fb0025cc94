// k_gemv.cu -- bandwidth-bound GEMV for batch-1 (and very small batch) decode.
// y[t][n] = sum_k X[t][k] * W[n][k], bf16 weights streamed once with 16-byte no-allocate loads,
// fp32 accumulate, four weight rows per CTA with K split over its warps, warp-shuffle + shared-memory reduction.
// Algorithmic bytes per launch: 2*N*K (weights) + 2*T*K (activations, L1/L2 resident) + 4*T*N.
#include "kernels.h"
#include "ptx.cuh"

namespace dis {

constexpr int kGemvWarps = 8;
constexpr int kGemvW = 4;      // weight rows streamed per CTA: 4 output rows, or 2 gate + their 2 up rows when kSilu
constexpr int kGemvDepth = 2;  // k-chunks whose weight loads are in flight per lane

__device__ __forceinline__ float dot8(const uint4 &w, const uint4 &x) {
  const float2 w0 = bf16x2_to_float2(w.x), w1 = bf16x2_to_float2(w.y), w2 = bf16x2_to_float2(w.z), w3 = bf16x2_to_float2(w.w);
  const float2 x0 = bf16x2_to_float2(x.x), x1 = bf16x2_to_float2(x.y), x2 = bf16x2_to_float2(x.z), x3 = bf16x2_to_float2(x.w);
  float a = w0.x * x0.x;
  a = fmaf(w0.y, x0.y, a); a = fmaf(w1.x, x1.x, a); a = fmaf(w1.y, x1.y, a);
  a = fmaf(w2.x, x2.x, a); a = fmaf(w2.y, x2.y, a); a = fmaf(w3.x, x3.x, a); a = fmaf(w3.y, x3.y, a);
  return a;
}

// One CTA = kW weight rows (4, or 8 for batch 1; half of them gate and half up rows when kSilu); its 8 warps split K in interleaved 256-element chunks (chunk c belongs
// to warp c % 8, so the CTA as a whole walks every weight row in contiguous 4 KB pieces).  A lane keeps
// kGemvDepth chunks x kW rows of 16-byte loads in flight (128-256 B per lane, 32-64 KB per CTA, 2-3 CTAs per SM): enough
// bytes in flight to cover HBM latency at N = 4096, where a row-per-warp mapping leaves the SM at 14 warps
// (measured: batch-1 step 4.70 -> 3.16 ms).  The first kGemvDepth chunks are requested BEFORE
// griddepcontrol.wait -- weights do not depend on the predecessor kernel -- so under programmatic dependent
// launch the stream starts while the norm / attention kernel before it is still running.  Partial sums meet in
// shared memory and are added in warp order (deterministic).
// Measured and rejected: a persistent one-wave variant walking (row group, chunk) items with the register ring
// kept full across row groups -- 4.15 ms per step: the per-item bookkeeping and a barrier per row group cost more
// issue slots than the CTA turnover they remove; 4 rows x depth 4 -- 3.62 ms (fewer CTAs resident); this kernel as ONE
// wave of CTAs each walking ~13 row groups (so that the dependent kernel is released at once) -- 4.18 ms: without
// loads in flight across the group boundary the memory pipe idles during every reduction.  Round 2: the residual add +
// RMSNorm in front of the QKV / gate-up / lm_head GEMV folded INTO the GEMV (every CTA rebuilds the 8 KB normalised
// activation for itself from x + the pending output, which removes both norm kernels of a layer) -- 5.46 ms per step
// against 3.18: 32 KB of L2 reads, two block barriers and an rsqrt in front of every one of 768-16032 short CTAs cost far
// more than the two 4 us kernels they replace.  The tcgen05 path at T = 1 takes 3.06 ms against this kernel's 3.18 (same
// box, alternating runs: profiles/r2_bench_b1_*.json): TMA keeps ~200 KB per SM in flight without registers or issue
// slots and one wave of CTAs has no turnover, where this kernel stages its loads in registers (256 B per lane) and pays a
// reduction tail per 8 rows.  The engine therefore uses the tcgen05 path from T = 1 by default; this kernel stays
// selectable (DIS_FLAG_FORCE_GEMV, DIS_GEMV_MAX_T) and parity-tested.
// kSilu: rows n < N are gate rows, rows N + n the matching up rows; output bf16 act[t][n].
template <int kT, bool kSilu, int kW = kGemvW, int kDepthT = kGemvDepth>
__global__ void __launch_bounds__(kGemvWarps * 32, (kT == 1 && kW * kDepthT <= 8) ? 3 : 2)
gemv_kernel(const uint16_t *__restrict__ X, const uint16_t *__restrict__ W, float *__restrict__ out,
            uint16_t *__restrict__ act, int N, int K) {
  constexpr int kRows = kSilu ? kW / 2 : kW, kGemvDepth = kDepthT;
  __shared__ float part[kGemvWarps][kW * kT];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * kRows;
  pdl_launch_dependents();
  const uint16_t *wrow[kW];
#pragma unroll
  for (int r = 0; r < kRows; r++) {
    const int n = min(n0 + r, N - 1);
    wrow[r] = W + (size_t)n * K;
    if (kSilu) wrow[kRows + r] = W + (size_t)(N + n) * K;
  }
  constexpr int kChunk = 256, kStride = kChunk * kGemvWarps;
  const int k0 = warp * kChunk + lane * 8;
  uint4 w[kGemvDepth][kW];
#pragma unroll
  for (int d = 0; d < kGemvDepth; d++) {
    const int k = k0 + d * kStride;
#pragma unroll
    for (int r = 0; r < kW; r++) w[d][r] = k < K ? ld_stream_16(wrow[r] + k) : make_uint4(0u, 0u, 0u, 0u);
  }
  pdl_wait();
  float acc[kW][kT];
#pragma unroll
  for (int r = 0; r < kW; r++)
#pragma unroll
    for (int t = 0; t < kT; t++) acc[r][t] = 0.f;

  for (int kb = k0; kb < K; kb += kGemvDepth * kStride) {
#pragma unroll
    for (int d = 0; d < kGemvDepth; d++) {
      const int k = kb + d * kStride;
      uint4 cur[kW];
#pragma unroll
      for (int r = 0; r < kW; r++) cur[r] = w[d][r];
      const int kn = k + kGemvDepth * kStride;  // refill this slot with the chunk kGemvDepth strides ahead
      if (kn < K) {
#pragma unroll
        for (int r = 0; r < kW; r++) w[d][r] = ld_stream_16(wrow[r] + kn);
      }
      if (k < K) {
#pragma unroll
        for (int t = 0; t < kT; t++) {
          const uint4 x = __ldg(reinterpret_cast<const uint4 *>(X + (size_t)t * K + k));
#pragma unroll
          for (int r = 0; r < kW; r++) acc[r][t] += dot8(cur[r], x);
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < kW; r++)
#pragma unroll
    for (int t = 0; t < kT; t++) {
      const float v = warp_sum(acc[r][t]);
      if (lane == 0) part[warp][r * kT + t] = v;
    }
  __syncthreads();
  if (threadIdx.x < kRows * kT) {
    const int r = threadIdx.x / kT, t = threadIdx.x % kT;
    const int n = n0 + r;
    if (n < N) {
      float g = 0.f, u = 0.f;
#pragma unroll
      for (int q = 0; q < kGemvWarps; q++) {
        g += part[q][r * kT + t];
        if (kSilu) u += part[q][(kRows + r) * kT + t];
      }
      if (kSilu) act[(size_t)t * N + n] = float_to_bf16_bits((g / (1.0f + expf(-g))) * u);
      else out[(size_t)t * N + n] = g;
    }
  }
}

template <bool kSilu>
static cudaError_t gemv_dispatch(const uint16_t *X, const uint16_t *W, float *out, uint16_t *act, int T, int N,
                                 int K, cudaStream_t st) {
  if (K % 8 != 0) { set_last_error("gemv: K must be a multiple of 8"); return cudaErrorInvalidValue; }
  if (T == 1) {  // batch-1 decode: 8 weight rows per CTA (half the CTA turnover; measured 3.30 -> 3.16 ms per step)
    const int r1 = kSilu ? 4 : 8;
    return launch_k(gemv_kernel<1, kSilu, 8, 2>, dim3((N + r1 - 1) / r1), dim3(kGemvWarps * 32), 0, st, X, W, out, act, N, K);
  }
  const int rows = kSilu ? kGemvW / 2 : kGemvW;
  const int blocks = (N + rows - 1) / rows;
  for (int t0 = 0; t0 < T; t0 += kGemvMaxT) {
    const int tn = min(kGemvMaxT, T - t0);
    const uint16_t *x = X + (size_t)t0 * K;
    float *o = out ? out + (size_t)t0 * N : nullptr;
    uint16_t *a = act ? act + (size_t)t0 * N : nullptr;
    switch (tn) {
      case 1: DIS_CUDA_TRY(launch_k(gemv_kernel<1, kSilu>, dim3(blocks), dim3(kGemvWarps * 32), 0, st, x, W, o, a, N, K)); break;
      case 2: DIS_CUDA_TRY(launch_k(gemv_kernel<2, kSilu>, dim3(blocks), dim3(kGemvWarps * 32), 0, st, x, W, o, a, N, K)); break;
      case 3: DIS_CUDA_TRY(launch_k(gemv_kernel<3, kSilu>, dim3(blocks), dim3(kGemvWarps * 32), 0, st, x, W, o, a, N, K)); break;
      default: DIS_CUDA_TRY(launch_k(gemv_kernel<4, kSilu>, dim3(blocks), dim3(kGemvWarps * 32), 0, st, x, W, o, a, N, K)); break;
    }
    DIS_CUDA_TRY(cudaGetLastError());
  }
  return cudaSuccess;
}

cudaError_t launch_gemv(const uint16_t *X, const uint16_t *W, float *out, int T, int N, int K, cudaStream_t st) {
  return gemv_dispatch<false>(X, W, out, nullptr, T, N, K, st);
}
cudaError_t launch_gemv_silu(const uint16_t *X, const uint16_t *Wgu, uint16_t *act, int T, int I, int K,
                             cudaStream_t st) {
  return gemv_dispatch<true>(X, Wgu, nullptr, act, T, I, K, st);
}

}  // namespace dis
