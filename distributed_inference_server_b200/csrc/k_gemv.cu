// k_gemv.cu -- bandwidth-bound GEMV for batch-1 (and very small batch) decode.
// y[t][n] = sum_k X[t][k] * W[n][k], bf16 weights streamed once with 16-byte no-allocate loads,
// fp32 accumulate, one warp per pair of output rows, warp-shuffle reduction.
// Algorithmic bytes per launch: 2*N*K (weights) + 2*T*K (activations, L1/L2 resident) + 4*T*N.
#include "kernels.h"
#include "ptx.cuh"

namespace dis {

constexpr int kGemvWarps = 8;
constexpr int kGemvRows = 2;  // rows per warp

__device__ __forceinline__ float dot8(const uint4 &w, const uint4 &x) {
  const float2 w0 = bf16x2_to_float2(w.x), w1 = bf16x2_to_float2(w.y), w2 = bf16x2_to_float2(w.z), w3 = bf16x2_to_float2(w.w);
  const float2 x0 = bf16x2_to_float2(x.x), x1 = bf16x2_to_float2(x.y), x2 = bf16x2_to_float2(x.z), x3 = bf16x2_to_float2(x.w);
  float a = w0.x * x0.x;
  a = fmaf(w0.y, x0.y, a); a = fmaf(w1.x, x1.x, a); a = fmaf(w1.y, x1.y, a);
  a = fmaf(w2.x, x2.x, a); a = fmaf(w2.y, x2.y, a); a = fmaf(w3.x, x3.x, a); a = fmaf(w3.y, x3.y, a);
  return a;
}

// kSilu: rows n < N are gate rows, rows N + n the matching up rows; output bf16 act[t][n].
template <int kT, bool kSilu>
__global__ void __launch_bounds__(kGemvWarps * 32)
gemv_kernel(const uint16_t *__restrict__ X, const uint16_t *__restrict__ W, float *__restrict__ out,
            uint16_t *__restrict__ act, int N, int K) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = (blockIdx.x * kGemvWarps + warp) * kGemvRows;
  pdl_launch_dependents();
  pdl_wait();
  if (n0 >= N) return;
  constexpr int kW = kSilu ? 2 * kGemvRows : kGemvRows;  // weight rows streamed per warp
  const uint16_t *wrow[kW];
#pragma unroll
  for (int r = 0; r < kGemvRows; r++) {
    const int n = min(n0 + r, N - 1);
    wrow[r] = W + (size_t)n * K;
    if (kSilu) wrow[kGemvRows + r] = W + (size_t)(N + n) * K;
  }
  float acc[kW][kT];
#pragma unroll
  for (int r = 0; r < kW; r++)
#pragma unroll
    for (int t = 0; t < kT; t++) acc[r][t] = 0.f;

#pragma unroll 2
  for (int k = lane * 8; k < K; k += 256) {
    uint4 w[kW];
#pragma unroll
    for (int r = 0; r < kW; r++) w[r] = ld_stream_16(wrow[r] + k);
#pragma unroll
    for (int t = 0; t < kT; t++) {
      const uint4 x = __ldg(reinterpret_cast<const uint4 *>(X + (size_t)t * K + k));
#pragma unroll
      for (int r = 0; r < kW; r++) acc[r][t] += dot8(w[r], x);
    }
  }
#pragma unroll
  for (int r = 0; r < kW; r++)
#pragma unroll
    for (int t = 0; t < kT; t++) acc[r][t] = warp_sum(acc[r][t]);
  if (lane == 0) {
#pragma unroll
    for (int r = 0; r < kGemvRows; r++) {
      const int n = n0 + r;
      if (n >= N) break;
#pragma unroll
      for (int t = 0; t < kT; t++) {
        if (kSilu) {
          const float g = acc[r][t], u = acc[kGemvRows + r][t];
          act[(size_t)t * N + n] = float_to_bf16_bits((g / (1.0f + expf(-g))) * u);
        } else {
          out[(size_t)t * N + n] = acc[r][t];
        }
      }
    }
  }
}

template <bool kSilu>
static cudaError_t gemv_dispatch(const uint16_t *X, const uint16_t *W, float *out, uint16_t *act, int T, int N,
                                 int K, cudaStream_t st) {
  if (K % 8 != 0) { set_last_error("gemv: K must be a multiple of 8"); return cudaErrorInvalidValue; }
  const int blocks = (N + kGemvWarps * kGemvRows - 1) / (kGemvWarps * kGemvRows);
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
