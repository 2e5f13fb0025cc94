// common.cuh -- shared device/host helpers for the sm_100a kernels.
#pragma once
#include <atomic>
#include <mutex>

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

namespace dis {

// ---- error plumbing ---------------------------------------------------------------------------
// Kernel launchers return cudaError_t; the C-ABI layer turns it into dis_status + message.
void set_last_error(const std::string &msg);
const char *get_last_error();

#define DIS_CUDA_TRY(expr)                                                                         \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) {                                                                       \
      ::dis::set_last_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + \
                            ":" + std::to_string(__LINE__) + ")");                                 \
      return _e;                                                                                   \
    }                                                                                              \
  } while (0)

constexpr int kNumSMs = 148;

// cudaFuncSetAttribute applies to the CURRENT device: a launcher that raises its kernel's dynamic shared memory
// limit must do so once per device, not once per process (a host may run one engine per GPU in one process).
struct DeviceOnce {
  std::atomic<unsigned long long> done{0};
  std::mutex mu;
  // Runs f() (the cudaFuncSetAttribute calls) once per device, with that device current.  The device's bit is
  // published only AFTER f() has succeeded, and under a mutex: a second host thread making its first launch on the
  // same device either sees the bit (attribute set) or waits for the first thread to finish setting it.
  template <class F>
  cudaError_t run(F &&f) {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d > 63) return f();
    const unsigned long long bit = 1ull << d;
    if (done.load(std::memory_order_acquire) & bit) return cudaSuccess;
    std::lock_guard<std::mutex> g(mu);
    if (done.load(std::memory_order_relaxed) & bit) return cudaSuccess;
    const cudaError_t e = f();
    if (e == cudaSuccess) done.fetch_or(bit, std::memory_order_release);
    return e;
  }
};

// ---- bf16 helpers -----------------------------------------------------------------------------
__host__ __device__ inline float bf16_bits_to_float(uint16_t b) {
  uint32_t u = (uint32_t)b << 16;
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}

__host__ __device__ inline uint16_t float_to_bf16_bits(float f) {  // round-nearest-even
  uint32_t u;
#ifdef __CUDA_ARCH__
  u = __float_as_uint(f);
#else
  memcpy(&u, &f, 4);
#endif
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40u);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

#ifdef __CUDACC__
// two packed bf16 (low = element 0) -> two floats
__device__ __forceinline__ float2 bf16x2_to_float2(uint32_t v) {
  return make_float2(__uint_as_float(v << 16), __uint_as_float(v & 0xffff0000u));
}
__device__ __forceinline__ uint32_t float2_to_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);  // .x = a (low half)
  return *reinterpret_cast<uint32_t *>(&h);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// streaming 16-byte load that does not allocate in L1 (weights / KV are read once)
__device__ __forceinline__ uint4 ld_stream_16(const void *p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// ---- counter-based hash shared by the synthetic-weight generator and the sampler --------------
__host__ __device__ inline uint64_t mix64(uint64_t seed, uint64_t key, uint64_t idx) {
  uint64_t z = seed * 0x9E3779B97F4A7C15ull + key * 0xD1B54A32D192ED03ull + idx + 0x632BE59BD9B4E019ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__host__ __device__ inline float hash_u01(uint64_t h) { return (float)(h >> 40) * (1.0f / 16777216.0f); }
#endif

}  // namespace dis
