// k_tp.cu -- tensor-parallel exchange fused with the compute around it, over NVLink peer memory
// (no reference counterpart: the reference specifies replicas only, requirements.md:100-110).
//
//   tp_add_rmsnorm : split-K reduce of this rank's O/down projection slices
//                    -> row PUSHED into every peer's symmetric buffer over NVLink (P2P stores), then a
//                       release flag (system scope) per (source rank, row)
//                    -> each rank polls its OWN memory for the peers' flags and reads the peers' rows locally
//                    -> summed IN RANK ORDER (every rank computes bit-identical results)
//                    -> residual add + RMSNorm -> bf16 GEMM input.        One kernel, one CTA per token row.
//   tp_argmax_merge: vocab-parallel greedy sampling: (max, id) candidates exchanged the same way.
//
// Three interchangeable kernels (same additions in the same order, bit-identical results; TpPeers.two_shot picks):
//   tp_add_rmsnorm_kernel     one-shot push, one flag hop                       (tp-1)   * T*H*4 bytes per rank
//   tp_add_rmsnorm2_kernel    reduce-scatter + all-gather of column blocks      2(tp-1)/tp * T*H*4, two flag hops
//   tp_add_rmsnorm_ll_kernel  the same two hops with the flag inside the data   4(tp-1)/tp * T*H*4, no fence, no flag
//                             (round 1 default: 3.28 / 2.64 / 2.42 ms per step at TP = 2 / 4 / 8, batch 64)
//   tp_add_rmsnorm_wide_kernel  the ll protocol on a (row, column block) grid that covers every SM (see below)
//
// The first is a one-shot all-reduce: every rank sends (tp-1) * T * H * 4 bytes; no NVLink round trip is on the
// critical path (stores are fire-and-forget, polling and reads are local).  Synchronisation is per ROW (a
// CTA only waits for the same row of each peer), so there is no grid-wide barrier and no CTA ever waits for
// another CTA of its own GPU.  Buffers are double-buffered by all-reduce parity: a rank can only reach
// all-reduce k+2 after every peer has finished all-reduce k (it needed their k+1 rows).
#include "kernels.h"
#include "ptx.cuh"

namespace dis {

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Every wait on a peer is bounded: a dead peer, or a rank whose step counter has drifted, must fail the launch (the
// engine then reports unhealthy) instead of hanging the GPU.  2^27 polls of an NVLink-visible word are tens of
// seconds; no legitimate wait in these kernels is longer than milliseconds.
__device__ __forceinline__ void wait_flag_sys(const uint32_t *f, uint32_t epoch) {
  long long spins = 0;
  while (ld_acquire_sys(f) != epoch)
    if (++spins > (1ll << 27)) __trap();
}
__device__ __forceinline__ float4 ld_volatile_f4(const float *p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

// explicit fma chain: both exchange kernels must round the sum of squares identically (left to the compiler, the
// contraction of x*x + y*y + ... differed between them and flipped a bf16 rounding once in ~1e5 elements)
__device__ __forceinline__ float sumsq4(float ss, const float4 &a) {
  ss = fmaf(a.x, a.x, ss); ss = fmaf(a.y, a.y, ss); ss = fmaf(a.z, a.z, ss); ss = fmaf(a.w, a.w, ss);
  return ss;
}

template <int kThreadsUnused = 0>
__device__ __forceinline__ float tp_block_sum(float v, float *sm) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int nw = (int)(blockDim.x >> 5);
  __syncthreads();
  if (l == 0) sm[w] = v;
  __syncthreads();
  float r = (l < nw) ? sm[l] : 0.f;
  return warp_sum(r);
}

__global__ void __launch_bounds__(1024)
tp_add_rmsnorm_kernel(TpPeers peers, float *__restrict__ x, const float *__restrict__ partial, int n_partial,
                      long long partial_stride, const float *__restrict__ gain, uint16_t *__restrict__ xn,
                      const uint32_t *__restrict__ step_ptr, uint32_t idx, int parity, int H, float eps) {
  __shared__ float red[32];
  extern __shared__ float row[];  // [2][H] (own contribution, running total) when a thread owns > 1 float4
  const int t = blockIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  const uint32_t epoch = *step_ptr * 256u + idx;
  const int step = blockDim.x * 4;
  const bool single = H <= step;
  const int me = peers.rank, tp = peers.tp;
  // slot of source rank r inside a rank's buffer: [parity][r][row t]
  const size_t slot_me = (((size_t)parity * tp + me) * peers.cap_rows + t) * (size_t)H;
  const size_t flag_me = ((size_t)parity * tp + me) * peers.cap_rows + t;

  // ---- local split-K reduce, pushed to every peer
  float4 keep = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = threadIdx.x * 4; i < H; i += step) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    const float *pp = partial + (size_t)t * H + i;
#pragma unroll 4
    for (int s = 0; s < n_partial; s++) {
      const float4 p = *reinterpret_cast<const float4 *>(pp + (size_t)s * partial_stride);
      v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
    }
    for (int r = 0; r < tp; r++)
      if (r != me) *reinterpret_cast<float4 *>(peers.data[r] + slot_me + i) = v;  // NVLink store
    if (single) keep = v; else *reinterpret_cast<float4 *>(row + i) = v;
  }
  __syncthreads();  // the release below is cumulative over the stores the barrier ordered before it
  if (threadIdx.x < tp && threadIdx.x != me) st_release_sys(peers.flags[threadIdx.x] + flag_me, epoch);
  // ---- wait for every peer's row (flags live in THIS rank's memory), one polling thread per source
  if (threadIdx.x < tp && threadIdx.x != me) {
    const uint32_t *f = peers.flags[me] + ((size_t)parity * tp + threadIdx.x) * peers.cap_rows + t;
    wait_flag_sys(f, epoch);
  }
  __syncthreads();
  // ---- sum in rank order, add the residual
  float ss = 0.f;
  for (int i = threadIdx.x * 4; i < H; i += step) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < tp; r++) {
      float4 v;
      if (r == me) v = single ? keep : *reinterpret_cast<const float4 *>(row + i);
      else v = ld_volatile_f4(peers.data[me] + (((size_t)parity * tp + r) * peers.cap_rows + t) * (size_t)H + i);
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    const float4 xv = *reinterpret_cast<const float4 *>(x + (size_t)t * H + i);
    a.x += xv.x; a.y += xv.y; a.z += xv.z; a.w += xv.w;
    *reinterpret_cast<float4 *>(x + (size_t)t * H + i) = a;
    if (single) keep = a; else *reinterpret_cast<float4 *>(row + H + i) = a;
    ss = sumsq4(ss, a);
  }
  ss = tp_block_sum(ss, red);
  const float inv = rsqrtf(ss / (float)H + eps);
  for (int i = threadIdx.x * 4; i < H; i += step) {
    const float4 v = single ? keep : *reinterpret_cast<const float4 *>(row + H + i);
    const float4 g = *reinterpret_cast<const float4 *>(gain + i);
    uint2 o;
    o.x = float2_to_bf16x2(v.x * inv * g.x, v.y * inv * g.y);
    o.y = float2_to_bf16x2(v.z * inv * g.z, v.w * inv * g.w);
    *reinterpret_cast<uint2 *>(xn + (size_t)t * H + i) = o;
  }
}

// Two-shot variant (tp >= 4): rank o OWNS columns [o*H/tp, (o+1)*H/tp) of every row.
//   hop 1 (reduce-scatter): each rank sums its split-K slices and pushes, per row, the owner's column block
//          into the owner's RS area; flag1 per (source, row).
//   the owner adds the tp contributions IN RANK ORDER and then the residual -- the same additions in the same
//          order as the one-shot kernel, so both variants produce identical bits --
//   hop 2 (all-gather): and pushes the finished block of x into every peer's AG area; flag2 per (owner, row).
//   every rank then holds the full new row and runs the RMSNorm.
// Bytes sent per rank: 2*(tp-1)/tp*T*H*4 instead of (tp-1)*T*H*4 -- 4x less at tp = 8 -- for one more flag hop.
//   RS area: data + [parity][source][cap_rows][H/tp]       AG area: data + 2*cap_rows*H + [parity][owner][cap_rows][H/tp]
//   flag1 = flags + [parity][source][cap_rows]              flag2 = flags + 2*tp*cap_rows + [parity][owner][cap_rows]
__global__ void __launch_bounds__(1024)
tp_add_rmsnorm2_kernel(TpPeers peers, float *__restrict__ x, const float *__restrict__ partial, int n_partial,
                       long long partial_stride, const float *__restrict__ gain, uint16_t *__restrict__ xn,
                       const uint32_t *__restrict__ step_ptr, uint32_t idx, int parity, int H, float eps) {
  __shared__ float red[32];
  extern __shared__ float row[];  // [H]: this rank's own column block, later the whole row when a thread owns > 1 float4
  const int t = blockIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  const uint32_t epoch = *step_ptr * 256u + idx;
  const int step = blockDim.x * 4;
  const bool single = H <= step;
  const int me = peers.rank, tp = peers.tp, cap = peers.cap_rows;
  const int Hs = H / tp;
  const size_t ag_base = (size_t)2 * cap * H;
  const size_t fl2 = (size_t)2 * tp * cap;
  const size_t blk_me = (((size_t)parity * tp + me) * cap + t) * (size_t)Hs;  // block written by this rank, row t
  const size_t flag_me = ((size_t)parity * tp + me) * cap + t;

  // ---- hop 1: split-K reduce, each column block goes to its owner
  float4 keep = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = threadIdx.x * 4; i < H; i += step) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    const float *pp = partial + (size_t)t * H + i;
#pragma unroll 4
    for (int s = 0; s < n_partial; s++) {
      const float4 p = *reinterpret_cast<const float4 *>(pp + (size_t)s * partial_stride);
      v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
    }
    const int owner = i / Hs;
    if (owner != me) *reinterpret_cast<float4 *>(peers.data[owner] + blk_me + (i - owner * Hs)) = v;  // NVLink store
    else *reinterpret_cast<float4 *>(row + i) = v;
  }
  __syncthreads();
  if (threadIdx.x < tp && threadIdx.x != me) {
    st_release_sys(peers.flags[threadIdx.x] + flag_me, epoch);
    const uint32_t *f = peers.flags[me] + ((size_t)parity * tp + threadIdx.x) * cap + t;
    wait_flag_sys(f, epoch);
  }
  __syncthreads();
  // ---- owner: sum in rank order, add the residual, hop 2: the finished block goes to everybody
  for (int i = me * Hs + threadIdx.x * 4; i < (me + 1) * Hs; i += step) {
    const int j = i - me * Hs;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < tp; r++) {
      float4 v;
      if (r == me) v = *reinterpret_cast<const float4 *>(row + i);
      else v = ld_volatile_f4(peers.data[me] + (((size_t)parity * tp + r) * cap + t) * (size_t)Hs + j);
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    const float4 xv = *reinterpret_cast<const float4 *>(x + (size_t)t * H + i);
    a.x += xv.x; a.y += xv.y; a.z += xv.z; a.w += xv.w;
    for (int r = 0; r < tp; r++)
      if (r != me) *reinterpret_cast<float4 *>(peers.data[r] + ag_base + blk_me + j) = a;  // NVLink store
    *reinterpret_cast<float4 *>(row + i) = a;
  }
  __syncthreads();
  if (threadIdx.x < tp && threadIdx.x != me) {
    st_release_sys(peers.flags[threadIdx.x] + fl2 + flag_me, epoch);
    const uint32_t *f = peers.flags[me] + fl2 + ((size_t)parity * tp + threadIdx.x) * cap + t;
    wait_flag_sys(f, epoch);
  }
  __syncthreads();
  // ---- full row: own block from shared memory, the others from the AG area; RMSNorm
  float ss = 0.f;
  for (int i = threadIdx.x * 4; i < H; i += step) {
    const int owner = i / Hs;
    float4 a;
    if (owner == me) a = *reinterpret_cast<const float4 *>(row + i);
    else {
      a = ld_volatile_f4(peers.data[me] + ag_base + (((size_t)parity * tp + owner) * cap + t) * (size_t)Hs + (i - owner * Hs));
      if (!single) *reinterpret_cast<float4 *>(row + i) = a;
    }
    *reinterpret_cast<float4 *>(x + (size_t)t * H + i) = a;
    if (single) keep = a;
    ss = sumsq4(ss, a);
  }
  ss = tp_block_sum(ss, red);
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

// Two-shot exchange with the flag INSIDE the data ("LL" protocol): every pair of fp32 values travels as one 16-byte
// store {v0, epoch, v1, epoch}.  An aligned 8-byte half is written atomically, so a receiver that sees the epoch
// of this exchange in a half also sees its value: no release fence (an NVLink round trip per hop in the kernels
// above), no separate flag store, no flag polling thread, no __syncthreads between the hops -- every thread spins on
// exactly the words it consumes.  Twice the bytes per hop, which the two-shot schedule can afford
// (4*(tp-1)/tp*T*H*4 per rank).  Same additions in the same order as the other two kernels: bit-identical results.
//   RS area: data + [parity][source][cap_rows][H/tp] pairs      AG area: data + 4*cap_rows*H + the same shape
// (a pair = 2 floats: value, epoch; areas are addressed in floats, so every index below is doubled)
__device__ __forceinline__ void st_ll(float *p, float v0, float v1, uint32_t e) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(__float_as_uint(v0)), "r"(e), "r"(__float_as_uint(v1)), "r"(e) : "memory");
}
__device__ __forceinline__ void ld_ll(const float *p, uint32_t e, float &v0, float &v1) {
  uint32_t a, fa, b, fb;
  long long spins = 0;
  do {
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(fa), "=r"(b), "=r"(fb) : "l"(p) : "memory");
    if (++spins > (1ll << 27)) __trap();  // a peer died or the protocol is broken: fail the launch, do not hang the GPU
  } while (fa != e || fb != e);
  v0 = __uint_as_float(a); v1 = __uint_as_float(b);
}

__global__ void __launch_bounds__(1024)
tp_add_rmsnorm_ll_kernel(TpPeers peers, float *__restrict__ x, const float *__restrict__ partial, int n_partial,
                         long long partial_stride, const float *__restrict__ gain, uint16_t *__restrict__ xn,
                         const uint32_t *__restrict__ step_ptr, uint32_t idx, int parity, int H, float eps) {
  __shared__ float red[32];
  extern __shared__ float row[];  // [H]: every element is written and read back by the same thread until the reduction
  const int t = blockIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  const uint32_t epoch = *step_ptr * 256u + idx;
  const int step = blockDim.x * 4;
  const int me = peers.rank, tp = peers.tp, cap = peers.cap_rows;
  const int Hs = H / tp;
  const size_t ag_base = (size_t)4 * cap * H;
  const size_t blk_me = (((size_t)parity * tp + me) * cap + t) * (size_t)Hs * 2;  // block written by this rank, row t

  // ---- hop 1: split-K reduce, each column block goes to its owner
  for (int i = threadIdx.x * 4; i < H; i += step) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    const float *pp = partial + (size_t)t * H + i;
#pragma unroll 4
    for (int s = 0; s < n_partial; s++) {
      const float4 p = *reinterpret_cast<const float4 *>(pp + (size_t)s * partial_stride);
      v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
    }
    const int owner = i / Hs;
    if (owner != me) {
      float *dst = peers.data[owner] + blk_me + (size_t)(i - owner * Hs) * 2;
      st_ll(dst, v.x, v.y, epoch);
      st_ll(dst + 4, v.z, v.w, epoch);
    } else {
      *reinterpret_cast<float4 *>(row + i) = v;
    }
  }
  // ---- owner (the same threads that kept their block): sum in rank order, add the residual, hop 2
  for (int i = threadIdx.x * 4; i < H; i += step) {
    if (i / Hs != me) continue;
    const int j = i - me * Hs;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < tp; r++) {
      float4 v;
      if (r == me) v = *reinterpret_cast<const float4 *>(row + i);
      else {
        const float *src = peers.data[me] + (((size_t)parity * tp + r) * cap + t) * (size_t)Hs * 2 + (size_t)j * 2;
        ld_ll(src, epoch, v.x, v.y);
        ld_ll(src + 4, epoch, v.z, v.w);
      }
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    const float4 xv = *reinterpret_cast<const float4 *>(x + (size_t)t * H + i);
    a.x += xv.x; a.y += xv.y; a.z += xv.z; a.w += xv.w;
    for (int r = 0; r < tp; r++)
      if (r != me) {
        float *dst = peers.data[r] + ag_base + blk_me + (size_t)j * 2;
        st_ll(dst, a.x, a.y, epoch);
        st_ll(dst + 4, a.z, a.w, epoch);
      }
    *reinterpret_cast<float4 *>(row + i) = a;
  }
  // ---- full row: own block from shared memory, the others as they arrive; RMSNorm
  float ss = 0.f;
  for (int i = threadIdx.x * 4; i < H; i += step) {
    const int owner = i / Hs;
    float4 a;
    if (owner == me) a = *reinterpret_cast<const float4 *>(row + i);
    else {
      const float *src = peers.data[me] + ag_base + (((size_t)parity * tp + owner) * cap + t) * (size_t)Hs * 2 + (size_t)(i - owner * Hs) * 2;
      ld_ll(src, epoch, a.x, a.y);
      ld_ll(src + 4, epoch, a.z, a.w);
      *reinterpret_cast<float4 *>(row + i) = a;
    }
    *reinterpret_cast<float4 *>(x + (size_t)t * H + i) = a;
    ss = sumsq4(ss, a);
  }
  ss = tp_block_sum(ss, red);
  const float inv = rsqrtf(ss / (float)H + eps);
  for (int i = threadIdx.x * 4; i < H; i += step) {
    const float4 v = *reinterpret_cast<const float4 *>(row + i);
    const float4 g = *reinterpret_cast<const float4 *>(gain + i);
    uint2 o;
    o.x = float2_to_bf16x2(v.x * inv * g.x, v.y * inv * g.y);
    o.y = float2_to_bf16x2(v.z * inv * g.z, v.w * inv * g.w);
    *reinterpret_cast<uint2 *>(xn + (size_t)t * H + i) = o;
  }
}

// Wide-grid variant of the flag-in-data exchange (two_shot == 3; the default for hidden sizes above 4096, see the
// measurements in Engine::tp_import): the same two hops, but
// one CTA per (row, column block) instead of one per row, so that at TP=8 a 64-row exchange runs on 512 small CTAs
// spread over every SM instead of 64 CTAs of 1024 threads that first walk 8-9 split-K slices of a whole row.
//   CTA (t, c), c = the owner rank of columns [c*Hs, (c+1)*Hs), one float4 per thread:
//     hop 1   sum the split-K slices of its 4 columns; c != me: send them to rank c's RS area, flag in the data
//     c == me sum the tp contributions IN RANK ORDER + residual, send the finished block to every peer's AG area (hop 2)
//     c != me take the finished block from the AG area as it arrives
//     all     reduce the block's sum of squares and publish it in THIS GPU's memory (one 8-byte {value, epoch} word per
//             (block, row) in the area the fenced kernels use for their flags); total = the tp block sums added in
//             block order (every rank, every CTA: same bits); RMSNorm of its 4 columns.
// The sum of squares is associated differently from the row kernels (per block, then over blocks), so this variant
// is deterministic and identical on every rank but NOT bit-identical to the other three.
// A CTA waits for the block sums of the other CTAs of its row on its own GPU, and a peer's owner CTA waits for the hop-1
// data of every CTA here: the launcher only uses this kernel when the whole grid is co-resident.
__device__ __forceinline__ void st_ll1(uint32_t *p, float v, uint32_t e) {
  asm volatile("st.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(e) : "memory");
}
__device__ __forceinline__ float ld_ll1(const uint32_t *p, uint32_t e) {
  uint32_t a, f;
  long long spins = 0;
  do {
    asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(a), "=r"(f) : "l"(p) : "memory");
    if (++spins > (1ll << 27)) __trap();
  } while (f != e);
  return __uint_as_float(a);
}

__global__ void __launch_bounds__(1024)
tp_add_rmsnorm_wide_kernel(TpPeers peers, float *__restrict__ x, const float *__restrict__ partial, int n_partial,
                           long long partial_stride, const float *__restrict__ gain, uint16_t *__restrict__ xn,
                           const uint32_t *__restrict__ step_ptr, uint32_t idx, int parity, int H, float eps) {
  __shared__ float red[32];
  __shared__ float s_inv;
  const int t = blockIdx.x;
  const int me = peers.rank, tp = peers.tp, cap = peers.cap_rows;
  const int c = (int)((blockIdx.y + (unsigned)me) % (unsigned)tp);  // blockIdx.y == 0: this rank's own block (scheduled first)
  pdl_launch_dependents();
  pdl_wait();
  const uint32_t epoch = *step_ptr * 256u + idx;
  const int Hs = H / tp;
  const int j = threadIdx.x * 4;          // column inside the block
  const int col = c * Hs + j;
  const size_t ag_base = (size_t)4 * cap * H;
  const size_t blk = (((size_t)parity * tp + me) * cap + t) * (size_t)Hs * 2;  // block written by this rank, row t (pairs)

  // ---- hop 1: split-K reduce of these 4 columns
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  {
    const float *pp = partial + (size_t)t * H + col;
#pragma unroll 4
    for (int s = 0; s < n_partial; s++) {
      const float4 p = *reinterpret_cast<const float4 *>(pp + (size_t)s * partial_stride);
      v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
    }
  }
  float4 a;
  if (c != me) {
    float *dst = peers.data[c] + blk + (size_t)j * 2;
    st_ll(dst, v.x, v.y, epoch);
    st_ll(dst + 4, v.z, v.w, epoch);
    // the finished block comes back from its owner
    const float *src = peers.data[me] + ag_base + (((size_t)parity * tp + c) * cap + t) * (size_t)Hs * 2 + (size_t)j * 2;
    ld_ll(src, epoch, a.x, a.y);
    ld_ll(src + 4, epoch, a.z, a.w);
  } else {
    a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < tp; r++) {
      float4 w;
      if (r == me) w = v;
      else {
        const float *src = peers.data[me] + (((size_t)parity * tp + r) * cap + t) * (size_t)Hs * 2 + (size_t)j * 2;
        ld_ll(src, epoch, w.x, w.y);
        ld_ll(src + 4, epoch, w.z, w.w);
      }
      a.x += w.x; a.y += w.y; a.z += w.z; a.w += w.w;
    }
    const float4 xv = *reinterpret_cast<const float4 *>(x + (size_t)t * H + col);
    a.x += xv.x; a.y += xv.y; a.z += xv.z; a.w += xv.w;
    for (int r = 0; r < tp; r++)
      if (r != me) {
        float *dst = peers.data[r] + ag_base + blk + (size_t)j * 2;
        st_ll(dst, a.x, a.y, epoch);
        st_ll(dst + 4, a.z, a.w, epoch);
      }
  }
  *reinterpret_cast<float4 *>(x + (size_t)t * H + col) = a;
  // ---- sum of squares of the row: every CTA reduces ITS block (the finished values are the same bits on every rank,
  // so no NVLink traffic is needed for this), publishes the block sum in this GPU's own memory as one 8-byte
  // {value, epoch} word, and adds the tp block sums of its row in block order
  const float ssb = tp_block_sum(sumsq4(0.f, a), red);
  if (threadIdx.x == 0) {
    uint32_t *mine = peers.flags[me] + (((size_t)parity * tp + c) * cap + t) * 2;
    st_ll1(mine, ssb, epoch);
    float ss = 0.f;
    for (int o = 0; o < tp; o++) ss += o == c ? ssb : ld_ll1(peers.flags[me] + (((size_t)parity * tp + o) * cap + t) * 2, epoch);
    s_inv = rsqrtf(ss / (float)H + eps);
  }
  __syncthreads();
  const float inv = s_inv;
  const float4 g = *reinterpret_cast<const float4 *>(gain + col);
  uint2 o;
  o.x = float2_to_bf16x2(a.x * inv * g.x, a.y * inv * g.y);
  o.y = float2_to_bf16x2(a.z * inv * g.z, a.w * inv * g.w);
  *reinterpret_cast<uint2 *>(xn + (size_t)t * H + col) = o;
}

// whether the wide-grid kernel may be used for T rows: every CTA of the grid must be resident at once
static bool tp_wide_fits(const TpPeers &peers, int T, int H) {
  const int Hs = H / peers.tp;
  if (H % peers.tp != 0 || Hs % 128 != 0 || Hs / 4 > 1024) return false;
  static int per_sm_cache[33] = {0};  // by threads / 32
  const int threads = Hs / 4;
  int &per_sm = per_sm_cache[threads / 32];
  if (per_sm == 0) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, tp_add_rmsnorm_wide_kernel, threads, 0) != cudaSuccess || n <= 0) { cudaGetLastError(); return false; }
    per_sm = n;
  }
  int dev = 0, sms = kNumSMs;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return (long long)T * peers.tp <= (long long)per_sm * sms;
}

// Streaming variant for prefill-sized passes (two_shot == 4): the exchange of thousands of rows is bandwidth-bound, so
// instead of one CTA per row with two release fences each (an NVLink round trip per fence: 8192 rows cost 451 ms of TTFT
// against NCCL's 323 at TP=2), a grid of resident CTAs -- one 1024-thread CTA per SM -- each takes every G-th row and goes through the
// two-shot schedule ONCE for all of its rows:
//   phase A  for each of its rows: sum the split-K slices, store each column block into its owner's RS area (plain
//            16-byte stores over NVLink; the own block into the local RS area); then ONE barrier + release flag per peer
//   phase B  wait for the peers' phase-A flags of the same CTA index (every rank walks the same rows in the same CTA);
//            for each row: the tp contributions of the own block IN RANK ORDER + residual -> x, and into every peer's
//            AG area; then one barrier + release flag per peer
//   phase C  wait for the phase-B flags; for each row: the finished blocks out of the AG area, x, RMSNorm -> bf16.
// Bytes on the wire 2 (tp-1)/tp * T*H*4 as in the fenced two-shot kernel, the same additions in the same order (bit-
// identical to it), two flag flights per CTA instead of per row.  Flags: peers.flags[rank][phase][parity][source][CTA].
__global__ void __launch_bounds__(1024, 1)
tp_add_rmsnorm_stream_kernel(TpPeers peers, float *__restrict__ x, const float *__restrict__ partial, int n_partial,
                             long long partial_stride, const float *__restrict__ gain, uint16_t *__restrict__ xn,
                             const uint32_t *__restrict__ step_ptr, uint32_t idx, int parity, int T, int H, float eps) {
  __shared__ float red[32];
  pdl_launch_dependents();
  pdl_wait();
  const uint32_t epoch = *step_ptr * 256u + idx;
  const int me = peers.rank, tp = peers.tp, cap = peers.cap_rows;
  const int Hs = H / tp, G = (int)gridDim.x, c = (int)blockIdx.x;
  const size_t ag_base = (size_t)2 * cap * H;
  auto rs_slot = [&](int src, int t) { return (((size_t)parity * tp + src) * cap + t) * (size_t)Hs; };
  auto flag_at = [&](int phase, int src) { return (((size_t)phase * 2 + parity) * tp + src) * (size_t)cap + c; };
  const int step = blockDim.x * 4;

  // ---- phase A
  for (int t = c; t < T; t += G) {
    for (int i = threadIdx.x * 4; i < H; i += step) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      const float *pp = partial + (size_t)t * H + i;
#pragma unroll 4
      for (int s = 0; s < n_partial; s++) {
        const float4 p = *reinterpret_cast<const float4 *>(pp + (size_t)s * partial_stride);
        v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
      }
      const int owner = i / Hs;
      *reinterpret_cast<float4 *>(peers.data[owner] + rs_slot(me, t) + (i - owner * Hs)) = v;  // NVLink store unless owner == me
    }
  }
  __syncthreads();  // the release below is cumulative over the stores the barrier ordered before it
  if ((int)threadIdx.x < tp && (int)threadIdx.x != me) {
    st_release_sys(peers.flags[threadIdx.x] + flag_at(0, me), epoch);
    wait_flag_sys(peers.flags[me] + flag_at(0, threadIdx.x), epoch);
  }
  __syncthreads();
  // ---- phase B: (row, 4 columns of the own block) items over all threads
  {
    const int per_row = Hs / 4;
    const int rows_here = c < T ? (T - 1 - c) / G + 1 : 0;
    for (int item = threadIdx.x; item < rows_here * per_row; item += blockDim.x) {
      const int t = c + (item / per_row) * G, j = (item % per_row) * 4;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < tp; r++) {
        const float4 v = ld_volatile_f4(peers.data[me] + rs_slot(r, t) + j);
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
      }
      float *xp = x + (size_t)t * H + me * Hs + j;
      const float4 xv = *reinterpret_cast<const float4 *>(xp);
      a.x += xv.x; a.y += xv.y; a.z += xv.z; a.w += xv.w;
      *reinterpret_cast<float4 *>(xp) = a;
      for (int r = 0; r < tp; r++)
        if (r != me) *reinterpret_cast<float4 *>(peers.data[r] + ag_base + rs_slot(me, t) + j) = a;  // NVLink store
    }
  }
  __syncthreads();
  if ((int)threadIdx.x < tp && (int)threadIdx.x != me) {
    st_release_sys(peers.flags[threadIdx.x] + flag_at(1, me), epoch);
    wait_flag_sys(peers.flags[me] + flag_at(1, threadIdx.x), epoch);
  }
  __syncthreads();
  // ---- phase C
  extern __shared__ float row[];  // [H] when a thread owns more than one float4
  const bool single = H <= step;
  for (int t = c; t < T; t += G) {
    float ss = 0.f;
    float4 keep = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = threadIdx.x * 4; i < H; i += step) {
      const int owner = i / Hs;
      float4 a;
      if (owner == me) a = *reinterpret_cast<const float4 *>(x + (size_t)t * H + i);
      else {
        a = ld_volatile_f4(peers.data[me] + ag_base + rs_slot(owner, t) + (i - owner * Hs));
        *reinterpret_cast<float4 *>(x + (size_t)t * H + i) = a;
      }
      if (single) keep = a; else *reinterpret_cast<float4 *>(row + i) = a;
      ss = sumsq4(ss, a);
    }
    ss = tp_block_sum(ss, red);
    const float inv = rsqrtf(ss / (float)H + eps);
    for (int i = threadIdx.x * 4; i < H; i += step) {
      const float4 v = single ? keep : *reinterpret_cast<const float4 *>(row + i);
      const float4 g = *reinterpret_cast<const float4 *>(gain + i);
      uint2 o;
      o.x = float2_to_bf16x2(v.x * inv * g.x, v.y * inv * g.y);
      o.y = float2_to_bf16x2(v.z * inv * g.z, v.w * inv * g.w);
      *reinterpret_cast<uint2 *>(xn + (size_t)t * H + i) = o;
    }
    __syncthreads();  // red[] and row[] are reused by the next row
  }
}

cudaError_t launch_tp_add_rmsnorm(const TpPeers &peers, float *x, const float *partial, int n_partial,
                                  long long partial_stride, const float *gain, uint16_t *xn, const uint32_t *step_ptr,
                                  uint32_t idx, int parity, int T, int H, float eps, cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  if (T > peers.cap_rows) { set_last_error("tp_add_rmsnorm: more rows than the symmetric buffer holds"); return cudaErrorInvalidValue; }
  if (H % 4 != 0) { set_last_error("tp_add_rmsnorm: H must be a multiple of 4"); return cudaErrorInvalidValue; }
  int threads = (H / 4 + 31) / 32 * 32;
  threads = threads > 1024 ? 1024 : threads;
  if (peers.two_shot == 4) {
    if (H % (4 * peers.tp) != 0) { set_last_error("tp_add_rmsnorm: the streaming exchange needs H % (4 * tp) == 0"); return cudaErrorInvalidValue; }
    int dev = 0, sms = kNumSMs;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = T < sms ? T : sms;  // every CTA must be resident: a CTA waits for the same CTA index on its peers
    const size_t smem4 = H <= threads * 4 ? 0 : (size_t)H * sizeof(float);
    if (smem4 > 48 * 1024) DIS_CUDA_TRY(cudaFuncSetAttribute(tp_add_rmsnorm_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4));
    return launch_k(tp_add_rmsnorm_stream_kernel, dim3(grid), dim3(threads), smem4, st, peers, x, partial, n_partial, partial_stride,
                    gain, xn, step_ptr, idx, parity, T, H, eps);
  }
  if (peers.two_shot == 3 && tp_wide_fits(peers, T, H)) {
    return launch_k(tp_add_rmsnorm_wide_kernel, dim3(T, peers.tp), dim3(H / peers.tp / 4), 0, st, peers, x, partial, n_partial,
                    partial_stride, gain, xn, step_ptr, idx, parity, H, eps);
  }
  if (peers.two_shot >= 2) {
    if (H % (4 * peers.tp) != 0) { set_last_error("tp_add_rmsnorm: the LL exchange needs H % (4 * tp) == 0"); return cudaErrorInvalidValue; }
    const size_t smem2 = (size_t)H * sizeof(float);
    if (smem2 > 48 * 1024) DIS_CUDA_TRY(cudaFuncSetAttribute(tp_add_rmsnorm_ll_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    return launch_k(tp_add_rmsnorm_ll_kernel, dim3(T), dim3(threads), smem2, st, peers, x, partial, n_partial, partial_stride,
                    gain, xn, step_ptr, idx, parity, H, eps);
  }
  if (peers.two_shot) {
    if (H % (4 * peers.tp) != 0) { set_last_error("tp_add_rmsnorm: two-shot exchange needs H % (4 * tp) == 0"); return cudaErrorInvalidValue; }
    const size_t smem2 = (size_t)H * sizeof(float);
    if (smem2 > 48 * 1024) DIS_CUDA_TRY(cudaFuncSetAttribute(tp_add_rmsnorm2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    return launch_k(tp_add_rmsnorm2_kernel, dim3(T), dim3(threads), smem2, st, peers, x, partial, n_partial, partial_stride,
                    gain, xn, step_ptr, idx, parity, H, eps);
  }
  const size_t smem = H <= threads * 4 ? 0 : (size_t)2 * H * sizeof(float);
  if (smem > 48 * 1024) DIS_CUDA_TRY(cudaFuncSetAttribute(tp_add_rmsnorm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  return launch_k(tp_add_rmsnorm_kernel, dim3(T), dim3(threads), smem, st, peers, x, partial, n_partial, partial_stride, gain,
                  xn, step_ptr, idx, parity, H, eps);
}

// One CTA.  Candidates are pushed into every peer's cand area: [parity][source rank][cap_rows][4] words
// {max bits, token id, sum exp(z - max) over the source's vocabulary slice, unused}.
__global__ void __launch_bounds__(1024)
tp_argmax_merge_kernel(TpPeers peers, const float *__restrict__ val, const uint32_t *__restrict__ tok,
                       const float *__restrict__ sumexp, const uint32_t *__restrict__ step_ptr, int T,
                       uint32_t *__restrict__ out, float *__restrict__ out_logprob) {
  pdl_wait();
  const uint32_t epoch = *step_ptr * 256u + 255u;
  const int parity = (int)(*step_ptr & 1u);  // one merge per forward pass: consecutive passes alternate
  const int t = threadIdx.x, me = peers.rank, tp = peers.tp;
  const size_t base_me = ((size_t)parity * tp + me) * peers.cap_rows;
  if (t < T) {
    const uint4 c = make_uint4(__float_as_uint(val[t]), tok[t], __float_as_uint(sumexp ? sumexp[t] : 0.f), 0u);
    for (int r = 0; r < tp; r++)
      if (r != me) *reinterpret_cast<uint4 *>(peers.cand[r] + (base_me + t) * 4) = c;
  }
  __syncthreads();
  if (t == 0) __threadfence_system();
  __syncthreads();
  if (t < tp && t != me) {
    st_release_sys(peers.cand_flags[t] + parity * tp + me, epoch);
    wait_flag_sys(peers.cand_flags[me] + parity * tp + t, epoch);
  }
  __syncthreads();
  if (t < T) {
    float bv = -INFINITY;
    uint32_t bt = 0xffffffffu;
    float mv[kTpMaxRanks], sv[kTpMaxRanks];
#pragma unroll
    for (int r = 0; r < kTpMaxRanks; r++) {
      mv[r] = -INFINITY; sv[r] = 0.f;
      if (r < tp) {
        uint32_t vb, k, sb;
        if (r == me) { vb = __float_as_uint(val[t]); k = tok[t]; sb = __float_as_uint(sumexp ? sumexp[t] : 0.f); }
        else {
          const volatile uint32_t *c = peers.cand[me] + (((size_t)parity * tp + r) * peers.cap_rows + t) * 4;
          vb = c[0]; k = c[1]; sb = c[2];
        }
        const float v = __uint_as_float(vb);
        mv[r] = v; sv[r] = __uint_as_float(sb);
        if (v > bv || (v == bv && k < bt)) { bv = v; bt = k; }
      }
    }
    out[t] = bt;
    if (out_logprob) {  // rank order: every rank computes the same bits
      float se = 0.f;
#pragma unroll
      for (int r = 0; r < kTpMaxRanks; r++)
        if (r < tp) se += sv[r] * expf(mv[r] - bv);
      out_logprob[t] = -logf(se);
    }
  }
}

cudaError_t launch_tp_argmax_merge(const TpPeers &peers, const float *val, const uint32_t *tok, const float *sumexp,
                                   const uint32_t *step_ptr, int T, uint32_t *out, float *out_logprob, cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  if (T > 1024 || T > peers.cap_rows) { set_last_error("tp_argmax_merge: too many rows"); return cudaErrorInvalidValue; }
  return launch_k(tp_argmax_merge_kernel, dim3(1), dim3((T + 31) / 32 * 32), 0, st, peers, val, tok, sumexp, step_ptr, T, out,
                  out_logprob);
}

}  // namespace dis
