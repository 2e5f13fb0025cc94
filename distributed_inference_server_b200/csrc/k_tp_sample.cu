// k_tp_sample.cu -- vocab-parallel sampling over NVLink peer memory (no reference counterpart: the reference
// specifies replicas only, requirements.md:100-110; the sampling controls are InferenceParams, design.md:666-671).
//
// Under tensor parallelism a rank holds V/tp columns of the logits.  Greedy rows need one exchange of (max, id,
// sum-exp) candidates (tp_argmax_merge_kernel, k_tp.cu).  Temperature / nucleus rows used to all-gather the full
// [R][V] fp32 logits through NCCL (16 MB at batch 32) and run the one-GPU sampler on every rank; this kernel keeps
// the logits where they are and exchanges a handful of scalars per row instead:
//
//   round 0      local (max, argmax id, sum exp(z - local max))            -> global max m, normaliser Z, greedy rows done
//   rounds 1..   16-ary search for the nucleus threshold tau: every rank sums its own p_i >= thr_j for 16 equally spaced
//                thresholds in [lo, hi) (one pass: a bucket per element, masses = suffix sums of the buckets), the
//                16 x tp partial masses are exchanged and added IN RANK ORDER (every rank computes the same bits and
//                narrows [lo, hi) identically); 2^30 bit patterns -> 8 rounds.  Same definition as sample_kernel
//                (k_elementwise.cu): the largest threshold whose kept mass reaches top_p.
//   draw         the ranks' kept masses are known from the last round: u * kept falls into exactly one rank's
//                slice (slices are in vocabulary order); that rank walks its kept tokens as sample_kernel does and
//   last round   broadcasts (token id, z of the token); every rank writes the same token and log-probability.
//
// One CTA per row; a CTA waits only for the same row's CTAs on the peers.  Every value travels as an 8-byte
// {bits, epoch} word written with one aligned store (atomic), so a reader that sees this call's epoch has the value:
// no fence, no separate flag (the protocol of tp_add_rmsnorm_ll_kernel).  Each round has its own slot, so nothing is
// overwritten inside a call; a rank can only start the next call's round k after every peer has finished this call
// (it needs their round k-1 values), which is what makes reusing the slots across calls safe.
//
// rank < 0 in TpSampPeers emulates the tp ranks on ONE GPU: grid (rows, tp), rank = blockIdx.y, every "peer" area in
// this GPU's memory.  The kernel-level parity test (tests/test_kernels_gpu.py) drives that mode against the oracle
// and against sample_kernel on a box with one GPU; the engine uses the real mode.
#include "kernels.h"
#include "ptx.cuh"

namespace dis {

namespace {

constexpr int kThreads = 1024;
constexpr int kVals = kTpSampleVals;          // values per rank and round
constexpr int kSearchRounds = kTpSampleRounds - 2;  // round 0: statistics, last round: the drawn token

struct Cand { float v; int i; };
__device__ __forceinline__ Cand better(Cand a, Cand b) { return (b.v > a.v || (b.v == a.v && b.i < a.i)) ? b : a; }

__device__ __forceinline__ void st_word(uint32_t *p, uint32_t bits, uint32_t epoch) {
  asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(bits), "r"(epoch) : "memory");
}
__device__ __forceinline__ uint32_t ld_word(const uint32_t *p, uint32_t epoch) {
  uint32_t bits, e;
  long long spins = 0;
  do {
    asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(bits), "=r"(e) : "l"(p) : "memory");
    if (++spins > (1ll << 25)) __trap();  // (tens of seconds) a peer died or the step counters drifted: fail the launch, do not hang the GPU
  } while (e != epoch);
  return bits;
}

struct Shared {
  float red[32][kVals];
  uint32_t mine[kVals];
  float bucket[kVals];
  uint32_t all[kTpMaxRanks][kVals];
  float scan[32];
  Cand am[32];
  int tok, last;
};

// All-gather of n values per rank for this row: mine[] -> all[r][] on every rank (own values copied locally).
// from_only >= 0: only that rank sends, the others only receive its values.
__device__ __forceinline__ void exchange(const TpSampPeers &sp, Shared &s, int me, int row, int round, int n, uint32_t epoch,
                                         int from_only = -1) {
  __syncthreads();  // mine[] is complete
  const int tp = sp.tp;
  const size_t slot0 = ((size_t)round * tp) * sp.cap_rows;
  for (int idx = threadIdx.x; idx < n * tp; idx += kThreads) {
    const int r = idx / n, j = idx - r * n;
    if (r == me) { if (from_only < 0 || from_only == me) s.all[me][j] = s.mine[j]; }
    else if (from_only < 0 || from_only == me)
      st_word(sp.buf[r] + ((slot0 + (size_t)me * sp.cap_rows + row) * kVals + j) * 2, s.mine[j], epoch);
  }
  for (int idx = threadIdx.x; idx < n * tp; idx += kThreads) {
    const int r = idx / n, j = idx - r * n;
    if (r != me && (from_only < 0 || from_only == r))
      s.all[r][j] = ld_word(sp.buf[me] + ((slot0 + (size_t)r * sp.cap_rows + row) * kVals + j) * 2, epoch);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kThreads)
tp_sample_kernel(TpSampPeers sp, const float *__restrict__ logits, long long ld, int Vl,
                 const float *__restrict__ temperature, const float *__restrict__ top_p,
                 const unsigned long long *__restrict__ seeds, const unsigned long long *__restrict__ steps,
                 const uint32_t *__restrict__ step_ptr, uint32_t step_imm, uint32_t *__restrict__ out,
                 float *__restrict__ out_logprob) {
  __shared__ Shared s;
  extern __shared__ float bins[];  // [kVals][kThreads]: a column per thread
  const int t = blockIdx.x, T = gridDim.x, tp = sp.tp;
  const bool emulated = sp.rank < 0;
  const int me = emulated ? (int)blockIdx.y : sp.rank;
  pdl_wait();
  const uint32_t call = step_ptr ? *step_ptr : step_imm;
  auto epoch = [&](int round) { return call * 16u + (uint32_t)round + 1u; };  // never 0, the value of an untouched word
  const float *row = logits + (size_t)t * ld + (emulated ? (size_t)me * Vl : 0);
  const int voff = me * Vl;
  uint32_t *o_tok = out + (emulated ? (size_t)me * T : 0) + t;
  float *o_lp = out_logprob ? out_logprob + (emulated ? (size_t)me * T : 0) + t : nullptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float temp = temperature ? temperature[t] : 0.f;
  const bool greedy = temp <= 0.f;
  const float inv_t = greedy ? 1.0f : 1.0f / temp;

  // ---- round 0: local argmax (lowest index wins ties), local sum of exp(z - local max), z = logit * inv_t
  Cand am{-INFINITY, 0x7fffffff};
  for (int i = threadIdx.x; i < Vl; i += kThreads) {
    const float v = row[i];
    if (v > am.v) { am.v = v; am.i = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    Cand b{__shfl_xor_sync(0xffffffffu, am.v, o), __shfl_xor_sync(0xffffffffu, am.i, o)};
    am = better(am, b);
  }
  if (lane == 0) s.am[warp] = am;
  __syncthreads();
  if (warp == 0) {
    am = s.am[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      Cand b{__shfl_xor_sync(0xffffffffu, am.v, o), __shfl_xor_sync(0xffffffffu, am.i, o)};
      am = better(am, b);
    }
    if (lane == 0) s.am[0] = am;
  }
  __syncthreads();
  am = s.am[0];
  const float m_loc = am.v * inv_t;
  float se = 0.f;
  for (int i = threadIdx.x; i < Vl; i += kThreads) se += expf(fmaf(row[i], inv_t, -m_loc));
  se = warp_sum(se);
  if (lane == 0) s.red[warp][0] = se;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
    for (int w = 0; w < kThreads / 32; w++) tot += s.red[w][0];  // warp order: deterministic
    s.mine[0] = __float_as_uint(m_loc);
    s.mine[1] = (uint32_t)(am.i + voff);
    s.mine[2] = __float_as_uint(tot);
    s.mine[3] = __float_as_uint(am.v);
  }
  exchange(sp, s, me, t, 0, 4, epoch(0));
  float m = -INFINITY;
  for (int r = 0; r < tp; r++) m = fmaxf(m, __uint_as_float(s.all[r][0]));
  float Z = 0.f;
  for (int r = 0; r < tp; r++) Z += __uint_as_float(s.all[r][2]) * expf(__uint_as_float(s.all[r][0]) - m);  // rank order
  if (greedy) {
    if (threadIdx.x == 0) {
      float bv = -INFINITY;
      uint32_t bt = 0xffffffffu;
      for (int r = 0; r < tp; r++) {
        const float v = __uint_as_float(s.all[r][3]);
        const uint32_t k = s.all[r][1];
        if (v > bv || (v == bv && k < bt)) { bv = v; bt = k; }
      }
      *o_tok = bt;
      if (o_lp) *o_lp = -logf(Z);  // the winner's z - m is 0
    }
    return;
  }

  // ---- nucleus threshold: 16-ary search over the bit patterns of positive floats (they are ordered).
  // A round's thresholds are thr_j = lo + j * 2^k with 16 * 2^k >= hi - lo, so an element's bucket is a shift:
  // b = min((bits - lo) >> k, 15).  Every thread adds its elements into its own column of bins[16][1024] in shared
  // memory (one add per element instead of sixteen compare-adds: the pass is then bound by the L2 read of the slice),
  // the columns are reduced in a fixed order, and mass(thr_j) = sum of buckets j..15 -- buckets are summed from the
  // top, so the masses cannot increase with j.  Bucket 15 also takes everything at or above lo + 15 * 2^k.
  const float inv_z = 1.0f / Z;
  const float tpv = top_p ? top_p[t] : 1.0f;
  const bool cut = tpv < 1.0f;  // top_p >= 1: no threshold, the one round only exchanges the slices' total masses
  uint32_t lo = 0u, hi = cut ? 0x3f800001u : 1u;  // invariant (cut): mass(lo) >= top_p > mass(hi), lo < hi
  int round = 1, jsel = 0;
  do {
    const uint32_t per = (hi - lo + kVals - 1) / kVals;  // >= 1
    const int k = per > 1u ? 32 - __clz(per - 1u) : 0;   // 2^k >= per
#pragma unroll
    for (int j = 0; j < kVals; j++) bins[j * kThreads + threadIdx.x] = 0.f;
    for (int i = threadIdx.x; i < Vl; i += kThreads) {
      const float p = expf(fmaf(row[i], inv_t, -m)) * inv_z;
      const uint32_t b = __float_as_uint(p);
      if (b >= lo) {
        const uint32_t q = (b - lo) >> k;
        bins[(q < (uint32_t)(kVals - 1) ? q : (uint32_t)(kVals - 1)) * kThreads + threadIdx.x] += p;  // own column: no race
      }
    }
#pragma unroll
    for (int j = 0; j < kVals; j++) {
      const float v = warp_sum(bins[j * kThreads + threadIdx.x]);
      if (lane == 0) s.red[warp][j] = v;
    }
    __syncthreads();
    if (threadIdx.x < kVals) {
      float tot = 0.f;
      for (int w = 0; w < kThreads / 32; w++) tot += s.red[w][threadIdx.x];  // warp order: deterministic
      s.bucket[threadIdx.x] = tot;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      float run = 0.f;
      for (int j = kVals - 1; j >= 0; j--) { run += s.bucket[j]; s.mine[j] = __float_as_uint(run); }
    }
    exchange(sp, s, me, t, round, kVals, epoch(round));
    jsel = 0;
    if (cut) {
      for (int j = 1; j < kVals; j++) {
        float mass = 0.f;
        for (int r = 0; r < tp; r++) mass += __uint_as_float(s.all[r][j]);  // rank order: the same bits on every rank
        // the last hit is the largest threshold that still reaches top_p; thresholds at or above hi are known to fall short
        if (mass >= tpv && (((uint32_t)j << k) >> k) == (uint32_t)j && lo + ((uint32_t)j << k) < hi) jsel = j;
      }
      const uint32_t next = lo + ((uint32_t)(jsel + 1) << k);
      if (jsel < kVals - 1 && next < hi) hi = next;
      lo += (uint32_t)jsel << k;
    }
    round++;
  } while (cut && hi - lo > 1u && round <= kSearchRounds);
  const uint32_t tau = lo;

  // ---- which rank's slice holds the draw
  float kept = 0.f;
  for (int r = 0; r < tp; r++) kept += __uint_as_float(s.all[r][jsel]);
  const float u = hash_u01(mix64(seeds ? seeds[t] : 0ull, 0xC0FFEEull, steps ? steps[t] : 0ull));
  const float target = u * kept;
  int owner = -1, last_nonempty = 0;
  float before = 0.f, before_owner = 0.f;
  for (int r = 0; r < tp; r++) {
    const float k = __uint_as_float(s.all[r][jsel]);
    if (k > 0.f) last_nonempty = r;
    if (owner < 0 && before + k > target) { owner = r; before_owner = before; }
    before += k;
  }
  if (owner < 0) {  // rounding left no interval: the last rank that kept anything, its last kept token
    owner = last_nonempty;
    before_owner = 0.f;
    for (int r = 0; r < owner; r++) before_owner += __uint_as_float(s.all[r][jsel]);
  }

  if (me == owner) {
    // first kept token of this slice, in index order, whose running mass exceeds the target (sample_kernel's walk)
    const float local_target = target - before_owner;
    const int chunk = (Vl + kThreads - 1) / kThreads;
    const int i0 = threadIdx.x * chunk, i1 = min(Vl, i0 + chunk);
    float local = 0.f;
    int my_last = -1;
    for (int i = i0; i < i1; i++) {
      const float p = expf(fmaf(row[i], inv_t, -m)) * inv_z;
      if (__float_as_uint(p) >= tau) { local += p; my_last = i; }
    }
    float incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float n = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += n;
    }
    if (threadIdx.x == 0) { s.tok = -1; s.last = -1; }
    if (lane == 31) s.scan[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const float w = s.scan[lane];
      float wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float n = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= o) wi += n;
      }
      s.scan[lane] = wi - w;
    }
    __syncthreads();
    const float excl = s.scan[warp] + incl - local;
    if (my_last >= 0) atomicMax(&s.last, my_last);
    if (local > 0.f && excl <= local_target && local_target < excl + local) {
      float run = excl;
      int tok = -1, lastk = -1;
      for (int i = i0; i < i1; i++) {
        const float p = expf(fmaf(row[i], inv_t, -m)) * inv_z;
        if (__float_as_uint(p) >= tau) { run += p; lastk = i; if (run > local_target) { tok = i; break; } }
      }
      if (tok < 0) tok = lastk;
      atomicMax(&s.tok, tok);  // the threads' intervals are disjoint: at most one gets here
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int tok = s.tok;
      if (tok < 0) tok = s.last >= 0 ? s.last : am.i;
      s.mine[0] = (uint32_t)(tok + voff);
      s.mine[1] = __float_as_uint(row[tok] * inv_t);
    }
  }
  exchange(sp, s, me, t, kTpSampleRounds - 1, 2, epoch(kTpSampleRounds - 1), owner);
  if (threadIdx.x == 0) {
    *o_tok = s.all[owner][0];
    // log-probability of the drawn token under softmax(logits / temperature), before the nucleus cut
    if (o_lp) *o_lp = __uint_as_float(s.all[owner][1]) - m - logf(Z);
  }
}

}  // namespace

size_t tp_sample_area_bytes(int tp, int cap_rows) {
  return (size_t)kTpSampleRounds * tp * cap_rows * kTpSampleVals * 8;
}

cudaError_t launch_tp_sample(const TpSampPeers &sp, const float *logits, long long ld, int T, int Vl,
                             const float *temperature, const float *top_p, const unsigned long long *seeds,
                             const unsigned long long *steps, const uint32_t *step_ptr, uint32_t step_imm,
                             uint32_t *out_tokens, float *out_logprob, cudaStream_t st) {
  if (T <= 0) return cudaSuccess;
  if (sp.tp < 1 || sp.tp > kTpMaxRanks || T > sp.cap_rows || Vl <= 0) {
    set_last_error("tp_sample: 1 <= tp <= 8, rows <= the exchange area's rows, and a non-empty vocabulary slice are required");
    return cudaErrorInvalidValue;
  }
  const bool emulated = sp.rank < 0;
  // emulated ranks wait for each other inside ONE grid: every CTA must be resident (one 1024-thread CTA per SM is certain)
  if (emulated && (long long)T * sp.tp > kNumSMs) {
    set_last_error("tp_sample (one-GPU emulation): rows * tp must not exceed the SM count");
    return cudaErrorInvalidValue;
  }
  constexpr size_t kBinBytes = (size_t)kVals * kThreads * sizeof(float);  // 64 KB of dynamic shared memory
  static DeviceOnce once;
  const cudaError_t e = once.run([] { return cudaFuncSetAttribute(tp_sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBinBytes); });
  if (e != cudaSuccess) { set_last_error("tp_sample: cannot reserve 64 KB of shared memory"); return e; }
  return launch_k(tp_sample_kernel, dim3(T, emulated ? sp.tp : 1), dim3(kThreads), kBinBytes, st, sp, logits, ld, Vl, temperature, top_p,
                  seeds, steps, step_ptr, step_imm, out_tokens, out_logprob);
}

}  // namespace dis
