// Shared helpers for libacai_b200.so (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/acai_b200.h"

namespace acai {

void set_error(const char* fmt, ...);

#define ACAI_CHECK_ARG(cond, ...)        \
  do {                                   \
    if (!(cond)) {                       \
      ::acai::set_error(__VA_ARGS__);    \
      return ACAI_ERR_INVALID;           \
    }                                    \
  } while (0)

#define ACAI_CUDA(call)                                                                 \
  do {                                                                                  \
    cudaError_t e__ = (call);                                                           \
    if (e__ != cudaSuccess) {                                                           \
      ::acai::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return ACAI_ERR_CUDA;                                                             \
    }                                                                                   \
  } while (0)

// ---- programmatic dependent launch (the decode step is ~150 kernels of a few microseconds each) -----------------------
// A kernel launched through launch_pdl may be scheduled while its predecessor in the stream still runs; it must execute
// pdl_wait() before it touches anything the predecessor writes (or writes anything the predecessor reads) - in this library
// always before the first global-memory access and before any early return, so that "this grid has finished" implies "its
// predecessors have finished".  pdl_launch_dependents() lets the NEXT kernel's CTAs be scheduled (they then sit in their
// own pdl_wait).  Without the launch attribute both are no-ops.  ACAI_PDL=0 launches everything fully serialised.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

inline int launch_status(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("launch of %s failed: %s", what, cudaGetErrorString(e));
    return ACAI_ERR_CUDA;
  }
  return ACAI_OK;
}

inline size_t dtype_size(int dt) { return dt == ACAI_BF16 ? 2 : 4; }
inline bool dtype_valid(int dt) { return dt == ACAI_F32 || dt == ACAI_BF16; }

// ---- runtime-typed element access for the generic (any-shape) kernels ----
__device__ __forceinline__ float ld_as_float(const void* p, int64_t i, int dt) {
  return dt == ACAI_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i])
                         : reinterpret_cast<const float*>(p)[i];
}
__device__ __forceinline__ void st_from_float(void* p, int64_t i, int dt, float v) {
  if (dt == ACAI_BF16) reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
  else reinterpret_cast<float*>(p)[i] = v;
}

__device__ __forceinline__ float gelu_erf(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

// Phi(z) = 0.5 (1 + erf(z / sqrt 2)) and e = exp(-z^2 / 2) for the bf16 training kernels: Abramowitz-Stegun 7.1.26
// (|error| < 1.5e-7, far below the bf16 output's rounding) costs two MUFU ops and ~10 FMAs, about a third of erff: with
// erff the 16-byte GELU kernels are instruction-bound (3.0 TB/s), with this they are HBM-bound.
__device__ __forceinline__ void phi_and_pdf(float z, float& cdf, float& e) {
  const float x = fabsf(z) * 0.70710678118654752440f;
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, x, 1.0f)));
  e = __expf(-x * x);
  const float poly = t * (0.254829592f + t * (-0.284496736f + t * (1.421413741f + t * (-1.453152027f + t * 1.061405429f))));
  const float q = 0.5f * poly * e;                     // 0.5 erfc(|z| / sqrt 2)
  cdf = z >= 0.0f ? 1.0f - q : q;
}

// GELU for the bf16 paths (GEMM epilogues, the 16-byte forward kernel): x * Phi(x) with Phi(x) = 0.5 * (1 + tanh(u(x))), u an odd degree-5 polynomial
// fitted so that |gelu_fast - exact erf GELU| <= 2.6e-5 before the MUFU's own error (tanh.approx.f32: rel 2^-11),
// i.e. well inside the bf16 rounding of the stored activation.  9 issue slots + 1 MUFU per element (the A&S erfc
// form it replaces cost 16 + 2 MUFU and made the GELU GEMM epilogue-bound).  u's polynomial turns over past
// |x| = 7.25, where tanh has long saturated, hence the clamp.
__device__ __forceinline__ float gelu_fast(float x) {
  const float xc = fminf(fmaxf(x, -7.0f), 7.0f);
  const float x2 = xc * xc;
  float p = fmaf(x2, -3.51708639e-04f, 3.70068271e-02f);
  p = fmaf(x2, p, 7.97506586e-01f);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(xc * p));
  const float hx = 0.5f * x;
  return fmaf(hx, t, hx);
}

// ---- counter-based dropout randomness ---------------------------------------------------------------------------
// One 32-bit avalanche hash (two multiply / xor-shift rounds) of (seed, index of an element PAIR) yields two 16-bit
// uniforms; an element is kept when its 16 bits are >= thr = round(p * 65536).  Stateless, so the backward kernels
// regenerate exactly the forward's mask from (seed, index) - no mask tensor is stored.
__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
// (split in two so that kernels hashing a run of pairs with the same high word - 4 pairs per 16-byte vector - pay the
// inner round once: dropout_bits(seed, pair) == dropout_bits_lo(dropout_bits_hi(seed, pair >> 32), (uint32_t)pair))
__host__ __device__ __forceinline__ uint32_t dropout_bits_hi(uint64_t seed, uint32_t pair_hi) {
  return mix32(pair_hi ^ (uint32_t)seed) ^ (uint32_t)(seed >> 32);
}
__host__ __device__ __forceinline__ uint32_t dropout_bits_lo(uint32_t hi_const, uint32_t pair_lo) { return mix32(pair_lo ^ hi_const); }
__host__ __device__ __forceinline__ uint32_t dropout_bits(uint64_t seed, uint64_t pair) {
  return dropout_bits_lo(dropout_bits_hi(seed, (uint32_t)(pair >> 32)), (uint32_t)pair);
}
__host__ __device__ __forceinline__ bool dropout_keep(uint64_t seed, uint64_t idx, uint32_t thr) {
  return ((dropout_bits(seed, idx >> 1) >> ((idx & 1) * 16)) & 0xffffu) >= thr;
}
__host__ __device__ __forceinline__ uint32_t dropout_threshold(float p) {
  const float t = p * 65536.0f + 0.5f;
  return t <= 0.0f ? 0u : (t >= 65536.0f ? 65536u : (uint32_t)t);
}

// Attention-probability dropout (nn.MultiheadAttention(dropout=p) in training mode): element (head h, global query row
// q, key position k inside its sequence) belongs to pair (rowkey << 15) + (k >> 1) with rowkey = h * total_q + q, and
// takes the low / high 16 bits of that pair's hash for even / odd k.  All attention kernels (tensor-core and SIMT,
// forward and both backward passes) derive the mask from these three functions, so they agree bit for bit.
struct AttnDropout { uint64_t seed; uint32_t thr; float rp; };        // thr = 0: no dropout; rp = 1 / (1 - p)
__host__ __device__ __forceinline__ uint32_t attn_drop_rowconst(uint64_t seed, uint64_t rowkey) {
  return mix32((uint32_t)((rowkey << 15) >> 32) ^ (uint32_t)seed) ^ (uint32_t)(seed >> 32);
}
// The per-pair hash sits in the inner loop of three issue-bound kernels, so it is two multiply-fold rounds (32 x 32 -> 64
// bit product, high ^ low word: IMAD.WIDE + LOP3 each) instead of mix32's two multiplies and three xor-shifts - 5
// instructions per pair instead of 10.  One round alone is a Weyl sequence on the consecutive (row, pair) counters
// (drops of neighbouring rows anti-correlated at -0.1); two rounds are statistically indistinguishable from mix32 here
// (drop rate, neighbour / row / column correlations and the variance of per-row rates, checked for p = 0.05 .. 0.25).
__host__ __device__ __forceinline__ uint32_t fold32(uint32_t x, uint32_t k) {
  const uint64_t m = (uint64_t)x * k;
  return (uint32_t)m ^ (uint32_t)(m >> 32);
}
__host__ __device__ __forceinline__ uint32_t attn_drop_mix(uint32_t x) { return fold32(fold32(x, 0x9E3779B1u), 0x85EBCA77u); }
__host__ __device__ __forceinline__ uint32_t attn_drop_bits(uint32_t rowconst, uint64_t rowkey, uint32_t pair_in_row) {
  return attn_drop_mix(((uint32_t)(rowkey << 15) + pair_in_row) ^ rowconst);
}
__host__ __device__ __forceinline__ bool attn_drop_keep(const AttnDropout& d, uint64_t rowkey, int k) {
  const uint32_t bits = attn_drop_bits(attn_drop_rowconst(d.seed, rowkey), rowkey, (uint32_t)k >> 1);
  return ((bits >> ((k & 1) * 16)) & 0xffffu) >= d.thr;
}

// 2^x on the MUFU pipe (ex2.approx.ftz): ~2 ulp, no denormal fix-up code around it
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// d/dz [z Phi(z)] = Phi(z) + z phi(z) with Phi from gelu_fast's tanh fit (|Phi_fit - Phi| < 1e-4) and phi from one ex2:
// 2 MUFU + 12 issue slots per element (the A&S form of phi_and_pdf: 2 MUFU + 18).  For the bf16 backward kernels, whose
// result is rounded to 8 bits of mantissa.
__device__ __forceinline__ float gelu_grad_fast(float z) {
  const float zc = fminf(fmaxf(z, -7.0f), 7.0f);
  const float z2 = zc * zc;
  float p = fmaf(z2, -3.51708639e-04f, 3.70068271e-02f);
  p = fmaf(z2, p, 7.97506586e-01f);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(zc * p));
  const float e = fast_exp2(z2 * -0.72134752044448170368f);          // exp(-z^2 / 2)
  return fmaf(zc * 0.39894228040143267794f, e, fmaf(0.5f, t, 0.5f));
}

// 2^x on the FMA / ALU pipes instead of the 16-lane MUFU unit: Cody-Waite split (round x to an integer n with the
// 1.5 * 2^23 trick, f = x - n in [-0.5, 0.5]), a degree-3 minimax polynomial for 2^f (relative error 7.5e-5, well under
// the 2e-3 half-ulp of the bf16 value it feeds) and n added straight into the exponent field.  x is clamped at -126 (the
// result is then ~1e-38, not 0: callers that need an exact zero for masked positions must not rely on it).
__device__ __forceinline__ float exp2_fma(float x) {
  x = fmaxf(x, -126.0f);
  const float xf = x + 12582912.0f;
  const float f = x - (xf - 12582912.0f);
  float p = fmaf(f, 0.055171653628349304f, 0.2426111251115799f);
  p = fmaf(p, f, 0.6932609677314758f);
  p = fmaf(p, f, 0.9999280571937561f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(xf) << 23));
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

// Block-wide sum for blockDim.x <= 1024 (multiple of 32); `red` is >= 32 floats of smem.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` from a previous use
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.0f;
  return warp_sum(r);
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : -INFINITY;
  return warp_max(r);
}

// Index of the sequence that packed row `row` belongs to: largest b with cu[b] <= row.
__device__ __forceinline__ int find_sequence(const int32_t* __restrict__ cu, int batch, int row) {
  int lo = 0, hi = batch;  // invariant: cu[lo] <= row < cu[hi]
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (__ldg(cu + mid) <= row) lo = mid; else hi = mid;
  }
  return lo;
}

int num_sms();

}  // namespace acai
