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

// Attention-probability dropout (nn.MultiheadAttention(dropout=p) in training mode), BIT-PARALLEL: the keep decisions of
// the 32 keys [32 kb, 32 kb + 32) of one probability row (rowkey = h * total_q + global query row) are the 32 bits of ONE
// word.  Sixteen pseudo-random words w_0 .. w_15 are hashed from (seed, rowkey, kb); bit position b of the sixteen words
// is a 16-bit uniform u_b (w_i supplies bit i), and the bit-sliced comparison u_b >= thr = round(p * 65536) runs on all 32
// positions at once - one LOP3 per word:  ge <- thr_i ? (w_i & ge) : (w_i | ge), from the least significant slice up.
// ~55 instructions per 32 decisions where the per-pair hash + 16-bit compare / select of rounds 1 - 2 took ~190; the
// same drop rate resolution (2^-16).  Words: two multiply-fold rounds of the block counter (the hash of rounds 1 - 2: a
// 32 x 32 -> 64 bit product, high ^ low word, IMAD.WIDE + LOP3; one round alone is a Weyl sequence on consecutive
// counters) and a third round per word with its own odd multiplier.  Drop rate, key / row / diagonal neighbour
// correlations (|r| < 2e-3 on 4 M samples), the variance of per-row, per-column and per-block rates against the binomial
// (ratio 0.92 - 1.06) checked for p = 0.05 .. 0.25 (tests/test_host.py restates the function in numpy).
// Bit layout: key 32 kb + 4 s + j (s = 0 .. 7, j = 0 .. 3) is bit 8 j + 7 - s, so that `word << s` has the four decisions
// of keys 4 s .. 4 s + 3 in the sign bits of its four bytes and ONE prmt (sign-replicating byte select) turns them into
// a 0 / ~0 mask for a packed bf16 pair or a whole fp32 word.  All attention kernels (tensor-core and SIMT, forward and both
// backward passes) derive the mask from these functions, so they agree bit for bit.
struct AttnDropout { uint64_t seed; uint32_t thr; float rp; };        // thr = 0: no dropout; rp = 1 / (1 - p)
__host__ __device__ __forceinline__ uint32_t fold32(uint32_t x, uint32_t k) {
  const uint64_t m = (uint64_t)x * k;
  return (uint32_t)m ^ (uint32_t)(m >> 32);
}
// block counter = (rowkey << 11) + kb: its low word is rowbase + kb, its high word goes through mix32 with the seed
__host__ __device__ __forceinline__ uint32_t attn_drop_rowconst(uint64_t seed, uint64_t rowkey) {
  return mix32((uint32_t)((rowkey << 11) >> 32) ^ (uint32_t)seed) ^ (uint32_t)(seed >> 32);
}
__host__ __device__ __forceinline__ uint32_t attn_drop_rowbase(uint64_t rowkey) { return (uint32_t)(rowkey << 11); }
__host__ __device__ __forceinline__ uint32_t attn_drop_word(uint32_t rowconst, uint32_t rowbase, uint32_t kb, uint32_t thr) {
  constexpr uint32_t K3[16] = {0xC2B2AE3Du, 0x27D4EB2Fu, 0x165667B1u, 0xD3A2646Du, 0xFD7046C5u, 0xB55A4F09u, 0x7FEB352Du, 0x846CA68Bu,
                               0xCC9E2D51u, 0x1B873593u, 0xE6546B65u, 0x2C1B3C6Du, 0x297A2D39u, 0x5BD1E995u, 0x68E31DA5u, 0xB5297A4Du};
  const uint32_t h = fold32(fold32((rowbase + kb) ^ rowconst, 0x9E3779B1u), 0x85EBCA77u);
  uint32_t ge = 0xffffffffu;
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
  for (int i = 0; i < 16; ++i) {
    const uint32_t w = fold32(h, K3[i]);
    const uint32_t t = 0u - ((thr >> i) & 1u);          // (uniform: thr is a kernel parameter)
    ge = (w & ge) | ((w | ge) & ~t);
  }
  return thr >= 65536u ? 0u : ge;
}
__host__ __device__ __forceinline__ int attn_drop_bit_of_key(int k) { return 8 * (k & 3) + 7 - ((k & 31) >> 2); }
__host__ __device__ __forceinline__ bool attn_drop_keep(const AttnDropout& d, uint64_t rowkey, int k) {
  const uint32_t w = attn_drop_word(attn_drop_rowconst(d.seed, rowkey), attn_drop_rowbase(rowkey), (uint32_t)k >> 5, d.thr);
  return (w >> attn_drop_bit_of_key(k)) & 1u;
}
#ifdef __CUDACC__
// 0 / ~0 masks out of a word shifted so that the wanted decisions are byte sign bits (ws = word << s for keys 4 s .. 4 s + 3):
// attn_drop_pair_mask(ws, 0 / 1): both 16-bit halves for keys (4 s, 4 s + 1) / (4 s + 2, 4 s + 3); attn_drop_full_mask(ws, j): key 4 s + j
// (pair / j are compile-time constants after unrolling at every call site)
// (inline PTX: the __byte_perm intrinsic only honours three bits per selector nibble, the sign-replicate flag is the fourth)
__device__ __forceinline__ uint32_t prmt_sign(uint32_t a, uint32_t sel) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(0u), "r"(sel));
  return d;
}
__device__ __forceinline__ uint32_t attn_drop_pair_mask(uint32_t ws, int pair) { return prmt_sign(ws, pair ? 0xBBAAu : 0x9988u); }
__device__ __forceinline__ uint32_t attn_drop_full_mask(uint32_t ws, int j) { return prmt_sign(ws, 0x8888u + 0x1111u * (uint32_t)j); }
// 32 x 32 bit transpose across a warp: lane b ends up with bit l = bit b of lane l's x.  Used by the kernels whose threads own a
// KEY (dK / dV orientation): lane l builds the word of query column l, the transpose hands every key its 32 query decisions.
__device__ __forceinline__ uint32_t warp_transpose32(uint32_t x, int lane) {
#pragma unroll
  for (int k = 16; k >= 1; k >>= 1) {
    const uint32_t m = k == 16 ? 0x0000ffffu : k == 8 ? 0x00ff00ffu : k == 4 ? 0x0f0f0f0fu : k == 2 ? 0x33333333u : 0x55555555u;
    const uint32_t y = __shfl_xor_sync(0xffffffffu, x, k);
    x = (lane & k) ? ((x & ~m) | ((y & ~m) >> k)) : ((x & m) | ((y & m) << k));
  }
  return x;
}
// Thread = key row (lane = key & 31, the warp's 32 keys are one aligned block kb), 32 query columns c = 0 .. 31 with hash
// constants rc[c] / rb[c] (attn_drop_rowconst / attn_drop_rowbase of the column's row): returns the word whose bit
// 8 j + 7 - s is the decision of (column 4 s + j, this thread's key) - the same layout along the columns as attn_drop_word has
// along the keys, so the callers use the same shift + prmt.  Lane l hashes column perm(l), perm(8 j + 7 - s) = 4 s + j.
__device__ __forceinline__ uint32_t attn_drop_word_transposed(const uint32_t* rc, const uint32_t* rb, uint32_t kb, uint32_t thr, int lane) {
  const int c = 4 * (7 - (lane & 7)) + (lane >> 3);
  const uint32_t w = attn_drop_word(rc[c], rb[c], kb, thr);              // bit 8 j + 7 - s: key 4 s + j of column c
  const uint32_t t = warp_transpose32(w, lane);                          // lane b: bit l = (column perm(l), key of bit b)
  return __shfl_sync(0xffffffffu, t, attn_drop_bit_of_key(lane));        // the lane that holds this thread's key
}
#endif

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
