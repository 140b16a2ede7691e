// Backward-pass building blocks of the teacher-forced / MAE training step.  The dense
// contractions of the backward (dX = dY W, dW = dY^T X) run on the same tcgen05 GEMM as the forward: this file supplies
// the operand transposes they need plus the memory-bound pieces (bias gradients, GELU, LayerNorm, cross-entropy,
// embedding / position scatter-adds, MAE glue).  All reductions that feed parameters have a fixed order or use fp32
// atomics on distinct rows only where the reference's own autograd is non-deterministic too (index_add).
#include <mutex>
#include "common.cuh"

namespace acai {

// out[c, r] = in[r, c] for r < rows, 0 for rows <= r < ld_out (zero padding so that the transposed matrix can be a
// K-major GEMM operand with K % 64 == 0).  32x32 tiles through padded shared memory: coalesced both ways.
template <typename T>
__global__ void __launch_bounds__(256)
transpose_pad_kernel(const T* __restrict__ in, int64_t ld_in, int rows, int cols, T* __restrict__ out, int64_t ld_out) {
  __shared__ T tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int r = r0 + ty + i, c = c0 + tx;
    tile[ty + i][tx] = (r < rows && c < cols) ? in[(int64_t)r * ld_in + c] : T(0);
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int c = c0 + ty + i, r = r0 + tx;
    if (c < cols && r < ld_out) out[(int64_t)c * ld_out + r] = tile[tx][ty + i];
  }
}

// Column sums in a fixed order: stage 1 sums 128-row slabs per column, stage 2 folds the slabs.
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const void* __restrict__ x, int dt, int64_t ld, int rows, int cols, float* __restrict__ part) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r0 = blockIdx.y * 128, r1 = min(rows, r0 + 128);
  if (c >= cols) return;
  float s = 0.0f;
  for (int r = r0; r < r1; ++r) s += ld_as_float(x, (int64_t)r * ld + c, dt);
  part[(int64_t)blockIdx.y * cols + c] = s;
}

// bf16 fast path: a CTA covers 256 columns x one slab of CS_SLAB rows as 32 column groups (one 16-byte load = 8 columns
// per thread) x 8 row lanes; the 8 row lanes are folded through shared memory in a fixed order.
constexpr int CS_SLAB = 256;
__global__ void __launch_bounds__(256)
colsum_partial_bf16x8_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld, int rows, int cols, float* __restrict__ part) {
  __shared__ float red[8][256 + 8];
  const int cg = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = blockIdx.x * 256 + cg * 8;
  const int r0 = blockIdx.y * CS_SLAB, r1 = min(rows, r0 + CS_SLAB);
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.0f;
  if (c < cols) {
#pragma unroll 4
    for (int r = r0 + ry; r < r1; r += 8) {
      const uint4 u = __ldcs(reinterpret_cast<const uint4*>(x + (int64_t)r * ld + c));
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int e = 0; e < 4; ++e) { const float2 f = __bfloat1622float2(h[e]); acc[2 * e] += f.x; acc[2 * e + 1] += f.y; }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) red[ry][cg * 8 + e] = acc[e];
  __syncthreads();
  const int cc = blockIdx.x * 256 + threadIdx.x;
  if (cc < cols) {
    float t = 0.0f;
#pragma unroll
    for (int y = 0; y < 8; ++y) t += red[y][threadIdx.x];
    part[(int64_t)blockIdx.y * cols + cc] = t;
  }
}
// stage 2, parallel over the slabs as well: 32 columns x 8 slab lanes per CTA, fixed summation order
__global__ void __launch_bounds__(256)
colsum_final_wide_kernel(const float* __restrict__ part, int slabs, int cols, float* __restrict__ out, int accumulate) {
  __shared__ float red[8][33];
  const int cx = threadIdx.x & 31, sy = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  float s = 0.0f;
  if (c < cols) {
#pragma unroll 4
    for (int i = sy; i < slabs; i += 8) s += part[(int64_t)i * cols + c];
  }
  red[sy][cx] = s;
  __syncthreads();
  if (sy == 0 && c < cols) {
    float t = 0.0f;
#pragma unroll
    for (int y = 0; y < 8; ++y) t += red[y][cx];
    out[c] = accumulate ? out[c] + t : t;
  }
}

// the same for TWO slab buffers in one launch (LayerNorm backward: dgamma and dbeta; blockIdx.y picks the pair): 62 launches
// fewer per teacher-forced step than two colsum_final_wide_kernel launches per LayerNorm
__global__ void __launch_bounds__(256)
colsum_final_pair_kernel(const float* __restrict__ part0, const float* __restrict__ part1, int slabs, int cols,
                         float* __restrict__ out0, float* __restrict__ out1, int accumulate) {
  __shared__ float red[8][33];
  const float* part = blockIdx.y ? part1 : part0;
  float* out = blockIdx.y ? out1 : out0;
  const int cx = threadIdx.x & 31, sy = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  float s = 0.0f;
  if (c < cols) {
#pragma unroll 4
    for (int i = sy; i < slabs; i += 8) s += part[(int64_t)i * cols + c];
  }
  red[sy][cx] = s;
  __syncthreads();
  if (sy == 0 && c < cols) {
    float t = 0.0f;
#pragma unroll
    for (int y = 0; y < 8; ++y) t += red[y][cx];
    out[c] = accumulate ? out[c] + t : t;
  }
}

// exact-erf GELU and its derivative  d/dz [z Phi(z)] = Phi(z) + z phi(z)
__device__ __forceinline__ float gelu_grad(float v) {
  const float cdf = 0.5f * (1.0f + erff(v * 0.70710678118654752440f));
  const float pdf = 0.39894228040143267794f * __expf(-0.5f * v * v);
  return cdf + v * pdf;
}
// Both take an optional dropout (thr != 0) applied to the activation: out = keep(i) * gelu(z) / (1 - p), the mask being a
// function of (seed, i) only, so the backward regenerates it (nn.Dropout after the activation of the feed-forward block).
__global__ void gelu_fwd_kernel(const void* __restrict__ z, int dt, void* __restrict__ out, int out_dt, int64_t n, uint32_t thr,
                                float rp, uint64_t seed) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = gelu_erf(ld_as_float(z, i, dt));
    if (thr) v = dropout_keep(seed, (uint64_t)i, thr) ? v * rp : 0.0f;
    st_from_float(out, i, out_dt, v);
  }
}
__global__ void gelu_bwd_kernel(const void* __restrict__ dy, int dy_dt, const void* __restrict__ z, int z_dt,
                                void* __restrict__ dz, int dz_dt, int64_t n, uint32_t thr, float rp, uint64_t seed) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float g = ld_as_float(dy, i, dy_dt) * gelu_grad(ld_as_float(z, i, z_dt));
    if (thr) g = dropout_keep(seed, (uint64_t)i, thr) ? g * rp : 0.0f;
    st_from_float(dz, i, dz_dt, g);
  }
}
// keep factors (rp or 0) of the 8 consecutive elements 8*i8 .. 8*i8+7 = hash pairs 4*i8 .. 4*i8+3
__device__ __forceinline__ void dropout_factors8(uint64_t seed, int64_t i8, uint32_t thr, float rp, float (&f)[8]) {
  const uint64_t pair0 = (uint64_t)(4 * i8);             // a multiple of 4: the four pairs share their high word
  const uint32_t hc = dropout_bits_hi(seed, (uint32_t)(pair0 >> 32));
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t bits = dropout_bits_lo(hc, (uint32_t)pair0 + k);
    f[2 * k] = (bits & 0xffffu) >= thr ? rp : 0.0f;
    f[2 * k + 1] = (bits >> 16) >= thr ? rp : 0.0f;
  }
}
// bf16 fast paths: 8 elements (16 bytes) per vector, TWO vectors per thread per step with both loads issued before any
// math (one 16-byte load in flight per thread left these kernels at ~50 % of HBM peak)
__device__ __forceinline__ uint4 gelu_fwd_vec(uint4 u, int64_t i, uint32_t thr, float rp, uint64_t seed) {
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
  float m[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) m[k] = 1.0f;
  if (thr) dropout_factors8(seed, i, thr, rp, m);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 f = __bfloat1622float2(h[k]);
    // tanh form (one MUFU + 9 issue slots, 2.6e-5 from the exact erf GELU: far inside the bf16 rounding of h); the backward
    // differentiates the exact form from the saved z
    h[k] = __floats2bfloat162_rn(gelu_fast(f.x) * m[2 * k], gelu_fast(f.y) * m[2 * k + 1]);
  }
  return u;
}
__device__ __forceinline__ uint4 gelu_bwd_vec(uint4 u, uint4 g, int64_t i, uint32_t thr, float rp, uint64_t seed) {
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
  const __nv_bfloat162* gh = reinterpret_cast<const __nv_bfloat162*>(&g);
  float m[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) m[k] = 1.0f;
  if (thr) dropout_factors8(seed, i, thr, rp, m);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 f = __bfloat1622float2(h[k]), d = __bfloat1622float2(gh[k]);
    h[k] = __floats2bfloat162_rn(d.x * gelu_grad_fast(f.x) * m[2 * k], d.y * gelu_grad_fast(f.y) * m[2 * k + 1]);
  }
  return u;
}
__device__ __forceinline__ uint4 dropout_vec(uint4 u, uint4 r, bool has_res, int64_t i, uint32_t thr, float rp, uint64_t seed) {
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
  const __nv_bfloat162* rh = reinterpret_cast<const __nv_bfloat162*>(&r);
  float m[8];
  dropout_factors8(seed, i, thr, rp, m);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 f = __bfloat1622float2(h[k]);
    const float2 g = has_res ? __bfloat1622float2(rh[k]) : make_float2(0.0f, 0.0f);
    h[k] = has_res ? __floats2bfloat162_rn(fmaf(f.x, m[2 * k], g.x), fmaf(f.y, m[2 * k + 1], g.y))
                   : __floats2bfloat162_rn(f.x * m[2 * k], f.y * m[2 * k + 1]);
  }
  return u;
}

__global__ void __launch_bounds__(256)
gelu_fwd_bf16x8_kernel(const uint4* __restrict__ z, uint4* __restrict__ out, int64_t n8, uint32_t thr, float rp, uint64_t seed) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += 2 * stride) {
    const int64_t j = i + stride;
    const bool two = j < n8;
    const uint4 u0 = z[i];
    const uint4 u1 = two ? z[j] : make_uint4(0u, 0u, 0u, 0u);
    out[i] = gelu_fwd_vec(u0, i, thr, rp, seed);
    if (two) out[j] = gelu_fwd_vec(u1, j, thr, rp, seed);
  }
}
__global__ void __launch_bounds__(256)
gelu_bwd_bf16x8_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ z, uint4* __restrict__ dz, int64_t n8, uint32_t thr,
                       float rp, uint64_t seed) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += 2 * stride) {
    const int64_t j = i + stride;
    const bool two = j < n8;
    const uint4 u0 = z[i], g0 = dy[i];
    const uint4 u1 = two ? z[j] : make_uint4(0u, 0u, 0u, 0u), g1 = two ? dy[j] : make_uint4(0u, 0u, 0u, 0u);
    dz[i] = gelu_bwd_vec(u0, g0, i, thr, rp, seed);
    if (two) dz[j] = gelu_bwd_vec(u1, g1, j, thr, rp, seed);
  }
}
// residual + dropout(x), bf16
__global__ void __launch_bounds__(256)
dropout_bf16x8_kernel(const uint4* __restrict__ x, const uint4* __restrict__ res, uint4* __restrict__ out, int64_t n8, uint32_t thr,
                      float rp, uint64_t seed) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const bool has_res = res != nullptr;
  const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += 2 * stride) {
    const int64_t j = i + stride;
    const bool two = j < n8;
    const uint4 u0 = x[i], r0 = has_res ? res[i] : zero;
    const uint4 u1 = two ? x[j] : zero, r1 = (two && has_res) ? res[j] : zero;
    out[i] = dropout_vec(u0, r0, has_res, i, thr, rp, seed);
    if (two) out[j] = dropout_vec(u1, r1, has_res, j, thr, rp, seed);
  }
}

// LayerNorm backward for y = (v - mu) * rstd * gamma + beta with v = x (+ residual):
//   dv = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * gamma ;  dgamma = sum_rows dy * xhat ;  dbeta = sum_rows dy
// One warp per row; each CTA (8 rows per pass, grid-stride) keeps per-column partial sums in registers and writes one
// slab of partials, folded by colsum_final_wide_kernel (fixed order).
template <int NV>   // cols <= NV * 32; lane owns columns lane + 32 * i
__global__ void __launch_bounds__(256)
layernorm_bwd_kernel(const void* __restrict__ dy, int dy_dt, const void* __restrict__ x, int x_dt,
                     const void* __restrict__ res, int res_dt, const float* __restrict__ gamma, float eps,
                     void* __restrict__ dv, int dv_dt, float* __restrict__ part_g, float* __restrict__ part_b,
                     int rows, int cols) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float ag[NV], ab[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) { ag[i] = 0.0f; ab[i] = 0.0f; }
  for (int row = blockIdx.x * 8 + warp; row < rows; row += gridDim.x * 8) {
    const int64_t base = (int64_t)row * cols;
    float v[NV], d[NV];
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      v[i] = 0.0f; d[i] = 0.0f;
      if (c < cols) {
        v[i] = ld_as_float(x, base + c, x_dt) + (res ? ld_as_float(res, base + c, res_dt) : 0.0f);
        d[i] = ld_as_float(dy, base + c, dy_dt);
        s += v[i];
      }
    }
    const float mu = warp_sum(s) / (float)cols;
    float q = 0.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i) if (lane + 32 * i < cols) { const float t = v[i] - mu; q += t * t; }
    const float rstd = rsqrtf(warp_sum(q) / (float)cols + eps);
    float m1 = 0.0f, m2 = 0.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < cols) {
        const float xh = (v[i] - mu) * rstd, g = d[i] * __ldg(gamma + c);
        m1 += g; m2 += g * xh;
        ag[i] += d[i] * xh; ab[i] += d[i];
        v[i] = xh; d[i] = g;
      }
    }
    m1 = warp_sum(m1) / (float)cols; m2 = warp_sum(m2) / (float)cols;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < cols) st_from_float(dv, base + c, dv_dt, rstd * (d[i] - m1 - v[i] * m2));
    }
  }
  const int64_t slab = (int64_t)(blockIdx.x * 8 + warp) * cols;     // one slab of partials per warp
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + 32 * i;
    if (c < cols) { part_g[slab + c] = ag[i]; part_b[slab + c] = ab[i]; }
  }
}

// ---- vectorised LayerNorm backward: TPR threads (a multiple of 32) own one row, 8 consecutive columns each (16-byte
// bf16 / 2 x 16-byte fp32 accesses); G = blockDim / TPR rows are in flight per CTA, each row group synchronising on its
// own named barrier.  Column partial sums stay in 16 registers per thread, are folded over the G groups through shared
// memory and leave the CTA as ONE slab (fixed order everywhere -> bitwise reproducible dgamma / dbeta).
__device__ __forceinline__ void load8(const void* p, int dt, int64_t i, float (&v)[8]) {
  if (dt == ACAI_BF16) {
    const uint4 u = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p) + i);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int k = 0; k < 4; ++k) { const float2 f = __bfloat1622float2(h[k]); v[2 * k] = f.x; v[2 * k + 1] = f.y; }
  } else {
    const float4 a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p) + i);
    const float4 b = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p) + i + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
}
__device__ __forceinline__ void store8(void* p, int dt, int64_t i, const float (&v)[8]) {
  if (dt == ACAI_BF16) {
    uint4 u; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
    *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p) + i) = u;
  } else {
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(p) + i) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(p) + i + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
}

// PRE (x, dy and the residual are all bf16): the NEXT row's 16-byte pieces are loaded before the current row's three
// group reductions, so every thread keeps two rows of HBM traffic in flight.
template <int TPR, bool PRE>
__global__ void __launch_bounds__(384)
layernorm_bwd_vec_kernel(const void* __restrict__ dy, int dy_dt, const void* __restrict__ x, int x_dt,
                         const void* __restrict__ res, int res_dt, const float* __restrict__ gamma, float eps,
                         void* __restrict__ dv, int dv_dt, float* __restrict__ part_g, float* __restrict__ part_b,
                         int rows, int cols) {
  constexpr int NW = TPR / 32;                       // warps per row group
  extern __shared__ float sm[];                      // [G][NW][6] reduction scratch, then [G][cols] x 2 for the final fold
  const int G = blockDim.x / TPR;
  const int g = threadIdx.x / TPR, t = threadIdx.x % TPR, wg = t >> 5, lane = threadIdx.x & 31;
  const int c0 = t * 8;
  const bool active = c0 < cols;
  float* red = sm + g * NW * 6;
  float gm[8], ag[8], ab[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) { gm[k] = active ? __ldg(gamma + c0 + k) : 0.0f; ag[k] = 0.0f; ab[k] = 0.0f; }
  auto group_sum2 = [&](float a, float b, int slot, float& ra, float& rb) {     // all TPR threads of the group take part
    a = warp_sum(a); b = warp_sum(b);
    if (NW == 1) { ra = a; rb = b; return; }
    if (lane == 0) { red[wg * 6 + slot] = a; red[wg * 6 + slot + 1] = b; }
    asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(TPR) : "memory");
    ra = 0.0f; rb = 0.0f;
#pragma unroll
    for (int w = 0; w < NW; ++w) { ra += red[w * 6 + slot]; rb += red[w * 6 + slot + 1]; }
  };
  const float inv_n = 1.0f / (float)cols;
  const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
  uint4 nx = zero4, nd = zero4, nr = zero4;
  auto fetch_raw = [&](int r) {
    const int64_t b8 = ((int64_t)r * cols + c0) >> 3;
    nx = __ldcs(reinterpret_cast<const uint4*>(x) + b8);
    nd = __ldcs(reinterpret_cast<const uint4*>(dy) + b8);
    if (res) nr = __ldcs(reinterpret_cast<const uint4*>(res) + b8);
  };
  auto unpack8 = [](const uint4& u, float (&f)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int k = 0; k < 4; ++k) { const float2 t2 = __bfloat1622float2(h[k]); f[2 * k] = t2.x; f[2 * k + 1] = t2.y; }
  };
  if (PRE && active && blockIdx.x * G + g < rows) fetch_raw(blockIdx.x * G + g);
  for (int row = blockIdx.x * G + g; row < rows; row += gridDim.x * G) {
    const int64_t base = (int64_t)row * cols + c0;
    float v[8], d[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { v[k] = 0.0f; d[k] = 0.0f; }
    if (PRE) {
      if (active) {
        unpack8(nx, v);
        unpack8(nd, d);
        if (res) { float r8[8]; unpack8(nr, r8);
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] += r8[k]; }
        const int next = row + gridDim.x * G;
        if (next < rows) fetch_raw(next);
      }
    } else if (active) {
      load8(x, x_dt, base, v);
      if (res) { float r8[8]; load8(res, res_dt, base, r8);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] += r8[k]; }
      load8(dy, dy_dt, base, d);
    }
    float s = 0.0f, dummy;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += v[k];
    float mu;
    group_sum2(s, 0.0f, 0, mu, dummy);
    mu *= inv_n;
    float q = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) { const float tt = active ? v[k] - mu : 0.0f; q += tt * tt; }
    float var;
    group_sum2(q, 0.0f, 2, var, dummy);
    const float rstd = rsqrtf(var * inv_n + eps);
    float m1 = 0.0f, m2 = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float xh = active ? (v[k] - mu) * rstd : 0.0f, gg = d[k] * gm[k];
      m1 += gg; m2 += gg * xh;
      ag[k] += d[k] * xh; ab[k] += d[k];
      v[k] = xh; d[k] = gg;
    }
    // three slot pairs: a warp that races ahead into the next row writes pair 0 while a slower one may still read pair 4
    group_sum2(m1, m2, 4, m1, m2);
    m1 *= inv_n; m2 *= inv_n;
    if (active) {
      float o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = rstd * (d[k] - m1 - v[k] * m2);
      store8(dv, dv_dt, base, o);
    }
  }
  // fold the G groups (fixed order) and write this CTA's slab
  __syncthreads();
  float* fg = sm;
  float* fbeta = sm + (size_t)G * cols;
  if (active) {
#pragma unroll
    for (int k = 0; k < 8; ++k) { fg[g * cols + c0 + k] = ag[k]; fbeta[g * cols + c0 + k] = ab[k]; }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < cols; c += blockDim.x) {
    float a = 0.0f, b2 = 0.0f;
    for (int gg = 0; gg < G; ++gg) { a += fg[gg * cols + c]; b2 += fbeta[gg * cols + c]; }
    part_g[(int64_t)blockIdx.x * cols + c] = a;
    part_b[(int64_t)blockIdx.x * cols + c] = b2;
  }
}

// Cross-entropy with ignore_index and label smoothing e (OMRCELoss, models.py:784-796 = nn.CrossEntropyLoss(ignore_index,
// label_smoothing)), one warp per row, with q = (1 - e) onehot(target) + e / V:
//   row_loss = -sum_c q_c log softmax(z)_c = lse - (1 - e) z[target] - (e / V) sum_c z_c   (0 for ignored rows)
//   dlogits  = softmax(z) - q                                                             (0 for ignored rows)
// dlogits is scaled by *scale_ptr (= upstream gradient / count, a device scalar).  `valid[row]` = 1 for counted rows.
__global__ void ce_fwd_bwd_kernel(const void* __restrict__ logits, int dt, int64_t ld, const int64_t* __restrict__ targets,
                                  int rows, int vocab, int ignore_index, float* __restrict__ row_loss,
                                  uint8_t* __restrict__ valid, void* __restrict__ dlogits, int dl_dt, int64_t ldd,
                                  const float* __restrict__ scale_ptr, float smoothing) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int64_t tgt = targets[row];
  const bool ok = tgt != ignore_index;
  float mx = -INFINITY;
  for (int c = lane; c < vocab; c += 32) mx = fmaxf(mx, ld_as_float(logits, (int64_t)row * ld + c, dt));
  mx = warp_max(mx);
  float s = 0.0f, zsum = 0.0f;
  for (int c = lane; c < vocab; c += 32) {
    const float z = ld_as_float(logits, (int64_t)row * ld + c, dt);
    s += expf(z - mx);
    zsum += z;
  }
  s = warp_sum(s);
  const float lse = mx + logf(s);
  const float uni = smoothing / (float)vocab;
  if (smoothing != 0.0f) zsum = warp_sum(zsum);
  if (lane == 0 && row_loss) {
    float l = 0.0f;
    if (ok) {
      l = lse - ld_as_float(logits, (int64_t)row * ld + tgt, dt);
      if (smoothing != 0.0f) l = (1.0f - smoothing) * l + uni * ((float)vocab * lse - zsum);
    }
    row_loss[row] = l;
    valid[row] = ok ? 1 : 0;
  }
  if (dlogits) {
    const float sc = *scale_ptr;
    for (int c = lane; c < vocab; c += 32) {
      float g = 0.0f;
      if (ok) g = sc * (expf(ld_as_float(logits, (int64_t)row * ld + c, dt) - lse) - (c == tgt ? 1.0f - smoothing : 0.0f) - uni);
      st_from_float(dlogits, (int64_t)row * ldd + c, dl_dt, g);
    }
  }
}

// dst[index[r], :] += src[r, :]  (fp32 atomics; rows with index < 0 are skipped).  Embedding / position-table
// gradients: the reference's index_add / embedding_dense_backward are atomic adds as well.
__global__ void scatter_add_rows_kernel(const void* __restrict__ src, int dt, int64_t ld, const int64_t* __restrict__ index,
                                        int rows, int dim, float* __restrict__ dst, float scale) {
  const int r = blockIdx.x;
  const int64_t d = index[r];
  if (d < 0) return;
  for (int c = threadIdx.x; c < dim; c += blockDim.x)
    atomicAdd(dst + d * dim + c, scale * ld_as_float(src, (int64_t)r * ld + c, dt));
}

// out[r, :] = index[r] >= 0 ? src[index[r], :] : 0   (gather of gradient rows, e.g. MAE restore backward)
__global__ void gather_rows_kernel(const void* __restrict__ src, int dt, const int64_t* __restrict__ index, int rows, int dim,
                                   void* __restrict__ out, int out_dt) {
  const int r = blockIdx.x;
  const int64_t s = index[r];
  for (int c = threadIdx.x; c < dim; c += blockDim.x)
    st_from_float(out, (int64_t)r * dim + c, out_dt, s >= 0 ? ld_as_float(src, s * dim + c, dt) : 0.0f);
}

// MAELoss backward: dpred = mask * scale * 2 (pred - tn) / F with tn the per-patch normalised target.
__global__ void mae_loss_bwd_kernel(const void* __restrict__ pred, int pred_dt, const float* __restrict__ target,
                                    const uint8_t* __restrict__ mask, int feats, const float* __restrict__ scale_ptr,
                                    void* __restrict__ dpred, int dp_dt) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  const int64_t base = (int64_t)row * feats;
  if (!mask[row]) { for (int c = threadIdx.x; c < feats; c += blockDim.x) st_from_float(dpred, base + c, dp_dt, 0.0f); return; }
  float s = 0.0f;
  for (int c = threadIdx.x; c < feats; c += blockDim.x) s += target[base + c];
  const float mean = block_sum(s, red) / (float)feats;
  float q = 0.0f;
  for (int c = threadIdx.x; c < feats; c += blockDim.x) { const float d = target[base + c] - mean; q += d * d; }
  const float rstd = 1.0f / sqrtf(block_sum(q, red) / (float)(feats - 1) + 1.0e-6f);
  const float k = 2.0f * (*scale_ptr) / (float)feats;
  for (int c = threadIdx.x; c < feats; c += blockDim.x)
    st_from_float(dpred, base + c, dp_dt, k * (ld_as_float(pred, base + c, pred_dt) - (target[base + c] - mean) * rstd));
}

}  // namespace acai

namespace acai {
__global__ void masked_sum_kernel(const float* __restrict__ row_loss, const uint8_t* __restrict__ mask, int rows, float* __restrict__ out);
}
using namespace acai;

static unsigned grid_for(int64_t n, int threads = 256) {
  const int64_t b = (n + threads - 1) / threads;
  return (unsigned)(b < 148 * 32 ? (b < 1 ? 1 : b) : 148 * 32);
}

extern "C" int acai_transpose(const void* in, int32_t dtype, int64_t ld_in, int32_t rows, int32_t cols, void* out,
                              int64_t ld_out, void* stream) {
  ACAI_CHECK_ARG(in && out && rows > 0 && cols > 0 && ld_in >= cols && ld_out >= rows && dtype_valid(dtype),
                 "acai_transpose: bad argument");
  dim3 grid((cols + 31) / 32, (unsigned)((ld_out + 31) / 32));
  if (dtype == ACAI_BF16)
    transpose_pad_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const __nv_bfloat16*>(in), ld_in, rows, cols, reinterpret_cast<__nv_bfloat16*>(out), ld_out);
  else
    transpose_pad_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float*>(in), ld_in, rows, cols,
                                                                     reinterpret_cast<float*>(out), ld_out);
  return launch_status("transpose_pad");
}

extern "C" size_t acai_colsum_workspace_bytes(int32_t rows, int32_t cols) {
  return (size_t)((rows + 127) / 128) * cols * sizeof(float);
}

extern "C" int acai_colsum(const void* x, int32_t dtype, int64_t ld, int32_t rows, int32_t cols, float* out,
                           int32_t accumulate, void* workspace, size_t workspace_bytes, void* stream) {
  ACAI_CHECK_ARG(x && out && workspace && rows > 0 && cols > 0 && dtype_valid(dtype), "acai_colsum: bad argument");
  ACAI_CHECK_ARG(workspace_bytes >= acai_colsum_workspace_bytes(rows, cols), "acai_colsum: workspace too small");
  if (dtype == ACAI_BF16 && cols % 8 == 0 && ld % 8 == 0 && !(reinterpret_cast<uintptr_t>(x) & 15)) {
    const int slabs = (rows + CS_SLAB - 1) / CS_SLAB;
    colsum_partial_bf16x8_kernel<<<dim3((cols + 255) / 256, slabs), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), ld, rows, cols, reinterpret_cast<float*>(workspace));
    int rc = launch_status("colsum_partial");
    if (rc) return rc;
    colsum_final_wide_kernel<<<(cols + 31) / 32, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<float*>(workspace), slabs, cols, out,
                                                                               accumulate);
    return launch_status("colsum_final");
  }
  const int slabs = (rows + 127) / 128;
  colsum_partial_kernel<<<dim3((cols + 255) / 256, slabs), 256, 0, (cudaStream_t)stream>>>(x, dtype, ld, rows, cols,
                                                                                        reinterpret_cast<float*>(workspace));
  int rc = launch_status("colsum_partial");
  if (rc) return rc;
  colsum_final_wide_kernel<<<(cols + 31) / 32, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<float*>(workspace), slabs, cols, out,
                                                                             accumulate);
  return launch_status("colsum_final");
}

// ---- acai_colsum_static: the same two kernels with the slab partials in a file-scope device array ---------------------
// For callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): no
// workspace argument, no cudaMalloc / cudaFree / allocating library call anywhere on this path.  The array is part of the
// module image: every GPU that loads the library holds its 4 MiB until the process exits.  In host code the NAME of a
// __device__ array is a host-side stand-in, not a device address, so the address each device gave the array is looked up
// once per device with cudaGetSymbolAddress (that device current) and passed to the kernels like any other workspace.
// All calls share the array: it keeps the last call's partials (nothing has to start at zero - every slot a call reads
// was written by that call's first kernel), and two calls that overlap on the GPU (different streams) race on it.
constexpr size_t COLSUM_STATIC_BYTES = size_t(4) << 20;
static __device__ __align__(256) float g_colsum_static_scratch[COLSUM_STATIC_BYTES / sizeof(float)];
constexpr int COLSUM_STATIC_MAX_DEVICES = 64;
static float* g_colsum_static_ptr[COLSUM_STATIC_MAX_DEVICES] = {};
static std::mutex g_colsum_static_mutex;

extern "C" size_t acai_colsum_static_max_workspace_bytes(void) { return COLSUM_STATIC_BYTES; }

extern "C" int acai_colsum_static(const void* x, int32_t dtype, int64_t ld, int32_t rows, int32_t cols, float* out,
                                  int32_t accumulate, void* stream) {
  ACAI_CHECK_ARG(x && out && rows > 0 && cols > 0 && dtype_valid(dtype), "acai_colsum_static: bad argument");
  const size_t need = acai_colsum_workspace_bytes(rows, cols);
  ACAI_CHECK_ARG(need <= COLSUM_STATIC_BYTES,
                 "acai_colsum_static: %d rows x %d columns need %zu bytes of slab partials, the static scratch holds %zu "
                 "(ceil(rows / 128) * cols * 4 <= 4 MiB); use acai_colsum with a caller workspace",
                 rows, cols, need, COLSUM_STATIC_BYTES);
  int dev = -1;
  ACAI_CUDA(cudaGetDevice(&dev));
  ACAI_CHECK_ARG(dev >= 0 && dev < COLSUM_STATIC_MAX_DEVICES, "acai_colsum_static: device ordinal %d out of range", dev);
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_colsum_static_mutex);
    if (!g_colsum_static_ptr[dev]) {               // set-up, once per device, with that device current
      void* sym = nullptr;
      ACAI_CUDA(cudaGetSymbolAddress(&sym, g_colsum_static_scratch));
      g_colsum_static_ptr[dev] = reinterpret_cast<float*>(sym);
    }
    scratch = g_colsum_static_ptr[dev];
  }
  return acai_colsum(x, dtype, ld, rows, cols, out, accumulate, scratch, COLSUM_STATIC_BYTES, stream);
}

extern "C" int acai_gelu_fwd(const void* z, int32_t z_dtype, void* out, int32_t out_dtype, int64_t n, float dropout_p,
                             uint64_t dropout_seed, void* stream) {
  ACAI_CHECK_ARG(z && out && n >= 0 && dtype_valid(z_dtype) && dtype_valid(out_dtype) && dropout_p >= 0.0f && dropout_p < 1.0f,
                 "acai_gelu_fwd: bad argument");
  if (n == 0) return ACAI_OK;
  const uint32_t thr = dropout_threshold(dropout_p);
  const float rp = 1.0f / (1.0f - dropout_p);
  if (z_dtype == ACAI_BF16 && out_dtype == ACAI_BF16 && n % 8 == 0 && !((reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(out)) & 15)) {
    gelu_fwd_bf16x8_kernel<<<grid_for(n / 8), 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const uint4*>(z), reinterpret_cast<uint4*>(out), n / 8,
                                                                            thr, rp, dropout_seed);
    return launch_status("gelu_fwd");
  }
  gelu_fwd_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(z, z_dtype, out, out_dtype, n, thr, rp, dropout_seed);
  return launch_status("gelu_fwd");
}

extern "C" int acai_gelu_bwd(const void* dy, int32_t dy_dtype, const void* z, int32_t z_dtype, void* dz, int32_t dz_dtype,
                             int64_t n, float dropout_p, uint64_t dropout_seed, void* stream) {
  ACAI_CHECK_ARG(dy && z && dz && n >= 0 && dtype_valid(dy_dtype) && dtype_valid(z_dtype) && dtype_valid(dz_dtype) &&
                 dropout_p >= 0.0f && dropout_p < 1.0f, "acai_gelu_bwd: bad argument");
  if (n == 0) return ACAI_OK;
  const uint32_t thr = dropout_threshold(dropout_p);
  const float rp = 1.0f / (1.0f - dropout_p);
  if (dy_dtype == ACAI_BF16 && z_dtype == ACAI_BF16 && dz_dtype == ACAI_BF16 && n % 8 == 0 &&
      !((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(dz)) & 15)) {
    gelu_bwd_bf16x8_kernel<<<grid_for(n / 8), 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const uint4*>(dy), reinterpret_cast<const uint4*>(z),
                                                                            reinterpret_cast<uint4*>(dz), n / 8, thr, rp, dropout_seed);
    return launch_status("gelu_bwd");
  }
  gelu_bwd_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(dy, dy_dtype, z, z_dtype, dz, dz_dtype, n, thr, rp, dropout_seed);
  return launch_status("gelu_bwd");
}

static int ln_bwd_grid(int rows) { const int g = (rows + 7) / 8; return g < 296 ? g : 296; }

extern "C" size_t acai_layernorm_bwd_workspace_bytes(int32_t rows, int32_t cols) {
  return (size_t)2 * ln_bwd_grid(rows) * 8 * cols * sizeof(float);
}

extern "C" int acai_layernorm_bwd(const void* dy, int32_t dy_dtype, const void* x, int32_t x_dtype, const void* residual,
                                  int32_t res_dtype, const float* gamma, float eps, void* dv, int32_t dv_dtype,
                                  float* dgamma, float* dbeta, int32_t accumulate, int32_t rows, int32_t cols,
                                  void* workspace, size_t workspace_bytes, void* stream) {
  ACAI_CHECK_ARG(dy && x && gamma && dv && dgamma && dbeta && workspace, "acai_layernorm_bwd: null pointer");
  ACAI_CHECK_ARG(rows > 0 && cols > 0 && cols <= 1024, "acai_layernorm_bwd: rows=%d cols=%d out of range (cols <= 1024)", rows, cols);
  ACAI_CHECK_ARG(dtype_valid(dy_dtype) && dtype_valid(x_dtype) && dtype_valid(dv_dtype) && (!residual || dtype_valid(res_dtype)),
                 "acai_layernorm_bwd: bad dtype");
  ACAI_CHECK_ARG(workspace_bytes >= acai_layernorm_bwd_workspace_bytes(rows, cols), "acai_layernorm_bwd: workspace too small");
  int grid = ln_bwd_grid(rows), slabs = grid * 8;
  float* pg = reinterpret_cast<float*>(workspace);
  cudaStream_t st = (cudaStream_t)stream;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (cols % 8 == 0 && al16(dy) && al16(x) && al16(dv) && (!residual || al16(residual))) {
    const int tpr = ((cols / 8) + 31) / 32 * 32;           // 32, 64, 96 or 128
    const int G = 384 / tpr;
    grid = (rows + G - 1) / G < 2 * num_sms() ? (rows + G - 1) / G : 2 * num_sms();
    float* pb2 = pg + (size_t)grid * cols;
    const size_t shm = sizeof(float) * (size_t)(2 * G * cols > G * 4 * 6 ? 2 * G * cols : G * 4 * 6);
    const bool pre = dy_dtype == ACAI_BF16 && x_dtype == ACAI_BF16 && (!residual || res_dtype == ACAI_BF16);
#define LNV2(T, P) layernorm_bwd_vec_kernel<T, P><<<grid, G * T, shm, st>>>(dy, dy_dtype, x, x_dtype, residual, res_dtype, gamma, \
                                                                          eps, dv, dv_dtype, pg, pb2, rows, cols)
#define LNV(T) do { if (pre) LNV2(T, true); else LNV2(T, false); } while (0)
    if (tpr == 32) LNV(32); else if (tpr == 64) LNV(64); else if (tpr == 96) LNV(96); else LNV(128);
#undef LNV
#undef LNV2
    int rc2 = launch_status("layernorm_bwd_vec");
    if (rc2) return rc2;
    colsum_final_pair_kernel<<<dim3((cols + 31) / 32, 2), 256, 0, st>>>(pg, pb2, grid, cols, dgamma, dbeta, accumulate);
    return launch_status("layernorm_bwd_reduce");
  }
  float* pb = pg + (size_t)slabs * cols;
#define LNB(NV) layernorm_bwd_kernel<NV><<<grid, 256, 0, st>>>(dy, dy_dtype, x, x_dtype, residual, res_dtype, gamma, eps, dv, \
                                                               dv_dtype, pg, pb, rows, cols)
  if (cols <= 128) LNB(4); else if (cols <= 256) LNB(8); else if (cols <= 512) LNB(16); else if (cols <= 768) LNB(24); else LNB(32);
#undef LNB
  int rc = launch_status("layernorm_bwd");
  if (rc) return rc;
  colsum_final_pair_kernel<<<dim3((cols + 31) / 32, 2), 256, 0, st>>>(pg, pb, slabs, cols, dgamma, dbeta, accumulate);
  return launch_status("layernorm_bwd_reduce");
}

extern "C" int acai_ce_loss(const void* logits, int32_t dtype, int64_t ld, const int64_t* targets, int32_t rows,
                            int32_t vocab, int32_t ignore_index, float* row_loss, uint8_t* valid, float* sums,
                            void* dlogits, int32_t dlogits_dtype, int64_t ldd, const float* scale, float label_smoothing,
                            void* stream) {
  ACAI_CHECK_ARG(logits && targets && rows > 0 && vocab > 0 && dtype_valid(dtype), "acai_ce_loss: bad argument");
  ACAI_CHECK_ARG(label_smoothing >= 0.0f && label_smoothing <= 1.0f, "acai_ce_loss: label_smoothing %f outside [0, 1]", label_smoothing);
  ACAI_CHECK_ARG((row_loss && valid && sums) || (dlogits && scale && dtype_valid(dlogits_dtype)), "acai_ce_loss: nothing to compute");
  cudaStream_t st = (cudaStream_t)stream;
  ce_fwd_bwd_kernel<<<(rows + 3) / 4, 128, 0, st>>>(logits, dtype, ld, targets, rows, vocab, ignore_index, row_loss, valid,
                                                    dlogits, dlogits_dtype, ldd, scale, label_smoothing);
  int rc = launch_status("ce_loss");
  if (rc || !row_loss) return rc;
  masked_sum_kernel<<<1, 1024, 0, st>>>(row_loss, valid, rows, sums);
  return launch_status("masked_sum");
}

extern "C" int acai_scatter_add_rows(const void* src, int32_t dtype, int64_t ld, const int64_t* index, int32_t rows,
                                     int32_t dim, float* dst, float scale, void* stream) {
  ACAI_CHECK_ARG(src && index && dst && rows >= 0 && dim > 0 && dtype_valid(dtype), "acai_scatter_add_rows: bad argument");
  if (rows == 0) return ACAI_OK;
  scatter_add_rows_kernel<<<rows, dim >= 512 ? 256 : 64, 0, (cudaStream_t)stream>>>(src, dtype, ld, index, rows, dim, dst, scale);
  return launch_status("scatter_add_rows");
}

extern "C" int acai_gather_rows(const void* src, int32_t dtype, const int64_t* index, int32_t rows, int32_t dim, void* out,
                                int32_t out_dtype, void* stream) {
  ACAI_CHECK_ARG(src && index && out && rows >= 0 && dim > 0 && dtype_valid(dtype) && dtype_valid(out_dtype),
                 "acai_gather_rows: bad argument");
  if (rows == 0) return ACAI_OK;
  gather_rows_kernel<<<rows, dim >= 512 ? 256 : 64, 0, (cudaStream_t)stream>>>(src, dtype, index, rows, dim, out, out_dtype);
  return launch_status("gather_rows");
}

extern "C" int acai_mae_loss_bwd(const void* pred, int32_t pred_dtype, const float* target, const uint8_t* loss_mask,
                                 int32_t rows, int32_t feats, const float* scale, void* dpred, int32_t dpred_dtype,
                                 void* stream) {
  ACAI_CHECK_ARG(pred && target && loss_mask && scale && dpred && rows > 0 && feats > 1 && dtype_valid(pred_dtype) &&
                 dtype_valid(dpred_dtype), "acai_mae_loss_bwd: bad argument");
  mae_loss_bwd_kernel<<<rows, feats >= 256 ? 128 : 64, 0, (cudaStream_t)stream>>>(pred, pred_dtype, target, loss_mask, feats,
                                                                                scale, dpred, dpred_dtype);
  return launch_status("mae_loss_bwd");
}

// Fused AdamW over one flat fp32 segment (torch.optim.AdamW semantics, decoupled weight decay; the reference's
// optimiser, omr_teacher_force_train.py:207).  Also refreshes the bf16 shadow copy the GEMMs read, so no per-weight
// cast is launched after a step.  `grad_scale` folds the 1/world_size of a summed all-reduce into the update.
__global__ void adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                             __nv_bfloat16* __restrict__ p_lp, int64_t n, float lr, float b1, float b2, float eps, float wd,
                             float bc1, float bc2_rsqrt, float grad_scale) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float gi = g[i] * grad_scale;
    float pi = p[i] * (1.0f - lr * wd);
    const float mi = b1 * m[i] + (1.0f - b1) * gi;
    const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    pi -= (lr / bc1) * mi / (sqrtf(vi) * bc2_rsqrt + eps);
    p[i] = pi;
    if (p_lp) p_lp[i] = __float2bfloat16(pi);
  }
}

extern "C" int acai_adamw_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, void* param_bf16,
                               int64_t n, float lr, float beta1, float beta2, float eps, float weight_decay, int32_t step,
                               float grad_scale, void* stream) {
  ACAI_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && n >= 0 && step >= 1, "acai_adamw_step: bad argument");
  if (n == 0) return ACAI_OK;
  const float bc1 = 1.0f - powf(beta1, (float)step), bc2 = 1.0f - powf(beta2, (float)step);
  const int64_t blocks = (n + 255) / 256;
  adamw_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, (cudaStream_t)stream>>>(
      param, grad, exp_avg, exp_avg_sq, (__nv_bfloat16*)param_bf16, n, lr, beta1, beta2, eps, weight_decay, bc1,
      1.0f / sqrtf(bc2), grad_scale);
  return launch_status("adamw_step");
}

namespace acai {
int wgrad_tcgen05(const void* dz, int64_t ldz, const void* x, int64_t ldx, float* dw, int64_t ldo, float* dbias, int n_out, int k_in,
                  int T, cudaStream_t st);
}

extern "C" int acai_wgrad_accumulate(const void* dz, int64_t ldz, const void* x, int64_t ldx, int32_t dtype, float* dw,
                                     int64_t ldw, int32_t n_out, int32_t k_in, int32_t tokens, void* stream) {
  ACAI_CHECK_ARG(dz && x && dw && n_out > 0 && k_in > 0 && tokens > 0, "acai_wgrad_accumulate: bad argument");
  if (dtype != ACAI_BF16) { set_error("acai_wgrad_accumulate: only bf16 operands run on the tensor-core path"); return ACAI_ERR_UNSUPPORTED; }
  int rc = acai::wgrad_tcgen05(dz, ldz, x, ldx, dw, ldw, nullptr, n_out, k_in, tokens, (cudaStream_t)stream);
  if (rc == ACAI_ERR_UNSUPPORTED) set_error("acai_wgrad_accumulate: operands are not TMA-addressable (16-byte aligned rows needed)");
  return rc;
}

extern "C" int acai_wgrad_bias_accumulate(const void* dz, int64_t ldz, const void* x, int64_t ldx, int32_t dtype, float* dw,
                                          int64_t ldw, float* dbias, int32_t n_out, int32_t k_in, int32_t tokens, void* stream) {
  ACAI_CHECK_ARG(dz && x && dw && dbias && n_out > 0 && k_in > 0 && tokens > 0, "acai_wgrad_bias_accumulate: bad argument");
  if (dtype != ACAI_BF16) { set_error("acai_wgrad_bias_accumulate: only bf16 operands run on the tensor-core path"); return ACAI_ERR_UNSUPPORTED; }
  int rc = acai::wgrad_tcgen05(dz, ldz, x, ldx, dw, ldw, dbias, n_out, k_in, tokens, (cudaStream_t)stream);
  if (rc == ACAI_ERR_UNSUPPORTED) set_error("acai_wgrad_bias_accumulate: operands are not TMA-addressable (16-byte aligned rows needed)");
  return rc;
}

// ------------------------------------------------------------------------------------------ scheduled sampling (models.py:798-838)
namespace acai {

// F.gumbel_softmax(logits, tau, hard) per row with the Gumbel noise supplied by the caller (so that a run can be replayed):
//   y_soft = softmax((logits + g) / tau);  y = hard ? onehot(argmax y_soft) : y_soft   (first maximum wins, as torch.max)
// y_soft is always written (the backward of both modes is that of y_soft: straight-through estimator).  One warp per row.
__global__ void gumbel_softmax_fwd_kernel(const void* __restrict__ logits, int dt, int64_t ld, const float* __restrict__ noise,
                                          int rows, int vocab, float inv_tau, int hard, float* __restrict__ y_soft,
                                          void* __restrict__ y, int y_dt) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int64_t base = (int64_t)row * vocab;
  float mx = -INFINITY; int arg = 0x7fffffff;
  for (int c = lane; c < vocab; c += 32) {
    const float v = (ld_as_float(logits, (int64_t)row * ld + c, dt) + noise[base + c]) * inv_tau;
    if (v > mx) { mx = v; arg = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, mx, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
  }
  float s = 0.0f;
  for (int c = lane; c < vocab; c += 32)
    s += expf((ld_as_float(logits, (int64_t)row * ld + c, dt) + noise[base + c]) * inv_tau - mx);
  s = warp_sum(s);
  const float inv = 1.0f / s;
  for (int c = lane; c < vocab; c += 32) {
    const float p = expf((ld_as_float(logits, (int64_t)row * ld + c, dt) + noise[base + c]) * inv_tau - mx) * inv;
    y_soft[base + c] = p;
    st_from_float(y, base + c, y_dt, hard ? (c == arg ? 1.0f : 0.0f) : p);
  }
}

// dlogits = (1 / tau) * y_soft * (dy - sum_c y_soft * dy)
__global__ void gumbel_softmax_bwd_kernel(const void* __restrict__ dy, int dy_dt, const float* __restrict__ y_soft, int rows,
                                          int vocab, float inv_tau, void* __restrict__ dlogits, int dl_dt) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int64_t base = (int64_t)row * vocab;
  float dot = 0.0f;
  for (int c = lane; c < vocab; c += 32) dot += y_soft[base + c] * ld_as_float(dy, base + c, dy_dt);
  dot = warp_sum(dot);
  for (int c = lane; c < vocab; c += 32)
    st_from_float(dlogits, base + c, dl_dt, inv_tau * y_soft[base + c] * (ld_as_float(dy, base + c, dy_dt) - dot));
}

// out[r, :] = (sel && sel[r] >= 0 ? a[sel[r], :] : b[r, :]) + (pos ? pos[r % period, :] : 0)
// (mixing expected and gold embeddings; adding the position table to embedding-space decoder inputs)
__global__ void mix_rows_kernel(const void* __restrict__ a, int a_dt, const int64_t* __restrict__ sel, const void* __restrict__ b,
                                int b_dt, const float* __restrict__ pos, int period, int dim, void* __restrict__ out, int out_dt) {
  const int r = blockIdx.x;
  const int64_t s = sel ? sel[r] : -1;
  for (int c = threadIdx.x; c < dim; c += blockDim.x) {
    float v = s >= 0 ? ld_as_float(a, s * dim + c, a_dt) : ld_as_float(b, (int64_t)r * dim + c, b_dt);
    if (pos) v += pos[(int64_t)(r % period) * dim + c];
    st_from_float(out, (int64_t)r * dim + c, out_dt, v);
  }
}

}  // namespace acai

extern "C" int acai_gumbel_softmax_fwd(const void* logits, int32_t dtype, int64_t ld, const float* noise, int32_t rows,
                                       int32_t vocab, float tau, int32_t hard, float* y_soft, void* y, int32_t y_dtype,
                                       void* stream) {
  ACAI_CHECK_ARG(logits && noise && y_soft && y && rows > 0 && vocab > 0 && tau > 0.0f && dtype_valid(dtype) && dtype_valid(y_dtype),
                 "acai_gumbel_softmax_fwd: bad argument");
  acai::gumbel_softmax_fwd_kernel<<<(rows + 3) / 4, 128, 0, (cudaStream_t)stream>>>(logits, dtype, ld, noise, rows, vocab,
                                                                                  1.0f / tau, hard, y_soft, y, y_dtype);
  return launch_status("gumbel_softmax_fwd");
}

extern "C" int acai_gumbel_softmax_bwd(const void* dy, int32_t dy_dtype, const float* y_soft, int32_t rows, int32_t vocab,
                                       float tau, void* dlogits, int32_t dlogits_dtype, void* stream) {
  ACAI_CHECK_ARG(dy && y_soft && dlogits && rows > 0 && vocab > 0 && tau > 0.0f && dtype_valid(dy_dtype) && dtype_valid(dlogits_dtype),
                 "acai_gumbel_softmax_bwd: bad argument");
  acai::gumbel_softmax_bwd_kernel<<<(rows + 3) / 4, 128, 0, (cudaStream_t)stream>>>(dy, dy_dtype, y_soft, rows, vocab, 1.0f / tau,
                                                                                  dlogits, dlogits_dtype);
  return launch_status("gumbel_softmax_bwd");
}

extern "C" int acai_mix_rows(const void* a, int32_t a_dtype, const int64_t* sel, const void* b, int32_t b_dtype, const float* pos,
                             int32_t period, int32_t rows, int32_t dim, void* out, int32_t out_dtype, void* stream) {
  ACAI_CHECK_ARG(b && out && rows >= 0 && dim > 0 && dtype_valid(b_dtype) && dtype_valid(out_dtype) && (!sel || (a && dtype_valid(a_dtype))) &&
                 (!pos || period > 0), "acai_mix_rows: bad argument");
  if (rows == 0) return ACAI_OK;
  acai::mix_rows_kernel<<<rows, dim >= 512 ? 256 : 64, 0, (cudaStream_t)stream>>>(a, a_dtype, sel, b, b_dtype, pos, period, dim, out, out_dtype);
  return launch_status("mix_rows");
}

// ------------------------------------------------------------------------------------------ dropout (hidden states)
namespace acai {
// out[i] = residual[i] + (keep(i) ? x[i] / (1 - p) : 0); the same call on dY (no residual) is the backward.
__global__ void __launch_bounds__(256)
dropout_kernel(const void* __restrict__ x, int dt, const void* __restrict__ res, int res_dt, int64_t n, uint32_t thr, float scale,
               uint64_t seed, void* __restrict__ out, int out_dt) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t pr = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; 2 * pr < n; pr += stride) {
    const uint32_t bits = dropout_bits(seed, (uint64_t)pr);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int64_t i = 2 * pr + j;
      if (i < n) {
        float v = ((bits >> (16 * j)) & 0xffffu) >= thr ? ld_as_float(x, i, dt) * scale : 0.0f;
        if (res) v += ld_as_float(res, i, res_dt);
        st_from_float(out, i, out_dt, v);
      }
    }
  }
}
__global__ void dropout_mask_kernel(int64_t n, uint32_t thr, uint64_t seed, uint8_t* __restrict__ keep) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    keep[i] = dropout_keep(seed, (uint64_t)i, thr) ? 1 : 0;
}
}  // namespace acai

extern "C" int acai_dropout(const void* x, int32_t dtype, const void* residual, int32_t res_dtype, int64_t n, float p,
                            uint64_t seed, void* out, int32_t out_dtype, void* stream) {
  ACAI_CHECK_ARG(x && out && n >= 0 && p >= 0.0f && p < 1.0f && dtype_valid(dtype) && dtype_valid(out_dtype) &&
                 (!residual || dtype_valid(res_dtype)), "acai_dropout: bad argument");
  if (n == 0) return ACAI_OK;
  auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  if (dtype == ACAI_BF16 && out_dtype == ACAI_BF16 && (!residual || res_dtype == ACAI_BF16) && n % 8 == 0 && al(x) && al(out) &&
      (!residual || al(residual))) {
    acai::dropout_bf16x8_kernel<<<grid_for(n / 8), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(x), reinterpret_cast<const uint4*>(residual), reinterpret_cast<uint4*>(out), n / 8,
        dropout_threshold(p), 1.0f / (1.0f - p), seed);
    return launch_status("dropout");
  }
  acai::dropout_kernel<<<grid_for((n + 1) / 2), 256, 0, (cudaStream_t)stream>>>(x, dtype, residual, res_dtype, n, dropout_threshold(p),
                                                                               1.0f / (1.0f - p), seed, out, out_dtype);
  return launch_status("dropout");
}

extern "C" int acai_dropout_mask(int64_t n, float p, uint64_t seed, uint8_t* keep, void* stream) {
  ACAI_CHECK_ARG(keep && n >= 0 && p >= 0.0f && p < 1.0f, "acai_dropout_mask: bad argument");
  if (n == 0) return ACAI_OK;
  acai::dropout_mask_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(n, dropout_threshold(p), seed, keep);
  return launch_status("dropout_mask");
}
