// K3 backward (first version: fp32 math on the CUDA cores, any head_dim <= 128, bf16 or fp32 storage).
// Flash-style recomputation from the forward's per-row log-sum-exp (log2 domain):
//     P = exp2(S * scale_log2 - lse),  dP = dO V^T,  dS = P * (dP - delta) * scale,  delta = rowsum(dO * O)
//     dQ = dS K        (attn_bwd_dq_kernel : query rows stationary, loops over key tiles)
//     dK = dS^T Q, dV = P^T dO   (attn_bwd_dkv_kernel: key rows stationary, loops over query tiles)
// Two lanes share a row (each holds half of the head dimension in registers), so no atomics are needed and the
// results are deterministic.  The tensor-core (tcgen05) version of this kernel is the next step;
// this one exists so that the training path is correct end to end first.
#include "common.cuh"

namespace acai {

constexpr int BR = 64;   // stationary rows per CTA (2 lanes each -> 128 threads)
constexpr int BT = 32;   // streamed rows per shared-memory tile

struct AttnBwdArgs {
  const void* q; const void* k; const void* v; int dt; int64_t qs, ks, vs;
  const void* o; int64_t os; const void* dout; int64_t dos;
  const float* lse; const float* delta;
  const int32_t* cu_q; const int32_t* cu_k; int total_q; int heads; int d; int causal; float scale, scale_log2;
  void* dq; void* dk; void* dv; int gdt; int64_t dqs, dks, dvs;
  AttnDropout drop;
};

// delta[h, row] = sum_e dO[row, h, e] * O[row, h, e]; one warp per (row, head)
__global__ void attn_delta_kernel(AttnBwdArgs a) {
  const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (w >= a.total_q * a.heads) return;
  const int row = w / a.heads, h = w - row * a.heads;
  float s = 0.0f;
  for (int e = lane; e < a.d; e += 32)
    s += ld_as_float(a.dout, (int64_t)row * a.dos + h * a.d + e, a.dt) * ld_as_float(a.o, (int64_t)row * a.os + h * a.d + e, a.dt);
  s = warp_sum(s);
  if (lane == 0) const_cast<float*>(a.delta)[(int64_t)h * a.total_q + row] = s;
}

// bf16 fast path: one 16-byte piece of a row per thread (a warp reads 512 contiguous bytes of dO and of O), the G = d / 8
// lanes of one (row, head) folded by shuffles; grid-stride.
template <int G>
__global__ void __launch_bounds__(256)
attn_delta_bf16_kernel(AttnBwdArgs a) {
  const int per_row = a.heads * G;
  const int64_t items = (int64_t)a.total_q * per_row;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < items; base += stride) {
    const int64_t i = base + (threadIdx.x & 31);
    float s = 0.0f;
    int row = 0, c8 = 0;
    if (i < items) {
      row = (int)(i / per_row); c8 = (int)(i - (int64_t)row * per_row);
      const uint4 u = __ldcs(reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(a.dout) + (int64_t)row * a.dos + c8 * 8));
      const uint4 w = __ldcs(reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(a.o) + (int64_t)row * a.os + c8 * 8));
      const __nv_bfloat162* hu = reinterpret_cast<const __nv_bfloat162*>(&u);
      const __nv_bfloat162* hw = reinterpret_cast<const __nv_bfloat162*>(&w);
#pragma unroll
      for (int e = 0; e < 4; ++e) { const float2 x = __bfloat1622float2(hu[e]), y = __bfloat1622float2(hw[e]); s = fmaf(x.x, y.x, fmaf(x.y, y.y, s)); }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (i < items && (c8 % G) == 0) const_cast<float*>(a.delta)[(int64_t)(c8 / G) * a.total_q + row] = s;
  }
}

template <int DMAX>
__device__ __forceinline__ float pair_dot(const float (&x)[DMAX / 2], const float* __restrict__ y) {
  float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
  for (int e = 0; e < DMAX / 2; e += 4) {
    const float4 t = *reinterpret_cast<const float4*>(y + e);
    s0 = fmaf(x[e], t.x, s0); s1 = fmaf(x[e + 1], t.y, s1); s0 = fmaf(x[e + 2], t.z, s0); s1 = fmaf(x[e + 3], t.w, s1);
  }
  const float s = s0 + s1;
  return s + __shfl_xor_sync(0xffffffffu, s, 1);
}

template <int DMAX>
__global__ void __launch_bounds__(2 * BR)
attn_bwd_dq_kernel(AttnBwdArgs a) {
  constexpr int HALF = DMAX / 2;
  __shared__ __align__(16) float Ks[BT][DMAX];
  __shared__ __align__(16) float Vs[BT][DMAX];
  const int b = blockIdx.z, h = blockIdx.y, tid = threadIdx.x, sub = tid & 1;
  const int q0 = a.cu_q[b], lq = a.cu_q[b + 1] - q0, k0 = a.cu_k[b], lk = a.cu_k[b + 1] - k0;
  if (blockIdx.x * BR >= lq) return;
  const int row = blockIdx.x * BR + (tid >> 1);
  const bool valid = row < lq;
  const int shift = lk - lq, d = a.d;
  float q[HALF], g[HALF], dq[HALF];
#pragma unroll
  for (int e = 0; e < HALF; ++e) {
    const int c = sub * HALF + e;
    const bool in = valid && c < d;
    q[e] = in ? ld_as_float(a.q, (int64_t)(q0 + row) * a.qs + h * d + c, a.dt) : 0.0f;
    g[e] = in ? ld_as_float(a.dout, (int64_t)(q0 + row) * a.dos + h * d + c, a.dt) : 0.0f;
    dq[e] = 0.0f;
  }
  const float lse = valid ? a.lse[(int64_t)h * a.total_q + q0 + row] : INFINITY;
  const float delta = valid ? a.delta[(int64_t)h * a.total_q + q0 + row] : 0.0f;
  int kv_end = lk;
  if (a.causal) { const int last = blockIdx.x * BR + BR + shift; kv_end = last < lk ? (last < 0 ? 0 : last) : lk; }
  const int my_end = a.causal ? min(lk, row + shift + 1) : lk;
  for (int t0 = 0; t0 < kv_end; t0 += BT) {
    __syncthreads();
    for (int i = tid; i < BT * DMAX; i += 2 * BR) {
      const int r = i / DMAX, c = i - r * DMAX;
      const bool in = t0 + r < lk && c < d;
      Ks[r][c] = in ? ld_as_float(a.k, (int64_t)(k0 + t0 + r) * a.ks + h * d + c, a.dt) : 0.0f;
      Vs[r][c] = in ? ld_as_float(a.v, (int64_t)(k0 + t0 + r) * a.vs + h * d + c, a.dt) : 0.0f;
    }
    __syncthreads();
#pragma unroll 2
    for (int r = 0; r < BT; ++r) {
      const float s = pair_dot<DMAX>(q, &Ks[r][sub * HALF]);
      const float dp = pair_dot<DMAX>(g, &Vs[r][sub * HALF]);
      const float p = (t0 + r < my_end) ? fast_exp2(s * a.scale_log2 - lse) : 0.0f;
      float mk = 1.0f;
      if (a.drop.thr) mk = attn_drop_keep(a.drop, (uint64_t)h * a.total_q + (uint64_t)(q0 + (valid ? row : 0)), t0 + r) ? a.drop.rp : 0.0f;
      const float ds = p * (mk * dp - delta) * a.scale;
#pragma unroll
      for (int e = 0; e < HALF; ++e) dq[e] = fmaf(ds, Ks[r][sub * HALF + e], dq[e]);
    }
  }
  if (valid) {
#pragma unroll
    for (int e = 0; e < HALF; ++e) {
      const int c = sub * HALF + e;
      if (c < d) st_from_float(a.dq, (int64_t)(q0 + row) * a.dqs + h * d + c, a.gdt, dq[e]);
    }
  }
}

template <int DMAX>
__global__ void __launch_bounds__(2 * BR)
attn_bwd_dkv_kernel(AttnBwdArgs a) {
  constexpr int HALF = DMAX / 2;
  __shared__ __align__(16) float Qs[BT][DMAX];
  __shared__ __align__(16) float Gs[BT][DMAX];
  __shared__ float Ls[BT], Ds[BT];
  const int b = blockIdx.z, h = blockIdx.y, tid = threadIdx.x, sub = tid & 1;
  const int q0 = a.cu_q[b], lq = a.cu_q[b + 1] - q0, k0 = a.cu_k[b], lk = a.cu_k[b + 1] - k0;
  if (blockIdx.x * BR >= lk) return;
  const int key = blockIdx.x * BR + (tid >> 1);
  const bool valid = key < lk;
  const int shift = lk - lq, d = a.d;
  float kk[HALF], vv[HALF], dk[HALF], dv[HALF];
#pragma unroll
  for (int e = 0; e < HALF; ++e) {
    const int c = sub * HALF + e;
    const bool in = valid && c < d;
    kk[e] = in ? ld_as_float(a.k, (int64_t)(k0 + key) * a.ks + h * d + c, a.dt) : 0.0f;
    vv[e] = in ? ld_as_float(a.v, (int64_t)(k0 + key) * a.vs + h * d + c, a.dt) : 0.0f;
    dk[e] = 0.0f; dv[e] = 0.0f;
  }
  // causal: query i sees key j iff j <= i + shift  <=>  i >= j - shift; the CTA's first key bounds the first query tile
  int q_begin = 0;
  if (a.causal) { q_begin = blockIdx.x * BR - shift; q_begin = q_begin < 0 ? 0 : (q_begin / BT) * BT; }
  const int my_first = a.causal ? key - shift : 0;     // first query row that sees this key
  for (int t0 = q_begin; t0 < lq; t0 += BT) {
    __syncthreads();
    for (int i = tid; i < BT * DMAX; i += 2 * BR) {
      const int r = i / DMAX, c = i - r * DMAX;
      const bool in = t0 + r < lq && c < d;
      Qs[r][c] = in ? ld_as_float(a.q, (int64_t)(q0 + t0 + r) * a.qs + h * d + c, a.dt) : 0.0f;
      Gs[r][c] = in ? ld_as_float(a.dout, (int64_t)(q0 + t0 + r) * a.dos + h * d + c, a.dt) : 0.0f;
    }
    if (tid < BT) {
      const bool in = t0 + tid < lq;
      Ls[tid] = in ? a.lse[(int64_t)h * a.total_q + q0 + t0 + tid] : INFINITY;
      Ds[tid] = in ? a.delta[(int64_t)h * a.total_q + q0 + t0 + tid] : 0.0f;
    }
    __syncthreads();
#pragma unroll 2
    for (int r = 0; r < BT; ++r) {
      const float s = pair_dot<DMAX>(kk, &Qs[r][sub * HALF]);
      const float dp = pair_dot<DMAX>(vv, &Gs[r][sub * HALF]);
      const float p = (valid && t0 + r < lq && t0 + r >= my_first) ? fast_exp2(s * a.scale_log2 - Ls[r]) : 0.0f;
      float mk = 1.0f;
      if (a.drop.thr) mk = attn_drop_keep(a.drop, (uint64_t)h * a.total_q + (uint64_t)(q0 + min(t0 + r, lq - 1)), valid ? key : 0) ? a.drop.rp : 0.0f;
      const float ds = p * (mk * dp - Ds[r]) * a.scale;
      const float pm = p * mk;
#pragma unroll
      for (int e = 0; e < HALF; ++e) {
        dv[e] = fmaf(pm, Gs[r][sub * HALF + e], dv[e]);
        dk[e] = fmaf(ds, Qs[r][sub * HALF + e], dk[e]);
      }
    }
  }
  if (valid) {
#pragma unroll
    for (int e = 0; e < HALF; ++e) {
      const int c = sub * HALF + e;
      if (c < d) {
        st_from_float(a.dk, (int64_t)(k0 + key) * a.dks + h * d + c, a.gdt, dk[e]);
        st_from_float(a.dv, (int64_t)(k0 + key) * a.dvs + h * d + c, a.gdt, dv[e]);
      }
    }
  }
}

int attn_bwd_tc(const void* q, const void* k, const void* v, int dtype, int64_t qs, int64_t ks, int64_t vs, const void* dout,
                int64_t dos, const float* lse, const float* delta, const int32_t* cu_q, const int32_t* cu_k, int batch,
                int max_q, int max_k, int64_t total_q, int64_t total_k, int heads, int head_dim, int causal, float scale,
                void* dq, void* dk, void* dv, int grad_dtype, int64_t dqs, int64_t dks, int64_t dvs, AttnDropout drop,
                cudaStream_t st);

}  // namespace acai

using namespace acai;

extern "C" int acai_attn_varlen_bwd(const void* q, const void* k, const void* v, int32_t dtype, int64_t q_stride,
                                    int64_t k_stride, int64_t v_stride, const void* out, int64_t o_stride,
                                    const void* dout, int64_t do_stride, const float* lse, const int32_t* cu_seqlens_q,
                                    const int32_t* cu_seqlens_k, int32_t batch, int32_t max_seqlen_q, int32_t max_seqlen_k,
                                    int32_t total_q, int32_t total_k, int32_t heads, int32_t head_dim, int32_t causal, float scale,
                                    void* dq, void* dk, void* dv, int32_t grad_dtype, int64_t dq_stride, int64_t dk_stride,
                                    int64_t dv_stride, float* delta_ws, float dropout_p, uint64_t dropout_seed, void* stream) {
  ACAI_CHECK_ARG(q && k && v && out && dout && lse && cu_seqlens_q && cu_seqlens_k && dq && dk && dv && delta_ws,
                 "acai_attn_varlen_bwd: null pointer");
  ACAI_CHECK_ARG(batch > 0 && heads > 0 && head_dim > 0 && total_q > 0 && batch <= 65535 && heads <= 65535,
                 "acai_attn_varlen_bwd: bad sizes");
  ACAI_CHECK_ARG(dtype_valid(dtype) && dtype_valid(grad_dtype), "acai_attn_varlen_bwd: bad dtype");
  if (head_dim > 128) { set_error("acai_attn_varlen_bwd: head_dim %d > 128 unsupported", head_dim); return ACAI_ERR_UNSUPPORTED; }
  ACAI_CHECK_ARG(dropout_p >= 0.0f && dropout_p < 1.0f, "acai_attn_varlen_bwd: dropout_p out of range");
  const AttnDropout drop{dropout_seed, dropout_threshold(dropout_p), 1.0f / (1.0f - dropout_p)};
  cudaStream_t st = (cudaStream_t)stream;
  AttnBwdArgs a{q, k, v, dtype, q_stride, k_stride, v_stride, out, o_stride, dout, do_stride, lse, delta_ws,
                cu_seqlens_q, cu_seqlens_k, total_q, heads, head_dim, causal, scale, scale * 1.4426950408889634f,
                dq, dk, dv, grad_dtype, dq_stride, dk_stride, dv_stride, drop};
  const int64_t nw = (int64_t)total_q * heads;
  const bool delta_fast = dtype == ACAI_BF16 && (head_dim == 32 || head_dim == 64 || head_dim == 128) && do_stride % 8 == 0 &&
                          o_stride % 8 == 0 && !((reinterpret_cast<uintptr_t>(dout) | reinterpret_cast<uintptr_t>(out)) & 15);
  if (delta_fast) {
    const int64_t items = nw * (head_dim / 8);
    const int64_t want = (items + 255) / 256, cap = (int64_t)acai::num_sms() * 8;
    const unsigned g = (unsigned)(want < cap ? want : cap);
    if (head_dim == 32) attn_delta_bf16_kernel<4><<<g, 256, 0, st>>>(a);
    else if (head_dim == 64) attn_delta_bf16_kernel<8><<<g, 256, 0, st>>>(a);
    else attn_delta_bf16_kernel<16><<<g, 256, 0, st>>>(a);
  } else
    attn_delta_kernel<<<(unsigned)((nw + 7) / 8), 256, 0, st>>>(a);
  int rc = launch_status("attn_delta");
  if (rc) return rc;
  rc = acai::attn_bwd_tc(q, k, v, dtype, q_stride, k_stride, v_stride, dout, do_stride, lse, delta_ws, cu_seqlens_q, cu_seqlens_k,
                   batch, max_seqlen_q, max_seqlen_k, total_q, total_k, heads, head_dim, causal, scale, dq, dk, dv, grad_dtype,
                   dq_stride, dk_stride, dv_stride, drop, st);
  if (rc != ACAI_ERR_UNSUPPORTED) return rc;
  dim3 gq((max_seqlen_q + BR - 1) / BR, heads, batch), gk((max_seqlen_k + BR - 1) / BR, heads, batch);
  if (head_dim <= 32) { attn_bwd_dq_kernel<32><<<gq, 2 * BR, 0, st>>>(a); attn_bwd_dkv_kernel<32><<<gk, 2 * BR, 0, st>>>(a); }
  else if (head_dim <= 64) { attn_bwd_dq_kernel<64><<<gq, 2 * BR, 0, st>>>(a); attn_bwd_dkv_kernel<64><<<gk, 2 * BR, 0, st>>>(a); }
  else { attn_bwd_dq_kernel<128><<<gq, 2 * BR, 0, st>>>(a); attn_bwd_dkv_kernel<128><<<gk, 2 * BR, 0, st>>>(a); }
  return launch_status("attn_varlen_bwd");
}
