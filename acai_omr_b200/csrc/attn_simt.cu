// Generic (any head_dim <= 128, fp32 or bf16 storage, fp32 math) attention kernels:
//   * attn_varlen_simt : flash-style online-softmax over packed ragged sequences (cu_seqlens),
//                        one thread per query row, K/V tiles staged in shared memory.
//   * decode_attn_simt : one query per (sequence, head) against a strided (B,H,T,d) cache.
// These are the exact-fp32 / odd-shape path.  The tensor-core (tcgen05) varlen kernel and the
// paged split-KV decode kernel for head_dim 64/32 live in attn_tc.cu / decode_paged.cu.
#include <cstdlib>

#include "common.cuh"

namespace acai {

constexpr int QB = 64;   // query rows per CTA (one per thread)
constexpr int KT = 32;   // keys per shared-memory tile
constexpr int KC = 8;    // keys per online-softmax rescale

template <int DMAX>
__global__ void __launch_bounds__(QB)
attn_varlen_simt(const void* __restrict__ q, const void* __restrict__ k, const void* __restrict__ v, int dt,
                 int64_t qs, int64_t ks, int64_t vs, const int32_t* __restrict__ cu_q,
                 const int32_t* __restrict__ cu_k, int d, int causal, float scale,
                 void* __restrict__ out, int out_dt, int64_t os, float* __restrict__ lse, int total_q, AttnDropout drop) {
  __shared__ float Ks[KT][DMAX];
  __shared__ float Vs[KT][DMAX];
  const int b = blockIdx.z, h = blockIdx.y, qb = blockIdx.x, tid = threadIdx.x;
  const int q0 = cu_q[b], lq = cu_q[b + 1] - q0, k0 = cu_k[b], lk = cu_k[b + 1] - k0;
  if (qb * QB >= lq) return;
  const int row = qb * QB + tid;
  const bool valid = row < lq;
  const int shift = lk - lq;  // causal: key j visible to query i iff j <= i + shift
  float qr[DMAX], acc[DMAX];
#pragma unroll
  for (int e = 0; e < DMAX; ++e) {
    qr[e] = (valid && e < d) ? ld_as_float(q, (int64_t)(q0 + row) * qs + (int64_t)h * d + e, dt) * scale : 0.0f;
    acc[e] = 0.0f;
  }
  // columns d..DMAX-1 of the tiles are never loaded; zero them once so q[e]=0 times them stays 0
  for (int i = tid; i < KT * DMAX; i += QB) { (&Ks[0][0])[i] = 0.0f; (&Vs[0][0])[i] = 0.0f; }
  __syncthreads();
  float m = -INFINITY, l = 0.0f;
  int kv_end = lk;
  if (causal) { int last = qb * QB + QB - 1 + shift + 1; kv_end = last < lk ? (last < 0 ? 0 : last) : lk; }
  const int my_end = causal ? min(lk, row + shift + 1) : lk;  // exclusive key bound for this row

  for (int t0 = 0; t0 < kv_end; t0 += KT) {
    const int nt = min(KT, kv_end - t0);
    for (int i = tid; i < KT * d; i += QB) {
      const int j = i / d, e = i - j * d;
      float kv = 0.0f, vv = 0.0f;
      if (j < nt) {
        kv = ld_as_float(k, (int64_t)(k0 + t0 + j) * ks + (int64_t)h * d + e, dt);
        vv = ld_as_float(v, (int64_t)(k0 + t0 + j) * vs + (int64_t)h * d + e, dt);
      }
      Ks[j][e] = kv; Vs[j][e] = vv;
    }
    __syncthreads();
    for (int c0 = 0; c0 < nt; c0 += KC) {
      float s[KC]; float cmax = -INFINITY;
#pragma unroll
      for (int c = 0; c < KC; ++c) {
        const int j = c0 + c;
        float dot = 0.0f;
#pragma unroll
        for (int e = 0; e < DMAX; ++e) dot = fmaf(qr[e], Ks[j < KT ? j : KT - 1][e], dot);
        s[c] = (j < nt && t0 + j < my_end) ? dot : -INFINITY;
        cmax = fmaxf(cmax, s[c]);
      }
      if (cmax > -INFINITY) {
        const float m_new = fmaxf(m, cmax), corr = expf(m - m_new);
        l *= corr;
#pragma unroll
        for (int e = 0; e < DMAX; ++e) acc[e] *= corr;
#pragma unroll
        for (int c = 0; c < KC; ++c) {
          const int j = c0 + c;
          const float p = (s[c] > -INFINITY) ? expf(s[c] - m_new) : 0.0f;
          l += p;                                   // the normaliser sums the un-dropped probabilities
          float pd = p;
          if (drop.thr && p > 0.0f && !attn_drop_keep(drop, (uint64_t)h * total_q + (uint64_t)(q0 + (valid ? row : 0)), t0 + j)) pd = 0.0f;
#pragma unroll
          for (int e = 0; e < DMAX; ++e) acc[e] = fmaf(pd, Vs[j < KT ? j : KT - 1][e], acc[e]);
        }
        m = m_new;
      }
    }
    __syncthreads();
  }
  if (valid) {
    const float inv = l > 0.0f ? (drop.thr ? drop.rp : 1.0f) / l : 0.0f;
#pragma unroll
    for (int e = 0; e < DMAX; ++e)
      if (e < d) st_from_float(out, (int64_t)(q0 + row) * os + (int64_t)h * d + e, out_dt, acc[e] * inv);
    if (lse) lse[(int64_t)h * total_q + q0 + row] = l > 0.0f ? (m + logf(l)) * 1.4426950408889634f : INFINITY;   // log2 domain
  }
}

// One CTA per (b, h).  Scores for all keys are kept in dynamic shared memory.
__global__ void __launch_bounds__(128)
decode_attn_simt(const void* __restrict__ q, int q_dt, int64_t q_bs, const void* __restrict__ k,
                 const void* __restrict__ v, int kv_dt, int64_t bs, int64_t hs, int64_t ts, int d, int kv_len,
                 const uint8_t* __restrict__ mask, int64_t mask_stride, float scale, void* __restrict__ out,
                 int out_dt, int64_t o_bs) {
  extern __shared__ float sm[];
  __shared__ float red[32];
  float* sc = sm;               // [kv_len]
  float* qs = sm + kv_len;      // [d]
  float* part = qs + d;         // [ng][d]
  const int b = blockIdx.x, h = blockIdx.y, tid = threadIdx.x;
  for (int e = tid; e < d; e += blockDim.x) qs[e] = ld_as_float(q, (int64_t)b * q_bs + (int64_t)h * d + e, q_dt) * scale;
  __syncthreads();
  const int64_t base = (int64_t)b * bs + (int64_t)h * hs;
  float lmax = -INFINITY;
  for (int j = tid; j < kv_len; j += blockDim.x) {
    float s = -INFINITY;
    if (!(mask && mask[(int64_t)b * mask_stride + j])) {
      s = 0.0f;
      for (int e = 0; e < d; ++e) s = fmaf(qs[e], ld_as_float(k, base + (int64_t)j * ts + e, kv_dt), s);
    }
    sc[j] = s; lmax = fmaxf(lmax, s);
  }
  const float mx = block_max(lmax, red);
  float lsum = 0.0f;
  for (int j = tid; j < kv_len; j += blockDim.x) {
    const float p = (sc[j] > -INFINITY) ? expf(sc[j] - mx) : 0.0f;
    sc[j] = p; lsum += p;
  }
  const float total = block_sum(lsum, red);  // includes the barrier that publishes sc[]
  const int ng = max(1, (int)blockDim.x / d), g = tid / d, e = tid - g * d;
  if (g < ng) {
    float a = 0.0f;
    for (int j = g; j < kv_len; j += ng) a = fmaf(sc[j], ld_as_float(v, base + (int64_t)j * ts + e, kv_dt), a);
    part[g * d + e] = a;
  }
  __syncthreads();
  for (int e2 = tid; e2 < d; e2 += blockDim.x) {
    float a = 0.0f;
    for (int g2 = 0; g2 < ng; ++g2) a += part[g2 * d + e2];
    st_from_float(out, (int64_t)b * o_bs + (int64_t)h * d + e2, out_dt, total > 0.0f ? a / total : 0.0f);
  }
}

int attn_varlen_tc(const void* q, const void* k, const void* v, int dtype, int64_t qs, int64_t ks, int64_t vs,
                   const int32_t* cu_q, const int32_t* cu_k, int batch, int max_q, int64_t total_q, int64_t total_k,
                   int heads, int head_dim, int causal, float scale, void* out, int out_dt, int64_t os, float* lse,
                   AttnDropout drop, cudaStream_t st);
}  // namespace acai

using namespace acai;

extern "C" int acai_attn_varlen_fwd(const void* q, const void* k, const void* v, int32_t dtype, int64_t q_stride,
                                    int64_t k_stride, int64_t v_stride, const int32_t* cu_seqlens_q,
                                    const int32_t* cu_seqlens_k, int32_t batch, int32_t max_seqlen_q, int32_t total_q,
                                    int32_t total_k, int32_t heads, int32_t head_dim, int32_t causal, float scale,
                                    void* out, int32_t out_dtype, int64_t o_stride, float* lse, float dropout_p,
                                    uint64_t dropout_seed, void* stream) {
  ACAI_CHECK_ARG(q && k && v && cu_seqlens_q && cu_seqlens_k && out, "acai_attn_varlen_fwd: null pointer");
  ACAI_CHECK_ARG(dropout_p >= 0.0f && dropout_p < 1.0f, "acai_attn_varlen_fwd: dropout_p out of range");
  const AttnDropout drop{dropout_seed, dropout_threshold(dropout_p), 1.0f / (1.0f - dropout_p)};
  ACAI_CHECK_ARG(batch > 0 && heads > 0 && head_dim > 0 && max_seqlen_q >= 0, "acai_attn_varlen_fwd: bad sizes");
  ACAI_CHECK_ARG(dtype_valid(dtype) && dtype_valid(out_dtype), "acai_attn_varlen_fwd: bad dtype");
  if (head_dim > 128) { set_error("acai_attn_varlen_fwd: head_dim %d > 128 unsupported", head_dim); return ACAI_ERR_UNSUPPORTED; }
  ACAI_CHECK_ARG(heads <= 65535 && batch <= 65535, "acai_attn_varlen_fwd: heads/batch too large for the grid");
  if (max_seqlen_q == 0 || total_q == 0) return ACAI_OK;
  cudaStream_t st = (cudaStream_t)stream;
  {
    int rc = attn_varlen_tc(q, k, v, dtype, q_stride, k_stride, v_stride, cu_seqlens_q, cu_seqlens_k, batch,
                            max_seqlen_q, total_q, total_k, heads, head_dim, causal, scale, out, out_dtype, o_stride, lse, drop, st);
    if (rc != ACAI_ERR_UNSUPPORTED) return rc;
  }
  dim3 grid((max_seqlen_q + QB - 1) / QB, heads, batch);
#define LAUNCH(D) attn_varlen_simt<D><<<grid, QB, 0, st>>>(q, k, v, dtype, q_stride, k_stride, v_stride, cu_seqlens_q, \
                                                           cu_seqlens_k, head_dim, causal, scale, out, out_dtype, o_stride, lse, total_q, drop)
  if (head_dim <= 16) LAUNCH(16);
  else if (head_dim <= 32) LAUNCH(32);
  else if (head_dim <= 64) LAUNCH(64);
  else LAUNCH(128);
#undef LAUNCH
  return launch_status("attn_varlen_simt");
}

extern "C" int acai_decode_attn(const void* q, int32_t q_dtype, int64_t q_batch_stride, const void* k, const void* v,
                                int32_t kv_dtype, int64_t kv_batch_stride, int64_t kv_head_stride,
                                int64_t kv_tok_stride, int32_t batch, int32_t heads, int32_t head_dim, int32_t kv_len,
                                const uint8_t* key_padding_mask, int64_t mask_stride, float scale, void* out,
                                int32_t out_dtype, int64_t o_batch_stride, void* workspace, size_t workspace_bytes,
                                void* stream) {
  (void)workspace; (void)workspace_bytes;
  ACAI_CHECK_ARG(q && k && v && out, "acai_decode_attn: null pointer");
  ACAI_CHECK_ARG(batch > 0 && heads > 0 && head_dim > 0 && kv_len > 0, "acai_decode_attn: bad sizes");
  ACAI_CHECK_ARG(dtype_valid(q_dtype) && dtype_valid(kv_dtype) && dtype_valid(out_dtype), "acai_decode_attn: bad dtype");
  ACAI_CHECK_ARG(heads <= 65535, "acai_decode_attn: too many heads");
  if (head_dim > 128) { set_error("acai_decode_attn: head_dim %d > 128 unsupported", head_dim); return ACAI_ERR_UNSUPPORTED; }
  const int ng = 128 / head_dim > 0 ? 128 / head_dim : 1;
  const size_t smem = sizeof(float) * ((size_t)kv_len + head_dim + (size_t)ng * head_dim);
  if (smem > 200 * 1024) { set_error("acai_decode_attn: kv_len %d too long for the generic kernel", kv_len); return ACAI_ERR_UNSUPPORTED; }
  if (smem > 48 * 1024)
    ACAI_CUDA(cudaFuncSetAttribute(decode_attn_simt, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  decode_attn_simt<<<dim3(batch, heads), 128, smem, (cudaStream_t)stream>>>(
      q, q_dtype, q_batch_stride, k, v, kv_dtype, kv_batch_stride, kv_head_stride, kv_tok_stride, head_dim, kv_len,
      key_padding_mask, mask_stride, scale, out, out_dtype, o_batch_stride);
  return launch_status("decode_attn_simt");
}


// Test / inspection aid: the keep decisions of the attention-probability dropout, keep[h, q, k] for global query rows q
// and key positions k < max_k inside their sequence - exactly what every attention kernel regenerates.
namespace acai {
__global__ void attn_dropout_mask_kernel(AttnDropout drop, int heads, int total_q, int max_k, uint8_t* __restrict__ keep) {
  const int64_t n = (int64_t)heads * total_q * max_k;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int k = (int)(i % max_k);
    const int64_t rowkey = i / max_k;                 // = h * total_q + q
    keep[i] = attn_drop_keep(drop, (uint64_t)rowkey, k) ? 1 : 0;
  }
}
}  // namespace acai

extern "C" int acai_attn_dropout_mask(float dropout_p, uint64_t dropout_seed, int32_t heads, int32_t total_q, int32_t max_k,
                                      uint8_t* keep, void* stream) {
  ACAI_CHECK_ARG(keep && heads > 0 && total_q > 0 && max_k > 0 && dropout_p >= 0.0f && dropout_p < 1.0f,
                 "acai_attn_dropout_mask: bad argument");
  const acai::AttnDropout drop{dropout_seed, acai::dropout_threshold(dropout_p), 1.0f / (1.0f - dropout_p)};
  acai::attn_dropout_mask_kernel<<<1184, 256, 0, (cudaStream_t)stream>>>(drop, heads, total_q, max_k, keep);
  return launch_status("attn_dropout_mask");
}
