// K4: the decode-step attention over a (paged) token-major KV cache - HBM-bound by design.
//
// Cache row layout (both the self-attention pool and the encoder-memory cache): one row per token,
//     [ K(head 0..H-1, 64 each) | V(head 0..H-1, 64 each) ]     = 2*H*64 elements
// which is exactly the [K|V] slice the fused QKV / cross-KV projections emit, so appending a token is a row copy and
// the cross cache IS the projection GEMM's output buffer.  Rows are addressed either through a page table
// (self cache: PAGE tokens per page) or as seq_start[b] + t (memory cache: packed, no padding rows).
//
// Grid (chunks, B): a CTA streams CHUNK consecutive tokens of one sequence for ALL heads (warp = head), so its
// reads are whole contiguous rows.  Inside a warp, 64/VEC lanes cover one token's 64-dim head slice with 16-byte
// loads and 32/(64/VEC) tokens are processed per instruction; each lane group keeps its own online-softmax state
// (warp shuffles only for the 64-dim dot product), groups are merged once at the end, and the per-chunk partial
// (acc, m, l) goes to a workspace that decode_combine folds.  The CTA that owns position `len` also appends the new
// token's K|V row to the cache (KVCache.update, kv_caching.py:57-109).
#include "common.cuh"

namespace acai {

constexpr int DCH = 256;    // tokens per CTA
constexpr int DPAGE = 64;   // tokens per page of the self-attention pool
constexpr int DHD = 64;     // head dim

template <typename T> struct Vec16;
template <> struct Vec16<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ static void load(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 raw = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
  }
  __device__ static void copy(__nv_bfloat16* dst, const __nv_bfloat16* src) {
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
  }
};
template <> struct Vec16<float> {
  static constexpr int N = 4;
  __device__ static void load(const float* p, float (&v)[4]) {
    const float4 f = *reinterpret_cast<const float4*>(p);
    v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
  }
  __device__ static void copy(float* dst, const float* src) {
    *reinterpret_cast<float4*>(dst) = *reinterpret_cast<const float4*>(src);
  }
};

struct DecodeArgs {
  const void* q; int64_t q_stride;              // [B, >= H*64] query rows
  const void* new_kv; int64_t new_kv_stride;    // [B, >= 2*H*64] this step's [K|V] rows, or null
  void* pool;                                   // cache rows, 2*H*64 elements each
  const int32_t* page_table; int table_stride;  // [B, table_stride] page ids, or null
  const int32_t* seq_start;                     // [B] first row of each sequence when page_table is null
  const int32_t* kv_lens;                       // [B] cached tokens per sequence, or null ...
  const int32_t* step_ptr; int len_offset;      // ... then len = *step_ptr + len_offset for every sequence
  int heads; float scale_log2;
  float* ws; int max_chunks;                    // partials: [B][max_chunks][H][64 + 2]
};

template <typename T>
__global__ void __launch_bounds__(512)
decode_attn_paged_kernel(DecodeArgs a) {
  constexpr int VEC = Vec16<T>::N, LPT = DHD / VEC, GROUPS = 32 / LPT, UNROLL = 4;
  const int b = blockIdx.y, chunk = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = blockDim.x >> 5;
  const int len = a.kv_lens ? a.kv_lens[b] : (*a.step_ptr + a.len_offset);
  const int total = len + (a.new_kv ? 1 : 0);
  const int t_begin = chunk * DCH;
  if (t_begin >= total) return;
  const int t_end = min(total, t_begin + DCH);
  const int g = lane / LPT, sub = lane % LPT;
  const int H = a.heads, ROW = 2 * H * DHD;
  T* pool = reinterpret_cast<T*>(a.pool);
  const T* newkv = a.new_kv ? reinterpret_cast<const T*>(a.new_kv) + (int64_t)b * a.new_kv_stride : nullptr;
  const int32_t* tbl = a.page_table ? a.page_table + (int64_t)b * a.table_stride : nullptr;
  const int64_t row_base = tbl ? 0 : (int64_t)a.seq_start[b];

  for (int h = warp; h < H; h += nwarps) {
    float q[VEC];
    Vec16<T>::load(reinterpret_cast<const T*>(a.q) + (int64_t)b * a.q_stride + h * DHD + sub * VEC, q);
#pragma unroll
    for (int e = 0; e < VEC; ++e) q[e] *= a.scale_log2;
    float m = -INFINITY, l = 0.0f, acc[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[e] = 0.0f;
    const int koff = h * DHD + sub * VEC, voff = H * DHD + koff;

    for (int tb = t_begin; tb < t_end; tb += GROUPS * UNROLL) {   // warp-uniform trip count (shuffles inside)
      const int t0 = tb + g;
      float kf[UNROLL][VEC], vf[UNROLL][VEC];
      bool ok[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int t = t0 + u * GROUPS;
        ok[u] = t < t_end;
        if (ok[u]) {
          if (newkv && t == len) {                       // the token being appended this step
            const int64_t row = tbl ? (int64_t)tbl[t / DPAGE] * DPAGE + (t % DPAGE) : row_base + t;
            Vec16<T>::load(newkv + koff, kf[u]);
            Vec16<T>::load(newkv + voff, vf[u]);
            Vec16<T>::copy(pool + row * ROW + koff, newkv + koff);
            Vec16<T>::copy(pool + row * ROW + voff, newkv + voff);
          } else {
            const int64_t row = tbl ? (int64_t)tbl[t / DPAGE] * DPAGE + (t % DPAGE) : row_base + t;
            Vec16<T>::load(pool + row * ROW + koff, kf[u]);
            Vec16<T>::load(pool + row * ROW + voff, vf[u]);
          }
        }
      }
      float s[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        float d = 0.0f;
        if (ok[u]) {
#pragma unroll
          for (int e = 0; e < VEC; ++e) d = fmaf(q[e], kf[u][e], d);
        }
#pragma unroll
        for (int o = LPT / 2; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        s[u] = ok[u] ? d : -INFINITY;
      }
      float mx = m;
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) mx = fmaxf(mx, s[u]);
      if (mx > -INFINITY) {
        const float corr = fast_exp2(m - mx);
        l *= corr;
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[e] *= corr;
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          if (ok[u]) {
            const float p = fast_exp2(s[u] - mx);
            l += p;
#pragma unroll
            for (int e = 0; e < VEC; ++e) acc[e] = fmaf(p, vf[u][e], acc[e]);
          }
        }
        m = mx;
      }
    }
    // merge the GROUPS lane groups (same dims, different tokens)
#pragma unroll
    for (int o = LPT; o < 32; o <<= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, m, o), l2 = __shfl_xor_sync(0xffffffffu, l, o);
      const float mx = fmaxf(m, m2);
      const float c1 = mx > -INFINITY ? fast_exp2(m - mx) : 0.0f, c2 = mx > -INFINITY ? fast_exp2(m2 - mx) : 0.0f;
      l = l * c1 + l2 * c2;
#pragma unroll
      for (int e = 0; e < VEC; ++e) acc[e] = acc[e] * c1 + __shfl_xor_sync(0xffffffffu, acc[e], o) * c2;
      m = mx;
    }
    if (g == 0) {
      float* w = a.ws + (((int64_t)b * a.max_chunks + chunk) * H + h) * (DHD + 2);
#pragma unroll
      for (int e = 0; e < VEC; ++e) w[sub * VEC + e] = acc[e];
      if (sub == 0) { w[DHD] = m; w[DHD + 1] = l; }
    }
  }
}

// Fold the per-chunk partials: out[b, h*64+e] = sum_c acc_c * 2^(m_c - M) / sum_c l_c * 2^(m_c - M).
__global__ void __launch_bounds__(256)
decode_combine_kernel(const float* __restrict__ ws, int max_chunks, int chunk_tokens, int heads,
                      const int32_t* __restrict__ kv_lens, const int32_t* __restrict__ step_ptr, int len_offset,
                      int has_new, void* __restrict__ out, int out_dt, int64_t o_stride) {
  const int b = blockIdx.x;
  const int len = (kv_lens ? kv_lens[b] : (*step_ptr + len_offset)) + has_new;
  const int n_chunks = min(max_chunks, (len + chunk_tokens - 1) / chunk_tokens);
  for (int i = threadIdx.x; i < heads * DHD; i += blockDim.x) {
    const int h = i / DHD, e = i - h * DHD;
    const float* w = ws + ((int64_t)b * max_chunks * heads + h) * (DHD + 2);
    const int64_t cs = (int64_t)heads * (DHD + 2);
    float M = -INFINITY;
    for (int c = 0; c < n_chunks; ++c) M = fmaxf(M, w[c * cs + DHD]);
    float num = 0.0f, den = 0.0f;
    for (int c = 0; c < n_chunks; ++c) {
      const float mc = w[c * cs + DHD];
      if (mc > -INFINITY) {
        const float f = fast_exp2(mc - M);
        num = fmaf(w[c * cs + e], f, num);
        den = fmaf(w[c * cs + DHD + 1], f, den);
      }
    }
    st_from_float(out, (int64_t)b * o_stride + i, out_dt, den > 0.0f ? num / den : 0.0f);
  }
}

// ---- per-step bookkeeping kernels of the greedy loop (models.py:575-583, 603-610), all driven by a device-side
// step counter so that one captured CUDA graph can be replayed for every step ----
__global__ void decode_embed_kernel(const int64_t* __restrict__ seqs, int64_t seq_stride, const int32_t* __restrict__ step_ptr,
                                    const float* __restrict__ table, const float* __restrict__ pos,
                                    void* __restrict__ out, int out_dt, void* __restrict__ out2, int out2_dt, int dim) {
  const int b = blockIdx.x, t = *step_ptr;
  const int64_t tok = seqs[(int64_t)b * seq_stride + t - 1];      // token at index t-1 ...
  for (int c = threadIdx.x; c < dim; c += blockDim.x) {           // ... gets pos_embedding[t] (reference quirk, models.py:576)
    const float v = __ldg(table + tok * dim + c) + __ldg(pos + (int64_t)t * dim + c);
    st_from_float(out, (int64_t)b * dim + c, out_dt, v);
    if (out2) st_from_float(out2, (int64_t)b * dim + c, out2_dt, v);   // fp32 residual stream + its bf16 GEMM operand
  }
}

__global__ void greedy_update_kernel(const void* __restrict__ logits, int dt, int64_t ld, int rows, int vocab,
                                     const int32_t* __restrict__ step_ptr, int64_t* __restrict__ seqs, int64_t seq_stride,
                                     float* __restrict__ logp, uint8_t* __restrict__ finished, int eos) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  float best = -INFINITY; int idx = 0x7fffffff;
  for (int c = lane; c < vocab; c += 32) {
    const float v = ld_as_float(logits, (int64_t)row * ld + c, dt);
    if (v > best || (v == best && c < idx)) { best = v; idx = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o); const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (ob > best || (ob == best && oi < idx)) { best = ob; idx = oi; }
  }
  float s = 0.0f;
  for (int c = lane; c < vocab; c += 32) s += expf(ld_as_float(logits, (int64_t)row * ld + c, dt) - best);
  s = warp_sum(s);
  if (lane == 0) {
    const int t = *step_ptr;
    seqs[(int64_t)row * seq_stride + t] = idx;
    logp[(int64_t)row * seq_stride + t] = -logf(s);
    if (idx == eos) finished[row] = 1;
  }
}

// GRPO rollout step (GRPOViTOMR.cached_forward_rollout_policy, models.py:1003-1029): keep the top_k logits of the row,
// draw the next token from softmax(top_k_logits / temperature), record log_softmax(top_k_logits)[token] (WITHOUT the
// temperature, as the reference's cached path does).  One warp per row; the row is staged in shared memory.  Top-k
// membership is exact: element c is kept when fewer than top_k elements beat it (larger value, or equal value and lower
// index).  The draw is an inverse-CDF walk over the kept tokens in vocabulary order with a uniform u in [0, 1) that is
// either supplied (uniforms[row, step]: parity tests) or hashed from (*seed_ptr, row, step) - any walk order samples the
// same distribution as torch.multinomial; its random stream is not reproducible bit for bit.
__global__ void topk_sample_update_kernel(const void* __restrict__ logits, int dt, int64_t ld, int rows, int vocab, int top_k,
                                          float inv_temperature, const float* __restrict__ uniforms, int64_t u_stride,
                                          const uint64_t* __restrict__ seed_ptr, const int32_t* __restrict__ step_ptr,
                                          int64_t* __restrict__ seqs, int64_t seq_stride, float* __restrict__ logp,
                                          uint8_t* __restrict__ finished, int eos) {
  extern __shared__ float zs[];                                 // [warps][vocab]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + warp;
  if (row >= rows) return;
  float* z = zs + (size_t)warp * vocab;
  float mx = -INFINITY;
  for (int c = lane; c < vocab; c += 32) { const float v = ld_as_float(logits, (int64_t)row * ld + c, dt); z[c] = v; mx = fmaxf(mx, v); }
  mx = warp_max(mx);
  __syncwarp();
  // pass 1: rank every element against the row; the element of rank top_k - 1 is the membership threshold
  float sum_t = 0.0f, sum_1 = 0.0f, thr_v = -INFINITY;
  int thr_c = -1;
  for (int c = lane; c < vocab; c += 32) {
    const float v = z[c];
    int rank = 0;
    for (int j = 0; j < vocab; ++j) { const float o = z[j]; rank += (o > v || (o == v && j < c)) ? 1 : 0; }
    if (rank < top_k) { sum_t += __expf((v - mx) * inv_temperature); sum_1 += __expf(v - mx); }
    if (rank == top_k - 1) { thr_v = v; thr_c = c; }
  }
  sum_t = warp_sum(sum_t); sum_1 = warp_sum(sum_1);
  {
    const unsigned owner = __ballot_sync(0xffffffffu, thr_c >= 0);       // exactly one lane holds the threshold element
    const int src = __ffs(owner) - 1;
    thr_v = __shfl_sync(0xffffffffu, thr_v, src);
    thr_c = __shfl_sync(0xffffffffu, thr_c, src);
  }
  const int t = *step_ptr;
  float u;
  if (uniforms) u = uniforms[(int64_t)row * u_stride + t];
  else u = (float)(dropout_bits(*seed_ptr, ((uint64_t)row << 24) + (uint64_t)t) >> 8) * (1.0f / 16777216.0f);
  const float target = u * sum_t;
  // pass 2: inverse-CDF walk in vocabulary order
  float base = 0.0f;
  int chosen = -1, last_kept = -1;
  for (int c0 = 0; c0 < vocab && chosen < 0; c0 += 32) {
    const int c = c0 + lane;
    const float v = c < vocab ? z[c] : -INFINITY;
    const bool keep = c < vocab && (v > thr_v || (v == thr_v && c <= thr_c));
    float w = keep ? __expf((v - mx) * inv_temperature) : 0.0f;
    float cum = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const float n = __shfl_up_sync(0xffffffffu, cum, o); if (lane >= o) cum += n; }
    cum += base;
    const unsigned hit = __ballot_sync(0xffffffffu, keep && cum > target);
    const unsigned kept = __ballot_sync(0xffffffffu, keep);
    if (kept) last_kept = c0 + 31 - __clz(kept);
    if (hit) chosen = c0 + __ffs(hit) - 1;
    base = __shfl_sync(0xffffffffu, cum, 31);
  }
  if (chosen < 0) chosen = last_kept;                            // u * sum rounded past the last bucket
  if (lane == 0) {
    seqs[(int64_t)row * seq_stride + t] = chosen;
    logp[(int64_t)row * seq_stride + t] = (z[chosen] - mx) - __logf(sum_1);
    if (chosen == eos) finished[row] = 1;
  }
}

__global__ void advance_step_kernel(int32_t* step_ptr) { *step_ptr += 1; }

}  // namespace acai

using namespace acai;

namespace acai {
int decode_attn_tma(const void* q, int64_t q_stride, const void* new_kv, int64_t new_kv_stride, void* kv_rows,
                    int64_t kv_rows_total, const int32_t* page_table, int table_stride, const int32_t* seq_start,
                    const int32_t* kv_lens, const int32_t* step_ptr, int len_offset, int max_kv_len, int batch, int heads,
                    float scale, float* ws, int max_chunks, int chunk_tokens, int32_t* counters, void* out, int out_dt,
                    int64_t o_stride, cudaStream_t st);
// the workspace starts with one int32 arrival counter per (sequence, head) for the in-kernel split-KV combine
static inline size_t counter_bytes(int batch, int heads) { return (((size_t)batch * heads * sizeof(int32_t)) + 255) & ~(size_t)255; }
constexpr int DCH_MIN = 64;   // smallest split-KV chunk (small batches: more CTAs)
}

extern "C" size_t acai_decode_workspace_bytes(int32_t batch, int32_t heads, int32_t max_kv_len) {
  const size_t chunks = (size_t)(max_kv_len + DCH_MIN - 1) / DCH_MIN;
  return counter_bytes(batch, heads) + (size_t)batch * chunks * heads * (DHD + 2) * sizeof(float);
}

extern "C" int acai_decode_attn_paged(const void* q, int64_t q_stride, const void* new_kv, int64_t new_kv_stride,
                                      void* kv_rows, int32_t dtype, const int32_t* page_table, int32_t table_stride,
                                      const int32_t* seq_start, const int32_t* kv_lens, const int32_t* step_ptr,
                                      int32_t len_offset, int32_t max_kv_len, int32_t batch, int32_t heads,
                                      int32_t head_dim, float scale, void* out, int32_t out_dtype, int64_t o_stride,
                                      void* workspace, size_t workspace_bytes, int64_t kv_rows_total, void* stream) {
  ACAI_CHECK_ARG(q && kv_rows && out && workspace, "acai_decode_attn_paged: null pointer");
  ACAI_CHECK_ARG((page_table != nullptr) != (seq_start != nullptr), "acai_decode_attn_paged: give a page table or seq_start");
  ACAI_CHECK_ARG((kv_lens != nullptr) != (step_ptr != nullptr), "acai_decode_attn_paged: give kv_lens or step_ptr");
  ACAI_CHECK_ARG(batch > 0 && heads > 0 && max_kv_len > 0 && batch <= 65535, "acai_decode_attn_paged: bad sizes");
  ACAI_CHECK_ARG(dtype_valid(dtype) && dtype_valid(out_dtype), "acai_decode_attn_paged: bad dtype");
  if (head_dim != DHD) { set_error("acai_decode_attn_paged: head_dim %d unsupported (64 only)", head_dim); return ACAI_ERR_UNSUPPORTED; }
  const int al = dtype == ACAI_BF16 ? 8 : 4;
  ACAI_CHECK_ARG(q_stride % al == 0 && (!new_kv || new_kv_stride % al == 0) && !((uintptr_t)q & 15) &&
                 !((uintptr_t)kv_rows & 15) && !((uintptr_t)new_kv & 15), "acai_decode_attn_paged: operands must be 16-byte aligned");
  ACAI_CHECK_ARG(workspace_bytes >= acai_decode_workspace_bytes(batch, heads, max_kv_len), "acai_decode_attn_paged: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  int32_t* counters = reinterpret_cast<int32_t*>(workspace);
  float* ws = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + counter_bytes(batch, heads));
  int chunk_tokens = DCH, rc = ACAI_ERR_UNSUPPORTED;
  if (dtype == ACAI_BF16 && kv_rows_total > 0) {
    // TMA-streaming kernel; chunk length: the longest of 256/128/64 that still gives every SM a few CTAs
    const int hg = (heads + 3) / 4;
    int ct = 256;
    while (ct > DCH_MIN && (int64_t)((max_kv_len + ct - 1) / ct) * hg * batch < 4 * num_sms()) ct >>= 1;
    rc = decode_attn_tma(q, q_stride, new_kv, new_kv_stride, kv_rows, kv_rows_total, page_table, table_stride, seq_start,
                         kv_lens, step_ptr, len_offset, max_kv_len, batch, heads, scale, ws, (max_kv_len + ct - 1) / ct, ct, counters, out,
                         out_dtype, o_stride, st);
    if (rc != ACAI_ERR_UNSUPPORTED) return rc;         // (the TMA kernel folds its own split-KV partials: no combine launch)
  }
  const int chunks = (max_kv_len + chunk_tokens - 1) / chunk_tokens;
  if (rc == ACAI_ERR_UNSUPPORTED) {
    DecodeArgs a{q, q_stride, new_kv, new_kv_stride, kv_rows, page_table, table_stride, seq_start, kv_lens, step_ptr,
                 len_offset, heads, scale * 1.4426950408889634f, ws, chunks};
    const int threads = 32 * (heads < 16 ? heads : 16);
    if (dtype == ACAI_BF16) decode_attn_paged_kernel<__nv_bfloat16><<<dim3(chunks, batch), threads, 0, st>>>(a);
    else decode_attn_paged_kernel<float><<<dim3(chunks, batch), threads, 0, st>>>(a);
    rc = launch_status("decode_attn_paged");
    if (rc) return rc;
  }
  decode_combine_kernel<<<batch, 256, 0, st>>>(ws, chunks, chunk_tokens, heads, kv_lens, step_ptr, len_offset,
                                               new_kv ? 1 : 0, out, out_dtype, o_stride);
  return launch_status("decode_combine");
}

extern "C" int acai_decode_embed(const int64_t* seqs, int64_t seq_stride, const int32_t* step_ptr, int32_t batch,
                                 const float* table, const float* pos, void* out, int32_t out_dtype, void* out2,
                                 int32_t out2_dtype, int32_t dim, void* stream) {
  ACAI_CHECK_ARG(seqs && step_ptr && table && pos && out && batch > 0 && dim > 0 && dtype_valid(out_dtype) &&
                 (!out2 || dtype_valid(out2_dtype)), "acai_decode_embed: bad argument");
  decode_embed_kernel<<<batch, dim >= 512 ? 256 : 64, 0, (cudaStream_t)stream>>>(seqs, seq_stride, step_ptr, table, pos,
                                                                              out, out_dtype, out2, out2_dtype, dim);
  return launch_status("decode_embed");
}

extern "C" int acai_greedy_update(const void* logits, int32_t dtype, int64_t ld, int32_t rows, int32_t vocab,
                                  const int32_t* step_ptr, int64_t* seqs, int64_t seq_stride, float* log_probs,
                                  uint8_t* finished, int32_t eos_idx, void* stream) {
  ACAI_CHECK_ARG(logits && step_ptr && seqs && log_probs && finished && rows > 0 && vocab > 0 && dtype_valid(dtype),
                 "acai_greedy_update: bad argument");
  greedy_update_kernel<<<(rows + 3) / 4, 128, 0, (cudaStream_t)stream>>>(logits, dtype, ld, rows, vocab, step_ptr, seqs,
                                                                        seq_stride, log_probs, finished, eos_idx);
  return launch_status("greedy_update");
}

extern "C" int acai_topk_sample_update(const void* logits, int32_t dtype, int64_t ld, int32_t rows, int32_t vocab, int32_t top_k,
                                       float temperature, const float* uniforms, int64_t uniforms_stride,
                                       const uint64_t* seed_ptr, const int32_t* step_ptr, int64_t* seqs, int64_t seq_stride,
                                       float* log_probs, uint8_t* finished, int32_t eos_idx, void* stream) {
  ACAI_CHECK_ARG(logits && step_ptr && seqs && log_probs && finished && rows > 0 && vocab > 0 && dtype_valid(dtype),
                 "acai_topk_sample_update: bad argument");
  ACAI_CHECK_ARG(top_k >= 1 && temperature > 0.0f && (uniforms || seed_ptr), "acai_topk_sample_update: top_k >= 1, temperature > 0 and a "
                 "source of randomness (uniforms or seed_ptr) are required");
  ACAI_CHECK_ARG(vocab <= 8192, "acai_topk_sample_update: vocab %d > 8192 does not fit the shared-memory row buffer", vocab);
  const int warps = vocab <= 2048 ? 4 : 1;
  topk_sample_update_kernel<<<(rows + warps - 1) / warps, warps * 32, (size_t)warps * vocab * sizeof(float), (cudaStream_t)stream>>>(
      logits, dtype, ld, rows, vocab, top_k < vocab ? top_k : vocab, 1.0f / temperature, uniforms, uniforms_stride, seed_ptr, step_ptr,
      seqs, seq_stride, log_probs, finished, eos_idx);
  return launch_status("topk_sample_update");
}

extern "C" int acai_advance_step(int32_t* step_ptr, void* stream) {
  ACAI_CHECK_ARG(step_ptr, "acai_advance_step: null pointer");
  advance_step_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_ptr);
  return launch_status("advance_step");
}
