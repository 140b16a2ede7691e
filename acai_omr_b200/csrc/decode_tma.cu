// K4 (bf16 path): decode-step attention that streams the token-major KV cache through TMA - HBM-bound by design.
//
// The register kernel in decode_paged.cu keeps only ~16 KB of loads in flight per SM (128 registers x 512 threads,
// one CTA per SM, loads issued and consumed by the same instructions) and measured 2.7 TB/s.  Here every warp owns
// one head of one (sequence, chunk) and runs its own 3-stage TMA ring: a stage = 32 tokens of that head's K slice and
// V slice (two [32 x 64] bf16 boxes, 128B-swizzled), so ~24 KB per warp / ~190 KB per SM are in flight independent of
// the math, and the math itself needs no shuffles: lane = token.  A lane reads its token's 128-byte K row from the
// swizzled tile (conflict-free: 16-byte piece c of row r sits at (c ^ (r & 7))), dots it with the head's query held
// in registers, keeps a private online-softmax state (lazy rescale: the running maximum only moves when a score beats
// it by more than 2^8) and accumulates p * V[token] into 64 private fp32 accumulators.  The 32 lane states are merged
// once per chunk through shared memory.  Split-KV combine: a sequence with one chunk writes its output directly; otherwise
// every warp publishes its chunk partial to the workspace, bumps an arrival counter of its (sequence, head), and the last
// one in folds the partials in chunk order (same arithmetic as decode_combine_kernel) - no second launch.
// The token appended this step (self-attention) is taken from the projection output, attended to, and written to the
// cache by the CTA that owns its position (KVCache.update, kv_caching.py:57-109).
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace acai {

int make_tma_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                     uint32_t box_cols);

namespace dt {
constexpr int HD = 64, TOK = 32, NST = 3, WARPS = 4, PAGE = 64;
constexpr uint32_t BOX = TOK * HD * 2;                      // 4 KB
constexpr uint32_t WARP_SMEM = NST * 2 * BOX;               // 24 KB: [stage][K|V]
constexpr size_t SMEM = WARPS * WARP_SMEM + 256;
constexpr float RESCALE = 8.0f;
}  // namespace dt

struct DecodeTmaArgs {
  const __nv_bfloat16* q; int64_t q_stride;
  const __nv_bfloat16* new_kv; int64_t new_kv_stride;
  __nv_bfloat16* pool;
  const int32_t* page_table; int table_stride;
  const int32_t* seq_start; const int32_t* kv_lens; const int32_t* step_ptr; int len_offset;
  int heads; float scale_log2; float* ws; int max_chunks; int chunk_tokens;
  // split-KV combine inside this kernel: the LAST warp of a (sequence, head) to publish its partial folds all of them
  int32_t* counters;            // [batch, heads], zero before the first launch; every launch leaves them zero
  void* out; int out_dt; int64_t o_stride;
};

__device__ __forceinline__ void bf16x8_to_f32(const uint4& w, float (&f)[8]) {
  f[0] = __uint_as_float(w.x << 16); f[1] = __uint_as_float(w.x & 0xffff0000u);
  f[2] = __uint_as_float(w.y << 16); f[3] = __uint_as_float(w.y & 0xffff0000u);
  f[4] = __uint_as_float(w.z << 16); f[5] = __uint_as_float(w.z & 0xffff0000u);
  f[6] = __uint_as_float(w.w << 16); f[7] = __uint_as_float(w.w & 0xffff0000u);
}

// One token for this lane: score, lazy online-softmax update, accumulate.  `kp` / `vp` return 16-byte piece c (8 dims).
template <class KP, class VP>
__device__ __forceinline__ void token_update(bool valid, const float (&q)[64], float& m, float& l, float (&acc)[64],
                                             KP kp, VP vp) {
  float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    float k[8];
    bf16x8_to_f32(kp(c), k);
    s0 = fmaf(q[c * 8 + 0], k[0], s0); s1 = fmaf(q[c * 8 + 1], k[1], s1);
    s2 = fmaf(q[c * 8 + 2], k[2], s2); s3 = fmaf(q[c * 8 + 3], k[3], s3);
    s0 = fmaf(q[c * 8 + 4], k[4], s0); s1 = fmaf(q[c * 8 + 5], k[5], s1);
    s2 = fmaf(q[c * 8 + 6], k[6], s2); s3 = fmaf(q[c * 8 + 7], k[7], s3);
  }
  const float s = valid ? (s0 + s1) + (s2 + s3) : -INFINITY;     // log2 units (q is pre-scaled)
  const bool grow = s > m + dt::RESCALE || (m == -INFINITY && s > -INFINITY);
  if (__any_sync(0xffffffffu, grow)) {
    const float m_new = grow ? s : m;
    const float corr = (m == -INFINITY) ? 0.0f : fast_exp2(m - m_new);   // 1 for lanes that keep their maximum
    l *= corr;
#pragma unroll
    for (int e = 0; e < 64; ++e) acc[e] *= corr;
    m = m_new;
  }
  if (valid) {
    const float p = fast_exp2(s - m);
    l += p;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      float v[8];
      bf16x8_to_f32(vp(c), v);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[c * 8 + e] = fmaf(p, v[e], acc[c * 8 + e]);
    }
  }
}

__global__ void __launch_bounds__(dt::WARPS * 32, 2)
decode_attn_tma_kernel(const __grid_constant__ CUtensorMap tma_kv, DecodeTmaArgs a) {
  using namespace dt;
  pdl_wait();
  pdl_launch_dependents();
  extern __shared__ __align__(1024) uint8_t smem[];
  // (the warp index through a shuffle: provably warp-uniform, so that everything the TMA issue below derives from it - head
  // column, ring slot, barrier - lives in uniform registers; see ptx::elect_one_sync)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int chunk = blockIdx.x, h = blockIdx.y * WARPS + warp, b = blockIdx.z;
  const int H = a.heads;
  uint8_t* wbuf = smem + warp * WARP_SMEM;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + WARPS * WARP_SMEM) + warp * NST;

  const int len = a.kv_lens ? a.kv_lens[b] : (*a.step_ptr + a.len_offset);     // tokens already in the cache
  const int total = len + (a.new_kv ? 1 : 0);
  const int t_begin = chunk * a.chunk_tokens;
  if (h >= H) return;                                                           // warp-uniform; no CTA-wide sync below
  if (t_begin >= total) {
    if (chunk == 0) {                                                           // no keys at all: the output row is zero
      st_from_float(a.out, (int64_t)b * a.o_stride + h * HD + lane, a.out_dt, 0.0f);
      st_from_float(a.out, (int64_t)b * a.o_stride + h * HD + 32 + lane, a.out_dt, 0.0f);
    }
    return;
  }
  const int t_end = min(total, t_begin + a.chunk_tokens);
  const int c_end = min(len, t_end);                                            // cached tokens of this chunk
  const int n_st = c_end > t_begin ? (c_end - t_begin + TOK - 1) / TOK : 0;
  const int32_t* tbl = a.page_table ? a.page_table + (int64_t)b * a.table_stride : nullptr;
  const int row0 = tbl ? 0 : a.seq_start[b];

  if (lane == 0) {
    for (int s = 0; s < NST; ++s) ptx::mbar_init(&full[s], 1);
    ptx::fence_barrier_init();
  }
  __syncwarp();
  auto issue = [&](int i) {                                                     // stage i -> ring slot i % NST; the whole warp calls
    const int t0 = t_begin + i * TOK;
    const int row = __shfl_sync(0xffffffffu, tbl ? tbl[t0 / PAGE] * PAGE + (t0 % PAGE) : row0 + t0, 0);
    if (ptx::elect_one_sync()) {
      uint8_t* dst = wbuf + (i % NST) * 2 * BOX;
      ptx::mbar_arrive_expect_tx(&full[i % NST], 2 * BOX);
      ptx::tma_load_2d(dst, &tma_kv, &full[i % NST], h * HD, row);
      ptx::tma_load_2d(dst + BOX, &tma_kv, &full[i % NST], (H + h) * HD, row);
    }
    __syncwarp();
  };
  {
    const int n_u = __shfl_sync(0xffffffffu, n_st, 0);
    for (int i = 0; i < NST && i < n_u; ++i) issue(i);
  }

  float q[64];
  {
    const uint4* qp = reinterpret_cast<const uint4*>(a.q + (int64_t)b * a.q_stride + h * HD);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      float f[8];
      bf16x8_to_f32(__ldg(qp + c), f);
#pragma unroll
      for (int e = 0; e < 8; ++e) q[c * 8 + e] = f[e] * a.scale_log2;
    }
  }
  float m = -INFINITY, l = 0.0f, acc[64];
#pragma unroll
  for (int e = 0; e < 64; ++e) acc[e] = 0.0f;
  const uint32_t sw = (uint32_t)(lane & 7);

  for (int i = 0; i < n_st; ++i) {
    ptx::mbar_wait(&full[i % NST], (i / NST) & 1);
    const uint8_t* kb = wbuf + (i % NST) * 2 * BOX + lane * 128;
    const uint8_t* vb = kb + BOX;
    const bool valid = t_begin + i * TOK + lane < c_end;
    token_update(valid, q, m, l, acc,
                 [&](int c) { return *reinterpret_cast<const uint4*>(kb + ((((uint32_t)c) ^ sw) << 4)); },
                 [&](int c) { return *reinterpret_cast<const uint4*>(vb + ((((uint32_t)c) ^ sw) << 4)); });
    __syncwarp();                                     // every lane is done with the slot before it is refilled
    if (i + NST < n_st) issue(i + NST);
  }
  if (a.new_kv && len >= t_begin && len < t_end) {    // the token appended this step: lane 0 attends to it ...
    const __nv_bfloat16* nk = a.new_kv + (int64_t)b * a.new_kv_stride + h * HD;
    const __nv_bfloat16* nv = nk + H * HD;
    token_update(lane == 0, q, m, l, acc,
                 [&](int c) { return __ldg(reinterpret_cast<const uint4*>(nk) + c); },
                 [&](int c) { return __ldg(reinterpret_cast<const uint4*>(nv) + c); });
    if (lane < 16) {                                   // ... and lanes 0-15 copy its K and V slices into the cache
      const int64_t row = tbl ? (int64_t)tbl[len / PAGE] * PAGE + (len % PAGE) : (int64_t)row0 + len;
      const int64_t col = (lane < 8 ? h * HD : (H + h) * HD) + (lane & 7) * 8;
      const __nv_bfloat16* src = (lane < 8 ? nk : nv) + (lane & 7) * 8;
      *reinterpret_cast<uint4*>(a.pool + row * (2 * H * HD) + col) = __ldg(reinterpret_cast<const uint4*>(src));
    }
  }
  // ---- merge the 32 lane states: rescale to the warp maximum, transpose through smem (pitch 65 floats: conflict-free
  // both ways), lane d sums dims d and d + 32
  const float M = warp_max(m);
  const float f = (m == -INFINITY) ? 0.0f : fast_exp2(m - M);
  const float lsum = warp_sum(l * f);
  __syncwarp();
  float* tr = reinterpret_cast<float*>(wbuf);          // 32 x 65 floats = 8320 B of the drained ring
#pragma unroll
  for (int e = 0; e < 64; ++e) tr[lane * 65 + e] = acc[e] * f;
  __syncwarp();
  float o0 = 0.0f, o1 = 0.0f;
#pragma unroll 8
  for (int r = 0; r < 32; ++r) { o0 += tr[r * 65 + lane]; o1 += tr[r * 65 + 32 + lane]; }
  const int n_chunks = min(a.max_chunks, (total + a.chunk_tokens - 1) / a.chunk_tokens);   // chunks that reach this point
  const int64_t orow = (int64_t)b * a.o_stride + h * HD;
  if (n_chunks == 1) {                                  // nothing to fold (same arithmetic as the fold of one partial)
    st_from_float(a.out, orow + lane, a.out_dt, lsum > 0.0f ? o0 / lsum : 0.0f);
    st_from_float(a.out, orow + 32 + lane, a.out_dt, lsum > 0.0f ? o1 / lsum : 0.0f);
    return;
  }
  float* w = a.ws + (((int64_t)b * a.max_chunks + chunk) * H + h) * (HD + 2);
  __stcg(w + lane, o0); __stcg(w + 32 + lane, o1);
  if (lane == 0) { __stcg(w + HD, M); __stcg(w + HD + 1, lsum); }
  __threadfence();                                      // this warp's partial is visible device-wide ...
  __syncwarp();
  int old = 0;
  if (lane == 0) old = atomicAdd(a.counters + b * H + h, 1);   // ... before it is counted
  old = __shfl_sync(0xffffffffu, old, 0);
  if (old != n_chunks - 1) return;
  // last one in: out[b, h*64 + e] = sum_c acc_c[e] 2^(m_c - M) / sum_c l_c 2^(m_c - M), chunks in ascending order
  __threadfence();
  if (lane == 0) a.counters[b * H + h] = 0;            // ready for the next launch
  const float* w0 = a.ws + ((int64_t)b * a.max_chunks * H + h) * (HD + 2);
  const int64_t cs = (int64_t)H * (HD + 2);
  float Mx = -INFINITY;
#pragma unroll 4
  for (int c = 0; c < n_chunks; ++c) Mx = fmaxf(Mx, __ldcg(w0 + c * cs + HD));
  float n0 = 0.0f, n1 = 0.0f, den = 0.0f;
#pragma unroll 4
  for (int c = 0; c < n_chunks; ++c) {
    const float mc = __ldcg(w0 + c * cs + HD), lc = __ldcg(w0 + c * cs + HD + 1);
    const float a0 = __ldcg(w0 + c * cs + lane), a1 = __ldcg(w0 + c * cs + 32 + lane);
    if (mc > -INFINITY) {
      const float fc = fast_exp2(mc - Mx);
      n0 = fmaf(a0, fc, n0); n1 = fmaf(a1, fc, n1); den = fmaf(lc, fc, den);
    }
  }
  st_from_float(a.out, orow + lane, a.out_dt, den > 0.0f ? n0 / den : 0.0f);
  st_from_float(a.out, orow + 32 + lane, a.out_dt, den > 0.0f ? n1 / den : 0.0f);
}

// returns ACAI_ERR_UNSUPPORTED when the problem does not qualify (the caller then runs the register kernel)
int decode_attn_tma(const void* q, int64_t q_stride, const void* new_kv, int64_t new_kv_stride, void* kv_rows,
                    int64_t kv_rows_total, const int32_t* page_table, int table_stride, const int32_t* seq_start,
                    const int32_t* kv_lens, const int32_t* step_ptr, int len_offset, int max_kv_len, int batch, int heads,
                    float scale, float* ws, int max_chunks, int chunk_tokens, int32_t* counters, void* out, int out_dt,
                    int64_t o_stride, cudaStream_t st) {
  using namespace dt;
  if (kv_rows_total <= 0 || chunk_tokens % TOK != 0 || batch > 65535) return ACAI_ERR_UNSUPPORTED;
  CUtensorMap tm;
  int rc = make_tma_2d_bf16(&tm, kv_rows, (uint64_t)kv_rows_total, (uint64_t)2 * heads * HD, (uint64_t)2 * heads * HD, TOK, HD);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    ACAI_CUDA(cudaFuncSetAttribute(decode_attn_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    attr_set = true;
  }
  DecodeTmaArgs a{reinterpret_cast<const __nv_bfloat16*>(q), q_stride, reinterpret_cast<const __nv_bfloat16*>(new_kv),
                  new_kv_stride, reinterpret_cast<__nv_bfloat16*>(kv_rows), page_table, table_stride, seq_start, kv_lens,
                  step_ptr, len_offset, heads, scale * 1.4426950408889634f, ws, max_chunks, chunk_tokens, counters, out, out_dt, o_stride};
  const int chunks = (max_kv_len + chunk_tokens - 1) / chunk_tokens;
  dim3 grid(chunks, (heads + WARPS - 1) / WARPS, batch);
  ACAI_CUDA(launch_pdl(decode_attn_tma_kernel, grid, dim3(WARPS * 32), SMEM, st, tm, a));
  return launch_status("decode_attn_tma");
}

}  // namespace acai
