// Row-wise memory-bound kernels: LayerNorm(+residual), token embedding, greedy epilogue,
// packed <-> padded conversion, KV append.  All HBM-bound; one CTA (or warp) per row,
// coalesced loads along the feature dimension.
#include <cstdlib>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace acai {

// y = LN(x + residual) * gamma + beta ; biased variance, computed as mean((v-mu)^2) (two-pass, as torch).
// The row is staged in dynamic shared memory as fp32 so x/residual are read exactly once.
__global__ void __launch_bounds__(256)
layernorm_kernel(const void* __restrict__ x, int x_dt, const void* __restrict__ res, int res_dt,
                 const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                 void* __restrict__ out, int out_dt, void* __restrict__ out2, int out2_dt, int cols) {
  extern __shared__ float rowbuf[];
  __shared__ float red[32];
  const int64_t base = (int64_t)blockIdx.x * cols;
  float s = 0.0f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) {
    float v = ld_as_float(x, base + c, x_dt);
    if (res) v += ld_as_float(res, base + c, res_dt);
    rowbuf[c] = v;
    s += v;
  }
  const float mu = block_sum(s, red) / (float)cols;
  float q = 0.0f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) { float d = rowbuf[c] - mu; q += d * d; }
  const float rstd = rsqrtf(block_sum(q, red) / (float)cols + eps);
  for (int c = threadIdx.x; c < cols; c += blockDim.x) {
    float y = (rowbuf[c] - mu) * rstd * __ldg(gamma + c) + __ldg(beta + c);
    st_from_float(out, base + c, out_dt, y);
    if (out2) st_from_float(out2, base + c, out2_dt, y);
  }
}

// ---- vectorised path: one warp per row, the row lives in registers (cols % 8 == 0, cols <= 8*32*NV) ----
__device__ __forceinline__ void load8(const void* p, int64_t idx, int dt, float (&v)[8]) {
  if (dt == ACAI_BF16) {
    uint4 raw = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p) + idx);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
  } else {
    const float4* q = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p) + idx);
    float4 a = q[0], b = q[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
}
__device__ __forceinline__ void store8(void* p, int64_t idx, int dt, const float (&v)[8]) {
  if (dt == ACAI_BF16) {
    uint4 raw; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p) + idx) = raw;
  } else {
    float4* q = reinterpret_cast<float4*>(reinterpret_cast<float*>(p) + idx);
    q[0] = make_float4(v[0], v[1], v[2], v[3]); q[1] = make_float4(v[4], v[5], v[6], v[7]);
  }
}

template <int NV>
__global__ void __launch_bounds__(256)
layernorm_warp_kernel(const void* __restrict__ x, int x_dt, const void* __restrict__ res, int res_dt,
                      const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                      void* __restrict__ out, int out_dt, void* __restrict__ out2, int out2_dt, int rows, int cols) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int64_t base = (int64_t)row * cols;
  float v[NV][8];
  float s = 0.0f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * 8;
    if (c < cols) {
      load8(x, base + c, x_dt, v[i]);
      if (res) {
        float r[8]; load8(res, base + c, res_dt, r);
#pragma unroll
        for (int e = 0; e < 8; ++e) v[i][e] += r[e];
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) s += v[i][e];
    }
  }
  const float mu = warp_sum(s) / (float)cols;
  float q = 0.0f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * 8;
    if (c < cols) {
#pragma unroll
      for (int e = 0; e < 8; ++e) { const float d = v[i][e] - mu; q += d * d; }
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)cols + eps);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * 8;
    if (c < cols) {
      float g[8], b[8], y[8];
      load8(gamma, c, ACAI_F32, g); load8(beta, c, ACAI_F32, b);
#pragma unroll
      for (int e = 0; e < 8; ++e) y[e] = (v[i][e] - mu) * rstd * g[e] + b[e];
      store8(out, base + c, out_dt, y);
      if (out2) store8(out2, base + c, out2_dt, y);
    }
  }
}

// bf16 -> bf16 fast path (every LayerNorm of the bf16 model): a persistent grid of warps strides over the rows, and the
// NEXT row's 16-byte loads are issued before the current row's two warp reductions, so each warp keeps two rows of HBM
// traffic in flight and no CTA launch sits between rows (the one-row-per-warp kernel above reached ~50% of HBM peak).
template <int NV>
__global__ void __launch_bounds__(256, NV <= 2 ? 4 : (NV == 3 ? 3 : 2))
layernorm_bf16_kernel(const uint4* __restrict__ x, const uint4* __restrict__ res, const float* __restrict__ gamma,
                      const float* __restrict__ beta, float eps, uint4* __restrict__ out, int rows, int cols) {
  const int lane = threadIdx.x & 31, nwarps = gridDim.x * (blockDim.x >> 5);
  const int c8 = cols >> 3;                                   // 16-byte pieces per row
  int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  uint4 cur[NV], cur_r[NV];
  auto fetch = [&](int r, uint4 (&a)[NV], uint4 (&b)[NV]) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = i * 32 + lane;
      if (c < c8) {
        a[i] = __ldcs(x + (int64_t)r * c8 + c);
        if (res) b[i] = __ldcs(res + (int64_t)r * c8 + c);
      }
    }
  };
  if (row < rows) fetch(row, cur, cur_r);
  const float inv_n = 1.0f / (float)cols;
  while (row < rows) {
    const int next = row + nwarps;
    uint4 nxt[NV], nxt_r[NV];
    if (next < rows) fetch(next, nxt, nxt_r);
    // the row stays in its packed bf16 form (few registers = many resident warps = many bytes in flight); each of the
    // three passes unpacks it again, which costs two ALU ops per pair
    auto unpack = [&](int i, float (&f)[8]) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&cur[i]);
      const __nv_bfloat162* hr = reinterpret_cast<const __nv_bfloat162*>(&cur_r[i]);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float2 t = __bfloat1622float2(h[e]);
        if (res) { const float2 g = __bfloat1622float2(hr[e]); t.x += g.x; t.y += g.y; }
        f[2 * e] = t.x; f[2 * e + 1] = t.y;
      }
    };
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      if (i * 32 + lane < c8) {
        float f[8]; unpack(i, f);
#pragma unroll
        for (int e = 0; e < 8; ++e) s += f[e];
      }
    }
    const float mu = warp_sum(s) * inv_n;
    float q = 0.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      if (i * 32 + lane < c8) {
        float f[8]; unpack(i, f);
#pragma unroll
        for (int e = 0; e < 8; ++e) { const float d = f[e] - mu; q += d * d; }
      }
    }
    const float rstd = rsqrtf(warp_sum(q) * inv_n + eps);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = i * 32 + lane;
      if (c < c8) {
        float f[8]; unpack(i, f);
        const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * c), g1 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * c + 1);
        const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta) + 2 * c), b1 = __ldg(reinterpret_cast<const float4*>(beta) + 2 * c + 1);
        uint4 o; __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
        ho[0] = __floats2bfloat162_rn((f[0] - mu) * rstd * g0.x + b0.x, (f[1] - mu) * rstd * g0.y + b0.y);
        ho[1] = __floats2bfloat162_rn((f[2] - mu) * rstd * g0.z + b0.z, (f[3] - mu) * rstd * g0.w + b0.w);
        ho[2] = __floats2bfloat162_rn((f[4] - mu) * rstd * g1.x + b1.x, (f[5] - mu) * rstd * g1.y + b1.y);
        ho[3] = __floats2bfloat162_rn((f[6] - mu) * rstd * g1.z + b1.z, (f[7] - mu) * rstd * g1.w + b1.w);
        __stcs(out + (int64_t)row * c8 + c, o);
      }
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) { cur[i] = nxt[i]; cur_r[i] = nxt_r[i]; }
    row = next;
  }
}

// bf16 -> bf16, no residual, TMA version: every warp runs a private ring of LNT_STAGES row buffers in shared memory.  Lane 0
// keeps LNT_STAGES - 1 bulk loads (cp.async.bulk, one whole row each) in flight, the warp normalises a landed row in place
// and lane 0 sends it back with a bulk store - ~4 rows per warp x 24 warps per SM in flight without holding them in registers
// (the register version above tops out at ~3.8 TB/s: its bytes in flight are bounded by registers x resident warps).
constexpr int LNT_WARPS = 8, LNT_MAX_STAGES = 12;   // stages are chosen by the host so that two CTAs fill an SM's shared memory
template <int NV>
__global__ void __launch_bounds__(LNT_WARPS * 32)
layernorm_bf16_tma_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                          float eps, __nv_bfloat16* __restrict__ out, int rows, int cols, int LNT_STAGES, int rpo) {
  // one ring slot = one bulk copy = `rpo` consecutive rows (small rows are paired: a 1.5 KB copy per mbarrier round trip leaves
  // the copy engine's per-operation overhead exposed)
  extern __shared__ __align__(128) uint8_t lnt_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t row_bytes = (uint32_t)cols * 2, slot_bytes = row_bytes * (uint32_t)rpo;
  uint8_t* ring = lnt_smem + (size_t)warp * LNT_STAGES * slot_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(lnt_smem + (size_t)LNT_WARPS * LNT_STAGES * slot_bytes) + warp * LNT_STAGES;
  const int nwarps = gridDim.x * LNT_WARPS, first = blockIdx.x * LNT_WARPS + warp;
  const int c8 = cols >> 3;
  if (lane == 0) {
    for (int s = 0; s < LNT_STAGES; ++s) ptx::mbar_init(&bars[s], 1);
    ptx::fence_barrier_init();
  }
  __syncwarp();
  const int items = (rows + rpo - 1) / rpo;           // an item = rpo rows (the last one may be short)
  const int n_mine = first < items ? (items - first + nwarps - 1) / nwarps : 0;
  auto item_rows = [&](int k) { const int r0 = (first + k * nwarps) * rpo; return rows - r0 < rpo ? rows - r0 : rpo; };
  auto issue = [&](int k, int s) {                   // k-th item of this warp into slot s = k % LNT_STAGES (lane 0 only)
    const uint32_t bytes = row_bytes * (uint32_t)item_rows(k);
    ptx::mbar_arrive_expect_tx(&bars[s], bytes);
    ptx::bulk_load_1d(ring + (size_t)s * slot_bytes, x + (int64_t)(first + (int64_t)k * nwarps) * rpo * cols, bytes, &bars[s]);
  };
  if (lane == 0) {
    for (int k = 0; k < LNT_STAGES - 1 && k < n_mine; ++k) issue(k, k);
  }
  // gamma / beta of this lane's columns stay in registers for every row
  float4 g[NV][2], b[NV][2];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = i * 32 + lane;
    if (c < c8) {
      g[i][0] = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * c); g[i][1] = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * c + 1);
      b[i][0] = __ldg(reinterpret_cast<const float4*>(beta) + 2 * c); b[i][1] = __ldg(reinterpret_cast<const float4*>(beta) + 2 * c + 1);
    }
  }
  const float inv_n = 1.0f / (float)cols;
  int s = 0; uint32_t phase = 0;                     // slot and barrier parity of row k
  for (int k = 0; k < n_mine; ++k) {
    uint8_t* const slot0 = ring + (size_t)s * slot_bytes;
    const int nr = item_rows(k);
    ptx::mbar_wait(&bars[s], phase);
    for (int rr = 0; rr < nr; ++rr) {
    uint4* slot = reinterpret_cast<uint4*>(slot0 + (size_t)rr * row_bytes);
    float f[NV][8];
    float sum = 0.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = i * 32 + lane;
      if (c < c8) {
        const uint4 v = slot[c];
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
        for (int e = 0; e < 4; ++e) { const float2 t = __bfloat1622float2(h[e]); f[i][2 * e] = t.x; f[i][2 * e + 1] = t.y; sum += t.x + t.y; }
      }
    }
    const float mu = warp_sum(sum) * inv_n;
    float q = 0.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      if (i * 32 + lane < c8) {
#pragma unroll
        for (int e = 0; e < 8; ++e) { const float d = f[i][e] - mu; q += d * d; }
      }
    }
    const float rstd = rsqrtf(warp_sum(q) * inv_n + eps);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = i * 32 + lane;
      if (c < c8) {
        uint4 o; __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
        ho[0] = __floats2bfloat162_rn((f[i][0] - mu) * rstd * g[i][0].x + b[i][0].x, (f[i][1] - mu) * rstd * g[i][0].y + b[i][0].y);
        ho[1] = __floats2bfloat162_rn((f[i][2] - mu) * rstd * g[i][0].z + b[i][0].z, (f[i][3] - mu) * rstd * g[i][0].w + b[i][0].w);
        ho[2] = __floats2bfloat162_rn((f[i][4] - mu) * rstd * g[i][1].x + b[i][1].x, (f[i][5] - mu) * rstd * g[i][1].y + b[i][1].y);
        ho[3] = __floats2bfloat162_rn((f[i][6] - mu) * rstd * g[i][1].z + b[i][1].z, (f[i][7] - mu) * rstd * g[i][1].w + b[i][1].w);
        slot[c] = o;
      }
    }
    }
    ptx::fence_proxy_async_smem();                   // the generic-proxy writes above, before the bulk store reads them
    __syncwarp();
    if (lane == 0) {
      ptx::bulk_store_1d(out + (int64_t)(first + (int64_t)k * nwarps) * rpo * cols, slot0, row_bytes * (uint32_t)nr);
      ptx::tma_store_commit();
      // slot (k - 1) % STAGES is refilled with row k - 1 + STAGES: its store (one behind this one) must have read the slot
      ptx::tma_store_wait_read<1>();
      // (k = 0 refills nothing: it takes the ring's last slot, which has not been used yet)
      if (k + LNT_STAGES - 1 < n_mine) issue(k + LNT_STAGES - 1, s == 0 ? LNT_STAGES - 1 : s - 1);
    }
    __syncwarp();
    if (++s == LNT_STAGES) { s = 0; phase ^= 1; }
  }
  if (lane == 0) ptx::tma_store_wait<0>();
}

__global__ void embed_tokens_kernel(const int64_t* __restrict__ ids, const float* __restrict__ table,
                                    const float* __restrict__ pos, const int32_t* __restrict__ pos_ids,
                                    int pos_offset, int period, void* __restrict__ out, int out_dt, int dim) {
  const int r = blockIdx.x;
  const int64_t tok = ids[r];
  const int p = pos_ids ? pos_ids[r] : pos_offset + (period > 0 ? r % period : 0);
  for (int c = threadIdx.x; c < dim; c += blockDim.x)
    st_from_float(out, (int64_t)r * dim + c, out_dt, __ldg(table + tok * dim + c) + __ldg(pos + (int64_t)p * dim + c));
}

// One warp per row: argmax (lowest index among equal maxima, as torch.argmax) and
// log_softmax(row)[argmax] = max - (max + log(sum exp(x - max))) = -log(sum exp(x - max)).
__global__ void argmax_logprob_kernel(const void* __restrict__ logits, int dt, int64_t ld, int rows, int vocab,
                                      int64_t* __restrict__ next_token, float* __restrict__ chosen) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  float best = -INFINITY; int idx = 0x7fffffff;
  for (int c = lane; c < vocab; c += 32) {
    float v = ld_as_float(logits, (int64_t)row * ld + c, dt);
    if (v > best || (v == best && c < idx)) { best = v; idx = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ob = __shfl_xor_sync(0xffffffffu, best, o); int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (ob > best || (ob == best && oi < idx)) { best = ob; idx = oi; }
  }
  float s = 0.0f;
  for (int c = lane; c < vocab; c += 32) s += expf(ld_as_float(logits, (int64_t)row * ld + c, dt) - best);
  s = warp_sum(s);
  if (lane == 0) { next_token[row] = idx; chosen[row] = -logf(s); }
}

__global__ void unpack_padded_kernel(const void* __restrict__ packed, int in_dt, const int32_t* __restrict__ cu,
                                     int max_len, int dim, const float* __restrict__ fill,
                                     void* __restrict__ out, int out_dt) {
  const int b = blockIdx.y, t = blockIdx.x;
  const int start = cu[b], len = cu[b + 1] - start;
  const int64_t o = ((int64_t)b * max_len + t) * dim;
  if (t < len) {
    const int64_t i = (int64_t)(start + t) * dim;
    for (int c = threadIdx.x; c < dim; c += blockDim.x) st_from_float(out, o + c, out_dt, ld_as_float(packed, i + c, in_dt));
  } else {
    for (int c = threadIdx.x; c < dim; c += blockDim.x) st_from_float(out, o + c, out_dt, fill ? __ldg(fill + c) : 0.0f);
  }
}

__global__ void pack_padded_kernel(const void* __restrict__ padded, int in_dt, const int32_t* __restrict__ cu,
                                   int max_len, int dim, void* __restrict__ out, int out_dt) {
  const int b = blockIdx.y, t = blockIdx.x;
  const int start = cu[b], len = cu[b + 1] - start;
  if (t >= len) return;
  const int64_t i = ((int64_t)b * max_len + t) * dim, o = (int64_t)(start + t) * dim;
  for (int c = threadIdx.x; c < dim; c += blockDim.x) st_from_float(out, o + c, out_dt, ld_as_float(padded, i + c, in_dt));
}

__global__ void kv_append_kernel(const void* __restrict__ k_new, const void* __restrict__ v_new, int src_dt,
                                 int64_t sbs, int64_t shs, int64_t sts, void* __restrict__ kc, void* __restrict__ vc,
                                 int c_dt, int64_t bs, int64_t hs, int64_t ts, int heads, int d, int pos, int n_new) {
  const int b = blockIdx.x, j = blockIdx.y;
  for (int i = threadIdx.x; i < heads * d; i += blockDim.x) {
    const int h = i / d, e = i - h * d;
    const int64_t src = (int64_t)b * sbs + (int64_t)h * shs + (int64_t)j * sts + e;
    const int64_t dst = (int64_t)b * bs + (int64_t)h * hs + (int64_t)(pos + j) * ts + e;
    st_from_float(kc, dst, c_dt, ld_as_float(k_new, src, src_dt));
    st_from_float(vc, dst, c_dt, ld_as_float(v_new, src, src_dt));
  }
}


// ---- MAE glue (models.py:219-241, 265-267, 271-288) ------------------------------------------------------------
// nn.Unfold patches of a ragged image list as packed rows [total, C*P*P] (feature order channel, row, col).
__global__ void patchify_kernel(const acai_image_desc* __restrict__ images, int num_images, const int32_t* __restrict__ cu,
                                int total, int P, int C, void* __restrict__ out, int out_dt) {
  const int K = C * P * P;
  const int64_t n = (int64_t)total * K;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < n; g += (int64_t)gridDim.x * blockDim.x) {
    const int tok = (int)(g / K), k = (int)(g - (int64_t)tok * K);
    const int b = find_sequence(cu, num_images, tok);
    const acai_image_desc d = images[b];
    const int pid = tok - cu[b], prow = pid / d.wp, pcol = pid - prow * d.wp;
    const int ch = k / (P * P), rem = k - ch * P * P, i = rem / P, j = rem - i * P;
    const float v = __ldg(d.pixels + (int64_t)ch * d.height * d.width + (int64_t)(prow * P + i) * d.width + pcol * P + j);
    st_from_float(out, g, out_dt, v);
  }
}

// prepare_for_decoder: row i of sequence b = (r < keep_b ? latent[cu_keep[b] + r] : mask_token) + dec_pe[i / wp, i % wp]
// with r = ids_restore[cu_full[b] + i].  One CTA per output row.
__global__ void mae_restore_kernel(const void* __restrict__ latent, int lat_dt, const int32_t* __restrict__ cu_keep,
                                   const int32_t* __restrict__ cu_full, const int64_t* __restrict__ ids_restore,
                                   const int32_t* __restrict__ wps, int batch, const float* __restrict__ mask_token,
                                   const float* __restrict__ dec_pe, int pe_w, void* __restrict__ out, int out_dt, int dim) {
  const int row = blockIdx.x;
  const int b = find_sequence(cu_full, batch, row);
  const int i = row - cu_full[b];
  const int r = (int)ids_restore[row];
  const int keep = cu_keep[b + 1] - cu_keep[b];
  const int wp = wps[b];
  const float* pe = dec_pe + ((int64_t)(i / wp) * pe_w + (i % wp)) * dim;
  const int64_t src = (int64_t)(cu_keep[b] + r) * dim;
  for (int c = threadIdx.x; c < dim; c += blockDim.x) {
    const float v = r < keep ? ld_as_float(latent, src + c, lat_dt) : __ldg(mask_token + c);
    st_from_float(out, (int64_t)row * dim + c, out_dt, v + __ldg(pe + c));
  }
}

// MAELoss per patch: target normalised by its own mean / unbiased variance (eps 1e-6), mean squared error vs pred.
__global__ void mae_row_loss_kernel(const void* __restrict__ pred, int pred_dt, const float* __restrict__ target,
                                    const uint8_t* __restrict__ mask, int feats, float* __restrict__ row_loss) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  if (!mask[row]) { if (threadIdx.x == 0) row_loss[row] = 0.0f; return; }
  const int64_t base = (int64_t)row * feats;
  float s = 0.0f;
  for (int c = threadIdx.x; c < feats; c += blockDim.x) s += target[base + c];
  const float mean = block_sum(s, red) / (float)feats;
  float q = 0.0f;
  for (int c = threadIdx.x; c < feats; c += blockDim.x) { const float d = target[base + c] - mean; q += d * d; }
  const float var = block_sum(q, red) / (float)(feats - 1);
  const float rstd = 1.0f / sqrtf(var + 1.0e-6f);
  float e = 0.0f;
  for (int c = threadIdx.x; c < feats; c += blockDim.x) {
    const float d = ld_as_float(pred, base + c, pred_dt) - (target[base + c] - mean) * rstd;
    e += d * d;
  }
  e = block_sum(e, red);
  if (threadIdx.x == 0) row_loss[row] = e / (float)feats;
}

// Deterministic masked sum: out[0] = sum row_loss[mask], out[1] = count(mask).  One CTA, fixed reduction order.
__global__ void __launch_bounds__(1024)
masked_sum_kernel(const float* __restrict__ row_loss, const uint8_t* __restrict__ mask, int rows, float* __restrict__ out) {
  __shared__ float red[32];
  float s = 0.0f, n = 0.0f;
  for (int r = threadIdx.x; r < rows; r += blockDim.x) if (mask[r]) { s += row_loss[r]; n += 1.0f; }
  s = block_sum(s, red);
  n = block_sum(n, red);
  if (threadIdx.x == 0) { out[0] = s; out[1] = n; }
}

}  // namespace acai

using namespace acai;

extern "C" int acai_patchify(const acai_image_desc* images, int32_t num_images, const int32_t* cu_seqlens,
                             int32_t total_tokens, int32_t patch_size, int32_t channels, void* out, int32_t out_dtype,
                             void* stream) {
  ACAI_CHECK_ARG(images && cu_seqlens && out && num_images > 0 && total_tokens >= 0 && patch_size > 0 && channels > 0 &&
                 dtype_valid(out_dtype), "acai_patchify: bad argument");
  if (total_tokens == 0) return ACAI_OK;
  const int64_t n = (int64_t)total_tokens * channels * patch_size * patch_size;
  const int64_t blocks = (n + 255) / 256;
  patchify_kernel<<<(unsigned)(blocks < 65535 * 16 ? blocks : 65535 * 16), 256, 0, (cudaStream_t)stream>>>(
      images, num_images, cu_seqlens, total_tokens, patch_size, channels, out, out_dtype);
  return launch_status("patchify");
}

extern "C" int acai_mae_restore(const void* latent, int32_t latent_dtype, const int32_t* cu_keep, const int32_t* cu_full,
                                const int64_t* ids_restore, const int32_t* patches_per_row, int32_t batch,
                                int32_t total_rows, const float* mask_token, const float* dec_pos_embedding, int32_t pe_w,
                                void* out, int32_t out_dtype, int32_t dim, void* stream) {
  ACAI_CHECK_ARG(latent && cu_keep && cu_full && ids_restore && patches_per_row && mask_token && dec_pos_embedding && out,
                 "acai_mae_restore: null pointer");
  ACAI_CHECK_ARG(batch > 0 && total_rows >= 0 && dim > 0 && dtype_valid(latent_dtype) && dtype_valid(out_dtype),
                 "acai_mae_restore: bad argument");
  if (total_rows == 0) return ACAI_OK;
  mae_restore_kernel<<<total_rows, dim >= 512 ? 256 : 64, 0, (cudaStream_t)stream>>>(
      latent, latent_dtype, cu_keep, cu_full, ids_restore, patches_per_row, batch, mask_token, dec_pos_embedding, pe_w, out,
      out_dtype, dim);
  return launch_status("mae_restore");
}

extern "C" int acai_mae_loss_fwd(const void* pred, int32_t pred_dtype, const float* target, const uint8_t* loss_mask,
                                 int32_t rows, int32_t feats, float* row_loss, float* sums, void* stream) {
  ACAI_CHECK_ARG(pred && target && loss_mask && row_loss && sums && rows > 0 && feats > 1 && dtype_valid(pred_dtype),
                 "acai_mae_loss_fwd: bad argument");
  mae_row_loss_kernel<<<rows, feats >= 256 ? 128 : 64, 0, (cudaStream_t)stream>>>(pred, pred_dtype, target, loss_mask, feats,
                                                                                row_loss);
  int rc = launch_status("mae_row_loss");
  if (rc) return rc;
  masked_sum_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(row_loss, loss_mask, rows, sums);
  return launch_status("masked_sum");
}


extern "C" int acai_layernorm_fwd(const void* x, int32_t x_dtype, const void* residual, int32_t res_dtype,
                                  const float* gamma, const float* beta, float eps, void* out, int32_t out_dtype,
                                  void* out2, int32_t out2_dtype, int32_t rows, int32_t cols, void* stream) {
  ACAI_CHECK_ARG(x && gamma && beta && out, "acai_layernorm_fwd: null pointer");
  ACAI_CHECK_ARG(rows >= 0 && cols > 0 && cols <= 12288, "acai_layernorm_fwd: rows=%d cols=%d out of range", rows, cols);
  ACAI_CHECK_ARG(dtype_valid(x_dtype) && dtype_valid(out_dtype) && (!residual || dtype_valid(res_dtype)) &&
                 (!out2 || dtype_valid(out2_dtype)), "acai_layernorm_fwd: bad dtype");
  if (rows == 0) return ACAI_OK;
  const bool aligned = cols % 8 == 0 && cols <= 2048 && !((uintptr_t)x & 15) && !((uintptr_t)out & 15) &&
                       !((uintptr_t)gamma & 15) && !((uintptr_t)beta & 15) && (!residual || !((uintptr_t)residual & 15)) &&
                       (!out2 || !((uintptr_t)out2 & 15));
  if (aligned) {
    const int nv = (cols / 8 + 31) / 32;
    dim3 grid((rows + 7) / 8);
    cudaStream_t st = (cudaStream_t)stream;
    if (x_dtype == ACAI_BF16 && out_dtype == ACAI_BF16 && !out2 && (!residual || res_dtype == ACAI_BF16) && nv <= 4) {
      // large row counts without a residual: the per-warp TMA rings (ACAI_LN_TMA=0 keeps the register kernel for A/B runs)
      static const bool use_tma = [] { const char* e = getenv("ACAI_LN_TMA"); return !(e && atoi(e) == 0); }();
      if (use_tma && !residual && rows >= 4096) {
        const int rpo = cols * 2 >= 2048 ? 1 : (cols * 2 >= 1024 ? 2 : 4);               // rows per bulk copy (>= 2 KB each)
        int stages = (int)((size_t)100 * 1024 / ((size_t)acai::LNT_WARPS * rpo * cols * 2));   // ~100 KB per CTA, two CTAs per SM
        stages = stages < 3 ? 3 : (stages > acai::LNT_MAX_STAGES ? acai::LNT_MAX_STAGES : stages);
        const size_t smem = (size_t)acai::LNT_WARPS * stages * ((size_t)rpo * cols * 2 + 8);
        const int per_sm = (int)(227 * 1024 / (smem + 1024));
        const int want = ((rows + rpo - 1) / rpo + acai::LNT_WARPS - 1) / acai::LNT_WARPS, capt = acai::num_sms() * (per_sm < 1 ? 1 : per_sm);
        dim3 tg(want < capt ? want : capt);
#define LNT(NV)                                                                                                          \
  do {                                                                                                                   \
    static bool attr_set = false;                                                                                        \
    if (!attr_set) {                                                                                                     \
      ACAI_CUDA(cudaFuncSetAttribute(acai::layernorm_bf16_tma_kernel<NV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); \
      attr_set = true;                                                                                                   \
    }                                                                                                                    \
    acai::layernorm_bf16_tma_kernel<NV><<<tg, acai::LNT_WARPS * 32, smem, st>>>((const __nv_bfloat16*)x, gamma, beta, eps, \
                                                                               (__nv_bfloat16*)out, rows, cols, stages, rpo); \
  } while (0)
        if (nv <= 1) LNT(1); else if (nv <= 2) LNT(2); else if (nv <= 3) LNT(3); else LNT(4);
#undef LNT
        return launch_status("layernorm_bf16_tma");
      }
      const int ctas = (rows + 7) / 8, cap = acai::num_sms() * 6;
      dim3 pg(ctas < cap ? ctas : cap);
#define LNB(NV) acai::layernorm_bf16_kernel<NV><<<pg, 256, 0, st>>>((const uint4*)x, (const uint4*)residual, gamma, beta, eps, \
                                                                   (uint4*)out, rows, cols)
      if (nv <= 1) LNB(1); else if (nv <= 2) LNB(2); else if (nv <= 3) LNB(3); else LNB(4);
#undef LNB
      return launch_status("layernorm_bf16");
    }
#define LN(NV) layernorm_warp_kernel<NV><<<grid, 256, 0, st>>>(x, x_dtype, residual, res_dtype, gamma, beta, eps, out, \
                                                              out_dtype, out2, out2_dtype, rows, cols)
    if (nv <= 1) LN(1); else if (nv <= 2) LN(2); else if (nv <= 3) LN(3); else if (nv <= 4) LN(4); else LN(8);
#undef LN
    return launch_status("layernorm_warp");
  }
  const int threads = cols >= 1024 ? 256 : (cols >= 256 ? 128 : 64);
  layernorm_kernel<<<rows, threads, cols * sizeof(float), (cudaStream_t)stream>>>(
      x, x_dtype, residual, res_dtype, gamma, beta, eps, out, out_dtype, out2, out2_dtype, cols);
  return launch_status("layernorm");
}

extern "C" int acai_embed_tokens(const int64_t* ids, int32_t rows, const float* table, const float* pos,
                                 const int32_t* pos_ids, int32_t pos_offset, int32_t period, void* out,
                                 int32_t out_dtype, int32_t dim, void* stream) {
  ACAI_CHECK_ARG(ids && table && pos && out && dim > 0 && rows >= 0, "acai_embed_tokens: bad argument");
  ACAI_CHECK_ARG(dtype_valid(out_dtype), "acai_embed_tokens: bad dtype");
  if (rows == 0) return ACAI_OK;
  embed_tokens_kernel<<<rows, dim >= 512 ? 256 : 64, 0, (cudaStream_t)stream>>>(ids, table, pos, pos_ids, pos_offset,
                                                                                period, out, out_dtype, dim);
  return launch_status("embed_tokens");
}

extern "C" int acai_argmax_logprob(const void* logits, int32_t dtype, int64_t ld, int32_t rows, int32_t vocab,
                                   int64_t* next_token, float* chosen_logprob, void* stream) {
  ACAI_CHECK_ARG(logits && next_token && chosen_logprob && vocab > 0 && rows >= 0 && dtype_valid(dtype),
                 "acai_argmax_logprob: bad argument");
  if (rows == 0) return ACAI_OK;
  argmax_logprob_kernel<<<(rows + 3) / 4, 128, 0, (cudaStream_t)stream>>>(logits, dtype, ld, rows, vocab, next_token,
                                                                          chosen_logprob);
  return launch_status("argmax_logprob");
}

extern "C" int acai_unpack_padded(const void* packed, int32_t in_dtype, const int32_t* cu_seqlens, int32_t batch,
                                  int32_t max_len, int32_t dim, const float* fill, void* out, int32_t out_dtype,
                                  void* stream) {
  ACAI_CHECK_ARG(packed && cu_seqlens && out && batch > 0 && max_len > 0 && dim > 0, "acai_unpack_padded: bad argument");
  ACAI_CHECK_ARG(dtype_valid(in_dtype) && dtype_valid(out_dtype) && batch <= 65535, "acai_unpack_padded: bad dtype/batch");
  unpack_padded_kernel<<<dim3(max_len, batch), dim >= 512 ? 256 : 64, 0, (cudaStream_t)stream>>>(
      packed, in_dtype, cu_seqlens, max_len, dim, fill, out, out_dtype);
  return launch_status("unpack_padded");
}

extern "C" int acai_pack_padded(const void* padded, int32_t in_dtype, const int32_t* cu_seqlens, int32_t batch,
                                int32_t max_len, int32_t dim, void* out, int32_t out_dtype, void* stream) {
  ACAI_CHECK_ARG(padded && cu_seqlens && out && batch > 0 && max_len > 0 && dim > 0, "acai_pack_padded: bad argument");
  ACAI_CHECK_ARG(dtype_valid(in_dtype) && dtype_valid(out_dtype) && batch <= 65535, "acai_pack_padded: bad dtype/batch");
  pack_padded_kernel<<<dim3(max_len, batch), dim >= 512 ? 256 : 64, 0, (cudaStream_t)stream>>>(
      padded, in_dtype, cu_seqlens, max_len, dim, out, out_dtype);
  return launch_status("pack_padded");
}

extern "C" int acai_kv_append(const void* k_new, const void* v_new, int32_t src_dtype, int64_t src_batch_stride,
                              int64_t src_head_stride, int64_t src_tok_stride, int32_t n_new,
                              void* k_cache, void* v_cache, int32_t cache_dtype, int64_t kv_batch_stride,
                              int64_t kv_head_stride, int64_t kv_tok_stride, int32_t batch, int32_t heads,
                              int32_t head_dim, int32_t pos, void* stream) {
  ACAI_CHECK_ARG(k_new && v_new && k_cache && v_cache && batch > 0 && heads > 0 && head_dim > 0 && pos >= 0 && n_new > 0,
                 "acai_kv_append: bad argument");
  ACAI_CHECK_ARG(dtype_valid(src_dtype) && dtype_valid(cache_dtype) && n_new <= 65535, "acai_kv_append: bad dtype");
  kv_append_kernel<<<dim3(batch, n_new), 256, 0, (cudaStream_t)stream>>>(
      k_new, v_new, src_dtype, src_batch_stride, src_head_stride, src_tok_stride, k_cache, v_cache, cache_dtype,
      kv_batch_stride, kv_head_stride, kv_tok_stride, heads, head_dim, pos, n_new);
  return launch_status("kv_append");
}
