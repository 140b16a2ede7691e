// Generic fp32-accumulating SIMT projections: any shape, any of {fp32, bf16} operands.
//   * gemm_nt_simt  : 64x64x16 register-tiled out = A . W^T with pluggable A loader / epilogue
//   * gemv_small_m  : M <= 8 rows (decode steps at small batch): one warp per output column,
//                     W streamed once with 128-bit loads  -> HBM-bound
// The patch-embedding kernel (K1) is the same GEMM with an im2col A loader and a
// position-embedding epilogue, so pixels are read straight from the ragged images and the
// [tokens, C*P*P] matrix is never materialised.
#include "common.cuh"

namespace acai {

constexpr int BM = 64, BN = 64, BK = 16, PADS = 4;

// ------------------------------------------------------------------------------------------
struct PlainA {
  const void* a; int dt; int64_t lda; int M, K;
  __device__ __forceinline__ void init(int, float*) const {}
  __device__ __forceinline__ float load(int row, int k, int) const {
    return (row < M && k < K) ? ld_as_float(a, (int64_t)row * lda + k, dt) : 0.0f;
  }
};

struct LinearEpi {
  const float* bias; const void* res; int res_dt; int64_t ldr; void* out; int out_dt; int64_t ldo; int act;
  __device__ __forceinline__ void init(int, float*) const {}
  __device__ __forceinline__ void store(int m, int n, int, float acc) const {
    float v = acc + (bias ? __ldg(bias + n) : 0.0f);
    if (act == ACAI_ACT_GELU) v = gelu_erf(v);
    if (res) v += ld_as_float(res, (int64_t)m * ldr + n, res_dt);
    st_from_float(out, (int64_t)m * ldo + n, out_dt, v);
  }
};

// Per-row patch bookkeeping kept in shared memory for the 64 rows of a CTA tile.
struct PatchRow {
  const float* p0;   // first pixel of the patch (channel 0)
  int32_t width;     // image row stride
  int32_t plane;     // H*W channel stride
  int32_t r0, r1, c0, c1;  // position-embedding taps
  float lr, lc;      // bilinear weights (0 when not interpolating)
  int32_t interp, valid;
};

struct PatchCtx {
  const acai_image_desc* images; int num_images; const int32_t* cu; const int32_t* patch_index;
  int total; int P; int C; int K;
  const float* bias; const float* pe; int pe_h, pe_w, interpolate;
  void* out; int out_dt; int E;

  // torch's area_pixel_compute_source_index for bilinear, align_corners=False
  __device__ static void taps(int dst, int n_out, int n_in, int& i0, int& i1, float& lam) {
    float scale = (float)n_in / (float)n_out;
    float src = scale * ((float)dst + 0.5f) - 0.5f;
    if (src < 0.0f) src = 0.0f;
    i0 = (int)src; if (i0 > n_in - 1) i0 = n_in - 1;
    i1 = i0 + ((i0 < n_in - 1) ? 1 : 0);
    lam = src - (float)i0; lam = fminf(fmaxf(lam, 0.0f), 1.0f);
  }

  __device__ void init_rows(int row0, PatchRow* rows) const {
    for (int r = threadIdx.x; r < BM; r += blockDim.x) {
      PatchRow pr{}; int tok = row0 + r;
      if (tok < total) {
        int b = find_sequence(cu, num_images, tok);
        acai_image_desc d = images[b];
        int pid = patch_index ? patch_index[tok] : tok - cu[b];
        int prow = pid / d.wp, pcol = pid % d.wp;
        pr.p0 = d.pixels + (int64_t)prow * P * d.width + (int64_t)pcol * P;
        pr.width = d.width; pr.plane = d.height * d.width; pr.valid = 1;
        if (interpolate && (d.hp > pe_h || d.wp > pe_w)) {
          pr.interp = 1;
          taps(prow, d.hp, pe_h, pr.r0, pr.r1, pr.lr);
          taps(pcol, d.wp, pe_w, pr.c0, pr.c1, pr.lc);
        } else {
          pr.r0 = pr.r1 = prow; pr.c0 = pr.c1 = pcol;
        }
      }
      rows[r] = pr;
    }
  }
};

struct PatchA {
  PatchCtx c;
  __device__ __forceinline__ void init(int row0, float* scratch) const { c.init_rows(row0, reinterpret_cast<PatchRow*>(scratch)); }
  __device__ __forceinline__ float load(int row, int k, const PatchRow* rows, int row0) const {
    const PatchRow& pr = rows[row - row0];
    if (!pr.valid || k >= c.K) return 0.0f;
    int pp = c.P * c.P;
    int ch = k / pp, rem = k - ch * pp, i = rem / c.P, j = rem - i * c.P;
    return __ldg(pr.p0 + (int64_t)ch * pr.plane + (int64_t)i * pr.width + j);
  }
};

struct PatchEpi {
  PatchCtx c;
  __device__ __forceinline__ void store(int m, int n, const PatchRow* rows, int row0, float acc) const {
    const PatchRow& pr = rows[m - row0];
    float v = acc + (c.bias ? __ldg(c.bias + n) : 0.0f);
    const float* pe = c.pe; const int64_t E = c.E, rs = (int64_t)c.pe_w * E;
    float pos;
    if (pr.interp) {
      float p00 = __ldg(pe + pr.r0 * rs + pr.c0 * E + n), p01 = __ldg(pe + pr.r0 * rs + pr.c1 * E + n);
      float p10 = __ldg(pe + pr.r1 * rs + pr.c0 * E + n), p11 = __ldg(pe + pr.r1 * rs + pr.c1 * E + n);
      float h1 = pr.lr, h0 = 1.0f - h1, w1 = pr.lc, w0 = 1.0f - w1;
      pos = h0 * (w0 * p00 + w1 * p01) + h1 * (w0 * p10 + w1 * p11);
    } else {
      pos = __ldg(pe + pr.r0 * rs + pr.c0 * E + n);
    }
    st_from_float(c.out, (int64_t)m * E + n, c.out_dt, v + pos);
  }
};

// ------------------------------------------------------------------------------------------
template <class ALoad, class Epi, bool kPatch>
__global__ void __launch_bounds__(256)
gemm_nt_simt(ALoad aload, const void* __restrict__ w, int w_dt, int64_t ldw, Epi epi, int M, int N, int K) {
  __shared__ __align__(16) float As[BK][BM + PADS];
  __shared__ __align__(16) float Ws[BK][BN + PADS];
  __shared__ __align__(16) float scratch[kPatch ? (BM * sizeof(PatchRow) / sizeof(float)) : 1];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int row0 = blockIdx.x * BM, col0 = blockIdx.y * BN;
  if constexpr (kPatch) { aload.init(row0, scratch); __syncthreads(); }
  const PatchRow* prow = reinterpret_cast<const PatchRow*>(scratch);

  float acc[4][4] = {};
  const int lk = tid & 15, lr = tid >> 4;
  for (int k0 = 0; k0 < K; k0 += BK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int r = lr + 16 * i;
      float av;
      if constexpr (kPatch) av = aload.load(row0 + r, k0 + lk, prow, row0);
      else av = aload.load(row0 + r, k0 + lk, 0);
      As[lk][r] = av;
      int n = col0 + r, k = k0 + lk;
      Ws[lk][r] = (n < N && k < K) ? ld_as_float(w, (int64_t)n * ldw + k, w_dt) : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      float4 b4 = *reinterpret_cast<const float4*>(&Ws[kk][tx * 4]);
      float a[4] = {a4.x, a4.y, a4.z, a4.w}, b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int m = row0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = col0 + tx * 4 + j;
      if (n >= N) continue;
      if constexpr (kPatch) epi.store(m, n, prow, row0, acc[i][j]);
      else epi.store(m, n, 0, acc[i][j]);
    }
  }
}

// out[m][n] for M <= 8: one warp per n, lanes stride K by 4 (fp32) / 8 (bf16) element vectors.
template <int MAXM>
__global__ void __launch_bounds__(256)
gemv_small_m(const void* __restrict__ a, int a_dt, int64_t lda, const void* __restrict__ w, int w_dt, int64_t ldw,
             LinearEpi epi, int M, int N, int K) {
  const int lane = threadIdx.x & 31;
  const int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (n >= N) return;
  float acc[MAXM];
#pragma unroll
  for (int m = 0; m < MAXM; ++m) acc[m] = 0.0f;
  const bool vec_ok = (K % 8 == 0) && (ldw % 8 == 0) && (lda % 8 == 0) &&
                      ((reinterpret_cast<uintptr_t>(w) & 15) == 0) && ((reinterpret_cast<uintptr_t>(a) & 15) == 0);
  if (vec_ok) {
    for (int k = lane * 8; k < K; k += 256) {
      float wv[8];
      if (w_dt == ACAI_BF16) {
        uint4 raw = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(w) + (int64_t)n * ldw + k));
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) { float2 f = __bfloat1622float2(h[i]); wv[2 * i] = f.x; wv[2 * i + 1] = f.y; }
      } else {
        const float4* p = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(w) + (int64_t)n * ldw + k);
        float4 x = __ldg(p), y = __ldg(p + 1);
        wv[0] = x.x; wv[1] = x.y; wv[2] = x.z; wv[3] = x.w; wv[4] = y.x; wv[5] = y.y; wv[6] = y.z; wv[7] = y.w;
      }
#pragma unroll
      for (int m = 0; m < MAXM; ++m) {
        if (m < M) {
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[m] = fmaf(ld_as_float(a, (int64_t)m * lda + k + i, a_dt), wv[i], acc[m]);
        }
      }
    }
  } else {
    for (int k = lane; k < K; k += 32) {
      float wv = ld_as_float(w, (int64_t)n * ldw + k, w_dt);
#pragma unroll
      for (int m = 0; m < MAXM; ++m)
        if (m < M) acc[m] = fmaf(ld_as_float(a, (int64_t)m * lda + k, a_dt), wv, acc[m]);
    }
  }
#pragma unroll
  for (int m = 0; m < MAXM; ++m) {
    float v = warp_sum(acc[m]);
    if (lane == 0 && m < M) epi.store(m, n, 0, v);
  }
}

// bf16 decode-step GEMV for M <= 32 rows (greedy decode at small batch): one warp per output column n, the weight row
// streamed exactly once with 16-byte loads that bypass L1 allocation, FOUR loads in flight per lane before the first
// FMA (a 2 KB weight row is otherwise a chain of dependent DRAM round trips); the M activation rows come from L1.
// All SMs take part (N / 8 CTAs), so a [1024 x 1024] projection costs one DRAM latency instead of the 16-CTA tensor-core
// launch it replaces (the tcgen05 tile is 128 rows tall: at M = 1 it moves 2 MB of weights through 16 SMs).
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float dot8_bf16(const uint4& a, const uint4& b, float acc) {
  const __nv_bfloat162* x = reinterpret_cast<const __nv_bfloat162*>(&a);
  const __nv_bfloat162* y = reinterpret_cast<const __nv_bfloat162*>(&b);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 fx = __bfloat1622float2(x[i]), fy = __bfloat1622float2(y[i]);
    acc = fmaf(fx.x, fy.x, acc);
    acc = fmaf(fx.y, fy.y, acc);
  }
  return acc;
}
template <int MAXM>
__global__ void __launch_bounds__(256)
gemv_bf16_small_m(const __nv_bfloat16* __restrict__ a, int64_t lda, const __nv_bfloat16* __restrict__ w, int64_t ldw,
                  LinearEpi epi, int M, int N, int K) {
  const int lane = threadIdx.x & 31;
  const int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (n >= N) return;
  float acc[MAXM];
#pragma unroll
  for (int m = 0; m < MAXM; ++m) acc[m] = 0.0f;
  const uint4* wp = reinterpret_cast<const uint4*>(w + (int64_t)n * ldw);
  const int nv = K >> 3;
  for (int v0 = lane; v0 < nv; v0 += 128) {
    uint4 wr[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) wr[u] = v0 + 32 * u < nv ? ldg_stream(wp + v0 + 32 * u) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int idx = v0 + 32 * u;
      if (idx < nv) {
#pragma unroll
        for (int m = 0; m < MAXM; ++m)
          if (m < M) acc[m] = dot8_bf16(__ldg(reinterpret_cast<const uint4*>(a + (int64_t)m * lda) + idx), wr[u], acc[m]);
      }
    }
  }
#pragma unroll
  for (int m = 0; m < MAXM; ++m) {
    if (m < M) {                                   // M is uniform: no divergence around the shuffles
      const float v = warp_sum(acc[m]);
      if (lane == 0) epi.store(m, n, 0, v);
    }
  }
}

// bf16 x bf16, M <= 32, 16-byte addressable rows: the decode-step projections at small batch.  ACAI_ERR_UNSUPPORTED otherwise.
int linear_gemv_bf16(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* res, int res_dt,
                     int64_t ldr, void* out, int out_dt, int64_t ldo, int M, int N, int K, int act, cudaStream_t st) {
  if (M > 32 || K % 8 || lda % 8 || ldw % 8 || (reinterpret_cast<uintptr_t>(a) & 15) || (reinterpret_cast<uintptr_t>(w) & 15))
    return ACAI_ERR_UNSUPPORTED;
  LinearEpi epi{bias, res, res_dt, ldr, out, out_dt, ldo, act};
  const __nv_bfloat16* ab = reinterpret_cast<const __nv_bfloat16*>(a);
  const __nv_bfloat16* wb = reinterpret_cast<const __nv_bfloat16*>(w);
  const dim3 grid((N + 7) / 8);
  if (M <= 2) gemv_bf16_small_m<2><<<grid, 256, 0, st>>>(ab, lda, wb, ldw, epi, M, N, K);
  else if (M <= 8) gemv_bf16_small_m<8><<<grid, 256, 0, st>>>(ab, lda, wb, ldw, epi, M, N, K);
  else if (M <= 16) gemv_bf16_small_m<16><<<grid, 256, 0, st>>>(ab, lda, wb, ldw, epi, M, N, K);
  else gemv_bf16_small_m<32><<<grid, 256, 0, st>>>(ab, lda, wb, ldw, epi, M, N, K);
  return launch_status("gemv_bf16_small_m");
}

int linear_simt(const void* a, int a_dt, int64_t lda, const void* w, int w_dt, int64_t ldw, const float* bias,
                const void* res, int res_dt, int64_t ldr, void* out, int out_dt, int64_t ldo,
                int M, int N, int K, int act, cudaStream_t st) {
  LinearEpi epi{bias, res, res_dt, ldr, out, out_dt, ldo, act};
  if (M <= 8) {
    dim3 grid((N + 7) / 8);
    gemv_small_m<8><<<grid, 256, 0, st>>>(a, a_dt, lda, w, w_dt, ldw, epi, M, N, K);
    return launch_status("gemv_small_m");
  }
  PlainA al{a, a_dt, lda, M, K};
  dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN);
  gemm_nt_simt<PlainA, LinearEpi, false><<<grid, 256, 0, st>>>(al, w, w_dt, ldw, epi, M, N, K);
  return launch_status("gemm_nt_simt");
}

}  // namespace acai

namespace acai {

int linear_tcgen05_ex(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* res,
                      int res_dt, int64_t ldr, const int32_t* res_rows, void* out, int out_dt, int64_t ldo, int M, int N,
                      int K, int act, cudaStream_t st);

// im2col for the tensor-core patch embedding: token t -> its C*P*P pixels as one bf16 row (feature order channel,
// row-in-patch, col-in-patch = nn.Unfold's), plus the row of the position grid it will add.  One thread moves 8
// consecutive pixels of one patch row (2 x 16-byte loads when the image rows allow it, one 16-byte store); a warp
// covers whole 32/64-byte pixel runs, so every sector fetched is used.
__global__ void __launch_bounds__(256)
patch_im2col_kernel(const acai_image_desc* __restrict__ images, int num_images, const int32_t* __restrict__ cu,
                    const int32_t* __restrict__ patch_index, int total, int P, int C, int K, int pe_w,
                    __nv_bfloat16* __restrict__ patches, int32_t* __restrict__ pe_rows) {
  const int pieces = K / 8;                                    // 16-byte output pieces per token
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (int64_t)total * pieces) return;
  const int tok = (int)(gid / pieces), piece = (int)(gid - (int64_t)tok * pieces);
  const int b = find_sequence(cu, num_images, tok);
  const acai_image_desc d = images[b];
  const int pid = patch_index ? patch_index[tok] : tok - cu[b];
  const int prow = pid / d.wp, pcol = pid - prow * d.wp;
  if (piece == 0) pe_rows[tok] = prow * pe_w + pcol;
  const int k = piece * 8, pp = P * P;
  const int ch = k / pp, rem = k - ch * pp, i = rem / P, j = rem - i * P;      // P % 8 == 0: the 8 pixels share a row
  const float* src = d.pixels + (int64_t)ch * d.height * d.width + (int64_t)(prow * P + i) * d.width + pcol * P + j;
  float v[8];
  if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(src)), c = __ldg(reinterpret_cast<const float4*>(src) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
  } else {
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = __ldg(src + e);
  }
  uint4 pk; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
  for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
  *reinterpret_cast<uint4*>(patches + (int64_t)tok * K + k) = pk;
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = __float2bfloat16_rn(in[i]);
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace acai

using namespace acai;

extern "C" size_t acai_patch_embed_workspace_bytes(int32_t total_tokens, int32_t patch_size, int32_t channels,
                                                   int32_t embed_dim) {
  const size_t K = (size_t)channels * patch_size * patch_size;
  return align256((size_t)total_tokens * K * 2) + align256((size_t)total_tokens * 4) + align256((size_t)embed_dim * K * 2);
}

extern "C" int acai_patch_embed_fwd(const acai_image_desc* images, int32_t num_images, const int32_t* cu_seqlens,
                                    const int32_t* patch_index, int32_t total_tokens, int32_t patch_size,
                                    int32_t channels, const float* weight, const float* bias,
                                    const float* pos_embedding, int32_t pe_h, int32_t pe_w, int32_t interpolate,
                                    void* out, int32_t out_dtype, int32_t embed_dim, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  ACAI_CHECK_ARG(images && cu_seqlens && weight && pos_embedding && out, "acai_patch_embed_fwd: null pointer");
  ACAI_CHECK_ARG(num_images > 0 && total_tokens >= 0 && patch_size > 0 && channels > 0 && embed_dim > 0,
                 "acai_patch_embed_fwd: bad sizes");
  ACAI_CHECK_ARG(dtype_valid(out_dtype), "acai_patch_embed_fwd: bad out dtype %d", out_dtype);
  if (total_tokens == 0) return ACAI_OK;
  const int K = channels * patch_size * patch_size;
  cudaStream_t st = (cudaStream_t)stream;
  if (workspace && out_dtype == ACAI_BF16 && !interpolate && K % 64 == 0 && patch_size % 8 == 0 && embed_dim % 8 == 0 &&
      !((uintptr_t)workspace & 255) && !((uintptr_t)out & 15) && !((uintptr_t)pos_embedding & 15) &&
      workspace_bytes >= acai_patch_embed_workspace_bytes(total_tokens, patch_size, channels, embed_dim)) {
    char* ws = reinterpret_cast<char*>(workspace);
    __nv_bfloat16* patches = reinterpret_cast<__nv_bfloat16*>(ws);
    int32_t* pe_rows = reinterpret_cast<int32_t*>(ws + align256((size_t)total_tokens * K * 2));
    __nv_bfloat16* wb = reinterpret_cast<__nv_bfloat16*>(ws + align256((size_t)total_tokens * K * 2) + align256((size_t)total_tokens * 4));
    const int64_t n_pieces = (int64_t)total_tokens * (K / 8);
    patch_im2col_kernel<<<(unsigned)((n_pieces + 255) / 256), 256, 0, st>>>(images, num_images, cu_seqlens, patch_index,
                                                                          total_tokens, patch_size, channels, K, pe_w,
                                                                          patches, pe_rows);
    int rc = launch_status("patch_im2col");
    if (rc) return rc;
    const int64_t nw = (int64_t)embed_dim * K;
    cast_f32_bf16_kernel<<<(unsigned)((nw + 255) / 256), 256, 0, st>>>(weight, wb, nw);
    rc = launch_status("cast_f32_bf16");
    if (rc) return rc;
    rc = linear_tcgen05_ex(patches, K, wb, K, bias, pos_embedding, ACAI_F32, embed_dim, pe_rows, out, ACAI_BF16, embed_dim,
                           total_tokens, embed_dim, K, ACAI_ACT_NONE, st);
    if (rc != ACAI_ERR_UNSUPPORTED) return rc;
  }
  PatchCtx c{images, num_images, cu_seqlens, patch_index, total_tokens, patch_size, channels, K,
             bias, pos_embedding, pe_h, pe_w, interpolate, out, out_dtype, embed_dim};
  PatchA al{c}; PatchEpi ep{c};
  dim3 grid((total_tokens + BM - 1) / BM, (embed_dim + BN - 1) / BN);
  gemm_nt_simt<PatchA, PatchEpi, true><<<grid, 256, 0, st>>>(
      al, weight, ACAI_F32, (int64_t)K, ep, total_tokens, embed_dim, K);
  return launch_status("patch_embed");
}
