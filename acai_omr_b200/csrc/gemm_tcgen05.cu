// K2: bf16 x bf16 -> fp32-accumulate GEMM on the 5th-gen tensor cores (tcgen05.mma, TMEM accumulators),
// operands staged by TMA into 128B-swizzled shared memory.
//
//   out[M,N] = act(A[M,K] . W[N,K]^T + bias) + residual          (both operands K-major, as nn.Linear)
//
// Persistent, warp-specialised CTA (one per SM):
//   warp 0      : TMA producer   (A tile 128x64, W tile BNx64 per stage; mbarrier full/empty ring)
//   warp 1      : MMA issuer     (one lane; 4 x tcgen05.mma 128xBNx16 per stage; tcgen05.commit frees the stage)
//   warps 2..9  : epilogue       (tcgen05.ld 32 lanes x 32 columns -> bias/GELU/residual -> global)
// The accumulator is double-buffered in TMEM (2 x BN columns), so the epilogue of tile i overlaps the MMAs
// of tile i+1.  Tiles are walked n-fastest so the CTAs running concurrently share their A rows in L2.
#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace acai {

namespace g5 {
constexpr int BM = 128;           // UMMA_M (cta_group::1)
constexpr int BK = 64;            // 64 bf16 = 128 bytes = one swizzle row
constexpr int UMMA_K = 16;
constexpr int EPI_WARPS = 8;
constexpr int THREADS = 64 + EPI_WARPS * 32;   // 320

struct Epi {
  const float* bias; const void* res; int res_dt; int64_t ldr; void* out; int out_dt; int64_t ldo; int act;
};

template <int BN> constexpr int stages() { return BN >= 256 ? 4 : (BN >= 128 ? 6 : 8); }
template <int BN> constexpr size_t smem_bytes() {
  return (size_t)stages<BN>() * (BM * BK * 2 + BN * BK * 2) + EPI_WARPS * 4096 /*epilogue staging*/ +
         1024 /*align slack*/ + 256 /*barriers*/;
}
}  // namespace g5

using namespace g5;

__device__ __forceinline__ void epi_store_chunk(const Epi& ep, const uint32_t (&acc)[32], int row, int col0, int N,
                                                bool vec_ok) {
  // this thread owns out[row, col0 .. col0+31]
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc[j]);
  if (ep.bias) {
#pragma unroll
    for (int j = 0; j < 32; ++j) if (col0 + j < N) v[j] += __ldg(ep.bias + col0 + j);
  }
  if (ep.act == ACAI_ACT_GELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
  }
  const bool full = vec_ok && (col0 + 32 <= N);
  if (ep.res) {
    if (full && ep.res_dt == ACAI_BF16) {
      const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(ep.res) + (int64_t)row * ep.ldr + col0);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 raw = __ldg(rp + q);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) { float2 f = __bfloat1622float2(h[i]); v[q * 8 + 2 * i] += f.x; v[q * 8 + 2 * i + 1] += f.y; }
      }
    } else if (full && ep.res_dt == ACAI_F32) {
      const float4* rp = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(ep.res) + (int64_t)row * ep.ldr + col0);
#pragma unroll
      for (int q = 0; q < 8; ++q) { float4 f = __ldg(rp + q); v[4 * q] += f.x; v[4 * q + 1] += f.y; v[4 * q + 2] += f.z; v[4 * q + 3] += f.w; }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) if (col0 + j < N) v[j] += ld_as_float(ep.res, (int64_t)row * ep.ldr + col0 + j, ep.res_dt);
    }
  }
  if (full && ep.out_dt == ACAI_BF16) {
    uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(ep.out) + (int64_t)row * ep.ldo + col0);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint4 pk;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
      for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[q * 8 + 2 * i], v[q * 8 + 2 * i + 1]);
      op[q] = pk;
    }
  } else if (full && ep.out_dt == ACAI_F32) {
    float4* op = reinterpret_cast<float4*>(reinterpret_cast<float*>(ep.out) + (int64_t)row * ep.ldo + col0);
#pragma unroll
    for (int q = 0; q < 8; ++q) op[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) if (col0 + j < N) st_from_float(ep.out, (int64_t)row * ep.ldo + col0 + j, ep.out_dt, v[j]);
  }
}

// exact-erf GELU: x * Phi(x) with erfc from Abramowitz-Stegun 7.1.26 (|abs err of erf| < 1.5e-7).
// 2 MUFU (rcp, ex2) + ~11 FMA-pipe ops per element; the negative branch keeps full relative accuracy.
__device__ __forceinline__ float gelu_fast(float x) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.0f)));
  float poly = fmaf(t, 1.061405429f, -1.453152027f);
  poly = fmaf(t, poly, 1.421413741f);
  poly = fmaf(t, poly, -0.284496736f);
  poly = fmaf(t, poly, 0.254829592f);
  const float zs = z * 1.2011224087864498f;                 // sqrt(log2 e) * z  ->  exp(-z^2) = 2^(-zs^2)
  const float h = 0.5f * poly * t * fast_exp2(-zs * zs);    // 0.5 * erfc(z)
  return x * (x >= 0.0f ? 1.0f - h : h);
}

// Residual prefetch for one 32x32 block: the 16-byte pieces this lane will add in epi_block_coalesced.
__device__ __forceinline__ void epi_prefetch_res(const Epi& ep, int lane, int row0, int col0, int M, int N, uint4 (&r)[8]) {
  const int c8 = (lane & 3) * 8;
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    const int grow = row0 + it * 8 + (lane >> 2), gcol = col0 + c8;
    if (grow < M && gcol < N) {
      if (ep.res_dt == ACAI_BF16) {
        r[it] = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(ep.res) + (int64_t)grow * ep.ldr + gcol));
      } else {
        const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const float*>(ep.res) + (int64_t)grow * ep.ldr + gcol);
        r[it] = __ldg(rp); r[4 + it] = __ldg(rp + 1);
      }
    }
  }
}

// Coalesced epilogue for one 32-row x 32-column accumulator block owned by a warp (thread = row):
// bias/GELU in registers -> fp32 staging tile in shared memory (16B chunks XOR-swizzled by row) -> re-read with
// 4 lanes per row so that every global access is a 64..128-byte row segment; the residual is added at that point,
// with the same coalesced pattern.  Requires N % 8 == 0 and 16-byte aligned rows (checked by the caller).
__device__ __forceinline__ void epi_block_coalesced(const Epi& ep, const uint32_t (&acc)[32], float* stage, int lane,
                                                    int row0, int col0, int M, int N, const uint4 (&rpre)[8]) {
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc[j]);
  if (ep.bias) {
    if (col0 + 32 <= N) {
      const float4* bp = reinterpret_cast<const float4*>(ep.bias + col0);
#pragma unroll
      for (int q = 0; q < 8; ++q) { const float4 b4 = __ldg(bp + q); v[4 * q] += b4.x; v[4 * q + 1] += b4.y; v[4 * q + 2] += b4.z; v[4 * q + 3] += b4.w; }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) if (col0 + j < N) v[j] += __ldg(ep.bias + col0 + j);
    }
  }
  if (ep.act == ACAI_ACT_GELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = gelu_fast(v[j]);
  }
  __syncwarp();   // the previous block's readers are done with the staging tile
  {
    float4* srow = reinterpret_cast<float4*>(stage + lane * 32);
#pragma unroll
    for (int q = 0; q < 8; ++q) srow[q ^ (lane & 7)] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
  }
  __syncwarp();
  const int sub = lane & 3, c8 = sub * 8;
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    const int r = it * 8 + (lane >> 2);
    const int grow = row0 + r, gcol = col0 + c8;
    const float4* srow = reinterpret_cast<const float4*>(stage + r * 32);
    const float4 x0 = srow[(2 * sub) ^ (r & 7)], x1 = srow[(2 * sub + 1) ^ (r & 7)];
    if (grow < M && gcol < N) {
      float y[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
      if (ep.res) {
        if (ep.res_dt == ACAI_BF16) {
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&rpre[it]);
#pragma unroll
          for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h[i]); y[2 * i] += f.x; y[2 * i + 1] += f.y; }
        } else {
          const float4 a = *reinterpret_cast<const float4*>(&rpre[it]), b = *reinterpret_cast<const float4*>(&rpre[4 + it]);
          y[0] += a.x; y[1] += a.y; y[2] += a.z; y[3] += a.w; y[4] += b.x; y[5] += b.y; y[6] += b.z; y[7] += b.w;
        }
      }
      if (ep.out_dt == ACAI_BF16) {
        uint4 pk; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
        for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(y[2 * i], y[2 * i + 1]);
        *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(ep.out) + (int64_t)grow * ep.ldo + gcol) = pk;
      } else {
        float4* op = reinterpret_cast<float4*>(reinterpret_cast<float*>(ep.out) + (int64_t)grow * ep.ldo + gcol);
        op[0] = make_float4(y[0], y[1], y[2], y[3]); op[1] = make_float4(y[4], y[5], y[6], y[7]);
      }
    }
  }
}

template <int BN>
__global__ void __launch_bounds__(THREADS, 1)
gemm_bf16_tcgen05(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, Epi ep,
                  int M, int N, int K) {
  constexpr int STAGES = stages<BN>();
  constexpr uint32_t A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;
  constexpr uint32_t TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;   // power of two for BN in {32,64,128,256}
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * (A_BYTES + B_BYTES));
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  float* staging = reinterpret_cast<float*>(smem + STAGES * (A_BYTES + B_BYTES) + 256);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_m = (M + BM - 1) / BM, num_n = (N + BN - 1) / BN, num_tiles = num_m * num_n, num_k = K / BK;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tma_a);
    ptx::prefetch_tensormap(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&full_bar[s], 1); ptx::mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(&tfull_bar[s], 1); ptx::mbar_init(&tempty_bar[s], EPI_WARPS); }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile / num_n) * BM, n0 = (tile % num_n) * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          ptx::mbar_arrive_expect_tx(&full_bar[stage], A_BYTES + B_BYTES);
          ptx::tma_load_2d(smem_a + stage * A_BYTES, &tma_a, &full_bar[stage], kb * BK, m0);
          ptx::tma_load_2d(smem_b + stage * B_BYTES, &tma_b, &full_bar[stage], kb * BK, n0);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t IDESC = ptx::idesc_bf16_f32(BM, BN);
      int stage = 0; uint32_t phase = 0; int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
        ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint64_t adesc = ptx::smem_desc_sw128(ptx::smem_u32(smem_a + stage * A_BYTES), 1024);
          const uint64_t bdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem_b + stage * B_BYTES), 1024);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)   // +32 bytes along K inside the 128B swizzle row = +2 in the address field
            ptx::umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, IDESC, (kb | k) != 0);
          ptx::umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        ptx::umma_commit(&tfull_bar[as]);
      }
    }
  } else {
    const int ew = warp - 2;             // 0..7
    const int lane_grp = warp & 3;       // TMEM lanes this warp may touch: 32*(warp%4) ..
    const int half = ew >> 2;            // which half of the BN columns
    constexpr int COLS_PER_WARP = BN / 2;
    const bool vec_ok = (N % 8 == 0) && (ep.ldo % 8 == 0) && ((reinterpret_cast<uintptr_t>(ep.out) & 15) == 0) &&
                        (!ep.res || ((ep.ldr % 8 == 0) && ((reinterpret_cast<uintptr_t>(ep.res) & 15) == 0))) &&
                        (!ep.bias || ((reinterpret_cast<uintptr_t>(ep.bias) & 15) == 0));
    float* stage = staging + ew * 1024;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      const int m0 = (tile / num_n) * BM, n0 = (tile % num_n) * BN;
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after();
      const int row = m0 + lane_grp * 32 + lane;
      const uint32_t taddr = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + as * BN + half * COLS_PER_WARP;
      const int brow0 = m0 + lane_grp * 32, bcol0 = n0 + half * COLS_PER_WARP;
      const bool active = vec_ok && brow0 < M;
      uint4 rcur[8], rnext[8];
      if (active && ep.res && bcol0 < N) epi_prefetch_res(ep, lane, brow0, bcol0, M, N, rcur);
#pragma unroll 1
      for (int c = 0; c < COLS_PER_WARP; c += 32) {
        const int col0 = bcol0 + c;
        if (active && ep.res && c + 32 < COLS_PER_WARP && col0 + 32 < N)
          epi_prefetch_res(ep, lane, brow0, col0 + 32, M, N, rnext);   // one block ahead: hides the L2/HBM latency
        uint32_t acc[32];
        ptx::tmem_ld_32x32(taddr + c, acc);
        ptx::tmem_ld_wait();
        if (vec_ok) {
          if (col0 < N && brow0 < M) epi_block_coalesced(ep, acc, stage, lane, brow0, col0, M, N, rcur);
        } else if (row < M && col0 < N) {
          epi_store_chunk(ep, acc, row, col0, N, false);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) rcur[i] = rnext[i];
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// 2-D bf16 row-major tensor [rows, cols] (ld elements between rows), box = [box_rows, 64 cols], 128B swizzle.
int make_tma_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                     uint32_t box_cols) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled is unavailable (driver too old?)"); return ACAI_ERR_CUDA; }
  struct Key { const void* p; uint64_t r, c, l; uint32_t br, bc; };
  struct Hash { size_t operator()(const Key& k) const {
    size_t h = reinterpret_cast<size_t>(k.p);
    h ^= k.r * 0x9E3779B97F4A7C15ull + (h << 6); h ^= k.c * 0xC2B2AE3D27D4EB4Full + (h >> 3);
    h ^= k.l * 0x165667B19E3779F9ull; h ^= ((size_t)k.br << 32 | k.bc) * 0x27D4EB2F165667C5ull; return h; } };
  struct Eq { bool operator()(const Key& a, const Key& b) const {
    return a.p == b.p && a.r == b.r && a.c == b.c && a.l == b.l && a.br == b.br && a.bc == b.bc; } };
  static std::unordered_map<Key, CUtensorMap, Hash, Eq> cache;
  static std::mutex mu;
  Key key{ptr, rows, cols, ld, box_rows, box_cols};
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *map = it->second; return ACAI_OK; }
  }
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return ACAI_ERR_CUDA; }
  std::lock_guard<std::mutex> g(mu);
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *map);
  return ACAI_OK;
}

template <int BN>
static int launch_gemm(const void* a, int64_t lda, const void* w, int64_t ldw, const Epi& ep, int M, int N, int K,
                       cudaStream_t st) {
  CUtensorMap ta, tb;
  int rc = make_tma_2d_bf16(&ta, a, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BM, BK);
  if (rc) return rc;
  rc = make_tma_2d_bf16(&tb, w, (uint64_t)N, (uint64_t)K, (uint64_t)ldw, BN, BK);
  if (rc) return rc;
  static bool attr_set = false;
  constexpr size_t smem = smem_bytes<BN>();
  if (!attr_set) {
    ACAI_CUDA(cudaFuncSetAttribute(gemm_bf16_tcgen05<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int grid = tiles < num_sms() ? tiles : num_sms();
  gemm_bf16_tcgen05<BN><<<grid, THREADS, smem, st>>>(ta, tb, ep, M, N, K);
  return launch_status("gemm_bf16_tcgen05");
}

int linear_tcgen05(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* res,
                   int res_dt, int64_t ldr, void* out, int out_dt, int64_t ldo, int M, int N, int K, int act,
                   cudaStream_t st) {
  // qualification: TMA needs 16-byte aligned bases and row pitches; the K loop walks whole 64-element blocks
  if (K % BK != 0 || lda % 8 != 0 || ldw % 8 != 0 || (reinterpret_cast<uintptr_t>(a) & 15) ||
      (reinterpret_cast<uintptr_t>(w) & 15) || M < 1 || N < 8)
    return ACAI_ERR_UNSUPPORTED;
  Epi ep{bias, res, res_dt, ldr, out, out_dt, ldo, act};
  // widest N tile that still yields about one tile per SM; skinny decode-step problems get narrow tiles so that
  // more SMs stream the weight matrix
  const int m_tiles = (M + BM - 1) / BM, sms = num_sms();
  auto tiles = [&](int bn) { return m_tiles * ((N + bn - 1) / bn); };
  if (N > 128 && tiles(256) >= sms) return launch_gemm<256>(a, lda, w, ldw, ep, M, N, K, st);
  if (N > 64 && tiles(128) >= sms) return launch_gemm<128>(a, lda, w, ldw, ep, M, N, K, st);
  return launch_gemm<64>(a, lda, w, ldw, ep, M, N, K, st);
}

}  // namespace acai
