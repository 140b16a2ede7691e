// K2: bf16 x bf16 -> fp32-accumulate GEMM on the 5th-gen tensor cores (tcgen05.mma, TMEM accumulators),
// operands staged by TMA into 128B-swizzled shared memory.
//
//   out[M,N] = act(A[M,K] . W[N,K]^T + bias) + residual          (both operands K-major, as nn.Linear)
//
// Persistent, warp-specialised CTA (one per SM):
//   warp 0      : TMA producer   (A tile 128x64, W tile BNx64 per stage; mbarrier full/empty ring)
//   warp 1      : MMA issuer     (one lane; 4 x tcgen05.mma 128xBNx16 per stage; tcgen05.commit frees the stage)
//   warps 2..9  : epilogue       (tcgen05.ld -> bias/GELU/residual in registers -> 128B-swizzled smem -> TMA store)
// The accumulator is double-buffered in TMEM (2 x BN columns), so the epilogue of tile i overlaps the MMAs
// of tile i+1.  Tiles are walked n-fastest so the CTAs running concurrently share their A rows in L2.
// Epilogue latency hiding: the tile's bias slice is fetched before the wait on the accumulator and broadcast from
// shared memory; the residual (thread = row, whole 128-byte lines) is prefetched one chunk ahead; output rows
// leave through TMA, which also clips the ragged M / N edges.  TMEM is released as soon as the last tcgen05.ld of
// the tile has landed, before the stores.  Problems whose output is not TMA-addressable use the generic epilogue.
#include <mutex>
#include <unordered_map>

#include <cstdlib>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace acai {

namespace g5 {
constexpr int BM = 128;           // UMMA_M (cta_group::1)
constexpr int BK = 64;            // 64 bf16 = 128 bytes = one swizzle row
constexpr int UMMA_K = 16;
constexpr int EPI_WARPS = 8;
#ifndef GEMM_EPI16_DEFAULT
#define GEMM_EPI16_DEFAULT 1
#endif
constexpr int THREADS = 64 + EPI_WARPS * 32;   // 320

// Training-only epilogue modes (internal; the public acai_act values are 0 / 1):
//   ACT_GELU_TRAIN: z = acc + bias is stored (bf16) through a second tensor map AND h = dropout(gelu_erf(z)) through the
//                   first - the feed-forward GELU (+ its nn.Dropout) without a separate elementwise pass;
//   ACT_GELU_BWD:   out = acc * gelu'(z) * dropout_mask, z read like a bf16 residual - the backward of the same.
constexpr int ACT_GELU_TRAIN = 2, ACT_GELU_BWD = 3;

struct Epi {
  const float* bias; const void* res; int res_dt; int64_t ldr; void* out; int out_dt; int64_t ldo; int act;
  const int32_t* res_rows;   // optional gather: residual row of output row m is res_rows[m] (position embeddings)
  void* aux; int64_t ld_aux; // ACT_GELU_TRAIN: the pre-activation output
  uint32_t drop_thr; float drop_rp; uint64_t drop_seed;   // dropout on the activation (thr = 0: none)
  // split-K (small-M decode projections, generic epilogue only): tile = (k slice, m, n); slice ks accumulates k blocks
  // [ks * kper, (ks + 1) * kper) and stores its raw fp32 accumulator to ws[ks][M][N]; splitk_reduce_kernel finishes the job
  int splitk; int kper; float* ws;
};

template <int BN> constexpr int stages() { return BN >= 256 ? 4 : (BN >= 128 ? 6 : 8); }
// EW = epilogue warps: 8, or 16 for the training epilogues (ACT_GELU_TRAIN / ACT_GELU_BWD), whose ~30 dependent instructions per
// element run at ~0.45 IPC on two warps per scheduler and outlast the K = 768 main loop: four warps per scheduler, 64-column slabs,
// three pipeline stages (the sixteen 4 KB staging tiles take the fourth stage's room)
template <int BN, int EW> constexpr int stages_ew() { return EW == 16 ? 3 : stages<BN>(); }
template <int BN, int EW = EPI_WARPS> constexpr size_t smem_bytes() {
  return (size_t)stages_ew<BN, EW>() * (BM * BK * 2 + BN * BK * 2) + EW * 4096 /*epilogue staging*/ +
         2 * BN * 4 /*bias slices*/ + 256 /*barriers*/;
}
}  // namespace g5

using namespace g5;

__device__ __forceinline__ void epi_store_chunk(const Epi& ep, const uint32_t (&acc)[32], int row, int64_t rrow, int col0,
                                                int N, bool vec_ok) {
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
    for (int j = 0; j < 32; ++j) v[j] = gelu_fast(v[j]);
  }
  const bool full = vec_ok && (col0 + 32 <= N);
  if (ep.res) {
    if (full && ep.res_dt == ACAI_BF16) {
      const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(ep.res) + rrow * ep.ldr + col0);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 raw = __ldg(rp + q);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) { float2 f = __bfloat1622float2(h[i]); v[q * 8 + 2 * i] += f.x; v[q * 8 + 2 * i + 1] += f.y; }
      }
    } else if (full && ep.res_dt == ACAI_F32) {
      const float4* rp = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(ep.res) + rrow * ep.ldr + col0);
#pragma unroll
      for (int q = 0; q < 8; ++q) { float4 f = __ldg(rp + q); v[4 * q] += f.x; v[4 * q + 1] += f.y; v[4 * q + 2] += f.z; v[4 * q + 3] += f.w; }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) if (col0 + j < N) v[j] += ld_as_float(ep.res, rrow * ep.ldr + col0 + j, ep.res_dt);
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

// ---- TMA-store epilogue ------------------------------------------------------------------------------------
// RES: 0 = no residual, 1 = bf16 residual, 2 = fp32 residual.  A "half" is 32 accumulator columns of this thread's row.
template <int RES> struct ResHalf { uint4 v[RES == 2 ? 8 : 4]; };

template <int RES>
__device__ __forceinline__ void res_prefetch(const Epi& ep, int64_t rrow, int col0, int N, ResHalf<RES>& rb) {
  if constexpr (RES != 0) {
    constexpr int NV = RES == 2 ? 8 : 4, EPV = RES == 2 ? 4 : 8;
    const uint4* base = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(ep.res) +
                                                       (rrow * ep.ldr + col0) * (RES == 2 ? 4 : 2));
#pragma unroll
    for (int i = 0; i < NV; ++i)
      rb.v[i] = (rrow >= 0 && col0 + i * EPV < N) ? __ldg(base + i) : make_uint4(0u, 0u, 0u, 0u);
  }
}

// bias (broadcast from smem) / GELU / residual on 32 columns, then the row's 16-byte pieces go to the 128B-swizzled
// staging tile: piece p of row r lives at r*128 + ((p ^ (r & 7)) << 4), conflict-free for thread = row.
template <bool OUT_BF16, int RES>
__device__ __forceinline__ void epi_half(const Epi& ep, int64_t elem0, const uint32_t (&acc)[32], const float* sb, const ResHalf<RES>& rb,
                                         uint8_t* srow, uint32_t sw, int piece0) {
  const int act = ep.act;
  float v[32];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const float4 b4 = reinterpret_cast<const float4*>(sb)[q];
    v[4 * q] = __uint_as_float(acc[4 * q]) + b4.x; v[4 * q + 1] = __uint_as_float(acc[4 * q + 1]) + b4.y;
    v[4 * q + 2] = __uint_as_float(acc[4 * q + 2]) + b4.z; v[4 * q + 3] = __uint_as_float(acc[4 * q + 3]) + b4.w;
  }
  if (act == ACAI_ACT_GELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = gelu_fast(v[j]);
  }
  if (ep.drop_thr) {   // out = residual + dropout(a w^T + bias): the post-norm blocks' nn.Dropout, mask counter = row * N + column
    const uint32_t pair0 = (uint32_t)(elem0 >> 1), hc = dropout_bits_hi(ep.drop_seed, (uint32_t)(elem0 >> 33));
#pragma unroll
    for (int k = 0; k < 16; ++k) {      // elem0 is a multiple of 32: the 16 pairs share their high word
      const uint32_t bits = dropout_bits_lo(hc, pair0 + k);
      v[2 * k] *= (bits & 0xffffu) >= ep.drop_thr ? ep.drop_rp : 0.0f;
      v[2 * k + 1] *= (bits >> 16) >= ep.drop_thr ? ep.drop_rp : 0.0f;
    }
  }
  if constexpr (RES == 1) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&rb.v[q]);
#pragma unroll
      for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h[i]); v[8 * q + 2 * i] += f.x; v[8 * q + 2 * i + 1] += f.y; }
    }
  } else if constexpr (RES == 2) {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float4 f = *reinterpret_cast<const float4*>(&rb.v[q]);
      v[4 * q] += f.x; v[4 * q + 1] += f.y; v[4 * q + 2] += f.z; v[4 * q + 3] += f.w;
    }
  }
  if constexpr (OUT_BF16) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint4 pk; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
      for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[8 * q + 2 * i], v[8 * q + 2 * i + 1]);
      *reinterpret_cast<uint4*>(srow + ((((uint32_t)(piece0 + q)) ^ sw) << 4)) = pk;
    }
  } else {
#pragma unroll
    for (int q = 0; q < 8; ++q)
      *reinterpret_cast<float4*>(srow + ((((uint32_t)(piece0 + q)) ^ sw) << 4)) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
  }
}

// Training epilogues on 32 columns of this thread's row, 8 columns (one 16-byte staging piece) at a time:
//   MODE 1 (ACT_GELU_TRAIN): pass 0 stages z = acc + bias, pass 1 stages h = dropout(gelu(bf16(z)));
//   MODE 2 (ACT_GELU_BWD):   stages acc * gelu'(z) * mask with z in the (bf16) residual registers.
// `elem0`: linear index (row * N + column, even) of the first of the 32 elements - the dropout mask's counter.
template <int MODE>
__device__ __forceinline__ void epi_half_train(const Epi& ep, int pass, int64_t elem0, const uint32_t (&acc)[32], const float* sb,
                                               const ResHalf<MODE == 2 ? 1 : 0>& rb, uint8_t* srow, uint32_t sw, int piece0) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float v[8];
    const float4 b0 = reinterpret_cast<const float4*>(sb)[2 * q], b1 = reinterpret_cast<const float4*>(sb)[2 * q + 1];
    v[0] = __uint_as_float(acc[8 * q]) + b0.x; v[1] = __uint_as_float(acc[8 * q + 1]) + b0.y;
    v[2] = __uint_as_float(acc[8 * q + 2]) + b0.z; v[3] = __uint_as_float(acc[8 * q + 3]) + b0.w;
    v[4] = __uint_as_float(acc[8 * q + 4]) + b1.x; v[5] = __uint_as_float(acc[8 * q + 5]) + b1.y;
    v[6] = __uint_as_float(acc[8 * q + 6]) + b1.z; v[7] = __uint_as_float(acc[8 * q + 7]) + b1.w;
    bool mask = false;
    if constexpr (MODE == 1) {
      if (pass == 1) {
        mask = ep.drop_thr != 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float z = __bfloat162float(__float2bfloat16_rn(v[j]));   // what the backward will read
          v[j] = gelu_fast(z);
        }
      }
    } else {
      mask = ep.drop_thr != 0;
      const __nv_bfloat162* zh = reinterpret_cast<const __nv_bfloat162*>(&rb.v[q]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 z = __bfloat1622float2(zh[i]);
        v[2 * i] *= gelu_grad_fast(z.x);
        v[2 * i + 1] *= gelu_grad_fast(z.y);
      }
    }
    if (mask) {
      // elem0 is a multiple of 8 (N % 8 == 0, 32-column halves): the 4 pairs of this piece share their high word
      const uint64_t pair0 = (uint64_t)(elem0 >> 1) + 4 * q;
      const uint32_t hc = dropout_bits_hi(ep.drop_seed, (uint32_t)(pair0 >> 32));
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t bits = dropout_bits_lo(hc, (uint32_t)pair0 + k);
        v[2 * k] *= (bits & 0xffffu) >= ep.drop_thr ? ep.drop_rp : 0.0f;
        v[2 * k + 1] *= (bits >> 16) >= ep.drop_thr ? ep.drop_rp : 0.0f;
      }
    }
    uint4 pk; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(srow + ((((uint32_t)(piece0 + q)) ^ sw) << 4)) = pk;
  }
}

template <int EW> __device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(EW * 32) : "memory"); }

// OUT: 0 = generic epilogue (any shape/alignment), 1 = bf16 output via TMA, 2 = fp32 output via TMA.
// MODE: 0 = inference / plain epilogue, 1 = ACT_GELU_TRAIN (OUT 1, RES 0), 2 = ACT_GELU_BWD (OUT 1, RES 1).
// BKN: the B operand is given as W[K, N] (contraction index = ROW: "MN-major" in UMMA terms) instead of W[N, K] - the dX = dZ W
// GEMMs of the backward read the forward's weight copy as it lies in HBM, no transposed copy per step.  TMA brings
// [64 contraction rows x 64 output columns] boxes, the descriptor walks them with LBO = 8 KB (next 64 output columns) and a
// K step is 16 rows of 128 bytes - exactly how gemm_wgrad_tcgen05 reads both of its operands.
template <int BN, int OUT, int RES, int MODE = 0, bool BKN = false, int EW = EPI_WARPS>
__global__ void __launch_bounds__(64 + EW * 32, 1)
gemm_bf16_tcgen05(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                  const __grid_constant__ CUtensorMap tma_out, const __grid_constant__ CUtensorMap tma_aux, Epi ep, int M, int N,
                  int K) {
  constexpr int STAGES = stages_ew<BN, EW>();
  constexpr uint32_t A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;
  constexpr uint32_t TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;   // power of two for BN in {32,64,128,256}
  extern __shared__ __align__(1024) uint8_t smem[];            // SWIZZLE_128B tiles need 1024-byte alignment
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_BYTES;
  uint8_t* staging = smem + STAGES * (A_BYTES + B_BYTES);                 // 8 warps x 4 KB
  float* sbias = reinterpret_cast<float*>(staging + EW * 4096);           // [2][BN]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sbias + 2 * BN);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // (provably warp-uniform: ptx::elect_one_sync)
  const int num_m = (M + BM - 1) / BM, num_n = (N + BN - 1) / BN, num_mn = num_m * num_n, num_k = K / BK;
  const int splitk = (OUT == 0 && ep.splitk > 1) ? ep.splitk : 1, kper = splitk > 1 ? ep.kper : num_k;
  const int num_tiles = num_mn * splitk;

  if (warp == 0 && lane == 0) {
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    ptx::prefetch_tensormap(&tma_a);
    ptx::prefetch_tensormap(&tma_b);
    if constexpr (OUT != 0) ptx::prefetch_tensormap(&tma_out);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&full_bar[s], 1); ptx::mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(&tfull_bar[s], 1); ptx::mbar_init(&tempty_bar[s], EW); }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                    // everything above (barriers, TMEM, tensor-map prefetch) may overlap the previous kernel's tail
  pdl_launch_dependents();

  // Producer and issuer warps walk their loops CONVERGED (every lane polls the barriers) and one elected lane issues
  // (ptx::elect_one_sync): the TMA / tcgen05 operands then come straight from uniform registers.  Under `if (lane == 0)`
  // every UTMALDG / UTCHMMA was a waterfall loop of ~100 clk - next to the 128 clk a 128 x 256 x 16 MMA takes.
  if (warp == 0) {
    int stage = 0; uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int mn = tile % num_mn, ks = tile / num_mn;
      const int m0 = (mn / num_n) * BM, n0 = (mn % num_n) * BN;
      const int kb0 = ks * kper, kb1 = kb0 + kper < num_k ? kb0 + kper : num_k;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
        if (ptx::elect_one_sync()) {
          ptx::mbar_arrive_expect_tx(&full_bar[stage], A_BYTES + B_BYTES);
          ptx::tma_load_2d(smem_a + stage * A_BYTES, &tma_a, &full_bar[stage], kb * BK, m0);
          if constexpr (BKN) {
#pragma unroll
            for (int i = 0; i < BN / 64; ++i)
              ptx::tma_load_2d(smem_b + stage * B_BYTES + i * 8192, &tma_b, &full_bar[stage], n0 + 64 * i, kb * BK);
          } else {
            ptx::tma_load_2d(smem_b + stage * B_BYTES, &tma_b, &full_bar[stage], kb * BK, n0);
          }
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t IDESC = ptx::idesc_bf16_f32(BM, BN, false, BKN);
    const uint32_t tb0 = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint64_t adesc0 = ptx::smem_desc_sw128(ptx::smem_u32(smem_a), 1024);
    const uint64_t bdesc0 = BKN ? ptx::smem_desc_sw128(ptx::smem_u32(smem_b), 1024, 8192) : ptx::smem_desc_sw128(ptx::smem_u32(smem_b), 1024);
    int stage = 0; uint32_t phase = 0; int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
      ptx::tc_fence_after();
      const uint32_t d_tmem = tb0 + as * BN;
      const int kb0 = (tile / num_mn) * kper, kb1 = kb0 + kper < num_k ? kb0 + kper : num_k;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        if (ptx::elect_one_sync()) {
          const uint64_t adesc = adesc0 + (uint64_t)(stage * (A_BYTES >> 4)), bdesc = bdesc0 + (uint64_t)(stage * (B_BYTES >> 4));
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)   // +32 bytes along K inside the 128B swizzle row = +2 in the address field
            ptx::umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + (BKN ? (uint64_t)k * (16 * 128 >> 4) : (uint64_t)(2 * k)), IDESC,
                             (kb != kb0 || k != 0) ? 1u : 0u);
          ptx::umma_commit(&empty_bar[stage]);
          if (kb == kb1 - 1) ptx::umma_commit(&tfull_bar[as]);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    const int ew = warp - 2;             // 0..EW-1
    const int lane_grp = warp & 3;       // TMEM lanes this warp may touch: 32*(warp%4) ..
    const int half = ew >> 2;            // which half (EW 16: quarter) of the BN columns
    constexpr int CPW = BN / (EW / 4);   // columns per warp
    int it = 0;
    if constexpr (OUT != 0) {
      constexpr bool OUT_BF16 = OUT == 1;
      constexpr int CW = OUT_BF16 ? 64 : 32;          // columns per chunk = 128 bytes of output per row
      static_assert(CPW % CW == 0, "a warp's column slab must hold whole 128-byte chunks");
      constexpr int HALVES = CW / 32;
      uint8_t* const stage = staging + ew * 4096;
      uint8_t* const srow = stage + lane * 128;
      const uint32_t sw = (uint32_t)(lane & 7);
      const int etid = (int)threadIdx.x - 64;         // 0..255
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
        const int m0 = (tile / num_n) * BM, n0 = (tile % num_n) * BN;   // (no split-K with the TMA-store epilogue)
        const int row0 = m0 + lane_grp * 32, grow = row0 + lane, col0 = n0 + half * CPW;
        // everything that comes from global memory is requested before the wait on the accumulator
        float bv = 0.0f;
        if (etid < BN && ep.bias && n0 + etid < N) bv = __ldg(ep.bias + n0 + etid);
        int64_t rrow = -1;
        if (RES != 0 && grow < M) rrow = ep.res_rows ? (int64_t)__ldg(ep.res_rows + grow) : (int64_t)grow;
        const int ncols = N - col0 < CPW ? N - col0 : CPW;                 // warp-uniform
        const int nch = (row0 < M && ncols > 0) ? (ncols + CW - 1) / CW : 0;
        ResHalf<RES> rb[HALVES];
        if (nch > 0) {
#pragma unroll
          for (int hh = 0; hh < HALVES; ++hh) res_prefetch<RES>(ep, rrow, col0 + hh * 32, N, rb[hh]);
        }
        if (etid < BN) sbias[as * BN + etid] = bv;
        epi_bar_sync<EW>();
        ptx::mbar_wait(&tfull_bar[as], aphase);
        ptx::tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + as * BN + half * CPW;
        const float* sb = sbias + as * BN + half * CPW;
        if (nch == 0) {
          ptx::tc_fence_before();
          if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
        }
#pragma unroll 1
        for (int c = 0; c < nch; ++c) {
          uint32_t acc[HALVES][32];
#pragma unroll
          for (int hh = 0; hh < HALVES; ++hh) ptx::tmem_ld_32x32(taddr + c * CW + hh * 32, acc[hh]);
          ptx::tmem_ld_wait();
          if (c == nch - 1) {                          // accumulator fully read: hand TMEM back before the stores
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
          }
          constexpr int PASSES = MODE == 1 ? 2 : 1;
#pragma unroll
          for (int ps = 0; ps < PASSES; ++ps) {
            if (lane == 0) ptx::tma_store_wait_read<0>();   // the staging tile's previous store has been read out
            __syncwarp();
#pragma unroll
            for (int hh = 0; hh < HALVES; ++hh) {
              if constexpr (MODE == 0)
                epi_half<OUT_BF16, RES>(ep, (int64_t)grow * N + col0 + c * CW + hh * 32, acc[hh], sb + c * CW + hh * 32, rb[hh], srow, sw,
                                        hh * (OUT_BF16 ? 4 : 0));
              else
                epi_half_train<MODE>(ep, ps, (int64_t)grow * N + col0 + c * CW + hh * 32, acc[hh], sb + c * CW + hh * 32, rb[hh],
                                     srow, sw, hh * 4);
              if (ps == PASSES - 1 && c + 1 < nch) res_prefetch<RES>(ep, rrow, col0 + (c + 1) * CW + hh * 32, N, rb[hh]);
            }
            ptx::fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              ptx::tma_store_2d(PASSES == 2 && ps == 0 ? &tma_aux : &tma_out, stage, col0 + c * CW, row0);
              ptx::tma_store_commit();
            }
          }
        }
      }
      if (lane == 0) ptx::tma_store_wait<0>();
    } else {
      const bool vec_ok = (N % 8 == 0) && (ep.ldo % 8 == 0) && ((reinterpret_cast<uintptr_t>(ep.out) & 15) == 0) &&
                          (!ep.res || ((ep.ldr % 8 == 0) && ((reinterpret_cast<uintptr_t>(ep.res) & 15) == 0)));
      const bool vec_ws = (N % 8 == 0) && ((reinterpret_cast<uintptr_t>(ep.ws) & 15) == 0);
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
        const int mn = tile % num_mn, ks = tile / num_mn;
        const int m0 = (mn / num_n) * BM, n0 = (mn % num_n) * BN;
        Epi et = ep;
        if (splitk > 1) {                              // raw fp32 partial sums; bias / residual / activation come in the reduce
          et.bias = nullptr; et.res = nullptr; et.act = ACAI_ACT_NONE;
          et.out = ep.ws + (int64_t)ks * M * N; et.out_dt = ACAI_F32; et.ldo = N;
        }
        ptx::mbar_wait(&tfull_bar[as], aphase);
        ptx::tc_fence_after();
        const int row = m0 + lane_grp * 32 + lane;
        const uint32_t taddr = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + as * BN + half * CPW;
        const int bcol0 = n0 + half * CPW;
        const int64_t rrow = (ep.res && ep.res_rows && row < M) ? (int64_t)__ldg(ep.res_rows + row) : (int64_t)row;
#pragma unroll 1
        for (int c = 0; c < CPW; c += 32) {
          const int col0 = bcol0 + c;
          uint32_t acc[32];
          ptx::tmem_ld_32x32(taddr + c, acc);
          ptx::tmem_ld_wait();
          if (row < M && col0 < N) epi_store_chunk(et, acc, row, rrow, col0, N, splitk > 1 ? vec_ws : vec_ok);
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---- split-K reduce (+ bias, activation, residual, LayerNorm) ---------------------------------------------------
// One warp per output row: v = sum_ks ws[ks][row][:] + bias, activation, + residual; with gamma != nullptr the row is then
// normalised (two-pass variance over the fp32 values, as acai_layernorm_fwd) - the decode step's out_proj / linear2 +
// residual + LayerNorm in one pass over data that is still in L2.  VPL = float4 pieces per lane (N = 128 * VPL) when the
// row has to stay in registers for the LayerNorm; VPL = 0 streams any N % 4 == 0 without it.
template <int VPL>
__global__ void __launch_bounds__(64)
splitk_reduce_kernel(const float* __restrict__ ws, int splitk, int M, int N, const float* __restrict__ bias, const void* res,
                     int res_dt, int64_t ldr, int act, const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                     void* out, int out_dt, int64_t ldo, void* out2, int out2_dt, int64_t ldo2) {
  pdl_wait();
  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 2 + warp;
  if (row >= M) return;
  const int64_t slab = (int64_t)M * N;
  const float* wrow = ws + (int64_t)row * N;
  auto load = [&](int col) {
    float4 a = *reinterpret_cast<const float4*>(wrow + col);
    for (int ks = 1; ks < splitk; ++ks) {
      const float4 b = *reinterpret_cast<const float4*>(wrow + ks * slab + col);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    if (bias) { const float4 b = __ldg(reinterpret_cast<const float4*>(bias + col)); a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
    if (act == ACAI_ACT_GELU) { a.x = gelu_fast(a.x); a.y = gelu_fast(a.y); a.z = gelu_fast(a.z); a.w = gelu_fast(a.w); }
    if (res) {
      const int64_t o = (int64_t)row * ldr + col;
      a.x += ld_as_float(res, o, res_dt); a.y += ld_as_float(res, o + 1, res_dt);
      a.z += ld_as_float(res, o + 2, res_dt); a.w += ld_as_float(res, o + 3, res_dt);
    }
    return a;
  };
  // out2 (optional): a second copy of the row in another dtype - the fp32 residual stream keeps `out` = fp32 and hands the
  // next GEMM its bf16 operand through `out2` (what torch.autocast does to a LayerNorm output that feeds a Linear)
  auto store = [&](int col, float4 a) {
    if (out2) {
      const int64_t o2 = (int64_t)row * ldo2 + col;
      if (out2_dt == ACAI_BF16) {
        const __nv_bfloat162 h0 = __floats2bfloat162_rn(a.x, a.y), h1 = __floats2bfloat162_rn(a.z, a.w);
        uint2 pk; pk.x = *reinterpret_cast<const uint32_t*>(&h0); pk.y = *reinterpret_cast<const uint32_t*>(&h1);
        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(out2) + o2) = pk;
      } else {
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(out2) + o2) = a;
      }
    }
    const int64_t o = (int64_t)row * ldo + col;
    if (out_dt == ACAI_BF16) {
      const __nv_bfloat162 h0 = __floats2bfloat162_rn(a.x, a.y), h1 = __floats2bfloat162_rn(a.z, a.w);
      uint2 pk; pk.x = *reinterpret_cast<const uint32_t*>(&h0); pk.y = *reinterpret_cast<const uint32_t*>(&h1);
      *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(out) + o) = pk;
    } else {
      st_from_float(out, o, out_dt, a.x); st_from_float(out, o + 1, out_dt, a.y);
      st_from_float(out, o + 2, out_dt, a.z); st_from_float(out, o + 3, out_dt, a.w);
    }
  };
  if constexpr (VPL == 0) {
    for (int col = lane * 4; col < N; col += 128) store(col, load(col));
  } else {
    // all loads of one K slice are issued before any is consumed (a per-piece loop over the slices would serialise
    // splitk * VPL L2 round trips per lane)
    float4 v[VPL];
#pragma unroll
    for (int c = 0; c < VPL; ++c) v[c] = *reinterpret_cast<const float4*>(wrow + (c * 32 + lane) * 4);
    if (splitk == 4) {                               // the decode step's case: every slice's loads in flight at once
      float4 t[3][VPL];
#pragma unroll
      for (int ks = 1; ks < 4; ++ks) {
#pragma unroll
        for (int c = 0; c < VPL; ++c) t[ks - 1][c] = *reinterpret_cast<const float4*>(wrow + ks * slab + (c * 32 + lane) * 4);
      }
#pragma unroll
      for (int ks = 1; ks < 4; ++ks) {               // (same summation order as the generic loop)
#pragma unroll
        for (int c = 0; c < VPL; ++c) { v[c].x += t[ks - 1][c].x; v[c].y += t[ks - 1][c].y; v[c].z += t[ks - 1][c].z; v[c].w += t[ks - 1][c].w; }
      }
    } else {
      for (int ks = 1; ks < splitk; ++ks) {
        float4 t[VPL];
#pragma unroll
        for (int c = 0; c < VPL; ++c) t[c] = *reinterpret_cast<const float4*>(wrow + ks * slab + (c * 32 + lane) * 4);
#pragma unroll
        for (int c = 0; c < VPL; ++c) { v[c].x += t[c].x; v[c].y += t[c].y; v[c].z += t[c].z; v[c].w += t[c].w; }
      }
    }
    if (bias) {
#pragma unroll
      for (int c = 0; c < VPL; ++c) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(bias + (c * 32 + lane) * 4));
        v[c].x += b.x; v[c].y += b.y; v[c].z += b.z; v[c].w += b.w;
      }
    }
    if (act == ACAI_ACT_GELU) {
#pragma unroll
      for (int c = 0; c < VPL; ++c) { v[c].x = gelu_fast(v[c].x); v[c].y = gelu_fast(v[c].y); v[c].z = gelu_fast(v[c].z); v[c].w = gelu_fast(v[c].w); }
    }
    if (res) {
      if (res_dt == ACAI_BF16 && ldr % 4 == 0 && (reinterpret_cast<uintptr_t>(res) & 7) == 0) {
        uint2 r[VPL];
#pragma unroll
        for (int c = 0; c < VPL; ++c)
          r[c] = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(res) + (int64_t)row * ldr + (c * 32 + lane) * 4));
#pragma unroll
        for (int c = 0; c < VPL; ++c) {
          const float2 f0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&r[c].x));
          const float2 f1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&r[c].y));
          v[c].x += f0.x; v[c].y += f0.y; v[c].z += f1.x; v[c].w += f1.y;
        }
      } else {
#pragma unroll
        for (int c = 0; c < VPL; ++c) {
          const int64_t o = (int64_t)row * ldr + (c * 32 + lane) * 4;
          v[c].x += ld_as_float(res, o, res_dt); v[c].y += ld_as_float(res, o + 1, res_dt);
          v[c].z += ld_as_float(res, o + 2, res_dt); v[c].w += ld_as_float(res, o + 3, res_dt);
        }
      }
    }
    if (!gamma) {
#pragma unroll
      for (int c = 0; c < VPL; ++c) store((c * 32 + lane) * 4, v[c]);
      return;
    }
    float sum = 0.0f;
#pragma unroll
    for (int c = 0; c < VPL; ++c) sum += (v[c].x + v[c].y) + (v[c].z + v[c].w);
    const float mean = warp_sum(sum) / (float)N;
    float sq = 0.0f;
#pragma unroll
    for (int c = 0; c < VPL; ++c) {
      const float a = v[c].x - mean, b = v[c].y - mean, cc = v[c].z - mean, d = v[c].w - mean;
      sq += (a * a + b * b) + (cc * cc + d * d);
    }
    const float rstd = rsqrtf(warp_sum(sq) / (float)N + eps);
#pragma unroll
    for (int c = 0; c < VPL; ++c) {
      const int col = (c * 32 + lane) * 4;
      const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + col)), b = __ldg(reinterpret_cast<const float4*>(beta + col));
      store(col, make_float4((v[c].x - mean) * rstd * g.x + b.x, (v[c].y - mean) * rstd * g.y + b.y,
                             (v[c].z - mean) * rstd * g.z + b.z, (v[c].w - mean) * rstd * g.w + b.w));
    }
  }
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

// 2-D row-major tensor [rows, cols] (ld elements between rows) of bf16 or fp32, box = [box_rows, box_cols], 128B swizzle
// (box_cols * element size must be 128 bytes).
int make_tma_2d(CUtensorMap* map, const void* ptr, int dtype, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                uint32_t box_cols) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled is unavailable (driver too old?)"); return ACAI_ERR_CUDA; }
  struct Key { const void* p; uint64_t r, c, l; uint32_t br, bc; int dt; };
  struct Hash { size_t operator()(const Key& k) const {
    size_t h = reinterpret_cast<size_t>(k.p);
    h ^= k.r * 0x9E3779B97F4A7C15ull + (h << 6); h ^= k.c * 0xC2B2AE3D27D4EB4Full + (h >> 3);
    h ^= k.l * 0x165667B19E3779F9ull; h ^= ((size_t)k.br << 32 | k.bc) * 0x27D4EB2F165667C5ull; h ^= (size_t)k.dt << 57; return h; } };
  struct Eq { bool operator()(const Key& a, const Key& b) const {
    return a.p == b.p && a.r == b.r && a.c == b.c && a.l == b.l && a.br == b.br && a.bc == b.bc && a.dt == b.dt; } };
  static std::unordered_map<Key, CUtensorMap, Hash, Eq> cache;
  static std::mutex mu;
  Key key{ptr, rows, cols, ld, box_rows, box_cols, dtype};
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *map = it->second; return ACAI_OK; }
  }
  const uint64_t es = dtype == ACAI_BF16 ? 2 : 4;
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * es};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, dtype == ACAI_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                   const_cast<void*>(ptr), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return ACAI_ERR_CUDA; }
  std::lock_guard<std::mutex> g(mu);
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *map);
  return ACAI_OK;
}

int make_tma_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                     uint32_t box_cols) {
  return make_tma_2d(map, ptr, ACAI_BF16, rows, cols, ld, box_rows, box_cols);
}

template <int BN, int OUT, int RES, int MODE = 0, bool BKN = false, int EW = EPI_WARPS>
static int launch_variant(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const CUtensorMap& taux,
                          const Epi& ep, int M, int N, int K, cudaStream_t st) {
  static bool attr_set = false;
  constexpr size_t smem = smem_bytes<BN, EW>();
  if (!attr_set) {
    ACAI_CUDA(cudaFuncSetAttribute(gemm_bf16_tcgen05<BN, OUT, RES, MODE, BKN, EW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int grid = tiles < num_sms() ? tiles : num_sms();
  ACAI_CUDA(launch_pdl(gemm_bf16_tcgen05<BN, OUT, RES, MODE, BKN, EW>, dim3(grid), dim3(64 + EW * 32), smem, st, ta, tb, to, taux, ep, M, N, K));
  return launch_status("gemm_bf16_tcgen05");
}

// Sixteen epilogue warps for the 256-column ACT_GELU_TRAIN epilogue (see smem_bytes).  Measured in one box
// (profiles/r2f_training_step.md): 63 664 x 3 072 x 768: 401 -> 325 us, 41 976 x 4 096 x 1 024: 373 -> 324 us - on by default.
// The same on ACT_GELU_BWD is SLOWER (401 -> 592 us with both accumulator halves live: 576 threads are allocated as 640, 96
// registers per thread, spills; 388 -> 413 us one half at a time, no spills) and on the output-dropout epilogue too
// (90 -> 125 us): both stay on eight warps.  ACAI_GEMM_EPI16=0 selects eight warps everywhere (A/B runs, read per call).
static bool epi16(int bit) {
  const char* e = getenv("ACAI_GEMM_EPI16");
  return (((e ? atoi(e) : GEMM_EPI16_DEFAULT) >> bit) & 1) != 0;
}

// dX = dZ[M, K] . W[K, N] (+ residual, or x gelu'(z) x mask): bf16 in, bf16 out through the TMA-store epilogue only
template <int BN>
static int launch_gemm_kn(const void* a, int64_t lda, const void* w, int64_t ldw, const Epi& ep, int M, int N, int K, cudaStream_t st) {
  CUtensorMap ta, tb, to;
  int rc = make_tma_2d_bf16(&ta, a, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BM, BK);
  if (rc) return rc;
  rc = make_tma_2d_bf16(&tb, w, (uint64_t)K, (uint64_t)N, (uint64_t)ldw, 64, 64);
  if (rc) return rc;
  rc = make_tma_2d(&to, ep.out, ACAI_BF16, (uint64_t)M, (uint64_t)N, (uint64_t)ep.ldo, 32, 64);
  if (rc) return rc;
  if (ep.act == ACT_GELU_BWD) return launch_variant<BN, 1, 1, 2, true>(ta, tb, to, to, ep, M, N, K, st);
  if (!ep.res) return launch_variant<BN, 1, 0, 0, true>(ta, tb, to, to, ep, M, N, K, st);
  return launch_variant<BN, 1, 1, 0, true>(ta, tb, to, to, ep, M, N, K, st);
}

template <int BN>
static int launch_gemm(const void* a, int64_t lda, const void* w, int64_t ldw, const Epi& ep, int M, int N, int K,
                       cudaStream_t st) {
  CUtensorMap ta, tb, to;
  int rc = make_tma_2d_bf16(&ta, a, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BM, BK);
  if (rc) return rc;
  rc = make_tma_2d_bf16(&tb, w, (uint64_t)N, (uint64_t)K, (uint64_t)ldw, BN, BK);
  if (rc) return rc;
  // TMA-store epilogue: 16-byte aligned output rows, whole 16-byte residual pieces
  const int oes = ep.out_dt == ACAI_BF16 ? 2 : 4, res_es = ep.res_dt == ACAI_BF16 ? 2 : 4;
  const bool tma_ok = N % 8 == 0 && (ep.ldo * oes) % 16 == 0 && (reinterpret_cast<uintptr_t>(ep.out) & 15) == 0 &&
                      (!ep.res || ((ep.ldr * res_es) % 16 == 0 && (reinterpret_cast<uintptr_t>(ep.res) & 15) == 0));
  // (BN = 64 serves small / skinny problems only; its 32-column warp slabs are narrower than a 128-byte bf16 chunk)
  const bool train_act = ep.act == ACT_GELU_TRAIN || ep.act == ACT_GELU_BWD;   // only the TMA-store epilogue has these
  if (BN < 128 || !tma_ok) return (train_act || ep.drop_thr) ? ACAI_ERR_UNSUPPORTED : launch_variant<BN, 0, 0>(ta, tb, ta, ta, ep, M, N, K, st);
  if constexpr (BN >= 128) {
  rc = make_tma_2d(&to, ep.out, ep.out_dt, (uint64_t)M, (uint64_t)N, (uint64_t)ep.ldo, 32, ep.out_dt == ACAI_BF16 ? 64 : 32);
  if (rc) return rc;
  const int res = !ep.res ? 0 : (ep.res_dt == ACAI_BF16 ? 1 : 2);
  if (ep.act == ACT_GELU_TRAIN) {
    if (ep.out_dt != ACAI_BF16 || res != 0 || (ep.ld_aux * 2) % 16 != 0 || (reinterpret_cast<uintptr_t>(ep.aux) & 15))
      return ACAI_ERR_UNSUPPORTED;
    CUtensorMap taux;
    rc = make_tma_2d(&taux, ep.aux, ACAI_BF16, (uint64_t)M, (uint64_t)N, (uint64_t)ep.ld_aux, 32, 64);
    if (rc) return rc;
    if constexpr (BN == 256) { if (epi16(0)) return launch_variant<BN, 1, 0, 1, false, 16>(ta, tb, to, taux, ep, M, N, K, st); }
    return launch_variant<BN, 1, 0, 1>(ta, tb, to, taux, ep, M, N, K, st);
  }
  if (ep.act == ACT_GELU_BWD) {
    if (ep.out_dt != ACAI_BF16 || res != 1 || ep.res_rows) return ACAI_ERR_UNSUPPORTED;
    return launch_variant<BN, 1, 1, 2>(ta, tb, to, to, ep, M, N, K, st);
  }
  if (ep.out_dt == ACAI_BF16) {
    if (res == 0) return launch_variant<BN, 1, 0>(ta, tb, to, to, ep, M, N, K, st);
    if (res == 1) return launch_variant<BN, 1, 1>(ta, tb, to, to, ep, M, N, K, st);   // (16 warps on the output-dropout epilogue: 1.08 -> 1.50 ms, spills)
    return launch_variant<BN, 1, 2>(ta, tb, to, to, ep, M, N, K, st);
  }
  if (res == 0) return launch_variant<BN, 2, 0>(ta, tb, to, to, ep, M, N, K, st);
  if (res == 1) return launch_variant<BN, 2, 1>(ta, tb, to, to, ep, M, N, K, st);
  return launch_variant<BN, 2, 2>(ta, tb, to, to, ep, M, N, K, st);
  }
  return ACAI_ERR_UNSUPPORTED;
}

static int dispatch_tiles(const void* a, int64_t lda, const void* w, int64_t ldw, const Epi& ep, int M, int N, int K,
                          cudaStream_t st);

int linear_tcgen05_ex(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* res,
                      int res_dt, int64_t ldr, const int32_t* res_rows, void* out, int out_dt, int64_t ldo, int M, int N,
                      int K, int act, cudaStream_t st) {
  // qualification: TMA needs 16-byte aligned bases and row pitches; the K loop walks whole 64-element blocks
  if (K % BK != 0 || lda % 8 != 0 || ldw % 8 != 0 || (reinterpret_cast<uintptr_t>(a) & 15) ||
      (reinterpret_cast<uintptr_t>(w) & 15) || M < 1 || N < 8)
    return ACAI_ERR_UNSUPPORTED;
  Epi ep{bias, res, res_dt, ldr, out, out_dt, ldo, act, res_rows, nullptr, 0, 0u, 1.0f, 0ull};
  return dispatch_tiles(a, lda, w, ldw, ep, M, N, K, st);
}

static int dispatch_tiles(const void* a, int64_t lda, const void* w, int64_t ldw, const Epi& ep, int M, int N, int K,
                          cudaStream_t st) {
  // widest N tile that still yields about one tile per SM; skinny decode-step problems get narrow tiles so that
  // more SMs stream the weight matrix
  const int m_tiles = (M + BM - 1) / BM, sms = num_sms();
  auto tiles = [&](int bn) { return m_tiles * ((N + bn - 1) / bn); };
  // (Round 2f tried a wave-quantisation rule here - 128-column tiles when the last round of 256-column tiles would be mostly
  // empty, e.g. MAE encoder 7 962 x 768: 189 tiles = 2 rounds at 64 % - and measured NO gain on the MAE step of the bench
  // (20.29 vs 19.85 ms, in-box A/B): the narrower tile re-reads A twice as often and pays twice the epilogues.  Not kept.)
  if (N > 128 && tiles(256) >= sms) return launch_gemm<256>(a, lda, w, ldw, ep, M, N, K, st);
  if ((N > 64 && tiles(128) >= sms) || ep.act >= ACT_GELU_TRAIN) return launch_gemm<128>(a, lda, w, ldw, ep, M, N, K, st);
  return launch_gemm<64>(a, lda, w, ldw, ep, M, N, K, st);
}

static bool gemm_operands_ok(const void* a, int64_t lda, const void* w, int64_t ldw, int M, int N, int K) {
  return !(K % BK != 0 || lda % 8 != 0 || ldw % 8 != 0 || (reinterpret_cast<uintptr_t>(a) & 15) ||
           (reinterpret_cast<uintptr_t>(w) & 15) || M < 1 || N < 8);
}

// z = a @ w^T + bias (bf16, kept for the backward) and h = dropout(gelu(z)) from ONE pass over the accumulators.
int linear_gelu_train_tcgen05(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, void* z, int64_t ldz,
                              void* h, int64_t ldh, int M, int N, int K, float dropout_p, uint64_t seed, cudaStream_t st) {
  if (!gemm_operands_ok(a, lda, w, ldw, M, N, K) || N % 2 != 0) return ACAI_ERR_UNSUPPORTED;
  const uint32_t thr = dropout_threshold(dropout_p);
  Epi ep{bias, nullptr, ACAI_BF16, 0, h, ACAI_BF16, ldh, ACT_GELU_TRAIN, nullptr, z, ldz, thr, thr ? 1.0f / (1.0f - dropout_p) : 1.0f,
         seed};
  if (ldh != N && thr) return ACAI_ERR_UNSUPPORTED;   // the dropout mask is indexed row * N + col like acai_gelu_fwd on a dense tensor
  return dispatch_tiles(a, lda, w, ldw, ep, M, N, K, st);
}

// dz = (dh @ wt^T) * gelu'(z) * mask: the feed-forward's second GEMM backward with the GELU backward in its epilogue.
int linear_gelu_bwd_tcgen05(const void* dy, int64_t ldd, const void* wt, int64_t ldw, const void* z, int64_t ldz, void* dz,
                            int64_t ldo, int M, int N, int K, float dropout_p, uint64_t seed, cudaStream_t st) {
  if (!gemm_operands_ok(dy, ldd, wt, ldw, M, N, K) || N % 2 != 0) return ACAI_ERR_UNSUPPORTED;
  const uint32_t thr = dropout_threshold(dropout_p);
  Epi ep{nullptr, z, ACAI_BF16, ldz, dz, ACAI_BF16, ldo, ACT_GELU_BWD, nullptr, nullptr, 0, thr,
         thr ? 1.0f / (1.0f - dropout_p) : 1.0f, seed};
  return dispatch_tiles(dy, ldd, wt, ldw, ep, M, N, K, st);
}

// dX[M, N] = dZ[M, K] . W[K, N] with W as the forward keeps it ([out = K, in = N] rows): + residual (bf16), or, with z,
// x gelu'(z) x dropout mask (the ACT_GELU_BWD epilogue).  ACAI_ERR_UNSUPPORTED unless TMA serves every operand.
int linear_dgrad_tcgen05(const void* dz, int64_t ldz, const void* w, int64_t ldw, const void* res, int64_t ldr, const void* z,
                         int64_t ldzz, void* out, int64_t ldo, int M, int N, int K, float dropout_p, uint64_t seed, cudaStream_t st) {
  auto mis = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) != 0; };
  if (K % BK != 0 || ldz % 8 != 0 || ldw % 8 != 0 || ldo % 8 != 0 || N % 8 != 0 || N < 128 || M < 1 || mis(dz) || mis(w) || mis(out) ||
      (res && (mis(res) || ldr % 8 != 0)) || (z && (mis(z) || ldzz % 8 != 0 || res)))
    return ACAI_ERR_UNSUPPORTED;
  const uint32_t thr = dropout_threshold(dropout_p);
  Epi ep{nullptr, z ? z : res, ACAI_BF16, z ? ldzz : ldr, out, ACAI_BF16, ldo, z ? ACT_GELU_BWD : ACAI_ACT_NONE, nullptr, nullptr, 0,
         z ? thr : 0u, (z && thr) ? 1.0f / (1.0f - dropout_p) : 1.0f, seed};
  const int m_tiles = (M + BM - 1) / BM;
  if (N > 128 && m_tiles * ((N + 255) / 256) >= num_sms()) return launch_gemm_kn<256>(dz, ldz, w, ldw, ep, M, N, K, st);
  return launch_gemm_kn<128>(dz, ldz, w, ldw, ep, M, N, K, st);
}

// out = residual + dropout(a @ w^T + bias), bf16, dense [M, N] output (the mask counter is row * N + column, the one
// acai_dropout uses on the same tensor).  ACAI_ERR_UNSUPPORTED unless the TMA-store epilogue serves the problem.
int linear_dropout_tcgen05(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* res, int res_dt,
                           int64_t ldr, void* out, int64_t ldo, int M, int N, int K, float dropout_p, uint64_t seed, cudaStream_t st) {
  if (!gemm_operands_ok(a, lda, w, ldw, M, N, K) || N % 32 != 0 || ldo != N) return ACAI_ERR_UNSUPPORTED;
  const uint32_t thr = dropout_threshold(dropout_p);
  Epi ep{bias, res, res_dt, ldr, out, ACAI_BF16, ldo, ACAI_ACT_NONE, nullptr, nullptr, 0, thr, thr ? 1.0f / (1.0f - dropout_p) : 1.0f, seed};
  return dispatch_tiles(a, lda, w, ldw, ep, M, N, K, st);
}

int linear_tcgen05(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* res,
                   int res_dt, int64_t ldr, void* out, int out_dt, int64_t ldo, int M, int N, int K, int act,
                   cudaStream_t st) {
  return linear_tcgen05_ex(a, lda, w, ldw, bias, res, res_dt, ldr, nullptr, out, out_dt, ldo, M, N, K, act, st);
}

// ---- small-M projections: split-K GEMM + fused reduce ----------------------------------------------------------------
static int splitk_factor(int M, int N, int K) {
  const int base = ((M + BM - 1) / BM) * ((N + 63) / 64), num_k = K / BK;
  int sk = num_sms() / (base > 0 ? base : 1);
  if (sk > 8) sk = 8;
  if (sk > num_k / 2) sk = num_k / 2;
  return sk < 1 ? 1 : sk;
}

size_t linear_splitk_workspace_bytes(int M, int N, int K) {
  return (size_t)splitk_factor(M, N, K) * (size_t)M * (size_t)N * sizeof(float);
}

// out = LN?( act(a . w^T + bias) + residual ) for M <= 256 rows.  ACAI_ERR_UNSUPPORTED (nothing written) when the problem
// does not qualify: operands not TMA-able, no K slices to gain, N % 4, LayerNorm rows that do not fit the register variant.
int linear_splitk_tcgen05(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* res, int res_dt,
                          int64_t ldr, const float* gamma, const float* beta, float eps, void* out, int out_dt, int64_t ldo,
                          void* out2, int out2_dt, int64_t ldo2, int M,
                          int N, int K, int act, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (!gemm_operands_ok(a, lda, w, ldw, M, N, K) || M > 256 || N % 4 != 0 || (reinterpret_cast<uintptr_t>(ws) & 15) ||
      (out_dt == ACAI_BF16 && ((ldo % 4) || (reinterpret_cast<uintptr_t>(out) & 7))))
    return ACAI_ERR_UNSUPPORTED;
  if (out2 && ((ldo2 % 4) || (reinterpret_cast<uintptr_t>(out2) & (out2_dt == ACAI_BF16 ? 7 : 15)))) return ACAI_ERR_UNSUPPORTED;
  const int sk = splitk_factor(M, N, K);
  const bool in_regs = N % 128 == 0 && N / 128 <= 8;     // the row fits the register-resident reduce variants
  const int vpl = in_regs ? N / 128 : 0;
  if (sk < 2 || (gamma && (!in_regs || !beta))) return ACAI_ERR_UNSUPPORTED;
  if (ws_bytes < (size_t)sk * M * N * sizeof(float)) return ACAI_ERR_UNSUPPORTED;
  const int num_k = K / BK, kper = (num_k + sk - 1) / sk;
  Epi ep{nullptr, nullptr, ACAI_BF16, 0, ws, ACAI_F32, N, ACAI_ACT_NONE, nullptr, nullptr, 0, 0u, 1.0f, 0ull,
         (num_k + kper - 1) / kper, kper, static_cast<float*>(ws)};
  CUtensorMap ta, tb;
  int rc = make_tma_2d_bf16(&ta, a, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BM, BK);
  if (rc) return rc;
  rc = make_tma_2d_bf16(&tb, w, (uint64_t)N, (uint64_t)K, (uint64_t)ldw, 64, BK);
  if (rc) return rc;
  {
    static bool attr_set = false;
    constexpr size_t smem = smem_bytes<64>();
    if (!attr_set) {
      ACAI_CUDA(cudaFuncSetAttribute(gemm_bf16_tcgen05<64, 0, 0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr_set = true;
    }
    const int tiles = ((M + BM - 1) / BM) * ((N + 63) / 64) * ep.splitk;
    const int grid = tiles < num_sms() ? tiles : num_sms();
    ACAI_CUDA(launch_pdl(gemm_bf16_tcgen05<64, 0, 0, 0>, dim3(grid), dim3(THREADS), smem, st, ta, tb, ta, ta, ep, M, N, K));
    rc = launch_status("gemm_bf16_tcgen05 (split-K)");
    if (rc) return rc;
  }
  const dim3 grid((M + 1) / 2);
#define ACAI_RED(V) ACAI_CUDA(launch_pdl(splitk_reduce_kernel<V>, grid, dim3(64), 0, st, static_cast<const float*>(ws), ep.splitk, M, N, \
                                        bias, res, res_dt, ldr, act, gamma, beta, eps, out, out_dt, ldo, out2, out2_dt, ldo2))
  switch (vpl) {
    case 0: ACAI_RED(0); break;
    case 1: ACAI_RED(1); break;
    case 2: ACAI_RED(2); break;
    case 3: ACAI_RED(3); break;
    case 4: ACAI_RED(4); break;
    case 5: ACAI_RED(5); break;
    case 6: ACAI_RED(6); break;
    case 7: ACAI_RED(7); break;
    default: ACAI_RED(8); break;
  }
#undef ACAI_RED
  return launch_status("splitk_reduce_kernel");
}

}  // namespace acai
