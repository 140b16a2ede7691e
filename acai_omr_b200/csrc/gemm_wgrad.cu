// Weight-gradient GEMM on tcgen05:   dW[N_out, K_in] += dZ[T, N_out]^T . X[T, K_in]      (bf16 operands, fp32 out)
//
// Both operands are read exactly as the forward / backward left them in HBM - token-major rows - so the contraction
// dimension (tokens) is the ROW index of both: "MN-major" A and B in UMMA terms.  TMA brings [64 tokens x 64 features]
// boxes (128-byte swizzled rows); one pipeline stage holds BM/64 such boxes of dZ and BN/64 of X, and the MMA
// descriptors walk them with LBO = 8 KB (next 64-feature block) and SBO = 1 KB (next 8 tokens).  No transposed copies
// of activations or gradients are ever materialised (the first version of the training step spent 12 % of its time
// transposing).
//
// The output is small (a weight matrix) and the contraction very long (all tokens of the batch), so the token range
// is split over CTAs (split-K) until every SM has work; partial tiles are added into the fp32 gradient buffer with
// TMA reduce-add (cp.reduce.async.bulk.tensor .add), which is also what makes gradient ACCUMULATION free: the
// destination is the parameter's slice of the flat gradient buffer itself.
// Ragged edges: TMA zero-fills token rows past T and feature columns past N_out / K_in on load and clips on store.
//   warp 0: TMA producer    warp 1: MMA issuer    warps 2..9: epilogue (TMEM -> swizzled smem -> TMA reduce-add)
//
// BIAS variant (acai_wgrad_bias_accumulate): the bias gradient db[n] = sum over tokens of dZ[t, n] is the column sum of
// the very dZ boxes the pipeline already holds in shared memory.  Four more warps (10..13) follow the stage ring; on the
// items of feature tile 0 (every (dZ row tile, token range) pair has exactly one) they add up the stage's [64 tokens x 128
// features] boxes - thread = (16-byte chunk of a 128-byte row, token row mod 8), which is conflict-free under the 128-byte
// swizzle - fold the eight row classes with shuffles when the item ends and atomicAdd 128 sums into db.  The stage is
// released by the MMAs' commit AND these warps (empty barrier count 5).  The separate column-sum pass over dZ (a whole
// extra read of every gradient tensor: 5.7 ms of the 158 ms teacher-forced step) disappears.
#include <cstdlib>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace acai {

int make_tma_2d(CUtensorMap* map, const void* ptr, int dtype, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                uint32_t box_cols);

namespace wg {
constexpr int BM = 128, BK = 64, EPI_WARPS = 8, THREADS = 64 + EPI_WARPS * 32, STAGES = 4, BIAS_WARPS = 4;
constexpr uint32_t BOX = 64 * 64 * 2;            // one [64 tokens x 64 features] bf16 box = 8 KB
template <int BN> constexpr size_t smem_bytes() { return (size_t)STAGES * (BM / 64 + BN / 64) * BOX + EPI_WARPS * 4096 + 256; }
}  // namespace wg

template <int BN, bool BIAS>
__global__ void __launch_bounds__(wg::THREADS + (BIAS ? wg::BIAS_WARPS * 32 : 0), 1)
gemm_wgrad_tcgen05(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                   const __grid_constant__ CUtensorMap tma_out, float* __restrict__ dbias, int M, int N, int T, int splits) {
  using namespace wg;
  constexpr uint32_t A_BYTES = (BM / 64) * BOX, B_BYTES = (BN / 64) * BOX;
  constexpr uint32_t TMEM_COLS = 2 * BN;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_BYTES;
  uint8_t* staging = smem + STAGES * (A_BYTES + B_BYTES);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + EPI_WARPS * 4096);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // (provably warp-uniform: ptx::elect_one_sync)
  const int num_m = (M + BM - 1) / BM, num_n = (N + BN - 1) / BN, num_tiles = num_m * num_n;
  const int nk = (T + BK - 1) / BK, kper = (nk + splits - 1) / splits;
  const int num_items = num_tiles * splits;          // item = split * num_tiles + tile: neighbours share a token range

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tma_a); ptx::prefetch_tensormap(&tma_b); ptx::prefetch_tensormap(&tma_out);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&full_bar[s], 1); ptx::mbar_init(&empty_bar[s], BIAS ? 1 + BIAS_WARPS : 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(&tfull_bar[s], 1); ptx::mbar_init(&tempty_bar[s], EPI_WARPS); }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // (producer / issuer: converged loops, one elected lane issues - see ptx::elect_one_sync)
  if (warp == 0) {
    int stage = 0; uint32_t phase = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int tile = item % num_tiles, split = item / num_tiles;
      const int m0 = (tile / num_n) * BM, n0 = (tile % num_n) * BN;
      const int kb0 = split * kper, kb1 = min(nk, kb0 + kper);
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
        if (ptx::elect_one_sync()) {
          ptx::mbar_arrive_expect_tx(&full_bar[stage], A_BYTES + B_BYTES);
#pragma unroll
          for (int i = 0; i < BM / 64; ++i)
            ptx::tma_load_2d(smem_a + stage * A_BYTES + i * BOX, &tma_a, &full_bar[stage], m0 + 64 * i, kb * BK);
#pragma unroll
          for (int i = 0; i < BN / 64; ++i)
            ptx::tma_load_2d(smem_b + stage * B_BYTES + i * BOX, &tma_b, &full_bar[stage], n0 + 64 * i, kb * BK);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t IDESC = ptx::idesc_bf16_f32(BM, BN, true, true);
    const uint32_t tb0 = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint64_t adesc0 = ptx::smem_desc_sw128(ptx::smem_u32(smem_a), 1024, BOX);
    const uint64_t bdesc0 = ptx::smem_desc_sw128(ptx::smem_u32(smem_b), 1024, BOX);
    int stage = 0; uint32_t phase = 0; int it = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int split = item / num_tiles;
      const int kb0 = split * kper, kb1 = min(nk, kb0 + kper);
      if (kb1 <= kb0) continue;                      // empty token range: nothing to add (all roles skip alike)
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      ++it;
      ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
      ptx::tc_fence_after();
      const uint32_t d_tmem = tb0 + as * BN;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        if (ptx::elect_one_sync()) {
          const uint64_t adesc = adesc0 + (uint64_t)(stage * (A_BYTES >> 4)), bdesc = bdesc0 + (uint64_t)(stage * (B_BYTES >> 4));
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)             // 16 tokens = 16 rows of 128 bytes
            ptx::umma_f16_ss(d_tmem, adesc + (uint64_t)k * (16 * 128 >> 4), bdesc + (uint64_t)k * (16 * 128 >> 4), IDESC,
                             (kb > kb0 || k != 0));
          ptx::umma_commit(&empty_bar[stage]);
          if (kb == kb1 - 1) ptx::umma_commit(&tfull_bar[as]);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (BIAS && warp >= 2 + EPI_WARPS) {
    // ---- bias-gradient column sums of the dZ boxes (see the header comment)
    const int tid = (int)threadIdx.x - THREADS;                  // 0..127
    const int rg = tid & 7, cc = tid >> 3;                        // token row class, 16-byte chunk: box cc >> 3, chunk cc & 7
    const uint32_t off = (uint32_t)(cc >> 3) * BOX + (uint32_t)rg * 128 + (uint32_t)(((cc & 7) ^ rg) << 4);
    int stage = 0; uint32_t phase = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int tile = item % num_tiles, split = item / num_tiles;
      const int kb0 = split * kper, kb1 = min(nk, kb0 + kper);
      if (kb1 <= kb0) continue;
      const bool mine = tile % num_n == 0;
      float acc[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = 0.0f;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase);
        if (mine) {
          const uint8_t* a = smem_a + stage * A_BYTES + off;
#pragma unroll
          for (int k = 0; k < 8; ++k) {                           // token rows 8 k + rg of the stage
            const uint4 v = *reinterpret_cast<const uint4*>(a + k * 1024);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              acc[2 * q] += __uint_as_float(w[q] << 16);
              acc[2 * q + 1] += __uint_as_float(w[q] & 0xffff0000u);
            }
          }
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&empty_bar[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (mine) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 1);
          acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 2);
          acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 4);
        }
        const int f0 = (tile / num_n) * BM + cc * 8;              // first of this thread's eight dZ features
        if (rg == 0) {
#pragma unroll
          for (int e = 0; e < 8; ++e)
            if (f0 + e < M) atomicAdd(dbias + f0 + e, acc[e]);
        }
      }
    }
  } else {
    const int ew = warp - 2, lane_grp = warp & 3, half = ew >> 2;
    constexpr int CPW = BN / 2;
    uint8_t* const stage_buf = staging + ew * 4096;
    uint8_t* const srow = stage_buf + lane * 128;
    const uint32_t sw = (uint32_t)(lane & 7);
    int it = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int tile = item % num_tiles, split = item / num_tiles;
      const int kb0 = split * kper, kb1 = min(nk, kb0 + kper);
      if (kb1 <= kb0) continue;
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      ++it;
      const int m0 = (tile / num_n) * BM, n0 = (tile % num_n) * BN;
      const int row0 = m0 + lane_grp * 32, col0 = n0 + half * CPW;
      const int ncols = N - col0 < CPW ? N - col0 : CPW;
      const int nch = (row0 < M && ncols > 0) ? (ncols + 31) / 32 : 0;
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + as * BN + half * CPW;
      if (nch == 0) {
        ptx::tc_fence_before();
        if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
      }
#pragma unroll 1
      for (int c = 0; c < nch; ++c) {
        uint32_t acc[32];
        ptx::tmem_ld_32x32(taddr + c * 32, acc);
        ptx::tmem_ld_wait();
        if (c == nch - 1) {
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
        }
        if (lane == 0) ptx::tma_store_wait_read<0>();
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 8; ++q)
          *reinterpret_cast<uint4*>(srow + ((((uint32_t)q) ^ sw) << 4)) = make_uint4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          ptx::tma_reduce_add_2d(&tma_out, stage_buf, col0 + c * 32, row0);
          ptx::tma_store_commit();
        }
      }
    }
    if (lane == 0) ptx::tma_store_wait<0>();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

template <int BN, bool BIAS>
static int launch_wgrad(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, float* dbias, int M, int N, int T,
                        cudaStream_t st) {
  using namespace wg;
  static bool attr_set = false;
  constexpr size_t smem = smem_bytes<BN>();
  if (!attr_set) {
    ACAI_CUDA(cudaFuncSetAttribute(gemm_wgrad_tcgen05<BN, BIAS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN), nk = (T + BK - 1) / BK, sms = num_sms();
  // How many token ranges?  The persistent CTAs take the (tile, range) items in rounds of #SMs and every item costs its
  // k-blocks plus an epilogue (a 128 x BN fp32 reduce-add, about a dozen k-blocks' worth): pick the split that minimises
  // rounds x (k-blocks per item + epilogue), each range at least 8 stages long.  (Round 1's rule - just over two waves of
  // items - left a third, almost empty round: 768 x 768 ran 306 items on 148 SMs, 69 % busy, 897 TFLOP/s against 1 300 for
  // the shapes that happened to divide well.)  ACAI_WGRAD_SPLIT=0 restores that rule for A/B runs.
  static const bool tuned = [] { const char* e = getenv("ACAI_WGRAD_SPLIT"); return !e || atoi(e) != 0; }();
  int splits = (2 * sms + tiles - 1) / tiles;
  if (splits > nk / 8) splits = nk / 8;
  if (splits < 1) splits = 1;
  if (tuned) {
    const int max_splits = nk / 8 < 1 ? 1 : nk / 8;
    constexpr int EPI = 12;
    long best_cost = -1;
    // (a range is ONE fp32 accumulation chain in TMEM; the tensor core's adds truncate, so the error against fp64 grows
    // linearly with the chain: 1.1e-5 relative per 10 k tokens, profiles/ubench/wgrad_precision.py - chains stay <= 32 k tokens,
    // <= 4e-5, forty times below one bf16 rounding of the same sum)
    constexpr int MAX_KPER = 512;
    for (int sp = 1; sp <= max_splits && (long)tiles * sp <= 4L * sms + tiles; ++sp) {
      const int kp = (nk + sp - 1) / sp, real = (nk + kp - 1) / kp;        // (empty trailing ranges are dropped)
      if (kp > MAX_KPER && sp < max_splits) continue;
      const long rounds = ((long)tiles * real + sms - 1) / sms, cost = rounds * (kp + EPI);
      if (best_cost < 0 || cost < best_cost) { best_cost = cost; splits = real; }
    }
  }
  const int kper = (nk + splits - 1) / splits;
  splits = (nk + kper - 1) / kper;                     // drop empty trailing ranges
  const int items = tiles * splits;
  gemm_wgrad_tcgen05<BN, BIAS><<<items < sms ? items : sms, THREADS + (BIAS ? BIAS_WARPS * 32 : 0), smem, st>>>(ta, tb, to, dbias, M, N, T, splits);
  return launch_status("gemm_wgrad_tcgen05");
}

// dW[M = n_out, N = k_in] (fp32, ld ldo) += dz[T, n_out]^T . x[T, k_in]; ACAI_ERR_UNSUPPORTED if TMA cannot address it.
// dbias (optional, fp32 [n_out]): += column sums of dz, from the same pass.
int wgrad_tcgen05(const void* dz, int64_t ldz, const void* x, int64_t ldx, float* dw, int64_t ldo, float* dbias, int n_out, int k_in,
                  int T, cudaStream_t st) {
  auto mis = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) != 0; };
  if (ldz % 8 || ldx % 8 || ldo % 4 || mis(dz) || mis(x) || mis(dw) || n_out < 1 || k_in < 8 || T < 1) return ACAI_ERR_UNSUPPORTED;
  CUtensorMap ta, tb, to;
  int rc = make_tma_2d(&ta, dz, ACAI_BF16, (uint64_t)T, (uint64_t)n_out, (uint64_t)ldz, 64, 64);
  if (rc) return rc;
  if ((rc = make_tma_2d(&tb, x, ACAI_BF16, (uint64_t)T, (uint64_t)k_in, (uint64_t)ldx, 64, 64))) return rc;
  if ((rc = make_tma_2d(&to, dw, ACAI_F32, (uint64_t)n_out, (uint64_t)k_in, (uint64_t)ldo, 32, 32))) return rc;
  if (dbias) return k_in > 128 ? launch_wgrad<256, true>(ta, tb, to, dbias, n_out, k_in, T, st) : launch_wgrad<128, true>(ta, tb, to, dbias, n_out, k_in, T, st);
  return k_in > 128 ? launch_wgrad<256, false>(ta, tb, to, nullptr, n_out, k_in, T, st) : launch_wgrad<128, false>(ta, tb, to, nullptr, n_out, k_in, T, st);
}

}  // namespace acai
