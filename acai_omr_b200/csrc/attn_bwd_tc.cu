// K3 backward (tensor-core path): packed variable-length attention gradients for head_dim 64, bf16 operands.
//
// The forward saved L = log2-domain log-sum-exp per (head, query) and `attn_delta_kernel` (attn_bwd.cu) gives
// delta = rowsum(dO . O), so no online softmax is needed here: P = exp2(S * scale_log2 - L) is recomputed exactly.
// Two kernels, both with the forward kernel's shape (warps 0-3 elementwise math on TMEM rows, warp 4 TMA producer,
// warp 5 MMA issuer, 256 TMEM columns -> two CTAs per SM so one CTA's MMAs overlap the other's exp2/FMA phase):
//
//   dQ kernel   CTA = (sequence, head, 128 queries), loop over 64-key blocks j:
//       S  = Q K_j^T , dP = dO V_j^T          tcgen05.mma SS -> TMEM columns 0..63 / 64..127
//       dS = P . (dP - delta)                  thread = query row; bf16 written back over S (TS operand)
//       dQ += dS K_j                           tcgen05.mma TS, B = K_j MN-major from the same smem tile
//   dK/dV kernel CTA = (sequence, head, 128 keys), loop over 64-query blocks i (transposed problem, thread = key row):
//       S^T = K Q_i^T , dP^T = V dO_i^T        SS
//       P^T, dS^T                               bf16 over S^T / dP^T
//       dV += P^T dO_i , dK += dS^T Q_i        TS, B = dO_i / Q_i MN-major
// Seven GEMMs instead of the five of a fused single pass, but no atomics and no cross-CTA reduction: every output
// element has exactly one writer and a fixed summation order (bitwise reproducible gradients).
#include <cstdlib>
#include <type_traits>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace acai {

int make_tma_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                     uint32_t box_cols);

namespace fb {
constexpr int BIG = 128, SMALL = 64, D = 64, STAGES = 4;
constexpr int THREADS = 320;           // warps 0-7 elementwise (4 lane quarters x 2 column halves), 8 TMA, 9 MMA
constexpr uint32_t BIGTILE = BIG * D * 2;           // 16 KB
constexpr uint32_t SMALLTILE = SMALL * D * 2;       // 8 KB
constexpr uint32_t OFF_A0 = 0, OFF_A1 = BIGTILE;    // the two resident 128-row tiles (Q,dO or K,V)
constexpr uint32_t OFF_B0 = 2 * BIGTILE, OFF_B1 = OFF_B0 + STAGES * SMALLTILE;   // streamed 64-row tiles (K,V or Q,dO)
constexpr uint32_t OFF_VEC = OFF_B1 + STAGES * SMALLTILE;      // 2 x (64 L + 64 delta + 64 dropout row constants + 64 row bases)
constexpr uint32_t OFF_BAR = OFF_VEC + 2 * 256 * 4;
constexpr size_t SMEM = OFF_BAR + 256;
constexpr uint32_t TMEM_COLS = 256;                 // 0..63 S, 64..127 dP, 128..255 accumulators
}  // namespace fb

struct AttnBwdParams {
  const int32_t* cu_q; const int32_t* cu_k;
  const float* lse; const float* delta; int total_q;
  int causal; float scale, scale_log2;
  __nv_bfloat16* dq; __nv_bfloat16* dk; __nv_bfloat16* dv; int64_t dqs, dks, dvs;
  AttnDropout drop;             // the forward's attention-probability dropout, regenerated here from (seed, head, row, key)
};

__device__ __forceinline__ void store_row_bf16(__nv_bfloat16* dst, const uint32_t (&o)[32], float mul) {
  uint4* op = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int q4 = 0; q4 < 4; ++q4) {
    uint4 v4; __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&v4);
#pragma unroll
    for (int i = 0; i < 4; ++i)
      hh[i] = __floats2bfloat162_rn(__uint_as_float(o[q4 * 8 + 2 * i]) * mul, __uint_as_float(o[q4 * 8 + 2 * i + 1]) * mul);
    op[q4] = v4;
  }
}

// ------------------------------------------------------------------------------------------ dQ
// TMEM: S[0] 0..63, S[1] 64..127 (double-buffered: S_{j+1} is computed while block j is in the elementwise phase),
// dP 128..191 (single buffer, handed back as soon as the elementwise warps have it in registers), dQ 192..255.
template <int DH, bool DROP>
__global__ void __launch_bounds__(fb::THREADS, 2)
attn_bwd_dq_tc(const __grid_constant__ CUtensorMap tma_q, const __grid_constant__ CUtensorMap tma_do,
               const __grid_constant__ CUtensorMap tma_k, const __grid_constant__ CUtensorMap tma_v, AttnBwdParams p) {
  using namespace fb;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* a_full = bars;                     // Q and dO tiles
  uint64_t* b_full = bars + 1;                 // STAGES
  uint64_t* b_empty = b_full + STAGES;         // STAGES
  uint64_t* sp_full = b_empty + STAGES;        // 2: S (buffer j&1) and dP of block j are in TMEM
  uint64_t* ds_full = sp_full + 2;             // 2 x 256 arrivals: dS of block j written over S buffer j&1
  uint64_t* dp_free = ds_full + 2;             // 256 arrivals: dP of block j is in registers
  uint64_t* acc_full = dp_free + 1;            // last dQ MMA landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, qb = blockIdx.x;
  const int q0 = p.cu_q[b], lq = p.cu_q[b + 1] - q0, k0 = p.cu_k[b], lk = p.cu_k[b + 1] - k0;
  if (qb * BIG >= lq) return;
  const int shift = lk - lq;
  int kv_end = lk;
  if (p.causal) { const int last = qb * BIG + BIG + shift; kv_end = last < lk ? (last < 0 ? 0 : last) : lk; }
  const int n_kv = (kv_end + SMALL - 1) / SMALL;

  if (warp == 9 && lane == 0) {
    ptx::mbar_init(a_full, 1);
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&b_full[s], 1); ptx::mbar_init(&b_empty[s], 1); }
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(&sp_full[i], 1); ptx::mbar_init(&ds_full[i], 256); }
    ptx::mbar_init(dp_free, 256); ptx::mbar_init(acc_full, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 8) {
    if (lane == 0) {
      ptx::prefetch_tensormap(&tma_q); ptx::prefetch_tensormap(&tma_do);
      ptx::prefetch_tensormap(&tma_k); ptx::prefetch_tensormap(&tma_v);
    }
    ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // (helper warps: converged loops, one elected lane issues - see ptx::elect_one_sync)
  const int u_nkv = __shfl_sync(0xffffffffu, n_kv, 0);
  if (warp == 8) {
    if (u_nkv > 0) {
      const int hcol = (h * DH) & ~63;
      const int qrow = __shfl_sync(0xffffffffu, q0 + qb * BIG, 0);
      int krow = __shfl_sync(0xffffffffu, k0, 0);
      if (ptx::elect_one_sync()) {
        ptx::mbar_arrive_expect_tx(a_full, 2 * BIGTILE);
        ptx::tma_load_2d(smem + OFF_A0, &tma_q, a_full, hcol, qrow);
        ptx::tma_load_2d(smem + OFF_A1, &tma_do, a_full, hcol, qrow);
      }
      __syncwarp();
      int s = 0; uint32_t ph = 0;
      for (int j = 0; j < u_nkv; ++j, krow += SMALL) {
        ptx::mbar_wait(&b_empty[s], ph ^ 1);
        if (ptx::elect_one_sync()) {
          ptx::mbar_arrive_expect_tx(&b_full[s], 2 * SMALLTILE);
          ptx::tma_load_2d(smem + OFF_B0 + s * SMALLTILE, &tma_k, &b_full[s], hcol, krow);
          ptx::tma_load_2d(smem + OFF_B1 + s * SMALLTILE, &tma_v, &b_full[s], hcol, krow);
        }
        __syncwarp();
        if (++s == STAGES) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 9) {
    if (u_nkv > 0) {
      constexpr uint32_t IDESC_S = ptx::idesc_bf16_f32(BIG, SMALL);
      constexpr uint32_t IDESC_ACC = ptx::idesc_bf16_f32(BIG, DH, false, true);     // B MN-major
      const uint32_t tb0 = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint64_t hoff = (uint64_t)(((h * DH) & 63) >> 3);    // DH = 32: the tile holds a head pair, 64-byte start offset
      const uint64_t qdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_A0), 1024) + hoff;
      const uint64_t dodesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_A1), 1024) + hoff;
      const uint64_t kdesc0 = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_B0), 1024) + hoff;
      const uint64_t vdesc0 = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_B1), 1024) + hoff;
      auto kdesc_of = [&](int j) { return kdesc0 + (uint64_t)((j % STAGES) * (SMALLTILE >> 4)); };
      auto vdesc_of = [&](int j) { return vdesc0 + (uint64_t)((j % STAGES) * (SMALLTILE >> 4)); };
      auto issue_s = [&](int j) {                  // S_j into buffer j & 1 (waits for K_j / V_j to land)
        ptx::mbar_wait(&b_full[j % STAGES], (j / STAGES) & 1);
        ptx::tc_fence_after();
        if (ptx::elect_one_sync()) {
          const uint64_t kd = kdesc_of(j);
#pragma unroll
          for (int k = 0; k < DH / 16; ++k) ptx::umma_f16_ss(tb0 + (j & 1) * 64, qdesc + 2 * k, kd + 2 * k, IDESC_S, k != 0);
        }
        __syncwarp();
      };
      auto issue_dp = [&](int j) {                 // dP_j, then S_j and dP_j are announced together
        if (ptx::elect_one_sync()) {
          const uint64_t vd = vdesc_of(j);
#pragma unroll
          for (int k = 0; k < DH / 16; ++k) ptx::umma_f16_ss(tb0 + 128, dodesc + 2 * k, vd + 2 * k, IDESC_S, k != 0);
          ptx::umma_commit(&sp_full[j & 1]);
        }
        __syncwarp();
      };
      ptx::mbar_wait(a_full, 0);
      issue_s(0);
      issue_dp(0);
      if (u_nkv > 1) issue_s(1);
      for (int j = 0; j < u_nkv; ++j) {
        if (j + 1 < u_nkv) {
          ptx::mbar_wait(dp_free, j & 1);          // block j's dP is in registers: the buffer can take dP_{j+1}
          ptx::tc_fence_after();
          issue_dp(j + 1);
        }
        ptx::mbar_wait(&ds_full[j & 1], (j >> 1) & 1);
        ptx::tc_fence_after();
        if (ptx::elect_one_sync()) {
          const uint64_t kd = kdesc_of(j);
#pragma unroll
          for (int kk = 0; kk < SMALL / 16; ++kk)
            ptx::umma_f16_ts(tb0 + 192, tb0 + (j & 1) * 64 + (kk >> 1) * 32 + (kk & 1) * 8,
                             kd + (uint64_t)kk * (16 * 128 >> 4), IDESC_ACC, (j | kk) != 0);
          ptx::umma_commit(&b_empty[j % STAGES]);
          if (j == u_nkv - 1) ptx::umma_commit(acc_full);
        }
        __syncwarp();
        if (j + 2 < u_nkv) issue_s(j + 2);          // overwrites dS_j: tcgen05 ops execute in issue order
      }
    }
  } else {
    const int r = (warp & 3) * 32 + lane, half = warp >> 2;       // this thread: query row r, key columns half*32..+31 of a block
    const int qi = qb * BIG + r;
    const bool valid = qi < lq;
    const int my_end = !valid ? 0 : (p.causal ? max(0, min(lk, qi + shift + 1)) : lk);
    const uint32_t lane_addr = (uint32_t)((warp & 3) * 32) << 16;
    const float sl2 = p.scale_log2;
    const float L = valid ? p.lse[(int64_t)h * p.total_q + q0 + qi] : INFINITY;
    const float dl = valid ? p.delta[(int64_t)h * p.total_q + q0 + qi] : 0.0f;
    const uint64_t rowkey = (uint64_t)h * p.total_q + (uint64_t)(q0 + (valid ? qi : 0));
    const uint32_t rowconst = DROP ? attn_drop_rowconst(p.drop.seed, rowkey) : 0u;
    const uint32_t rowbase = DROP ? attn_drop_rowbase(rowkey) : 0u;
    const float rp = p.drop.rp;
    for (int j = 0; j < n_kv; ++j) {
      const uint32_t tmem_s = tmem_base + lane_addr + (j & 1) * 64 + half * 32;
      ptx::mbar_wait(&sp_full[j & 1], (j >> 1) & 1);
      ptx::tc_fence_after();
      uint32_t s[32], dp[32], pk[16];
      ptx::tmem_ld_32x32(tmem_base + lane_addr + 128 + half * 32, dp);
      ptx::tmem_ld_32x32(tmem_s, s);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      ptx::mbar_arrive(dp_free);
      const int kbase = j * SMALL + half * 32;
      // the fp32 arithmetic runs on packed pairs (FFMA2 / FADD2 / FMUL2: the same fma.rn / add.rn / mul.rn per element in half
      // the issue slots - the elementwise warps of this kernel are issue bound)
      const uint64_t sl2x = ptx::pack_f32x2(sl2, sl2), nLx = ptx::pack_f32x2(-L, -L), ndlx = ptx::pack_f32x2(-dl, -dl);
      auto exp_pair = [&](int i, float& p0, float& p1) {
        float x0, x1;
        ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), sl2x, nLx), x0, x1);
        p0 = fast_exp2(x0); p1 = fast_exp2(x1);
      };
      auto ds_pair = [&](float p0, float p1, uint64_t t2) -> uint32_t {   // bf16 pair of P . (t - delta)
        float d0, d1;
        ptx::unpack_f32x2(ptx::mul_f32x2(ptx::pack_f32x2(p0, p1), ptx::add_f32x2(t2, ndlx)), d0, d1);
        const __nv_bfloat162 hh = __floats2bfloat162_rn(d0, d1);
        return *reinterpret_cast<const uint32_t*>(&hh);
      };
      // Columns [vis, 32) of this thread's half are masked (causal diagonal, rows / keys past the sequence end).  A warp whose
      // 32 rows all see the whole half - every block but the diagonal and the last one - runs the variant without the
      // compare / select instructions.  With dropout: dS = P . (M dP / (1 - p) - delta), M = the forward's keep mask.
      const int vis = my_end - kbase;
      // the forward's 32 keep decisions of this row's keys kbase .. kbase + 31 (attn_drop_word: one prmt + and per element)
      const uint32_t dw = DROP ? attn_drop_word(rowconst, rowbase, (uint32_t)(kbase >> 5), p.drop.thr) : 0u;
      auto block_math = [&](auto masked_tag) {
        constexpr bool MASKED = decltype(masked_tag)::value;
        const uint64_t rpx = ptx::pack_f32x2(rp, rp);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float p0, p1;
          exp_pair(i, p0, p1);
          if (MASKED) {
            if (2 * i >= vis) p0 = 0.0f;
            if (2 * i + 1 >= vis) p1 = 0.0f;
          }
          if (DROP) {
            const uint32_t ws = dw << (i >> 1);
            const float k0 = __uint_as_float(dp[2 * i] & attn_drop_full_mask(ws, 2 * (i & 1)));
            const float k1 = __uint_as_float(dp[2 * i + 1] & attn_drop_full_mask(ws, 2 * (i & 1) + 1));
            pk[i] = ds_pair(p0, p1, ptx::mul_f32x2(ptx::pack_f32x2(k0, k1), rpx));
          } else {
            pk[i] = ds_pair(p0, p1, ptx::pack_f32x2(__uint_as_float(dp[2 * i]), __uint_as_float(dp[2 * i + 1])));
          }
        }
      };
      if (__all_sync(0xffffffffu, vis >= 32)) block_math(std::false_type{});
      else block_math(std::true_type{});
      // dS (bf16, 16 columns) goes over the START of this warp's own S half, never over the half the sibling warp reads
      ptx::tmem_st_32x16(tmem_s, pk);
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      ptx::mbar_arrive(&ds_full[j & 1]);
    }
    if (n_kv > 0) {
      ptx::mbar_wait(acc_full, 0);
      ptx::tc_fence_after();
    }
    if (half < DH / 32) {
      const int c = half;
      uint32_t o[32];
      if (n_kv > 0) {
        ptx::tmem_ld_32x32(tmem_base + lane_addr + 192 + c * 32, o);
        ptx::tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = 0u;
      }
      if (valid) store_row_bf16(p.dq + (int64_t)(q0 + qi) * p.dqs + (int64_t)h * DH + c * 32, o, p.scale);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 8) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------------------ dK / dV
// Shared memory of this kernel: K, V (resident), 2 stages of (Q_i, dO_i), and the bf16 P^T / dS^T tiles the elementwise
// warps hand to the dV / dK MMAs.  P^T and dS^T go through shared memory (not back into TMEM over S^T / dP^T) so that the
// S^T / dP^T columns are free as soon as they have been read into registers: the MMAs of block i + 1 then run underneath
// the exponentials of block i even though TMEM only has room for one S^T / dP^T pair next to the two accumulators.
namespace fkv {
constexpr int STAGES = 2;
constexpr uint32_t OFF_B0 = 2 * fb::BIGTILE, OFF_B1 = OFF_B0 + STAGES * fb::SMALLTILE;
constexpr uint32_t OFF_P = OFF_B1 + STAGES * fb::SMALLTILE, OFF_DS = OFF_P + fb::BIGTILE;
constexpr uint32_t OFF_VEC = OFF_DS + fb::BIGTILE, OFF_BAR = OFF_VEC + 2 * 256 * 4;
constexpr size_t SMEM = OFF_BAR + 256;
static_assert(SMEM == fb::SMEM, "both backward kernels are launched with fb::SMEM bytes of shared memory");
}  // namespace fkv

template <int DH, bool DROP>
__global__ void __launch_bounds__(fb::THREADS, 2)
attn_bwd_dkv_tc(const __grid_constant__ CUtensorMap tma_k, const __grid_constant__ CUtensorMap tma_v,
                  const __grid_constant__ CUtensorMap tma_q, const __grid_constant__ CUtensorMap tma_do, AttnBwdParams p) {
  using namespace fb;
  constexpr int STAGES = fkv::STAGES;
  constexpr uint32_t OFF_B0 = fkv::OFF_B0, OFF_B1 = fkv::OFF_B1, OFF_P = fkv::OFF_P, OFF_DS = fkv::OFF_DS, OFF_VEC = fkv::OFF_VEC,
                     OFF_BAR = fkv::OFF_BAR;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* a_full = bars;                     // K and V tiles
  uint64_t* b_full = bars + 1;                 // STAGES: Q_i and dO_i
  uint64_t* b_empty = b_full + STAGES;
  uint64_t* sp_full = b_empty + STAGES;
  uint64_t* ds_full = sp_full + 1;
  uint64_t* acc_full = ds_full + 1;
  uint64_t* sp_free = acc_full + 1;            // S^T / dP^T of the current block are in registers (256 arrivals)
  uint64_t* pd_free = sp_free + 1;             // the dV / dK MMAs have consumed the P^T / dS^T tiles
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pd_free + 1);
  float* vec = reinterpret_cast<float*>(smem + OFF_VEC);          // [2][128]: L of 64 queries, then delta of 64 queries

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, kb = blockIdx.x;
  const int q0 = p.cu_q[b], lq = p.cu_q[b + 1] - q0, k0 = p.cu_k[b], lk = p.cu_k[b + 1] - k0;
  if (kb * BIG >= lk) return;
  const int shift = lk - lq;
  // causal: key kr is seen by queries q >= kr - shift; the first query block that can see this key tile
  int i0 = 0;
  if (p.causal) { const int first = kb * BIG - shift; i0 = first > 0 ? first / SMALL : 0; }
  const int n_q = (lq + SMALL - 1) / SMALL;
  const int n_it = n_q > i0 ? n_q - i0 : 0;

  if (warp == 9 && lane == 0) {
    ptx::mbar_init(a_full, 1);
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&b_full[s], 1); ptx::mbar_init(&b_empty[s], 1); }
    ptx::mbar_init(sp_full, 1); ptx::mbar_init(ds_full, 256); ptx::mbar_init(acc_full, 1);
    ptx::mbar_init(sp_free, 256); ptx::mbar_init(pd_free, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 8) {
    if (lane == 0) {
      ptx::prefetch_tensormap(&tma_q); ptx::prefetch_tensormap(&tma_do);
      ptx::prefetch_tensormap(&tma_k); ptx::prefetch_tensormap(&tma_v);
    }
    ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // (helper warps: converged loops, one elected lane issues - see ptx::elect_one_sync)
  const int u_nit = __shfl_sync(0xffffffffu, n_it, 0);
  if (warp == 8) {
    if (u_nit > 0) {
      const int hcol = (h * DH) & ~63;
      const int krow = __shfl_sync(0xffffffffu, k0 + kb * BIG, 0);
      int qrow = __shfl_sync(0xffffffffu, q0 + i0 * SMALL, 0);
      if (ptx::elect_one_sync()) {
        ptx::mbar_arrive_expect_tx(a_full, 2 * BIGTILE);
        ptx::tma_load_2d(smem + OFF_A0, &tma_k, a_full, hcol, krow);
        ptx::tma_load_2d(smem + OFF_A1, &tma_v, a_full, hcol, krow);
      }
      __syncwarp();
      int s = 0; uint32_t ph = 0;
      for (int it = 0; it < u_nit; ++it, qrow += SMALL) {
        ptx::mbar_wait(&b_empty[s], ph ^ 1);
        if (ptx::elect_one_sync()) {
          ptx::mbar_arrive_expect_tx(&b_full[s], 2 * SMALLTILE);
          ptx::tma_load_2d(smem + OFF_B0 + s * SMALLTILE, &tma_q, &b_full[s], hcol, qrow);
          ptx::tma_load_2d(smem + OFF_B1 + s * SMALLTILE, &tma_do, &b_full[s], hcol, qrow);
        }
        __syncwarp();
        if (++s == STAGES) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 9) {
    if (u_nit > 0) {
      constexpr uint32_t IDESC_S = ptx::idesc_bf16_f32(BIG, SMALL);
      constexpr uint32_t IDESC_ACC = ptx::idesc_bf16_f32(BIG, DH, false, true);
      const uint32_t tb0 = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint64_t hoff = (uint64_t)(((h * DH) & 63) >> 3);
      const uint64_t kdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_A0), 1024) + hoff;
      const uint64_t vdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_A1), 1024) + hoff;
      const uint64_t pdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_P), 1024);
      const uint64_t dsdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_DS), 1024);
      const uint64_t qdesc0 = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_B0), 1024) + hoff;
      const uint64_t dodesc0 = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_B1), 1024) + hoff;
      auto issue_sp = [&](int it) {                // S^T_it and dP^T_it (waits for Q_it / dO_it to land)
        const int st = it % STAGES;
        ptx::mbar_wait(&b_full[st], (it / STAGES) & 1);
        ptx::tc_fence_after();
        if (ptx::elect_one_sync()) {
          const uint64_t qdesc = qdesc0 + (uint64_t)(st * (SMALLTILE >> 4)), dodesc = dodesc0 + (uint64_t)(st * (SMALLTILE >> 4));
#pragma unroll
          for (int k = 0; k < DH / 16; ++k) ptx::umma_f16_ss(tb0, kdesc + 2 * k, qdesc + 2 * k, IDESC_S, k != 0);       // S^T
#pragma unroll
          for (int k = 0; k < DH / 16; ++k) ptx::umma_f16_ss(tb0 + 64, vdesc + 2 * k, dodesc + 2 * k, IDESC_S, k != 0); // dP^T
          ptx::umma_commit(sp_full);
        }
        __syncwarp();
      };
      ptx::mbar_wait(a_full, 0);
      issue_sp(0);
      for (int it = 0; it < u_nit; ++it) {
        const int st = it % STAGES;
        if (it + 1 < u_nit) {                         // issued BEFORE dV / dK of block it: runs under block it's exponentials
          ptx::mbar_wait(sp_free, it & 1);
          ptx::tc_fence_after();
          issue_sp(it + 1);
        }
        ptx::mbar_wait(ds_full, it & 1);
        ptx::tc_fence_after();
        if (ptx::elect_one_sync()) {
          const uint64_t qdesc = qdesc0 + (uint64_t)(st * (SMALLTILE >> 4)), dodesc = dodesc0 + (uint64_t)(st * (SMALLTILE >> 4));
#pragma unroll
          for (int kk = 0; kk < SMALL / 16; ++kk)      // dV += P^T dO_i   (A = P^T [128 keys x 64 queries] K-major in smem)
            ptx::umma_f16_ss(tb0 + 128, pdesc + 2 * kk, dodesc + (uint64_t)kk * (16 * 128 >> 4), IDESC_ACC, (it | kk) != 0);
#pragma unroll
          for (int kk = 0; kk < SMALL / 16; ++kk)      // dK += dS^T Q_i
            ptx::umma_f16_ss(tb0 + 192, dsdesc + 2 * kk, qdesc + (uint64_t)kk * (16 * 128 >> 4), IDESC_ACC, (it | kk) != 0);
          ptx::umma_commit(&b_empty[st]);
          ptx::umma_commit(pd_free);
          if (it == u_nit - 1) ptx::umma_commit(acc_full);
        }
        __syncwarp();
      }
    }
  } else {
    const int r = (warp & 3) * 32 + lane, half = warp >> 2;       // key row r, query columns half*32..+31 of a block
    const int tid = threadIdx.x;                                   // 0..255 here
    const int kr = kb * BIG + r;
    const bool valid = kr < lk;
    const uint32_t lane_addr = (uint32_t)((warp & 3) * 32) << 16;
    const float sl2 = p.scale_log2;
    // Per query block the 64 L / delta values (and, with dropout, the 64 hash row constants / row bases) are staged in
    // shared memory for broadcast reads.  The global loads for block it + 1 are issued BEFORE block it's math and land in
    // the other half of `vec` after it, so their latency never sits between two blocks.
    auto stage_load = [&](int it) -> uint32_t {
      const int qbase = (i0 + it) * SMALL;
      if (tid < 128) {
        const int q = qbase + (tid & 63);
        const float* src = tid < 64 ? p.lse : p.delta;
        return __float_as_uint(q < lq ? __ldg(src + (int64_t)h * p.total_q + q0 + q) : (tid < 64 ? INFINITY : 0.0f));
      }
      if (DROP) {                                  // per query column: hash row constant (128..191) and low row base (192..255)
        const uint64_t rowkey = (uint64_t)h * p.total_q + (uint64_t)(q0 + qbase + (tid & 63));
        return tid < 192 ? attn_drop_rowconst(p.drop.seed, rowkey) : attn_drop_rowbase(rowkey);
      }
      return 0u;
    };
    if (n_it > 0) {
      reinterpret_cast<uint32_t*>(vec)[tid] = stage_load(0);
      asm volatile("bar.sync 1, 256;" ::: "memory");
    }
    for (int it = 0; it < n_it; ++it) {
      const int qbase = (i0 + it) * SMALL;
      float* vb = vec + (it & 1) * 256;
      const uint32_t staged_next = it + 1 < n_it ? stage_load(it + 1) : 0u;
      // first query column this key row may attend to (causal), all columns masked for rows past the sequence end
      const int cmin = !valid ? SMALL : (p.causal ? kr - shift - qbase : 0);
      ptx::mbar_wait(sp_full, it & 1);
      ptx::tc_fence_after();
      uint32_t s[32], dp[32], pk[16], dk[16];
      ptx::tmem_ld_32x32(tmem_base + lane_addr + half * 32, s);
      ptx::tmem_ld_32x32(tmem_base + lane_addr + 64 + half * 32, dp);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      ptx::mbar_arrive(sp_free);                   // the MMA warp may overwrite S^T / dP^T with block it + 1
      const uint64_t sl2x = ptx::pack_f32x2(sl2, sl2);
      const float4* vl = reinterpret_cast<const float4*>(vb + half * 32);
      const float4* vd = reinterpret_cast<const float4*>(vb + 64 + half * 32);
      // Columns [0, cm) of this thread's 32 are masked.  Only the diagonal blocks of a causal problem and the last key tile
      // of a sequence mask anything: every other warp takes the variant without the 64 compare / select instructions
      // (the elementwise warps are issue bound, their time follows the instruction count).
      const int cm = cmin - half * 32;
      // dropout: the keep decisions of (this thread's 32 query columns, this key) as one word - lane l hashes the key-block
      // word of query column perm(l), a 32 x 32 bit transpose across the warp turns rows into columns (common.cuh)
      const uint32_t dw = DROP ? attn_drop_word_transposed(reinterpret_cast<const uint32_t*>(vb + 128 + half * 32),
                                                           reinterpret_cast<const uint32_t*>(vb + 192 + half * 32),
                                                           (uint32_t)kr >> 5, p.drop.thr, lane) : 0u;
      auto block_math = [&](auto masked_tag) {
        constexpr bool MASKED = decltype(masked_tag)::value;
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          const float4 l4 = vl[i4], d4 = vd[i4];
          // (packed fp32 pairs, see the dQ kernel)
          float x0, x1, x2, x3;
          ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(s[4 * i4]), __uint_as_float(s[4 * i4 + 1])), sl2x,
                                           ptx::pack_f32x2(-l4.x, -l4.y)), x0, x1);
          ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(s[4 * i4 + 2]), __uint_as_float(s[4 * i4 + 3])), sl2x,
                                           ptx::pack_f32x2(-l4.z, -l4.w)), x2, x3);
          float p0 = fast_exp2(x0), p1 = fast_exp2(x1), p2 = fast_exp2(x2), p3 = fast_exp2(x3);
          if (MASKED) {
            if (4 * i4 < cm) p0 = 0.0f;
            if (4 * i4 + 1 < cm) p1 = 0.0f;
            if (4 * i4 + 2 < cm) p2 = 0.0f;
            if (4 * i4 + 3 < cm) p3 = 0.0f;
          }
          const uint64_t pa = ptx::pack_f32x2(p0, p1), pb = ptx::pack_f32x2(p2, p3);
          const uint64_t dpa = ptx::pack_f32x2(__uint_as_float(dp[4 * i4]), __uint_as_float(dp[4 * i4 + 1]));
          const uint64_t dpb = ptx::pack_f32x2(__uint_as_float(dp[4 * i4 + 2]), __uint_as_float(dp[4 * i4 + 3]));
          const uint64_t nda = ptx::pack_f32x2(-d4.x, -d4.y), ndb = ptx::pack_f32x2(-d4.z, -d4.w);
          float e0, e1, e2, e3, f0, f1, f2, f3;
          if (DROP) {                                                 // M / (1 - p) of (query column, this key)
            const uint32_t ws = dw << i4, rpb = __float_as_uint(p.drop.rp);
            const float m0 = __uint_as_float(rpb & attn_drop_full_mask(ws, 0)), m1 = __uint_as_float(rpb & attn_drop_full_mask(ws, 1));
            const float m2 = __uint_as_float(rpb & attn_drop_full_mask(ws, 2)), m3 = __uint_as_float(rpb & attn_drop_full_mask(ws, 3));
            const uint64_t ma = ptx::pack_f32x2(m0, m1), mb = ptx::pack_f32x2(m2, m3);
            ptx::unpack_f32x2(ptx::mul_f32x2(pa, ma), e0, e1);
            ptx::unpack_f32x2(ptx::mul_f32x2(pb, mb), e2, e3);
            ptx::unpack_f32x2(ptx::mul_f32x2(pa, ptx::fma_f32x2(ma, dpa, nda)), f0, f1);
            ptx::unpack_f32x2(ptx::mul_f32x2(pb, ptx::fma_f32x2(mb, dpb, ndb)), f2, f3);
          } else {                                                    // P^T as it is, dS^T = P^T (dP^T - delta)
            e0 = p0; e1 = p1; e2 = p2; e3 = p3;
            ptx::unpack_f32x2(ptx::mul_f32x2(pa, ptx::add_f32x2(dpa, nda)), f0, f1);
            ptx::unpack_f32x2(ptx::mul_f32x2(pb, ptx::add_f32x2(dpb, ndb)), f2, f3);
          }
          const __nv_bfloat162 hp0 = __floats2bfloat162_rn(e0, e1), hp1 = __floats2bfloat162_rn(e2, e3);
          const __nv_bfloat162 hd0 = __floats2bfloat162_rn(f0, f1), hd1 = __floats2bfloat162_rn(f2, f3);
          pk[2 * i4] = *reinterpret_cast<const uint32_t*>(&hp0); pk[2 * i4 + 1] = *reinterpret_cast<const uint32_t*>(&hp1);
          dk[2 * i4] = *reinterpret_cast<const uint32_t*>(&hd0); dk[2 * i4 + 1] = *reinterpret_cast<const uint32_t*>(&hd1);
        }
      };
      if (__any_sync(0xffffffffu, cm > 0)) block_math(std::true_type{});
      else block_math(std::false_type{});
      // P^T / dS^T rows (this thread: key row r, 32 query columns = four 16-byte pieces) into the 128B-swizzled K-major
      // tiles the dV / dK MMAs read as their A operand; block it - 1's MMAs must have finished with the tiles
      if (it > 0) ptx::mbar_wait(pd_free, (it - 1) & 1);
      {
        uint8_t* prow = smem + OFF_P + r * 128;
        uint8_t* drow = smem + OFF_DS + r * 128;
        const uint32_t sw = (uint32_t)(r & 7);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const uint32_t off = (((uint32_t)(half * 4 + c)) ^ sw) << 4;
          *reinterpret_cast<uint4*>(prow + off) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
          *reinterpret_cast<uint4*>(drow + off) = make_uint4(dk[4 * c], dk[4 * c + 1], dk[4 * c + 2], dk[4 * c + 3]);
        }
      }
      ptx::fence_proxy_async_smem();
      ptx::mbar_arrive(ds_full);
      // the other half of `vec` was last read in block it - 1, which every thread left before this block's barrier
      reinterpret_cast<uint32_t*>(vec + ((it + 1) & 1) * 256)[tid] = staged_next;
      asm volatile("bar.sync 1, 256;" ::: "memory");
    }
    if (n_it > 0) {
      ptx::mbar_wait(acc_full, 0);
      ptx::tc_fence_after();
    }
    // half 0 stores dV (TMEM columns 128..), half 1 stores dK (192..)
#pragma unroll
    for (int cc = 0; cc < DH / 32; ++cc) {
      const bool is_k = half == 1;
      uint32_t o[32];
      if (n_it > 0) {
        ptx::tmem_ld_32x32(tmem_base + lane_addr + (is_k ? 192 : 128) + cc * 32, o);
        ptx::tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = 0u;
      }
      if (valid) {
        if (!is_k) store_row_bf16(p.dv + (int64_t)(k0 + kr) * p.dvs + (int64_t)h * DH + cc * 32, o, 1.0f);
        else store_row_bf16(p.dk + (int64_t)(k0 + kr) * p.dks + (int64_t)h * DH + cc * 32, o, p.scale);
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 8) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------------------ grouped variant (head_dim 64)
// Round 2.  The two kernels above are bound by none of their pipes: their elementwise warps sit in hand-off latencies
// (profiles/r1i_attention_backward.md) and two CTAs per SM give the SM only two independent chains of 64 columns.  What
// limits the work in flight is TMEM (fp32 S and dP cost 128 columns per 64-wide block) and the shared-memory operand
// traffic of small MMAs.  This variant:
//   * keeps the CTA's 128 resident rows (K and V for dK/dV, Q and dO for dQ) in TMEM as the A operands (TS form, 32 columns
//     each, written once by the elementwise threads): an S MMA then reads only its 1 KB B slice from shared memory, which
//     is what makes N = 32 MMAs affordable (A from shared memory would be 5 KB per 16 clk of math);
//   * streams the other operand in 32-row sub-blocks; sub-block s belongs to group s % G = four warps (thread = resident
//     row, 32 columns) with its OWN TMEM columns: S, dP (fp32) and the bf16 operands P^T / dS^T it hands back;
//   * splits the MMA issue over two warps: one issues S / dP of a group's NEXT sub-block as soon as the group has the
//     current one in registers (so the tensor-pipe round trip runs under the group's own exponentials), the other issues
//     the accumulating MMAs (dV / dK or dQ) when the bf16 operands are written; every true dependency between the two is
//     an mbarrier, none relies on issue order across threads;
//   * touches shared memory only for the streamed tiles (4 KB each, read twice) and the 32 L / delta of a sub-block.
// TMEM (one CTA per SM): resident rows 0..63; group g at 64 + 96 g: S, dP (32 + 32), P^T, dS^T (16 + 16); accumulators after
// the groups.  MODE 0 = dK/dV: 3 groups, 480 columns;  MODE 1 = dQ: 4 groups, 512 columns.
// Per 128 x 64 block-equivalent: tensor pipe 512 clk (dK/dV) / 384 clk (dQ), MUFU 512 clk, shared memory 256 clk.
namespace fg {
constexpr int BIG = 128, SUB = 32, D = 64;
constexpr int groups(int mode) { return mode == 0 ? 3 : 4; }
constexpr int stages(int mode) { return 4 * groups(mode); }      // a multiple of 2 G: see the helper warps
constexpr int issuers(int mode) { return mode == 0 ? 1 : 2; }             // S / dP issuer warps
constexpr int threads(int mode) { return (groups(mode) * 4 + 2 + issuers(mode)) * 32; }   // + TMA producer, S / dP issuers, accumulate issuer
constexpr uint32_t SUBTILE = SUB * D * 2;           // 4 KB
constexpr int MAX_STAGES = 16;
constexpr uint32_t OFF_B0 = 0, OFF_B1 = MAX_STAGES * SUBTILE, OFF_VEC = 2 * MAX_STAGES * SUBTILE;
constexpr uint32_t OFF_BAR = OFF_VEC + 4 * 2 * 128 * 4;   // per group 2 x (32 L, 32 delta, 32 hash row constants, 32 row bases)
constexpr size_t SMEM = OFF_BAR + 512;
constexpr uint32_t TMEM_COLS = 512, COL_R0 = 0, COL_R1 = 32, COL_G = 64, GSTRIDE = 96;
constexpr uint32_t col_acc0(int mode) { return COL_G + groups(mode) * GSTRIDE; }
}  // namespace fg

struct G4Rows { const __nv_bfloat16* r0; const __nv_bfloat16* r1; int64_t s0, s1; };   // the resident operands and their row strides

template <int MODE, bool DROP>
__global__ void __launch_bounds__(fg::threads(MODE), 1)
attn_bwd_g4(const __grid_constant__ CUtensorMap tma_b0, const __grid_constant__ CUtensorMap tma_b1, AttnBwdParams p, G4Rows rr) {
  using namespace fg;
  constexpr int G = groups(MODE), EW = G * 4, STAGES = stages(MODE), NI = issuers(MODE);
  static_assert(STAGES <= MAX_STAGES && STAGES % (2 * G) == 0, "stage ring");
  constexpr uint32_t COL_ACC0 = col_acc0(MODE), COL_ACC1 = COL_ACC0 + 64;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* b_full = bars;                     // STAGES: streamed sub-block s (both tiles) has landed
  uint64_t* b_empty = b_full + MAX_STAGES;     // STAGES: the last MMA reading it has completed
  uint64_t* r_ready = b_empty + MAX_STAGES;    // the resident rows are in TMEM (8 warps)
  uint64_t* sp_full = r_ready + 1;             // 4: S and dP of the group's sub-block are in TMEM
  uint64_t* sp_free = sp_full + 4;             // 4: ... and in the group's registers (4 warps)
  uint64_t* ds_full = sp_free + 4;             // 4: the group's bf16 operands are in TMEM (4 warps)
  uint64_t* pd_free = ds_full + 4;             // 4: the accumulating MMAs have consumed them
  uint64_t* acc_full = pd_free + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);
  float* vec = reinterpret_cast<float*>(smem + OFF_VEC);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, tb = blockIdx.x;
  const int q0 = p.cu_q[b], lq = p.cu_q[b + 1] - q0, k0 = p.cu_k[b], lk = p.cu_k[b + 1] - k0;
  const int rows_len = MODE == 0 ? lk : lq;
  if (tb * BIG >= rows_len) return;
  const int shift = lk - lq;
  int s0 = 0, n_sub;                           // first streamed sub-block and how many this tile needs
  if (MODE == 0) {                             // causal: key kr is seen by queries q >= kr - shift
    if (p.causal) { const int first = tb * BIG - shift; s0 = first > 0 ? first / SUB : 0; }
    const int n_all = (lq + SUB - 1) / SUB;
    n_sub = n_all > s0 ? n_all - s0 : 0;
  } else {
    int kv_end = lk;
    if (p.causal) { const int last = tb * BIG + BIG + shift; kv_end = last < lk ? (last < 0 ? 0 : last) : lk; }
    n_sub = (kv_end + SUB - 1) / SUB;
  }
  const int res_base = MODE == 0 ? k0 : q0, str_base = MODE == 0 ? q0 : k0;

  if (warp == EW + NI + 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&b_full[s], 1); ptx::mbar_init(&b_empty[s], 1); }
    ptx::mbar_init(r_ready, 8);
    for (int g = 0; g < 4; ++g) {
      ptx::mbar_init(&sp_full[g], 1); ptx::mbar_init(&sp_free[g], 4); ptx::mbar_init(&ds_full[g], 4); ptx::mbar_init(&pd_free[g], 1);
    }
    ptx::mbar_init(acc_full, 1);
    ptx::fence_barrier_init();
  }
  if (warp == EW) {
    if (lane == 0) { ptx::prefetch_tensormap(&tma_b0); ptx::prefetch_tensormap(&tma_b1); }
    ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= EW) {
    // The three helper warps walk their loops CONVERGED (every lane polls the barriers) and one elected lane issues; all
    // operands are provably warp-uniform (shuffles of lane 0), see ptx::elect_one_sync.  The loops are unrolled over one
    // round of the stage ring (STAGES = a multiple of 2 G): stage, group and every group-barrier parity are then compile-time
    // constants and the descriptors constant offsets from two uniform registers - with run-time stage / group arithmetic one
    // trip of the S / dP loop cost ~500 clk of (dependent, single-warp) issue and was the bound of the whole kernel.
    const int nsub = __shfl_sync(0xffffffffu, n_sub, 0);
    const uint32_t tb0 = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint64_t d0_base = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_B0), 1024);
    const uint64_t d1_base = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_B1), 1024);
    if (warp == EW) {
      // ---------------------------------------------------------------- TMA producer
      int row = __shfl_sync(0xffffffffu, str_base + s0 * SUB, 0);
      uint32_t ph = 0;
      for (int sb = 0; sb < nsub; sb += STAGES, ph ^= 1) {
#pragma unroll
        for (int u = 0; u < STAGES; ++u) {
          if (sb + u >= nsub) break;
          ptx::mbar_wait(&b_empty[u], ph ^ 1);
          if (ptx::elect_one_sync()) {
            ptx::mbar_arrive_expect_tx(&b_full[u], 2 * SUBTILE);
            ptx::tma_load_2d(smem + OFF_B0 + u * SUBTILE, &tma_b0, &b_full[u], h * D, row);
            ptx::tma_load_2d(smem + OFF_B1 + u * SUBTILE, &tma_b1, &b_full[u], h * D, row);
          }
          __syncwarp();
          row += SUB;
        }
      }
    } else if (warp <= EW + NI) {
      // ---------------------------------------------------------------- S / dP issuers: NI warps, warp i serves the groups
      // g % NI == i (one trip costs ~400 clk of dependent single-warp issue, more than a sub-block's share of any pipe).
      // A group's sub-blocks must stay with ONE issuer: its sp_free parity wait is only safe in program order.
      if (nsub > 0) {
        constexpr uint32_t IDESC_S = ptx::idesc_bf16_f32(BIG, SUB);
        const int mine = warp - (EW + 1);
        ptx::mbar_wait(r_ready, 0);
        ptx::tc_fence_after();
        uint32_t ph = 0;
        for (int sb = 0; sb < nsub; sb += STAGES, ph ^= 1) {
#pragma unroll
          for (int u = 0; u < STAGES; ++u) {
            if ((u % G) % NI != mine) continue;
            if (sb + u >= nsub) break;
            const int g = u % G;                           // (compile-time after unrolling; STAGES / G is even, so is the parity)
            if (sb + u >= G) ptx::mbar_wait(&sp_free[g], (uint32_t)((u / G - 1) & 1));   // the group holds its previous S / dP in registers
            ptx::mbar_wait(&b_full[u], ph);
            ptx::tc_fence_after();
            if (ptx::elect_one_sync()) {
              const uint64_t d0 = d0_base + (uint64_t)(u * (SUBTILE >> 4)), d1 = d1_base + (uint64_t)(u * (SUBTILE >> 4));
              const uint32_t cs = tb0 + COL_G + g * GSTRIDE;
#pragma unroll
              for (int k = 0; k < D / 16; ++k) {           // S and dP k-steps alternate
                ptx::umma_f16_ts(cs, tb0 + COL_R0 + k * 8, d0 + 2 * k, IDESC_S, k != 0);
                ptx::umma_f16_ts(cs + 32, tb0 + COL_R1 + k * 8, d1 + 2 * k, IDESC_S, k != 0);
              }
              ptx::umma_commit(&sp_full[g]);
            }
            __syncwarp();
          }
        }
      }
    } else {
      // ---------------------------------------------------------------- accumulate issuer
      if (nsub > 0) {
        constexpr uint32_t IDESC_ACC = ptx::idesc_bf16_f32(BIG, D, false, true);        // B MN-major
        uint32_t ph = 0;
        for (int sb = 0; sb < nsub; sb += STAGES, ph ^= 1) {
#pragma unroll
          for (int u = 0; u < STAGES; ++u) {
            if (sb + u >= nsub) break;
            const int g = u % G;
            ptx::mbar_wait(&ds_full[g], (uint32_t)((u / G) & 1));
            ptx::mbar_wait(&b_full[u], ph);                // (completed long ago: orders this thread after the TMA writes)
            ptx::tc_fence_after();
            if (ptx::elect_one_sync()) {
              const uint64_t d0 = d0_base + (uint64_t)(u * (SUBTILE >> 4)), d1 = d1_base + (uint64_t)(u * (SUBTILE >> 4));
              const uint32_t cp = tb0 + COL_G + g * GSTRIDE + 64;    // P^T (16 columns), dS^T / dS at + 16
              const uint32_t acc = (sb + u) != 0;
              if (MODE == 0) {
#pragma unroll
                for (int kk = 0; kk < SUB / 16; ++kk) {
                  // dV += P^T dO_s   (16 queries = 8 packed columns per step)
                  ptx::umma_f16_ts(tb0 + COL_ACC0, cp + kk * 8, d1 + (uint64_t)kk * (16 * 128 >> 4), IDESC_ACC, kk ? 1u : acc);
                  // dK += dS^T Q_s
                  ptx::umma_f16_ts(tb0 + COL_ACC1, cp + 16 + kk * 8, d0 + (uint64_t)kk * (16 * 128 >> 4), IDESC_ACC, kk ? 1u : acc);
                }
              } else {
#pragma unroll
                for (int kk = 0; kk < SUB / 16; ++kk)      // dQ += dS K_s
                  ptx::umma_f16_ts(tb0 + COL_ACC0, cp + 16 + kk * 8, d0 + (uint64_t)kk * (16 * 128 >> 4), IDESC_ACC, kk ? 1u : acc);
              }
              ptx::umma_commit(&b_empty[u]);
              ptx::umma_commit(&pd_free[g]);
              if (sb + u == nsub - 1) ptx::umma_commit(acc_full);
            }
            __syncwarp();
          }
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ elementwise warps (thread = resident row)
    const int qtr = warp & 3, g = warp >> 2, r = qtr * 32 + lane;
    const int row = tb * BIG + r;
    const bool valid = row < rows_len;
    const uint32_t lane_addr = (uint32_t)(qtr * 32) << 16;
    const float sl2 = p.scale_log2;
    if (n_sub > 0 && g < 2) {                    // groups 0 / 1 bring the two resident operands into TMEM (bf16 pairs per column)
      const __nv_bfloat16* base = g == 0 ? rr.r0 : rr.r1;
      const int64_t stride = g == 0 ? rr.s0 : rr.s1;
      const uint4* src = reinterpret_cast<const uint4*>(base + (int64_t)(res_base + (valid ? row : 0)) * stride + (int64_t)h * D);
      uint32_t rv[32];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const uint4 v4 = valid ? __ldg(src + c) : make_uint4(0u, 0u, 0u, 0u);
        rv[4 * c] = v4.x; rv[4 * c + 1] = v4.y; rv[4 * c + 2] = v4.z; rv[4 * c + 3] = v4.w;
      }
      ptx::tmem_st_32x32(tmem_base + lane_addr + (g == 0 ? COL_R0 : COL_R1), rv);
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(r_ready);
    }
    const uint32_t tmem_s = tmem_base + lane_addr + COL_G + g * GSTRIDE;      // S, dP at + 32, P^T at + 64, dS^T / dS at + 80
    // MODE 1: per-row constants
    float L = INFINITY, dl = 0.0f;
    int my_end = 0;
    uint64_t rowkey = 0; uint32_t rowconst = 0, rowbase = 0;
    if (MODE == 1) {
      if (valid) {
        L = p.lse[(int64_t)h * p.total_q + q0 + row];
        dl = p.delta[(int64_t)h * p.total_q + q0 + row];
        my_end = p.causal ? max(0, min(lk, row + shift + 1)) : lk;
      }
      rowkey = (uint64_t)h * p.total_q + (uint64_t)(q0 + (valid ? row : 0));
      if (DROP) { rowconst = attn_drop_rowconst(p.drop.seed, rowkey); rowbase = attn_drop_rowbase(rowkey); }
    }
    // MODE 0: the sub-block's 32 L / delta (and hash constants) are staged per group, one value per thread, one sub-block ahead
    float* vg = vec + g * 256;
    auto stage_load = [&](int s) -> uint32_t {
      const int q = (s0 + s) * SUB + (r & 31), which = r >> 5;
      if (which == 0) return __float_as_uint(q < lq ? __ldg(p.lse + (int64_t)h * p.total_q + q0 + q) : INFINITY);
      if (which == 1) return __float_as_uint(q < lq ? __ldg(p.delta + (int64_t)h * p.total_q + q0 + q) : 0.0f);
      if (DROP) {
        const uint64_t rk = (uint64_t)h * p.total_q + (uint64_t)(q0 + q);
        return which == 2 ? attn_drop_rowconst(p.drop.seed, rk) : attn_drop_rowbase(rk);
      }
      return 0u;
    };
    auto group_sync = [&]() { asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory"); };
    if (MODE == 0 && g < n_sub) {
      reinterpret_cast<uint32_t*>(vg)[r] = stage_load(g);
      group_sync();
    }
    int n = 0;
    for (int s = g; s < n_sub; s += G, ++n) {
      const int cbase = (s0 + s) * SUB;            // first streamed row of this sub-block (query for MODE 0, key for MODE 1)
      const float* vb = vg + (n & 1) * 128;
      uint32_t staged_next = 0u;
      if (MODE == 0 && s + G < n_sub) staged_next = stage_load(s + G);
      ptx::mbar_wait(&sp_full[g], n & 1);
      ptx::tc_fence_after();
      uint32_t sv[32], dp[32], pk[16], dk[16];
      ptx::tmem_ld_32x32(tmem_s, sv);
      ptx::tmem_ld_32x32(tmem_s + 32, dp);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&sp_free[g]);   // S / dP of this group's next sub-block may be issued
      const uint64_t sl2x = ptx::pack_f32x2(sl2, sl2);
      if (MODE == 0) {
        // columns [0, cm) are masked: queries before the causal diagonal of this key row; everything for rows past the end
        const int cm = !valid ? SUB : (p.causal ? row - shift - cbase : 0);
        const float4* vl = reinterpret_cast<const float4*>(vb);
        const float4* vd = reinterpret_cast<const float4*>(vb + 32);
        // dropout: (32 query columns, this key) decisions as one word (attn_drop_word_transposed, common.cuh)
        const uint32_t dw = DROP ? attn_drop_word_transposed(reinterpret_cast<const uint32_t*>(vb + 64), reinterpret_cast<const uint32_t*>(vb + 96),
                                                             (uint32_t)row >> 5, p.drop.thr, lane) : 0u;
        auto block_math = [&](auto masked_tag) {
          constexpr bool MASKED = decltype(masked_tag)::value;
#pragma unroll
          for (int i4 = 0; i4 < 8; ++i4) {
            const float4 l4 = vl[i4], d4 = vd[i4];
            float x0, x1, x2, x3;
            ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(sv[4 * i4]), __uint_as_float(sv[4 * i4 + 1])), sl2x,
                                             ptx::pack_f32x2(-l4.x, -l4.y)), x0, x1);
            ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(sv[4 * i4 + 2]), __uint_as_float(sv[4 * i4 + 3])), sl2x,
                                             ptx::pack_f32x2(-l4.z, -l4.w)), x2, x3);
            float p0 = fast_exp2(x0), p1 = fast_exp2(x1), p2 = fast_exp2(x2), p3 = fast_exp2(x3);
            if (MASKED) {
              if (4 * i4 < cm) p0 = 0.0f;
              if (4 * i4 + 1 < cm) p1 = 0.0f;
              if (4 * i4 + 2 < cm) p2 = 0.0f;
              if (4 * i4 + 3 < cm) p3 = 0.0f;
            }
            const uint64_t pa = ptx::pack_f32x2(p0, p1), pb = ptx::pack_f32x2(p2, p3);
            const uint64_t dpa = ptx::pack_f32x2(__uint_as_float(dp[4 * i4]), __uint_as_float(dp[4 * i4 + 1]));
            const uint64_t dpb = ptx::pack_f32x2(__uint_as_float(dp[4 * i4 + 2]), __uint_as_float(dp[4 * i4 + 3]));
            const uint64_t nda = ptx::pack_f32x2(-d4.x, -d4.y), ndb = ptx::pack_f32x2(-d4.z, -d4.w);
            float e0, e1, e2, e3, f0, f1, f2, f3;
            if (DROP) {                                                 // M / (1 - p) of (query column, this key)
              const uint32_t ws = dw << i4, rpb = __float_as_uint(p.drop.rp);
              const float m0 = __uint_as_float(rpb & attn_drop_full_mask(ws, 0)), m1 = __uint_as_float(rpb & attn_drop_full_mask(ws, 1));
              const float m2 = __uint_as_float(rpb & attn_drop_full_mask(ws, 2)), m3 = __uint_as_float(rpb & attn_drop_full_mask(ws, 3));
              const uint64_t ma = ptx::pack_f32x2(m0, m1), mb = ptx::pack_f32x2(m2, m3);
              ptx::unpack_f32x2(ptx::mul_f32x2(pa, ma), e0, e1);
              ptx::unpack_f32x2(ptx::mul_f32x2(pb, mb), e2, e3);
              ptx::unpack_f32x2(ptx::mul_f32x2(pa, ptx::fma_f32x2(ma, dpa, nda)), f0, f1);
              ptx::unpack_f32x2(ptx::mul_f32x2(pb, ptx::fma_f32x2(mb, dpb, ndb)), f2, f3);
            } else {
              e0 = p0; e1 = p1; e2 = p2; e3 = p3;
              ptx::unpack_f32x2(ptx::mul_f32x2(pa, ptx::add_f32x2(dpa, nda)), f0, f1);
              ptx::unpack_f32x2(ptx::mul_f32x2(pb, ptx::add_f32x2(dpb, ndb)), f2, f3);
            }
            const __nv_bfloat162 hp0 = __floats2bfloat162_rn(e0, e1), hp1 = __floats2bfloat162_rn(e2, e3);
            const __nv_bfloat162 hd0 = __floats2bfloat162_rn(f0, f1), hd1 = __floats2bfloat162_rn(f2, f3);
            pk[2 * i4] = *reinterpret_cast<const uint32_t*>(&hp0); pk[2 * i4 + 1] = *reinterpret_cast<const uint32_t*>(&hp1);
            dk[2 * i4] = *reinterpret_cast<const uint32_t*>(&hd0); dk[2 * i4 + 1] = *reinterpret_cast<const uint32_t*>(&hd1);
          }
        };
        if (__any_sync(0xffffffffu, cm > 0)) block_math(std::true_type{});
        else block_math(std::false_type{});
      } else {
        // columns [vis, 32) are masked: keys past this query row's causal bound / the sequence end
        const int vis = my_end - cbase;
        const uint64_t nLx = ptx::pack_f32x2(-L, -L), ndlx = ptx::pack_f32x2(-dl, -dl), rpx = ptx::pack_f32x2(p.drop.rp, p.drop.rp);
        const uint32_t dw = DROP ? attn_drop_word(rowconst, rowbase, (uint32_t)(cbase >> 5), p.drop.thr) : 0u;
        auto block_math = [&](auto masked_tag) {
          constexpr bool MASKED = decltype(masked_tag)::value;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x0, x1;
            ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(sv[2 * i]), __uint_as_float(sv[2 * i + 1])), sl2x, nLx), x0, x1);
            float p0 = fast_exp2(x0), p1 = fast_exp2(x1);
            if (MASKED) {
              if (2 * i >= vis) p0 = 0.0f;
              if (2 * i + 1 >= vis) p1 = 0.0f;
            }
            uint64_t t2;
            if (DROP) {
              const uint32_t ws = dw << (i >> 1);
              const float k0_ = __uint_as_float(dp[2 * i] & attn_drop_full_mask(ws, 2 * (i & 1)));
              const float k1_ = __uint_as_float(dp[2 * i + 1] & attn_drop_full_mask(ws, 2 * (i & 1) + 1));
              t2 = ptx::mul_f32x2(ptx::pack_f32x2(k0_, k1_), rpx);
            } else {
              t2 = ptx::pack_f32x2(__uint_as_float(dp[2 * i]), __uint_as_float(dp[2 * i + 1]));
            }
            float d0, d1;
            ptx::unpack_f32x2(ptx::mul_f32x2(ptx::pack_f32x2(p0, p1), ptx::add_f32x2(t2, ndlx)), d0, d1);
            const __nv_bfloat162 hh = __floats2bfloat162_rn(d0, d1);
            dk[i] = *reinterpret_cast<const uint32_t*>(&hh);
          }
        };
        if (__all_sync(0xffffffffu, vis >= SUB)) block_math(std::false_type{});
        else block_math(std::true_type{});
      }
      // the accumulating MMAs of this group's previous sub-block have read the bf16 operand columns
      if (n > 0) { ptx::mbar_wait(&pd_free[g], (n - 1) & 1); ptx::tc_fence_after(); }
      if (MODE == 0) ptx::tmem_st_32x16(tmem_s + 64, pk);   // P^T
      ptx::tmem_st_32x16(tmem_s + 80, dk);                  // dS^T (MODE 0) / dS (MODE 1)
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&ds_full[g]);
      if (MODE == 0) {
        // the other half of the group's `vec` was last read in sub-block n - 1, which every thread of the group has left
        reinterpret_cast<uint32_t*>(vg + ((n + 1) & 1) * 128)[r] = staged_next;
        group_sync();
      }
    }
    if (n_sub > 0) {
      ptx::mbar_wait(acc_full, 0);
      ptx::tc_fence_after();
    }
    // 32-column output pieces: MODE 0: dV 0..31, dV 32..63, dK 0..31, dK 32..63 (piece i to group i % 3); MODE 1: dQ 0..31, 32..63
    constexpr int PIECES = MODE == 0 ? 4 : 2;
    for (int piece = g; piece < PIECES; piece += G) {
      const bool second = MODE == 0 && piece >= 2;
      uint32_t o[32];
      if (n_sub > 0) {
        ptx::tmem_ld_32x32(tmem_base + lane_addr + (second ? COL_ACC1 : COL_ACC0) + (piece & 1) * 32, o);
        ptx::tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = 0u;
      }
      if (valid) {
        if (MODE == 1) store_row_bf16(p.dq + (int64_t)(q0 + row) * p.dqs + (int64_t)h * D + (piece & 1) * 32, o, p.scale);
        else if (!second) store_row_bf16(p.dv + (int64_t)(k0 + row) * p.dvs + (int64_t)h * D + (piece & 1) * 32, o, 1.0f);
        else store_row_bf16(p.dk + (int64_t)(k0 + row) * p.dks + (int64_t)h * D + (piece & 1) * 32, o, p.scale);
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == EW) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

// returns ACAI_ERR_UNSUPPORTED when the problem does not qualify (the caller then runs the SIMT kernels)
int attn_bwd_tc(const void* q, const void* k, const void* v, int dtype, int64_t qs, int64_t ks, int64_t vs, const void* dout,
                int64_t dos, const float* lse, const float* delta, const int32_t* cu_q, const int32_t* cu_k, int batch,
                int max_q, int max_k, int64_t total_q, int64_t total_k, int heads, int head_dim, int causal, float scale,
                void* dq, void* dk, void* dv, int grad_dtype, int64_t dqs, int64_t dks, int64_t dvs, AttnDropout drop,
                cudaStream_t st) {
  using namespace fb;
  auto misaligned = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) != 0; };
  if (dtype != ACAI_BF16 || grad_dtype != ACAI_BF16 || (head_dim != 64 && head_dim != 32) || qs % 8 || ks % 8 || vs % 8 || dos % 8 || dqs % 8 ||
      dks % 8 || dvs % 8 || total_q <= 0 || total_k <= 0 || misaligned(q) || misaligned(k) || misaligned(v) ||
      misaligned(dout) || misaligned(dq) || misaligned(dk) || misaligned(dv))
    return ACAI_ERR_UNSUPPORTED;
  const uint64_t cols = (uint64_t)heads * head_dim;
  CUtensorMap q_big, do_big, k_small, v_small, k_big, v_big, q_small, do_small;
  int rc = 0;
  if ((rc = make_tma_2d_bf16(&q_big, q, (uint64_t)total_q, cols, (uint64_t)qs, BIG, D))) return rc;
  if ((rc = make_tma_2d_bf16(&do_big, dout, (uint64_t)total_q, cols, (uint64_t)dos, BIG, D))) return rc;
  if ((rc = make_tma_2d_bf16(&k_small, k, (uint64_t)total_k, cols, (uint64_t)ks, SMALL, D))) return rc;
  if ((rc = make_tma_2d_bf16(&v_small, v, (uint64_t)total_k, cols, (uint64_t)vs, SMALL, D))) return rc;
  if ((rc = make_tma_2d_bf16(&k_big, k, (uint64_t)total_k, cols, (uint64_t)ks, BIG, D))) return rc;
  if ((rc = make_tma_2d_bf16(&v_big, v, (uint64_t)total_k, cols, (uint64_t)vs, BIG, D))) return rc;
  if ((rc = make_tma_2d_bf16(&q_small, q, (uint64_t)total_q, cols, (uint64_t)qs, SMALL, D))) return rc;
  if ((rc = make_tma_2d_bf16(&do_small, dout, (uint64_t)total_q, cols, (uint64_t)dos, SMALL, D))) return rc;
  static bool attr_set = false;
  if (!attr_set) {
#define ACAI_SET(K) ACAI_CUDA(cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM))
    ACAI_SET((attn_bwd_dq_tc<64, false>)); ACAI_SET((attn_bwd_dkv_tc<64, false>));
    ACAI_SET((attn_bwd_dq_tc<32, false>)); ACAI_SET((attn_bwd_dkv_tc<32, false>));
    ACAI_SET((attn_bwd_dq_tc<64, true>)); ACAI_SET((attn_bwd_dkv_tc<64, true>));
    ACAI_SET((attn_bwd_dq_tc<32, true>)); ACAI_SET((attn_bwd_dkv_tc<32, true>));
#undef ACAI_SET
    attr_set = true;
  }
  AttnBwdParams p{cu_q, cu_k, lse, delta, (int)total_q, causal, scale, scale * 1.4426950408889634f,
                  (__nv_bfloat16*)dq, (__nv_bfloat16*)dk, (__nv_bfloat16*)dv, dqs, dks, dvs, drop};
  const dim3 gq((max_q + BIG - 1) / BIG, heads, batch), gk((max_k + BIG - 1) / BIG, heads, batch);
  const bool dr = drop.thr != 0;
  // head_dim 64: the grouped kernels (attn_bwd_g4); ACAI_ATTN_BWD_G4 = 0 / 1 / 2 / 3 (bit 0: dK/dV, bit 1: dQ) for A/B runs.
  // Default 1: the grouped dK/dV kernel is 1.2 - 1.3x the two-CTA kernel on sequences of ~1000 tokens and level at ~650 causal;
  // the grouped dQ kernel is not faster than attn_bwd_dq_tc (profiles/r2d_issue_path.md).
  // Unset: dK/dV grouped for non-causal problems (encoder self-attention 1 211 -> 1 080 us, decoder cross-attention 1 036 -> 987 us;
  // causal decoder self-attention 621 -> 644 us: its key tiles see too few query sub-blocks to amortise a one-CTA-per-SM prologue).
  // (read on every call - the tests switch it inside one process)
  const int g4 = [] { const char* e = getenv("ACAI_ATTN_BWD_G4"); const int v = e ? atoi(e) : -1; return v < 0 || v > 3 ? -1 : v; }();
  const int use_g4 = head_dim != 64 ? 0 : (g4 >= 0 ? g4 : (causal ? 0 : 1));
  CUtensorMap q_sub, do_sub, k_sub, v_sub;
  if (use_g4 & 1) {
    if ((rc = make_tma_2d_bf16(&q_sub, q, (uint64_t)total_q, cols, (uint64_t)qs, fg::SUB, D))) return rc;
    if ((rc = make_tma_2d_bf16(&do_sub, dout, (uint64_t)total_q, cols, (uint64_t)dos, fg::SUB, D))) return rc;
  }
  if (use_g4 & 2) {
    if ((rc = make_tma_2d_bf16(&k_sub, k, (uint64_t)total_k, cols, (uint64_t)ks, fg::SUB, D))) return rc;
    if ((rc = make_tma_2d_bf16(&v_sub, v, (uint64_t)total_k, cols, (uint64_t)vs, fg::SUB, D))) return rc;
  }
  static bool g4_attr_set = false;
  if (use_g4 && !g4_attr_set) {
#define ACAI_SET(K) ACAI_CUDA(cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fg::SMEM))
    ACAI_SET((attn_bwd_g4<0, false>)); ACAI_SET((attn_bwd_g4<0, true>)); ACAI_SET((attn_bwd_g4<1, false>)); ACAI_SET((attn_bwd_g4<1, true>));
#undef ACAI_SET
    g4_attr_set = true;
  }
  const G4Rows rows_q{(const __nv_bfloat16*)q, (const __nv_bfloat16*)dout, qs, dos}, rows_k{(const __nv_bfloat16*)k, (const __nv_bfloat16*)v, ks, vs};
#define ACAI_LAUNCH(K, G, ...) do { if (head_dim == 64) { if (dr) K<64, true><<<G, THREADS, SMEM, st>>>(__VA_ARGS__); else K<64, false><<<G, THREADS, SMEM, st>>>(__VA_ARGS__); } \
                                    else { if (dr) K<32, true><<<G, THREADS, SMEM, st>>>(__VA_ARGS__); else K<32, false><<<G, THREADS, SMEM, st>>>(__VA_ARGS__); } } while (0)
  if (use_g4 & 2) {
    if (dr) attn_bwd_g4<1, true><<<gq, fg::threads(1), fg::SMEM, st>>>(k_sub, v_sub, p, rows_q);
    else attn_bwd_g4<1, false><<<gq, fg::threads(1), fg::SMEM, st>>>(k_sub, v_sub, p, rows_q);
  } else {
    ACAI_LAUNCH(attn_bwd_dq_tc, gq, q_big, do_big, k_small, v_small, p);
  }
  rc = launch_status("attn_bwd_dq");
  if (rc) return rc;
  if (use_g4 & 1) {
    if (dr) attn_bwd_g4<0, true><<<gk, fg::threads(0), fg::SMEM, st>>>(q_sub, do_sub, p, rows_k);
    else attn_bwd_g4<0, false><<<gk, fg::threads(0), fg::SMEM, st>>>(q_sub, do_sub, p, rows_k);
  } else {
    ACAI_LAUNCH(attn_bwd_dkv_tc, gk, k_big, v_big, q_small, do_small, p);
  }
#undef ACAI_LAUNCH
  return launch_status("attn_bwd_dkv");
}

}  // namespace acai
