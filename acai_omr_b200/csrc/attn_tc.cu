// K3 (tensor-core path): packed variable-length flash attention for head_dim 64, bf16 operands.
//
// One CTA = one (sequence, head, 128-query tile); two CTAs are resident per SM (256 of the SM's 512 TMEM columns
// each), so one CTA's prologue / output store overlaps the other's steady state and the MUFU pipe - the real bound
// here: 64 exp2 per row per 64-key block against 256 MMA cycles - always has a softmax warp with work.
// Per 64-key block j:
//     S_j = Q K_j^T            tcgen05.mma 128x64x16 x4 (SS), fp32 accumulator in TMEM, DOUBLE-buffered (2 x 64 columns):
//                              S_{j+1} is already finished when the softmax of block j ends, no MMA round trip in between
//     P_j = exp2(S_j - m)      4 softmax warps, thread = query row: one tcgen05.ld pass (64 values in registers),
//                              bf16 pairs written back OVER S_j with tcgen05.st (32 columns) - P never touches smem
//     O  += P_j V_j            tcgen05.mma 128x64x16 x4 with A = P from TMEM (TS), B = V MN-major from smem;
//                              O (64 columns) stays in TMEM for the whole sequence
// Online-softmax rescaling of O is lazy: m is only advanced (and O multiplied by 2^(m_old-m_new) through a
// tcgen05.ld/st round trip) when some row of the warp saw its maximum grow by more than 2^8; otherwise the stale
// maximum keeps being used, which is exact after the final division by l.  The rescale needs O to hold P_{j-1} V_{j-1}:
// the issuer commits every PV to an `o_step` barrier that the softmax warps consult only on that (rare) path.
// Q/K/V tiles arrive by TMA straight out of the fused [T, 3*H*64] QKV buffer (or separate Q and KV buffers for cross
// attention); ragged sequence ends are masked in the softmax, never padded.
//   warps 0-3: softmax + output     warp 4: TMA producer (+TMEM alloc)     warp 5: MMA issuer
#include <cstdlib>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace acai {

int make_tma_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                     uint32_t box_cols);

namespace fa {
constexpr int BQ = 128, BKV = 64, D = 64;
constexpr uint32_t QTILE = BQ * D * 2;            // 16 KB
constexpr uint32_t KTILE = BKV * D * 2;           // 8 KB
constexpr float RESCALE_LOG2 = 8.0f;
// NBUF = number of S buffers in TMEM, CTAS = resident CTAs per SM.
//   NBUF 2, CTAS 2: 256 TMEM columns (S[0] 0..63, S[1] 64..127, O 128..191), 4 K/V stages.
//   NBUF 1, CTAS 3 / 4: 128 columns (S 0..63, O 64..127), 3 / 2 K/V stages.  A CTA's tensor-pipe round trip
//   (P_j -> PV_j, S_{j+1}) is then exposed to its own softmax warps, but every scheduler holds three / four softmax warps
//   of independent CTAs instead of two, which is what keeps its MUFU unit fed.  These variants launch 256 threads (setmaxnreg
//   is a WARPGROUP-wide instruction, so the helper warps 4-7 form a full warpgroup; 6 and 7 idle): 64 / 80 registers per
//   thread at launch, the helper warpgroup hands its share back (setmaxnreg 32 / 40) and the softmax warpgroup takes 96 / 120.
//   NBUF 2, CTAS 4, SB 32: the softmax works on HALF key blocks (SB = 32 keys: S is 128 x 32, two buffers at columns 0..31 and
//   32..63, O 64..127), so S is double-buffered again inside 128 columns: the tensor-pipe round trip leaves the softmax
//   warps' critical path and a row's 32 values fit the register budget in one pass.
//   QT 1: Q lives in TMEM as the A operand of the S MMAs (TS form, like P for the PV MMAs) instead of shared memory: an S MMA
//   then reads only its K slice from shared memory (the SS form re-reads 4 KB of Q per K = 16 step, which makes it
//   operand-bandwidth bound).  The softmax warps load their own Q row from global memory once and tcgen05.st it.
//     QT 1, NBUF 1, CTAS 4, SB 32: 128 columns = Q 0..31, S 32..63, O 64..127 (single S buffer);
//     QT 1, NBUF 2, CTAS 3, SB 32: 128 columns = S 0..31 / 32..63, O 64..127, plus a SECOND 32-column allocation for Q (shared
//     memory is sized so that a fourth CTA can never be resident: 4 x 128 columns would leave no room for anybody's Q).
template <int NBUF, int CTAS_, int SB_, int QT_ = 0> struct Cfg {
  static constexpr int CTAS = CTAS_, SB = SB_, QT = QT_;
  static constexpr int THREADS = CTAS == 2 ? 192 : 256;
  static constexpr int KV_STAGES = QT ? (CTAS == 3 ? 4 : 3) : (CTAS == 2 ? 4 : (CTAS == 3 ? 3 : 2));
  static constexpr uint32_t OFF_Q = 0, OFF_K = QT ? 0 : QTILE, OFF_V = OFF_K + KV_STAGES * KTILE, OFF_BAR = OFF_V + KV_STAGES * KTILE;
  static constexpr size_t SMEM = OFF_BAR + 256;
  static constexpr bool Q_SECOND_ALLOC = QT && NBUF * SB + 64 + 32 > 128;
  static constexpr uint32_t Q_COL = 0;                                       // (inside its own allocation when Q_SECOND_ALLOC)
  static constexpr uint32_t S_COL = (QT && !Q_SECOND_ALLOC) ? 32 : 0;
  static constexpr uint32_t TMEM_COLS = S_COL + NBUF * SB + 64 > 128 ? 256 : 128;
  static constexpr uint32_t O_COL = S_COL + NBUF * SB;
  static_assert(CTAS * (TMEM_COLS + (Q_SECOND_ALLOC ? 32 : 0)) <= 512 && (SB == 32 || SB == 64), "TMEM budget");
  static_assert(!Q_SECOND_ALLOC || (CTAS + 1) * (SMEM + 1024) > 233472, "a surplus CTA could take TMEM columns and deadlock the Q allocations");
};
}  // namespace fa
using namespace fa;

struct AttnParams {
  const int32_t* cu_q; const int32_t* cu_k;
  void* out; int out_dt; int64_t o_stride;
  int causal; float scale_log2;
  float* lse; int total_q;      // optional [heads, total_q] log2-domain log-sum-exp of every query row (for the backward)
  const __nv_bfloat16* q; int64_t q_stride;   // raw Q rows (the QT variants load Q themselves)
  AttnDropout drop;             // attention-probability dropout (training): P is masked before P V, l sums the unmasked P
};

// DH = head_dim (64 or 32).  Shared-memory tiles are always 64 columns (one 128-byte swizzle row) wide: for DH = 32 a
// tile holds a PAIR of heads and this CTA's head is addressed by a 64-byte start offset in the MMA descriptors.
// POLY: how many of every 4 exponentials are evaluated with exp2_fma on the FMA pipe instead of MUFU.EX2 (0, 1 or 2).
// The MUFU unit (16 results / clk / SM) is this kernel's busiest pipe; the FMA pipe is mostly idle.
// SPEC: the exponentials of block j use the reference maximum as block j - 1 left it and do NOT wait for block j's own row
// maximum (which runs on the ALU pipe beside the MUFU work): any reference gives the same softmax after the division by l, the
// maximum only has to be advanced before the values overflow.  The check moves behind the P hand-off: a row whose maximum grew by
// more than 2^8 has O / l rescaled once PV_j has landed, before block j + 1; a jump of more than 2^60 inside ONE block (which
// would overflow bf16 P) recomputes the block the old way.  Takes the max -> vote -> branch chain out of every block's critical path.
template <int DH, bool DROP, int POLY, int NBUF, int CTAS, int SB, int QT, int PIPE = 0, int SPEC = 0>
__global__ void __launch_bounds__((Cfg<NBUF, CTAS, SB, QT>::THREADS), CTAS)
attn_fwd_tc(const __grid_constant__ CUtensorMap tma_q, const __grid_constant__ CUtensorMap tma_k,
              const __grid_constant__ CUtensorMap tma_v, AttnParams p) {
  using C = Cfg<NBUF, CTAS, SB, QT>;
  constexpr int R = BKV / SB;                        // softmax blocks per K/V tile
  constexpr int KV_STAGES = C::KV_STAGES;
  constexpr uint32_t OFF_Q = C::OFF_Q, OFF_K = C::OFF_K, OFF_V = C::OFF_V, OFF_BAR = C::OFF_BAR, TMEM_COLS = C::TMEM_COLS,
                     O_COL = C::O_COL, S_COL = C::S_COL;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* q_full = bars;                          // 1
  uint64_t* kv_full = bars + 1;                     // KV_STAGES
  uint64_t* kv_empty = kv_full + KV_STAGES;         // KV_STAGES
  uint64_t* s_full = kv_empty + KV_STAGES;          // 2 (per S buffer)
  uint64_t* p_full = s_full + 2;                    // 2
  uint64_t* o_step = p_full + 2;                    // 1: one phase per finished PV (consulted on the rescale path)
  uint64_t* o_full = o_step + 1;                    // 1: the last PV has landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, qb = blockIdx.x;
  const int q0 = p.cu_q[b], lq = p.cu_q[b + 1] - q0, k0 = p.cu_k[b], lk = p.cu_k[b + 1] - k0;
  if (qb * BQ >= lq) return;                       // uniform for the whole CTA
  const int shift = lk - lq;
  int kv_end = lk;
  if (p.causal) { const int last = qb * BQ + BQ + shift; kv_end = last < lk ? (last < 0 ? 0 : last) : lk; }
  const int n_kv = (kv_end + BKV - 1) / BKV;       // K/V tiles (TMA granularity)
  const int n_sb = (kv_end + SB - 1) / SB;         // softmax blocks (S / P granularity)

  if (warp == 5 && lane == 0) {
    ptx::mbar_init(q_full, QT ? 128 : 1);
    for (int s = 0; s < KV_STAGES; ++s) { ptx::mbar_init(&kv_full[s], 1); ptx::mbar_init(&kv_empty[s], 1); }
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(&s_full[i], 1); ptx::mbar_init(&p_full[i], 128); }
    ptx::mbar_init(o_step, 1); ptx::mbar_init(o_full, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 4) {
    if (lane == 0) { ptx::prefetch_tensormap(&tma_q); ptx::prefetch_tensormap(&tma_k); ptx::prefetch_tensormap(&tma_v); }
    if constexpr (C::Q_SECOND_ALLOC) ptx::tmem_alloc_keep_permit(tmem_slot + 1, 32);
    ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_q = (C::Q_SECOND_ALLOC ? tmem_slot[1] : tmem_base) + C::Q_COL;

  // (CTAS 4: the register budget moves from the two helper warps to the softmax warps inside each role's own branch -
  // ptxas sizes a branch by the setmaxnreg that dominates it)
  constexpr int REG_LO = CTAS == 4 ? 32 : 40, REG_HI = CTAS == 4 ? 96 : 120;
  if (warp >= 4) {
  if constexpr (CTAS > 2) ptx::setmaxnreg_dec<REG_LO>();   // one instruction for the whole helper warpgroup
  // Both helper warps walk their loops CONVERGED (every lane polls the barriers) and one elected lane issues: with the
  // operands provably warp-uniform (shuffles of lane 0) the TMA / tcgen05 instructions take them from uniform registers
  // directly; under `if (lane == 0)` every one of them is a waterfall loop of ~100 clk (ptx::elect_one_sync).
  const int u_nkv = __shfl_sync(0xffffffffu, n_kv, 0), u_nsb = __shfl_sync(0xffffffffu, n_sb, 0);
  if (warp == 4) {
    // ------------------------------------------------------------------ TMA producer
    if (u_nkv > 0) {
      const int hcol = (h * DH) & ~63;
      const int qrow = __shfl_sync(0xffffffffu, q0 + qb * BQ, 0);
      int krow = __shfl_sync(0xffffffffu, k0, 0);
      if constexpr (!QT) {
        if (ptx::elect_one_sync()) {
          ptx::mbar_arrive_expect_tx(q_full, QTILE);
          ptx::tma_load_2d(smem + OFF_Q, &tma_q, q_full, hcol, qrow);
        }
        __syncwarp();
      }
      int s = 0; uint32_t ph = 0;
      for (int j = 0; j < u_nkv; ++j, krow += BKV) {
        ptx::mbar_wait(&kv_empty[s], ph ^ 1);
        if (ptx::elect_one_sync()) {
          ptx::mbar_arrive_expect_tx(&kv_full[s], 2 * KTILE);
          ptx::tma_load_2d(smem + OFF_K + s * KTILE, &tma_k, &kv_full[s], hcol, krow);
          ptx::tma_load_2d(smem + OFF_V + s * KTILE, &tma_v, &kv_full[s], hcol, krow);
        }
        __syncwarp();
        if (++s == KV_STAGES) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 5) {
    // ------------------------------------------------------------------ MMA issuer
    if (u_nkv > 0) {
      constexpr uint32_t IDESC_S = ptx::idesc_bf16_f32(BQ, SB);
      constexpr uint32_t IDESC_O = ptx::idesc_bf16_f32(BQ, DH, false, true);  // B = V is MN-major
      const uint32_t tb0 = __shfl_sync(0xffffffffu, tmem_base, 0), tq0 = __shfl_sync(0xffffffffu, tmem_q, 0);
      const uint64_t hoff = (uint64_t)(((h * DH) & 63) >> 3);                 // this head's 16-byte units inside the 128 B row
      const uint64_t qdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_Q), 1024) + hoff;
      const uint64_t kdesc0 = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_K), 1024) + hoff;
      const uint64_t vdesc0 = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_V), 1024) + hoff;
      auto issue_s = [&](int i) {                    // S_i into buffer i % NBUF; its keys are rows (i % R) * SB.. of tile i / R
        const int t = i / R, st = t % KV_STAGES;
        ptx::mbar_wait(&kv_full[st], (t / KV_STAGES) & 1);
        ptx::tc_fence_after();
        if (ptx::elect_one_sync()) {
          const uint64_t kdesc = kdesc0 + (uint64_t)(st * (KTILE >> 4)) + (uint64_t)((i % R) * (SB * 128 >> 4));
#pragma unroll
          for (int k = 0; k < DH / 16; ++k) {
            if constexpr (QT) ptx::umma_f16_ts(tb0 + S_COL + (i % NBUF) * SB, tq0 + k * 8, kdesc + 2 * k, IDESC_S, k != 0);
            else ptx::umma_f16_ss(tb0 + S_COL + (i % NBUF) * SB, qdesc + 2 * k, kdesc + 2 * k, IDESC_S, k != 0);
          }
          ptx::umma_commit(&s_full[i % NBUF]);
        }
        __syncwarp();
      };
      ptx::mbar_wait(q_full, 0);
      issue_s(0);
      if (NBUF == 2 && u_nsb > 1) issue_s(1);
      for (int i = 0; i < u_nsb; ++i) {
        const int st = (i / R) % KV_STAGES;
        ptx::mbar_wait(&p_full[i % NBUF], (i / NBUF) & 1);
        ptx::tc_fence_after();
        if (ptx::elect_one_sync()) {
          const uint64_t vdesc = vdesc0 + (uint64_t)(st * (KTILE >> 4));
#pragma unroll
          for (int kk = 0; kk < SB / 16; ++kk)        // A: 16 keys = 8 packed TMEM columns; B: 16 V rows = 2048 bytes
            ptx::umma_f16_ts(tb0 + O_COL, tb0 + S_COL + (i % NBUF) * SB + kk * 8,
                             vdesc + (uint64_t)((i % R) * (SB / 16) + kk) * (16 * 128 >> 4), IDESC_O, (i | kk) != 0);
          if (i % R == R - 1 || i == u_nsb - 1) ptx::umma_commit(&kv_empty[st]);   // last reader of this K/V tile
          ptx::umma_commit(o_step);
          if (i == u_nsb - 1) ptx::umma_commit(o_full);
        }
        __syncwarp();
        if (i + NBUF < u_nsb) issue_s(i + NBUF);      // overwrites P_i: tcgen05 ops execute in issue order
      }
    }
  }
  } else {
    // ------------------------------------------------------------------ softmax warps (thread = query row)
    if constexpr (CTAS > 2) ptx::setmaxnreg_inc<REG_HI>();
    const int r = warp * 32 + lane;
    const int qi = qb * BQ + r;
    const bool valid = qi < lq;
    const int my_end = p.causal ? max(0, min(lk, qi + shift + 1)) : lk;     // exclusive key bound of this row
    const int warp_min_end = __shfl_sync(0xffffffffu, my_end, 0);          // rows ascend inside a warp
    const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
    const uint32_t tmem_o = tmem_base + lane_addr + O_COL;
    const float scale_log2 = p.scale_log2;
    float m = -INFINITY, l = 0.0f;
    const uint64_t rowkey = (uint64_t)h * p.total_q + (uint64_t)(q0 + (valid ? qi : 0));
    const uint32_t rowconst = DROP ? attn_drop_rowconst(p.drop.seed, rowkey) : 0u;
    const uint32_t rowbase = DROP ? attn_drop_rowbase(rowkey) : 0u;
    if constexpr (QT) {                             // this thread's Q row -> TMEM (bf16 pairs per column, the TS A-operand layout)
      if (n_kv > 0) {
        uint32_t qr[DH / 2];
        const uint4* src = reinterpret_cast<const uint4*>(p.q + (int64_t)(q0 + (valid ? qi : 0)) * p.q_stride + (int64_t)h * DH);
#pragma unroll
        for (int c = 0; c < DH / 8; ++c) {
          const uint4 v4 = valid ? __ldg(src + c) : make_uint4(0u, 0u, 0u, 0u);
          qr[4 * c] = v4.x; qr[4 * c + 1] = v4.y; qr[4 * c + 2] = v4.z; qr[4 * c + 3] = v4.w;
        }
        if constexpr (DH == 64) ptx::tmem_st_32x32(tmem_q + lane_addr, qr);
        else ptx::tmem_st_32x16(tmem_q + lane_addr, qr);
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        ptx::mbar_arrive(q_full);
      }
    }

    if constexpr (PIPE) {
      // Software-pipelined variant (NBUF 2, SB 32): S_{j+1} is requested from TMEM (the other S buffer is already complete)
      // BEFORE the exponentials of block j, so that the tcgen05.ld round trip and the wait for S never sit between two
      // blocks' MUFU phases.  Costs 32 more live registers per thread.
      static_assert(!PIPE || (NBUF == 2 && SB == 32 && QT == 0), "pipelined softmax: two 32-key S buffers");
      uint32_t sa[32], sb2[32];
      auto request = [&](int j, uint32_t (&sv)[32]) {          // wait for S_j, start its tcgen05.ld (no wait)
        ptx::mbar_wait(&s_full[j % NBUF], (j / NBUF) & 1);
        ptx::tc_fence_after();
        ptx::tmem_ld_32x32(tmem_base + lane_addr + S_COL + (j % NBUF) * SB, sv);
      };
      auto block = [&](int j, uint32_t (&sv)[32], uint32_t (&sn)[32]) {
        const uint32_t tmem_s = tmem_base + lane_addr + S_COL + (j % NBUF) * SB;
        const int kbase = j * SB;
        if (kbase + SB > warp_min_end) {
#pragma unroll
          for (int i = 0; i < 32; ++i) if (kbase + i >= my_end) sv[i] = 0xff800000u;   // -inf
        }
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          mx0 = fmaxf(mx0, __uint_as_float(sv[i])); mx1 = fmaxf(mx1, __uint_as_float(sv[i + 1]));
          mx2 = fmaxf(mx2, __uint_as_float(sv[i + 2])); mx3 = fmaxf(mx3, __uint_as_float(sv[i + 3]));
        }
        const float m_new = fmaxf(m, fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)));
        const bool grow = (m_new - m) * scale_log2 > RESCALE_LOG2 || (m == -INFINITY && m_new > -INFINITY);
        if (__any_sync(0xffffffffu, grow)) {
          const float m_use = m_new == -INFINITY ? 0.0f : m_new;
          const float corr = m == -INFINITY ? 0.0f : fast_exp2((m - m_use) * scale_log2);
          l *= corr;
          if (j > 0) {
            ptx::mbar_wait(o_step, (j - 1) & 1);      // O holds P_{j-1} V_{j-1}
            ptx::tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < DH / 16; ++c) {
              uint32_t o[16];
              ptx::tmem_ld_32x16(tmem_o + c * 16, o);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * corr);
              ptx::tmem_st_32x16(tmem_o + c * 16, o);
            }
          }
          m = m_new;
        }
        const float moff = (m == -INFINITY ? 0.0f : m) * scale_log2;
        const uint64_t sc2 = ptx::pack_f32x2(scale_log2, scale_log2), mo2 = ptx::pack_f32x2(-moff, -moff);
        uint64_t l2 = ptx::pack_f32x2(0.0f, 0.0f);
        uint32_t pk[16];
        const uint32_t dw = DROP ? attn_drop_word(rowconst, rowbase, (uint32_t)(kbase >> 5), p.drop.thr) : 0u;   // 32 keep decisions
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float x0, x1;
          ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(sv[2 * i]), __uint_as_float(sv[2 * i + 1])), sc2, mo2), x0, x1);
          const float p0 = fast_exp2(x0);
          const float p1 = (POLY == 2 || (POLY == 1 && (i & 1))) ? exp2_fma(x1) : fast_exp2(x1);
          l2 = ptx::add_f32x2(l2, ptx::pack_f32x2(p0, p1));
          const __nv_bfloat162 hh = __floats2bfloat162_rn(p0, p1);
          pk[i] = *reinterpret_cast<const uint32_t*>(&hh);
          if (DROP) pk[i] &= attn_drop_pair_mask(dw << (i >> 1), i & 1);
        }
        ptx::tmem_st_32x16(tmem_s, pk);
        if (j + 1 < n_sb) request(j + 1, sn);            // S_{j+1}'s round trip runs under the store / fence / arrive of P_j
        float l0, l1;
        ptx::unpack_f32x2(l2, l0, l1);
        l += l0 + l1;
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        ptx::mbar_arrive(&p_full[j % NBUF]);
        if (j + 1 < n_sb) ptx::tmem_ld_wait();
      };
      if (n_sb > 0) { request(0, sa); ptx::tmem_ld_wait(); }
      for (int j = 0; j < n_sb; j += 2) {
        block(j, sa, sb2);
        if (j + 1 < n_sb) block(j + 1, sb2, sa);
      }
    } else
    for (int j = 0; j < n_sb; ++j) {
      const uint32_t tmem_s = tmem_base + lane_addr + S_COL + (j % NBUF) * SB;
      ptx::mbar_wait(&s_full[j % NBUF], (j / NBUF) & 1);
      ptx::tc_fence_after();
      // (SB 64 with four CTAs: S is read from TMEM twice - maximum, then exponentials - so that <= 32 values are live)
      constexpr bool TWOPASS = CTAS == 4 && SB == 64;
      constexpr int NS = TWOPASS ? 32 : SB;          // values held in registers at a time
      const int kbase = j * SB;
      const bool edge = kbase + SB > warp_min_end;   // ragged end / causal diagonal: mask keys >= my_end
      uint32_t s[NS];
      if constexpr (SPEC && !TWOPASS && SB == 32) {
        if (j > 0 && !__any_sync(0xffffffffu, m == -INFINITY)) {
          ptx::tmem_ld_32x32(tmem_s, *reinterpret_cast<uint32_t(*)[32]>(&s[0]));
          ptx::tmem_ld_wait();
          if (edge) {
#pragma unroll
            for (int i = 0; i < 32; ++i) if (kbase + i >= my_end) s[i] = 0xff800000u;   // -inf
          }
          const float moff = m * scale_log2;
          const uint64_t sc2 = ptx::pack_f32x2(scale_log2, scale_log2), mo2 = ptx::pack_f32x2(-moff, -moff);
          uint64_t l2 = ptx::pack_f32x2(0.0f, 0.0f);
          const uint32_t dw0 = DROP ? attn_drop_word(rowconst, rowbase, (uint32_t)(kbase >> 5), p.drop.thr) : 0u;
          uint32_t pk[16];
          float a0 = -INFINITY, a1 = -INFINITY, a2 = -INFINITY, a3 = -INFINITY;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x0, x1;
            ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), sc2, mo2), x0, x1);
            const float p0 = fast_exp2(x0);
            const float p1 = (POLY == 2 || (POLY == 1 && (i & 1))) ? exp2_fma(x1) : fast_exp2(x1);
            l2 = ptx::add_f32x2(l2, ptx::pack_f32x2(p0, p1));
            const __nv_bfloat162 hh = __floats2bfloat162_rn(p0, p1);
            pk[i] = *reinterpret_cast<const uint32_t*>(&hh);
            if (DROP) pk[i] &= attn_drop_pair_mask(dw0 << (i >> 1), i & 1);
            if (i & 1) { a2 = fmaxf(a2, __uint_as_float(s[2 * i])); a3 = fmaxf(a3, __uint_as_float(s[2 * i + 1])); }
            else { a0 = fmaxf(a0, __uint_as_float(s[2 * i])); a1 = fmaxf(a1, __uint_as_float(s[2 * i + 1])); }
          }
          const float m_new = fmaxf(m, fmaxf(fmaxf(a0, a1), fmaxf(a2, a3)));
          const float growth = (m_new - m) * scale_log2;
          if (!__any_sync(0xffffffffu, growth > 60.0f)) {     // (else: fall through to the block computed under its own maximum)
            ptx::tmem_st_32x16(tmem_s, pk);
            float l0, l1;
            ptx::unpack_f32x2(l2, l0, l1);
            l += l0 + l1;
            ptx::tmem_st_wait();
            ptx::tc_fence_before();
            ptx::mbar_arrive(&p_full[j % NBUF]);
            if (__any_sync(0xffffffffu, growth > RESCALE_LOG2)) {   // advance the reference for the blocks to come
              const float corr = fast_exp2((m - m_new) * scale_log2);
              l *= corr;
              // O holds P_j V_j: o_step completes one phase per PV and is only looked at here, so a parity wait is only safe one
              // phase ahead of what is KNOWN to have completed - s_full(j) implies PV_(j-2); step through PV_(j-1), then PV_j
              ptx::mbar_wait(o_step, (j - 1) & 1);
              ptx::mbar_wait(o_step, j & 1);
              ptx::tc_fence_after();
#pragma unroll 1
              for (int c = 0; c < DH / 16; ++c) {
                uint32_t o[16];
                ptx::tmem_ld_32x16(tmem_o + c * 16, o);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * corr);
                ptx::tmem_st_32x16(tmem_o + c * 16, o);
              }
              ptx::tmem_st_wait();
              ptx::tc_fence_before();
              m = m_new;
            }
            continue;
          }
        }
      }
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int hf = 0; hf < SB / NS; ++hf) {        // (one trip unless TWOPASS)
#pragma unroll
        for (int c = 0; c < NS / 32; ++c)
          ptx::tmem_ld_32x32(tmem_s + hf * NS + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&s[c * 32]));
        ptx::tmem_ld_wait();
        if (edge) {
#pragma unroll
          for (int i = 0; i < NS; ++i) if (kbase + hf * NS + i >= my_end) s[i] = 0xff800000u;   // -inf
        }
#pragma unroll
        for (int i = 0; i < NS; i += 4) {
          mx0 = fmaxf(mx0, __uint_as_float(s[i])); mx1 = fmaxf(mx1, __uint_as_float(s[i + 1]));
          mx2 = fmaxf(mx2, __uint_as_float(s[i + 2])); mx3 = fmaxf(mx3, __uint_as_float(s[i + 3]));
        }
      }
      const float m_new = fmaxf(m, fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)));
      const bool grow = (m_new - m) * scale_log2 > RESCALE_LOG2 || (m == -INFINITY && m_new > -INFINITY);
      if (__any_sync(0xffffffffu, grow)) {          // advance the reference maximum (warp-uniform: TMEM ops are collective)
        const float m_use = m_new == -INFINITY ? 0.0f : m_new;
        const float corr = m == -INFINITY ? 0.0f : fast_exp2((m - m_use) * scale_log2);
        l *= corr;
        if (j > 0) {
          ptx::mbar_wait(o_step, (j - 1) & 1);      // O holds P_{j-1} V_{j-1}
          ptx::tc_fence_after();
          if constexpr (CTAS == 2) {
#pragma unroll
            for (int c = 0; c < DH / 32; ++c) {
              uint32_t o[32];
              ptx::tmem_ld_32x32(tmem_o + c * 32, o);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * corr);
              ptx::tmem_st_32x32(tmem_o + c * 32, o);
            }
          } else {                                    // tighter register budget: 16 columns at a time (rare path)
#pragma unroll 1
            for (int c = 0; c < DH / 16; ++c) {
              uint32_t o[16];
              ptx::tmem_ld_32x16(tmem_o + c * 16, o);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * corr);
              ptx::tmem_st_32x16(tmem_o + c * 16, o);
            }
          }
        }
        m = m_new;
      }
      const float moff = (m == -INFINITY ? 0.0f : m) * scale_log2;
      // x = s * scale - m * scale and the two running sums are packed fp32 pairs (FFMA2 / FADD2: the same fma.rn / add.rn
      // per element in half the issue slots - this loop issues ~0.9 instructions per MUFU cycle, see profiles/r1h_attention.md)
      const uint64_t sc2 = ptx::pack_f32x2(scale_log2, scale_log2), mo2 = ptx::pack_f32x2(-moff, -moff);
      uint64_t l2 = ptx::pack_f32x2(0.0f, 0.0f);
      uint32_t dw[SB / 32];                           // keep decisions of this block's keys, 32 per word
#pragma unroll
      for (int c = 0; c < SB / 32; ++c) dw[c] = DROP ? attn_drop_word(rowconst, rowbase, (uint32_t)(kbase >> 5) + c, p.drop.thr) : 0u;
      auto exp_pair = [&](uint32_t s0, uint32_t s1, int i) -> uint32_t {   // keys kbase + 2i, kbase + 2i + 1
        float x0, x1;
        ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(s0), __uint_as_float(s1)), sc2, mo2), x0, x1);
        const float p0 = fast_exp2(x0);
        const float p1 = (POLY == 2 || (POLY == 1 && (i & 1))) ? exp2_fma(x1) : fast_exp2(x1);
        l2 = ptx::add_f32x2(l2, ptx::pack_f32x2(p0, p1));
        const __nv_bfloat162 hh = __floats2bfloat162_rn(p0, p1);
        uint32_t pkv = *reinterpret_cast<const uint32_t*>(&hh);
        if (DROP) pkv &= attn_drop_pair_mask(dw[i >> 4] << ((i & 15) >> 1), i & 1);   // (see attn_drop_word: one prmt per pair)
        return pkv;
      };
#pragma unroll
      for (int hf = 0; hf < SB / NS; ++hf) {        // P columns written by trip hf only cover S columns already in registers
        if constexpr (TWOPASS) {
          ptx::tmem_ld_32x32(tmem_s + hf * 32, *reinterpret_cast<uint32_t(*)[32]>(&s[0]));
          ptx::tmem_ld_wait();
          if (edge) {
#pragma unroll
            for (int i = 0; i < 32; ++i) if (kbase + hf * 32 + i >= my_end) s[i] = 0xff800000u;
          }
        }
        uint32_t pk[NS / 2];
#pragma unroll
        for (int i = 0; i < NS / 2; ++i) pk[i] = exp_pair(s[2 * i], s[2 * i + 1], hf * (NS / 2) + i);
        if constexpr (NS == 64) ptx::tmem_st_32x32(tmem_s, *reinterpret_cast<const uint32_t(*)[32]>(&pk[0]));
        else ptx::tmem_st_32x16(tmem_s + hf * 16, *reinterpret_cast<const uint32_t(*)[16]>(&pk[0]));
      }
      float l0, l1;
      ptx::unpack_f32x2(l2, l0, l1);
      l += l0 + l1;
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      ptx::mbar_arrive(&p_full[j % NBUF]);
    }
    if (n_kv > 0) {
      ptx::mbar_wait(o_full, 0);
      ptx::tc_fence_after();
    }
    const bool any_key = l > 0.0f && m > -INFINITY;   // (exp2_fma never returns an exact 0: a row without keys has m = -inf)
    const float inv = any_key ? (DROP ? p.drop.rp : 1.0f) / l : 0.0f;
    if (p.lse && valid) p.lse[(int64_t)h * p.total_q + q0 + qi] = any_key ? m * scale_log2 + log2f(l) : INFINITY;
    const int64_t off = (int64_t)(q0 + qi) * p.o_stride + (int64_t)h * DH;
#pragma unroll
    for (int c = 0; c < DH / 32; ++c) {
      uint32_t o[32];
      if (n_kv > 0) {
        ptx::tmem_ld_32x32(tmem_o + c * 32, o);
        ptx::tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = 0u;
      }
      if (valid) {
        if (p.out_dt == ACAI_BF16) {
          uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + off + c * 32);
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            uint4 v4; __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&v4);
#pragma unroll
            for (int i = 0; i < 4; ++i)
              hh[i] = __floats2bfloat162_rn(__uint_as_float(o[q4 * 8 + 2 * i]) * inv, __uint_as_float(o[q4 * 8 + 2 * i + 1]) * inv);
            op[q4] = v4;
          }
        } else {
          float4* op = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + off + c * 32);
#pragma unroll
          for (int q4 = 0; q4 < 8; ++q4)
            op[q4] = make_float4(__uint_as_float(o[4 * q4]) * inv, __uint_as_float(o[4 * q4 + 1]) * inv,
                                 __uint_as_float(o[4 * q4 + 2]) * inv, __uint_as_float(o[4 * q4 + 3]) * inv);
        }
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    ptx::tmem_dealloc(tmem_base, TMEM_COLS);
    if constexpr (C::Q_SECOND_ALLOC) ptx::tmem_dealloc(tmem_slot[1], 32);
  }
}

template <int DH, bool DROP, int POLY, int NBUF = 2, int CTAS = 2, int SB = 64, int QT = 0, int PIPE = 0, int SPEC = 0>
static int launch_fwd(dim3 grid, const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnParams& p,
                      cudaStream_t st) {
  static bool attr_set = false;
  constexpr size_t SMEM = Cfg<NBUF, CTAS, SB, QT>::SMEM;
  if (!attr_set) {
    ACAI_CUDA(cudaFuncSetAttribute(attn_fwd_tc<DH, DROP, POLY, NBUF, CTAS, SB, QT, PIPE, SPEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    ACAI_CUDA(cudaFuncSetAttribute(attn_fwd_tc<DH, DROP, POLY, NBUF, CTAS, SB, QT, PIPE, SPEC>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    attr_set = true;
  }
  attn_fwd_tc<DH, DROP, POLY, NBUF, CTAS, SB, QT, PIPE, SPEC><<<grid, Cfg<NBUF, CTAS, SB, QT>::THREADS, SMEM, st>>>(tq, tk, tv, p);
  return ACAI_OK;
}

// returns ACAI_ERR_UNSUPPORTED when the problem does not qualify
int attn_varlen_tc(const void* q, const void* k, const void* v, int dtype, int64_t qs, int64_t ks, int64_t vs,
                   const int32_t* cu_q, const int32_t* cu_k, int batch, int max_q, int64_t total_q, int64_t total_k,
                   int heads, int head_dim, int causal, float scale, void* out, int out_dt, int64_t os, float* lse,
                   AttnDropout drop, cudaStream_t st) {
  if (dtype != ACAI_BF16 || (head_dim != 64 && head_dim != 32) || qs % 8 || ks % 8 || vs % 8 || os % 8 || total_q <= 0 || total_k <= 0 ||
      (reinterpret_cast<uintptr_t>(q) & 15) || (reinterpret_cast<uintptr_t>(k) & 15) ||
      (reinterpret_cast<uintptr_t>(v) & 15) || (reinterpret_cast<uintptr_t>(out) & 15))
    return ACAI_ERR_UNSUPPORTED;
  CUtensorMap tq, tk, tv;
  const uint64_t cols = (uint64_t)heads * head_dim;
  int rc = make_tma_2d_bf16(&tq, q, (uint64_t)total_q, cols, (uint64_t)qs, BQ, D);
  if (rc) return rc;
  rc = make_tma_2d_bf16(&tk, k, (uint64_t)total_k, cols, (uint64_t)ks, BKV, D);
  if (rc) return rc;
  rc = make_tma_2d_bf16(&tv, v, (uint64_t)total_k, cols, (uint64_t)vs, BKV, D);
  if (rc) return rc;
  // share of the exponentials moved to the FMA pipe (see POLY); ACAI_ATTN_POLY overrides the tuned default for A/B runs
  static const int poly = [] { const char* e = getenv("ACAI_ATTN_POLY"); const int v = e ? atoi(e) : 0; return v < 0 ? 0 : (v > 2 ? 2 : v); }();
  // resident CTAs per SM (see Cfg): 4 is the measured best (profiles/r1h_attention.md); ACAI_ATTN_CTAS = 2 / 3 for A/B runs
  static const int ctas = [] { const char* e = getenv("ACAI_ATTN_CTAS"); const int v = e ? atoi(e) : 4; return v == 2 || v == 3 ? v : 4; }();
  // softmax block (keys per S / P round) of the four-CTA variant: 32 (half blocks, S double-buffered) is the measured best;
  // ACAI_ATTN_SB = 64 selects the single-buffer two-pass variant for A/B runs
  static const int sb = [] { const char* e = getenv("ACAI_ATTN_SB"); return e && atoi(e) == 64 ? 64 : 32; }();
  // Q as a TMEM-resident MMA operand (see Cfg): ACAI_ATTN_QT = 1 (four CTAs, single S buffer) / 2 (three CTAs, two S buffers)
  static const int qt = [] { const char* e = getenv("ACAI_ATTN_QT"); const int v = e ? atoi(e) : 0; return v == 1 || v == 2 ? v : 0; }();
  AttnParams p{cu_q, cu_k, out, out_dt, os, causal, scale * 1.4426950408889634f, lse, (int)total_q,
               static_cast<const __nv_bfloat16*>(q), qs, drop};
  dim3 grid((max_q + BQ - 1) / BQ, heads, batch);
  // speculative reference maximum (SPEC, see the kernel).  Measured (kernel_bench, two pairs in one box, gpurun_out/r2h_e.log):
  // encoder layer 339.5 / 338.2 -> 333.3 / 333.0 us, causal decoder 163.3 -> 162.2 us; with dropout 448.9 -> 456.3 us and for
  // head_dim 32 123.6 -> 124.7 us - so it is the default for head_dim 64 without dropout only.  ACAI_ATTN_SPEC = 0 / 1 forces it
  // off / on everywhere (A/B runs, read per call).
  const int spec_env = [] { const char* e = getenv("ACAI_ATTN_SPEC"); return e ? (atoi(e) != 0 ? 1 : 0) : -1; }();   // (per call: tests toggle it)
  const bool spec = spec_env >= 0 ? spec_env != 0 : (head_dim == 64 && drop.thr == 0);
#define ACAI_FWD32(DH_, DROP_)                                                                                    \
  (poly == 0 && spec ? launch_fwd<DH_, DROP_, 0, 2, 4, 32, 0, 0, 1>(grid, tq, tk, tv, p, st) :                    \
   poly == 0 ? launch_fwd<DH_, DROP_, 0, 2, 4, 32>(grid, tq, tk, tv, p, st) : poly == 1 ? launch_fwd<DH_, DROP_, 1, 2, 4, 32>(grid, tq, tk, tv, p, st) \
                                                                                        : launch_fwd<DH_, DROP_, 2, 2, 4, 32>(grid, tq, tk, tv, p, st))
  // software-pipelined softmax (PIPE, see the kernel), for A/B runs: ACAI_ATTN_PIPE = 1 three CTAs per SM / 2 four CTAs per SM.  Measured
  // on one box, encoder shape: 337.0 us off, 335.0 us with 2 (three runs each), 355 us with 1; with dropout the extra live registers
  // spill (564 -> 590 us).  S_{j+1} requested BEFORE the exponentials of block j is slower (374 us): S is not ready that early.
  static const int pipe = [] { const char* e = getenv("ACAI_ATTN_PIPE"); const int v = e ? atoi(e) : 0; return v == 1 || v == 2 ? v : 0; }();
#define ACAI_FWD(DH_, DROP_)                                                                                      \
  (pipe == 1 ? launch_fwd<DH_, DROP_, 0, 2, 3, 32, 0, 1>(grid, tq, tk, tv, p, st) : pipe == 2 ? launch_fwd<DH_, DROP_, 0, 2, 4, 32, 0, 1>(grid, tq, tk, tv, p, st) : \
   qt == 1 ? launch_fwd<DH_, DROP_, 0, 1, 4, 32, 1>(grid, tq, tk, tv, p, st) : qt == 2 ? launch_fwd<DH_, DROP_, 0, 2, 3, 32, 1>(grid, tq, tk, tv, p, st) : \
   ctas == 4 && sb == 32 ? ACAI_FWD32(DH_, DROP_) : ctas == 4 ? launch_fwd<DH_, DROP_, 0, 1, 4>(grid, tq, tk, tv, p, st) : ctas == 3 ? launch_fwd<DH_, DROP_, 0, 1, 3>(grid, tq, tk, tv, p, st)       \
             : launch_fwd<DH_, DROP_, 0, 2, 2>(grid, tq, tk, tv, p, st))
  if (drop.thr == 0) rc = head_dim == 64 ? ACAI_FWD(64, false) : ACAI_FWD(32, false);
  else rc = head_dim == 64 ? ACAI_FWD(64, true) : ACAI_FWD(32, true);
#undef ACAI_FWD
#undef ACAI_FWD32
  if (rc) return rc;
  return launch_status("attn_fwd_tc");
}

}  // namespace acai
