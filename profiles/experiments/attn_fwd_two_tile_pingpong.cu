// EXPERIMENT, NOT BUILT (round 1g).  Result on B200, 12-head encoder layer (T = 65 233 tokens, 64 sequences): 494.6 us
// against 402.2 us for attn_tc.cu; 16-head causal decoder layer: 247.9 us against 212.2 us.  Numerically identical
// (the whole -m gpu suite passed with this kernel switched in).  The hand-over removes the MUFU sharing between the two
// softmax groups, but one 512-column CTA per SM exposes every CTA's prologue (TMEM alloc, Q load) and epilogue (O store),
// which two independent CTAs per SM overlap for free; a persistent variant that prefetches the next work item's Q and
// stores O asynchronously would be needed to come out ahead.  Kept as the record of the experiment.
//
// K3, two-tile variant: the forward attention kernel of attn_tc.cu with TWO 128-query tiles per CTA and an explicit
// ping-pong between their softmax warp groups.
//
// Why: with two independent CTAs per SM (attn_tc.cu) the two softmax warps that share a scheduler also share its MUFU
// unit "fairly" - when both are in their exponential phase each runs at half rate, they finish together, do their
// non-MUFU work (TMEM load, row max, TMEM store, barriers) together, and come back together: the pipe idles ~40 % of the
// time (ncu: XU 60 %).  Here one CTA owns both tiles and a pair of named barriers hands the exponential phase back and
// forth, so one group's TMEM / barrier latencies always sit under the other group's exponentials.  K and V tiles are
// loaded once for both query tiles.
//
//   warps 0-3: softmax of tile 0      warps 4-7: softmax of tile 1      warp 8: TMA producer (+TMEM alloc)
//   warp 9: MMA issuer of tile 0      warp 10: MMA issuer of tile 1
// TMEM (512 columns): tile t at 256 t: S[0] +0, S[1] +64, O +128.  Per-tile math, masking, lazy rescaling, dropout and
// the log-sum-exp output are exactly those of attn_fwd_tc (same results bit for bit).
#include <cstdlib>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace acai {

int make_tma_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                     uint32_t box_cols);

namespace fa2 {
constexpr int BQ = 128, BKV = 64, D = 64, KV_STAGES = 4, TILES = 2;
constexpr int THREADS = 352;
constexpr uint32_t QTILE = BQ * D * 2;            // 16 KB
constexpr uint32_t KTILE = BKV * D * 2;           // 8 KB
constexpr uint32_t OFF_Q = 0, OFF_K = TILES * QTILE, OFF_V = OFF_K + KV_STAGES * KTILE, OFF_BAR = OFF_V + KV_STAGES * KTILE;
constexpr size_t SMEM = OFF_BAR + 512;
constexpr uint32_t TMEM_COLS = 512;
constexpr float RESCALE_LOG2 = 8.0f;
}  // namespace fa2

struct AttnParams2 {
  const int32_t* cu_q; const int32_t* cu_k;
  void* out; int out_dt; int64_t o_stride;
  int causal; float scale_log2;
  float* lse; int total_q;
  AttnDropout drop;
};

template <int DH, bool DROP>
__global__ void __launch_bounds__(fa2::THREADS, 1)
attn_fwd_tc2(const __grid_constant__ CUtensorMap tma_q, const __grid_constant__ CUtensorMap tma_k,
             const __grid_constant__ CUtensorMap tma_v, AttnParams2 p) {
  using namespace fa2;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* q_full = bars;                          // 1
  uint64_t* kv_full = bars + 1;                     // KV_STAGES
  uint64_t* kv_empty = kv_full + KV_STAGES;         // KV_STAGES (one arrival per active tile)
  uint64_t* s_full = kv_empty + KV_STAGES;          // [tile][2]
  uint64_t* p_full = s_full + 2 * TILES;            // [tile][2]
  uint64_t* o_step = p_full + 2 * TILES;            // [tile]
  uint64_t* o_full = o_step + TILES;                // [tile]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + TILES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, qb0 = blockIdx.x * TILES;
  const int q0 = p.cu_q[b], lq = p.cu_q[b + 1] - q0, k0 = p.cu_k[b], lk = p.cu_k[b + 1] - k0;
  if (qb0 * BQ >= lq) return;                      // uniform for the whole CTA
  const int shift = lk - lq;
  const int ntiles = (qb0 + 1) * BQ < lq ? 2 : 1;
  auto n_kv_of = [&](int t) {
    int kv_end = lk;
    if (p.causal) { const int last = (qb0 + t) * BQ + BQ + shift; kv_end = last < lk ? (last < 0 ? 0 : last) : lk; }
    return (kv_end + BKV - 1) / BKV;
  };
  const int n_kv0 = n_kv_of(0), n_kv1 = ntiles == 2 ? n_kv_of(1) : 0;
  const int n_kv_max = n_kv0 > n_kv1 ? n_kv0 : n_kv1;          // (causal: the later tile sees at least as many keys)

  if (warp == 9 && lane == 0) {
    ptx::mbar_init(q_full, 1);
    for (int s = 0; s < KV_STAGES; ++s) { ptx::mbar_init(&kv_full[s], 1); ptx::mbar_init(&kv_empty[s], ntiles); }
    for (int i = 0; i < 2 * TILES; ++i) { ptx::mbar_init(&s_full[i], 1); ptx::mbar_init(&p_full[i], 128); }
    for (int t = 0; t < TILES; ++t) { ptx::mbar_init(&o_step[t], 1); ptx::mbar_init(&o_full[t], 1); }
    ptx::fence_barrier_init();
  }
  if (warp == 8) {
    if (lane == 0) { ptx::prefetch_tensormap(&tma_q); ptx::prefetch_tensormap(&tma_k); ptx::prefetch_tensormap(&tma_v); }
    ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_cta = *tmem_slot;

  if (warp == 8) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0 && n_kv_max > 0) {
      ptx::mbar_arrive_expect_tx(q_full, ntiles * QTILE);
      const int hcol = (h * DH) & ~63;
      for (int t = 0; t < ntiles; ++t) ptx::tma_load_2d(smem + OFF_Q + t * QTILE, &tma_q, q_full, hcol, q0 + (qb0 + t) * BQ);
      int s = 0; uint32_t ph = 0;
      for (int j = 0; j < n_kv_max; ++j) {
        ptx::mbar_wait(&kv_empty[s], ph ^ 1);
        ptx::mbar_arrive_expect_tx(&kv_full[s], 2 * KTILE);
        ptx::tma_load_2d(smem + OFF_K + s * KTILE, &tma_k, &kv_full[s], hcol, k0 + j * BKV);
        ptx::tma_load_2d(smem + OFF_V + s * KTILE, &tma_v, &kv_full[s], hcol, k0 + j * BKV);
        if (++s == KV_STAGES) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp >= 9) {
    // ------------------------------------------------------------------ MMA issuer of tile t
    const int t = warp - 9;
    const int n_kv = t == 0 ? n_kv0 : n_kv1;
    if (lane == 0 && t < ntiles && n_kv_max > 0) {
      const uint32_t tmem_base = tmem_cta + t * 256;
      constexpr uint32_t IDESC_S = ptx::idesc_bf16_f32(BQ, BKV);
      constexpr uint32_t IDESC_O = ptx::idesc_bf16_f32(BQ, DH, false, true);  // B = V is MN-major
      const uint64_t hoff = (uint64_t)(((h * DH) & 63) >> 3);
      const uint64_t qdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_Q + t * QTILE), 1024) + hoff;
      uint64_t* sf = s_full + 2 * t;
      uint64_t* pf = p_full + 2 * t;
      auto issue_s = [&](int j) {
        const int st = j % KV_STAGES;
        ptx::mbar_wait(&kv_full[st], (j / KV_STAGES) & 1);
        ptx::tc_fence_after();
        const uint64_t kdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_K + st * KTILE), 1024) + hoff;
#pragma unroll
        for (int k = 0; k < DH / 16; ++k)
          ptx::umma_f16_ss(tmem_base + (j & 1) * 64, qdesc + 2 * k, kdesc + 2 * k, IDESC_S, k != 0);
        ptx::umma_commit(&sf[j & 1]);
      };
      ptx::mbar_wait(q_full, 0);
      if (n_kv > 0) issue_s(0);
      if (n_kv > 1) issue_s(1);
      for (int j = 0; j < n_kv; ++j) {
        const int st = j % KV_STAGES;
        ptx::mbar_wait(&pf[j & 1], (j >> 1) & 1);
        ptx::tc_fence_after();
        const uint64_t vdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_V + st * KTILE), 1024) + hoff;
#pragma unroll
        for (int kk = 0; kk < BKV / 16; ++kk)
          ptx::umma_f16_ts(tmem_base + 128, tmem_base + (j & 1) * 64 + kk * 8, vdesc + (uint64_t)kk * (16 * 128 >> 4),
                           IDESC_O, (j | kk) != 0);
        ptx::umma_commit(&kv_empty[st]);
        ptx::umma_commit(&o_step[t]);
        if (j == n_kv - 1) ptx::umma_commit(&o_full[t]);
        if (j + 2 < n_kv) issue_s(j + 2);
      }
      // key blocks only the other tile needs (causal): this tile still owes the producer its share of the release
      for (int j = n_kv; j < n_kv_max; ++j) {
        const int st = j % KV_STAGES;
        ptx::mbar_wait(&kv_full[st], (j / KV_STAGES) & 1);
        ptx::mbar_arrive(&kv_empty[st]);
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax warps (thread = query row)
    const int t = warp >> 2, wq = warp & 3;        // tile, lane quarter
    const int qb = qb0 + t;
    const int n_kv = t == 0 ? n_kv0 : n_kv1;
    const bool pingpong = ntiles == 2;             // both groups alive: hand the exponential phase back and forth
    const uint32_t tmem_base = tmem_cta + t * 256;
    uint64_t* sf = s_full + 2 * t;
    uint64_t* pf = p_full + 2 * t;
    const int r = wq * 32 + lane;
    const int qi = qb * BQ + r;
    const bool valid = qi < lq;
    const int my_end = p.causal ? max(0, min(lk, qi + shift + 1)) : lk;     // exclusive key bound of this row
    const int warp_min_end = __shfl_sync(0xffffffffu, my_end, 0);          // rows ascend inside a warp
    const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
    const uint32_t tmem_o = tmem_base + lane_addr + 128;
    const float scale_log2 = p.scale_log2;
    float m = -INFINITY, l = 0.0f;
    const uint64_t rowkey = (uint64_t)h * p.total_q + (uint64_t)(q0 + (valid ? qi : 0));
    const uint32_t rowconst = DROP ? attn_drop_rowconst(p.drop.seed, rowkey) : 0u;

    if (pingpong && t == 1 && n_kv_max > 0) asm volatile("bar.arrive 1, 256;" ::: "memory");
    for (int j = 0; j < n_kv; ++j) {
      const uint32_t tmem_s = tmem_base + lane_addr + (j & 1) * 64;
      ptx::mbar_wait(&sf[j & 1], (j >> 1) & 1);
      ptx::tc_fence_after();
      uint32_t s[64];
      ptx::tmem_ld_32x32(tmem_s, *reinterpret_cast<uint32_t(*)[32]>(&s[0]));
      ptx::tmem_ld_32x32(tmem_s + 32, *reinterpret_cast<uint32_t(*)[32]>(&s[32]));
      ptx::tmem_ld_wait();
      const int kbase = j * BKV;
      if (kbase + BKV > warp_min_end) {             // ragged end / causal diagonal: mask keys >= my_end
#pragma unroll
        for (int i = 0; i < 64; ++i) if (kbase + i >= my_end) s[i] = 0xff800000u;   // -inf
      }
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int i = 0; i < 64; i += 4) {
        mx0 = fmaxf(mx0, __uint_as_float(s[i])); mx1 = fmaxf(mx1, __uint_as_float(s[i + 1]));
        mx2 = fmaxf(mx2, __uint_as_float(s[i + 2])); mx3 = fmaxf(mx3, __uint_as_float(s[i + 3]));
      }
      const float m_new = fmaxf(m, fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)));
      const bool grow = (m_new - m) * scale_log2 > RESCALE_LOG2 || (m == -INFINITY && m_new > -INFINITY);
      if (__any_sync(0xffffffffu, grow)) {          // advance the reference maximum (warp-uniform: TMEM ops are collective)
        const float m_use = m_new == -INFINITY ? 0.0f : m_new;
        const float corr = m == -INFINITY ? 0.0f : fast_exp2((m - m_use) * scale_log2);
        l *= corr;
        if (j > 0) {
          ptx::mbar_wait(&o_step[t], (j - 1) & 1);      // O holds P_{j-1} V_{j-1}
          ptx::tc_fence_after();
#pragma unroll
          for (int c = 0; c < DH / 32; ++c) {
            uint32_t o[32];
            ptx::tmem_ld_32x32(tmem_o + c * 32, o);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * corr);
            ptx::tmem_st_32x32(tmem_o + c * 32, o);
          }
        }
        m = m_new;
      }
      const float moff = (m == -INFINITY ? 0.0f : m) * scale_log2;
      // my turn on the MUFU pipe: named barrier 1 + t completes with this group's 128 syncs + the other group's 128
      // arrivals after ITS exponential phase (tile 1 arrives once up front so that tile 0 goes first)
      if (pingpong) asm volatile("bar.sync %0, 256;" ::"r"(1 + t) : "memory");
      float l0 = 0.0f, l1 = 0.0f;
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float x0 = fmaf(__uint_as_float(s[2 * i]), scale_log2, -moff), x1 = fmaf(__uint_as_float(s[2 * i + 1]), scale_log2, -moff);
        const float p0 = fast_exp2(x0);
        const float p1 = fast_exp2(x1);
        l0 += p0; l1 += p1;
        float d0 = p0, d1 = p1;
        if (DROP) {                                   // keys kbase + 2i, kbase + 2i + 1 = one hash pair
          const uint32_t bits = attn_drop_bits(rowconst, rowkey, (uint32_t)(j * (BKV / 2) + i));
          if ((bits & 0xffffu) < p.drop.thr) d0 = 0.0f;
          if ((bits >> 16) < p.drop.thr) d1 = 0.0f;
        }
        const __nv_bfloat162 hh = __floats2bfloat162_rn(d0, d1);
        pk[i] = *reinterpret_cast<const uint32_t*>(&hh);
      }
      l += l0 + l1;
      if (pingpong && !(t == 1 && j == n_kv_max - 1)) asm volatile("bar.arrive %0, 256;" ::"r"(2 - t) : "memory");
      ptx::tmem_st_32x32(tmem_s, pk);
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      ptx::mbar_arrive(&pf[j & 1]);
    }
    // (causal) the other tile has more key blocks: keep the hand-over going without doing work
    if (pingpong) {
      for (int j = n_kv; j < n_kv_max; ++j) {
        asm volatile("bar.sync %0, 256;" ::"r"(1 + t) : "memory");
        if (!(t == 1 && j == n_kv_max - 1)) asm volatile("bar.arrive %0, 256;" ::"r"(2 - t) : "memory");
      }
    }
    if (n_kv > 0) {
      ptx::mbar_wait(&o_full[t], 0);
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
  if (warp == 8) ptx::tmem_dealloc(tmem_cta, TMEM_COLS);
}

template <int DH, bool DROP>
static int launch_fwd2(dim3 grid, const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnParams2& p,
                       cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    ACAI_CUDA(cudaFuncSetAttribute(attn_fwd_tc2<DH, DROP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fa2::SMEM));
    attr_set = true;
  }
  attn_fwd_tc2<DH, DROP><<<grid, fa2::THREADS, fa2::SMEM, st>>>(tq, tk, tv, p);
  return ACAI_OK;
}

// Same contract as attn_varlen_tc (attn_tc.cu); returns ACAI_ERR_UNSUPPORTED when the problem does not qualify.
int attn_varlen_tc2(const void* q, const void* k, const void* v, int dtype, int64_t qs, int64_t ks, int64_t vs,
                    const int32_t* cu_q, const int32_t* cu_k, int batch, int max_q, int64_t total_q, int64_t total_k,
                    int heads, int head_dim, int causal, float scale, void* out, int out_dt, int64_t os, float* lse,
                    AttnDropout drop, cudaStream_t st) {
  using namespace fa2;
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
  AttnParams2 p{cu_q, cu_k, out, out_dt, os, causal, scale * 1.4426950408889634f, lse, (int)total_q, drop};
  dim3 grid((max_q + TILES * BQ - 1) / (TILES * BQ), heads, batch);
  if (drop.thr == 0) rc = head_dim == 64 ? launch_fwd2<64, false>(grid, tq, tk, tv, p, st) : launch_fwd2<32, false>(grid, tq, tk, tv, p, st);
  else rc = head_dim == 64 ? launch_fwd2<64, true>(grid, tq, tk, tv, p, st) : launch_fwd2<32, true>(grid, tq, tk, tv, p, st);
  if (rc) return rc;
  return launch_status("attn_fwd_tc2");
}

}  // namespace acai
