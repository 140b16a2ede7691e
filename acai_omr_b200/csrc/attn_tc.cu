// K3 (tensor-core path): packed variable-length flash attention for head_dim 64, bf16 operands.
//
// One CTA = one (sequence, head, 128-query block); 2 CTAs are resident per SM so that one CTA's softmax
// overlaps the other's MMAs.  Per 128-key block j:
//     S_j = Q K_j^T            tcgen05.mma 128x128x16 x4, accumulator in TMEM (128 columns)
//     P_j = exp2(S_j - m)      4 softmax warps, thread = query row: tcgen05.ld -> registers -> bf16 -> swizzled smem
//     O_j = P_j V_j            tcgen05.mma 128x64x16 x8 (V is the MN-major B operand), TMEM (64 columns)
//     acc = acc * corr + O_j   in registers (fp32), so TMEM never needs a read-modify-write
// Q/K/V tiles come in by TMA straight out of the fused [T, 3*H*64] QKV buffer (or separate Q and KV buffers for
// cross attention); ragged sequence ends are masked in the softmax, never padded.
//   warps 0-3: softmax + output     warp 4: TMA producer (+TMEM alloc)     warp 5: MMA issuer
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace acai {

int make_tma_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                     uint32_t box_cols);

namespace fa {
constexpr int BQ = 128, BKV = 128, D = 64;
constexpr int THREADS = 192;
constexpr uint32_t TILE = 128 * 64 * 2;           // 16 KB: one 128-row x 64-col bf16 tile
constexpr uint32_t OFF_Q = 0, OFF_K = TILE, OFF_V = 3 * TILE, OFF_PHI = 5 * TILE, OFF_BAR = 6 * TILE;
constexpr size_t SMEM = 6 * TILE + 1024 + 128;
constexpr uint32_t TMEM_COLS = 256;               // S: 128 columns, O: 64 columns
}  // namespace fa
using namespace fa;

struct AttnParams {
  const int32_t* cu_q; const int32_t* cu_k;
  void* out; int out_dt; int64_t o_stride;
  int causal; float scale_log2;
};

__global__ void __launch_bounds__(THREADS, 2)
attn_fwd_tc64(const __grid_constant__ CUtensorMap tma_q, const __grid_constant__ CUtensorMap tma_k,
              const __grid_constant__ CUtensorMap tma_v, AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* q_full = bars;            // 1
  uint64_t* kv_full = bars + 1;       // 2
  uint64_t* kv_empty = bars + 3;      // 2
  uint64_t* s_full = bars + 5;
  uint64_t* p_full = bars + 6;
  uint64_t* o_full = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, qb = blockIdx.x;
  const int q0 = p.cu_q[b], lq = p.cu_q[b + 1] - q0, k0 = p.cu_k[b], lk = p.cu_k[b + 1] - k0;
  if (qb * BQ >= lq) return;                       // uniform for the whole CTA
  const int shift = lk - lq;
  int kv_end = lk;
  if (p.causal) { const int last = qb * BQ + BQ + shift; kv_end = last < lk ? (last < 0 ? 0 : last) : lk; }
  const int n_kv = (kv_end + BKV - 1) / BKV;

  if (warp == 5 && lane == 0) {
    ptx::mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(&kv_full[s], 1); ptx::mbar_init(&kv_empty[s], 1); }
    ptx::mbar_init(s_full, 1); ptx::mbar_init(p_full, 128); ptx::mbar_init(o_full, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 4) {
    if (lane == 0) { ptx::prefetch_tensormap(&tma_q); ptx::prefetch_tensormap(&tma_k); ptx::prefetch_tensormap(&tma_v); }
    ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_s = tmem_base, tmem_o = tmem_base + 128;

  if (warp == 4) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0 && n_kv > 0) {
      ptx::mbar_arrive_expect_tx(q_full, TILE);
      ptx::tma_load_2d(smem + OFF_Q, &tma_q, q_full, h * D, q0 + qb * BQ);
      for (int j = 0; j < n_kv; ++j) {
        const int s = j & 1;
        ptx::mbar_wait(&kv_empty[s], ((j >> 1) & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(&kv_full[s], 2 * TILE);
        ptx::tma_load_2d(smem + OFF_K + s * TILE, &tma_k, &kv_full[s], h * D, k0 + j * BKV);
        ptx::tma_load_2d(smem + OFF_V + s * TILE, &tma_v, &kv_full[s], h * D, k0 + j * BKV);
      }
    }
  } else if (warp == 5) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0 && n_kv > 0) {
      constexpr uint32_t IDESC_S = ptx::idesc_bf16_f32(BQ, BKV);
      constexpr uint32_t IDESC_O = ptx::idesc_bf16_f32(BQ, D, false, true);   // B = V is MN-major
      const uint64_t qdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_Q), 1024);
      const uint64_t phidesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_PHI), 1024);
      auto issue_s = [&](int j) {
        const int s = j & 1;
        ptx::mbar_wait(&kv_full[s], (j >> 1) & 1);
        ptx::tc_fence_after();
        const uint64_t kdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_K + s * TILE), 1024);
#pragma unroll
        for (int k = 0; k < D / 16; ++k) ptx::umma_f16_ss(tmem_s, qdesc + 2 * k, kdesc + 2 * k, IDESC_S, k != 0);
        ptx::umma_commit(s_full);
      };
      ptx::mbar_wait(q_full, 0);
      issue_s(0);
      for (int j = 0; j < n_kv; ++j) {
        const int s = j & 1;
        ptx::mbar_wait(p_full, j & 1);
        ptx::tc_fence_after();
        const uint64_t plodesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_K + s * TILE), 1024);  // P[:, 0:64] aliases K_j
        const uint64_t vdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + OFF_V + s * TILE), 1024);
#pragma unroll
        for (int kk = 0; kk < BKV / 16; ++kk) {
          const uint64_t adesc = (kk < 4 ? plodesc : phidesc) + 2 * (kk & 3);
          ptx::umma_f16_ss(tmem_o, adesc, vdesc + (uint64_t)kk * (16 * 128 >> 4), IDESC_O, kk != 0);
        }
        ptx::umma_commit(o_full);
        ptx::umma_commit(&kv_empty[s]);
        if (j + 1 < n_kv) issue_s(j + 1);
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax + output (thread = query row)
    const int r = warp * 32 + lane;
    const int qi = qb * BQ + r;
    const bool valid = qi < lq;
    const int my_end = p.causal ? min(lk, qi + shift + 1) : lk;     // exclusive key bound of this row
    const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
    float acc[D];
#pragma unroll
    for (int e = 0; e < D; ++e) acc[e] = 0.0f;
    float m = -INFINITY, l = 0.0f;
    const float scale_log2 = p.scale_log2;
    const uint32_t row_off = (uint32_t)(r >> 3) * 1024 + (uint32_t)(r & 7) * 128;
    const uint32_t sw = (uint32_t)(r & 7);

    for (int j = 0; j < n_kv; ++j) {
      if (j > 0) {                      // fold O_{j-1} (scale m_{j-1}) into the running accumulator
        ptx::mbar_wait(o_full, (j - 1) & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t o[32];
          ptx::tmem_ld_32x32(tmem_o + lane_addr + c * 32, o);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[c * 32 + i] += __uint_as_float(o[i]);
        }
      }
      ptx::mbar_wait(s_full, j & 1);
      ptx::tc_fence_after();
      const int kbase = j * BKV;
      const bool full = kbase + BKV <= my_end;      // no key of this block is masked for this row
      float rmax = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; c += 2) {
        uint32_t s0[32], s1[32];
        ptx::tmem_ld_32x32(tmem_s + lane_addr + c * 32, s0);
        ptx::tmem_ld_32x32(tmem_s + lane_addr + c * 32 + 32, s1);
        ptx::tmem_ld_wait();
        if (full) {
#pragma unroll
          for (int i = 0; i < 32; ++i) rmax = fmaxf(rmax, fmaxf(__uint_as_float(s0[i]), __uint_as_float(s1[i])));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            if (kbase + c * 32 + i < my_end) rmax = fmaxf(rmax, __uint_as_float(s0[i]));
            if (kbase + c * 32 + 32 + i < my_end) rmax = fmaxf(rmax, __uint_as_float(s1[i]));
          }
        }
      }
      const float m_new = fmaxf(m, rmax);
      const float m_use = m_new == -INFINITY ? 0.0f : m_new;
      const float corr = fast_exp2((m - m_use) * scale_log2);        // m = -inf -> 0
      l *= corr;
#pragma unroll
      for (int e = 0; e < D; ++e) acc[e] *= corr;
      const float moff = m_use * scale_log2;
      uint8_t* const p_lo = smem + OFF_K + (j & 1) * TILE;
      uint8_t* const p_hi = smem + OFF_PHI;
      float lsum0 = 0.0f, lsum1 = 0.0f;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t s[32];
        ptx::tmem_ld_32x32(tmem_s + lane_addr + c * 32, s);
        ptx::tmem_ld_wait();
        uint32_t pk[16];
        if (full) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float p0 = fast_exp2(fmaf(__uint_as_float(s[2 * i]), scale_log2, -moff));
            const float p1 = fast_exp2(fmaf(__uint_as_float(s[2 * i + 1]), scale_log2, -moff));
            lsum0 += p0; lsum1 += p1;
            __nv_bfloat162 hh = __floats2bfloat162_rn(p0, p1);
            pk[i] = *reinterpret_cast<uint32_t*>(&hh);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int kk = kbase + c * 32 + 2 * i;
            const float p0 = kk < my_end ? fast_exp2(fmaf(__uint_as_float(s[2 * i]), scale_log2, -moff)) : 0.0f;
            const float p1 = kk + 1 < my_end ? fast_exp2(fmaf(__uint_as_float(s[2 * i + 1]), scale_log2, -moff)) : 0.0f;
            lsum0 += p0; lsum1 += p1;
            __nv_bfloat162 hh = __floats2bfloat162_rn(p0, p1);
            pk[i] = *reinterpret_cast<uint32_t*>(&hh);
          }
        }
        uint8_t* base = (c < 2 ? p_lo : p_hi) + row_off;
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          const uint32_t chunk = (uint32_t)((c & 1) * 4 + q4);
          *reinterpret_cast<uint4*>(base + ((chunk ^ sw) << 4)) = make_uint4(pk[4 * q4], pk[4 * q4 + 1], pk[4 * q4 + 2], pk[4 * q4 + 3]);
        }
      }
      l += lsum0 + lsum1;
      m = m_new;
      ptx::fence_proxy_async_smem();   // make the generic-proxy smem writes visible to the tensor core
      ptx::tc_fence_before();
      ptx::mbar_arrive(p_full);
    }
    if (n_kv > 0) {
      ptx::mbar_wait(o_full, (n_kv - 1) & 1);
      ptx::tc_fence_after();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t o[32];
        ptx::tmem_ld_32x32(tmem_o + lane_addr + c * 32, o);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[c * 32 + i] += __uint_as_float(o[i]);
      }
    }
    if (valid) {
      const float inv = l > 0.0f ? 1.0f / l : 0.0f;
      const int64_t off = (int64_t)(q0 + qi) * p.o_stride + (int64_t)h * D;
      if (p.out_dt == ACAI_BF16) {
        uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + off);
#pragma unroll
        for (int q4 = 0; q4 < 8; ++q4) {
          uint4 v4; __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&v4);
#pragma unroll
          for (int i = 0; i < 4; ++i) hh[i] = __floats2bfloat162_rn(acc[q4 * 8 + 2 * i] * inv, acc[q4 * 8 + 2 * i + 1] * inv);
          op[q4] = v4;
        }
      } else {
        float4* op = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + off);
#pragma unroll
        for (int q4 = 0; q4 < 16; ++q4)
          op[q4] = make_float4(acc[4 * q4] * inv, acc[4 * q4 + 1] * inv, acc[4 * q4 + 2] * inv, acc[4 * q4 + 3] * inv);
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 4) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

// returns ACAI_ERR_UNSUPPORTED when the problem does not qualify
int attn_varlen_tc(const void* q, const void* k, const void* v, int dtype, int64_t qs, int64_t ks, int64_t vs,
                   const int32_t* cu_q, const int32_t* cu_k, int batch, int max_q, int64_t total_q, int64_t total_k,
                   int heads, int head_dim, int causal, float scale, void* out, int out_dt, int64_t os, cudaStream_t st) {
  if (dtype != ACAI_BF16 || head_dim != D || qs % 8 || ks % 8 || vs % 8 || os % 8 || total_q <= 0 || total_k <= 0 ||
      (reinterpret_cast<uintptr_t>(q) & 15) || (reinterpret_cast<uintptr_t>(k) & 15) ||
      (reinterpret_cast<uintptr_t>(v) & 15) || (reinterpret_cast<uintptr_t>(out) & 15))
    return ACAI_ERR_UNSUPPORTED;
  CUtensorMap tq, tk, tv;
  int rc = make_tma_2d_bf16(&tq, q, (uint64_t)total_q, (uint64_t)heads * D, (uint64_t)qs, BQ, D);
  if (rc) return rc;
  rc = make_tma_2d_bf16(&tk, k, (uint64_t)total_k, (uint64_t)heads * D, (uint64_t)ks, BKV, D);
  if (rc) return rc;
  rc = make_tma_2d_bf16(&tv, v, (uint64_t)total_k, (uint64_t)heads * D, (uint64_t)vs, BKV, D);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    ACAI_CUDA(cudaFuncSetAttribute(attn_fwd_tc64, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    attr_set = true;
  }
  AttnParams p{cu_q, cu_k, out, out_dt, os, causal, scale * 1.4426950408889634f};
  dim3 grid((max_q + BQ - 1) / BQ, heads, batch);
  attn_fwd_tc64<<<grid, THREADS, SMEM, st>>>(tq, tk, tv, p);
  return launch_status("attn_fwd_tc64");
}

}  // namespace acai
