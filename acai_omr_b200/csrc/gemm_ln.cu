// K2+K5 fused: the post-norm sub-layer tail of every transformer block as ONE kernel
//
//   x' = LayerNorm( A[M,K] . W[N,K]^T + bias + residual ) * gamma + beta          (out_proj / linear2 + residual + norm)
//
// (nn.TransformerEncoderLayer / DecoderLayer with norm_first = False, models.py:31,423 in the reference:
// x = norm1(x + sa(x)), x = norm2(x + ff(x)).)  The separate pipeline writes the pre-norm sum y to HBM, reads it back in a
// LayerNorm kernel and writes x' - for the fp32 residual stream that is 18 bytes per element against 10 here.
//
// A LayerNorm row spans the whole N = 512 / 768 / 1024 columns, more than one CTA's 256-column accumulator, so the row
// statistics are combined ACROSS CTAs: a thread-block cluster of CL = N / 256 CTAs owns one 128-row M tile, CTA r of the
// cluster computes columns [256 r, 256 r + 256) exactly like gemm_bf16_tcgen05<256> (TMA producer warp / single-thread
// tcgen05.mma issuer / 8 epilogue warps, fp32 accumulator double-buffered in TMEM), and the epilogue does
//   pass A   v = acc + bias + residual, written back over the accumulator in TMEM (tcgen05.st); sum and sum of squares of this
//            thread's 128 columns around a pivot (the row's first value), converted to (mean, centred M2)
//   combine  the two column halves of the CTA through shared memory, then (mean, M2) of the CTA's 256 columns is PUSHED to
//            every CTA of the cluster through distributed shared memory (st.shared::cluster + a remote mbarrier arrive);
//            Chan's parallel-variance formula merges the CL partial moments: the two-pass variance of acai_layernorm_fwd up
//            to fp32 rounding
//   pass B   second TMEM read, normalise, scale / shift, stage 128-byte row pieces in swizzled shared memory and TMA-store
//            the bf16 copy (the next GEMM's operand) and, for an fp32 residual stream, the fp32 copy.
// y never exists in HBM.  Persistent: one CTA per SM, cluster c walks M tiles c, c + #clusters, ...; every CTA of a
// cluster runs the same tile sequence, so the push / wait pairs line up without any other synchronisation.
#include <mutex>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace acai {

int make_tma_2d(CUtensorMap* map, const void* ptr, int dtype, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                uint32_t box_cols);

#ifndef GLN_RDEPTH
#define GLN_RDEPTH 2
#endif
#ifndef GLN_SBUF
#define GLN_SBUF 1
#endif

namespace gln {
constexpr int BM = 128, BN = 256, BK = 64, UMMA_K = 16, EPI_WARPS = 8, THREADS = 64 + EPI_WARPS * 32, STAGES = 3;
constexpr uint32_t A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;
constexpr int CPW = BN / 2;      // columns per epilogue thread (thread = row, two column halves per CTA)

struct Params {
  const float* bias; const void* res; int64_t ldr; const float* gamma; const float* beta; float eps;
};

template <int CL> constexpr size_t smem_bytes() {
  return (size_t)STAGES * (A_BYTES + B_BYTES) + GLN_SBUF * EPI_WARPS * 4096 /*store staging*/ + 3 * BN * 4 /*bias, gamma, beta*/ +
         2 * BM * 8 /*half combine*/ + 2 * CL * BM * 8 /*cluster moments, double-buffered*/ + 256 /*barriers*/;
}

// ---- cluster-scope PTX -------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cluster_id() { uint32_t r; asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t num_clusters() { uint32_t r; asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r)); return r; }
// address of the same shared-memory location in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
  uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank)); return r;
}
// asynchronous remote store: the 8 bytes land in the peer's shared memory and complete 8 bytes of the peer's mbarrier
// transaction count when they do - the data IS the signal, no fence and no separate arrive (a release-arrive at cluster
// scope costs a MEMBAR.GPU per peer)
__device__ __forceinline__ void st_async_f32x2(uint32_t cluster_addr, float a, float b, uint32_t cluster_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];"
               ::"r"(cluster_addr), "f"(a), "f"(b), "r"(cluster_bar) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(ptx::smem_u32(bar)), "r"(parity), "r"(0x989680u) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory"); }

// The residual chunk of a warp: 32 rows x 32 columns (RES: 1 = bf16 -> 64 bytes per row, 2 = fp32 -> 128 bytes per row).
// Loaded COALESCED - the PPR lanes that share a row read its consecutive 16-byte pieces, so one warp instruction touches
// 32 / PPR rows instead of 32 (thread = row loads are 32 cache lines per instruction and made the LSU the bottleneck of
// the epilogue: 49 of 160 us, profiles/r2b_linear_layernorm.md) - and transposed to thread = row through the warp's
// swizzled staging tile.
template <int RES> struct ResChunk { uint4 v[RES == 2 ? 8 : 4]; };
template <int RES>
__device__ __forceinline__ void res_load(const Params& p, int row0, int M, int col, int lane, ResChunk<RES>& rc) {
  constexpr int PPR = RES == 2 ? 8 : 4, RPI = 32 / PPR;          // pieces per row, rows per warp instruction
  const int r = lane / PPR, pc = lane % PPR;
#pragma unroll
  for (int j = 0; j < PPR; ++j) {
    const int row = row0 + j * RPI + r;
#ifdef GLN_ABL_NORES
    rc.v[j] = make_uint4(0u, 0u, 0u, 0u);
#else
    rc.v[j] = row < M ? __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(p.res) +
                                                              ((int64_t)row * p.ldr + col) * (RES == 2 ? 4 : 2)) + pc)
                      : make_uint4(0u, 0u, 0u, 0u);
#endif
  }
}
// staging tile <- the chunk (row rr, piece pc at rr * 128 + ((pc ^ (rr & 7)) << 4)); then v += this thread's row
template <int RES>
__device__ __forceinline__ void res_add(const ResChunk<RES>& rc, uint8_t* stage, int lane, float (&v)[32]) {
  constexpr int PPR = RES == 2 ? 8 : 4, RPI = 32 / PPR;
  const int r = lane / PPR, pc = lane % PPR;
#pragma unroll
  for (int j = 0; j < PPR; ++j) {
    const int rr = j * RPI + r;
    *reinterpret_cast<uint4*>(stage + rr * 128 + ((pc ^ (rr & 7)) << 4)) = rc.v[j];
  }
  __syncwarp();
  const uint8_t* mine = stage + lane * 128;
  const int sw = lane & 7;
  if constexpr (RES == 1) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint4 raw = *reinterpret_cast<const uint4*>(mine + ((q ^ sw) << 4));
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
      for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h[i]); v[8 * q + 2 * i] += f.x; v[8 * q + 2 * i + 1] += f.y; }
    }
  } else {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float4 f = *reinterpret_cast<const float4*>(mine + ((q ^ sw) << 4));
      v[4 * q] += f.x; v[4 * q + 1] += f.y; v[4 * q + 2] += f.z; v[4 * q + 3] += f.w;
    }
  }
  __syncwarp();
}

// RES: 1 = bf16 residual, 2 = fp32 residual.  OUT32: also store the normalised rows in fp32 (the residual stream copy).
template <int CL, int RES, bool OUT32>
__global__ void __launch_bounds__(THREADS, 1)
gemm_ln_tcgen05(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                const __grid_constant__ CUtensorMap tma_out16, const __grid_constant__ CUtensorMap tma_out32, Params p, int M, int K) {
  constexpr int N = CL * BN;
  constexpr uint32_t TMEM_COLS = 2 * BN;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_BYTES;
  uint8_t* staging = smem + STAGES * (A_BYTES + B_BYTES);                 // 8 warps x 2 x 4 KB
  float* sbias = reinterpret_cast<float*>(staging + GLN_SBUF * EPI_WARPS * 4096);    // [BN]
  float* sgamma = sbias + BN;
  float* sbeta = sgamma + BN;
  float2* comb = reinterpret_cast<float2*>(sbeta + BN);                   // [2 halves][BM] (mean, M2) of 128 columns
  float2* moments = comb + 2 * BM;                                        // [2 buffers][CL ranks][BM] (mean, M2) of 256 columns
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(moments + 2 * CL * BM);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* stat_bar = tempty_bar + 2;                                    // [2]: CL * BM * 8 bytes of peer moments per tile
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(stat_bar + 2);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // (provably warp-uniform: ptx::elect_one_sync)
  const uint32_t rank = cluster_ctarank();
  const int n0 = (int)rank * BN;
  const int num_m = (M + BM - 1) / BM, num_k = K / BK;
  const int first = (int)cluster_id(), step = (int)num_clusters();

  if (warp == 0 && lane == 0) {
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    ptx::prefetch_tensormap(&tma_a);
    ptx::prefetch_tensormap(&tma_b);
    ptx::prefetch_tensormap(&tma_out16);
    if constexpr (OUT32) ptx::prefetch_tensormap(&tma_out32);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&full_bar[s], 1); ptx::mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tfull_bar[s], 1); ptx::mbar_init(&tempty_bar[s], EPI_WARPS); ptx::mbar_init(&stat_bar[s], 1);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  cluster_sync_all();            // every CTA's barriers exist before anyone arrives on them remotely
  // Programmatic dependent launch, one way only (profiles/r2b_linear_layernorm.md, 12-layer encoder step): THIS kernel is
  // launched in plain stream order - started early, its CTAs take whole SMs and all 512 TMEM columns while the previous
  // kernel (attention: several waves of 4 CTAs per SM) still has CTAs to place: 17.7 vs 16.4 ms - but it lets ITS dependents
  // start early (they can only land on the SMs no cluster could use): 16.4 -> 16.15 ms.  -DGLN_PDL_WAIT restores the former.
#ifdef GLN_PDL_WAIT
  pdl_wait();
#endif
  pdl_launch_dependents();

  // (producer / issuer: converged loops, one elected lane issues - see ptx::elect_one_sync)
  if (warp == 0) {
    int stage = 0; uint32_t phase = 0;
    for (int mt = first; mt < num_m; mt += step) {
      for (int kb = 0; kb < num_k; ++kb) {
        ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
        if (ptx::elect_one_sync()) {
          ptx::mbar_arrive_expect_tx(&full_bar[stage], A_BYTES + B_BYTES);
          ptx::tma_load_2d(smem_a + stage * A_BYTES, &tma_a, &full_bar[stage], kb * BK, mt * BM);
          ptx::tma_load_2d(smem_b + stage * B_BYTES, &tma_b, &full_bar[stage], kb * BK, n0);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t IDESC = ptx::idesc_bf16_f32(BM, BN);
    const uint32_t tb0 = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint64_t adesc0 = ptx::smem_desc_sw128(ptx::smem_u32(smem_a), 1024);
    const uint64_t bdesc0 = ptx::smem_desc_sw128(ptx::smem_u32(smem_b), 1024);
    int stage = 0; uint32_t phase = 0; int it = 0;
    for (int mt = first; mt < num_m; mt += step, ++it) {
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
      ptx::tc_fence_after();
      const uint32_t d_tmem = tb0 + as * BN;
      for (int kb = 0; kb < num_k; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        if (ptx::elect_one_sync()) {
          const uint64_t adesc = adesc0 + (uint64_t)(stage * (A_BYTES >> 4)), bdesc = bdesc0 + (uint64_t)(stage * (B_BYTES >> 4));
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            ptx::umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, IDESC, (kb != 0 || k != 0) ? 1u : 0u);
          ptx::umma_commit(&empty_bar[stage]);
          if (kb == num_k - 1) ptx::umma_commit(&tfull_bar[as]);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    const int ew = warp - 2;             // 0..7
    const int lane_grp = warp & 3;       // TMEM lane quarter of this warp
    const int half = ew >> 2;            // column half of the CTA's 256 columns
    const int trow = lane_grp * 32 + lane;                       // row inside the tile, 0..127
    const int etid = (int)threadIdx.x - 64;                      // 0..255
    uint8_t* const stage0 = staging + ew * (GLN_SBUF * 4096);                 // two 4 KB tiles: one is being read out by TMA while the
    const uint32_t sw = (uint32_t)(lane & 7);                    // other is filled
    int sbuf = 0;
    // the CTA's columns never change: parameters are staged once
    sbias[etid] = p.bias ? __ldg(p.bias + n0 + etid) : 0.0f;
    sgamma[etid] = __ldg(p.gamma + n0 + etid);
    sbeta[etid] = __ldg(p.beta + n0 + etid);
    epi_bar_sync();
    const float* sb = sbias + half * CPW;
    const float* sg = sgamma + half * CPW;
    const float* sbt = sbeta + half * CPW;
    const int ccol0 = n0 + half * CPW;                            // first global column of this thread
    int it = 0;
    for (int mt = first; mt < num_m; mt += step, ++it) {
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      const int row0 = mt * BM + lane_grp * 32, grow = row0 + lane;
      (void)grow;
#ifndef GLN_ABL_NOSTAT
      if (etid == 0) ptx::mbar_arrive_expect_tx(&stat_bar[as], CL * BM * 8);   // this tile's moments: 8 bytes per row and peer
#endif
      constexpr int RDEPTH = RES == 2 ? GLN_RDEPTH : (GLN_RDEPTH == 3 ? 4 : GLN_RDEPTH);   // residual chunks in flight (register budget: 168)
      ResChunk<RES> rc[RDEPTH];
#pragma unroll
      for (int c = 0; c < RDEPTH; ++c) res_load<RES>(p, row0, M, ccol0 + c * 32, lane, rc[c]);     // issued under the MMAs
      if (lane == 0) ptx::tma_store_wait_read<0>();               // the staging tile doubles as the transpose buffer: the previous
      __syncwarp();                                               // tile's last store must have been read out of it
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + as * BN + half * CPW;
      // ---- pass A1: v = acc + bias + residual -> TMEM; moments of the thread's 128 columns around a pivot (its first value:
      // within a few standard deviations of the row mean, so S2 - S1^2 / n loses a bit or two, not the cancellation of raw
      // E[x^2] - mean^2) - one pass, where the centred second moment used to cost a second TMEM read of the row
      float s1 = 0.0f, s2 = 0.0f, pivot = 0.0f;
#pragma unroll
      for (int ci = 0; ci < CPW / 32; ++ci) {
        const int c = ci * 32;
        uint32_t acc[32];
        ptx::tmem_ld_32x32(taddr + c, acc);
        ptx::tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 b4 = reinterpret_cast<const float4*>(sb + c)[q];
          v[4 * q] = __uint_as_float(acc[4 * q]) + b4.x; v[4 * q + 1] = __uint_as_float(acc[4 * q + 1]) + b4.y;
          v[4 * q + 2] = __uint_as_float(acc[4 * q + 2]) + b4.z; v[4 * q + 3] = __uint_as_float(acc[4 * q + 3]) + b4.w;
        }
        res_add<RES>(rc[ci % RDEPTH], stage0, lane, v);
        if (ci + RDEPTH < CPW / 32) res_load<RES>(p, row0, M, ccol0 + (ci + RDEPTH) * 32, lane, rc[ci % RDEPTH]);
        if (ci == 0) pivot = v[0];
#pragma unroll
        for (int j = 0; j < 32; ++j) { const float d = v[j] - pivot; s1 += d; s2 = fmaf(d, d, s2); acc[j] = __float_as_uint(v[j]); }
        ptx::tmem_st_32x32(taddr + c, acc);
      }
      ptx::tmem_st_wait();
      const float mean_t = pivot + s1 * (1.0f / CPW);
      const float m2 = fmaxf(s2 - s1 * s1 * (1.0f / CPW), 0.0f);
      // ---- the two halves of the CTA -> (mean, M2) over 256 columns; half 0 pushes it to every CTA of the cluster
      comb[half * BM + trow] = make_float2(mean_t, m2);
      epi_bar_sync();
      if (half == 0) {
        const float2 o = comb[BM + trow];
        const float d = o.x - mean_t;
        const float mean_c = 0.5f * (mean_t + o.x);
        const float m2_c = m2 + o.y + d * d * (0.5f * CPW);       // Chan: n_a n_b / (n_a + n_b) = 64
        const uint32_t slot = ptx::smem_u32(&moments[(as * CL + (int)rank) * BM + trow]);
        const uint32_t bar = ptx::smem_u32(&stat_bar[as]);
#ifndef GLN_ABL_NOSTAT
#pragma unroll
        for (int r = 0; r < CL; ++r) st_async_f32x2(mapa(slot, (uint32_t)r), mean_c, m2_c, mapa(bar, (uint32_t)r));
#else
        (void)slot; (void)bar; moments[(as * CL) * BM + trow] = make_float2(mean_c, m2_c);
#endif
      }
#ifndef GLN_ABL_NOSTAT
      mbar_wait_cluster(&stat_bar[as], aphase);
#endif
      float mean = 0.0f;
      float2 mo[CL];
#pragma unroll
      for (int r = 0; r < CL; ++r) { mo[r] = moments[(as * CL + r) * BM + trow]; mean += mo[r].x; }
      mean *= (1.0f / CL);
      float M2 = 0.0f;
#pragma unroll
      for (int r = 0; r < CL; ++r) { const float d = mo[r].x - mean; M2 += mo[r].y + (float)BN * d * d; }
      const float rstd = rsqrtf(M2 * (1.0f / N) + p.eps);
      // ---- pass B: normalise 64 columns at a time; bf16 piece = one 128-byte row chunk, fp32 = two
#pragma unroll 1
      for (int c = 0; c < CPW; c += 64) {
        uint32_t acc[2][32];
        ptx::tmem_ld_32x32(taddr + c, acc[0]);
        ptx::tmem_ld_32x32(taddr + c + 32, acc[1]);
        ptx::tmem_ld_wait();
        if (c + 64 == CPW) {                                      // accumulator fully consumed: the MMAs of tile it + 2 may start
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
        }
        float y[2][32];
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 g4 = reinterpret_cast<const float4*>(sg + c + hh * 32)[q];
            const float4 b4 = reinterpret_cast<const float4*>(sbt + c + hh * 32)[q];
            y[hh][4 * q] = (__uint_as_float(acc[hh][4 * q]) - mean) * rstd * g4.x + b4.x;
            y[hh][4 * q + 1] = (__uint_as_float(acc[hh][4 * q + 1]) - mean) * rstd * g4.y + b4.y;
            y[hh][4 * q + 2] = (__uint_as_float(acc[hh][4 * q + 2]) - mean) * rstd * g4.z + b4.z;
            y[hh][4 * q + 3] = (__uint_as_float(acc[hh][4 * q + 3]) - mean) * rstd * g4.w + b4.w;
          }
        }
        // bf16 copy: 64 columns = 128 bytes of this thread's row
        uint8_t* stage = stage0 + sbuf * 4096;
        uint8_t* srow = stage + lane * 128;
        sbuf ^= (GLN_SBUF - 1);
        if (lane == 0) ptx::tma_store_wait_read<GLN_SBUF - 1>();               // the store that last used THIS tile has been read out
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          uint4 pk; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
          const float* src = &y[q >> 2][(q & 3) * 8];
#pragma unroll
          for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(src[2 * i], src[2 * i + 1]);
          *reinterpret_cast<uint4*>(srow + ((((uint32_t)q) ^ sw) << 4)) = pk;
        }
        ptx::fence_proxy_async_smem();
        __syncwarp();
        #ifndef GLN_ABL_NOSTORE
        if (lane == 0) { ptx::tma_store_2d(&tma_out16, stage, ccol0 + c, row0); ptx::tma_store_commit(); }
#endif
        if constexpr (OUT32) {
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            stage = stage0 + sbuf * 4096;
            srow = stage + lane * 128;
            sbuf ^= (GLN_SBUF - 1);
            if (lane == 0) ptx::tma_store_wait_read<GLN_SBUF - 1>();
            __syncwarp();
#pragma unroll
            for (int q = 0; q < 8; ++q)
              *reinterpret_cast<float4*>(srow + ((((uint32_t)q) ^ sw) << 4)) =
                  make_float4(y[hh][4 * q], y[hh][4 * q + 1], y[hh][4 * q + 2], y[hh][4 * q + 3]);
            ptx::fence_proxy_async_smem();
            __syncwarp();
            #ifndef GLN_ABL_NOSTORE
            if (lane == 0) { ptx::tma_store_2d(&tma_out32, stage, ccol0 + c + hh * 32, row0); ptx::tma_store_commit(); }
#endif
          }
        }
      }
    }
    if (lane == 0) ptx::tma_store_wait<0>();
  }
  ptx::tc_fence_before();
  __syncthreads();
  cluster_sync_all();            // nobody leaves while a peer may still push into its shared memory
  if (warp == 2) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}
}  // namespace gln

using namespace gln;

// How many clusters of CL CTAs of this kernel can be resident at once (clusters live inside one GPC, so this is less than
// #SMs / CL on parts whose GPCs do not hold a multiple of CL SMs).  Queried once per instantiation.
template <int CL, int RES, bool OUT32>
static int resident_clusters() {
  static int cached = 0;
  static std::once_flag once;
  std::call_once(once, [] {
    auto kernel = gemm_ln_tcgen05<CL, RES, OUT32>;
    constexpr size_t smem = smem_bytes<CL>();
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(num_sms() / CL * CL)); cfg.blockDim = dim3(THREADS); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) == cudaSuccess && n > 0) cached = n;
    else cached = num_sms() / CL;
    (void)cudaGetLastError();
  });
  return cached;
}

template <int CL, int RES, bool OUT32>
static int launch_ln(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& t16, const CUtensorMap& t32, const Params& p,
                     int M, int K, cudaStream_t st) {
  const int resident = resident_clusters<CL, RES, OUT32>();
  if (resident <= 0) { set_error("gemm_ln_tcgen05: no cluster of %d CTAs fits this device", CL); return ACAI_ERR_CUDA; }
  const int num_m = (M + BM - 1) / BM;
  const int clusters = num_m < resident ? num_m : resident;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(clusters * CL)); cfg.blockDim = dim3(THREADS); cfg.dynamicSmemBytes = smem_bytes<CL>(); cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
#ifdef GLN_PDL_WAIT
  if (pdl_enabled()) cfg.numAttrs = 2;
#endif
  ACAI_CUDA(cudaLaunchKernelEx(&cfg, gemm_ln_tcgen05<CL, RES, OUT32>, ta, tb, t16, t32, p, M, K));
  return launch_status("gemm_ln_tcgen05");
}

template <int CL>
static int dispatch_ln(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& t16, const CUtensorMap& t32, const Params& p,
                       int res_dt, bool out32, int M, int K, cudaStream_t st) {
  if (res_dt == ACAI_BF16) return out32 ? launch_ln<CL, 1, true>(ta, tb, t16, t32, p, M, K, st) : launch_ln<CL, 1, false>(ta, tb, t16, t32, p, M, K, st);
  return out32 ? launch_ln<CL, 2, true>(ta, tb, t16, t32, p, M, K, st) : launch_ln<CL, 2, false>(ta, tb, t16, t32, p, M, K, st);
}

}  // namespace acai

using namespace acai;

// Diagnostics: how many clusters of `cluster_size` CTAs (2, 3 or 4) of the kernel can be resident on this device (0: none).
extern "C" int acai_linear_layernorm_resident_clusters(int32_t cluster_size) {
  switch (cluster_size) {
    case 2: return resident_clusters<2, 2, true>();
    case 3: return resident_clusters<3, 2, true>();
    case 4: return resident_clusters<4, 2, true>();
    default: return 0;
  }
}

// C ABI: see include/acai_b200.h.  ACAI_ERR_UNSUPPORTED (nothing launched) when the problem does not qualify; the caller
// then runs acai_linear_fwd + acai_layernorm_fwd.
extern "C" int acai_linear_layernorm(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias,
                                     const void* residual, int32_t res_dtype, int64_t ldr, const float* gamma, const float* beta,
                                     float eps, void* out_bf16, int64_t ldo16, void* out_f32, int64_t ldo32, int32_t M, int32_t N,
                                     int32_t K, void* stream) {
  ACAI_CHECK_ARG(a && w && residual && gamma && beta && out_bf16, "acai_linear_layernorm: null pointer");
  ACAI_CHECK_ARG(M >= 0 && N > 0 && K > 0 && lda >= K && ldw >= K && ldr >= N && ldo16 >= N && (!out_f32 || ldo32 >= N),
                 "acai_linear_layernorm: bad sizes M=%d N=%d K=%d", M, N, K);
  ACAI_CHECK_ARG(dtype_valid(res_dtype), "acai_linear_layernorm: bad residual dtype %d", res_dtype);
  if (M == 0) return ACAI_OK;
  const int cl = N / gln::BN;
  const int res_es = res_dtype == ACAI_BF16 ? 2 : 4;
  auto mis = [](const void* ptr) { return (reinterpret_cast<uintptr_t>(ptr) & 15) != 0; };
  if (N % gln::BN != 0 || cl < 2 || cl > 4 || K % gln::BK != 0 || lda % 8 || ldw % 8 || mis(a) || mis(w) || mis(residual) ||
      (ldr * res_es) % 16 || mis(out_bf16) || ldo16 % 8 || (out_f32 && (mis(out_f32) || ldo32 % 4)))
    return ACAI_ERR_UNSUPPORTED;
  CUtensorMap ta, tb, t16, t32;
  int rc = make_tma_2d(&ta, a, ACAI_BF16, (uint64_t)M, (uint64_t)K, (uint64_t)lda, gln::BM, gln::BK);
  if (rc) return rc;
  rc = make_tma_2d(&tb, w, ACAI_BF16, (uint64_t)N, (uint64_t)K, (uint64_t)ldw, gln::BN, gln::BK);
  if (rc) return rc;
  rc = make_tma_2d(&t16, out_bf16, ACAI_BF16, (uint64_t)M, (uint64_t)N, (uint64_t)ldo16, 32, 64);
  if (rc) return rc;
  t32 = t16;
  if (out_f32) {
    rc = make_tma_2d(&t32, out_f32, ACAI_F32, (uint64_t)M, (uint64_t)N, (uint64_t)ldo32, 32, 32);
    if (rc) return rc;
  }
  gln::Params p{bias, residual, ldr, gamma, beta, eps};
  cudaStream_t st = (cudaStream_t)stream;
  switch (cl) {
    case 2: return dispatch_ln<2>(ta, tb, t16, t32, p, res_dtype, out_f32 != nullptr, M, K, st);
    case 3: return dispatch_ln<3>(ta, tb, t16, t32, p, res_dtype, out_f32 != nullptr, M, K, st);
    default: return dispatch_ln<4>(ta, tb, t16, t32, p, res_dtype, out_f32 != nullptr, M, K, st);
  }
}
