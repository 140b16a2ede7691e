// Micro-benchmark: how long a tcgen05.mma takes as a function of N, of where A lives (shared memory / TMEM) and of how many
// INDEPENDENT accumulators the issue order alternates between.  One CTA on one SM, one elected thread issues `iters` rounds
// of `chains` MMAs (chain c accumulates into its own N columns), then one commit; the time is clock64 around issue + wait.
// A dependent chain (chains = 1) shows the accumulate latency of one instruction, many chains its throughput.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../acai_omr_b200/csrc mma_chain.cu -o mma_chain
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda.h>
#include "common.cuh"
#include "ptx_sm100.cuh"
using namespace acai;

template <int N, bool TS, int CHAINS>
__global__ void __launch_bounds__(128, 1) mma_chain(int iters, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];       // A tile 128 x 64 bf16 (16 KB) + B tile N x 64 bf16
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (16384 + N * 128) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_barrier_init(); }
  if (warp == 0) ptx::tmem_alloc(&slot, 512);
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tb = __shfl_sync(0xffffffffu, slot, 0);
  if (warp == 0) {
    const uint64_t adesc = ptx::smem_desc_sw128(ptx::smem_u32(smem), 1024);
    const uint64_t bdesc = ptx::smem_desc_sw128(ptx::smem_u32(smem + 16384), 1024);
    constexpr uint32_t IDESC = ptx::idesc_bf16_f32(128, N);
    const uint32_t a_tmem = tb + 480;                      // 32 columns of whatever is there: timing only
    long long t0 = 0, t1 = 0;
    for (int rep = 0; rep < 2; ++rep) {                    // rep 0 warms up
      __syncwarp();
      t0 = clock64();
      if (ptx::elect_one_sync()) {
        for (int it = 0; it < iters; it += 4) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int c = 0; c < CHAINS; ++c) {
              const uint32_t d = tb + (uint32_t)(c * N);
              if (TS) ptx::umma_f16_ts(d, a_tmem + k * 8, bdesc + 2 * k, IDESC, 1u);
              else ptx::umma_f16_ss(d, adesc + 2 * k, bdesc + 2 * k, IDESC, 1u);
            }
          }
        }
        ptx::umma_commit(&bar);
      }
      __syncwarp();
      ptx::mbar_wait_spin(&bar, rep & 1);
      t1 = clock64();
    }
    if (threadIdx.x == 0) out[0] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tb, 512);
}

template <int N, bool TS, int CHAINS>
static void run(long long* dout, int iters = 512) {
  constexpr int chains = CHAINS;
  const size_t smem = 16384 + N * 128 + 1024;
  cudaFuncSetAttribute(mma_chain<N, TS, CHAINS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  mma_chain<N, TS, CHAINS><<<1, 128, smem>>>(iters, dout);
  cudaError_t e = cudaDeviceSynchronize();
  long long clk = 0;
  cudaMemcpy(&clk, dout, sizeof(clk), cudaMemcpyDeviceToHost);
  const double per = (double)clk / (iters * chains);
  if (iters >= 64)
    printf("N=%3d A=%s chains=%d : %7.1f clk per MMA  (floor %d)  %s\n", N, TS ? "tmem" : "smem", chains, per, N / 2,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
  else
    printf("N=%3d A=%s: %d MMAs + commit -> barrier seen by the issuing warp after %lld clk (floor %d)\n", N, TS ? "tmem" : "smem",
           iters * chains, clk, iters * chains * N / 2);
}

template <int CHAINS>
static void sweep(long long* dout) {
  if constexpr (CHAINS * 32 <= 448) { run<32, true, CHAINS>(dout); run<32, false, CHAINS>(dout); }
  if constexpr (CHAINS * 64 <= 448) { run<64, true, CHAINS>(dout); run<64, false, CHAINS>(dout); }
  if constexpr (CHAINS * 128 <= 448) { run<128, true, CHAINS>(dout); run<128, false, CHAINS>(dout); }
  if constexpr (CHAINS * 256 <= 448) { run<256, false, CHAINS>(dout); }
}

int main() {
  long long* dout;
  cudaMalloc(&dout, 64);
  sweep<1>(dout); sweep<2>(dout); sweep<4>(dout); sweep<8>(dout);
  // latency of a short burst: issue -> tcgen05.commit -> mbarrier phase visible to a polling warp
  run<32, true, 1>(dout, 4); run<32, true, 2>(dout, 4); run<32, true, 2>(dout, 8); run<64, true, 1>(dout, 4); run<64, false, 1>(dout, 4);
  run<64, true, 2>(dout, 4); run<128, false, 1>(dout, 4); run<256, false, 1>(dout, 4);
  return 0;
}
