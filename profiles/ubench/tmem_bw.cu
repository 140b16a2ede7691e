// Micro-benchmark: tcgen05.ld / tcgen05.st throughput per SM as a function of the number of warps issuing
// (one CTA on one SM, 512 TMEM columns).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../acai_omr_b200/csrc
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda.h>
#include "common.cuh"
#include "ptx_sm100.cuh"
using namespace acai;

// consume every loaded register with STATIC indices (a dynamic index would send the arrays to local memory and the
// benchmark would time STL traffic instead of tcgen05.ld - the first version of this file did exactly that)
template <int N> __device__ __forceinline__ uint32_t fold(const uint32_t (&a)[N]) {
  uint32_t x = 0;
#pragma unroll
  for (int i = 0; i < N; ++i) x ^= a[i];
  return x;
}

__global__ void __launch_bounds__(512, 1) tmem_bw(int iters, int mode, long long* out, uint32_t* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) ptx::tmem_alloc(&slot, 512);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;
  uint32_t r[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) r[i] = threadIdx.x + i;
  ptx::tmem_st_32x32(base, r);
  ptx::tmem_st_32x32(base + 32, r);
  ptx::tmem_st_wait();
  __syncthreads();
  const long long t0 = clock64();
  uint32_t acc = 0;
  for (int it = 0; it < iters; ++it) {
    if (mode == 0) {            // loads: 2 x 32 columns, then wait
      uint32_t a[32], b[32];
      ptx::tmem_ld_32x32(base, a);
      ptx::tmem_ld_32x32(base + 32, b);
      ptx::tmem_ld_wait();
      acc += fold(a) ^ fold(b);
    } else if (mode == 2) {     // 4 x 32-column loads in flight before one wait
      uint32_t a[32], b[32], c[32], d[32];
      ptx::tmem_ld_32x32(base, a);
      ptx::tmem_ld_32x32(base + 32, b);
      ptx::tmem_ld_32x32(base, c);
      ptx::tmem_ld_32x32(base + 32, d);
      ptx::tmem_ld_wait();
      acc += fold(a) ^ fold(b) ^ fold(c) ^ fold(d);
    } else if (mode == 3) {     // 16x256b.x8: 16 lanes x 64 columns per instruction (32 registers), two per iteration = same bytes
      uint32_t a[32], b[32];
      asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                   : "=r"(a[0]),"=r"(a[1]),"=r"(a[2]),"=r"(a[3]),"=r"(a[4]),"=r"(a[5]),"=r"(a[6]),"=r"(a[7]),"=r"(a[8]),"=r"(a[9]),"=r"(a[10]),"=r"(a[11]),"=r"(a[12]),"=r"(a[13]),"=r"(a[14]),"=r"(a[15]),"=r"(a[16]),"=r"(a[17]),"=r"(a[18]),"=r"(a[19]),"=r"(a[20]),"=r"(a[21]),"=r"(a[22]),"=r"(a[23]),"=r"(a[24]),"=r"(a[25]),"=r"(a[26]),"=r"(a[27]),"=r"(a[28]),"=r"(a[29]),"=r"(a[30]),"=r"(a[31]) : "r"(base));
      asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                   : "=r"(b[0]),"=r"(b[1]),"=r"(b[2]),"=r"(b[3]),"=r"(b[4]),"=r"(b[5]),"=r"(b[6]),"=r"(b[7]),"=r"(b[8]),"=r"(b[9]),"=r"(b[10]),"=r"(b[11]),"=r"(b[12]),"=r"(b[13]),"=r"(b[14]),"=r"(b[15]),"=r"(b[16]),"=r"(b[17]),"=r"(b[18]),"=r"(b[19]),"=r"(b[20]),"=r"(b[21]),"=r"(b[22]),"=r"(b[23]),"=r"(b[24]),"=r"(b[25]),"=r"(b[26]),"=r"(b[27]),"=r"(b[28]),"=r"(b[29]),"=r"(b[30]),"=r"(b[31]) : "r"(base + (16u << 16)));
      ptx::tmem_ld_wait();
      acc += fold(a) ^ fold(b);
    } else if (mode == 4) {     // 32x32b.x8 four times (narrow loads)
      uint32_t a[8], b[8], c[8], d[8];
#define LD8(arr, addr) asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(arr[0]),"=r"(arr[1]),"=r"(arr[2]),"=r"(arr[3]),"=r"(arr[4]),"=r"(arr[5]),"=r"(arr[6]),"=r"(arr[7]) : "r"(addr))
      LD8(a, base); LD8(b, base + 8); LD8(c, base + 16); LD8(d, base + 24);
      ptx::tmem_ld_wait();
      LD8(a, base + 32); LD8(b, base + 40); LD8(c, base + 48); LD8(d, base + 56);
      ptx::tmem_ld_wait();
      acc += fold(a) ^ fold(b) ^ fold(c) ^ fold(d);
    } else if (mode == 5) {     // 32x32b.x16 twice per wait, two waits
      uint32_t a[16], b[16];
      ptx::tmem_ld_32x16(base, a); ptx::tmem_ld_32x16(base + 16, b);
      ptx::tmem_ld_wait();
      acc += fold(a) ^ fold(b);
      ptx::tmem_ld_32x16(base + 32, a); ptx::tmem_ld_32x16(base + 48, b);
      ptx::tmem_ld_wait();
      acc += fold(a) ^ fold(b);
    } else if (mode == 6) {     // 32x32b.x8, all 8 (64 columns) in flight before one wait
      uint32_t a[8], b[8], c[8], d[8], e[8], f[8], g[8], h[8];
      LD8(a, base); LD8(b, base + 8); LD8(c, base + 16); LD8(d, base + 24);
      LD8(e, base + 32); LD8(f, base + 40); LD8(g, base + 48); LD8(h, base + 56);
      ptx::tmem_ld_wait();
      acc += fold(a) ^ fold(b) ^ fold(c) ^ fold(d) ^ fold(e) ^ fold(f) ^ fold(g) ^ fold(h);
    } else if (mode == 7) {     // 32x32b.x4, 16 of them (64 columns) before one wait
      uint32_t a[16][4];
#pragma unroll
      for (int q = 0; q < 16; ++q)
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(a[q][0]),"=r"(a[q][1]),"=r"(a[q][2]),"=r"(a[q][3]) : "r"(base + q * 4));
      ptx::tmem_ld_wait();
#pragma unroll
      for (int q = 0; q < 16; ++q) acc += fold(a[q]);
    } else {                    // stores
      r[0] = acc + it;
      ptx::tmem_st_32x32(base, r);
      ptx::tmem_st_32x32(base + 32, r);
      ptx::tmem_st_wait();
    }
  }
  const long long t1 = clock64();
  if ((threadIdx.x & 31) == 0) out[warp] = t1 - t0;
  sink[threadIdx.x] = acc;
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(slot, 512);
}

int main() {
  long long* out; uint32_t* sink;
  cudaMalloc(&out, 16 * sizeof(long long)); cudaMalloc(&sink, 512 * 4);
  const int iters = 2000;
  for (int mode : {0, 2, 3, 4, 5, 6, 7, 1})
    for (int warps : {1, 4, 8, 16}) {
      tmem_bw<<<1, warps * 32, 0>>>(iters, mode, out, sink);
      long long h[16];
      cudaError_t e = cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      long long mx = 0;
      for (int w = 0; w < warps; ++w) mx = h[w] > mx ? h[w] : mx;
      const double bytes = (double)warps * iters * (mode == 2 ? 4 : 2) * 32 * 32 * 4;
      printf("mode %d %s warps=%2d  %8lld clk  %.1f B/clk/SM  (%.1f clk per 32x32b.x32)\n", mode, mode == 1 ? "st" : "ld", warps, mx, bytes / mx,
             (double)mx / (iters * 2));
    }
  return 0;
}
