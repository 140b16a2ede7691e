// Input pipeline, the step before K1: DynamicResize (utils/utils.py:334-370) = torchvision F.resize(BICUBIC, antialias=True)
// (+ optional centre crop, clamp to [0, 1]) for a ragged batch of single-channel images, straight from uint8 scans
// (v2.ToImage + v2.ToDtype(float32, scale=True) = x / 255 is folded into the first pass) or from fp32 tensors.
//
// The arithmetic is torch's antialiased separable resize (aten/native/cpu/UpSampleKernel.cpp, HelperInterpCubic::aa_filter,
// _compute_indices_min_size_weights_aa), restated:
//     scale = in / out,  support = 2 * max(scale, 1),  centre = scale * (i + 0.5),  invscale = 1 / max(scale, 1)
//     xmin = max(0, int(centre - support + 0.5)),  xsize = min(in, int(centre + support + 0.5)) - xmin
//     w_j = cubic_{a = -0.5}((j + xmin - centre + 0.5) * invscale) / sum_j w_j
// width first (fp32 intermediate), then height.  Two launches for the whole ragged batch.  The kernels that normally run
// tabulate the normalised weights of a CTA's 128 output columns (16 output rows) in shared memory and reuse them over a
// strip of 16 rows (128 columns): 64 uint8 scans of 0.5-1 MB -> encoder inputs in 140 + 92 us (the per-thread-weight
// kernels, kept for windows too wide for shared memory, take 388 + 212 us).  Instruction bound, not HBM bound: a 1 MB scan
// is ~10 us of traffic; at 90 k systems/s the pass is 1-2 % of an encoder step.
#include "common.cuh"

namespace acai {

__device__ __forceinline__ float cubic_aa(float x) {          // Keys cubic, a = -0.5
  x = fabsf(x);
  const float a = -0.5f;
  if (x < 1.0f) return ((a + 2.0f) * x - (a + 3.0f)) * x * x + 1.0f;
  if (x < 2.0f) return (((x - 5.0f) * x + 8.0f) * x - 4.0f) * a;
  return 0.0f;
}

struct AaWindow { int xmin, xsize; float center, invscale; };

__device__ __forceinline__ AaWindow aa_window(int i, int in_size, int out_size) {
  const float scale = (float)in_size / (float)out_size;
  const float support = scale >= 1.0f ? 2.0f * scale : 2.0f;
  AaWindow w;
  w.center = scale * ((float)i + 0.5f);
  w.invscale = scale >= 1.0f ? 1.0f / scale : 1.0f;
  const int lo = (int)(w.center - support + 0.5f), hi = (int)(w.center + support + 0.5f);
  w.xmin = lo > 0 ? lo : 0;
  w.xsize = (hi < in_size ? hi : in_size) - w.xmin;
  return w;
}

// pass 1: tmp[y, x] = sum_j w_j src[y, xmin + j]   for every source row y and every target column x (before the crop's
// column window is applied: only the columns the crop keeps are computed)
__global__ void __launch_bounds__(128)
resize_rows_kernel(const acai_resize_desc* __restrict__ descs) {
  const acai_resize_desc d = descs[blockIdx.z];
  const int y = blockIdx.y, xo = blockIdx.x * blockDim.x + threadIdx.x;
  if (y >= d.src_h || xo >= d.out_w) return;
  const AaWindow w = aa_window(xo + d.crop_left, d.src_w, d.dst_w);
  float acc = 0.0f, tot = 0.0f;
  if (d.src_dtype == ACAI_RESIZE_U8) {
    const uint8_t* row = static_cast<const uint8_t*>(d.src) + (int64_t)y * d.src_stride + w.xmin;
    for (int j = 0; j < w.xsize; ++j) {
      const float wt = cubic_aa(((float)(j + w.xmin) - w.center + 0.5f) * w.invscale);
      acc = fmaf(wt, (float)row[j] * (1.0f / 255.0f), acc);
      tot += wt;
    }
  } else {
    const float* row = static_cast<const float*>(d.src) + (int64_t)y * d.src_stride + w.xmin;
    for (int j = 0; j < w.xsize; ++j) {
      const float wt = cubic_aa(((float)(j + w.xmin) - w.center + 0.5f) * w.invscale);
      acc = fmaf(wt, row[j], acc);
      tot += wt;
    }
  }
  d.tmp[(int64_t)y * d.out_w + xo] = acc / tot;
}

// pass 2: dst[y, x] = clamp(sum_k w_k tmp[ymin + k, x], 0, 1)
__global__ void __launch_bounds__(128)
resize_cols_kernel(const acai_resize_desc* __restrict__ descs) {
  const acai_resize_desc d = descs[blockIdx.z];
  const int yo = blockIdx.y, xo = blockIdx.x * blockDim.x + threadIdx.x;
  if (yo >= d.out_h || xo >= d.out_w) return;
  const AaWindow w = aa_window(yo + d.crop_top, d.src_h, d.dst_h);
  const float* col = d.tmp + (int64_t)w.xmin * d.out_w + xo;
  float acc = 0.0f, tot = 0.0f;
  for (int k = 0; k < w.xsize; ++k) {
    const float wt = cubic_aa(((float)(k + w.xmin) - w.center + 0.5f) * w.invscale);
    acc = fmaf(wt, col[(int64_t)k * d.out_w], acc);
    tot += wt;
  }
  d.dst[(int64_t)yo * d.dst_stride + xo] = fminf(fmaxf(acc / tot, 0.0f), 1.0f);
}

// ---- table variants (the ones that normally run): a CTA computes the NORMALISED weights of its output coordinates once
// (as torch does: w_j / sum w) into shared memory and reuses them for a strip of rows, so the inner loop is a shared-memory
// weight, a pixel and an FMA per tap instead of the cubic polynomial per tap and pixel.  `taps` = table width (host: the
// widest window of the batch); the kernels above remain the path for windows too wide for shared memory.
constexpr int RS_COLS = 128, RS_ROWS = 16;

__global__ void __launch_bounds__(RS_COLS)
resize_rows_table_kernel(const acai_resize_desc* __restrict__ descs, int taps) {
  extern __shared__ float rs_w[];                        // [taps][RS_COLS]
  const acai_resize_desc d = descs[blockIdx.z];
  const int y0 = blockIdx.y * RS_ROWS, xo = blockIdx.x * RS_COLS + threadIdx.x;
  if (y0 >= d.src_h || blockIdx.x * RS_COLS >= d.out_w) return;          // CTA-uniform
  const bool live = xo < d.out_w;
  const AaWindow w = aa_window(live ? xo + d.crop_left : d.crop_left, d.src_w, d.dst_w);
  float tot = 0.0f;
  for (int j = 0; j < w.xsize; ++j) {
    const float wt = cubic_aa(((float)(j + w.xmin) - w.center + 0.5f) * w.invscale);
    rs_w[j * RS_COLS + threadIdx.x] = wt;
    tot += wt;
  }
  const float inv = 1.0f / tot;
  for (int j = 0; j < w.xsize; ++j) rs_w[j * RS_COLS + threadIdx.x] *= inv;   // (own column only: no barrier needed)
  if (!live) return;
  const int y1 = y0 + RS_ROWS < d.src_h ? y0 + RS_ROWS : d.src_h;
  if (d.src_dtype == ACAI_RESIZE_U8) {
    for (int y = y0; y < y1; ++y) {
      const uint8_t* row = static_cast<const uint8_t*>(d.src) + (int64_t)y * d.src_stride + w.xmin;
      float acc = 0.0f;
      for (int j = 0; j < w.xsize; ++j) acc = fmaf(rs_w[j * RS_COLS + threadIdx.x], (float)row[j] * (1.0f / 255.0f), acc);
      d.tmp[(int64_t)y * d.out_w + xo] = acc;
    }
  } else {
    for (int y = y0; y < y1; ++y) {
      const float* row = static_cast<const float*>(d.src) + (int64_t)y * d.src_stride + w.xmin;
      float acc = 0.0f;
      for (int j = 0; j < w.xsize; ++j) acc = fmaf(rs_w[j * RS_COLS + threadIdx.x], row[j], acc);
      d.tmp[(int64_t)y * d.out_w + xo] = acc;
    }
  }
}

__global__ void __launch_bounds__(RS_COLS)
resize_cols_table_kernel(const acai_resize_desc* __restrict__ descs, int taps) {
  extern __shared__ float rs_w[];                        // [RS_ROWS][taps] weights, then [RS_ROWS] ymin, [RS_ROWS] ysize
  int* rs_min = reinterpret_cast<int*>(rs_w + RS_ROWS * taps);
  int* rs_size = rs_min + RS_ROWS;
  const acai_resize_desc d = descs[blockIdx.z];
  const int y0 = blockIdx.y * RS_ROWS, xo = blockIdx.x * RS_COLS + threadIdx.x;
  if (y0 >= d.out_h || blockIdx.x * RS_COLS >= d.out_w) return;           // CTA-uniform
  if (threadIdx.x < RS_ROWS && y0 + threadIdx.x < d.out_h) {              // one thread per output row of the strip
    const AaWindow w = aa_window(y0 + threadIdx.x + d.crop_top, d.src_h, d.dst_h);
    float* wr = rs_w + threadIdx.x * taps;
    float tot = 0.0f;
    for (int k = 0; k < w.xsize; ++k) {
      const float wt = cubic_aa(((float)(k + w.xmin) - w.center + 0.5f) * w.invscale);
      wr[k] = wt;
      tot += wt;
    }
    const float inv = 1.0f / tot;
    for (int k = 0; k < w.xsize; ++k) wr[k] *= inv;
    rs_min[threadIdx.x] = w.xmin; rs_size[threadIdx.x] = w.xsize;
  }
  __syncthreads();
  if (xo >= d.out_w) return;
  const int y1 = y0 + RS_ROWS < d.out_h ? y0 + RS_ROWS : d.out_h;
  for (int yo = y0; yo < y1; ++yo) {
    const float* wr = rs_w + (yo - y0) * taps;
    const int n = rs_size[yo - y0];
    const float* col = d.tmp + (int64_t)rs_min[yo - y0] * d.out_w + xo;
    float acc = 0.0f;
    for (int k = 0; k < n; ++k) acc = fmaf(wr[k], col[(int64_t)k * d.out_w], acc);
    d.dst[(int64_t)yo * d.dst_stride + xo] = fminf(fmaxf(acc, 0.0f), 1.0f);
  }
}

}  // namespace acai

extern "C" int acai_dynamic_resize(const acai_resize_desc* descs, int32_t num_images, int32_t max_src_h, int32_t max_out_h,
                                   int32_t max_out_w, int32_t max_taps, void* stream) {
  ACAI_CHECK_ARG(descs && num_images > 0 && num_images <= 65535, "acai_dynamic_resize: bad descriptor table");
  ACAI_CHECK_ARG(max_src_h > 0 && max_src_h <= 65535 && max_out_h > 0 && max_out_h <= 65535 && max_out_w > 0,
                 "acai_dynamic_resize: bad maxima (rows are grid.y: <= 65535)");
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned gx = (unsigned)((max_out_w + 127) / 128);
  using namespace acai;
  // max_taps: the widest filter window of the batch on either axis (host: 2 * ceil(2 * max(in / out, 1)) + 2 bounds it);
  // 0 or too wide for 48 KB of shared memory -> the per-thread kernels
  const size_t smem_rows = (size_t)max_taps * RS_COLS * sizeof(float);
  const size_t smem_cols = (size_t)RS_ROWS * max_taps * sizeof(float) + 2 * RS_ROWS * sizeof(int);
  if (max_taps > 0 && smem_rows <= 48 * 1024 && smem_cols <= 48 * 1024) {
    resize_rows_table_kernel<<<dim3(gx, (unsigned)((max_src_h + RS_ROWS - 1) / RS_ROWS), (unsigned)num_images), RS_COLS, smem_rows, st>>>(descs, max_taps);
    int rc = launch_status("resize_rows_table");
    if (rc) return rc;
    resize_cols_table_kernel<<<dim3(gx, (unsigned)((max_out_h + RS_ROWS - 1) / RS_ROWS), (unsigned)num_images), RS_COLS, smem_cols, st>>>(descs, max_taps);
    return launch_status("resize_cols_table");
  }
  resize_rows_kernel<<<dim3(gx, (unsigned)max_src_h, (unsigned)num_images), 128, 0, st>>>(descs);
  int rc = launch_status("resize_rows");
  if (rc) return rc;
  resize_cols_kernel<<<dim3(gx, (unsigned)max_out_h, (unsigned)num_images), 128, 0, st>>>(descs);
  return launch_status("resize_cols");
}
