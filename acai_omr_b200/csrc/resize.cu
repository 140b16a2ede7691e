// Input pipeline, the step before K1: DynamicResize (utils/utils.py:334-370) = torchvision F.resize(BICUBIC, antialias=True)
// (+ optional centre crop, clamp to [0, 1]) for a ragged batch of single-channel images, straight from uint8 scans
// (v2.ToImage + v2.ToDtype(float32, scale=True) = x / 255 is folded into the first pass) or from fp32 tensors.
//
// The arithmetic is torch's antialiased separable resize (aten/native/cpu/UpSampleKernel.cpp, HelperInterpCubic::aa_filter,
// _compute_indices_min_size_weights_aa), restated:
//     scale = in / out,  support = 2 * max(scale, 1),  centre = scale * (i + 0.5),  invscale = 1 / max(scale, 1)
//     xmin = max(0, int(centre - support + 0.5)),  xsize = min(in, int(centre + support + 0.5)) - xmin
//     w_j = cubic_{a = -0.5}((j + xmin - centre + 0.5) * invscale) / sum_j w_j
// width first (fp32 intermediate), then height.  Two launches for the whole batch: thread = one output pixel of a pass,
// neighbouring threads read overlapping input windows (L1 hits), weights are recomputed per thread (a handful of FMAs per
// tap; a table would cost a third launch).  HBM-bound by a wide margin: a 1 MB scan is ~10 us of traffic.
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

}  // namespace acai

extern "C" int acai_dynamic_resize(const acai_resize_desc* descs, int32_t num_images, int32_t max_src_h, int32_t max_out_h,
                                   int32_t max_out_w, void* stream) {
  ACAI_CHECK_ARG(descs && num_images > 0 && num_images <= 65535, "acai_dynamic_resize: bad descriptor table");
  ACAI_CHECK_ARG(max_src_h > 0 && max_src_h <= 65535 && max_out_h > 0 && max_out_h <= 65535 && max_out_w > 0,
                 "acai_dynamic_resize: bad maxima (rows are grid.y: <= 65535)");
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned gx = (unsigned)((max_out_w + 127) / 128);
  acai::resize_rows_kernel<<<dim3(gx, (unsigned)max_src_h, (unsigned)num_images), 128, 0, st>>>(descs);
  int rc = acai::launch_status("resize_rows");
  if (rc) return rc;
  acai::resize_cols_kernel<<<dim3(gx, (unsigned)max_out_h, (unsigned)num_images), 128, 0, st>>>(descs);
  return acai::launch_status("resize_cols");
}
