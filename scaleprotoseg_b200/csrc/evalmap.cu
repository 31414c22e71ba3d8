// Evaluation epilogue (segmentation/eval_valid_multiscale.py:229-234): bilinear upsampling of a
// low-resolution map (logits (K,h,w) or distances (P,h,w)) to the annotation size followed by argmax / argmin over the
// channel axis.  The reference materialises the upsampled map (P x 1024 x 2048 fp32 = 6.4 GB at P = 760) and reduces it
// afterwards; here every output pixel interpolates channel by channel in registers and only the winning index (and,
// optionally, its value) is written: the source map stays L2-resident and HBM sees H*W*4 (or 8) bytes.
#include "common.cuh"

namespace sps {

namespace {

// Source coordinate of F.interpolate(mode="bilinear", align_corners=False): src = (dst + 0.5) * in/out - 0.5, clamped
// at 0; upper neighbour clamped to the last row / column.
__device__ __forceinline__ void source_of(int dst, float scale, int in, int* i0, int* i1, float* lam) {
  float src = ((float)dst + 0.5f) * scale - 0.5f;
  src = src < 0.0f ? 0.0f : src;
  int lo = (int)src;
  lo = lo > in - 1 ? in - 1 : lo;
  *i0 = lo;
  *i1 = lo + (lo < in - 1 ? 1 : 0);
  *lam = src - (float)lo;
}

template <bool ARGMIN>
__global__ void upsample_argext_kernel(const float* __restrict__ maps, int C, int h, int w, int H, int W,
                                       int32_t* __restrict__ out_idx, float* __restrict__ out_val) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  const int b = blockIdx.z;
  if (x >= W || y >= H) return;
  int y0, y1, x0, x1;
  float ly, lx;
  source_of(y, (float)h / (float)H, h, &y0, &y1, &ly);
  source_of(x, (float)w / (float)W, w, &x0, &x1, &lx);
  const float hy = 1.0f - ly, hx = 1.0f - lx;
  const float* m = maps + (size_t)b * C * h * w;
  const int o00 = y0 * w + x0, o01 = y0 * w + x1, o10 = y1 * w + x0, o11 = y1 * w + x1;
  float best = ARGMIN ? INFINITY : -INFINITY;
  int arg = 0;
  const size_t hw = (size_t)h * w;
#pragma unroll 4
  for (int c = 0; c < C; ++c) {
    const float* mc = m + (size_t)c * hw;
    // same association as the reference's kernel: rows first (hx/lx), then the two rows blended with hy/ly
    const float v = hy * (hx * mc[o00] + lx * mc[o01]) + ly * (hx * mc[o10] + lx * mc[o11]);
    const bool better = ARGMIN ? (v < best) : (v > best);   // strict: first channel wins ties, as torch.argmax/argmin
    if (better) {
      best = v;
      arg = c;
    }
  }
  const size_t o = ((size_t)b * H + y) * W + x;
  out_idx[o] = arg;
  if (out_val != nullptr) out_val[o] = best;
}


// Strip variant for upsampling factors of 3.5 and more along x (the evaluation's 8x): one thread owns 8 consecutive
// output pixels of a row.  Their bilinear footprints fall inside 4 consecutive source columns, so a channel costs
// 8 loads (2 rows x 4 columns, blended vertically once) instead of 32, and the per-pixel column picks are register
// selects fixed before the channel loop.  The per-pixel kernel above is LSU-bound on exactly those loads.
constexpr int STRIP = 8;
constexpr int STRIP_COLS = 4;

__device__ __forceinline__ float pick4(const float (&v)[STRIP_COLS], int r) {
  const float lo = r == 0 ? v[0] : v[1], hi = r == 2 ? v[2] : v[3];
  return r < 2 ? lo : hi;
}

template <bool ARGMIN>
__global__ void upsample_argext_strip_kernel(const float* __restrict__ maps, int C, int h, int w, int H, int W,
                                             int32_t* __restrict__ out_idx, float* __restrict__ out_val) {
  const int xs = (blockIdx.x * blockDim.x + threadIdx.x) * STRIP;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  const int b = blockIdx.z;
  if (xs >= W || y >= H) return;
  int y0, y1;
  float ly;
  source_of(y, (float)h / (float)H, h, &y0, &y1, &ly);
  const float hy = 1.0f - ly;
  int r0[STRIP], r1[STRIP];
  float lx[STRIP], hx[STRIP];
  int xb = 0;
#pragma unroll
  for (int i = 0; i < STRIP; ++i) {
    const int x = xs + i < W ? xs + i : W - 1;
    int x0, x1;
    source_of(x, (float)w / (float)W, w, &x0, &x1, &lx[i]);
    if (i == 0) xb = x0;
    r0[i] = x0 - xb;
    r1[i] = x1 - xb;
    hx[i] = 1.0f - lx[i];
  }
  int col[STRIP_COLS];
#pragma unroll
  for (int c = 0; c < STRIP_COLS; ++c) col[c] = xb + c < w ? xb + c : w - 1;
  float best[STRIP];
  int arg[STRIP];
#pragma unroll
  for (int i = 0; i < STRIP; ++i) {
    best[i] = ARGMIN ? INFINITY : -INFINITY;
    arg[i] = 0;
  }
  const float* m0 = maps + (size_t)b * C * h * w + (size_t)y0 * w;
  const float* m1 = maps + (size_t)b * C * h * w + (size_t)y1 * w;
  const size_t hw = (size_t)h * w;
#pragma unroll 2
  for (int c = 0; c < C; ++c) {
    float v[STRIP_COLS];
#pragma unroll
    for (int k = 0; k < STRIP_COLS; ++k) v[k] = hy * m0[(size_t)c * hw + col[k]] + ly * m1[(size_t)c * hw + col[k]];
#pragma unroll
    for (int i = 0; i < STRIP; ++i) {
      const float val = hx[i] * pick4(v, r0[i]) + lx[i] * pick4(v, r1[i]);
      const bool better = ARGMIN ? (val < best[i]) : (val > best[i]);
      if (better) {
        best[i] = val;
        arg[i] = c;
      }
    }
  }
  const size_t o = ((size_t)b * H + y) * W + xs;
#pragma unroll
  for (int i = 0; i < STRIP; ++i)
    if (xs + i < W) {
      out_idx[o + i] = arg[i];
      if (out_val != nullptr) out_val[o + i] = best[i];
    }
}

}  // namespace

int launch_upsample_argext(const float* maps, int B, int C, int h, int w, int H, int W, int argmin, int32_t* out_idx,
                           float* out_val, cudaStream_t st) {
  if (B == 0 || H == 0 || W == 0) return SPS_OK;
  // 8 outputs per thread need 7 * w/W + 2 <= 4 source columns
  if (7.0f * (float)w / (float)W + 2.0f <= (float)STRIP_COLS - 1e-3f) {
    dim3 block(32, 4), grid((W + 32 * STRIP - 1) / (32 * STRIP), (H + 3) / 4, B);
    if (argmin)
      upsample_argext_strip_kernel<true><<<grid, block, 0, st>>>(maps, C, h, w, H, W, out_idx, out_val);
    else
      upsample_argext_strip_kernel<false><<<grid, block, 0, st>>>(maps, C, h, w, H, W, out_idx, out_val);
    count_launch(1);
    SPS_CUDA_TRY(cudaGetLastError());
    return SPS_OK;
  }
  dim3 block(32, 8), grid((W + 31) / 32, (H + 7) / 8, B);
  if (argmin)
    upsample_argext_kernel<true><<<grid, block, 0, st>>>(maps, C, h, w, H, W, out_idx, out_val);
  else
    upsample_argext_kernel<false><<<grid, block, 0, st>>>(maps, C, h, w, H, W, out_idx, out_val);
  count_launch(1);
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

}  // namespace sps
