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

}  // namespace

int launch_upsample_argext(const float* maps, int B, int C, int h, int w, int H, int W, int argmin, int32_t* out_idx,
                           float* out_val, cudaStream_t st) {
  if (B == 0 || H == 0 || W == 0) return SPS_OK;
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
