// Probe (not product code): TMA boxes out of an fp32 NCHW map with an ODD number of pixels per channel (row pitch not a
// multiple of 16 bytes), through four tensor maps, one per channel residue mod 4 (tma.cuh: make_tmap_nchw_f32).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../../scaleprotoseg_b200/csrc/tma.cuh"
using namespace sps;
using namespace sps::tc;

__global__ void __launch_bounds__(128) probe(const __grid_constant__ TmapNchwF32 tm, int rows, int pix0, int row0, int b, float* out) {
  extern __shared__ unsigned char raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  float* stg = reinterpret_cast<float*>(raw + (base - smem_u32(raw)));
  __shared__ __align__(8) unsigned long long barmem;
  const uint32_t bar = smem_u32(&barmem);
  if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_fence_init(); }
  fence_proxy_async();
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(bar, (uint32_t)(4 * rows * 132 * 4));
    for (int r = 0; r < 4; ++r) tma_load_3d(base + r * rows * 528, &tm.map[r], pix0, row0, b, bar);
  }
  mbar_wait(bar, 0);
  for (int c = 0; c < 4 * rows; ++c) out[c * 128 + threadIdx.x] = stg[((c & 3) * rows + (c >> 2)) * 132 + threadIdx.x + tm.shift[c & 3]];
}

int main() {
  const int B = 3, C = 256, H = 13, W = 21, HW = H * W, D = 64;   // 273 pixels: tiles 0, 1 full, tile 2 partial
  std::vector<float> h((size_t)B * C * HW);
  for (size_t i = 0; i < h.size(); ++i) h[i] = (float)i;
  float *d, *o;
  cudaMalloc(&d, h.size() * 4 + 64); cudaMalloc(&o, 64 * 128 * 4);
  for (int off : {0, 1}) {   // off = 1: a map that is only 4-byte aligned
    float* z = d + off;
    cudaMemcpy(z, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    TmapNchwF32 tm;
    int rc = make_tmap_nchw_f32(&tm, z, B, C, HW, 128, D / 4);
    printf("make rc %d shifts %d %d %d %d\n", rc, tm.shift[0], tm.shift[1], tm.shift[2], tm.shift[3]);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 40000);
    for (int sc : {0, 3}) for (int b : {0, 2}) for (int pix0 : {0, 128, 256}) {
      probe<<<1, 128, 40000>>>(tm, D / 4, pix0, sc * D / 4, b, o);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
      std::vector<float> r(64 * 128);
      cudaMemcpy(r.data(), o, r.size() * 4, cudaMemcpyDeviceToHost);
      int bad = 0;
      for (int c = 0; c < 64; ++c)
        for (int m = 0; m < 128; ++m) {
          const float want = pix0 + m < HW ? h[((size_t)b * C + sc * 64 + c) * HW + pix0 + m] : 0.0f;
          if (pix0 + m < HW && r[c * 128 + m] != want && bad++ < 5) printf("  off %d sc %d b %d c %d m %d: got %g want %g\n", off, sc, b, c, m, r[c * 128 + m], want);
        }
      printf("off %d scale %d map %d pix0 %d: %d mismatches\n", off, sc, b, pix0, bad);
    }
  }
  return 0;
}
