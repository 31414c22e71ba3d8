// Does MUFU work overlap with FP32 issue on one SM sub-partition?  (not product code)
// W warps per sub-partition run NM MUFU + NF FFMA per iteration on independent registers.
#include <cstdio>
#include <cuda_runtime.h>
template <int NM, int NF>
__global__ void k(float* out, int iters, long long* clk) {
  float m[8], f[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { m[i] = 1.0f + threadIdx.x * 1e-3f + i; f[i] = 0.5f + i; }
  const float a = 1.0001f, b = 0.001f;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int i = 0; i < NM; ++i) asm volatile("tanh.approx.f32 %0, %0;" : "+f"(m[(r * NM + i) & 7]));
#pragma unroll
      for (int i = 0; i < NF; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[(r * NF + i) & 7]) : "f"(a), "f"(b));
    }
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += m[i] + f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = t1 - t0;
}
template <int NM, int NF> void run(int warps) {
  float* o; long long* c; cudaMalloc(&o, 4 * 1024 * 148); cudaMalloc(&c, 8);
  const int iters = 2000;
  k<NM, NF><<<148, warps * 32>>>(o, iters, c); k<NM, NF><<<148, warps * 32>>>(o, iters, c);
  cudaDeviceSynchronize();
  long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
  printf("warps/CTA %2d  per 8 rounds: %d MUFU + %d FFMA per round -> %.1f clk per round (per warp)\n", warps, NM, NF, (double)h / iters / 8);
  cudaFree(o); cudaFree(c);
}
int main() {
  for (int w : {4, 8, 16}) {
    run<1, 0>(w); run<0, 6>(w); run<1, 6>(w); run<2, 12>(w); run<1, 12>(w); run<2, 0>(w);
    printf("\n");
  }
  return 0;
}
