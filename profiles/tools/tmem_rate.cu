// Probe (not product code): sustained tcgen05.ld rate of one SM. W warps (4, 8 or 16) read their own 32-lane quarter of
// TMEM over and over with 32x32b.x32 loads; prints bytes per clock per SM. Decides whether the push scan's read-out of
// every (position, prototype) distance is what bounds it.
#include <cstdio>
#include <cstdlib>
#include "../../scaleprotoseg_b200/csrc/tc_common.cuh"
using namespace sps::tc;

template <int DEPTH>
__global__ void __launch_bounds__(512) probe(int iters, long long* out, float* sink) {
  __shared__ uint32_t slot;
  if (threadIdx.x < 32) tmem_alloc(smem_u32(&slot), 512);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tbase = slot;
  const int warp = threadIdx.x >> 5;
  const uint32_t lane_base = (uint32_t)(warp & 3) * 32u;
  float acc = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int c = 0; c < 512; c += 32 * DEPTH) {
      uint32_t r[DEPTH][32];
#pragma unroll
      for (int d = 0; d < DEPTH; ++d) tmem_ld32(tbase + (lane_base << 16) + (uint32_t)(c + 32 * d), r[d]);
      tmem_ld_wait();
#pragma unroll
      for (int d = 0; d < DEPTH; ++d)
#pragma unroll
        for (int j = 0; j < 32; j += 8) acc += __uint_as_float(r[d][j]);
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  if (acc == 123.456f) sink[0] = acc;
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tbase, 512);
}

int main() {
  long long* out; float* sink;
  cudaMalloc(&out, 8 * 148); cudaMalloc(&sink, 4);
  const int iters = 200;
  for (int warps : {4, 8, 16}) {
    for (int depth : {1, 2}) {
      for (int rep = 0; rep < 2; ++rep) {
        if (depth == 1) probe<1><<<148, warps * 32>>>(iters, out, sink); else probe<2><<<148, warps * 32>>>(iters, out, sink);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      }
      long long h[148]; cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
      const double bytes = (double)warps * iters * 512.0 * 32 * 4;   // every warp reads 32 lanes x 512 columns x 4 B per iteration
      printf("warps/CTA %2d, loads in flight per warp %d: %lld clk -> %.1f B/clk/SM (%.1f B/clk per warp)\n", warps, depth, h[0],
             bytes / h[0], bytes / h[0] / warps);
    }
  }
  return 0;
}
