// Probe (not product code): tcgen05.mma with the A operand in TMEM (written there by tcgen05.st, row m in lane m, K
// elements 2j, 2j+1 in 32-bit column j) and B in shared memory (K-major, 128-byte swizzle).  Checks the product against
// the CPU and times a stream of M128 x N x K16 MMAs for N = 64 / 128 against the same MMAs with A in shared memory.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_fp16.h>
#include "../../scaleprotoseg_b200/csrc/tc_common.cuh"
using namespace sps::tc;

__device__ __forceinline__ void mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0],"
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16,"
      " %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
        "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
        "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}

// A: 128 x 64 fp16 (row-major), B: N x 64 fp16 (row-major = K-major), D: 128 x N fp32
template <bool TS>
__global__ void __launch_bounds__(160) probe(int N, int reps, const __half* A, const __half* B, float* D, long long* clk) {
  extern __shared__ unsigned char raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  unsigned char* g = raw + (base - smem_u32(raw));
  unsigned char* sa = g;            // 128 rows x 128 B (SS mode)
  unsigned char* sb = g + 16384;    // N rows x 128 B
  __shared__ uint32_t slot;
  __shared__ __align__(8) unsigned long long barmem;
  const uint32_t bar = smem_u32(&barmem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 128 * 64; i += blockDim.x) {
    const int m = i / 64, k = i % 64;
    *reinterpret_cast<__half*>(sa + sw128_off(m, k)) = A[i];
  }
  for (int i = threadIdx.x; i < N * 64; i += blockDim.x) {
    const int n = i / 64, k = i % 64;
    *reinterpret_cast<__half*>(sb + sw128_off(n, k)) = B[i];
  }
  if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_fence_init(); }
  fence_proxy_async();
  if (warp == 4) tmem_alloc(smem_u32(&slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  const uint32_t TA = 384;          // A at columns [384, 416)
  if (warp < 4) {
    const int m = warp * 32 + lane;
    uint32_t r[32];
    for (int j = 0; j < 32; ++j) r[j] = reinterpret_cast<const uint32_t*>(A)[m * 32 + j];   // halfs 2j, 2j+1 of row m
    tmem_st32(tmem + ((uint32_t)(warp * 32) << 16) + TA, r);
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 4) {
    const uint32_t ID = idesc_f16(0 /*f16*/, 128, N);
    const uint64_t db0 = desc_sw128(smem_u32(sb)), da0 = desc_sw128(smem_u32(sa));
    const uint32_t ta0 = tmem + TA;
    if (TS) mma_f16_ts(tmem, ta0, db0, ID, 0u); else mma_f16(tmem, da0, db0, ID, 0u);
#pragma unroll
    for (int k = 1; k < 4; ++k) {
      if (TS) mma_f16_ts(tmem, ta0 + 8 * k, db0 + 2 * k, ID, 1u); else mma_f16(tmem, da0 + 2 * k, db0 + 2 * k, ID, 1u);
    }
    const long long t0 = clock64();
    for (int r = 1; r < reps; ++r) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (TS) mma_f16_ts(tmem + 256, ta0 + 8 * k, db0 + 2 * k, ID, 1u); else mma_f16(tmem + 256, da0 + 2 * k, db0 + 2 * k, ID, 1u);
      }
    }
    mma_commit(bar);
    __syncwarp();
    mbar_wait(bar, 0);
    const long long t1 = clock64();
    if (lane == 0) clk[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp < 4) {
    const int m = warp * 32 + lane;
    for (int c = 0; c < N; c += 32) {
      uint32_t r[32];
      tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c, r);
      tmem_ld_wait();
      for (int j = 0; j < 32; ++j) D[m * N + c + j] = __uint_as_float(r[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main() {
  std::vector<__half> hA(128 * 64), hB(256 * 64);
  srand(1);
  for (auto& v : hA) v = __float2half((rand() % 200 - 100) / 64.0f);
  for (auto& v : hB) v = __float2half((rand() % 200 - 100) / 64.0f);
  __half *dA, *dB; float* dD; long long* dC;
  cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, 128 * 256 * 4); cudaMalloc(&dC, 8);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536); cudaFuncSetAttribute(probe<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  for (int N : {64, 128}) {
    for (int ts : {0, 1}) {
      if (ts) probe<true><<<1, 160, 65536>>>(N, 1, dA, dB, dD, dC); else probe<false><<<1, 160, 65536>>>(N, 1, dA, dB, dD, dC);
      if (cudaDeviceSynchronize() != cudaSuccess) { printf("N=%d ts=%d: %s\n", N, ts, cudaGetErrorString(cudaGetLastError())); return 1; }
      std::vector<float> hD(128 * N);
      cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
      double maxerr = 0;
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
          double ref = 0;
          for (int k = 0; k < 64; ++k) ref += (double)__half2float(hA[m * 64 + k]) * __half2float(hB[n * 64 + k]);
          maxerr = fmax(maxerr, fabs(ref - hD[m * N + n]));
        }
      if (ts) probe<true><<<1, 160, 65536>>>(N, 65, dA, dB, dD, dC); else probe<false><<<1, 160, 65536>>>(N, 65, dA, dB, dD, dC);
      cudaDeviceSynchronize();
      long long c; cudaMemcpy(&c, dC, 8, cudaMemcpyDeviceToHost);
      printf("M128 N%-3d K16, A in %s: max |err| %.3g, %.1f clk per MMA (256 MMAs back to back)\n", N, ts ? "TMEM" : "smem", maxerr, c / 256.0);
    }
  }
  return 0;
}
