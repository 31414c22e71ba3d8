// Probe of tcgen05.mma operand layouts used by the backward kernel (not product code).
// Each test fills shared-memory tiles from logical fp16 matrices, issues the MMAs, reads D from TMEM and
// compares with a CPU product.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "../scaleprotoseg_b200/csrc/tc_common.cuh"

using namespace sps::tc;

enum Layout { KMAJ_SW128 = 0, MNMAJ_SW128 = 1, KMAJ_SW64 = 2, MNMAJ_SW64 = 3, MNMAJ_NOSW = 4 };

// byte offset of logical element (mn, k) in a tile; tiles are 1024-aligned.
// KMAJ_SW128: rows mn, 64 k per 128-B row (k < 64).           mn-groups of 8 rows: 1024 B
// MNMAJ_SW128: rows k, 64 mn per 128-B row; mn >= 64 -> next block at `lbo`
// KMAJ_SW64: rows mn, 32 k per 64-B row (k < 32); 8 rows: 512 B; 16-B chunk index ^= (row>>1)&3
// MNMAJ_SW64: rows k, 32 mn per 64-B row; mn >= 32 -> +lbo
__host__ __device__ inline uint32_t elem_off(int layout, int mn, int k, uint32_t lbo) {
  switch (layout) {
    case KMAJ_SW128: return (uint32_t)mn * 128u + ((((uint32_t)k >> 3) ^ ((uint32_t)mn & 7u)) << 4) + ((k & 7) << 1);
    case MNMAJ_SW128: {
      uint32_t blk = mn >> 6, m = mn & 63;
      return blk * lbo + (uint32_t)k * 128u + (((m >> 3) ^ ((uint32_t)k & 7u)) << 4) + ((m & 7) << 1);
    }
    case KMAJ_SW64: return (uint32_t)mn * 64u + ((((uint32_t)k >> 3) ^ (((uint32_t)mn >> 1) & 3u)) << 4) + ((k & 7) << 1);
    case MNMAJ_NOSW: return (uint32_t)(k & 7) * 16u + (uint32_t)(k >> 3) * 256u + (uint32_t)(mn >> 3) * 128u + (uint32_t)(mn & 7) * 2u;
    default: {
      uint32_t blk = mn >> 5, m = mn & 31;
      return blk * lbo + (uint32_t)k * 64u + (((m >> 3) ^ (((uint32_t)k >> 1) & 3u)) << 4) + ((m & 7) << 1);
    }
  }
}

__device__ inline uint64_t make_desc(int layout, uint32_t addr, uint32_t lbo) {
  uint64_t d = (uint64_t)((addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 46);
  switch (layout) {
    case KMAJ_SW128: return d | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)2 << 61);
    case MNMAJ_SW128: return d | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)2 << 61);
    case KMAJ_SW64: return d | ((uint64_t)1 << 16) | ((uint64_t)(512 >> 4) << 32) | ((uint64_t)4 << 61);
    case MNMAJ_NOSW: return d | ((uint64_t)(256 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32);
    default: return d | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(512 >> 4) << 32) | ((uint64_t)4 << 61);
  }
}
// start-address advance (bytes) per 16-element K step
__device__ inline uint32_t kstep_bytes(int layout) {
  switch (layout) {
    case KMAJ_SW128: return 32;
    case MNMAJ_SW128: return 16 * 128;
    case KMAJ_SW64: return 32;
    case MNMAJ_NOSW: return 512;
    default: return 16 * 64;
  }
}

struct Test {
  int la, lb, M, N, K;
  uint32_t lbo_a, lbo_b;
  int lane_off;  // TMEM lane offset of D (M=64 interleave probe)
};

__global__ void __launch_bounds__(128) probe(Test t, const __half* A /*M x K*/, const __half* B /*N x K*/, float* D /*M x N*/) {
  extern __shared__ unsigned char raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  unsigned char* g = raw + (base - smem_u32(raw));
  unsigned char* sa = g;
  unsigned char* sb = g + 65536;
  __shared__ uint32_t slot;
  __shared__ __align__(8) unsigned long long barmem;
  const uint32_t bar = smem_u32(&barmem);
  for (int i = threadIdx.x; i < 65536 / 4; i += 128) {
    reinterpret_cast<uint32_t*>(sa)[i] = 0;
    reinterpret_cast<uint32_t*>(sb)[i] = 0;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < t.M * t.K; i += 128) {
    int m = i / t.K, k = i % t.K;
    *reinterpret_cast<__half*>(sa + elem_off(t.la, m, k, t.lbo_a)) = A[i];
  }
  for (int i = threadIdx.x; i < t.N * t.K; i += 128) {
    int n = i / t.K, k = i % t.K;
    *reinterpret_cast<__half*>(sb + elem_off(t.lb, n, k, t.lbo_b)) = B[i];
  }
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  fence_proxy_async();
  if (threadIdx.x < 32) tmem_alloc(smem_u32(&slot), 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    // idesc: fp32 accum, f16, a_major bit15, b_major bit16
    uint32_t id = (1u << 4) | ((uint32_t)(t.N >> 3) << 17) | ((uint32_t)(t.M >> 4) << 24);
    if (t.la == MNMAJ_SW128 || t.la == MNMAJ_SW64) id |= 1u << 15;
    if (t.lb == MNMAJ_SW128 || t.lb == MNMAJ_SW64 || t.lb == MNMAJ_NOSW) id |= 1u << 16;
    for (int ks = 0; ks < t.K / 16; ++ks) {
      // K-major tiles hold at most 64 (SW128) / 32 (SW64) k per row; the probe keeps K within one tile for those
      uint64_t da = make_desc(t.la, base + ks * kstep_bytes(t.la), t.lbo_a);
      uint64_t db = make_desc(t.lb, base + 65536 + ks * kstep_bytes(t.lb), t.lbo_b);
      mma_f16(tmem + ((uint32_t)t.lane_off << 16), da, db, id, ks > 0);
    }
    mma_commit(bar);
  }
  mbar_wait(bar, 0);
  tc_fence_after();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c0 = 0; c0 < t.N; c0 += 16) {
    uint32_t r[16];
    tmem_ld16(tmem + (((uint32_t)warp * 32) << 16) + c0, r);
    tmem_ld_wait();
    const int row = warp * 32 + lane;
    for (int j = 0; j < 16; ++j)
      if (c0 + j < t.N) D[(size_t)row * t.N + c0 + j] = __uint_as_float(r[j]);  // raw lane dump (128 lanes)
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tmem, 256);
}

static void run(const char* name, Test t) {
  std::vector<__half> A(t.M * t.K), B(t.N * t.K);
  std::vector<float> Af(t.M * t.K), Bf(t.N * t.K);
  for (size_t i = 0; i < A.size(); ++i) { float v = (float)((rand() % 17) - 8) / 8.0f; A[i] = __float2half(v); Af[i] = v; }
  for (size_t i = 0; i < B.size(); ++i) { float v = (float)((rand() % 17) - 8) / 8.0f; B[i] = __float2half(v); Bf[i] = v; }
  __half *dA, *dB; float* dD;
  cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2); cudaMalloc(&dD, 128 * t.N * 4);
  cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0, 128 * t.N * 4);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 65536 + 2048);
  probe<<<1, 128, 2 * 65536 + 2048>>>(t, dA, dB, dD);
  cudaError_t e = cudaDeviceSynchronize();
  std::vector<float> D(128 * t.N);
  cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
  // reference and, for M=64, try to discover the lane mapping
  double maxerr = 0; int bad = 0;
  for (int m = 0; m < t.M; ++m)
    for (int n = 0; n < t.N; ++n) {
      float ref = 0; for (int k = 0; k < t.K; ++k) ref += Af[m * t.K + k] * Bf[n * t.K + k];
      int lane = t.M == 64 ? ((m / 16) * 32 + m % 16 + t.lane_off) : m;
      double err = fabs(D[(size_t)lane * t.N + n] - ref);
      if (err > maxerr) maxerr = err;
      if (err > 1e-3) ++bad;
    }
  printf("%-44s M=%d N=%d K=%d  cuda=%s  maxerr=%.3g  bad=%d/%d\n", name, t.M, t.N, t.K, cudaGetErrorString(e), maxerr, bad, t.M * t.N);
  cudaFree(dA); cudaFree(dB); cudaFree(dD);
}

int main() {
  srand(1);
  run("T0 K-major SW128 A,B (forward GEMM)", Test{KMAJ_SW128, KMAJ_SW128, 128, 64, 64, 0, 0, 0});
  run("T1 A K-major SW128, B MN-major SW128 (dz)", Test{KMAJ_SW128, MNMAJ_SW128, 128, 64, 64, 0, 8192, 0});
  run("T2 A MN-major SW128 (2 blocks), B MN-major SW128", Test{MNMAJ_SW128, MNMAJ_SW128, 128, 64, 128, 16384, 16384, 0});
  run("T3 A,B K-major SW64 (dA GEMM)", Test{KMAJ_SW64, KMAJ_SW64, 128, 128, 32, 0, 0, 0});
  run("T4 A MN-major SW128, B MN-major SW64 (dW)", Test{MNMAJ_SW128, MNMAJ_SW64, 128, 32, 128, 16384, 8192, 0});
  run("T5 M=64 lane-mapping probe (K-major)", Test{KMAJ_SW128, KMAJ_SW128, 64, 64, 64, 0, 0, 0});
  run("T7 M=64 lane offset 16, MN-major A (dW style)", Test{MNMAJ_SW128, MNMAJ_SW64, 64, 32, 128, 16384, 8192, 16});
  run("T8 M=64 lane offset 0, MN-major A, B SW128", Test{MNMAJ_SW128, MNMAJ_SW128, 64, 64, 128, 16384, 16384, 0});
  run("T9 M=64 lane 16, B MN-major no-swizzle N=8 (csum)", Test{MNMAJ_SW128, MNMAJ_NOSW, 64, 8, 128, 16384, 0, 16});
  run("T6 A K-major SW128 N=128 B (2nd half rows)", Test{KMAJ_SW128, KMAJ_SW128, 128, 128, 64, 0, 0, 0});
  return 0;
}
