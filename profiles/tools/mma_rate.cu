// Issue-rate probe for tcgen05.mma shapes/layouts used by the kernels (not product code).
// One CTA, one issuing thread; clk per MMA = (commit observed - first issue) / count.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../scaleprotoseg_b200/csrc/tc_common.cuh"
using namespace sps::tc;

enum Layout { KMAJ_SW128 = 0, MNMAJ_SW128 = 1, KMAJ_SW64 = 2, MNMAJ_SW64 = 3, MNMAJ_NOSW = 4 };

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
__device__ inline uint32_t kstep_bytes(int layout) {
  switch (layout) {
    case KMAJ_SW128: return 32;
    case MNMAJ_SW128: return 16 * 128;
    case KMAJ_SW64: return 32;
    case MNMAJ_NOSW: return 512;
    default: return 16 * 64;
  }
}

struct Op { int la, lb, M, N, ksteps, dcol, dlane; uint32_t lbo_a, lbo_b; };
struct Test { Op op[4]; int nops, reps; };

__global__ void __launch_bounds__(128) rate(Test t, long long* out) {
  extern __shared__ unsigned char raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  __shared__ uint32_t slot;
  __shared__ __align__(8) unsigned long long barmem;
  const uint32_t bar = smem_u32(&barmem);
  for (int i = threadIdx.x; i < (2 * 65536) / 4; i += 128) reinterpret_cast<uint32_t*>(raw + (base - smem_u32(raw)))[i] = 0;
  if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_fence_init(); }
  fence_proxy_async();
  if (threadIdx.x < 32) tmem_alloc(smem_u32(&slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  __shared__ uint64_t s_da[256], s_db[256];
  __shared__ uint32_t s_id[256], s_d[256];
  __shared__ int s_cnt;
  if (threadIdx.x == 0) {
    int cnt = 0;
    for (int r = 0; r < t.reps; ++r) {
      int kmax = 0;
      for (int o = 0; o < t.nops; ++o) kmax = t.op[o].ksteps > kmax ? t.op[o].ksteps : kmax;
      for (int ks = 0; ks < kmax; ++ks)
        for (int o = 0; o < t.nops; ++o) {
          const Op& p = t.op[o];
          if (ks >= p.ksteps || cnt >= 256) continue;
          uint32_t id = (1u << 4) | ((uint32_t)(p.N >> 3) << 17) | ((uint32_t)(p.M >> 4) << 24);
          if (p.la == MNMAJ_SW128 || p.la == MNMAJ_SW64) id |= 1u << 15;
          if (p.lb == MNMAJ_SW128 || p.lb == MNMAJ_SW64 || p.lb == MNMAJ_NOSW) id |= 1u << 16;
          s_da[cnt] = make_desc(p.la, base + (ks & 3) * kstep_bytes(p.la), p.lbo_a);
          s_db[cnt] = make_desc(p.lb, base + 65536 + (ks & 3) * kstep_bytes(p.lb), p.lbo_b);
          s_id[cnt] = id;
          s_d[cnt] = tmem + ((uint32_t)p.dlane << 16) + p.dcol;
          ++cnt;
        }
    }
    s_cnt = cnt;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int cnt = s_cnt;
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < cnt; ++i) mma_f16(s_d[i], s_da[i], s_db[i], s_id[i], 1);
    long long t1 = clock64();
    mma_commit(bar);
    mbar_wait(bar, 0);
    long long t2 = clock64();
    out[0] = t1 - t0; out[1] = t2 - t0; out[2] = cnt;
  }
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x < 32) tmem_dealloc(tmem, 512);
}


__device__ __forceinline__ void mma_f16_elect(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// whole warp 0 runs the loop on warp-uniform values; TMEMZERO: the 512-column allocation starts at column 0
template <bool TMEMZERO>
__global__ void __launch_bounds__(128) rate_uniform(Test t, long long* out) {
  extern __shared__ unsigned char raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  __shared__ uint32_t slot;
  __shared__ __align__(8) unsigned long long barmem;
  const uint32_t bar = smem_u32(&barmem);
  for (int i = threadIdx.x; i < (2 * 65536) / 4; i += 128) reinterpret_cast<uint32_t*>(raw + (base - smem_u32(raw)))[i] = 0;
  if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_fence_init(); }
  fence_proxy_async();
  if (threadIdx.x < 32) tmem_alloc(smem_u32(&slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (TMEMZERO && slot != 0) __trap();
  const uint32_t tmem = TMEMZERO ? 0u : slot;
  if (threadIdx.x < 32) {
    const Op p = t.op[0];
    uint32_t id = (1u << 4) | ((uint32_t)(p.N >> 3) << 17) | ((uint32_t)(p.M >> 4) << 24);
    if (p.la == MNMAJ_SW128 || p.la == MNMAJ_SW64) id |= 1u << 15;
    if (p.lb == MNMAJ_SW128 || p.lb == MNMAJ_SW64 || p.lb == MNMAJ_NOSW) id |= 1u << 16;
    const uint64_t da0 = make_desc(p.la, base, p.lbo_a), db0 = make_desc(p.lb, base + 65536, p.lbo_b);
    const uint32_t sa = kstep_bytes(p.la) >> 4, sb = kstep_bytes(p.lb) >> 4;
    const uint32_t d = tmem + ((uint32_t)p.dlane << 16) + p.dcol;
    const int reps = t.reps;
    long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) mma_f16_elect(d, da0 + ks * sa, db0 + ks * sb, id, 1);
    }
    long long t1 = clock64();
    if (elect_one()) mma_commit(bar);
    mbar_wait(bar, 0);
    long long t2 = clock64();
    if (threadIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; out[2] = reps * 4; }
  }
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x < 32) tmem_dealloc(tmem, 512);
}

template <bool Z>
static void run_u(const char* name, Test t) {
  long long* d; cudaMalloc(&d, 64);
  cudaFuncSetAttribute(rate_uniform<Z>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 65536 + 2048);
  long long h[3];
  for (int it = 0; it < 2; ++it) {
    rate_uniform<Z><<<1, 128, 2 * 65536 + 2048>>>(t, d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%-56s cuda=%s\n", name, cudaGetErrorString(e)); exit(1); }
  }
  cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
  printf("%-44s %s n=%4lld  issue %6.1f clk/MMA  complete %6.1f clk/MMA\n", name, Z ? "[uniform, tmem=0]" : "[uniform, tmem=LDS]", h[2], (double)h[0] / h[2], (double)h[1] / h[2]);
  cudaFree(d);
}

static void run(const char* name, Test t) {
  long long* d; cudaMalloc(&d, 64);
  cudaFuncSetAttribute(rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 65536 + 2048);
  long long h[3];
  for (int it = 0; it < 2; ++it) {
    rate<<<1, 128, 2 * 65536 + 2048>>>(t, d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%-56s cuda=%s\n", name, cudaGetErrorString(e)); exit(1); }
  }
  cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
  printf("%-56s n=%4lld  issue %6.1f clk/MMA  complete %6.1f clk/MMA\n", name, h[2], (double)h[0] / h[2], (double)h[1] / h[2]);
  cudaFree(d);
}

int main() {
  const int R = 16;
  auto one = [&](Op o) { Test t{}; t.op[0] = o; t.nops = 1; t.reps = R; return t; };
  run("K/K SW128 M128 N64 (fwd, distance)", one(Op{KMAJ_SW128, KMAJ_SW128, 128, 64, 4, 0, 0, 0, 0}));
  run("K/K SW128 M128 N128", one(Op{KMAJ_SW128, KMAJ_SW128, 128, 128, 4, 0, 0, 0, 0}));
  run("K/K SW128 M128 N192", one(Op{KMAJ_SW128, KMAJ_SW128, 128, 192, 4, 0, 0, 0, 0}));
  run("K/K SW128 M128 N256", one(Op{KMAJ_SW128, KMAJ_SW128, 128, 256, 4, 0, 0, 0, 0}));
  run("K/K SW128 M128 N32", one(Op{KMAJ_SW128, KMAJ_SW128, 128, 32, 4, 0, 0, 0, 0}));
  run("K/K SW128 M128 N16", one(Op{KMAJ_SW128, KMAJ_SW128, 128, 16, 4, 0, 0, 0, 0}));
  run("K/K SW128 M64 N64", one(Op{KMAJ_SW128, KMAJ_SW128, 64, 64, 4, 0, 0, 0, 0}));
  run("K/K SW64 M128 N64 (dA-type)", one(Op{KMAJ_SW64, KMAJ_SW64, 128, 64, 2, 64, 0, 0, 0}));
  run("A K SW128, B MN SW128 M128 N64 (dz)", one(Op{KMAJ_SW128, MNMAJ_SW128, 128, 64, 4, 0, 0, 0, 8192}));
  run("A MN, B MN SW128 M64 N64 (dp)", one(Op{MNMAJ_SW128, MNMAJ_SW128, 64, 64, 4, 0, 0, 16384, 16384}));
  run("A MN, B MN SW128 M128 N64 (dp stacked)", one(Op{MNMAJ_SW128, MNMAJ_SW128, 128, 64, 4, 0, 0, 16384, 16384}));
  run("A MN, B MN SW128 M128 N128", one(Op{MNMAJ_SW128, MNMAJ_SW128, 128, 128, 4, 0, 0, 16384, 16384}));
  run("A MN SW128, B MN SW64 M64 N32 (dW)", one(Op{MNMAJ_SW128, MNMAJ_SW64, 64, 32, 4, 64, 16, 16384, 8192}));
  run("A MN SW128, B MN nosw M64 N8 (csum)", one(Op{MNMAJ_SW128, MNMAJ_NOSW, 64, 8, 4, 96, 16, 16384, 0}));
  run("A MN SW128, B MN SW128 M64 N72", one(Op{MNMAJ_SW128, MNMAJ_SW128, 64, 72, 4, 0, 0, 16384, 16384}));
  run("A MN SW128, B MN SW128 M64 N96", one(Op{MNMAJ_SW128, MNMAJ_SW128, 64, 96, 4, 0, 0, 16384, 16384}));
  {
    Test t{}; t.nops = 3; t.reps = R;
    t.op[0] = Op{MNMAJ_SW128, MNMAJ_SW128, 64, 64, 4, 0, 0, 16384, 16384};
    t.op[1] = Op{MNMAJ_SW128, MNMAJ_SW64, 64, 32, 4, 0, 16, 16384, 8192};
    t.op[2] = Op{MNMAJ_SW128, MNMAJ_NOSW, 64, 8, 4, 32, 16, 16384, 0};
    run("I2 mix dp + dW + csum (as issued today)", t);
  }
  {
    Test t{}; t.nops = 2; t.reps = R;
    t.op[0] = Op{KMAJ_SW128, KMAJ_SW128, 128, 64, 4, 0, 0, 0, 0};
    t.op[1] = Op{KMAJ_SW128, KMAJ_SW128, 128, 64, 4, 64, 0, 0, 0};
    run("two accumulators alternating, K/K M128 N64", t);
  }
  printf("---- warp-uniform issue ----\n");
  Op ops[] = {Op{KMAJ_SW128, KMAJ_SW128, 128, 64, 4, 0, 0, 0, 0}, Op{KMAJ_SW128, KMAJ_SW128, 128, 16, 4, 0, 0, 0, 0},
              Op{KMAJ_SW128, KMAJ_SW128, 128, 128, 4, 0, 0, 0, 0}, Op{KMAJ_SW128, KMAJ_SW128, 128, 192, 4, 0, 0, 0, 0},
              Op{KMAJ_SW128, KMAJ_SW128, 128, 256, 4, 0, 0, 0, 0}, Op{KMAJ_SW128, MNMAJ_SW128, 128, 64, 4, 0, 0, 0, 8192},
              Op{MNMAJ_SW128, MNMAJ_SW128, 64, 64, 4, 0, 0, 16384, 16384}, Op{MNMAJ_SW128, MNMAJ_SW128, 128, 64, 4, 0, 0, 16384, 16384},
              Op{MNMAJ_SW128, MNMAJ_SW64, 64, 32, 4, 64, 16, 16384, 8192}, Op{MNMAJ_SW128, MNMAJ_NOSW, 64, 8, 4, 96, 16, 16384, 0},
              Op{KMAJ_SW64, KMAJ_SW64, 128, 64, 2, 64, 0, 0, 0}};
  const char* names[] = {"K/K M128 N64", "K/K M128 N16", "K/K M128 N128", "K/K M128 N192", "K/K M128 N256", "K/MN M128 N64 (dz)",
                         "MN/MN M64 N64 (dp)", "MN/MN M128 N64", "MN/MN64 M64 N32 (dW)", "MN/nosw M64 N8 (csum)", "K/K SW64 M128 N64 (dA)"};
  for (int i = 0; i < 11; ++i) { run_u<false>(names[i], one(ops[i])); run_u<true>(names[i], one(ops[i])); }
  return 0;
}
