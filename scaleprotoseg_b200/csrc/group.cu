// Stage-2 grouping head: per-class simplex projections + exp + last_layer_group, forward and backward.
// Follows compute_group / run_last_layer of segmentation/model/model_multiscale_group.py:283-308:
//   groups[n, J*k'+j] = exp( sum_q act[n, members_k'[q]] * G_k'[j, q] ),  logits = groups . Wg^T
// The reference walks the classes in a Python loop with one host sync and one gather copy per class;
// here one CTA handles 16 positions and every class at once.
#include "common.cuh"

namespace sps {

constexpr int GTM = 16;   // positions per CTA
constexpr int GNT = 256;  // threads

// gs[m][J*k'+j] for the tile starting at n0 (zero rows beyond N)
__device__ __forceinline__ void group_tile(const HeadShape& s, const SpsGroupDesc& g, const float* __restrict__ act,
                                           long long n0, float* gs, int Gt) {
  for (int idx = threadIdx.x; idx < GTM * g.Kc; idx += GNT) {
    const int m = idx / g.Kc, kc = idx % g.Kc;
    const long long n = n0 + m;
    const int off = g.class_off[kc], nk = g.class_off[kc + 1] - off;
    float sum[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) sum[j] = 0.f;
    if (n < s.N) {
      const float* arow = act + n * s.P;
      const float* Gk = g.G + (long long)g.J * off;
      for (int q = 0; q < nk; ++q) {
        const float a = __ldg(arow + g.members[off + q]);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j < g.J) sum[j] = fmaf(a, __ldg(Gk + j * nk + q), sum[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (j < g.J) gs[m * Gt + g.J * kc + j] = n < s.N ? expf(sum[j]) : 0.f;
  }
}

__global__ void __launch_bounds__(GNT) group_fwd_kernel(HeadShape s, SpsGroupDesc g, const float* __restrict__ act,
                                                         float* __restrict__ groups, float* __restrict__ logits) {
  extern __shared__ __align__(16) float gs[];
  const int Gt = g.J * g.Kc;
  const long long n0 = (long long)blockIdx.x * GTM;
  group_tile(s, g, act, n0, gs, Gt);
  __syncthreads();
  if (groups != nullptr) {
    for (int idx = threadIdx.x; idx < GTM * Gt; idx += GNT) {
      const long long n = n0 + idx / Gt;
      if (n < s.N) groups[n * Gt + idx % Gt] = gs[idx];
    }
  }
  if (logits != nullptr) {
    for (int idx = threadIdx.x; idx < GTM * s.K; idx += GNT) {
      const int m = idx / s.K, k = idx % s.K;
      const long long n = n0 + m;
      if (n >= s.N) continue;
      const float* w = g.Wg + (long long)k * Gt;
      float a = 0.f;
      for (int gg = 0; gg < Gt; ++gg) a = fmaf(gs[m * Gt + gg], __ldg(w + gg), a);
      logits[n * s.K + k] = a;
    }
  }
}

int launch_group_fwd(const HeadShape& s, const SpsGroupDesc& g, const float* act, float* groups, float* logits,
                     cudaStream_t st) {
  if (s.N == 0 || g.Kc == 0) return SPS_OK;
  const int Gt = g.J * g.Kc;
  size_t smem = sizeof(float) * GTM * Gt;
  SPS_REQUIRE(smem <= 200 * 1024, SPS_ERR_UNSUPPORTED, "group head: J*Kc=%d too large", Gt);
  SPS_CUDA_TRY(cudaFuncSetAttribute(group_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long tiles = (s.N + GTM - 1) / GTM;
  group_fwd_kernel<<<(unsigned)tiles, GNT, smem, st>>>(s, g, act, groups, logits);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

// Backward pass 1: recompute groups, dpre = (dlogits Wg + dgroups) * groups, dact; spill groups and
// dpre to the workspace for the two reductions over n.
__global__ void __launch_bounds__(GNT) group_bwd_pos_kernel(HeadShape s, SpsGroupDesc g, const float* __restrict__ act,
                                                             const float* __restrict__ dlogits,
                                                             const float* __restrict__ dgroups,
                                                             float* __restrict__ dact, float* __restrict__ ws_groups,
                                                             float* __restrict__ ws_dpre) {
  extern __shared__ __align__(16) float smem[];
  const int Gt = g.J * g.Kc;
  float* gs = smem;
  float* dps = smem + GTM * Gt;
  const long long n0 = (long long)blockIdx.x * GTM;
  group_tile(s, g, act, n0, gs, Gt);
  __syncthreads();
  for (int idx = threadIdx.x; idx < GTM * Gt; idx += GNT) {
    const int m = idx / Gt, gg = idx % Gt;
    const long long n = n0 + m;
    float v = 0.f;
    if (n < s.N) {
      float up = dgroups != nullptr ? __ldg(dgroups + n * Gt + gg) : 0.f;
      if (dlogits != nullptr)
        for (int k = 0; k < s.K; ++k) up = fmaf(__ldg(dlogits + n * s.K + k), __ldg(g.Wg + (long long)k * Gt + gg), up);
      v = up * gs[idx];
      ws_groups[n * Gt + gg] = gs[idx];
      ws_dpre[n * Gt + gg] = v;
    }
    dps[idx] = v;
  }
  __syncthreads();
  if (dact != nullptr) {
    for (int idx = threadIdx.x; idx < GTM * s.P; idx += GNT) {
      const int m = idx / s.P, p = idx % s.P;
      const long long n = n0 + m;
      if (n >= s.N) continue;
      const int kc = g.proto_slot[p], col = g.proto_col[p];
      float a = 0.f;
      if (kc >= 0) {
        const int off = g.class_off[kc], nk = g.class_off[kc + 1] - off;
        const float* Gk = g.G + (long long)g.J * off;
        for (int j = 0; j < g.J; ++j) a = fmaf(dps[m * Gt + g.J * kc + j], __ldg(Gk + j * nk + col), a);
      }
      dact[n * s.P + p] = a;
    }
  }
}

// dG_k'[j, col(p)] += sum_n dpre[n, J k'+j] * act[n, p]  over this CTA's slice of n.  One thread per prototype p
// (blockIdx.x * 256 + threadIdx.x): the activation map is read along p, coalesced, exactly once.
constexpr int GDR = 16;  // dpre rows staged per step
__global__ void __launch_bounds__(256) group_bwd_dG_kernel(HeadShape s, SpsGroupDesc g, const float* __restrict__ act,
                                                           const float* __restrict__ ws_dpre, float* __restrict__ dG) {
  extern __shared__ __align__(16) float dps[];   // [GDR][Gt]
  const int Gt = g.J * g.Kc;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int kc = p < s.P ? g.proto_slot[p] : -1;
  const long long per = (s.N + gridDim.y - 1) / gridDim.y;
  const long long lo = per * blockIdx.y, hi = min(s.N, lo + per);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (long long r0 = lo; r0 < hi; r0 += GDR) {
    const int rows = (int)min((long long)GDR, hi - r0);
    __syncthreads();
    for (int idx = threadIdx.x; idx < rows * Gt; idx += blockDim.x) dps[idx] = __ldg(ws_dpre + r0 * Gt + idx);
    __syncthreads();
    if (kc >= 0) {
      for (int m = 0; m < rows; ++m) {
        const float a = __ldg(act + (r0 + m) * s.P + p);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j < g.J) acc[j] = fmaf(dps[m * Gt + g.J * kc + j], a, acc[j]);
      }
    }
  }
  if (kc >= 0) {
    const int off = g.class_off[kc], nk = g.class_off[kc + 1] - off, col = g.proto_col[p];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (j < g.J) atomicAdd(dG + (long long)g.J * off + j * nk + col, acc[j]);
  }
}

// dWg[k,g] += sum_n dlogits[n,k] * groups[n,g]
__global__ void group_bwd_dWg_kernel(HeadShape s, int Gt, const float* __restrict__ dlogits,
                                     const float* __restrict__ ws_groups, float* __restrict__ dWg) {
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= s.K * Gt) return;
  const int k = o / Gt, gg = o % Gt;
  const long long per = (s.N + gridDim.y - 1) / gridDim.y;
  const long long lo = per * blockIdx.y, hi = min(s.N, lo + per);
  float a = 0.f;
  for (long long n = lo; n < hi; ++n) a = fmaf(__ldg(dlogits + n * s.K + k), __ldg(ws_groups + n * Gt + gg), a);
  atomicAdd(dWg + o, a);
}

int launch_group_bwd(const HeadShape& s, const SpsGroupDesc& g, const float* act, const float* dlogits,
                     const float* dgroups, float* dact, float* dG, float* dWg, float* ws, cudaStream_t st) {
  const int Gt = g.J * g.Kc;
  if (dG != nullptr) SPS_CUDA_TRY(cudaMemsetAsync(dG, 0, sizeof(float) * (size_t)g.J * s.P, st));
  if (dWg != nullptr) SPS_CUDA_TRY(cudaMemsetAsync(dWg, 0, sizeof(float) * (size_t)s.K * Gt, st));
  if (s.N == 0 || g.Kc == 0) {
    if (dact != nullptr && s.N > 0) SPS_CUDA_TRY(cudaMemsetAsync(dact, 0, sizeof(float) * (size_t)s.N * s.P, st));
    return SPS_OK;
  }
  size_t smem = sizeof(float) * 2 * GTM * Gt;
  SPS_REQUIRE(smem <= 200 * 1024, SPS_ERR_UNSUPPORTED, "group head: J*Kc=%d too large", Gt);
  float* ws_groups = ws;
  float* ws_dpre = ws + (size_t)s.N * Gt;
  SPS_CUDA_TRY(cudaFuncSetAttribute(group_bwd_pos_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long tiles = (s.N + GTM - 1) / GTM;
  group_bwd_pos_kernel<<<(unsigned)tiles, GNT, smem, st>>>(s, g, act, dlogits, dgroups, dact, ws_groups, ws_dpre);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  long long split = (s.N + 255) / 256;
  if (split > 64) split = 64;
  if (split < 1) split = 1;
  if (dG != nullptr) {
    long long rsplit = (s.N + 127) / 128;   // >= 128 rows per CTA; enough CTAs to fill the machine
    const long long want = (long long)sm_count() * 8 / ((s.P + 255) / 256) + 1;
    if (rsplit > want) rsplit = want;
    if (rsplit < 1) rsplit = 1;
    dim3 grid((s.P + 255) / 256, (unsigned)rsplit);
    group_bwd_dG_kernel<<<grid, 256, sizeof(float) * GDR * Gt, st>>>(s, g, act, ws_dpre, dG);
    count_launch();
    SPS_CUDA_TRY(cudaGetLastError());
  }
  if (dWg != nullptr && dlogits != nullptr) {
    dim3 grid((s.K * Gt + 255) / 256, (unsigned)split);
    group_bwd_dWg_kernel<<<grid, 256, 0, st>>>(s, Gt, dlogits, ws_groups, dWg);
    count_launch();
    SPS_CUDA_TRY(cudaGetLastError());
  }
  return SPS_OK;
}

}  // namespace sps
