// Stage-2 grouping head: per-class simplex projections + exp + last_layer_group, forward and backward.
// Follows compute_group / run_last_layer of segmentation/model/model_multiscale_group.py:283-308:
//   groups[n, J*k'+j] = exp( sum_q act[n, members_k'[q]] * G_k'[j, q] ),  logits = groups . Wg^T
// The reference walks the classes in a Python loop with one host sync and one gather copy per class;
// here one CTA handles 16 positions and every class at once.
#include "common.cuh"

namespace sps {

constexpr int GTM = 16;   // positions per CTA
constexpr int GNT = 512;  // max threads per CTA (the launch uses the (row, class) pair count rounded up to a warp)

// Per-CTA shared copies of the packed projections and of the member lists, each class block padded by one word:
// the lanes of a warp work on DIFFERENT classes (class index fastest), and without the padding their reads of
// G_k'[j,q] / members_k'[q] (class blocks J*n_k' / n_k' words apart) collide in a few banks — as global loads they
// cost one L1 wavefront per lane and dominated the kernel.
__device__ __forceinline__ int gpad_floats(const HeadShape& s, const SpsGroupDesc& g) { return g.J * s.P + g.Kc; }
__device__ __forceinline__ void stage_tables(const HeadShape& s, const SpsGroupDesc& g, float* Gs, int* Ms) {
  for (int kc = 0; kc < g.Kc; ++kc) {
    const int off = g.class_off[kc], nk = g.class_off[kc + 1] - off;
    for (int i = threadIdx.x; i < g.J * nk; i += blockDim.x) Gs[g.J * off + kc + i] = __ldg(g.G + (long long)g.J * off + i);
    for (int i = threadIdx.x; i < nk; i += blockDim.x) Ms[off + kc + i] = __ldg(g.members + off + i);
  }
  __syncthreads();
}

// gs[m][J*k'+j] for the tile starting at n0 (zero rows beyond N)
__device__ __forceinline__ void group_tile(const HeadShape& s, const SpsGroupDesc& g, const float* __restrict__ act,
                                           long long n0, float* gs, const float* Gs, const int* Ms, int Gt) {
  for (int idx = threadIdx.x; idx < GTM * g.Kc; idx += blockDim.x) {
    const int m = idx / g.Kc, kc = idx % g.Kc;
    const long long n = n0 + m;
    const int off = g.class_off[kc], nk = g.class_off[kc + 1] - off;
    float sum[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) sum[j] = 0.f;
    if (n < s.N) {
      const float* arow = act + n * s.P;
      const float* Gk = Gs + g.J * off + kc;
      const int* Mk = Ms + off + kc;
      // batches of 8 members: the 8 activation loads (the only ones that go to DRAM) are in flight together
      int q = 0;
      for (; q + 8 <= nk; q += 8) {
        float a[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) a[u] = __ldg(arow + Mk[q + u]);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (j < g.J) sum[j] = fmaf(a[u], Gk[j * nk + q + u], sum[j]);
        }
      }
      for (; q < nk; ++q) {
        const float a = __ldg(arow + Mk[q]);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j < g.J) sum[j] = fmaf(a, Gk[j * nk + q], sum[j]);
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
  float* Gs = gs + GTM * Gt;
  int* Ms = reinterpret_cast<int*>(Gs + gpad_floats(s, g));
  stage_tables(s, g, Gs, Ms);
  const long long n0 = (long long)blockIdx.x * GTM;
  group_tile(s, g, act, n0, gs, Gs, Ms, Gt);
  __syncthreads();
  if (groups != nullptr) {
    for (int idx = threadIdx.x; idx < GTM * Gt; idx += blockDim.x) {
      const long long n = n0 + idx / Gt;
      if (n < s.N) groups[n * Gt + idx % Gt] = gs[idx];
    }
  }
  if (logits != nullptr) {
    for (int idx = threadIdx.x; idx < GTM * s.K; idx += blockDim.x) {
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

static int group_threads(const SpsGroupDesc& g) {
  int t = (GTM * g.Kc + 31) / 32 * 32;   // one (row, class) pair per thread in a single round
  return t < 128 ? 128 : (t > GNT ? GNT : t);
}

int launch_group_fwd(const HeadShape& s, const SpsGroupDesc& g, const float* act, float* groups, float* logits,
                     cudaStream_t st) {
  if (s.N == 0 || g.Kc == 0) return SPS_OK;
  const int Gt = g.J * g.Kc;
  size_t smem = sizeof(float) * ((size_t)GTM * Gt + (size_t)g.J * s.P + g.Kc + s.P + g.Kc);
  SPS_REQUIRE(smem <= 200 * 1024, SPS_ERR_UNSUPPORTED, "group head: J*Kc=%d, J*P=%d need %zu bytes of shared memory", Gt, g.J * s.P, smem);
  SPS_CUDA_TRY(cudaFuncSetAttribute(group_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long tiles = (s.N + GTM - 1) / GTM;
  group_fwd_kernel<<<(unsigned)tiles, group_threads(g), smem, st>>>(s, g, act, groups, logits);
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
  float* Gs = smem + 2 * GTM * Gt;
  int* Ms = reinterpret_cast<int*>(Gs + gpad_floats(s, g));
  stage_tables(s, g, Gs, Ms);
  const long long n0 = (long long)blockIdx.x * GTM;
  group_tile(s, g, act, n0, gs, Gs, Ms, Gt);
  __syncthreads();
  for (int idx = threadIdx.x; idx < GTM * Gt; idx += blockDim.x) {
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
    for (int idx = threadIdx.x; idx < GTM * s.P; idx += blockDim.x) {
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

// dWg[k,g] += sum_n dlogits[n,k] * groups[n,g] over this CTA's slice of rows.  The two row tiles are staged in shared
// memory (coalesced), every thread accumulates its (k,g) outputs over the slice, one atomicAdd per output and CTA.
constexpr int GWR = 64;  // rows per staging step
__global__ void __launch_bounds__(256) group_bwd_dWg_kernel(HeadShape s, int Gt, const float* __restrict__ dlogits,
                                                            const float* __restrict__ ws_groups, float* __restrict__ dWg) {
  extern __shared__ __align__(16) float sm[];   // dl [GWR][K] | groups [GWR][Gt]
  float* dls = sm;
  float* grs = sm + GWR * s.K;
  const long long per = (s.N + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x, hi = min(s.N, lo + per);
  const int n_out = s.K * Gt;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (long long r0 = lo; r0 < hi; r0 += GWR) {
    const int rows = (int)min((long long)GWR, hi - r0);
    __syncthreads();
    for (int i = threadIdx.x; i < rows * s.K; i += blockDim.x) dls[i] = __ldg(dlogits + r0 * s.K + i);
    for (int i = threadIdx.x; i < rows * Gt; i += blockDim.x) grs[i] = __ldg(ws_groups + r0 * Gt + i);
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int o = (blockIdx.y * 8 + i) * blockDim.x + threadIdx.x;
      if (o < n_out) {
        const int k = o / Gt, gg = o - k * Gt;
        float a = acc[i];
        for (int m = 0; m < rows; ++m) a = fmaf(dls[m * s.K + k], grs[m * Gt + gg], a);
        acc[i] = a;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int o = (blockIdx.y * 8 + i) * blockDim.x + threadIdx.x;
    if (o < n_out && lo < hi) atomicAdd(dWg + o, acc[i]);
  }
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
  size_t smem = sizeof(float) * (2 * (size_t)GTM * Gt + (size_t)g.J * s.P + g.Kc + s.P + g.Kc);
  SPS_REQUIRE(smem <= 200 * 1024, SPS_ERR_UNSUPPORTED, "group head: J*Kc=%d, J*P=%d need %zu bytes of shared memory", Gt, g.J * s.P, smem);
  float* ws_groups = ws;
  float* ws_dpre = ws + (size_t)s.N * Gt;
  SPS_CUDA_TRY(cudaFuncSetAttribute(group_bwd_pos_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long tiles = (s.N + GTM - 1) / GTM;
  group_bwd_pos_kernel<<<(unsigned)tiles, group_threads(g), smem, st>>>(s, g, act, dlogits, dgroups, dact, ws_groups, ws_dpre);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
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
    long long wsplit = (s.N + 4 * GWR - 1) / (4 * GWR);     // >= 256 rows per CTA
    if (wsplit > 4LL * sm_count()) wsplit = 4LL * sm_count();
    if (wsplit < 1) wsplit = 1;
    dim3 grid((unsigned)wsplit, (unsigned)((s.K * Gt + 8 * 256 - 1) / (8 * 256)));   // y: chunks of 2048 outputs
    group_bwd_dWg_kernel<<<grid, 256, sizeof(float) * GWR * (s.K + Gt), st>>>(s, Gt, dlogits, ws_groups, dWg);
    count_launch();
    SPS_CUDA_TRY(cudaGetLastError());
  }
  return SPS_OK;
}

}  // namespace sps
