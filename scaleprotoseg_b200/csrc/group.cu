// Stage-2 grouping head: per-class simplex projections + exp + last_layer_group, forward and backward.
// Follows compute_group / run_last_layer of segmentation/model/model_multiscale_group.py:283-308:
//   groups[n, J*k'+j] = exp( sum_q act[n, members_k'[q]] * G_k'[j, q] ),  logits = groups . Wg^T
// The reference walks the classes in a Python loop with one host sync and one gather copy per class;
// here one CTA handles 16 positions and every class at once.
#include "common.cuh"

namespace sps {

constexpr int GTM = 16;   // positions per CTA
#ifndef SPS_GROUP_TPB
#define SPS_GROUP_TPB 2
#endif
constexpr int GTPB = SPS_GROUP_TPB;   // consecutive 16-position tiles per CTA (tables staged once per CTA)
constexpr int GNT = 512;  // max threads per CTA (the launch uses the (row, class) pair count rounded up to a warp)

// Per-CTA shared copies of the packed projections and of the member lists, each class block padded by one word:
// the lanes of a warp work on DIFFERENT classes (class index fastest), and without the padding their reads of
// G_k'[j,q] / members_k'[q] (class blocks J*n_k' / n_k' words apart) collide in a few banks — as global loads they
// cost one L1 wavefront per lane and dominated the kernel.
__device__ __forceinline__ int gpad_floats(const HeadShape& s, const SpsGroupDesc& g) { return g.J * s.P + g.Kc; }
// The thread block is cut into Kc teams (one per class, >= 1 thread each) that copy their class's block of G and of the
// member list in parallel: no search, no division per element, and the loads of a thread are independent.  (A loop
// per class with the whole block made 2*Kc dependent round trips to L2 per CTA - 38 at Kc = 19 - and that latency,
// not the activation traffic, set the kernel time.)
__device__ __forceinline__ void stage_tables(const HeadShape& s, const SpsGroupDesc& g, float* Gs, int* Ms) {
  const int team = blockDim.x / g.Kc > 0 ? blockDim.x / g.Kc : 1;   // threads per class
  for (int kc = threadIdx.x / team; kc < g.Kc; kc += (blockDim.x + team - 1) / team) {
    const int off = g.class_off[kc], nk = g.class_off[kc + 1] - off, t = threadIdx.x % team;
    const float* src = g.G + (long long)g.J * off;
    float* dst = Gs + g.J * off + kc;
    // 4-byte asynchronous copies (the padded destination is not 16-byte aligned): no register staging, and they
    // complete together with the activation rows (wait_rows)
    for (int i = t; i < g.J * nk; i += team)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst + i)), "l"(src + i) : "memory");
    for (int i = t; i < nk; i += team)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(Ms + off + kc + i)),
                   "l"(g.members + off + i) : "memory");
  }
}

// Rows [n0, n0 + GTM) of the activation map into shared memory As[GTM][P] with 16-byte asynchronous copies (issued
// before the tables are staged, waited for after).  Gathering the members straight from global memory cost one L1
// wavefront per lane: the lanes of a warp sit on different classes, i.e. on different sectors of the row.
__device__ __forceinline__ void stage_rows(const HeadShape& s, const float* __restrict__ act, long long n0, float* As) {
  const long long left = s.N - n0;
  const int rows = left < GTM ? (int)left : GTM;
  if ((s.P & 3) == 0 && (reinterpret_cast<uintptr_t>(act) & 15) == 0) {
    const int per_row = s.P >> 2, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
    for (int m = warp; m < rows; m += nwarp) {          // one warp per row: consecutive lanes, consecutive 16-byte pieces
      const float* src = act + (n0 + m) * s.P;
      const uint32_t dst = (uint32_t)__cvta_generic_to_shared(As + m * s.P);
      for (int c = lane; c < per_row; c += 32)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * c), "l"(src + 4 * c) : "memory");
    }
  } else {
    for (int idx = threadIdx.x; idx < rows * s.P; idx += blockDim.x) As[idx] = __ldg(act + n0 * s.P + idx);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void wait_rows() {
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncthreads();
}

// gs[m][J*k'+j] for the tile starting at n0 (zero rows beyond N).  STAGED: the rows are in As (stage_rows).
// JT: groups per class as a compile-time constant (3 = the reference's num_groups), or 0 = read g.J at run time: with a
// run-time J the 8-wide unrolled inner loops issue 8 predicated LDS + FFMA per member instead of J.
template <bool STAGED, int JT>
__device__ __forceinline__ void group_tile_j(const HeadShape& s, const SpsGroupDesc& g, const float* __restrict__ act,
                                             long long n0, float* gs, const float* Gs, const int* Ms, int Gt, const float* As) {
  const int J = JT ? JT : g.J;
  for (int idx = threadIdx.x; idx < GTM * g.Kc; idx += blockDim.x) {
    const int m = idx / g.Kc, kc = idx % g.Kc;
    const long long n = n0 + m;
    const int off = g.class_off[kc], nk = g.class_off[kc + 1] - off;
    float sum[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) sum[j] = 0.f;
    if (n < s.N) {
      const float* arow = STAGED ? As + m * s.P : act + n * s.P;
      const float* Gk = Gs + J * off + kc;
      const int* Mk = Ms + off + kc;
      // batches of 8 members: the 8 activation loads (the only ones that go to DRAM) are in flight together
      int q = 0;
      for (; q + 8 <= nk; q += 8) {
        float a[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) a[u] = STAGED ? arow[Mk[q + u]] : __ldg(arow + Mk[q + u]);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (j < J) sum[j] = fmaf(a[u], Gk[j * nk + q + u], sum[j]);
        }
      }
      for (; q < nk; ++q) {
        const float a = STAGED ? arow[Mk[q]] : __ldg(arow + Mk[q]);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j < J) sum[j] = fmaf(a, Gk[j * nk + q], sum[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (j < J) gs[m * Gt + J * kc + j] = n < s.N ? expf(sum[j]) : 0.f;
  }
}

template <bool STAGED>
__device__ __forceinline__ void group_tile(const HeadShape& s, const SpsGroupDesc& g, const float* __restrict__ act,
                                           long long n0, float* gs, const float* Gs, const int* Ms, int Gt, const float* As) {
  if (g.J == 3) group_tile_j<STAGED, 3>(s, g, act, n0, gs, Gs, Ms, Gt, As);
  else group_tile_j<STAGED, 0>(s, g, act, n0, gs, Gs, Ms, Gt, As);
}

// shared memory: [rows As: GTM*P floats, 16-byte aligned, STAGED only][gs][Gs][Ms]
template <bool STAGED>
__global__ void __launch_bounds__(GNT) group_fwd_kernel(HeadShape s, SpsGroupDesc g, const float* __restrict__ act,
                                                         float* __restrict__ groups, float* __restrict__ logits, int ws_smem) {
  extern __shared__ __align__(16) float smem_f[];
  const int Gt = g.J * g.Kc;
  float* As = smem_f;
  float* gs = smem_f + (STAGED ? GTM * s.P : 0);
  float* Gs = gs + GTM * Gt;
  int* Ms = reinterpret_cast<int*>(Gs + gpad_floats(s, g));
  float* Ws = reinterpret_cast<float*>(Ms + s.P + g.Kc);   // last_layer_group [K][Gt] (when it fits: ws_smem)
  const float* Wp = ws_smem ? Ws : g.Wg;
  // a CTA walks GTPB consecutive tiles: the tables are staged once, the rows of the first tile fly meanwhile
  long long n0 = (long long)blockIdx.x * GTPB * GTM;
  if (STAGED) stage_rows(s, act, n0, As);
  if (logits != nullptr && ws_smem)
    for (int i = threadIdx.x; i < s.K * Gt; i += blockDim.x) Ws[i] = __ldg(g.Wg + i);
  stage_tables(s, g, Gs, Ms);
  for (int t = 0; t < GTPB && n0 < s.N; ++t, n0 += GTM) {
    if (STAGED && t > 0) stage_rows(s, act, n0, As);
    if (STAGED || t == 0) wait_rows();
    group_tile<STAGED>(s, g, act, n0, gs, Gs, Ms, Gt, As);
    __syncthreads();
    if (groups != nullptr) {   // the tile's rows are one contiguous range of the (N x Gt) output
      const long long left = s.N - n0;
      const int cnt = (left < GTM ? (int)left : GTM) * Gt;
      float* dst = groups + n0 * Gt;
      for (int idx = threadIdx.x; idx < cnt; idx += blockDim.x) dst[idx] = gs[idx];
    }
    if (logits != nullptr) {
      for (int idx = threadIdx.x; idx < GTM * s.K; idx += blockDim.x) {
        const int m = idx / s.K, k = idx % s.K;
        const long long n = n0 + m;
        if (n >= s.N) continue;
        const float* w = Wp + (long long)k * Gt;
        const float* grow = gs + m * Gt;
        float a0 = 0.f, a1 = 0.f;
        int gg = 0;
        for (; gg + 2 <= Gt; gg += 2) {
          a0 = fmaf(grow[gg], w[gg], a0);
          a1 = fmaf(grow[gg + 1], w[gg + 1], a1);
        }
        if (gg < Gt) a0 = fmaf(grow[gg], w[gg], a0);
        logits[n * s.K + k] = a0 + a1;
      }
    }
    __syncthreads();   // gs and As are rewritten by the next tile
  }
}

constexpr size_t kGroupStagedMax = 75 * 1024;   // staged rows only while three CTAs per SM fit (227 KB / 3)

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
  const size_t ws_bytes = sizeof(float) * (size_t)s.K * Gt;    // last_layer_group in shared memory only when it is small
  const int ws_smem = smem + ws_bytes <= 32 * 1024 ? 1 : 0;
  if (ws_smem) smem += ws_bytes;
  const size_t rows_bytes = sizeof(float) * (size_t)GTM * s.P;
  const bool staged = smem + rows_bytes <= kGroupStagedMax;   // three CTAs per SM still fit
  long long tiles = (s.N + GTM - 1) / GTM;
  if (staged) {
    smem += rows_bytes;
    SPS_CUDA_TRY(cudaFuncSetAttribute(group_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    group_fwd_kernel<true><<<(unsigned)((tiles + GTPB - 1) / GTPB), group_threads(g), smem, st>>>(s, g, act, groups, logits, ws_smem);
  } else {
    SPS_CUDA_TRY(cudaFuncSetAttribute(group_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    group_fwd_kernel<false><<<(unsigned)((tiles + GTPB - 1) / GTPB), group_threads(g), smem, st>>>(s, g, act, groups, logits, ws_smem);
  }
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

// Backward pass 1: recompute groups, dpre = (dlogits Wg + dgroups) * groups, dact; spill groups and
// dpre to the workspace for the two reductions over n.
// groups_in: the group activations the forward returned (saved by the caller).  With them this pass reads nothing of size
// N x P: dpre and dact depend on the activations only through the groups, so the 205 MB activation map is not re-read
// and the projections are not recomputed (the dG kernel is then the only reader of the map in the whole backward).
template <bool STAGED>
__global__ void __launch_bounds__(GNT) group_bwd_pos_kernel(HeadShape s, SpsGroupDesc g, const float* __restrict__ act,
                                                             const float* __restrict__ groups_in,
                                                             const float* __restrict__ dlogits,
                                                             const float* __restrict__ dgroups,
                                                             float* __restrict__ dact, float* __restrict__ ws_groups,
                                                             float* __restrict__ ws_dpre, int ws_smem) {
  extern __shared__ __align__(16) float smem[];
  const int Gt = g.J * g.Kc;
  float* As = smem;
  float* gs = smem + (STAGED ? GTM * s.P : 0);
  float* dps = gs + GTM * Gt;
  float* Gs = gs + 2 * GTM * Gt;
  int* Ms = reinterpret_cast<int*>(Gs + gpad_floats(s, g));
  float* Ws = reinterpret_cast<float*>(Ms + s.P + g.Kc);   // last_layer_group [K][Gt] (when it fits: ws_smem)
  float* dls = Ws + (ws_smem ? s.K * Gt : 0);              // dlogits rows of the tile [GTM][K]
  const float* Wp = ws_smem ? Ws : g.Wg;
  long long n0 = (long long)blockIdx.x * GTPB * GTM;
  const bool recompute = groups_in == nullptr;
  if (STAGED && recompute) stage_rows(s, act, n0, As);
  if (dlogits != nullptr && ws_smem)
    for (int i = threadIdx.x; i < s.K * Gt; i += blockDim.x) Ws[i] = __ldg(g.Wg + i);
  stage_tables(s, g, Gs, Ms);
  for (int t = 0; t < GTPB && n0 < s.N; ++t, n0 += GTM) {
    if (dlogits != nullptr)
      for (int i = threadIdx.x; i < GTM * s.K; i += blockDim.x) dls[i] = n0 * s.K + i < s.N * s.K ? __ldg(dlogits + n0 * s.K + i) : 0.f;
    if (recompute) {
      if (STAGED && t > 0) stage_rows(s, act, n0, As);
      __syncthreads();           // the dlogits tile
      if (STAGED || t == 0) wait_rows();
      group_tile<STAGED>(s, g, act, n0, gs, Gs, Ms, Gt, As);
    } else {
      if (t == 0) wait_rows();   // the tables (cp.async)
      for (int idx = threadIdx.x; idx < GTM * Gt; idx += blockDim.x) gs[idx] = n0 * Gt + idx < s.N * Gt ? __ldg(groups_in + n0 * Gt + idx) : 0.f;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < GTM * Gt; idx += blockDim.x) {
      const int m = idx / Gt, gg = idx % Gt;
      const long long n = n0 + m;
      float v = 0.f;
      if (n < s.N) {
        float up = dgroups != nullptr ? __ldg(dgroups + n * Gt + gg) : 0.f;
        if (dlogits != nullptr)
          for (int k = 0; k < s.K; ++k) up = fmaf(dls[m * s.K + k], Wp[(long long)k * Gt + gg], up);
        v = up * gs[idx];
        if (recompute) ws_groups[n * Gt + gg] = gs[idx];
        ws_dpre[n * Gt + gg] = v;
      }
      dps[idx] = v;
    }
    __syncthreads();
    if (dact != nullptr) {
      // a thread owns prototypes p = tid, tid + blockDim, ...: its class slot and its J projection weights are fetched
      // once and reused for the 16 rows (a flat (row, prototype) loop paid a division, four index loads and J weight
      // loads per ELEMENT)
      for (int p = threadIdx.x; p < s.P; p += blockDim.x) {
        const int kc = g.proto_slot[p];
        float w[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] = 0.f;
        if (kc >= 0) {
          const int off = g.class_off[kc], nk = g.class_off[kc + 1] - off, col = g.proto_col[p];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (j < g.J) w[j] = Gs[g.J * off + kc + j * nk + col];
        }
        const float* dp = dps + (kc >= 0 ? g.J * kc : 0);
        for (int m = 0; m < GTM; ++m) {
          if (n0 + m >= s.N) break;
          float a = 0.f;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (j < g.J) a = fmaf(dp[m * Gt + j], w[j], a);
          dact[(n0 + m) * s.P + p] = kc >= 0 ? a : 0.f;
        }
      }
    }
    __syncthreads();   // gs, dps, dls and As are rewritten by the next tile
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
constexpr int GWR = 64;   // rows per staging step
constexpr int GWT = 256;  // threads = (4 classes x 2 groups) output tiles per CTA
// Each thread keeps a 4 (classes) x 2 (groups) tile of dWg in registers: per staged row one 16-byte and one 8-byte
// shared-memory read feed 8 FMAs (one output per thread needed two reads per FMA and the kernel sat on the LSU).
// Rows are staged zero-padded to Kp = 4*ceil(K/4) and Gp = 2*ceil(Gt/2) floats.
__global__ void __launch_bounds__(GWT) group_bwd_dWg_kernel(HeadShape s, int Gt, int gwr, const float* __restrict__ dlogits,
                                                            const float* __restrict__ ws_groups, float* __restrict__ dWg) {
  extern __shared__ __align__(16) float sm[];   // dl [gwr][Kp] | groups [gwr][Gp]   (gwr <= GWR rows per staging step)
  const int Kt = (s.K + 3) >> 2, Gh = (Gt + 1) >> 1, Kp = Kt << 2, Gp = Gh << 1;
  float* dls = sm;
  float* grs = sm + gwr * Kp;
  const long long per = (s.N + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x, hi = min(s.N, lo + per);
  const int tile = blockIdx.y * GWT + threadIdx.x;
  const bool live = tile < Kt * Gh;
  const int kt = live ? tile / Gh : 0, g2 = live ? tile - kt * Gh : 0;
  float acc[4][2];
#pragma unroll
  for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = 0.f;
  for (long long r0 = lo; r0 < hi; r0 += gwr) {
    const int rows = (int)min((long long)gwr, hi - r0);
    __syncthreads();
    for (int i = threadIdx.x; i < rows * Kp; i += blockDim.x) {
      const int m = i / Kp, k = i - m * Kp;
      dls[i] = k < s.K ? __ldg(dlogits + (r0 + m) * s.K + k) : 0.f;
    }
    for (int i = threadIdx.x; i < rows * Gp; i += blockDim.x) {
      const int m = i / Gp, gg = i - m * Gp;
      grs[i] = gg < Gt ? __ldg(ws_groups + (r0 + m) * Gt + gg) : 0.f;
    }
    __syncthreads();
    if (live) {
      const float4* dp = reinterpret_cast<const float4*>(dls) + kt;
      const float2* gp = reinterpret_cast<const float2*>(grs) + g2;
#pragma unroll 4
      for (int m = 0; m < rows; ++m) {
        const float4 d = dp[m * Kt];
        const float2 gv = gp[m * Gh];
        acc[0][0] = fmaf(d.x, gv.x, acc[0][0]), acc[0][1] = fmaf(d.x, gv.y, acc[0][1]);
        acc[1][0] = fmaf(d.y, gv.x, acc[1][0]), acc[1][1] = fmaf(d.y, gv.y, acc[1][1]);
        acc[2][0] = fmaf(d.z, gv.x, acc[2][0]), acc[2][1] = fmaf(d.z, gv.y, acc[2][1]);
        acc[3][0] = fmaf(d.w, gv.x, acc[3][0]), acc[3][1] = fmaf(d.w, gv.y, acc[3][1]);
      }
    }
  }
  if (live && lo < hi) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int k = 4 * kt + i, gg = 2 * g2 + j;
        if (k < s.K && gg < Gt) atomicAdd(dWg + (long long)k * Gt + gg, acc[i][j]);
      }
  }
}

int launch_group_bwd(const HeadShape& s, const SpsGroupDesc& g, const float* act, const float* groups_in, const float* dlogits,
                     const float* dgroups, float* dact, float* dG, float* dWg, float* ws, cudaStream_t st) {
  const int Gt = g.J * g.Kc;
  if (dG != nullptr) SPS_CUDA_TRY(cudaMemsetAsync(dG, 0, sizeof(float) * (size_t)g.J * s.P, st));
  if (dWg != nullptr) SPS_CUDA_TRY(cudaMemsetAsync(dWg, 0, sizeof(float) * (size_t)s.K * Gt, st));
  if (s.N == 0 || g.Kc == 0) {
    if (dact != nullptr && s.N > 0) SPS_CUDA_TRY(cudaMemsetAsync(dact, 0, sizeof(float) * (size_t)s.N * s.P, st));
    return SPS_OK;
  }
  size_t smem = sizeof(float) * (2 * (size_t)GTM * Gt + (size_t)g.J * s.P + g.Kc + s.P + g.Kc + (size_t)GTM * s.K);
  const size_t ws_bytes = sizeof(float) * (size_t)s.K * Gt;    // last_layer_group in shared memory only when it is small
  const int ws_smem = smem + ws_bytes <= 32 * 1024 ? 1 : 0;
  if (ws_smem) smem += ws_bytes;
  SPS_REQUIRE(smem <= 200 * 1024, SPS_ERR_UNSUPPORTED, "group head: J*Kc=%d, J*P=%d need %zu bytes of shared memory", Gt, g.J * s.P, smem);
  float* ws_groups = ws;
  float* ws_dpre = ws + (size_t)s.N * Gt;
  const float* groups_rd = groups_in != nullptr ? groups_in : ws_groups;   // what the dWg kernel reads
  long long tiles = (s.N + GTM - 1) / GTM;
  const size_t rows_bytes = sizeof(float) * (size_t)GTM * s.P;
  if (groups_in == nullptr && smem + rows_bytes <= kGroupStagedMax) {
    smem += rows_bytes;
    SPS_CUDA_TRY(cudaFuncSetAttribute(group_bwd_pos_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    group_bwd_pos_kernel<true><<<(unsigned)((tiles + GTPB - 1) / GTPB), group_threads(g), smem, st>>>(s, g, act, groups_in, dlogits, dgroups, dact, ws_groups, ws_dpre, ws_smem);
  } else {
    SPS_CUDA_TRY(cudaFuncSetAttribute(group_bwd_pos_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    group_bwd_pos_kernel<false><<<(unsigned)((tiles + GTPB - 1) / GTPB), group_threads(g), smem, st>>>(s, g, act, groups_in, dlogits, dgroups, dact, ws_groups, ws_dpre, ws_smem);
  }
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
    const int Kt = (s.K + 3) / 4, Gh = (Gt + 1) / 2;
    const unsigned ny = (unsigned)((Kt * Gh + GWT - 1) / GWT);              // y: chunks of 256 output tiles
    long long wsplit = (s.N + 2 * GWR - 1) / (2 * GWR);                      // >= 128 rows per CTA
    const long long cap = (4LL * sm_count() + ny - 1) / ny;                  // about four CTAs per SM in all
    if (wsplit > cap) wsplit = cap;
    if (wsplit < 1) wsplit = 1;
    dim3 grid((unsigned)wsplit, ny);
    const size_t row_bytes = sizeof(float) * (size_t)(4 * Kt + 2 * Gh);
    int gwr = GWR;                                                           // fewer rows per step for wide heads (COCO-Stuff)
    while (gwr > 8 && gwr * row_bytes > 96 * 1024) gwr >>= 1;
    SPS_CUDA_TRY(cudaFuncSetAttribute(group_bwd_dWg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(gwr * row_bytes)));
    group_bwd_dWg_kernel<<<grid, GWT, gwr * row_bytes, st>>>(s, Gt, gwr, dlogits, groups_rd, dWg);
    count_launch();
    SPS_CUDA_TRY(cudaGetLastError());
  }
  return SPS_OK;
}

}  // namespace sps
