// Exact-operand (fp32 FFMA) prototype head: forward.  SPS_MATH_FP32.
//
// One CTA owns a tile of TM=64 positions and walks every scale and every 64-prototype chunk of that
// scale: smem-tiled z.p contraction -> d = relu(|z|^2 - 2 z.p + |p|^2) -> a = log((d+1)/(d+eps))
// -> logits += a . W^T, with optional coalesced stores of d (NCHW) and a (N x P).
// Follows segmentation/model/model_multiscale.py:256-281 (distance), :283-317 (scale split),
// :324-330 (similarity), :243-244 (last layer) of the reference.
#include "simt_tile.cuh"

namespace sps {

template <int KACC>
__global__ void __launch_bounds__(NT) head_fwd_simt_kernel(HeadShape s, const float* __restrict__ z,
                                                            const float* __restrict__ protos,
                                                            const float* __restrict__ last_w,
                                                            float* __restrict__ logits, float* __restrict__ dist,
                                                            float* __restrict__ act) {
  extern __shared__ __align__(16) float smem_raw[];
  FwdSmem sm = carve(smem_raw, s.K);
  const int t = threadIdx.x, tx = t % 16, ty = t / 16;
  const long long n0 = (long long)blockIdx.x * TM;
  init_offsets(s, n0, sm);

  float lacc[KACC];
#pragma unroll
  for (int q = 0; q < KACC; ++q) lacc[q] = 0.f;
  const int lm = t % TM, lkq = t / TM;  // logits: position lm, classes lkq + 4*q

  for (int scale = 0; scale < s.S; ++scale) {
    const int lo = s.scale_lo[scale], hi = s.scale_lo[scale + 1];
    if (hi <= lo) continue;
    __syncthreads();
    load_z_tile(s, z, scale, sm);
    for (int p0 = lo; p0 < hi; p0 += TN) {
      __syncthreads();  // previous chunk's readers of pt/at/dt/ws are done
      load_p_chunk(s, protos, p0, hi, sm);
      if (last_w != nullptr) {
        for (int idx = t; idx < s.K * TN; idx += NT) {
          int k = idx / TN, j = idx % TN;
          sm.ws[idx] = (p0 + j < hi) ? __ldg(last_w + (long long)k * s.P + p0 + j) : 0.0f;
        }
      }
      __syncthreads();
      float acc[4][4];
      zp_tile(s, sm, tx, ty, acc);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int m = tx * 4 + i;
        const float znm = sm.zn[m];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int jj = ty * 4 + j;
          float r = znm + (-2.0f * acc[i][j] + sm.pn[jj]);
          float d = r > 0.f ? r : 0.f;
          sm.at[m * (TN + 1) + jj] = act_exact(d, s.act_kind, s.eps);
          sm.dt[jj * (TM + 1) + m] = d;
        }
      }
      __syncthreads();
      const int nvalid = min(TN, hi - p0);
      if (act != nullptr) {
        for (int idx = t; idx < TM * TN; idx += NT) {
          int m = idx / TN, j = idx % TN;
          if (j < nvalid && n0 + m < s.N) act[(n0 + m) * s.P + p0 + j] = sm.at[m * (TN + 1) + j];
        }
      }
      if (dist != nullptr) {
        for (int idx = t; idx < TM * TN; idx += NT) {
          int j = idx / TM, m = idx % TM;
          long long off = sm.doff[m];
          if (j < nvalid && off >= 0) dist[off + (long long)(p0 + j) * s.HW] = sm.dt[j * (TM + 1) + m];
        }
      }
      if (last_w != nullptr) {
#pragma unroll
        for (int q = 0; q < KACC; ++q) {
          const int k = lkq + 4 * q;
          if (k < s.K) {
            float a = lacc[q];
            const float* wrow = sm.ws + k * TN;
            const float* arow = sm.at + lm * (TN + 1);
#pragma unroll 8
            for (int j = 0; j < TN; ++j) a = fmaf(arow[j], wrow[j], a);
            lacc[q] = a;
          }
        }
      }
    }
  }
  if (last_w != nullptr && n0 + lm < s.N) {
#pragma unroll
    for (int q = 0; q < KACC; ++q) {
      const int k = lkq + 4 * q;
      if (k < s.K) logits[(n0 + lm) * s.K + k] = lacc[q];
    }
  }
}

template <int KACC>
static int launch_fwd_t(const HeadShape& s, const float* z, const float* protos, const float* last_w, float* logits,
                        float* dist, float* act, cudaStream_t st) {
  size_t smem = fwd_smem_bytes(s.K);
  SPS_CUDA_TRY(cudaFuncSetAttribute(head_fwd_simt_kernel<KACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long tiles = (s.N + TM - 1) / TM;
  head_fwd_simt_kernel<KACC><<<(unsigned)tiles, NT, smem, st>>>(s, z, protos, last_w, logits, dist, act);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

int launch_head_fwd_simt(const HeadShape& s, const float* z, const float* protos, const float* last_w, float* logits,
                         float* dist, float* act, cudaStream_t st) {
  SPS_REQUIRE(s.D <= MAXD && s.D % 4 == 0, SPS_ERR_UNSUPPORTED, "fp32 head: D=%d must be a multiple of 4, <= %d", s.D, MAXD);
  SPS_REQUIRE(s.K <= 192, SPS_ERR_UNSUPPORTED, "fp32 head: K=%d > 192 classes", s.K);
  if (s.N == 0 || s.P == 0) return SPS_OK;
  if (s.K <= 20) return launch_fwd_t<5>(s, z, protos, last_w, logits, dist, act, st);
  if (s.K <= 64) return launch_fwd_t<16>(s, z, protos, last_w, logits, dist, act, st);
  return launch_fwd_t<48>(s, z, protos, last_w, logits, dist, act, st);
}


// ------------------------------------------------------------------------------------------------
// Backward (what autograd derives from model_multiscale.py:256-281, 324-330, 243-244).
//   r = |z|^2 - 2 z.p + |p|^2, d = relu(r), a = act(d), logits = a W^T
//   gamma[n,p] = [r>0] * ((dlogits W + dact)[n,p] * a'(d) + ddist[n,p])
//   dz_s[n]  = 2 (z_s[n] * sum_p gamma - sum_p gamma p_p)          (position-stationary kernel)
//   dp_p     = 2 (p_p * sum_n gamma - sum_n gamma z_s[n])          (prototype-stationary kernel)
//   dW[k,p]  = sum_n dlogits[n,k] a[n,p]
// Reductions over n go through per-split partials in the workspace and a fixed-order second pass,
// so the result is deterministic.
// ------------------------------------------------------------------------------------------------
struct BwdSmem {
  FwdSmem f;
  float* dls;  // [K][ZS]  dlogits tile transposed: dls[k][m]
  float* gts;  // [TN][ZS] gamma tile: gts[j][m]
};

__host__ __device__ inline size_t bwd_smem_bytes(int K) {
  return fwd_smem_bytes(K) + sizeof(float) * ((size_t)K * ZS + TN * ZS);
}

__device__ __forceinline__ BwdSmem carve_bwd(float* base, int K) {
  BwdSmem b;
  b.f = carve(base, K);
  b.dls = b.f.pn + TN;
  b.gts = b.dls + (size_t)K * ZS;
  return b;
}

__device__ __forceinline__ void load_dl_tile(const HeadShape& s, const float* __restrict__ dlogits, long long n0,
                                             const BwdSmem& sm) {
  for (int idx = threadIdx.x; idx < TM * s.K; idx += NT) {
    int m = idx / s.K, k = idx % s.K;
    sm.dls[k * ZS + m] = (n0 + m < s.N) ? __ldg(dlogits + (n0 + m) * s.K + k) : 0.0f;
  }
}

__device__ __forceinline__ void load_w_chunk(const HeadShape& s, const float* __restrict__ last_w, int p0, int hi,
                                             const FwdSmem& sm) {
  for (int idx = threadIdx.x; idx < s.K * TN; idx += NT) {
    int k = idx / TN, j = idx % TN;
    sm.ws[idx] = (p0 + j < hi) ? __ldg(last_w + (long long)k * s.P + p0 + j) : 0.0f;
  }
}

// Recomputes the (TM x TN) tile and leaves gamma in gts[j][m]; optionally a in at[m][j].
__device__ __forceinline__ void gamma_tile(const HeadShape& s, const BwdSmem& sm, long long n0, int p0, int nvalid,
                                           bool has_w, const float* __restrict__ ddist,
                                           const float* __restrict__ dact, bool want_act, int tx, int ty) {
  float acc[4][4], dA[4][4];
  zp_tile(s, sm.f, tx, ty, acc);
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) dA[i][j] = 0.f;
  if (has_w) {
    for (int k = 0; k < s.K; ++k) {
      float4 dv = *reinterpret_cast<const float4*>(sm.dls + k * ZS + tx * 4);
      float4 wv = *reinterpret_cast<const float4*>(sm.f.ws + k * TN + ty * 4);
      float da[4] = {dv.x, dv.y, dv.z, dv.w}, wa[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dA[i][j] = fmaf(da[i], wa[j], dA[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = tx * 4 + i;
    const bool mv = n0 + m < s.N;
    const float znm = sm.f.zn[m];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int jj = ty * 4 + j;
      float g = 0.f, a = 0.f;
      if (mv && jj < nvalid) {
        float r = znm + (-2.0f * acc[i][j] + sm.f.pn[jj]);
        float d = r > 0.f ? r : 0.f;
        a = act_exact(d, s.act_kind, s.eps);
        if (r > 0.f) {
          float up = dA[i][j];
          if (dact != nullptr) up += __ldg(dact + (n0 + m) * s.P + p0 + jj);
          g = up * dact_dd(d, s.act_kind, s.eps);
          if (ddist != nullptr) g += __ldg(ddist + sm.f.doff[m] + (long long)(p0 + jj) * s.HW);
        }
      }
      sm.gts[jj * ZS + m] = g;
      if (want_act) sm.f.at[m * (TN + 1) + jj] = a;
    }
  }
}

__global__ void __launch_bounds__(NT) head_bwd_dz_kernel(HeadShape s, const float* __restrict__ z,
                                                          const float* __restrict__ protos,
                                                          const float* __restrict__ last_w,
                                                          const float* __restrict__ dlogits,
                                                          const float* __restrict__ ddist,
                                                          const float* __restrict__ dact, float* __restrict__ dz) {
  extern __shared__ __align__(16) float smem_raw[];
  BwdSmem sm = carve_bwd(smem_raw, s.K);
  const int t = threadIdx.x, tx = t % 16, ty = t / 16;
  const long long n0 = (long long)blockIdx.x * TM;
  const bool has_w = last_w != nullptr && dlogits != nullptr;
  for (int idx = t; idx < MAXD * TN; idx += NT) sm.f.pt[idx] = 0.f;  // rows c >= D stay zero
  init_offsets(s, n0, sm.f);
  if (has_w) load_dl_tile(s, dlogits, n0, sm);

  for (int scale = 0; scale < s.S; ++scale) {
    const int lo = s.scale_lo[scale], hi = s.scale_lo[scale + 1];
    __syncthreads();
    load_z_tile(s, z, scale, sm.f);
    float dzacc[4][4], sg[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      sg[i] = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) dzacc[i][j] = 0.f;
    }
    for (int p0 = lo; p0 < hi; p0 += TN) {
      __syncthreads();
      load_p_chunk(s, protos, p0, hi, sm.f);
      if (has_w) load_w_chunk(s, last_w, p0, hi, sm.f);
      __syncthreads();
      gamma_tile(s, sm, n0, p0, min(TN, hi - p0), has_w, ddist, dact, false, tx, ty);
      __syncthreads();
      // dzacc[i][cc] += sum_j gamma[m_i][j] * p[j][c_cc],  c = ty*4+cc
      for (int j = 0; j < TN; ++j) {
        float4 gv = *reinterpret_cast<const float4*>(sm.gts + j * ZS + tx * 4);
        float ga[4] = {gv.x, gv.y, gv.z, gv.w};
        float pc[4];
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) pc[cc] = sm.f.pt[(ty * 4 + cc) * TN + j];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          sg[i] += ga[i];
#pragma unroll
          for (int cc = 0; cc < 4; ++cc) dzacc[i][cc] = fmaf(ga[i], pc[cc], dzacc[i][cc]);
        }
      }
    }
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      const int c = ty * 4 + cc;
      if (c < s.D) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int m = tx * 4 + i;
          long long off = sm.f.zoff[m];
          if (off >= 0)
            dz[off + (long long)(scale * s.D + c) * s.HW] = 2.0f * (sm.f.zt[c * ZS + m] * sg[i] - dzacc[i][cc]);
        }
      }
    }
  }
}

// blockIdx.x -> (scale, p0) of the chunk this CTA owns
__device__ __forceinline__ bool chunk_of(const HeadShape& s, int chunk, int& scale, int& p0, int& hi) {
  for (int sc = 0; sc < s.S; ++sc) {
    int lo = s.scale_lo[sc], h = s.scale_lo[sc + 1];
    int nch = (h - lo + TN - 1) / TN;
    if (chunk < nch) {
      scale = sc;
      p0 = lo + chunk * TN;
      hi = h;
      return true;
    }
    chunk -= nch;
  }
  return false;
}

inline int count_chunks(const HeadShape& s) {
  int n = 0;
  for (int sc = 0; sc < s.S; ++sc) n += (s.scale_lo[sc + 1] - s.scale_lo[sc] + TN - 1) / TN;
  return n;
}

template <int KACC>
__global__ void __launch_bounds__(NT) head_bwd_dp_kernel(HeadShape s, const float* __restrict__ z,
                                                          const float* __restrict__ protos,
                                                          const float* __restrict__ last_w,
                                                          const float* __restrict__ dlogits,
                                                          const float* __restrict__ ddist,
                                                          const float* __restrict__ dact, float* __restrict__ part_dp,
                                                          float* __restrict__ part_dw) {
  extern __shared__ __align__(16) float smem_raw[];
  BwdSmem sm = carve_bwd(smem_raw, s.K);
  const int t = threadIdx.x, tx = t % 16, ty = t / 16;
  int scale, p0, hi;
  if (!chunk_of(s, blockIdx.x, scale, p0, hi)) return;
  const int nvalid = min(TN, hi - p0);
  const bool has_w = last_w != nullptr && dlogits != nullptr;
  const bool want_dw = has_w && part_dw != nullptr;
  for (int idx = t; idx < MAXD * TN; idx += NT) sm.f.pt[idx] = 0.f;
  for (int idx = t; idx < MAXD * ZS; idx += NT) sm.f.zt[idx] = 0.f;  // rows c >= D stay zero
  __syncthreads();
  load_p_chunk(s, protos, p0, hi, sm.f);
  if (has_w) load_w_chunk(s, last_w, p0, hi, sm.f);

  float dpacc[4][4], sgp[4], wacc[KACC];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    sgp[j] = 0.f;
#pragma unroll
    for (int c = 0; c < 4; ++c) dpacc[j][c] = 0.f;
  }
#pragma unroll
  for (int q = 0; q < KACC; ++q) wacc[q] = 0.f;
  const int wj = t % TN, wkq = t / TN;

  for (long long n0 = (long long)blockIdx.y * TM; n0 < s.N; n0 += (long long)gridDim.y * TM) {
    __syncthreads();
    init_offsets(s, n0, sm.f);
    load_z_tile(s, z, scale, sm.f);
    if (has_w) load_dl_tile(s, dlogits, n0, sm);
    __syncthreads();
    gamma_tile(s, sm, n0, p0, nvalid, has_w, ddist, dact, want_dw, tx, ty);
    __syncthreads();
    // dpacc[jj][ii] += sum_m gamma[m][ty*4+jj] * z[m][tx+16*ii]
    for (int m4 = 0; m4 < TM / 4; ++m4) {
      float4 gv[4], zv[4];
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) gv[jj] = *reinterpret_cast<const float4*>(sm.gts + (ty * 4 + jj) * ZS + m4 * 4);
#pragma unroll
      for (int ii = 0; ii < 4; ++ii) zv[ii] = *reinterpret_cast<const float4*>(sm.f.zt + (tx + 16 * ii) * ZS + m4 * 4);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        sgp[jj] += (gv[jj].x + gv[jj].y) + (gv[jj].z + gv[jj].w);
#pragma unroll
        for (int ii = 0; ii < 4; ++ii) {
          float a = dpacc[jj][ii];
          a = fmaf(gv[jj].x, zv[ii].x, a);
          a = fmaf(gv[jj].y, zv[ii].y, a);
          a = fmaf(gv[jj].z, zv[ii].z, a);
          a = fmaf(gv[jj].w, zv[ii].w, a);
          dpacc[jj][ii] = a;
        }
      }
    }
    if (want_dw) {
#pragma unroll
      for (int q = 0; q < KACC; ++q) {
        const int k = wkq + 4 * q;
        if (k < s.K) {
          float a = wacc[q];
          for (int m = 0; m < TM; ++m) a = fmaf(sm.dls[k * ZS + m], sm.f.at[m * (TN + 1) + wj], a);
          wacc[q] = a;
        }
      }
    }
  }
  if (part_dp != nullptr) {
    float* out = part_dp + (long long)blockIdx.y * s.P * s.D;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const int j = ty * 4 + jj;
      if (j < nvalid) {
#pragma unroll
        for (int ii = 0; ii < 4; ++ii) {
          const int c = tx + 16 * ii;
          if (c < s.D) out[(long long)(p0 + j) * s.D + c] = 2.0f * (sm.f.pt[c * TN + j] * sgp[jj] - dpacc[jj][ii]);
        }
      }
    }
  }
  if (want_dw && wj < nvalid) {
    float* out = part_dw + (long long)blockIdx.y * s.K * s.P;
#pragma unroll
    for (int q = 0; q < KACC; ++q) {
      const int k = wkq + 4 * q;
      if (k < s.K) out[(long long)k * s.P + p0 + wj] = wacc[q];
    }
  }
}

__global__ void reduce_partials_kernel(const float* __restrict__ part, int nsplit, long long count,
                                       float* __restrict__ out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  float a = 0.f;
  for (int sp = 0; sp < nsplit; ++sp) a += part[(long long)sp * count + i];
  out[i] = a;
}

static int bwd_splits(const HeadShape& s) {
  int nch = count_chunks(s);
  if (nch == 0) return 1;
  long long tiles = (s.N + TM - 1) / TM;
  long long want = (2LL * sm_count() + nch - 1) / nch;
  if (want > tiles) want = tiles;
  if (want < 1) want = 1;
  return (int)want;
}

size_t head_bwd_simt_workspace(const HeadShape& s, int /*flags*/) {
  return sizeof(float) * (size_t)bwd_splits(s) * ((size_t)s.P * s.D + (size_t)s.K * s.P);
}

template <int KACC>
static int launch_dp_t(const HeadShape& s, int nsplit, const float* z, const float* protos, const float* last_w,
                       const float* dlogits, const float* ddist, const float* dact, float* part_dp, float* part_dw,
                       cudaStream_t st) {
  size_t smem = bwd_smem_bytes(s.K);
  SPS_CUDA_TRY(cudaFuncSetAttribute(head_bwd_dp_kernel<KACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(count_chunks(s), nsplit);
  head_bwd_dp_kernel<KACC><<<grid, NT, smem, st>>>(s, z, protos, last_w, dlogits, ddist, dact, part_dp, part_dw);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

int launch_head_bwd_simt(const HeadShape& s, int flags, const float* z, const float* protos, const float* last_w,
                         const float* dlogits, const float* ddist, const float* dact, float* dz, float* dprotos,
                         float* dlast_w, void* ws, size_t ws_bytes, cudaStream_t st) {
  SPS_REQUIRE(s.D <= MAXD && s.D % 4 == 0, SPS_ERR_UNSUPPORTED, "fp32 head: D=%d must be a multiple of 4, <= %d", s.D, MAXD);
  SPS_REQUIRE(s.K <= 192, SPS_ERR_UNSUPPORTED, "fp32 head: K=%d > 192 classes", s.K);
  if (s.N == 0 || s.P == 0) return SPS_OK;
  size_t smem = bwd_smem_bytes(s.K);
  if (dz != nullptr) {
    SPS_CUDA_TRY(cudaFuncSetAttribute(head_bwd_dz_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long tiles = (s.N + TM - 1) / TM;
    head_bwd_dz_kernel<<<(unsigned)tiles, NT, smem, st>>>(s, z, protos, last_w, dlogits, ddist, dact, dz);
    count_launch();
    SPS_CUDA_TRY(cudaGetLastError());
  }
  const bool want_dw = dlast_w != nullptr && last_w != nullptr && dlogits != nullptr;
  if (dlast_w != nullptr && !want_dw) SPS_CUDA_TRY(cudaMemsetAsync(dlast_w, 0, sizeof(float) * (size_t)s.K * s.P, st));
  if (dprotos != nullptr || want_dw) {
    const int nsplit = bwd_splits(s);
    size_t need = head_bwd_simt_workspace(s, flags);
    SPS_REQUIRE(ws != nullptr && ws_bytes >= need, SPS_ERR_WORKSPACE, "head_bwd workspace: need %zu bytes, got %zu", need,
                ws_bytes);
    float* part_dp = reinterpret_cast<float*>(ws);
    float* part_dw = part_dp + (size_t)nsplit * s.P * s.D;
    int rc;
    float* pdp = dprotos != nullptr ? part_dp : nullptr;
    float* pdw = want_dw ? part_dw : nullptr;
    if (s.K <= 20) rc = launch_dp_t<5>(s, nsplit, z, protos, last_w, dlogits, ddist, dact, pdp, pdw, st);
    else if (s.K <= 64) rc = launch_dp_t<16>(s, nsplit, z, protos, last_w, dlogits, ddist, dact, pdp, pdw, st);
    else rc = launch_dp_t<48>(s, nsplit, z, protos, last_w, dlogits, ddist, dact, pdp, pdw, st);
    if (rc != SPS_OK) return rc;
    if (dprotos != nullptr) {
      long long cnt = (long long)s.P * s.D;
      reduce_partials_kernel<<<(unsigned)((cnt + 255) / 256), 256, 0, st>>>(part_dp, nsplit, cnt, dprotos);
      count_launch();
    }
    if (want_dw) {
      long long cnt = (long long)s.K * s.P;
      reduce_partials_kernel<<<(unsigned)((cnt + 255) / 256), 256, 0, st>>>(part_dw, nsplit, cnt, dlast_w);
      count_launch();
    }
    SPS_CUDA_TRY(cudaGetLastError());
  }
  return SPS_OK;
}

}  // namespace sps
