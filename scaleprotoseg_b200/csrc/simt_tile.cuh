// Shared-memory tile helpers of the fp32 (CUDA-core) kernels: 64 positions x 64 prototypes per step.
#pragma once
#include "common.cuh"

namespace sps {

constexpr int TM = 64;   // positions per tile
constexpr int TN = 64;   // prototypes per chunk
constexpr int NT = 256;  // threads
constexpr int ZS = 68;   // row stride of the z tile (floats): 16B aligned, spreads banks for column reads
constexpr int MAXD = 64;

struct FwdSmem {
  float* zt;   // [MAXD][ZS]    zt[c][m]
  float* pt;   // [MAXD][TN]    pt[c][j]
  float* at;   // [TM][TN+1]    at[m][j]
  float* dt;   // [TN][TM+1]    dt[j][m]
  float* ws;   // [K][TN]
  float* zn;   // [TM]
  float* pn;   // [TN]
  long long* zoff;  // [TM] offset of (b, pix) into z: b*C*HW + pix ; -1 when n >= N
  long long* doff;  // [TM] offset into dist: b*P*HW + pix
};

__host__ __device__ inline size_t fwd_smem_bytes(int K) {
  return sizeof(float) * (MAXD * ZS + MAXD * TN + TM * (TN + 1) + TN * (TM + 1) + (size_t)K * TN + TM + TN) +
         sizeof(long long) * 2 * TM + 16;
}

__device__ __forceinline__ FwdSmem carve(float* base, int K) {
  FwdSmem s;
  s.zoff = reinterpret_cast<long long*>(base);
  s.doff = s.zoff + TM;
  float* f = reinterpret_cast<float*>(s.doff + TM);
  s.zt = f;
  s.pt = s.zt + MAXD * ZS;
  s.at = s.pt + MAXD * TN;
  s.dt = s.at + TM * (TN + 1);
  s.ws = s.dt + TN * (TM + 1);
  s.zn = s.ws + (size_t)K * TN;
  s.pn = s.zn + TM;
  return s;
}

// Loads the z tile of one scale (zero beyond N) and the squared norms of its rows.
__device__ __forceinline__ void load_z_tile(const HeadShape& s, const float* __restrict__ z, int scale,
                                            const FwdSmem& sm) {
  const int t = threadIdx.x;
  for (int idx = t; idx < s.D * TM; idx += NT) {
    int c = idx / TM, m = idx % TM;
    long long off = sm.zoff[m];
    sm.zt[c * ZS + m] = off >= 0 ? __ldg(z + off + (long long)(scale * s.D + c) * s.HW) : 0.0f;
  }
  __syncthreads();
  if (t < TM) {
    float acc = 0.f;
    for (int c = 0; c < s.D; ++c) {
      float v = sm.zt[c * ZS + t];
      acc = fmaf(v, v, acc);
    }
    sm.zn[t] = acc;
  }
}

// Loads prototype rows [p0, p0+TN) clipped to hi (zero beyond) transposed into pt[c][j], and norms.
__device__ __forceinline__ void load_p_chunk(const HeadShape& s, const float* __restrict__ protos, int p0, int hi,
                                             const FwdSmem& sm) {
  const int t = threadIdx.x;
  for (int idx = t; idx < s.D * TN; idx += NT) {
    int j = idx % TN, c = idx / TN;
    sm.pt[c * TN + j] = (p0 + j < hi) ? __ldg(protos + (long long)(p0 + j) * s.D + c) : 0.0f;
  }
  __syncthreads();
  if (t < TN) {
    float acc = 0.f;
    for (int c = 0; c < s.D; ++c) {
      float v = sm.pt[c * TN + t];
      acc = fmaf(v, v, acc);
    }
    sm.pn[t] = acc;
  }
}

// acc[i][j] = sum_c zt[c][tx*4+i] * pt[c][ty*4+j]
__device__ __forceinline__ void zp_tile(const HeadShape& s, const FwdSmem& sm, int tx, int ty, float acc[4][4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int c = 0; c < s.D; ++c) {
    float4 zv = *reinterpret_cast<const float4*>(sm.zt + c * ZS + tx * 4);
    float4 pv = *reinterpret_cast<const float4*>(sm.pt + c * TN + ty * 4);
    float za[4] = {zv.x, zv.y, zv.z, zv.w}, pa[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(za[i], pa[j], acc[i][j]);
  }
}

__device__ __forceinline__ void init_offsets(const HeadShape& s, long long n0, const FwdSmem& sm) {
  if (threadIdx.x < TM) {
    long long n = n0 + threadIdx.x;
    if (n < s.N) {
      long long b = n / s.HW, pix = n % s.HW;
      sm.zoff[threadIdx.x] = b * s.C * s.HW + pix;
      sm.doff[threadIdx.x] = b * s.P * s.HW + pix;
    } else {
      sm.zoff[threadIdx.x] = -1;
      sm.doff[threadIdx.x] = -1;
    }
  }
  __syncthreads();
}

}  // namespace sps
