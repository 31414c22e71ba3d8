// Push / nearest-patch projection: masked per-prototype argmin per map, dataset argmin, gather.
// Follows segmentation/push_multiscale_optimization.py:68-91 (compute_distances), :135-137
// (min_across_dataset) and :161-188 (global_min) of the reference.
//
// Key packing: (float_bits(d) << 32) | index.  d >= 0 after the relu, so unsigned order == float
// order, and the low word breaks ties towards the smallest index: exactly `.min(dim=-1).indices`
// (first occurrence) inside a map and `argmin(dim=0)` (first map) across the dataset.
// A position whose label does not match the prototype's class gets d + 1e10 in the reference, which
// rounds to 1e10f for every d < 512 (D <= 64 and z, p in [0,1] give d <= 64), so unmatched positions
// never beat the initial key (bits(1e10f) << 32) | 0 — also the reference's answer (value 1e10,
// index 0) for a class that is absent from the map.
#include <cuda_fp16.h>

#include "simt_tile.cuh"

namespace sps {

__device__ __forceinline__ unsigned long long absent_key() {
  return ((unsigned long long)__float_as_uint(1e10f)) << 32;
}

__global__ void fill_keys_kernel(unsigned long long* keys, long long count, unsigned long long v) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) keys[i] = v;
}

// grid: (tiles over H*W, B).  Tile = 64 positions of ONE map.
__global__ void __launch_bounds__(NT) push_scan_simt_kernel(HeadShape s, const float* __restrict__ z,
                                                             const float* __restrict__ protos,
                                                             const int16_t* __restrict__ labels,
                                                             const int32_t* __restrict__ proto_class,
                                                             unsigned long long* __restrict__ keys) {
  extern __shared__ __align__(16) float smem_raw[];
  FwdSmem sm = carve(smem_raw, 0);
  __shared__ int lab[TM];
  __shared__ unsigned long long kmin[TN];
  const int t = threadIdx.x, tx = t % 16, ty = t / 16;
  const int b = blockIdx.y;
  const int pix0 = blockIdx.x * TM;
  if (t < TM) {
    const int pix = pix0 + t;
    if (pix < s.HW) {
      sm.zoff[t] = (long long)b * s.C * s.HW + pix;
      lab[t] = labels[(long long)b * s.HW + pix];
    } else {
      sm.zoff[t] = -1;
      lab[t] = 0;  // void: never matches
    }
  }
  __syncthreads();
  for (int scale = 0; scale < s.S; ++scale) {
    const int lo = s.scale_lo[scale], hi = s.scale_lo[scale + 1];
    if (hi <= lo) continue;
    __syncthreads();
    load_z_tile(s, z, scale, sm);
    for (int p0 = lo; p0 < hi; p0 += TN) {
      __syncthreads();
      load_p_chunk(s, protos, p0, hi, sm);
      if (t < TN) kmin[t] = absent_key();
      __syncthreads();
      float acc[4][4];
      zp_tile(s, sm, tx, ty, acc);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int jj = ty * 4 + j;
        const int cls = (p0 + jj < hi) ? __ldg(proto_class + p0 + jj) + 1 : -1;
        unsigned long long best = absent_key();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int m = tx * 4 + i;
          if (lab[m] == cls) {
            float r = sm.zn[m] + (-2.0f * acc[i][j] + sm.pn[jj]);
            float d = r > 0.f ? r : 0.f;
            unsigned long long k = ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)(pix0 + m);
            best = k < best ? k : best;
          }
        }
        // the 16 lanes that share ty hold the other positions of this prototype
#pragma unroll
        for (int sh = 8; sh >= 1; sh >>= 1) {
          unsigned long long o = __shfl_xor_sync(0xffffffffu, best, sh);
          best = o < best ? o : best;
        }
        if (tx == 0) kmin[jj] = best;
      }
      __syncthreads();
      if (t < TN && p0 + t < hi && kmin[t] < absent_key()) atomicMin(keys + (long long)b * s.P + p0 + t, kmin[t]);
    }
  }
}

int launch_push_fill(unsigned long long* keys, long long count, cudaStream_t st) {
  if (count == 0) return SPS_OK;
  const unsigned long long init = ((unsigned long long)0x501502F9u) << 32;  // bits(1e10f)
  fill_keys_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(keys, count, init);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

int launch_push_scan_simt(const HeadShape& s, const float* z, const float* protos, const int16_t* labels,
                          const int32_t* proto_class, unsigned long long* keys, cudaStream_t st) {
  SPS_REQUIRE(s.D <= MAXD && s.D % 4 == 0, SPS_ERR_UNSUPPORTED, "fp32 push: D=%d must be a multiple of 4, <= %d", s.D, MAXD);
  long long count = (long long)s.B * s.P;
  if (count == 0) return SPS_OK;
  const unsigned long long init = ((unsigned long long)0x501502F9u) << 32;  // bits(1e10f)
  fill_keys_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(keys, count, init);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  if (s.HW == 0) return SPS_OK;
  size_t smem = fwd_smem_bytes(0);
  SPS_CUDA_TRY(cudaFuncSetAttribute(push_scan_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((s.HW + TM - 1) / TM, s.B);
  push_scan_simt_kernel<<<grid, NT, smem, st>>>(s, z, protos, labels, proto_class, keys);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

// best[p] = min(best[p], min_m (keys[m][p] & ~0xffffffff) | (map_base + m))
__global__ void push_reduce_kernel(const unsigned long long* __restrict__ keys, int n_maps, int P, uint32_t map_base,
                                   unsigned long long* __restrict__ best) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const int per = (n_maps + gridDim.y - 1) / gridDim.y;
  const int lo = per * blockIdx.y, hi = min(n_maps, lo + per);
  unsigned long long b = ~0ULL;
  for (int m = lo; m < hi; ++m) {
    unsigned long long k = (keys[(long long)m * P + p] & 0xFFFFFFFF00000000ULL) | (unsigned long long)(map_base + m);
    b = k < b ? k : b;
  }
  if (hi > lo) atomicMin(best + p, b);
}

int launch_push_reduce(const unsigned long long* keys, int n_maps, int P, uint32_t map_base,
                       unsigned long long* best, cudaStream_t st) {
  if (n_maps == 0 || P == 0) return SPS_OK;
  int split = (n_maps + 63) / 64;
  if (split > 128) split = 128;
  dim3 grid((P + 127) / 128, split);
  push_reduce_kernel<<<grid, 128, 0, st>>>(keys, n_maps, P, map_base, best);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

// k nearest maps per prototype (find_nearest.py:89-175: one heap of size k per prototype, fed with every image's
// closest patch).  One thread per prototype walks the maps of this shard in order and keeps a sorted list of the k
// smallest (distance, global map index) pairs with their pixel; the list is MERGED into what top_key / top_pix hold
// on entry (pre-filled with ~0 / 0, or a previous shard's result), so shards and ranks chain.
constexpr int TOPK_MAX = 16;

__global__ void push_topk_kernel(const unsigned long long* __restrict__ keys, int n_maps, int P, uint32_t map_base,
                                 int k, unsigned long long* __restrict__ top_key, int32_t* __restrict__ top_pix) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  unsigned long long kk[TOPK_MAX];
  int32_t pp[TOPK_MAX];
#pragma unroll
  for (int i = 0; i < TOPK_MAX; ++i) {
    kk[i] = i < k ? top_key[(long long)p * k + i] : 0ULL;   // slots beyond k never accept a candidate
    pp[i] = i < k ? top_pix[(long long)p * k + i] : 0;
  }
  for (int m = 0; m < n_maps; ++m) {
    const unsigned long long key = keys[(long long)m * P + p];
    unsigned long long cand = (key & 0xFFFFFFFF00000000ULL) | (unsigned long long)(map_base + m);
    int32_t pix = (int32_t)(key & 0xFFFFFFFFULL);
#pragma unroll
    for (int i = 0; i < TOPK_MAX; ++i) {   // one pass: the candidate sinks in, the displaced entry rides along
      if (i < k && cand < kk[i]) {
        const unsigned long long tk = kk[i];
        const int32_t tp = pp[i];
        kk[i] = cand;
        pp[i] = pix;
        cand = tk;
        pix = tp;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < TOPK_MAX; ++i)
    if (i < k) {
      top_key[(long long)p * k + i] = kk[i];
      top_pix[(long long)p * k + i] = pp[i];
    }
}

int launch_push_topk(const unsigned long long* keys, int n_maps, int P, uint32_t map_base, int k,
                     unsigned long long* top_key, int32_t* top_pix, cudaStream_t st) {
  SPS_REQUIRE(k >= 1 && k <= TOPK_MAX, SPS_ERR_UNSUPPORTED, "push_topk: k=%d outside [1,%d]", k, TOPK_MAX);
  if (n_maps == 0 || P == 0) return SPS_OK;
  push_topk_kernel<<<(P + 63) / 64, 64, 0, st>>>(keys, n_maps, P, map_base, k, top_key, top_pix);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

// ZT: 0 = fp32 NCHW maps, 1 = fp16 channels-last, 2 = bf16 channels-last (out is always fp32)
template <int ZT>
__global__ void push_gather_kernel(HeadShape s, const void* __restrict__ zv, const unsigned long long* __restrict__ keys,
                                   const unsigned long long* __restrict__ best, int n_maps, uint32_t map_base,
                                   const int32_t* __restrict__ proto_scale, float* __restrict__ out) {
  const int p = blockIdx.x;
  const unsigned long long w = best[p];
  const unsigned long long gmap = w & 0xFFFFFFFFULL;
  if (gmap < map_base || gmap >= (unsigned long long)map_base + n_maps) return;
  const long long m = (long long)(gmap - map_base);
  const long long pix = (long long)(keys[m * s.P + p] & 0xFFFFFFFFULL);
  const int sc = proto_scale[p];
  for (int c = threadIdx.x; c < s.D; c += blockDim.x) {
    float v;
    if (ZT == 0) {
      v = reinterpret_cast<const float*>(zv)[(m * s.C + sc * s.D + c) * s.HW + pix];
    } else {
      const unsigned short h = reinterpret_cast<const unsigned short*>(zv)[(m * s.HW + pix) * s.C + sc * s.D + c];
      v = ZT == 1 ? __half2float(__ushort_as_half(h)) : __uint_as_float((unsigned)h << 16);
    }
    out[(long long)p * s.D + c] = v;
  }
}

int launch_push_gather(const HeadShape& s, const void* z, const unsigned long long* keys,
                       const unsigned long long* best, int n_maps, uint32_t map_base, const int32_t* proto_scale,
                       float* out, cudaStream_t st) {
  if (s.P == 0 || n_maps == 0) return SPS_OK;
  if (s.z_dtype == SPS_Z_F16) push_gather_kernel<1><<<s.P, 64, 0, st>>>(s, z, keys, best, n_maps, map_base, proto_scale, out);
  else if (s.z_dtype == SPS_Z_BF16) push_gather_kernel<2><<<s.P, 64, 0, st>>>(s, z, keys, best, n_maps, map_base, proto_scale, out);
  else push_gather_kernel<0><<<s.P, 64, 0, st>>>(s, z, keys, best, n_maps, map_base, proto_scale, out);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

}  // namespace sps
