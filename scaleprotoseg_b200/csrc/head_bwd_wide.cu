// Tensor-core prototype head (sm_100a): ONE-PASS fused backward for wide last layers (more than 32 classes; COCO-Stuff: 171).
//
// head_bwd_tc.cu keeps the dW^T accumulator of its chunks in 32 TMEM columns; with K classes it takes K columns per chunk,
// so a CTA can own a single chunk of 64 prototypes and dz of a (tile, scale) would be spread over P_s / 64 CTAs.  Instead
// of class slabs (each recomputing every distance, dz by atomics) the work is cut along its two natural axes:
//
//   wide_dz_kernel   TILE-stationary.  One persistent CTA per SM walks position tiles; per tile the scaled dlogits tile
//                    (128 x Kp, 16 bit) is loaded once, per (tile, scale) the feature tile, and the prototype chunks of the
//                    scale stream through a ring:  S = |z|^2 - 2 z.p + |p|^2,  dA = dlogits . W,
//                    gamma = [S>0] dA a'(d)  ->  dz_s += gamma . (-2 p)  accumulated in TMEM over the chunks of the scale.
//                    dz is complete inside the CTA: plain stores, no atomics.
//   wide_dpw_kernel  CHUNK-stationary.  A CTA owns one chunk and a contiguous range of tiles; dp^T (64 x 64), dW^T (64 x Kp)
//                    and the column sums of gamma stay in TMEM for the whole range:  dp^T += gamma^T . z,
//                    dW^T += a^T . dlogits,  csum += gamma^T . 1.  Partials per (tile range, chunk) are summed in a fixed
//                    order by bwd_reduce_kernel (deterministic).
//
// Both recompute S and dA (the price of having no N x P tensor in HBM: 17 of the 21 / 41 MMAs per unit).  Everything is
// TMA-fed: the feature map is the 16-bit channels-last array [N][C] (the user's fp16 / bf16 tensor, or the re-lay of an fp32
// map), dlogits is re-laid once per call as a scaled 16-bit [N][192] array (dl_pack_kernel), so both land in shared memory
// in the SWIZZLE_128B layout the MMAs read; what is left for the SM is the row norm.  Distances are recomputed with
// single-term 16-bit products (the 3-term operands of F16X3 do not fit next to a 48 KB dlogits tile).
#include <type_traits>

#include "tma.cuh"

namespace sps {
namespace tw {
using namespace sps::tc;

constexpr int KPW = 192;                            // classes, padded (three 64-wide SW128 sub-tiles)
constexpr int DL_TILE = TILE_M * KPW * 2;           // 48 KB
constexpr int W3_TILE = CN * KPW * 2;               // 24 KB
constexpr int WOFF_BH = 0, WOFF_BN = CN * 128, WOFF_W3 = CN * 128 + 2048;
constexpr int WIMG = WOFF_W3 + W3_TILE;             // 34 KB per chunk: -2p tile | norm K step | W3
static_assert(WIMG % 1024 == 0 && WOFF_W3 % 1024 == 0, "chunk image alignment");

// power-of-two scale that brings max|dlogits| to [2^-5, 2^-4) (as head_bwd_tc.cu)
__device__ __forceinline__ float grad_scale(unsigned int amax_bits) {
  const float a = __uint_as_float(amax_bits);
  if (!(a > 0.f) || !isfinite(a)) return 1.0f;
  int e;
  frexpf(a, &e);
  int g = -4 - e;
  g = g > 100 ? 100 : (g < -100 ? -100 : g);
  return ldexpf(1.0f, g);
}

// ---------------------------------------------------------------------------------------------
// Pre-passes
// ---------------------------------------------------------------------------------------------
// 16 consecutive channels [kq, kq+16) of operand row j, value scale * src[k], rounded to one 16-bit term, into the SW128
// K-major tile; returns the sum of (0.5 * p_eff)^2 (pack_row16 of tc_common.cuh without the lo tile)
template <typename X>
__device__ __forceinline__ float pack_row16_hi(const float* __restrict__ src, bool live, int D, float scale, unsigned char* t_hi, int j,
                                               int kq) {
  float part = 0.f;
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    uint32_t hw[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k0 = kq + 8 * g + 2 * q;
      const float v0 = (live && k0 < D) ? scale * __ldg(src + k0) : 0.0f, v1 = (live && k0 + 1 < D) ? scale * __ldg(src + k0 + 1) : 0.0f;
      hw[q] = Pack2<X>::pack(v0, v1);
      const float2 e = Pack2<X>::unpack(hw[q]);
      part = fmaf(0.5f * e.x, 0.5f * e.x, part);
      part = fmaf(0.5f * e.y, 0.5f * e.y, part);
    }
    *reinterpret_cast<uint4*>(t_hi + sw128_off(j, kq + 8 * g)) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
  }
  return part;
}

// blocks [0, n_chunks): chunk images; the rest: max |dlogits| (one atomicMax per warp)
template <typename X>
__global__ void __launch_bounds__(256) wide_prep_kernel(const __grid_constant__ ChunkTable ct, int D, int P, int K,
                                                          const float* __restrict__ protos, const float* __restrict__ last_w,
                                                          unsigned char* __restrict__ img_base, const float* __restrict__ dlogits,
                                                          long long dl_count, unsigned int* __restrict__ amax_out) {
  if ((int)blockIdx.x >= ct.n_chunks) {
    const long long nb = (long long)gridDim.x - ct.n_chunks, b = (long long)blockIdx.x - ct.n_chunks;
    float m = 0.f;
    if ((reinterpret_cast<uintptr_t>(dlogits) & 15) == 0) {   // 16-byte loads, enough blocks to stream at HBM rate
      const long long n4 = dl_count >> 2;
      const float4* x4 = reinterpret_cast<const float4*>(dlogits);
      for (long long i = b * blockDim.x + threadIdx.x; i < n4; i += nb * blockDim.x) {
        const float4 v = __ldg(x4 + i);
        m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
      }
      for (long long i = (n4 << 2) + b * blockDim.x + threadIdx.x; i < dl_count; i += nb * blockDim.x) m = fmaxf(m, fabsf(__ldg(dlogits + i)));
    } else {
      for (long long i = b * blockDim.x + threadIdx.x; i < dl_count; i += nb * blockDim.x) m = fmaxf(m, fabsf(__ldg(dlogits + i)));
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(amax_out, __float_as_uint(m));
    return;
  }
  const int c = blockIdx.x, t = threadIdx.x;
  const int p0 = ct.p0[c], n = ct.n[c];
  unsigned char* img = img_base + (size_t)c * WIMG;
  __shared__ float pn[CN];
  {
    const int j = t >> 2, kq = (t & 3) * 16;
    float part = pack_row16_hi<X>(protos + (size_t)(p0 + (j < n ? j : 0)) * D, j < n, D, -2.0f, img + WOFF_BH, j, kq);
    part += __shfl_xor_sync(0xffffffffu, part, 1);
    part += __shfl_xor_sync(0xffffffffu, part, 2);
    if ((t & 3) == 0) pn[j] = part;
  }
  __syncthreads();
  if (t < CN) store_proto_norm<X>(img + WOFF_BN, t, pn[t], t < n);
  for (int idx = t; idx < CN * KPW; idx += 256) {
    const int k = idx / CN, j = idx % CN;
    const float v = (k < K && j < n) ? __ldg(last_w + (size_t)k * P + p0 + j) : 0.0f;
    *reinterpret_cast<uint16_t*>(img + WOFF_W3 + (k >> 6) * (CN * 128) + sw128_off(j, k & 63)) = bits16(Fmt<X>::from(v));
  }
}

// dlogits [N][K] fp32 -> G * dlogits as 16-bit [N][192] (zero padded): the TMA source of the dlogits tiles
template <typename X>
__global__ void __launch_bounds__(256) dl_pack_kernel(const float* __restrict__ dl, long long N, int K, const unsigned int* __restrict__ amax_bits,
                                                       uint4* __restrict__ out) {
  const float G = grad_scale(*amax_bits);
  const long long total = N * (KPW / 8);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long n = i / (KPW / 8);
    const int k0 = (int)(i % (KPW / 8)) * 8;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = k0 + e < K ? G * __ldg(dl + n * K + k0 + e) : 0.0f;
    out[i] = make_uint4(Pack2<X>::pack(v[0], v[1]), Pack2<X>::pack(v[2], v[3]), Pack2<X>::pack(v[4], v[5]), Pack2<X>::pack(v[6], v[7]));
  }
}

// ---------------------------------------------------------------------------------------------
// Tile-stationary dz kernel
// ---------------------------------------------------------------------------------------------
struct DzParams {
  CUtensorMap tm_z;     // [N][C] 16-bit feature map
  CUtensorMap tm_dl;    // [N][192] 16-bit scaled dlogits
  HeadShape s;
  ChunkTable ct;
  const unsigned char* img;
  const unsigned int* amax_bits;
  void* dz;
  int n_tiles;
};

constexpr int DZ_THREADS = 512;    // 8 epilogue warps (2 sets), feature / dlogits producer + norms, chunk producer, 2 issuers,
                                   // 4 warps that move the dlogits tile into TMEM (warps 12..15: lane quarters 0..3)
// The chunk operands travel in TWO rings.  The -2p tile (8 KB) is read by the distance GEMM and, an epilogue later, by the dz
// GEMM: a long life, so a deep ring of small slots.  The norm K step + W3 tile (26 KB) are read by issuer 1 only and free
// as soon as the dA GEMM has completed.  (One ring of whole 34 KB images held every slot for the full chain: 2 960 clk per unit.)
constexpr int DZ_NSH = 6, DZ_NSW = 2;
constexpr int DZ_WSLOT = 2048 + W3_TILE;           // norm K step | W3
constexpr int DZ_A = TILE_M * 128 + TILE_M * 32;   // feature tile + norm K step
constexpr int DZ_SM_DL = 0, DZ_SM_A = DL_TILE, DZ_SM_G = DZ_SM_A + 2 * DZ_A, DZ_SM_H = DZ_SM_G + 2 * TILE_M * 128,
              DZ_SM_W = DZ_SM_H + DZ_NSH * (CN * 128), DZ_SM_SG = DZ_SM_W + DZ_NSW * DZ_WSLOT, DZ_SM_BAR = DZ_SM_SG + 2 * 2 * 128 * 4,
              DZ_SM_TOTAL = DZ_SM_BAR + 512 + 1024;
static_assert(DZ_SM_TOTAL <= 227 * 1024 && DZ_WSLOT % 1024 == 0, "dz kernel shared memory");
enum { Z_TFULL = 0, Z_AFULL = 2, Z_AEMPTY = 4, Z_DLFULL = 6, Z_DLEMPTY = 7, Z_HFULL = 8, Z_HEMPTY = 16, Z_WFULL = 24, Z_WEMPTY = 26,
       Z_SFULL = 28, Z_SEMPTY = 30, Z_GFULL = 32, Z_GEMPTY = 34, Z_DZFULL = 36, Z_DZEMPTY = 38, Z_DLTFULL = 40, Z_DLTEMPTY = 41, Z_COUNT = 42 };
// The dlogits tile (128 positions x 192 classes) is the A operand of the dA GEMM of EVERY unit of its tile (12 of the 21 MMAs
// of a unit): it is copied once per tile from its TMA landing buffer into 96 TMEM columns (lane = position, column j =
// classes 2j, 2j+1), and those MMAs take A from TMEM: 32.8 clk instead of 48.9 for an M128 x N64 x K16 step
// (profiles/r02_probe_mma_ts.txt) - the kernel is bound by the time its MMAs occupy the tensor pipe.
constexpr int ZT_S = 0, ZT_DA = 128, ZT_DZ = 256, ZT_DL = 384;

// DZ16: dz is written in the 16-bit channels-last layout of the feature map; otherwise fp32 NCHW
template <typename X, bool DZ16>
__global__ void __launch_bounds__(DZ_THREADS, 1) wide_dz_kernel(const __grid_constant__ DzParams prm) {
  extern __shared__ unsigned char smem_raw[];
  const HeadShape& s = prm.s;
  const ChunkTable& ct = prm.ct;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* gbase = smem_raw + (base - smem_u32(smem_raw));
  auto bar = [&](int i) { return base + DZ_SM_BAR + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + DZ_SM_BAR + 8 * Z_COUNT);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int NC = ct.n_chunks;
  constexpr uint32_t FM = Fmt<X>::id;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(Z_TFULL + i), 1);
      mbar_init(bar(Z_AFULL + i), 1);
      mbar_init(bar(Z_AEMPTY + i), 9);     // commit of the pair's last distance GEMM + the 8 epilogue warps (dz read-out reads z)
      mbar_init(bar(Z_SFULL + i), 1);
      mbar_init(bar(Z_SEMPTY + i), 4);
      mbar_init(bar(Z_GFULL + i), 4);
      mbar_init(bar(Z_GEMPTY + i), 1);
      mbar_init(bar(Z_DZFULL + i), 1);
      mbar_init(bar(Z_DZEMPTY + i), 8);
    }
    mbar_init(bar(Z_DLFULL), 1);
    mbar_init(bar(Z_DLEMPTY), 4);          // the landing buffer is free once the four copy warps have read it
    mbar_init(bar(Z_DLTFULL), 4);          // TMEM copy of the dlogits tile written
    mbar_init(bar(Z_DLTEMPTY), 1);         // commit after the last dA GEMM of the tile
    for (int i = 0; i < DZ_NSH; ++i) {
      mbar_init(bar(Z_HFULL + i), 1);
      mbar_init(bar(Z_HEMPTY + i), 2);     // both issuers read the -2p tile (distance GEMM, dz GEMM)
    }
    for (int i = 0; i < DZ_NSW; ++i) {
      mbar_init(bar(Z_WFULL + i), 1);
      mbar_init(bar(Z_WEMPTY + i), 1);
    }
    mbar_fence_init();
  }
  if (warp == 10) tmem_alloc(smem_u32((const void*)tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const int my_tiles = prm.n_tiles > (int)blockIdx.x ? (prm.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int U = my_tiles * NC;
  auto first_of_scale = [&](int c) { return c == 0 || ct.scale[c] != ct.scale[c - 1]; };
  auto last_of_scale = [&](int c) { return c == NC - 1 || ct.scale[c + 1] != ct.scale[c]; };

  if (warp == 8) {
    // ================= feature tiles (per tile and scale), dlogits tile (per tile), row norms =================
    int pair = 0;
    for (int ti = 0; ti < my_tiles; ++ti) {
      const int row0 = (int)(((long long)blockIdx.x + (long long)ti * gridDim.x) * TILE_M);
      mbar_wait(bar(Z_DLEMPTY), (ti & 1) ^ 1);
      if (lane == 0) {
        mbar_arrive_expect_tx(bar(Z_DLFULL), DL_TILE);
#pragma unroll
        for (int t = 0; t < 3; ++t) tma_load_2d(base + DZ_SM_DL + t * (TILE_M * 128), &prm.tm_dl, 64 * t, row0, bar(Z_DLFULL));
      }
      __syncwarp();
      for (int c = 0; c < NC; ++c) {
        if (!first_of_scale(c)) continue;
        const uint32_t ab = pair & 1, aph = (pair >> 1) & 1;
        mbar_wait(bar(Z_AEMPTY + ab), aph ^ 1);
        if (lane == 0) {
          mbar_arrive_expect_tx(bar(Z_TFULL + ab), TILE_M * 128);
          tma_load_2d(base + DZ_SM_A + ab * DZ_A, &prm.tm_z, ct.scale[c] * s.D, row0, bar(Z_TFULL + ab));
        }
        __syncwarp();
        mbar_wait(bar(Z_TFULL + ab), aph);
        unsigned char* a0 = gbase + DZ_SM_A + ab * DZ_A;
#pragma unroll
        for (int rr = 0; rr < 4; ++rr) {
          const int r = lane + 32 * rr;
          ZStage<X, 1>::store_norm(r, a0 + TILE_M * 128, tile_row_norm<X, false>(a0, a0, r));
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(Z_AFULL + ab));
        ++pair;
      }
    }
  } else if (warp == 9) {
    // ================= prototype-chunk rings =================
    for (int i = 0; i < U; ++i) {
      const int c = i % NC, sh = i % DZ_NSH, sw = i % DZ_NSW;
      const unsigned char* src = prm.img + (size_t)c * WIMG;
      mbar_wait(bar(Z_HEMPTY + sh), ((i / DZ_NSH) & 1) ^ 1);
      if (lane == 0) {
        mbar_arrive_expect_tx(bar(Z_HFULL + sh), CN * 128);
        bulk_g2s(base + DZ_SM_H + sh * (CN * 128), src + WOFF_BH, CN * 128, bar(Z_HFULL + sh));
      }
      __syncwarp();
      mbar_wait(bar(Z_WEMPTY + sw), ((i / DZ_NSW) & 1) ^ 1);
      if (lane == 0) {
        mbar_arrive_expect_tx(bar(Z_WFULL + sw), DZ_WSLOT);
        bulk_g2s(base + DZ_SM_W + sw * DZ_WSLOT, src + WOFF_BN, DZ_WSLOT, bar(Z_WFULL + sw));
      }
      __syncwarp();
    }
  } else if (warp == 10) {
    // ================= issuer 1: distance GEMM + dA GEMM =================
    constexpr uint32_t ID = idesc_f16(FM, TILE_M, CN);
    int pair = -1, c = 0, ti = 0;
    for (int u = 0; u < U; ++u) {
      const uint32_t sh = u % DZ_NSH, sw = u % DZ_NSW, sb = u & 1;
      if (first_of_scale(c)) {
        ++pair;
        mbar_wait(bar(Z_AFULL + (pair & 1)), (pair >> 1) & 1);
      }
      if (c == 0) mbar_wait(bar(Z_DLTFULL), ti & 1);
      mbar_wait(bar(Z_HFULL + sh), (u / DZ_NSH) & 1);
      mbar_wait(bar(Z_WFULL + sw), (u / DZ_NSW) & 1);
      mbar_wait(bar(Z_SEMPTY + sb), ((u >> 1) & 1) ^ 1);
      tc_fence_after();
      {
        const uint32_t a0 = base + DZ_SM_A + (pair & 1) * DZ_A, bh = base + DZ_SM_H + sh * (CN * 128), bw = base + DZ_SM_W + sw * DZ_WSLOT;
        const uint32_t dS = tmem + ZT_S + sb * CN, dA = tmem + ZT_DA + sb * CN, tdl = tmem + ZT_DL;
        mma_f16(dS, desc_noswz(a0 + TILE_M * 128, NORM_LBO, NORM_SBO), desc_noswz(bw, NORM_LBO, NORM_SBO), ID, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_f16(dS, desc_sw128(a0 + 32 * k), desc_sw128(bh + 32 * k), ID, 1);
#pragma unroll
        for (int k = 0; k < KPW / 16; ++k)
          mma_f16_ts(dA, tdl + 8 * k, desc_sw128(bw + 2048 + (k >> 2) * (CN * 128) + 32 * (k & 3)), ID, k);
        mma_commit(bar(Z_SFULL + sb));
        mma_commit(bar(Z_HEMPTY + sh));
        mma_commit(bar(Z_WEMPTY + sw));
        if (last_of_scale(c)) mma_commit(bar(Z_AEMPTY + (pair & 1)));
        if (c == NC - 1) mma_commit(bar(Z_DLTEMPTY));
      }
      __syncwarp();
      if (++c == NC) { c = 0; ++ti; }
    }
  } else if (warp >= 12) {
    // ================= dlogits tile: landing buffer -> TMEM (one row per thread) =================
    const int q = warp - 12, m = q * 32 + lane;
    const uint32_t lane_base = ((uint32_t)(q * 32)) << 16;
    for (int ti = 0; ti < my_tiles; ++ti) {
      mbar_wait(bar(Z_DLFULL), ti & 1);
      mbar_wait(bar(Z_DLTEMPTY), (ti & 1) ^ 1);                   // the dA GEMMs of the previous tile have read the TMEM copy
      tc_fence_after();
#pragma unroll 1
      for (int t = 0; t < 3; ++t) {
        uint32_t v[32];
        const uint4* row = reinterpret_cast<const uint4*>(gbase + DZ_SM_DL + t * (TILE_M * 128) + m * 128);
#pragma unroll
        for (int cc = 0; cc < 8; ++cc) {
          const uint4 w = row[cc ^ (m & 7)];
          v[4 * cc] = w.x; v[4 * cc + 1] = w.y; v[4 * cc + 2] = w.z; v[4 * cc + 3] = w.w;
        }
        tmem_st32(tmem + lane_base + ZT_DL + 32 * t, v);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(Z_DLEMPTY));                // the next tile's dlogits may land (a whole tile ahead of their use)
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(Z_DLTFULL));
    }
  } else if (warp == 11) {
    // ================= issuer 2: dz_s += gamma . (-2 p) =================
    constexpr uint32_t ID4 = idesc_f16_major(FM, TILE_M, CN, 0, 1);
    int pair = -1, c = 0;
    for (int u = 0; u < U; ++u) {
      const uint32_t sh = u % DZ_NSH, gb = u & 1;
      const bool fos = first_of_scale(c);
      if (fos) {
        ++pair;
        mbar_wait(bar(Z_DZEMPTY + (pair & 1)), ((pair >> 1) & 1) ^ 1);
      }
      mbar_wait(bar(Z_GFULL + gb), (u >> 1) & 1);
      tc_fence_after();
      {
        const uint32_t gam = base + DZ_SM_G + gb * (TILE_M * 128), bh = base + DZ_SM_H + sh * (CN * 128);
        const uint32_t d = tmem + ZT_DZ + (pair & 1) * CN;
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(gam + 32 * k), desc_mn_sw128(bh + 2048 * k, 0), ID4, (!fos || k > 0) ? 1 : 0);
        mma_commit(bar(Z_GEMPTY + gb));
        mma_commit(bar(Z_HEMPTY + sh));
        if (last_of_scale(c)) mma_commit(bar(Z_DZFULL + (pair & 1)));
      }
      __syncwarp();
      if (++c == NC) c = 0;
    }
  } else if (warp < 8) {
    // ================= epilogue: gamma tiles, row sums, dz read-out =================
    const int q = warp & 3, set = warp >> 2;
    const int m = q * 32 + lane;
    const uint32_t lane_base = ((uint32_t)(q * 32)) << 16;
    const float eps = s.eps, ome = 1.0f - s.eps;
    const bool is_log = s.act_kind == SPS_ACT_LOG;
    const float G = grad_scale(*prm.amax_bits), invG = 1.0f / G;
    float* sgbuf = reinterpret_cast<float*>(gbase + DZ_SM_SG);
    // pairs (tile, scale) are numbered in processing order: pair = tile * live scales + rank of the scale among the live ones
    int n_scales_live = 0;
    int live_scale[SPS_MAX_SCALES];
    for (int c = 0; c < NC; ++c)
      if (first_of_scale(c)) live_scale[n_scales_live++] = ct.scale[c];
    float sg = 0.f;
    int next_readout = 0;
    auto readout = [&](int pr) {
      const int tj = pr / n_scales_live;
      int sc = live_scale[0];
#pragma unroll
      for (int k = 1; k < SPS_MAX_SCALES; ++k)
        if (k == pr % n_scales_live) sc = live_scale[k];      // register select: no local-memory indexing
      const long long n = ((long long)blockIdx.x + (long long)tj * gridDim.x) * TILE_M + m;
      const bool valid = n < s.N;
      const uint32_t ab = pr & 1;
      float* sgb = sgbuf + (pr & 1) * 256;
      sgb[set * 128 + m] = sg;
      sg = 0.f;
      named_bar_sync(1, 256);
      const float tot2 = 2.0f * (sgb[m] + sgb[128 + m]);
      uint32_t rd[32];
      mbar_wait(bar(Z_DZFULL + ab), (pr >> 1) & 1);
      tc_fence_after();
      tmem_ld32(tmem + lane_base + ZT_DZ + ab * CN + set * 32, rd);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(Z_DZEMPTY + ab));
      const unsigned char* zrow = gbase + DZ_SM_A + ab * DZ_A + (uint32_t)m * 128u;
      float dv[32];
#pragma unroll
      for (int gq = 0; gq < 4; ++gq) {
        const uint32_t zo = (((uint32_t)(set * 4 + gq)) ^ ((uint32_t)m & 7u)) << 4;
        const uint4 zz = *reinterpret_cast<const uint4*>(zrow + zo);
        const uint32_t zw[4] = {zz.x, zz.y, zz.z, zz.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 zf = Pack2<X>::unpack(zw[e]);
          dv[gq * 8 + 2 * e] = fmaf(zf.x, tot2, __uint_as_float(rd[gq * 8 + 2 * e])) * invG;
          dv[gq * 8 + 2 * e + 1] = fmaf(zf.y, tot2, __uint_as_float(rd[gq * 8 + 2 * e + 1])) * invG;
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(Z_AEMPTY + ab));
      if (prm.dz != nullptr && valid) {
        if (DZ16) {
          uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<unsigned char*>(prm.dz) + ((size_t)n * s.C + (size_t)sc * s.D + set * 32) * 2);
#pragma unroll
          for (int gq = 0; gq < 4; ++gq)
            dst[gq] = make_uint4(Pack2<X>::pack(dv[gq * 8], dv[gq * 8 + 1]), Pack2<X>::pack(dv[gq * 8 + 2], dv[gq * 8 + 3]),
                                 Pack2<X>::pack(dv[gq * 8 + 4], dv[gq * 8 + 5]), Pack2<X>::pack(dv[gq * 8 + 6], dv[gq * 8 + 7]));
        } else {
          const long long b = n / s.HW, pix = n % s.HW;
          float* dzp = reinterpret_cast<float*>(prm.dz) + (b * (long long)s.C + (long long)sc * s.D + set * 32) * s.HW + pix;
          const size_t pitch = (size_t)s.HW;
#pragma unroll
          for (int k = 0; k < 32; ++k) {
            *dzp = dv[k];
            dzp += pitch;
          }
        }
      }
    };
    // (chunk, pair) of this set's current unit, advanced two units at a time (an O(chunks) search per unit cost more
    // than the unit's arithmetic at 108 chunks)
    int cu = 0, pr = 0;
    auto step_unit = [&]() {
      if (++cu == NC) cu = 0;
      if (first_of_scale(cu)) ++pr;
    };
    if (set == 1) step_unit();
    for (int u = set; u < U; u += 2) {
      while (next_readout < pr) readout(next_readout++);     // both sets run every read-out, in the same order
      const uint32_t sb = set, gb = set, ph = (u >> 1) & 1;
      uint32_t gw[32];
      mbar_wait(bar(Z_SFULL + sb), ph);
      tc_fence_after();
#pragma unroll 1
      for (int hf = 0; hf < 2; ++hf) {
        uint32_t rs[32], ra[32];
        tmem_ld32(tmem + lane_base + ZT_S + sb * CN + hf * 32, rs);
        tmem_ld32(tmem + lane_base + ZT_DA + sb * CN + hf * 32, ra);
        tmem_ld_wait();
        if (hf == 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(Z_SEMPTY + sb));
        }
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          float gm[2];
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const float sv = __uint_as_float(rs[2 * e + h]), dav = __uint_as_float(ra[2 * e + h]);
            const float d = fmaxf(sv, 0.0f);
            const float ap = is_log ? -ome * rcp_approx((d + 1.0f) * (d + eps)) : -1.0f;
            gm[h] = sv > 0.0f ? dav * ap : 0.0f;
            sg += gm[h];
          }
          gw[hf * 16 + e] = Pack2<X>::pack(gm[0], gm[1]);
        }
      }
      mbar_wait(bar(Z_GEMPTY + gb), ph ^ 1);
      unsigned char* grow = gbase + DZ_SM_G + gb * (TILE_M * 128) + (uint32_t)m * 128u;
#pragma unroll
      for (int g8 = 0; g8 < 8; ++g8) {
        const uint32_t off = (((uint32_t)g8) ^ ((uint32_t)m & 7u)) << 4;
        *reinterpret_cast<uint4*>(grow + off) = make_uint4(gw[4 * g8], gw[4 * g8 + 1], gw[4 * g8 + 2], gw[4 * g8 + 3]);
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(Z_GFULL + gb));
      step_unit();
      step_unit();
    }
    const int n_pairs = my_tiles * n_scales_live;
    while (next_readout < n_pairs) readout(next_readout++);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 10) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// ---------------------------------------------------------------------------------------------
// Chunk-stationary dp / dW kernel
// ---------------------------------------------------------------------------------------------
struct DpwParams {
  CUtensorMap tm_z;
  CUtensorMap tm_dl;
  HeadShape s;
  ChunkTable ct;
  const unsigned char* img;
  const unsigned int* amax_bits;
  float* part_dp;   // [nsplit][P*D]
  float* part_dw;   // [nsplit][K*P]
  float* part_cs;   // [nsplit][P]
  int n_tiles;
  int nsplit;
};

constexpr int PW_THREADS = 480;
constexpr int PW_A = TILE_M * 128 + TILE_M * 32 + DL_TILE;     // feature tile | norm | dlogits tile = 68 KB
constexpr int PW_AOFF_N = TILE_M * 128, PW_AOFF_DL = TILE_M * 128 + TILE_M * 32;
constexpr int PW_SM_RES = 0, PW_SM_A = WIMG, PW_SM_G = PW_SM_A + 2 * PW_A, PW_SM_BAR = PW_SM_G + 2 * TILE_M * 128,
              PW_SM_TOTAL = PW_SM_BAR + 256 + 1024;
static_assert(PW_SM_TOTAL <= 227 * 1024 && PW_AOFF_DL % 1024 == 0 && PW_A % 1024 == 0, "dp/dW kernel shared memory");
enum { W_TFULL = 0, W_AFULL = 2, W_AEMPTY = 4, W_SFULL = 6, W_SEMPTY = 8, W_GFULL = 10, W_GEMPTY = 11, W_RES = 12, W_ACC = 13, W_COUNT = 14 };
constexpr int WT_S = 0, WT_DA = 128, WT_ACC = 320, WT_CS = 384;

template <typename X>
__global__ void __launch_bounds__(PW_THREADS, 1) wide_dpw_kernel(const __grid_constant__ DpwParams prm) {
  extern __shared__ unsigned char smem_raw[];
  const HeadShape& s = prm.s;
  const ChunkTable& ct = prm.ct;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* gbase = smem_raw + (base - smem_u32(smem_raw));
  auto bar = [&](int i) { return base + PW_SM_BAR + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + PW_SM_BAR + 8 * W_COUNT);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t FM = Fmt<X>::id;

  // tile range index is the slow grid index: the CTAs running together sweep the same ranges (the tiles stay hot in L2)
  const int c = (int)blockIdx.x % ct.n_chunks, sp = (int)blockIdx.x / ct.n_chunks;
  const int sc = ct.scale[c], p0 = ct.p0[c], nc = ct.n[c];
  const int t_lo = (int)((long long)prm.n_tiles * sp / prm.nsplit), t_hi = (int)((long long)prm.n_tiles * (sp + 1) / prm.nsplit);
  const int U = t_hi - t_lo;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(W_TFULL + i), 1);
      mbar_init(bar(W_AFULL + i), 4);
      mbar_init(bar(W_AEMPTY + i), 1);
      mbar_init(bar(W_SFULL + i), 1);
      mbar_init(bar(W_SEMPTY + i), 8);
    }
    mbar_init(bar(W_GFULL), 8);
    mbar_init(bar(W_GEMPTY), 1);
    mbar_init(bar(W_RES), 1);
    mbar_init(bar(W_ACC), 1);
    mbar_fence_init();
  }
  if (warp == 12) tmem_alloc(smem_u32((const void*)tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 14) {
    if (lane == 0) {
      mbar_arrive_expect_tx(bar(W_RES), WIMG);
      bulk_g2s(base + PW_SM_RES, prm.img + (size_t)c * WIMG, WIMG, bar(W_RES));
    }
  } else if (warp == 12) {
    // ================= issuer 1: distance GEMM + dA GEMM =================
    constexpr uint32_t ID = idesc_f16(FM, TILE_M, CN);
    mbar_wait(bar(W_RES), 0);
    for (int u = 0; u < U; ++u) {
      const uint32_t ab = u & 1, sb = u & 1, ph = (u >> 1) & 1;
      mbar_wait(bar(W_AFULL + ab), ph);
      mbar_wait(bar(W_SEMPTY + sb), ph ^ 1);
      tc_fence_after();
      {
        const uint32_t a0 = base + PW_SM_A + ab * PW_A, img = base + PW_SM_RES, dl = a0 + PW_AOFF_DL;
        const uint32_t dS = tmem + WT_S + sb * CN, dA = tmem + WT_DA + sb * CN;
        mma_f16(dS, desc_noswz(a0 + PW_AOFF_N, NORM_LBO, NORM_SBO), desc_noswz(img + WOFF_BN, NORM_LBO, NORM_SBO), ID, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_f16(dS, desc_sw128(a0 + 32 * k), desc_sw128(img + WOFF_BH + 32 * k), ID, 1);
#pragma unroll
        for (int k = 0; k < KPW / 16; ++k)
          mma_f16(dA, desc_sw128(dl + (k >> 2) * (TILE_M * 128) + 32 * (k & 3)), desc_sw128(img + WOFF_W3 + (k >> 2) * (CN * 128) + 32 * (k & 3)),
                  ID, k);
        mma_commit(bar(W_SFULL + sb));
      }
      __syncwarp();
    }
  } else if (warp == 13) {
    // ================= issuer 2: dp^T, dW^T, csum =================
    constexpr uint32_t ID5A = idesc_f16_major(FM, 64, 64, 1, 1);
    constexpr uint32_t ID5B = idesc_f16_major(FM, 64, KPW, 1, 1);
    constexpr uint32_t ID5C = idesc_f16_major(FM, 64, 8, 1, 1);
    mbar_wait(bar(W_RES), 0);
    for (int u = 0; u < U; ++u) {
      const uint32_t ab = u & 1;
      mbar_wait(bar(W_GFULL), u & 1);
      tc_fence_after();
      {
        const uint32_t a0 = base + PW_SM_A + ab * PW_A, dl = a0 + PW_AOFF_DL;
        const uint32_t gam = base + PW_SM_G, act = gam + TILE_M * 128;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t accum = (u > 0 || k > 0) ? 1 : 0;
          mma_f16(tmem + WT_ACC, desc_mn_sw128(gam + 2048 * k, 0), desc_mn_sw128(a0 + 2048 * k, 0), ID5A, accum);
          mma_f16(tmem + WT_ACC + (16u << 16), desc_mn_sw128(act + 2048 * k, 0), desc_mn_sw128(dl + 2048 * k, TILE_M * 128), ID5B, accum);
          mma_f16(tmem + WT_CS, desc_mn_sw128(gam + 2048 * k, 0),
                  (uint64_t)(((a0 + PW_AOFF_N + 512 * k) & 0x3FFFF) >> 4) | ((uint64_t)(256 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) |
                      ((uint64_t)1 << 46),
                  ID5C, accum);
        }
        mma_commit(bar(W_GEMPTY));
        mma_commit(bar(W_AEMPTY + ab));
        if (u == U - 1) mma_commit(bar(W_ACC));
      }
      __syncwarp();
    }
  } else if (warp >= 8 && warp < 12) {
    // ================= tile producers: TMA of the feature + dlogits tiles, row norms =================
    const int m = threadIdx.x - 8 * 32;
    for (int u = 0; u < U; ++u) {
      const uint32_t ab = u & 1, ph = (u >> 1) & 1;
      mbar_wait(bar(W_AEMPTY + ab), ph ^ 1);
      unsigned char* a0 = gbase + PW_SM_A + ab * PW_A;
      if (m == 0) {
        const int row0 = (t_lo + u) * TILE_M;
        const uint32_t dst = base + PW_SM_A + ab * PW_A;
        mbar_arrive_expect_tx(bar(W_TFULL + ab), TILE_M * 128 + DL_TILE);
        tma_load_2d(dst, &prm.tm_z, sc * s.D, row0, bar(W_TFULL + ab));
#pragma unroll
        for (int t = 0; t < 3; ++t) tma_load_2d(dst + PW_AOFF_DL + t * (TILE_M * 128), &prm.tm_dl, 64 * t, row0, bar(W_TFULL + ab));
      }
      mbar_wait(bar(W_TFULL + ab), ph);
      ZStage<X, 1>::store_norm(m, a0 + PW_AOFF_N, tile_row_norm<X, false>(a0, a0, m));
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(W_AFULL + ab));
    }
  } else if (warp < 8) {
    // ================= epilogue: gamma / a tiles, then the accumulator read-out =================
    const int q = warp & 3, half = warp >> 2;
    const int m = q * 32 + lane;
    const uint32_t lane_base = ((uint32_t)(q * 32)) << 16;
    const float eps = s.eps, ome = 1.0f - s.eps;
    const bool is_log = s.act_kind == SPS_ACT_LOG;
    const float G = grad_scale(*prm.amax_bits), invG = 1.0f / G;
    for (int u = 0; u < U; ++u) {
      const uint32_t sb = u & 1, ph = (u >> 1) & 1;
      uint32_t rs[32], ra[32];
      mbar_wait(bar(W_SFULL + sb), ph);
      tc_fence_after();
      tmem_ld32(tmem + lane_base + WT_S + sb * CN + half * 32, rs);
      tmem_ld32(tmem + lane_base + WT_DA + sb * CN + half * 32, ra);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(W_SEMPTY + sb));
      uint32_t gw[16], aw[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        float gm[2], a2[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const float sv = __uint_as_float(rs[2 * e + h]), dav = __uint_as_float(ra[2 * e + h]);
          const float d = fmaxf(sv, 0.0f);
          float ap, av;
          if (is_log) {
            const float s1 = d + 1.0f, ru = rcp_approx(s1 * (d + eps));
            ap = -ome * ru;
            av = lg2_approx(s1 * s1 * ru);
          } else {
            ap = -1.0f;
            av = -d;
          }
          gm[h] = sv > 0.0f ? dav * ap : 0.0f;
          a2[h] = av;
        }
        gw[e] = Pack2<X>::pack(gm[0], gm[1]);
        aw[e] = Pack2<X>::pack(a2[0], a2[1]);
      }
      mbar_wait(bar(W_GEMPTY), (u & 1) ^ 1);
      unsigned char* grow = gbase + PW_SM_G + (uint32_t)m * 128u;
#pragma unroll
      for (int gq = 0; gq < 4; ++gq) {
        const uint32_t off = (((uint32_t)(half * 4 + gq)) ^ ((uint32_t)m & 7u)) << 4;
        *reinterpret_cast<uint4*>(grow + off) = make_uint4(gw[4 * gq], gw[4 * gq + 1], gw[4 * gq + 2], gw[4 * gq + 3]);
        *reinterpret_cast<uint4*>(grow + TILE_M * 128 + off) = make_uint4(aw[4 * gq], aw[4 * gq + 1], aw[4 * gq + 2], aw[4 * gq + 3]);
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(W_GFULL));
    }
    if (U > 0) {
      // M = 64 accumulators: prototype row pl = q*16 + l sits in TMEM lane q*32 + l (dp^T: 64 channels at WT_ACC, csum at
      // WT_CS) and lane q*32 + 16 + l (dW^T: 192 classes at WT_ACC).  Warp (q, half) reads three 32-column blocks of its
      // half of [WT_ACC, WT_ACC + 192): the upper lanes take them as classes, the lower lanes as channels / column sum.
      mbar_wait(bar(W_ACC), 0);
      tc_fence_after();
      const float wsc = (is_log ? 0.69314718055994531f : 1.0f) * invG, m2 = -2.0f * invG;
      const int pl = q * 16 + (lane & 15);
      const bool live = pl < nc;
      float* o_dp = prm.part_dp + (size_t)sp * s.P * s.D + (size_t)(p0 + pl) * s.D;
      float* o_dw = prm.part_dw + (size_t)sp * s.K * s.P + p0 + pl;
#pragma unroll 1
      for (int blk = 0; blk < 3; ++blk) {
        const int col = half * 96 + blk * 32;       // column offset inside the 192-wide accumulator block
        uint32_t r[32];
        tmem_ld32(tmem + lane_base + WT_ACC + col, r);
        tmem_ld_wait();
        if (live) {
          if (lane >= 16) {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (col + e < s.K) o_dw[(size_t)(col + e) * s.P] = __uint_as_float(r[e]) * wsc;
          } else if (col < 64) {                      // channels [col, col + 32) (both blocks belong to the first warp set)
#pragma unroll
            for (int e4 = 0; e4 < 8; ++e4)
              *reinterpret_cast<float4*>(o_dp + col + 4 * e4) =
                  make_float4(__uint_as_float(r[4 * e4]) * m2, __uint_as_float(r[4 * e4 + 1]) * m2, __uint_as_float(r[4 * e4 + 2]) * m2,
                              __uint_as_float(r[4 * e4 + 3]) * m2);
          } else if (col == 64) {                     // WT_CS = WT_ACC + 64: column sum of gamma
            prm.part_cs[(size_t)sp * s.P + p0 + pl] = -__uint_as_float(r[0]) * m2;
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 12) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

}  // namespace tw

// ---------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------
int build_chunk_table(const HeadShape& s, bool with_last_layer, tc::ChunkTable* ct);
int launch_bwd_reduce(const float* part_dp, long long c1, float* dprotos, const float* part_dw, long long c2, float* dlast_w, int nsplit,
                      const float* part_cs, const float* protos, int D, int P, cudaStream_t st);

static int wide_nsplit(int n_chunks, int n_tiles) {
  // tile ranges per chunk: the grid (n_chunks * ranges CTAs, one per SM at a time) should fill whole waves, with at least two
  // waves (the tail of one range overlaps the start of another) but few of them: every CTA pays a prologue (resident
  // operands, first tile) and a 61 KB partial flush, and every range adds a partial buffer to the reduction
  const int sms = sm_count();
  int best = 1;
  double best_score = -1.0;
  for (int ns = 1; ns <= 8 && ns <= n_tiles; ++ns) {
    const int ctas = n_chunks * ns, waves = (ctas + sms - 1) / sms;
    const double eff = (double)ctas / ((double)waves * sms);
    const double score = eff - (waves < 2 ? 0.25 : 0.0) - 0.01 * waves;
    if (score > best_score + 1e-9) best_score = score, best = ns;
  }
  return best;
}

bool head_bwd_wide_supported(const HeadShape& s, int math, const float* last_w, const float* dlogits, const float* ddist,
                             const float* dact) {
  if (math == SPS_MATH_FP32 || last_w == nullptr || dlogits == nullptr || ddist != nullptr || dact != nullptr) return false;
  if (s.D != 64 || s.K > tw::KPW || s.N <= 0 || s.N >= (1ll << 31) - tc::TILE_M || s.P <= 0) return false;
  int nc = 0;
  for (int sc = 0; sc < s.S; ++sc) nc += (s.scale_lo[sc + 1] - s.scale_lo[sc] + tc::CN - 1) / tc::CN;
  return nc <= tc::MAX_CHUNKS;
}

size_t head_bwd_wide_workspace(const HeadShape& s) {
  tc::ChunkTable ct;
  if (build_chunk_table(s, true, &ct) != SPS_OK) return 0;
  const int n_tiles = (int)((s.N + tc::TILE_M - 1) / tc::TILE_M);
  const size_t nsplit = (size_t)wide_nsplit(ct.n_chunks, n_tiles > 0 ? n_tiles : 1);
  const size_t dl16 = ((size_t)s.N * tw::KPW * 2 + 1023) / 1024 * 1024;
  return 1024 + (size_t)ct.n_chunks * tw::WIMG + 1024 + dl16 + sizeof(float) * nsplit * ((size_t)s.P * s.D + (size_t)s.K * s.P + (size_t)s.P);
}

template <typename X>
static int launch_bwd_wide_t(const HeadShape& s, const void* z16, bool dz16, const float* protos, const float* last_w,
                             const float* dlogits, void* dz, float* dprotos, float* dlast_w, void* ws, size_t ws_bytes, cudaStream_t st) {
  tw::DzParams pz;
  int rc = build_chunk_table(s, true, &pz.ct);
  if (rc != SPS_OK) return rc;
  const int n_tiles = (int)((s.N + tc::TILE_M - 1) / tc::TILE_M);
  const int nsplit = wide_nsplit(pz.ct.n_chunks, n_tiles);
  const size_t need = head_bwd_wide_workspace(s);
  SPS_REQUIRE(ws != nullptr && ws_bytes >= need, SPS_ERR_WORKSPACE, "head_bwd workspace: need %zu bytes, got %zu", need, ws_bytes);
  unsigned char* img = reinterpret_cast<unsigned char*>(((uintptr_t)ws + 1023) & ~(uintptr_t)1023);
  unsigned int* amax = reinterpret_cast<unsigned int*>(img + (size_t)pz.ct.n_chunks * tw::WIMG);
  unsigned char* dl16 = reinterpret_cast<unsigned char*>(amax) + 1024;
  const size_t dl16_bytes = ((size_t)s.N * tw::KPW * 2 + 1023) / 1024 * 1024;
  float* part_dp = reinterpret_cast<float*>(dl16 + dl16_bytes);
  float* part_dw = part_dp + (size_t)nsplit * s.P * s.D;
  float* part_cs = part_dw + (size_t)nsplit * s.K * s.P;
  SPS_CUDA_TRY(cudaMemsetAsync(amax, 0, 1024, st));
  const bool bf = std::is_same<X, __nv_bfloat16>::value;
  // a scale without prototypes has no chunk: its block of dz is zero
  bool empty_scale = false;
  for (int sc = 0; sc < s.S; ++sc) empty_scale |= s.scale_lo[sc + 1] == s.scale_lo[sc];
  if (dz != nullptr && empty_scale) SPS_CUDA_TRY(cudaMemsetAsync(dz, 0, (dz16 ? 2 : 4) * (size_t)s.N * s.C, st));

  tw::wide_prep_kernel<X><<<pz.ct.n_chunks + 16 * sm_count(), 256, 0, st>>>(pz.ct, s.D, s.P, s.K, protos, last_w, img, dlogits,
                                                                        (long long)s.N * s.K, amax);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  tw::dl_pack_kernel<X><<<sm_count() * 8, 256, 0, st>>>(dlogits, s.N, s.K, amax, reinterpret_cast<uint4*>(dl16));
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());

  rc = make_tmap_2d_16bit(&pz.tm_z, z16, bf, s.N, s.C, tc::TILE_M, 64, 128);
  if (rc != SPS_OK) return rc;
  rc = make_tmap_2d_16bit(&pz.tm_dl, dl16, bf, s.N, tw::KPW, tc::TILE_M, 64, 128);
  if (rc != SPS_OK) return rc;
  if (dz != nullptr) {
    pz.s = s;
    pz.img = img;
    pz.amax_bits = amax;
    pz.dz = dz;
    pz.n_tiles = n_tiles;
    const int grid = n_tiles < sm_count() ? n_tiles : sm_count();
    if (dz16) {
      SPS_CUDA_TRY(cudaFuncSetAttribute(tw::wide_dz_kernel<X, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tw::DZ_SM_TOTAL));
      tw::wide_dz_kernel<X, true><<<grid, tw::DZ_THREADS, tw::DZ_SM_TOTAL, st>>>(pz);
    } else {
      SPS_CUDA_TRY(cudaFuncSetAttribute(tw::wide_dz_kernel<X, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, tw::DZ_SM_TOTAL));
      tw::wide_dz_kernel<X, false><<<grid, tw::DZ_THREADS, tw::DZ_SM_TOTAL, st>>>(pz);
    }
    count_launch();
    SPS_CUDA_TRY(cudaGetLastError());
  }
  if (dprotos != nullptr || dlast_w != nullptr) {
    tw::DpwParams pw;
    pw.tm_z = pz.tm_z;
    pw.tm_dl = pz.tm_dl;
    pw.s = s;
    pw.ct = pz.ct;
    pw.img = img;
    pw.amax_bits = amax;
    pw.part_dp = part_dp;
    pw.part_dw = part_dw;
    pw.part_cs = part_cs;
    pw.n_tiles = n_tiles;
    pw.nsplit = nsplit;
    SPS_CUDA_TRY(cudaFuncSetAttribute(tw::wide_dpw_kernel<X>, cudaFuncAttributeMaxDynamicSharedMemorySize, tw::PW_SM_TOTAL));
    tw::wide_dpw_kernel<X><<<pz.ct.n_chunks * nsplit, tw::PW_THREADS, tw::PW_SM_TOTAL, st>>>(pw);
    count_launch();
    SPS_CUDA_TRY(cudaGetLastError());
    const long long c1 = dprotos != nullptr ? (long long)s.P * s.D : 0, c2 = dlast_w != nullptr ? (long long)s.K * s.P : 0;
    rc = launch_bwd_reduce(part_dp, c1, dprotos, part_dw, c2, dlast_w, nsplit, part_cs, protos, s.D, s.P, st);
    if (rc != SPS_OK) return rc;
  }
  return SPS_OK;
}

// z16: 16-bit channels-last [N][C] map (the user's tensor, or the hi part of a re-lay).  dz16: dz in that layout too
// (otherwise fp32 NCHW).
int launch_head_bwd_wide(const HeadShape& s, int math, const void* z16, bool dz16, const float* protos, const float* last_w,
                         const float* dlogits, void* dz, float* dprotos, float* dlast_w, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (math == SPS_MATH_BF16 || math == SPS_MATH_BF16X3)
    return launch_bwd_wide_t<__nv_bfloat16>(s, z16, dz16, protos, last_w, dlogits, dz, dprotos, dlast_w, ws, ws_bytes, st);
  return launch_bwd_wide_t<__half>(s, z16, dz16, protos, last_w, dlogits, dz, dprotos, dlast_w, ws, ws_bytes, st);
}

}  // namespace sps
