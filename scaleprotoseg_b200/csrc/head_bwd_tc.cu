// Tensor-core prototype head (sm_100a): fused backward for the fused-mode step (upstream gradient = dlogits).
// What autograd derives from segmentation/model/model_multiscale.py:256-281, 324-330, 243-244 of the reference;
// d and a are recomputed from z and the prototypes instead of being read back from saved (N x P) maps.
//
//   S   = |z|^2 - 2 z.p + |p|^2            distance GEMM, as in the forward                     (tcgen05, K-major)
//   dA  = dlogits . W                       [128 n x 64 p] = DL[128 x 32k] . W3[64 x 32k]^T     (tcgen05, SW64 tiles)
//   d = max(S,0);  a' = -(1-eps)/((d+1)(d+eps));  a = log((d+1)/(d+eps));  gamma = [S>0] dA a'   (epilogue)
//   dz_s  += gamma . (-2 p)                 [128 n x 64 c], K = prototypes: B = the SAME packed prototype tile,
//                                           read as an MN-major operand                           (tcgen05)
//   dp^T  += gamma^T . z_s                  [64 p x 64 c], K = positions: gamma and z tiles read MN-major (M = 64)
//   dW^T  += a^T . dlogits                  [64 p x 32 k], K = positions                         (M = 64, lanes +16)
//   csum  += gamma^T . 1                    [64 p x 8]: the ones live in the z tile's norm columns
//   dz = (acc + 2 z sum_p gamma)/G,  dp = (2 p csum - 2 acc)/G,  dW = ln2 acc / G
//
// One CTA owns ONE scale (up to 3 prototype chunks of 64) and a strided set of position tiles, so the dp / dW /
// csum accumulators of its chunks stay in TMEM for the whole kernel and dz of a (tile, scale) is complete inside
// the CTA: no atomics on dz.  Partial dp / dW of the CTAs of a scale go to the workspace and a fixed-order
// second pass adds them (deterministic).  dlogits is scaled by a power of two G (from its max) so that the 16-bit
// gamma / dlogits tiles stay in the normal range; G is divided out at the end (exact).
//
// TMEM (512 columns): S x2 [0,128), dA x2 [128,256), dz [256,320), chunk j accumulators [320+64j, 384+64j):
//   lanes 0-15 (+32q): dp^T 64 columns; lanes 16-31 (+32q): dW^T columns [0,32), csum columns [32,40).
#include <type_traits>
#include "tma.cuh"

namespace sps {
namespace tcb {
using namespace sps::tc;

#ifdef SPS_TIMELINE
// Debug build only: clock stamps of CTA 0 (role r, event e) -> g_tlb[r * 1024 + e]; read by sps_debug_timeline_bwd.
__device__ long long g_tlb[8 * 1024];
#define TLB(role, ev)                                                                      \
  do {                                                                                     \
    if (blockIdx.x == 0 && (ev) < 1024) g_tlb[(role) * 1024 + (ev)] = clock64();            \
  } while (0)
// every CTA: global-timer stamps (ns) of kernel entry / main loop end / exit -> roles 5, 6, 7 indexed by blockIdx.x
__device__ __forceinline__ long long gtimer() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define TLB_CTA(role) do { if (blockIdx.x < 1024) g_tlb[(role) * 1024 + blockIdx.x] = gtimer(); } while (0)
#else
#define TLB(role, ev) do { } while (0)
#define TLB_CTA(role) do { } while (0)
#endif

constexpr int GROUP_CHUNKS = 3;   // chunks per CTA (TMEM budget)
constexpr int MAX_GROUPS = 64;
constexpr int KP_BWD = 32;        // classes per launch, padded (wider last layers run as slabs of 32 classes)
// per chunk: BH 8 KB | BN 2 KB | W3 4 KB | BL 8 KB (lo part of -2p; copied to shared memory only in the split mode)
constexpr int BIMG = CN * 128 + CN * 32 + CN * 64 + CN * 128;
constexpr int BOFF_BH = 0, BOFF_BN = CN * 128, BOFF_W3 = CN * 128 + CN * 32, BOFF_BL = CN * 128 + CN * 32 + CN * 64;
template <int NSPLIT> __host__ __device__ constexpr int bimg_smem() { return NSPLIT == 3 ? BIMG : BOFF_BL; }

struct GroupTable {
  int n_groups;
  int c0[MAX_GROUPS];        // first chunk
  short nch[MAX_GROUPS];     // chunks in the group (1..3)
  short scale[MAX_GROUPS];
  short cta0[MAX_GROUPS];    // first CTA of the group
  short ncta[MAX_GROUPS];    // CTAs working on the group
  short multi[MAX_GROUPS];   // 1: the scale is shared by several groups -> dz is accumulated with atomics
};

struct BwdParams {
  CUtensorMap tm_hi;            // ZM_TMA16 / ZM_TMAHL: [N][C] 16-bit map (or its hi part)
  CUtensorMap tm_lo;            // ZM_TMAHL in a split mode: lo part
  HeadShape s;
  ChunkTable ct;
  GroupTable gt;
  const float* z;               // the map (ZM_LDG: fp32 NCHW; TMA modes: the 16-bit [N][C] array / hi part)
  const void* z_lo;             // ZM_TMAHL in a split mode: lo part
  const float* protos;
  const float* dlogits;
  const float* ddist;           // [B, P, H, W] gradient arriving through the distance map (KLD loss), or nullptr
  const unsigned char* img;     // packed chunk images (BIMG each)
  const unsigned int* amax_bits;  // max |dlogits| as float bits
  float* dz;
  float* part_dp;               // [nsplit][P*D]
  float* part_dw;               // [nsplit][K*P]
  int dl_stride;                // floats between dlogits rows (the launch may cover a slab of <= 32 of the K classes)
  int dz_acc;                   // 1: dz is accumulated with atomics (several class slabs add into it)
  float* part_cs;               // [nsplit][P]   2/G * column sums of gamma (the 2 p csum term is applied by the reduce)
  int n_tiles;
};

template <typename X, int NSPLIT>
__global__ void __launch_bounds__(256) bwd_prep_kernel(const __grid_constant__ ChunkTable ct, int D, int P, int K,
                                                        const float* __restrict__ protos, const float* __restrict__ last_w,
                                                        unsigned char* __restrict__ img_base, const float* __restrict__ dlogits,
                                                        long long dl_count, const float* __restrict__ ddist, long long dd_count,
                                                        unsigned int* __restrict__ amax_out) {
  pdl_launch_dependents();  // the main kernel may start its prologue and feature loads; its readers of our output wait
  if ((int)blockIdx.x >= ct.n_chunks) {
    // blocks past the chunk images: max |dlogits| and max |ddist| (float4 loads when the base is 16-B aligned), one
    // atomicMax per warp
    const long long nb = (long long)gridDim.x - ct.n_chunks, b = (long long)blockIdx.x - ct.n_chunks;
    float m = 0.f;
    auto scan = [&](const float* __restrict__ x, long long count) {
      if (x == nullptr) return;
      if ((reinterpret_cast<uintptr_t>(x) & 15) == 0) {
        const long long n4 = count >> 2;
        const float4* x4 = reinterpret_cast<const float4*>(x);
        for (long long i = b * blockDim.x + threadIdx.x; i < n4; i += nb * blockDim.x) {
          const float4 v = __ldg(x4 + i);
          m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
        }
        for (long long i = (n4 << 2) + b * blockDim.x + threadIdx.x; i < count; i += nb * blockDim.x) m = fmaxf(m, fabsf(__ldg(x + i)));
      } else {
        for (long long i = b * blockDim.x + threadIdx.x; i < count; i += nb * blockDim.x) m = fmaxf(m, fabsf(__ldg(x + i)));
      }
    };
    scan(dlogits, dl_count);
    scan(ddist, dd_count);
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    // one atomic per BLOCK: the main kernel's prologue waits for this pre-pass, so it runs on many small blocks (a block
    // scans a few KB in one or two rounds of loads), and their atomics all hit one address
    __shared__ float wmax[8];
    if ((threadIdx.x & 31) == 0) wmax[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
      for (int w = 1; w < 8; ++w) m = fmaxf(m, wmax[w]);
      if (m > 0.f) atomicMax(amax_out, __float_as_uint(m));
    }
    return;
  }
  const int c = blockIdx.x, t = threadIdx.x;
  const int p0 = ct.p0[c], n = ct.n[c];
  unsigned char* img = img_base + (size_t)c * BIMG;
  __shared__ float pn[CN];
  {
    const int j = t >> 2, kq = (t & 3) * 16;
    float part = pack_row16<X, NSPLIT>(protos + (size_t)(p0 + (j < n ? j : 0)) * D, j < n, D, -2.0f, img + BOFF_BH, img + BOFF_BL, j, kq);
    part += __shfl_xor_sync(0xffffffffu, part, 1);
    part += __shfl_xor_sync(0xffffffffu, part, 2);
    if ((t & 3) == 0) pn[j] = part;
  }
  __syncthreads();
  if (t < CN) store_proto_norm<X>(img + BOFF_BN, t, pn[t], t < n);
  // all loads of a thread first, then the stores (a load -> store round per iteration was a chain of 8 dependent L2
  // round trips); prototype index fastest across threads so the loads are coalesced along the last-layer rows
  {
    constexpr int PER = CN * KP_BWD / 256;
    float v[PER];
#pragma unroll
    for (int u = 0; u < PER; ++u) {
      const int idx = t + u * 256, k = idx / CN, j = idx % CN;
      v[u] = (k < K && j < n) ? __ldg(last_w + (size_t)k * P + p0 + j) : 0.0f;
    }
#pragma unroll
    for (int u = 0; u < PER; ++u) {
      const int idx = t + u * 256, k = idx / CN, j = idx % CN;
      *reinterpret_cast<uint16_t*>(img + BOFF_W3 + sw64_off(j, k)) = bits16(Fmt<X>::from(v[u]));
    }
  }
}

// power-of-two scale that brings max|dlogits| to [2^-5, 2^-4)
__device__ __forceinline__ float grad_scale(unsigned int amax_bits) {
  const float a = __uint_as_float(amax_bits);
  if (!(a > 0.f) || !isfinite(a)) return 1.0f;
  int e;
  frexpf(a, &e);  // a = f * 2^e, f in [0.5, 1)
  int g = -4 - e;
  g = g > 100 ? 100 : (g < -100 ? -100 : g);
  return ldexpf(1.0f, g);
}

constexpr int NTHREADS = 480;
constexpr int WARP_LOAD0 = 8, WARP_I1 = 12, WARP_I2 = 13, WARP_RES = 14;
// A buffer: z tile 16 KB | norm 4 KB | DL tile 8 KB | z lo tile 16 KB (split mode)
constexpr int AOFF_Z = 0, AOFF_N = TILE_M * 128, AOFF_DL = TILE_M * 128 + TILE_M * 32, AOFF_ZL = TILE_M * 128 + TILE_M * 32 + TILE_M * 64;
constexpr int G_BUF = 2 * TILE_M * 128;                            // gamma tile 16 KB | a tile 16 KB
template <int NSPLIT> struct Lay {
  static constexpr int A_BUF = AOFF_ZL + (NSPLIT == 3 ? TILE_M * 128 : 0);
  static constexpr int RES = bimg_smem<NSPLIT>();                  // resident bytes per chunk
  static constexpr int SM_RES = 0;
  static constexpr int SM_A = (GROUP_CHUNKS * RES + 1023) / 1024 * 1024;
  static constexpr int SM_G = SM_A + 2 * A_BUF;
  static constexpr int SM_SG = SM_G + 2 * G_BUF;                   // sgbuf[4 tiles in flight][2 half / set][128] floats
  static constexpr int SM_BAR = SM_SG + 4 * 2 * 128 * 4;
  static constexpr int SM_TOTAL = SM_BAR + 256 + 1024;
};
constexpr int T_S = 0, T_DA = 128, T_DZ = 256, T_ACC = 320;
enum { B_AFULL = 0, B_AEMPTY = 2, B_SFULL = 4, B_SEMPTY = 6, B_GFULL = 8, B_GEMPTY = 10, B_DZFULL = 12, B_DZEMPTY = 13,
       B_RES = 14, B_ACC = 15, B_TFULL = 16, B_SGFULL = 18, B_COUNT = 22 };

// HAS_DD: a gradient also arrives through the dense distance map (prm.ddist; the KLD loss of stage-1 training).
template <typename X, int NSPLIT, bool HAS_DD, int ZMODE>
__global__ void __launch_bounds__(NTHREADS, 1) head_bwd_tc_kernel(const __grid_constant__ BwdParams prm) {
  using L = Lay<NSPLIT>;
  // the feature tile has a lo part: fp32 source in a split mode (a 16-bit map is exact in one term)
  constexpr bool ZLO = NSPLIT == 3 && ZMODE != ZM_TMA16;
  constexpr bool RD_BY_TILE_WARPS = ZMODE != ZM_LDG;
  constexpr int A_BUF = L::A_BUF, RES = L::RES, SM_RES = L::SM_RES, SM_A = L::SM_A, SM_G = L::SM_G, SM_SG = L::SM_SG,
                SM_BAR = L::SM_BAR;
  extern __shared__ unsigned char smem_raw[];
  const HeadShape& s = prm.s;
  const ChunkTable& ct = prm.ct;
  const GroupTable& gt = prm.gt;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* gbase = smem_raw + (base - smem_u32(smem_raw));
  auto bar = [&](int i) { return base + SM_BAR + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + SM_BAR + 8 * B_COUNT);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // which group (scale + up to 3 chunks) and which tiles this CTA owns
  int g = 0;
  while (g + 1 < gt.n_groups && (int)blockIdx.x >= gt.cta0[g + 1]) ++g;
  const int cta_local = (int)blockIdx.x - gt.cta0[g], ncta = gt.ncta[g];
  const int nch = gt.nch[g], c0 = gt.c0[g], sc = gt.scale[g];
  const bool multi = gt.multi[g] != 0 || prm.dz_acc != 0;
  const int my_tiles = prm.n_tiles > cta_local ? (prm.n_tiles - 1 - cta_local) / ncta + 1 : 0;
  const int U = my_tiles * nch;

  if (threadIdx.x == 0) {
    TLB_CTA(5);
    // TMA modes: the dz read-out (TMEM -> + 2 z sum(gamma) -> HBM) runs on the four tile warps, which have little else to
    // do once the tile arrives by TMA; on the epilogue warps it was 2 350 of the 9 000 clk tile period
    // (profiles/r02_timeline_bwd_tma16_before.txt).  The loader-fed mode keeps it on the epilogue warps: its tile warps
    // hold the next tile's 64 prefetched values in registers.
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(B_AFULL + i), 4);
      mbar_init(bar(B_AEMPTY + i), RD_BY_TILE_WARPS ? 1 : 9);   // commit of the tile's last GEMMs (+ the 8 epilogue warps whose read-out reads the z tile)
      mbar_init(bar(B_SGFULL + i), 8);   // row sums of gamma of a tile written by the 8 epilogue warps
      mbar_init(bar(B_SGFULL + 2 + i), 8);
      mbar_init(bar(B_SFULL + i), 1);
      // TMA modes: the two epilogue sets take ALTERNATE units (4 warps per unit); loader-fed mode: column halves of one unit
      mbar_init(bar(B_SEMPTY + i), RD_BY_TILE_WARPS ? 4 : 8);
      mbar_init(bar(B_GFULL + i), RD_BY_TILE_WARPS ? 4 : 8);
      mbar_init(bar(B_GEMPTY + i), 1);
    }
    mbar_init(bar(B_TFULL), 1);      // TMA modes: arrive.expect_tx + the tile bytes
    mbar_init(bar(B_TFULL + 1), 1);
    mbar_init(bar(B_DZFULL), 1);
    mbar_init(bar(B_DZEMPTY), RD_BY_TILE_WARPS ? 4 : 8);
    mbar_init(bar(B_RES), 1);
    mbar_init(bar(B_ACC), 1);
    mbar_fence_init();
  }
  if (warp == WARP_I1) tmem_alloc(smem_u32((const void*)tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  constexpr uint32_t FM = Fmt<X>::id;

  if (warp == WARP_RES) {
    pdl_wait();  // chunk images come from the pre-pass launched just before this kernel
    if (lane == 0) {
      mbar_arrive_expect_tx(bar(B_RES), (uint32_t)nch * RES);
      for (int j = 0; j < nch; ++j)
        bulk_g2s(base + SM_RES + j * RES, prm.img + (size_t)(c0 + j) * BIMG, RES, bar(B_RES));
    }
  } else if (warp == WARP_I1) {
    // ================= issuer 1: distance GEMM + dA GEMM of every unit =================
    constexpr uint32_t ID = idesc_f16(FM, TILE_M, CN);
    mbar_wait(bar(B_RES), 0);
    int j = 0, ti = 0;
    for (int u = 0; u < U; ++u) {
      const uint32_t ab = ti & 1, sb = u & 1;
      if (lane == 0) TLB(0, 4 * u);
      if (j == 0) mbar_wait(bar(B_AFULL + ab), (ti >> 1) & 1);
      mbar_wait(bar(B_SEMPTY + sb), ((u >> 1) & 1) ^ 1);
      if (lane == 0) TLB(0, 4 * u + 1);
      tc_fence_after();
      {
        const uint32_t ab0 = base + SM_A + ab * A_BUF, img = base + SM_RES + j * RES;
        const uint32_t dS = tmem + T_S + sb * CN, dA = tmem + T_DA + sb * CN;
        mma_f16(dS, desc_noswz(ab0 + AOFF_N, NORM_LBO, NORM_SBO), desc_noswz(img + BOFF_BN, NORM_LBO, NORM_SBO), ID, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_f16(dS, desc_sw128(ab0 + AOFF_Z + 32 * k), desc_sw128(img + BOFF_BH + 32 * k), ID, 1);
        if (NSPLIT == 3) {
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_f16(dS, desc_sw128(ab0 + AOFF_Z + 32 * k), desc_sw128(img + BOFF_BL + 32 * k), ID, 1);
          if (ZLO) {
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_f16(dS, desc_sw128(ab0 + AOFF_ZL + 32 * k), desc_sw128(img + BOFF_BH + 32 * k), ID, 1);
          }
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) mma_f16(dA, desc_k_sw64(ab0 + AOFF_DL + 32 * k), desc_k_sw64(img + BOFF_W3 + 32 * k), ID, k);
        mma_commit(bar(B_SFULL + sb));
        if (lane == 0) TLB(0, 4 * u + 2);
      }
      __syncwarp();
      if (++j == nch) { j = 0; ++ti; }
    }
  } else if (warp == WARP_I2) {
    // ================= issuer 2: dz, dp^T, dW^T, csum GEMMs (consume the gamma / a tiles) =================
    constexpr uint32_t ID4 = idesc_f16_major(FM, TILE_M, CN, 0, 1);
    constexpr uint32_t ID5A = idesc_f16_major(FM, 64, 64, 1, 1);
    constexpr uint32_t ID5B = idesc_f16_major(FM, 64, KP_BWD, 1, 1);
    constexpr uint32_t ID5C = idesc_f16_major(FM, 64, 8, 1, 1);
    mbar_wait(bar(B_RES), 0);
    int j = 0, ti = 0;
    for (int u = 0; u < U; ++u) {
      const uint32_t ab = ti & 1, gb = u & 1;
      if (lane == 0) TLB(1, 4 * u);
      mbar_wait(bar(B_GFULL + gb), (u >> 1) & 1);
      if (j == 0) mbar_wait(bar(B_DZEMPTY), (ti & 1) ^ 1);
      if (lane == 0) TLB(1, 4 * u + 1);
      tc_fence_after();
      {
        const uint32_t ab0 = base + SM_A + ab * A_BUF, img = base + SM_RES + j * RES;
        const uint32_t gam = base + SM_G + gb * G_BUF, act = gam + TILE_M * 128;
        const uint32_t acc = tmem + T_ACC + 64 * j;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          mma_f16(tmem + T_DZ, desc_sw128(gam + 32 * k), desc_mn_sw128(img + BOFF_BH + 2048 * k, 0), ID4, (j > 0 || k > 0) ? 1 : 0);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t accum = (ti > 0 || k > 0) ? 1 : 0;
          mma_f16(acc, desc_mn_sw128(gam + 2048 * k, 0), desc_mn_sw128(ab0 + AOFF_Z + 2048 * k, 0), ID5A, accum);
          mma_f16(acc + (16u << 16), desc_mn_sw128(act + 2048 * k, 0), desc_mn_sw64(ab0 + AOFF_DL + 1024 * k, 0), ID5B, accum);
          mma_f16(acc + (16u << 16) + 32, desc_mn_sw128(gam + 2048 * k, 0),
                  (uint64_t)(((ab0 + AOFF_N + 512 * k) & 0x3FFFF) >> 4) | ((uint64_t)(256 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) |
                      ((uint64_t)1 << 46),
                  ID5C, accum);
        }
        mma_commit(bar(B_GEMPTY + gb));
        if (lane == 0) TLB(1, 4 * u + 2);
        if (j == nch - 1) {
          mma_commit(bar(B_DZFULL));
          mma_commit(bar(B_AEMPTY + ab));
        }
        if (u == U - 1) mma_commit(bar(B_ACC));
      }
      __syncwarp();
      if (++j == nch) { j = 0; ++ti; }
    }
  } else if (warp >= WARP_LOAD0 && warp < WARP_I1) {
    // ================= loaders: z_s tile (16-bit, K-major swizzled) + norms + scaled dlogits tile =================
    // Software pipeline: the feature loads of tile ti+1 are issued as soon as tile ti has been handed over, and fly
    // while this warp waits for the next buffer.
    const int m = threadIdx.x - WARP_LOAD0 * 32;
    const uint32_t hw = (uint32_t)s.HW;
    auto row_of = [&](int ti) -> long long { return ((long long)cta_local + (long long)ti * ncta) * TILE_M + m; };
    auto row_ptr = [&](long long n) -> const float* {
      if (n >= s.N) n = s.N - 1;  // rows past the end read a valid row; their dlogits are zero and nothing is stored
      const uint32_t b = (uint32_t)n / hw, pix = (uint32_t)n % hw;   // N < 2^31 (checked by head_bwd_tc_supported)
      return prm.z + ((size_t)b * s.C + (size_t)sc * s.D) * s.HW + pix;
    };
    // raw dlogits row (zeros past the end / past K); scaled by G when it is packed
    auto load_dl = [&](long long n, float (&dl)[KP_BWD]) {
      const float* src = prm.dlogits + (n < s.N ? n : 0) * prm.dl_stride;
      const int kl = n < s.N ? s.K : 0;
#pragma unroll
      for (int k = 0; k < KP_BWD; ++k) dl[k] = k < kl ? __ldg(src + k) : 0.0f;
    };
    float G_loader = 1.0f;
    // dz of (tile tj, this scale) by the tile warps (TMA modes): acc + 2 z sum_p gamma, row m of the tile per thread.
    // Two steps with the NEXT tile's load between them: readout_begin empties the TMEM accumulator (the issuer may start
    // the next tile's dz GEMMs), readout_finish does the arithmetic and the stores.  z comes back from global memory (the
    // row is one 128-byte line, L2-resident: the tile was loaded from it a few microseconds earlier), so the shared-memory
    // stage is free as soon as the tile's GEMMs have completed and the next load never waits for a read-out.
    uint32_t r0[32], r1[32];
    auto readout_begin = [&](int tj) {
      const uint32_t lane_base = ((uint32_t)((warp & 3) * 32)) << 16;
      mbar_wait(bar(B_DZFULL), tj & 1);
      tc_fence_after();
      tmem_ld32(tmem + lane_base + T_DZ, r0);
      tmem_ld32(tmem + lane_base + T_DZ + 32, r1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(B_DZEMPTY));
    };
    auto readout_finish = [&](int tj) {
      const long long n = row_of(tj);
      if (prm.dz == nullptr || n >= s.N) return;
      const float invG_ = 1.0f / G_loader;
      mbar_wait(bar(B_SGFULL + (tj & 3)), (tj >> 2) & 1);
      const float* sgb = reinterpret_cast<const float*>(gbase + SM_SG) + (tj & 3) * 256;
      const float tot2 = 2.0f * (sgb[m] + sgb[128 + m]);
      const uint4* zg = reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned char*>(prm.z) + ((size_t)n * s.C + (size_t)sc * s.D) * 2);
      const uint4* zgl = reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned char*>(prm.z_lo) + ((size_t)n * s.C + (size_t)sc * s.D) * 2);
      const uint32_t bimg = (uint32_t)n / hw, pix = (uint32_t)n % hw;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint4 zq[4], zl[4];
#pragma unroll
        for (int gq = 0; gq < 4; ++gq) {
          zq[gq] = __ldg(zg + 4 * h + gq);
          if (ZLO) zl[gq] = __ldg(zgl + 4 * h + gq);
        }
        float dv[32];
#pragma unroll
        for (int gq = 0; gq < 4; ++gq) {
          const uint32_t zw[4] = {zq[gq].x, zq[gq].y, zq[gq].z, zq[gq].w};
          const uint32_t zlw[4] = {ZLO ? zl[gq].x : 0u, ZLO ? zl[gq].y : 0u, ZLO ? zl[gq].z : 0u, ZLO ? zl[gq].w : 0u};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float2 zf = Pack2<X>::unpack(zw[e]);
            if (ZLO) {
              const float2 zl2 = Pack2<X>::unpack(zlw[e]);
              zf.x += zl2.x;
              zf.y += zl2.y;
            }
            const int k = gq * 8 + 2 * e;
            dv[k] = fmaf(zf.x, tot2, __uint_as_float(h == 0 ? r0[k] : r1[k])) * invG_;
            dv[k + 1] = fmaf(zf.y, tot2, __uint_as_float(h == 0 ? r0[k + 1] : r1[k + 1])) * invG_;
          }
        }
        if (ZMODE == ZM_TMA16) {
          uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<unsigned char*>(prm.dz) + ((size_t)n * s.C + (size_t)sc * s.D + 32 * h) * 2);
#pragma unroll
          for (int gq = 0; gq < 4; ++gq)
            dst[gq] = make_uint4(Pack2<X>::pack(dv[gq * 8], dv[gq * 8 + 1]), Pack2<X>::pack(dv[gq * 8 + 2], dv[gq * 8 + 3]),
                                 Pack2<X>::pack(dv[gq * 8 + 4], dv[gq * 8 + 5]), Pack2<X>::pack(dv[gq * 8 + 6], dv[gq * 8 + 7]));
        } else {
          float* dzp = prm.dz + ((size_t)bimg * s.C + (size_t)sc * s.D + 32 * h) * s.HW + pix;
          const size_t pitch = (size_t)s.HW;
#pragma unroll
          for (int k = 0; k < 32; ++k) {
            if (multi) atomicAdd(dzp, dv[k]); else *dzp = dv[k];
            dzp += pitch;
          }
        }
      }
    };
    ZStage<X, NSPLIT> zst;
    float dl[KP_BWD];
    if (my_tiles > 0) {
      const long long n0 = row_of(0);
      if (ZMODE == ZM_LDG) {
        const float* p0 = row_ptr(n0);
        zst.issue_half(0, p0, hw, s.D);
        zst.issue_half(1, p0, hw, s.D);
      }
      if (ZMODE == ZM_LDG) load_dl(n0, dl);
    }
    pdl_wait();  // max |dlogits| comes from the pre-pass launched just before this kernel (the loads above are in flight)
    const float G = grad_scale(*reinterpret_cast<const volatile unsigned int*>(prm.amax_bits));
    G_loader = G;
    auto load_tile = [&](int ti) {
      const uint32_t ab = ti & 1;
      if (ZMODE != ZM_LDG) load_dl(row_of(ti), dl);   // TMA modes: the tile warps run two tiles ahead of the read-outs, no prefetch across tiles
      if (m == 0) TLB(2, 4 * ti);
      mbar_wait(bar(B_AEMPTY + ab), ((ti >> 1) & 1) ^ 1);
      if (m == 0) TLB(2, 4 * ti + 1);
      unsigned char* a0 = gbase + SM_A + ab * A_BUF;
      if (ZMODE != ZM_LDG && m == 0) {   // the buffer is free: one thread sends the tile's TMA (hi [+ lo] part)
        const int row0 = (int)(((long long)cta_local + (long long)ti * ncta) * TILE_M);
        mbar_arrive_expect_tx(bar(B_TFULL + ab), (ZLO ? 2u : 1u) * TILE_M * 128u);
        tma_load_2d(base + SM_A + ab * A_BUF + AOFF_Z, &prm.tm_hi, sc * s.D, row0, bar(B_TFULL + ab));
        if (ZLO) tma_load_2d(base + SM_A + ab * A_BUF + AOFF_ZL, &prm.tm_lo, sc * s.D, row0, bar(B_TFULL + ab));
      }
      // dlogits first: its 32 registers are dead while the feature row is converted
#pragma unroll
      for (int gq = 0; gq < 4; ++gq) {
        uint32_t w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) w[q] = Pack2<X>::pack(dl[gq * 8 + 2 * q] * G, dl[gq * 8 + 2 * q + 1] * G);
        *reinterpret_cast<uint4*>(a0 + AOFF_DL + (uint32_t)m * 64u + ((((uint32_t)gq) ^ (((uint32_t)m >> 1) & 3u)) << 4)) =
            make_uint4(w[0], w[1], w[2], w[3]);
      }
      float zn = 0.f;
      if (ZMODE == ZM_LDG) {
        zst.template convert_half<4>(0, m, a0 + AOFF_Z, a0 + AOFF_ZL, zn);
        if (m == 0) TLB(6, 512 + 2 * ti);
        zst.template convert_half<4>(1, m, a0 + AOFF_Z, a0 + AOFF_ZL, zn);
        if (m == 0) TLB(6, 513 + 2 * ti);
      } else {
        // the tile has landed in its final layout; what is left is the squared norm of this thread's row
        mbar_wait(bar(B_TFULL + ab), (ti >> 1) & 1);
        zn = tile_row_norm<X, ZLO>(a0 + AOFF_Z, a0 + AOFF_ZL, m);
      }
      ZStage<X, NSPLIT>::store_norm(m, a0 + AOFF_N, zn);
      if (m == 0) TLB(2, 4 * ti + 3);
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(B_AFULL + ab));
      if (m == 0) TLB(2, 4 * ti + 2);
      // the next tile's loads go out AFTER the arrive: mbarrier.arrive has release semantics and would wait for them
      if (ZMODE == ZM_LDG && ti + 1 < my_tiles) {
        const long long nn = row_of(ti + 1);
        const float* pn = row_ptr(nn);
        zst.issue_half(0, pn, hw, s.D);
        zst.issue_half(1, pn, hw, s.D);
        load_dl(nn, dl);
      }
    };
    if (!RD_BY_TILE_WARPS) {
      for (int ti = 0; ti < my_tiles; ++ti) load_tile(ti);
    } else {
      // Loads run two tiles ahead of the read-outs: a stage is free again when the GEMMs of its tile have completed (the
      // read-out takes z from L2, not from the stage), which is also when that tile's dz accumulator is complete.
      //   begin(t): wait for the accumulator, move it to registers, release it to the issuer
      //   load(t+2): into the stage tile t has just left
      //   finish(t): 2 z sum(gamma) + acc, stores
      if (my_tiles > 0) load_tile(0);
      if (my_tiles > 1) load_tile(1);
      for (int t = 0; t < my_tiles; ++t) {
        readout_begin(t);
        if (t + 2 < my_tiles) load_tile(t + 2);
        readout_finish(t);
      }
    }
  }
  else if (warp < WARP_LOAD0) {
    // ================= epilogue: gamma / a tiles, row sums, dz read-out, final accumulator read-out =================
    const int q = warp & 3, half = warp >> 2;
    const int m = q * 32 + lane;
    const uint32_t lane_base = ((uint32_t)(q * 32)) << 16;
    const float eps = s.eps, ome = 1.0f - s.eps;
    const bool is_log = s.act_kind == SPS_ACT_LOG;
    pdl_wait();  // max |dlogits| comes from the pre-pass launched just before this kernel
    const float G = grad_scale(*reinterpret_cast<const volatile unsigned int*>(prm.amax_bits)), invG = 1.0f / G;
    float* sgbuf = reinterpret_cast<float*>(gbase + SM_SG);
    // dz of (tile tj, this scale): acc + 2 z sum_p gamma.  Called one unit into the NEXT tile, so the epilogue never
    // sits waiting for the last dz GEMM of a tile (the issuer only needs the read-out before it overwrites T_DZ).
    auto dz_readout = [&](int tj) {
      const long long n = ((long long)cta_local + (long long)tj * ncta) * TILE_M + m;
      const bool valid = n < s.N;
      const long long b = valid ? n / s.HW : 0, pix = valid ? n % s.HW : 0;
      const uint32_t ab = tj & 1;
      const int ti = tj;
      float* sgb = sgbuf + (ti & 1) * 256;
      named_bar_sync(1, 256);
      const float tot = sgb[m] + sgb[128 + m];
      uint32_t rd[32];
      if (threadIdx.x == 0) TLB(4, 4 * tj);
      mbar_wait(bar(B_DZFULL), ti & 1);
      if (threadIdx.x == 0) TLB(4, 4 * tj + 1);
      tc_fence_after();
      tmem_ld32(tmem + lane_base + T_DZ + half * 32, rd);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(B_DZEMPTY));
      const unsigned char* zrow = gbase + SM_A + ab * A_BUF + AOFF_Z + (uint32_t)m * 128u;
      // values first (registers), then one tight store loop: the channel pointer advances by the map pitch
      float dv[32];
      const float tot2 = 2.0f * tot;
#pragma unroll
      for (int gq = 0; gq < 4; ++gq) {
        const uint32_t zo = (((uint32_t)(half * 4 + gq)) ^ ((uint32_t)m & 7u)) << 4;
        const uint4 zz = *reinterpret_cast<const uint4*>(zrow + zo);
        const uint32_t zw[4] = {zz.x, zz.y, zz.z, zz.w};
        uint32_t zlw[4] = {0u, 0u, 0u, 0u};
        if (ZLO) {
          const uint4 zl = *reinterpret_cast<const uint4*>(zrow + (AOFF_ZL - AOFF_Z) + zo);
          zlw[0] = zl.x, zlw[1] = zl.y, zlw[2] = zl.z, zlw[3] = zl.w;
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float2 zf = Pack2<X>::unpack(zw[e]);
          if (ZLO) {
            const float2 zl2 = Pack2<X>::unpack(zlw[e]);
            zf.x += zl2.x;
            zf.y += zl2.y;
          }
          dv[gq * 8 + 2 * e] = fmaf(zf.x, tot2, __uint_as_float(rd[gq * 8 + 2 * e])) * invG;
          dv[gq * 8 + 2 * e + 1] = fmaf(zf.y, tot2, __uint_as_float(rd[gq * 8 + 2 * e + 1])) * invG;
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(B_AEMPTY + ab));   // the shared-memory z tile has been read
      if (ZMODE == ZM_TMA16) {
        // dz mirrors z: 16-bit channels-last.  This thread owns 32 consecutive channels of row n: 64 contiguous bytes.
        if (prm.dz != nullptr && valid) {
          uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<unsigned char*>(prm.dz) +
                                                ((size_t)n * s.C + (size_t)sc * s.D + half * 32) * 2);
          if (!multi) {
#pragma unroll
            for (int gq = 0; gq < 4; ++gq)
              dst[gq] = make_uint4(Pack2<X>::pack(dv[gq * 8], dv[gq * 8 + 1]), Pack2<X>::pack(dv[gq * 8 + 2], dv[gq * 8 + 3]),
                                   Pack2<X>::pack(dv[gq * 8 + 4], dv[gq * 8 + 5]), Pack2<X>::pack(dv[gq * 8 + 6], dv[gq * 8 + 7]));
          }
        }
      } else if (prm.dz != nullptr && valid) {
        const int ch0 = half * 32;
        const int nlive = s.D - ch0 < 32 ? s.D - ch0 : 32;   // channels of this half that exist (D may be < 64)
        float* dzp = prm.dz + (b * (long long)s.C + (long long)sc * s.D + ch0) * s.HW + pix;
        const size_t pitch = (size_t)s.HW;
        if (!multi && nlive == 32) {
#pragma unroll
          for (int k = 0; k < 32; ++k) {
            *dzp = dv[k];
            dzp += pitch;
          }
        } else {
#pragma unroll
          for (int k = 0; k < 32; ++k) {
            if (k < nlive) {
              if (multi) atomicAdd(dzp, dv[k]); else *dzp = dv[k];
            }
            dzp += pitch;
          }
        }
      }
      if (threadIdx.x == 0) TLB(4, 4 * tj + 2);
    };
    if (RD_BY_TILE_WARPS) {
      // ---- TMA modes.  The two warp sets take ALTERNATE units, each walking both 32-column halves of its unit: the two
      // warps of an SM sub-partition are then out of phase (one computes while the other waits for its accumulators or
      // stores its tiles) instead of stalling together.  dz is read out by the tile warps; a set publishes its row sums
      // of gamma for a tile as soon as it has finished its last unit of that tile.
      const int set = half;
      int ti = set / nch, j = set % nch;
      float sg = 0.f;
      auto publish = [&](int t_from, int t_to, float v) {   // tiles [t_from, t_to): the first gets v, the rest (no unit of this set) 0
        for (int t = t_from; t < t_to; ++t) {
          sgbuf[(t & 3) * 256 + set * 128 + m] = t == t_from ? v : 0.f;
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(B_SGFULL + (t & 3)));
        }
      };
      if (U <= set) publish(0, my_tiles, 0.f);      // no unit at all for this set
      else if (set == 1) publish(0, ti, 0.f);      // (one chunk per tile: the second set has no unit in tile 0)
      for (int u = set; u < U; u += 2) {
        const uint32_t sb = u & 1, gb = u & 1, ph = (u >> 1) & 1;
        const long long n_row = ((long long)cta_local + (long long)ti * ncta) * TILE_M + m;
        const bool row_valid = n_row < s.N;
        const uint32_t row_b = row_valid ? (uint32_t)n_row / (uint32_t)s.HW : 0u, row_pix = row_valid ? (uint32_t)n_row % (uint32_t)s.HW : 0u;
        unsigned char* grow = gbase + SM_G + gb * G_BUF + (uint32_t)m * 128u;
#pragma unroll 1
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t rs[32], ra[32];
          float dd[HAS_DD ? 32 : 1];
          if (HAS_DD) {
            const int c = c0 + j, nlive = row_valid ? ct.n[c] - hf * 32 : 0;
            const float* src = prm.ddist + ((size_t)row_b * s.P + ct.p0[c] + hf * 32) * s.HW + row_pix;
#pragma unroll
            for (int k = 0; k < 32; ++k) dd[k] = k < nlive ? __ldg(src + (size_t)k * s.HW) : 0.0f;
          }
          if (hf == 0) {
            mbar_wait(bar(B_SFULL + sb), ph);
            tc_fence_after();
          }
          tmem_ld32(tmem + lane_base + T_S + sb * CN + hf * 32, rs);
          tmem_ld32(tmem + lane_base + T_DA + sb * CN + hf * 32, ra);
          tmem_ld_wait();
          if (hf == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(B_SEMPTY + sb));
          }
          uint32_t gw[16], aw[16];
          auto gamma_act = [&](auto log_t) {
            constexpr bool LOG = decltype(log_t)::value;
#pragma unroll
            for (int e = 0; e < 16; ++e) {
              float gm[2], a2[2];
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                const float sv = __uint_as_float(rs[2 * e + h]), dav = __uint_as_float(ra[2 * e + h]);
                const float d = fmaxf(sv, 0.0f);
                float ap, av;
                if (LOG) {
                  const float s1 = d + 1.0f, ru = rcp_approx(s1 * (d + eps));
                  ap = -ome * ru;
                  av = lg2_approx(s1 * s1 * ru);
                } else {
                  ap = -1.0f;
                  av = -d;
                }
                gm[h] = sv > 0.0f ? (HAS_DD ? fmaf(dav, ap, dd[2 * e + h] * G) : dav * ap) : 0.0f;
                a2[h] = av;
                sg += gm[h];
              }
              gw[e] = Pack2<X>::pack(gm[0], gm[1]);
              aw[e] = Pack2<X>::pack(a2[0], a2[1]);
            }
          };
          if (is_log) gamma_act(std::true_type{}); else gamma_act(std::false_type{});
          if (hf == 0) mbar_wait(bar(B_GEMPTY + gb), ph ^ 1);
#pragma unroll
          for (int gq = 0; gq < 4; ++gq) {
            const uint32_t off = (((uint32_t)(hf * 4 + gq)) ^ ((uint32_t)m & 7u)) << 4;
            *reinterpret_cast<uint4*>(grow + off) = make_uint4(gw[4 * gq], gw[4 * gq + 1], gw[4 * gq + 2], gw[4 * gq + 3]);
            *reinterpret_cast<uint4*>(grow + TILE_M * 128 + off) = make_uint4(aw[4 * gq], aw[4 * gq + 1], aw[4 * gq + 2], aw[4 * gq + 3]);
          }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(B_GFULL + gb));
        // next unit of this set; leaving the tile publishes its row sums
        int nti = ti, nj = j + 2;
        while (nj >= nch) { nj -= nch; ++nti; }
        const bool more = u + 2 < U;
        if (!more) nti = my_tiles;
        if (nti != ti) {
          publish(ti, nti, sg);
          sg = 0.f;
        }
        ti = nti;
        j = nj;
      }
    } else {
    int u = 0;
    for (int ti = 0; ti < my_tiles; ++ti) {
      float sg = 0.f;
      const long long n_row = ((long long)cta_local + (long long)ti * ncta) * TILE_M + m;
      const bool row_valid = n_row < s.N;
      const uint32_t row_b = row_valid ? (uint32_t)n_row / (uint32_t)s.HW : 0u, row_pix = row_valid ? (uint32_t)n_row % (uint32_t)s.HW : 0u;
      for (int j = 0; j < nch; ++j, ++u) {
        const uint32_t sb = u & 1, gb = u & 1;
        uint32_t rs[32], ra[32];
        float dd[HAS_DD ? 32 : 1];
        if (HAS_DD) {   // issued before the wait for the accumulators: the loads fly while the GEMMs finish
          const int c = c0 + j, nlive = row_valid ? ct.n[c] - half * 32 : 0;
          const float* src = prm.ddist + ((size_t)row_b * s.P + ct.p0[c] + half * 32) * s.HW + row_pix;
#pragma unroll
          for (int k = 0; k < 32; ++k) dd[k] = k < nlive ? __ldg(src + (size_t)k * s.HW) : 0.0f;
        }
        if (threadIdx.x == 0) TLB(3, 8 * u);
        mbar_wait(bar(B_SFULL + sb), (u >> 1) & 1);
        if (threadIdx.x == 0) TLB(3, 8 * u + 1);
        tc_fence_after();
        tmem_ld32(tmem + lane_base + T_S + sb * CN + half * 32, rs);
        tmem_ld32(tmem + lane_base + T_DA + sb * CN + half * 32, ra);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(B_SEMPTY + sb));
        if (threadIdx.x == 0) TLB(3, 8 * u + 2);
        uint32_t gw[16], aw[16];
        // the activation kind is resolved per unit, not per element: as a run-time select inside the loop both
        // formulas were evaluated for every element (20 instructions per element instead of 12)
        auto gamma_act = [&](auto log_t) {
          constexpr bool LOG = decltype(log_t)::value;
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            float gm[2], a2[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const float sv = __uint_as_float(rs[2 * e + h]), dav = __uint_as_float(ra[2 * e + h]);
              const float d = fmaxf(sv, 0.0f);
              float ap, av;
              if (LOG) {
                const float s1 = d + 1.0f, ru = rcp_approx(s1 * (d + eps));
                ap = -ome * ru;
                av = lg2_approx(s1 * s1 * ru);
              } else {
                ap = -1.0f;
                av = -d;
              }
              gm[h] = sv > 0.0f ? (HAS_DD ? fmaf(dav, ap, dd[2 * e + h] * G) : dav * ap) : 0.0f;
              a2[h] = av;
              sg += gm[h];
            }
            gw[e] = Pack2<X>::pack(gm[0], gm[1]);
            aw[e] = Pack2<X>::pack(a2[0], a2[1]);
          }
        };
        if (is_log) gamma_act(std::true_type{}); else gamma_act(std::false_type{});
        if (threadIdx.x == 0) TLB(3, 8 * u + 3);
        mbar_wait(bar(B_GEMPTY + gb), ((u >> 1) & 1) ^ 1);
        if (threadIdx.x == 0) TLB(3, 8 * u + 4);
        unsigned char* grow = gbase + SM_G + gb * G_BUF + (uint32_t)m * 128u;
#pragma unroll
        for (int gq = 0; gq < 4; ++gq) {
          const uint32_t off = (((uint32_t)(half * 4 + gq)) ^ ((uint32_t)m & 7u)) << 4;
          *reinterpret_cast<uint4*>(grow + off) = make_uint4(gw[4 * gq], gw[4 * gq + 1], gw[4 * gq + 2], gw[4 * gq + 3]);
          *reinterpret_cast<uint4*>(grow + TILE_M * 128 + off) = make_uint4(aw[4 * gq], aw[4 * gq + 1], aw[4 * gq + 2], aw[4 * gq + 3]);
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(B_GFULL + gb));
        if (threadIdx.x == 0) TLB(3, 8 * u + 5);
        if (j == 0 && ti > 0) dz_readout(ti - 1);
      }
      sgbuf[(ti & 1) * 256 + half * 128 + m] = sg;   // row sums of gamma of this tile (read after the named barrier in dz_readout)
    }
    if (my_tiles > 0) dz_readout(my_tiles - 1);
    }
    if (threadIdx.x == 0) TLB_CTA(6);
    pdl_launch_dependents();   // the reduce kernel may be scheduled as the CTAs drain (it waits for the whole grid)
    // ---- partial dp / dW of this CTA ----
    if (U > 0) {
      if (threadIdx.x == 0) TLB(4, 960);
      mbar_wait(bar(B_ACC), 0);
      if (threadIdx.x == 0) TLB(4, 961);
      tc_fence_after();
      // M = 64 accumulators: prototype row pl = q*16 + l sits in TMEM lane q*32 + l (dp^T, 64 channels) and lane
      // q*32 + 16 + l (dW^T in columns 0..31, the column sum of gamma in column 32).  Lanes 0-15 write -2/G * acc as
      // 16-byte pieces; lanes 16-31 write the dW columns (first warp set) or the column sums (second set) - the
      // 2 p csum term of dp is added by the reduce kernel, which reads the prototypes coalesced.  An earlier version
      // fetched them here, one dependent load per scalar store, and spent 25 us (20 % of the kernel) in this tail.
      const float wsc = (is_log ? 0.69314718055994531f : 1.0f) * invG;
      const int nlive = s.D - half * 32 < 32 ? s.D - half * 32 : 32;   // channels of this half that exist
      const float m2 = -2.0f * invG;
#pragma unroll
      for (int j = 0; j < GROUP_CHUNKS; ++j) {
        if (j < nch) {
          uint32_t racc[32];
          tmem_ld32(tmem + lane_base + T_ACC + 64 * j + half * 32, racc);
          tmem_ld_wait();
          const int c = c0 + j, p0 = ct.p0[c], nc = ct.n[c];
          if (lane < 16) {
            const int pl = q * 16 + lane;
            if (pl < nc) {
              float* out = prm.part_dp + (size_t)cta_local * s.P * s.D + (size_t)(p0 + pl) * s.D + half * 32;
#pragma unroll
              for (int e4 = 0; e4 < 8; ++e4)
                if (4 * e4 < nlive)   // D is a multiple of 8
                  *reinterpret_cast<float4*>(out + 4 * e4) =
                      make_float4(__uint_as_float(racc[4 * e4]) * m2, __uint_as_float(racc[4 * e4 + 1]) * m2,
                                  __uint_as_float(racc[4 * e4 + 2]) * m2, __uint_as_float(racc[4 * e4 + 3]) * m2);
            }
          } else {
            const int pl = q * 16 + lane - 16;
            if (pl < nc) {
              if (half == 0) {
                float* out = prm.part_dw + (size_t)cta_local * s.K * s.P + p0 + pl;
#pragma unroll
                for (int e = 0; e < 32; ++e)
                  if (e < s.K) out[(size_t)e * s.P] = __uint_as_float(racc[e]) * wsc;
              } else {
                prm.part_cs[(size_t)cta_local * s.P + p0 + pl] = -__uint_as_float(racc[0]) * m2;
              }
            }
          }
        }
      }
      if (threadIdx.x == 0) TLB(4, 963);
    }
  }

  if (lane == 0) TLB(4, 900 + warp);
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) TLB(4, 964);
  if (threadIdx.x == 0) TLB_CTA(7);
  if (warp == WARP_I1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// out1[i] = sum over the CTA partials of part1 (i < count1) + protos[i] * sum of the column-sum partials of its
// prototype (the 2 p csum term of dp); out2 likewise (plain sum) for the following count2 threads.  Either may be empty.
// Fixed order: deterministic.
__global__ void bwd_reduce_kernel(const float* __restrict__ part1, long long count1, float* __restrict__ out1,
                                  const float* __restrict__ part2, long long count2, float* __restrict__ out2, int nsplit,
                                  const float* __restrict__ part_cs, const float* __restrict__ protos, int D, int P,
                                  int accumulate_dp) {
  pdl_wait();   // launched while the main kernel drains: its partials are complete (and visible) after this
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const float* part = part1;
  float* out = out1;
  long long count = count1;
  bool is_dp = true;
  if (i >= count1) {
    i -= count1;
    part = part2, out = out2, count = count2;
    is_dp = false;
    if (i >= count2) return;
  }
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  int sp = 0;
  for (; sp + 4 <= nsplit; sp += 4) {
    a0 += part[(long long)sp * count + i];
    a1 += part[(long long)(sp + 1) * count + i];
    a2 += part[(long long)(sp + 2) * count + i];
    a3 += part[(long long)(sp + 3) * count + i];
  }
  for (; sp < nsplit; ++sp) a0 += part[(long long)sp * count + i];
  float r = (a0 + a1) + (a2 + a3);
  if (is_dp) {
    const long long p = i / D;
    float c0 = 0.f, c1 = 0.f;
    for (sp = 0; sp + 2 <= nsplit; sp += 2) {
      c0 += part_cs[(long long)sp * P + p];
      c1 += part_cs[(long long)(sp + 1) * P + p];
    }
    for (; sp < nsplit; ++sp) c0 += part_cs[(long long)sp * P + p];
    r = fmaf(protos[i], c0 + c1, r);
    if (accumulate_dp) r += out[i];   // a later class slab adds to the prototypes' gradient of the earlier ones
  }
  out[i] = r;
}


}  // namespace tcb

#ifdef SPS_TIMELINE
extern "C" int sps_debug_timeline_bwd(long long* host, int count) {
  return (int)cudaMemcpyFromSymbol(host, tcb::g_tlb, sizeof(long long) * (size_t)count);
}
#endif

// ---------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------
int build_chunk_table(const HeadShape& s, bool with_last_layer, tc::ChunkTable* ct);

static int build_group_table(const tc::ChunkTable& ct, int n_tiles, int n_sm, tcb::GroupTable* gt) {
  int ng = 0;
  for (int c = 0; c < ct.n_chunks;) {
    int e = c + 1;
    while (e < ct.n_chunks && ct.scale[e] == ct.scale[c] && e - c < tcb::GROUP_CHUNKS) ++e;
    SPS_REQUIRE(ng < tcb::MAX_GROUPS, SPS_ERR_UNSUPPORTED, "tensor-core backward: more than %d chunk groups", tcb::MAX_GROUPS);
    gt->c0[ng] = c;
    gt->nch[ng] = (short)(e - c);
    gt->scale[ng] = ct.scale[c];
    ++ng;
    c = e;
  }
  gt->n_groups = ng;
  for (int g = 0; g < ng; ++g) {
    bool multi = false;
    for (int h = 0; h < ng; ++h) multi |= (h != g && gt->scale[h] == gt->scale[g]);
    gt->multi[g] = multi ? 1 : 0;
  }
  // CTAs per group proportional to its chunks, at least 1, at most one per tile
  SPS_REQUIRE(ng <= n_sm, SPS_ERR_UNSUPPORTED, "tensor-core backward: %d chunk groups exceed %d SMs", ng, n_sm);
  int left = n_sm - ng, total = ct.n_chunks, next = 0;
  for (int g = 0; g < ng; ++g) {
    int extra = (int)((long long)(n_sm - ng) * gt->nch[g] / total);
    if (extra > left) extra = left;
    int n = 1 + extra;
    if (n > n_tiles) n = n_tiles;
    if (n < 1) n = 1;
    left -= (n - 1);
    gt->cta0[g] = (short)next;
    gt->ncta[g] = (short)n;
    next += n;
  }
  return next;  // grid size
}

static int bwd_nsplit(const tcb::GroupTable& gt) {
  int m = 1;
  for (int g = 0; g < gt.n_groups; ++g) m = gt.ncta[g] > m ? gt.ncta[g] : m;
  return m;
}

// plain launch of the fixed-order reduction of per-CTA partials (also used by head_bwd_wide.cu)
int launch_bwd_reduce(const float* part_dp, long long c1, float* dprotos, const float* part_dw, long long c2, float* dlast_w, int nsplit,
                      const float* part_cs, const float* protos, int D, int P, cudaStream_t st) {
  if (c1 + c2 <= 0) return SPS_OK;
  tcb::bwd_reduce_kernel<<<(unsigned)((c1 + c2 + 255) / 256), 256, 0, st>>>(part_dp, c1, dprotos, part_dw, c2, dlast_w, nsplit, part_cs,
                                                                           protos, D, P, 0);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

bool head_bwd_tc_supported(const HeadShape& s, int math, const float* last_w, const float* dlogits, const float* ddist,
                           const float* dact) {
  if (!(math != SPS_MATH_FP32 && last_w != nullptr && dlogits != nullptr && dact == nullptr && s.K <= 8 * tcb::KP_BWD &&
        s.D <= 64 && s.D % 8 == 0 && s.N > 0 && s.N < (1ll << 31) && s.P > 0))
    return false;
  // the chunk / group tables must fit as well (the forward checks the same limits): otherwise the fp32 kernels run
  tc::ChunkTable ct;
  int nc = 0;
  for (int sc = 0; sc < s.S; ++sc) nc += (s.scale_lo[sc + 1] - s.scale_lo[sc] + tc::CN - 1) / tc::CN;
  if (nc > tc::MAX_CHUNKS) return false;
  int ng = 0;
  for (int sc = 0; sc < s.S; ++sc) {
    const int c = (s.scale_lo[sc + 1] - s.scale_lo[sc] + tc::CN - 1) / tc::CN;
    ng += (c + tcb::GROUP_CHUNKS - 1) / tcb::GROUP_CHUNKS;
  }
  (void)ct;
  return ng <= tcb::MAX_GROUPS && ng <= sm_count();
}

size_t head_bwd_tc_workspace(const HeadShape& s) {
  tc::ChunkTable ct;
  if (build_chunk_table(s, true, &ct) != SPS_OK) return 0;
  tcb::GroupTable gt;
  const int n_tiles = (int)((s.N + tc::TILE_M - 1) / tc::TILE_M);
  if (build_group_table(ct, n_tiles > 0 ? n_tiles : 1, sm_count(), &gt) <= 0) return 0;
  const size_t nsplit = (size_t)bwd_nsplit(gt);
  return 1024 + (size_t)ct.n_chunks * tcb::BIMG + 256 + sizeof(float) * nsplit * ((size_t)s.P * s.D + (size_t)s.K * s.P + (size_t)s.P);
}

// One launch sequence (pre-pass, main kernel, reduce) over a SLAB of <= KP_BWD classes [k0, k0 + s.K) of the caller's
// K_total: gamma, dz and dprototypes are linear in dlogits, so a wider last layer is the sum of its slabs (dz by
// atomics, dprototypes by the reduce kernel's accumulate flag); every dW row belongs to exactly one slab.
//   s        : the head shape with K = the slab width
//   last_w   : row k0 of the (K_total x P) last layer;   dlast_w: row k0 of its gradient (or null)
//   dlogits  : column k0 of the (N x K_total) upstream gradient, rows dl_stride = K_total floats apart
//   dl_all / dl_all_count : the whole dlogits buffer (scanned for max |.|)
template <typename X, int NSPLIT>
static int launch_bwd_slab(const HeadShape& s, int zmode, const void* z, const void* z_lo, const float* protos, const float* last_w,
                           const float* dlogits, int dl_stride, const float* dl_all, long long dl_all_count, const float* ddist,
                           void* dz, float* dprotos, float* dlast_w, bool first_slab, bool several_slabs, void* ws, size_t ws_bytes,
                           cudaStream_t st) {
  tcb::BwdParams prm;
  int rc = build_chunk_table(s, true, &prm.ct);
  if (rc != SPS_OK) return rc;
  prm.n_tiles = (int)((s.N + tc::TILE_M - 1) / tc::TILE_M);
  const int grid = build_group_table(prm.ct, prm.n_tiles, sm_count(), &prm.gt);
  if (grid <= 0) return SPS_ERR_UNSUPPORTED;
  const int nsplit = bwd_nsplit(prm.gt);
  const size_t need = head_bwd_tc_workspace(s);
  SPS_REQUIRE(ws != nullptr && ws_bytes >= need, SPS_ERR_WORKSPACE, "head_bwd workspace: need %zu bytes, got %zu", need, ws_bytes);
  unsigned char* img = reinterpret_cast<unsigned char*>(((uintptr_t)ws + 1023) & ~(uintptr_t)1023);
  unsigned int* amax = reinterpret_cast<unsigned int*>(img + (size_t)prm.ct.n_chunks * tcb::BIMG);
  float* part_dp = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(amax) + 256);
  float* part_dw = part_dp + (size_t)nsplit * s.P * s.D;
  float* part_cs = part_dw + (size_t)nsplit * s.K * s.P;
  const size_t part_bytes = sizeof(float) * (size_t)nsplit * ((size_t)s.P * s.D + (size_t)s.K * s.P + (size_t)s.P);
  // every partial slot is written by exactly one CTA when all groups run the same number of CTAs and each CTA owns at
  // least one tile (the usual case); only then can the 9 MB clear of the partial buffers be skipped
  bool all_written = true;
  for (int g = 0; g < prm.gt.n_groups; ++g) all_written &= prm.gt.ncta[g] == nsplit && prm.n_tiles >= prm.gt.ncta[g];
  SPS_CUDA_TRY(cudaMemsetAsync(amax, 0, all_written ? 256 : 256 + part_bytes, st));
  bool any_multi = false;
  for (int g = 0; g < prm.gt.n_groups; ++g) any_multi |= prm.gt.multi[g] != 0;
  const bool dz_atomic = any_multi || several_slabs;
  SPS_REQUIRE(!(zmode == tc::ZM_TMA16 && dz_atomic && dz != nullptr), SPS_ERR_UNSUPPORTED,
              "16-bit channels-last backward: a scale with more than %d prototypes or more than %d classes needs dz accumulation, "
              "which the 16-bit dz layout does not support (P=%d, K=%d)", tcb::GROUP_CHUNKS * tc::CN, tcb::KP_BWD, s.P, s.K);
  bool empty_scale = false;   // a scale without prototypes has no chunk group: its block of dz is zero
  for (int sc = 0; sc < s.S; ++sc) empty_scale |= s.scale_lo[sc + 1] == s.scale_lo[sc];
  if (dz != nullptr && first_slab && (dz_atomic || empty_scale))
    SPS_CUDA_TRY(cudaMemsetAsync(dz, 0, (zmode == tc::ZM_TMA16 ? 2 : 4) * (size_t)s.B * s.C * s.HW, st));
  if (zmode != tc::ZM_LDG) {
    const bool bf = std::is_same<X, __nv_bfloat16>::value;
    rc = make_tmap_2d_16bit(&prm.tm_hi, z, bf, s.N, s.C, tc::TILE_M, 64, 128);
    if (rc != SPS_OK) return rc;
    rc = make_tmap_2d_16bit(&prm.tm_lo, z_lo != nullptr ? z_lo : z, bf, s.N, s.C, tc::TILE_M, 64, 128);
    if (rc != SPS_OK) return rc;
  }
  // one launch: chunk images (n_chunks blocks) + max |dlogits| (the remaining blocks)
  // (the amax blocks also scan the dense ddist map when there is one: enough of them to stream it at HBM rate)
  const int amax_blocks = sm_count() * (ddist != nullptr ? 16 : 4);
  tcb::bwd_prep_kernel<X, NSPLIT><<<prm.ct.n_chunks + amax_blocks, 256, 0, st>>>(prm.ct, s.D, s.P, s.K, protos, last_w, img, dl_all,
                                                                                dl_all_count, ddist,
                                                                                (long long)s.N * s.P, amax);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  prm.s = s;
  prm.z = reinterpret_cast<const float*>(z);
  prm.z_lo = z_lo != nullptr ? z_lo : z;
  prm.protos = protos;
  prm.dlogits = dlogits;
  prm.dl_stride = dl_stride;
  prm.dz_acc = several_slabs ? 1 : 0;
  prm.ddist = ddist;
  prm.img = img;
  prm.amax_bits = amax;
  prm.dz = reinterpret_cast<float*>(dz);
  prm.part_dp = part_dp;
  prm.part_dw = part_dw;
  prm.part_cs = part_cs;
  constexpr int SMEM = tcb::Lay<NSPLIT>::SM_TOTAL;
  static_assert(SMEM <= 227 * 1024, "backward kernel shared memory");
#define SPS_BWD_LAUNCH(DD, ZM)                                                                                                     \
  do {                                                                                                                             \
    SPS_CUDA_TRY(cudaFuncSetAttribute(tcb::head_bwd_tc_kernel<X, NSPLIT, DD, ZM>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM)); \
    SPS_CUDA_TRY(launch_pdl(tcb::head_bwd_tc_kernel<X, NSPLIT, DD, ZM>, dim3(grid), dim3(tcb::NTHREADS), (size_t)SMEM, st, prm));      \
  } while (0)
  if (ddist != nullptr) {
    if (zmode == tc::ZM_LDG) SPS_BWD_LAUNCH(true, tc::ZM_LDG);
    else if (zmode == tc::ZM_TMA16) SPS_BWD_LAUNCH(true, tc::ZM_TMA16);
    else SPS_BWD_LAUNCH(true, tc::ZM_TMAHL);
  } else {
    if (zmode == tc::ZM_LDG) SPS_BWD_LAUNCH(false, tc::ZM_LDG);
    else if (zmode == tc::ZM_TMA16) SPS_BWD_LAUNCH(false, tc::ZM_TMA16);
    else SPS_BWD_LAUNCH(false, tc::ZM_TMAHL);
  }
#undef SPS_BWD_LAUNCH
  count_launch();
  {
    const long long c1 = dprotos != nullptr ? (long long)s.P * s.D : 0, c2 = dlast_w != nullptr ? (long long)s.K * s.P : 0;
    if (c1 + c2 > 0) {
      SPS_CUDA_TRY(launch_pdl(tcb::bwd_reduce_kernel, dim3((unsigned)((c1 + c2 + 255) / 256)), dim3(256), (size_t)0, st,
                              (const float*)part_dp, c1, dprotos, (const float*)part_dw, c2, dlast_w, nsplit,
                              (const float*)part_cs, protos, s.D, s.P, first_slab ? 0 : 1));
      count_launch();
    }
  }
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

template <typename X, int NSPLIT>
static int launch_bwd_tc_t(const HeadShape& s, int zmode, const void* z, const void* z_lo, const float* protos, const float* last_w,
                           const float* dlogits, const float* ddist, void* dz, float* dprotos, float* dlast_w, void* ws,
                           size_t ws_bytes, cudaStream_t st) {
  const int n_slabs = (s.K + tcb::KP_BWD - 1) / tcb::KP_BWD;
  for (int sl = 0; sl < n_slabs; ++sl) {
    const int k0 = sl * tcb::KP_BWD;
    HeadShape ss = s;
    ss.K = s.K - k0 < tcb::KP_BWD ? s.K - k0 : tcb::KP_BWD;
    const int rc = launch_bwd_slab<X, NSPLIT>(ss, zmode, z, z_lo, protos, last_w + (size_t)k0 * s.P, dlogits + k0, s.K, dlogits,
                                              (long long)s.N * s.K, sl == 0 ? ddist : nullptr, dz, dprotos,
                                              dlast_w != nullptr ? dlast_w + (size_t)k0 * s.P : nullptr, sl == 0, n_slabs > 1, ws,
                                              ws_bytes, st);
    if (rc != SPS_OK) return rc;
  }
  return SPS_OK;
}

// zmode: tc::ZM_LDG (z = fp32 NCHW map), ZM_TMA16 (z = 16-bit channels-last map, dz likewise), ZM_TMAHL (z / z_lo = hi / lo
// re-lay of an fp32 map, dz fp32 NCHW)
int launch_head_bwd_tc(const HeadShape& s, int math, int zmode, const void* z, const void* z_lo, const float* protos,
                       const float* last_w, const float* dlogits, const float* ddist, void* dz, float* dprotos, float* dlast_w,
                       void* ws, size_t ws_bytes, cudaStream_t st) {
  SPS_REQUIRE(zmode == tc::ZM_LDG || s.D == 64, SPS_ERR_UNSUPPORTED, "TMA-fed backward: D=%d, only 64 channels per scale", s.D);
  // the split modes recompute the distances with the 3-term product (a' = da/dd is as sensitive to d as a is);
  // the gradient GEMMs themselves run single-term 16-bit operands with fp32 accumulation
  switch (math) {
    case SPS_MATH_BF16: return launch_bwd_tc_t<__nv_bfloat16, 1>(s, zmode, z, z_lo, protos, last_w, dlogits, ddist, dz, dprotos, dlast_w, ws, ws_bytes, st);
    case SPS_MATH_BF16X3: return launch_bwd_tc_t<__nv_bfloat16, 3>(s, zmode, z, z_lo, protos, last_w, dlogits, ddist, dz, dprotos, dlast_w, ws, ws_bytes, st);
    case SPS_MATH_F16X3: return launch_bwd_tc_t<__half, 3>(s, zmode, z, z_lo, protos, last_w, dlogits, ddist, dz, dprotos, dlast_w, ws, ws_bytes, st);
    default: return launch_bwd_tc_t<__half, 1>(s, zmode, z, z_lo, protos, last_w, dlogits, ddist, dz, dprotos, dlast_w, ws, ws_bytes, st);
  }
}

}  // namespace sps
