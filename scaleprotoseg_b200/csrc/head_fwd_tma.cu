// Tensor-core prototype head (sm_100a): fused forward FED BY TMA from 16-bit channels-last feature maps.
//
// Same arithmetic as head_tc.cu (distance GEMM with the squared norms riding in one extra K step, log-similarity
// epilogue, logits GEMM accumulated in TMEM over every chunk of a tile; segmentation/model/model_multiscale.py:256-317,
// 324-330, 243-244 of the reference), but the feature tile is not built by loader warps: the map is a 16-bit
// [N positions][C channels] array (the user's fp16 / bf16 channels-last tensor, or the hi / lo re-lay of an fp32 NCHW map
// written by z_pack_kernel), and a {64 channels x 128 positions} box with the 128-byte swizzle is exactly the K-major
// SWIZZLE_128B operand tile of the distance GEMM: one cp.async.bulk.tensor per tile, no conversion work on the SM.
// What is left of the loader's job is the row norm |z_s(n)|^2 (four warps, one row per thread, read back from the landed
// tile), written into the 16-wide norm K step as before.  The four loader warps of head_tc.cu are gone, so a third set
// of epilogue warps fits in the register file: three warps per SM sub-partition take units round-robin (the epilogue -
// two MUFU per element - is what bounds this kernel).
//
// F32 mode: the same kernel on the reference's own fp32 NCHW maps.  The pair's tile arrives as four TMA boxes (channel
// residues mod 4, tma.cuh) in one staging buffer; the four tile warps convert their position's channels to hi / lo 16-bit
// pairs and write them, with the squared norm, into a TMEM operand buffer (tcgen05.st: lane = position), and the distance
// MMAs take their A operand from TMEM.  Position tiles do not cross maps in this mode.  
//
// Warp roles (NSETS = 3 fused, 2 dense): warps [0, 4*NSETS) epilogue (warp % 4 = TMEM lane quarter, warp / 4 = set; set s
// takes units u = s mod NSETS), then four feature-tile warps (TMA issue + row norms / conversion), the prototype-chunk producer, the
// distance-GEMM issuer, the logits-GEMM issuer and the last-layer tile producer.  Every set owns ITS distance accumulator (TMEM) and activation tile
// (shared memory): a barrier is then only ever waited on by one set, one phase after the other (a parity wait cannot
// tell phase k from phase k + 2, so a set must never skip a phase of a barrier it shares).
#include "tma.cuh"

namespace sps {
namespace tf {
using namespace sps::tc;

#ifdef SPS_TIMELINE
// Debug build only: clock stamps of CTA 0, row = role * 8 + event kind, column = unit / pair index (profiles/tools/timeline_tma.py)
__device__ long long g_tl2[64 * 1024];
#define TL2(row, idx)                                                                                   \
  do {                                                                                                  \
    if (blockIdx.x == 0 && (idx) >= 0 && (idx) < 1024) g_tl2[(row) * 1024 + (idx)] = clock64();          \
  } while (0)
#else
#define TL2(row, idx) do { } while (0)
#endif

struct FwdTmaParams {
  CUtensorMap tm_hi;   // [N][C] 16-bit: the map itself, or the hi part of an fp32 map
  CUtensorMap tm_lo;   // lo part (ZLO only)
  TmapNchwF32 tm_f32;  // F32 mode: the fp32 NCHW map itself, four tensor maps (tma.cuh)
  int tiles_per_map;   // F32 mode: position tiles do not cross maps (a TMA box belongs to one map)
  HeadShape s;
  ChunkTable ct;
  const unsigned char* img;
  float* logits;
  float* dist;
  float* act;
  int n_tiles;
};

constexpr int MAX_NST = 8;
constexpr int A_TILE = TILE_M * 128;          // 16 KB: 128 positions x 64 channels, 16 bit
constexpr int A_NORM = TILE_M * 32;           // 4 KB norm K step
constexpr int P_BYTES = TILE_M * 128;
constexpr int TMEM_S0 = 0;
constexpr int MAX_NSTW = 4;   // last-layer tile ring: as deep as it is cheap (small K), the chunk producer runs that far ahead
constexpr int ACT_SCRATCH = 32 * 33 * 4;

// F32 mode (fp32 NCHW source): the "A stage" in shared memory is ONE staging buffer for the four fp32 boxes of a (tile,
// scale) pair (64 channel rows of 132 floats, tma.cuh); the operand itself lives in TMEM (two buffers), written by the
// tile warps with tcgen05.st after the fp32 -> hi / lo conversion.
constexpr int STG_PX = TILE_M + 4, STG_BYTES = 64 * STG_PX * 4;   // 33 KB
template <bool ZLO, bool F32> __host__ __device__ constexpr int a_stage_bytes() { return F32 ? STG_BYTES : (ZLO ? 2 : 1) * A_TILE + A_NORM; }
template <bool ZLO, bool F32> __host__ __device__ constexpr int a_stages() { return F32 ? 1 : (ZLO ? 2 : 3); }
template <int NSPLIT> __host__ __device__ constexpr int a_tmem_cols() { return NSPLIT == 3 ? 72 : 40; }   // hi 32 [+ lo 32] + norm 8

struct Plan {
  int a_stride, na, p0, p_stride, b0, b_stride, nst, nstw, w0, w_stride, scr0, bar_off, total;
  __host__ __device__ int a_off(int i) const { return i * a_stride; }
  __host__ __device__ int p_off(int i) const { return p0 + i * p_stride; }
  __host__ __device__ int b_off(int i) const { return b0 + i * b_stride; }
  __host__ __device__ int w_off(int i) const { return w0 + i * w_stride; }
};
// LSPLIT: terms of the activation / last-layer operands of the logits GEMM (3 = hi + lo as the distance products, 1 = one
// 16-bit term: wide last layers, whose hi + lo tiles do not fit shared memory; logits then carry ~5e-4 of sum |a W|)
template <int NSPLIT, bool ZLO, int NSETS, int LSPLIT, bool F32 = false> __host__ __device__ inline Plan make_plan(int kp, int scratch_warps) {
  Plan pl;
  pl.a_stride = a_stage_bytes<ZLO, F32>();
  pl.na = a_stages<ZLO, F32>();
  pl.p0 = pl.na * pl.a_stride;
  pl.p_stride = (LSPLIT == 3 ? 2 : 1) * P_BYTES;
  pl.b0 = pl.p0 + NSETS * pl.p_stride;
  pl.b_stride = ((NSPLIT == 3 ? 2 : 1) * CN * 128 + CN * 32 + 1023) / 1024 * 1024;
  pl.w_stride = ((LSPLIT == 3 ? 2 : 1) * kp * 128 + 1023) / 1024 * 1024;
  pl.nstw = pl.w_stride <= 8 * 1024 ? MAX_NSTW : 2;
  const int w_total = kp > 0 ? pl.nstw * pl.w_stride : 0;
  const int scr_total = scratch_warps * ACT_SCRATCH;
  int nst = (227 * 1024 - 512 - 1024 - pl.b0 - w_total - scr_total) / pl.b_stride;
  pl.nst = nst > MAX_NST ? MAX_NST : nst;
  pl.w0 = pl.b0 + (pl.nst > 0 ? pl.nst : 0) * pl.b_stride;
  pl.scr0 = pl.w0 + w_total;
  pl.bar_off = pl.scr0 + scr_total;
  pl.total = pl.bar_off + 512 + 1024;
  return pl;
}

enum { BAR_TFULL = 0, BAR_AFULL = 4, BAR_AEMPTY = 8, BAR_BFULL = 12, BAR_BEMPTY = 12 + MAX_NST, BAR_SFULL = 12 + 2 * MAX_NST,
       BAR_SEMPTY = 16 + 2 * MAX_NST, BAR_PFULL = 20 + 2 * MAX_NST, BAR_PEMPTY = 24 + 2 * MAX_NST, BAR_LFULL = 28 + 2 * MAX_NST,
       BAR_LEMPTY = 30 + 2 * MAX_NST, BAR_WFULL = 32 + 2 * MAX_NST, BAR_WEMPTY = 36 + 2 * MAX_NST, BAR_COUNT = 40 + 2 * MAX_NST };

constexpr int NZ = 4;   // feature-tile warps: one issues the TMA, all four compute the row norms (one row per thread)
__host__ __device__ constexpr int n_threads(int nsets) { return (4 * nsets + NZ + 4) * 32; }

template <typename X, int NSPLIT, bool ZLO, bool DENSE, int NSETS, int LSPLIT, bool F32 = false>
__global__ void __launch_bounds__(n_threads(NSETS), 1) head_fwd_tma_kernel(const __grid_constant__ FwdTmaParams prm) {
  static_assert(!F32 || ZLO == (NSPLIT == 3), "an fp32 source has a lo part exactly in the split modes");
  constexpr int NEPI = 4 * NSETS;
  constexpr int TMEM_L0 = NSETS * CN;
  constexpr int WARP_Z = NEPI, WARP_PROD = NEPI + NZ, WARP_MMA = NEPI + NZ + 1, WARP_MMA2 = NEPI + NZ + 2, WARP_PRODW = NEPI + NZ + 3;
  constexpr int NA = F32 ? 2 : a_stages<ZLO, false>();     // F32: operand buffers in TMEM
  constexpr int A_COLS = a_tmem_cols<NSPLIT>();
  static_assert(WARP_Z % 4 == 0, "tile warp w writes TMEM lanes 32 (w % 4) ..");
  extern __shared__ unsigned char smem_raw[];
  const HeadShape& s = prm.s;
  const ChunkTable& ct = prm.ct;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* gbase = smem_raw + (base - smem_u32(smem_raw));
  const Plan pl = make_plan<NSPLIT, ZLO, NSETS, LSPLIT, F32>(ct.kp, prm.act != nullptr ? NEPI : 0);
  const uint32_t bar0 = base + pl.bar_off;
  auto bar = [&](int i) { return bar0 + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + pl.bar_off + 8 * BAR_COUNT);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int NC = ct.n_chunks, KP = ct.kp, NST = pl.nst, NSTW = pl.nstw;
  const bool has_w = KP > 0;
  uint32_t tmem_cols = 32;
  const uint32_t TMEM_A0 = TMEM_L0 + 2 * KP;   // F32: two operand buffers of A_COLS columns
  while ((int)tmem_cols < TMEM_L0 + 2 * KP + (F32 ? 2 * A_COLS : 0)) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NA; ++i) {
      mbar_init(bar(BAR_TFULL + i), 1);    // arrive.expect_tx by the producer lane + the TMA bytes
      mbar_init(bar(BAR_AFULL + i), NZ);   // norm K step written (one arrive per feature-tile warp)
      mbar_init(bar(BAR_AEMPTY + i), 1);   // tcgen05.commit
    }
    if (F32) mbar_init(bar(BAR_TFULL + 2), NZ);   // F32: BAR_TFULL+0 = staging buffer landed, +2 = read out by the four tile warps
    for (int i = 0; i < NSETS; ++i) {
      mbar_init(bar(BAR_SFULL + i), 1);
      mbar_init(bar(BAR_SEMPTY + i), 4);   // the four warps of the set that owns the buffer
      mbar_init(bar(BAR_PFULL + i), 4);
      mbar_init(bar(BAR_PEMPTY + i), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(BAR_LFULL + i), 1);
      mbar_init(bar(BAR_LEMPTY + i), 4);
    }
    for (int i = 0; i < MAX_NSTW; ++i) {
      mbar_init(bar(BAR_WFULL + i), 1);
      mbar_init(bar(BAR_WEMPTY + i), 1);
    }
    for (int i = 0; i < MAX_NST; ++i) {
      mbar_init(bar(BAR_BFULL + i), 1);
      mbar_init(bar(BAR_BEMPTY + i), 1);
    }
    mbar_fence_init();
  }
  if (warp == WARP_MMA) tmem_alloc(smem_u32((const void*)tmem_slot), tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const int my_tiles = prm.n_tiles > (int)blockIdx.x ? (prm.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  // tile tj of this CTA: global position of its row 0 and the number of its rows that exist
  auto tile_rows = [&](int tj, long long& n0) -> int {
    const long long T = (long long)blockIdx.x + (long long)tj * gridDim.x;
    if (F32) {
      const long long b = T / prm.tiles_per_map;
      const int pix0 = (int)(T - b * prm.tiles_per_map) * TILE_M;
      n0 = b * s.HW + pix0;
      return s.HW - pix0 < TILE_M ? s.HW - pix0 : TILE_M;
    }
    n0 = T * TILE_M;
    return s.N - n0 < TILE_M ? (int)(s.N - n0) : TILE_M;
  };

  if (warp == WARP_MMA) {
    // ================================ distance-GEMM issuer ================================
    constexpr uint32_t ID1 = idesc_f16(Fmt<X>::id, TILE_M, CN);
    const int U = my_tiles * NC;
    uint32_t acount = 0, ab = 0, aph = 0, st = 0, bph = 0, sb = 0, sph = 0;
    int c = 0;
    for (int i = 0; i < U; ++i) {
      const bool first_of_scale = c == 0 || ct.scale[c] != ct.scale[c - 1];
      const bool last_of_scale = c == NC - 1 || ct.scale[c + 1] != ct.scale[c];
      if (lane == 0) TL2(0, i);
      if (first_of_scale) {
        if (!F32) mbar_wait(bar(BAR_TFULL + ab), aph);
        mbar_wait(bar(BAR_AFULL + ab), aph);
      }
      if (lane == 0) TL2(1, i);
      mbar_wait(bar(BAR_BFULL + st), bph);
      if (lane == 0) TL2(2, i);
      mbar_wait(bar(BAR_SEMPTY + sb), sph ^ 1);
      if (lane == 0) TL2(3, i);
      tc_fence_after();
      {
        const uint32_t a_hi = base + pl.a_off(ab), a_lo = a_hi + A_TILE;
        const uint32_t a_n = a_hi + (ZLO ? 2 : 1) * A_TILE;
        const uint32_t b_hi = base + pl.b_off(st) + OFF_BHI, b_lo = base + pl.b_off(st) + OFF_BLO;
        const uint32_t b_n = base + pl.b_off(st) + OFF_BN;
        const uint32_t d = tmem + TMEM_S0 + sb * CN;
        if (F32) {
          const uint32_t th = tmem + TMEM_A0 + ab * A_COLS, tl = th + 32, tn = th + (NSPLIT == 3 ? 64 : 32);
          mma_f16_ts(d, tn, desc_noswz(b_n, NORM_LBO, NORM_SBO), ID1, 0);
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_f16_ts(d, th + 8 * k, desc_sw128(b_hi + 32 * k), ID1, 1);
          if (NSPLIT == 3) {
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_f16_ts(d, th + 8 * k, desc_sw128(b_lo + 32 * k), ID1, 1);
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_f16_ts(d, tl + 8 * k, desc_sw128(b_hi + 32 * k), ID1, 1);
          }
        } else {
        mma_f16(d, desc_noswz(a_n, NORM_LBO, NORM_SBO), desc_noswz(b_n, NORM_LBO, NORM_SBO), ID1, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(a_hi + 32 * k), desc_sw128(b_hi + 32 * k), ID1, 1);
        if (NSPLIT == 3) {
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(a_hi + 32 * k), desc_sw128(b_lo + 32 * k), ID1, 1);
          if (ZLO) {
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(a_lo + 32 * k), desc_sw128(b_hi + 32 * k), ID1, 1);
          }
        }
        }
        mma_commit(bar(BAR_SFULL + sb));
        if (last_of_scale) mma_commit(bar(BAR_AEMPTY + ab));
        mma_commit(bar(BAR_BEMPTY + st));
        if (lane == 0) TL2(4, i);
      }
      __syncwarp();
      if (last_of_scale) {
        ++acount;
        if (++ab == NA) { ab = 0; aph ^= 1; }
      }
      if (++st == (uint32_t)NST) { st = 0; bph ^= 1; }
      if (++sb == NSETS) { sb = 0; sph ^= 1; }
      if (++c == NC) c = 0;
    }
  } else if (warp == WARP_MMA2) {
    // ================================ logits-GEMM issuer ================================
    if (has_w) {
      const uint32_t ID2 = idesc_f16(Fmt<X>::id, TILE_M, KP);
      const int U = my_tiles * NC;
      int c = 0, tl = 0;
      uint32_t pb = 0, pph = 0, sw = 0, wph = 0;
      for (int u = 0; u < U; ++u) {
        const uint32_t lb = tl & 1;
        if (lane == 0) TL2(8, u);
        if (c == 0) mbar_wait(bar(BAR_LEMPTY + lb), ((tl >> 1) & 1) ^ 1);
        mbar_wait(bar(BAR_WFULL + sw), wph);
        if (lane == 0) TL2(9, u);
        mbar_wait(bar(BAR_PFULL + pb), pph);
        if (lane == 0) TL2(10, u);
        tc_fence_after();
        {
          const uint32_t pa = base + pl.p_off(pb), w2 = base + pl.w_off(sw);
          const uint32_t d = tmem + TMEM_L0 + lb * KP;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            mma_f16(d, desc_sw128(pa + 32 * k), desc_sw128(w2 + 32 * k), ID2, (c > 0 || k > 0) ? 1 : 0);
          if (LSPLIT == 3) {
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(pa + 32 * k), desc_sw128(w2 + KP * 128 + 32 * k), ID2, 1);
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(pa + P_BYTES + 32 * k), desc_sw128(w2 + 32 * k), ID2, 1);
          }
          mma_commit(bar(BAR_PEMPTY + pb));
          mma_commit(bar(BAR_WEMPTY + sw));
          if (c == NC - 1) mma_commit(bar(BAR_LFULL + lb));
          if (lane == 0) TL2(11, u);
        }
        __syncwarp();
        if (++pb == NSETS) { pb = 0; pph ^= 1; }
        if (++sw == (uint32_t)NSTW) { sw = 0; wph ^= 1; }
        if (++c == NC) { c = 0; ++tl; }
      }
    }
  } else if (warp == WARP_PROD) {
    // ================================ prototype-chunk producer ================================
    // One bulk copy per unit (hi | norm [| lo] are contiguous in the packed image), running counters instead of
    // divisions, and the last-layer tiles have a producer warp of their own: one warp doing both with `i % NC`,
    // `i / NST` arithmetic took ~750 clk per copy and delivered every chunk just in time (MMA issuer waiting ~500 clk per
    // unit for it, profiles/r02_timeline_fwd_tma_*.txt).
    pdl_wait();  // the packed prototype images come from the pre-pass launched just before this kernel
    const int U = my_tiles * NC;
    const uint32_t bytes_b = (NSPLIT == 3 ? 2 : 1) * CN * 128 + CN * 32;
    const unsigned char* img = prm.img;
    const int img_bytes = ct.img_bytes;
    int c = 0, st = 0;
    uint32_t ph = 1;
    const unsigned char* src = img;
    for (int i = 0; i < U; ++i) {
      if (lane == 0) TL2(16, i);
      mbar_wait(bar(BAR_BEMPTY + st), ph);
      if (lane == 0) {
        TL2(17, i);
        mbar_arrive_expect_tx(bar(BAR_BFULL + st), bytes_b);
        bulk_g2s(base + pl.b_off(st), src, bytes_b, bar(BAR_BFULL + st));
      }
      __syncwarp();
      src += img_bytes;
      if (++c == NC) { c = 0; src = img; }
      if (++st == NST) { st = 0; ph ^= 1; }
    }
  } else if (warp == WARP_PRODW) {
    // ================================ last-layer tile producer ================================
    if (has_w) {
      pdl_wait();
      const int U = my_tiles * NC;
      const uint32_t bytes_w = (LSPLIT == 3 ? 2 : 1) * KP * 128;
      const unsigned char* img = prm.img + OFF_W2;
      const int img_bytes = ct.img_bytes;
      int c = 0, sw = 0;
      uint32_t ph = 1;
      const unsigned char* src = img;
      for (int i = 0; i < U; ++i) {
        if (lane == 0) TL2(18, i);
        mbar_wait(bar(BAR_WEMPTY + sw), ph);
        if (lane == 0) {
          TL2(19, i);
          mbar_arrive_expect_tx(bar(BAR_WFULL + sw), bytes_w);
          bulk_g2s(base + pl.w_off(sw), src, bytes_w, bar(BAR_WFULL + sw));
        }
        __syncwarp();
        src += img_bytes;
        if (++c == NC) { c = 0; src = img; }
        if (++sw == NSTW) { sw = 0; ph ^= 1; }
      }
    }
  } else if (warp >= WARP_Z && warp < WARP_Z + NZ) {
    // ================================ feature-tile producer + row norms ================================
    // Work list: (tile, scale) pairs in the order the distance-GEMM issuer consumes them.  The first warp keeps up to NA
    // tiles in flight (a stage is re-armed as soon as a NON-blocking probe sees it free); every thread of the four warps
    // then computes the squared norm of one row of the landed tile.  (One warp for all 128 rows took 2 200 clk per pair, and
    // a try_wait used as a probe suspended it ~3 000 clk per pair: together 5 200 clk per pair, the period of the whole
    // kernel - profiles/r02_timeline_fwd_tma_before.txt.)
    // (a 16-bit re-lay of an fp32 map was written by a kernel that completed before the prototype pre-pass started)
    auto advance = [&](int& ti, int& c) {
      const int sc = ct.scale[c];
      do { ++c; } while (c < NC && ct.scale[c] == sc);
      if (c == NC) { c = 0; ++ti; }
    };
    constexpr uint32_t TX = (ZLO ? 2 : 1) * A_TILE;
    const bool issuer = warp == WARP_Z;
    const int m = (warp - WARP_Z) * 32 + lane;
    if (F32) {
      // fp32 NCHW source.  Per (tile, scale) pair: four TMA boxes {132 px, 16 channels} (channel residues mod 4) into the
      // staging buffer; every thread then reads the 64 channels of ITS position (row m), releases the buffer (the next
      // pair's boxes start right away), converts to hi / lo 16-bit pairs and writes them with the squared norm into the
      // TMEM operand buffer of the pair: lane = position, column j = channels 2j, 2j+1.
      const int sh0 = prm.tm_f32.shift[0], sh1 = prm.tm_f32.shift[1], sh2 = prm.tm_f32.shift[2], sh3 = prm.tm_f32.shift[3];
      const uint32_t lane_base = ((uint32_t)((warp - WARP_Z) * 32)) << 16;
      int iti = 0, ic = 0;
      uint32_t issued = 0;
      auto issue = [&]() {
        if (iti >= my_tiles) return;
        mbar_wait(bar(BAR_TFULL + 2), (issued & 1) ^ 1);   // all four warps have read the previous pair out
        if (lane == 0) {
          const long long T = (long long)blockIdx.x + (long long)iti * gridDim.x;
          const int b = (int)(T / prm.tiles_per_map), pix0 = (int)(T - (long long)b * prm.tiles_per_map) * TILE_M;
          const int row0 = ct.scale[ic] * 16;
          TL2(27, issued);
          mbar_arrive_expect_tx(bar(BAR_TFULL), (uint32_t)STG_BYTES);
#pragma unroll
          for (int r = 0; r < 4; ++r)
            tma_load_3d(base + pl.a_off(0) + (uint32_t)(r * 16) * (STG_PX * 4), &prm.tm_f32.map[r], pix0, row0, b, bar(BAR_TFULL));
        }
        __syncwarp();
        ++issued;
        advance(iti, ic);
      };
      if (issuer) issue();
      int ti = 0, c = 0;
      uint32_t acount = 0;
      const float* stg = reinterpret_cast<const float*>(gbase + pl.a_off(0)) + m;
      while (ti < my_tiles) {
        const uint32_t ab = acount & 1, aph = (acount >> 1) & 1;
        if (issuer && lane == 0) TL2(24, acount);
        mbar_wait(bar(BAR_TFULL), acount & 1);
        if (issuer && lane == 0) TL2(25, acount);
        float v[64];
#pragma unroll
        for (int cc = 0; cc < 64; ++cc) {
          const int sh = (cc & 3) == 0 ? sh0 : (cc & 3) == 1 ? sh1 : (cc & 3) == 2 ? sh2 : sh3;
          v[cc] = stg[((cc & 3) * 16 + (cc >> 2)) * STG_PX + sh];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(BAR_TFULL + 2));
        if (issuer) issue();
        mbar_wait(bar(BAR_AEMPTY + ab), aph ^ 1);          // the MMAs that read this operand buffer two pairs ago are done
        tc_fence_after();
        const uint32_t ta = tmem + lane_base + TMEM_A0 + ab * A_COLS;
        float zn = 0.f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t hw_[16], lw_[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float v0 = v[h * 32 + 2 * j], v1 = v[h * 32 + 2 * j + 1];
            hw_[j] = Pack2<X>::pack(v0, v1);
            float2 e = Pack2<X>::unpack(hw_[j]);
            if (NSPLIT == 3) {
              lw_[j] = Pack2<X>::pack(v0 - e.x, v1 - e.y);
              const float2 el = Pack2<X>::unpack(lw_[j]);
              e.x += el.x;
              e.y += el.y;
            }
            zn = fmaf(e.x, e.x, zn);
            zn = fmaf(e.y, e.y, zn);
          }
          tmem_st16(ta + h * 16, hw_);
          if (NSPLIT == 3) tmem_st16(ta + 32 + h * 16, lw_);
        }
        {
          X nh, nm, nl;
          split3<X>(zn, nh, nm, nl);
          const uint32_t one = bits16(Fmt<X>::from(1.0f));
          const uint32_t nw[8] = {one | (one << 16), one | ((uint32_t)bits16(nh) << 16), (uint32_t)bits16(nm) | ((uint32_t)bits16(nl) << 16),
                                  0u, 0u, 0u, 0u, 0u};
          tmem_st8(ta + (NSPLIT == 3 ? 64 : 32), nw);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(BAR_AFULL + ab));
        if (issuer && lane == 0) TL2(26, acount);
        ++acount;
        advance(ti, c);
      }
    } else {
    int iti = 0, ic = 0;        // next pair to issue
    uint32_t issued = 0;
    auto issue = [&](bool blocking) -> bool {
      if (iti >= my_tiles) return true;
      const uint32_t ab = issued % NA, aph = (issued / NA) & 1;
      if (blocking) {
        mbar_wait(bar(BAR_AEMPTY + ab), aph ^ 1);
      } else {   // warp-uniform outcome of the probe
        const bool ok = mbar_test(bar(BAR_AEMPTY + ab), aph ^ 1);
        if (!__all_sync(0xffffffffu, ok)) return false;
      }
      if (lane == 0) {
        const int row0 = (int)(((long long)blockIdx.x + (long long)iti * gridDim.x) * TILE_M);
        const int c0 = ct.scale[ic] * s.D;
        TL2(27, issued);
        mbar_arrive_expect_tx(bar(BAR_TFULL + ab), TX);
        tma_load_2d(base + pl.a_off(ab), &prm.tm_hi, c0, row0, bar(BAR_TFULL + ab));
        if (ZLO) tma_load_2d(base + pl.a_off(ab) + A_TILE, &prm.tm_lo, c0, row0, bar(BAR_TFULL + ab));
      }
      __syncwarp();
      ++issued;
      advance(iti, ic);
      return true;
    };
    int ti = 0, c = 0;
    uint32_t acount = 0;
    while (ti < my_tiles) {
      const uint32_t ab = acount % NA, aph = (acount / NA) & 1;
      if (issuer) {
        if (issued <= acount) issue(true);                                   // the pair waited for next must be on its way
        while (issued < acount + NA && iti < my_tiles && issue(false)) {     // keep up to NA pairs in flight
        }
      }
      if (issuer && lane == 0) TL2(24, acount);
      mbar_wait(bar(BAR_TFULL + ab), aph);
      if (issuer && lane == 0) TL2(25, acount);
      unsigned char* a_hi = gbase + pl.a_off(ab);
      unsigned char* a_lo = a_hi + A_TILE;
      unsigned char* a_n = a_hi + (ZLO ? 2 : 1) * A_TILE;
      ZStage<X, NSPLIT>::store_norm(m, a_n, tile_row_norm<X, ZLO>(a_hi, a_lo, m));
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(BAR_AFULL + ab));
      if (issuer && lane == 0) TL2(26, acount);
      ++acount;
      advance(ti, c);
    }
    }
  } else {
    // ================================ epilogue ================================
    const int q = warp & 3, set = warp >> 2;
    const int m = q * 32 + lane;
    const uint32_t lane_base = ((uint32_t)(q * 32)) << 16;
    const float eps = s.eps, ome = 1.0f - s.eps;
    const bool is_log = s.act_kind == SPS_ACT_LOG;
    auto store_logits = [&](int tj) {
      long long nn0;
      const int lim = tile_rows(tj, nn0);
      const long long nn = nn0 + m;
      const uint32_t lb = tj & 1;
      mbar_wait(bar(BAR_LFULL + lb), (tj >> 1) & 1);
      tc_fence_after();
      for (int k0 = 0; k0 < KP; k0 += 16) {
        uint32_t lr[16];
        tmem_ld16(tmem + lane_base + TMEM_L0 + lb * KP + k0, lr);
        tmem_ld_wait();
        if (m < lim) {
#pragma unroll
          for (int k = 0; k < 16; ++k)
            if (k0 + k < s.K) prm.logits[nn * s.K + k0 + k] = __uint_as_float(lr[k]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(BAR_LEMPTY + lb));
    };
    const int defer_c = NC > 1 ? 1 : 0;
    int next_store = 0;   // set 0: next tile whose logits go to HBM
    const int U = my_tiles * NC;
    int ti = 0, c = set;
    while (c >= NC && NC > 0) { c -= NC; ++ti; }
    int cur_ti = -1, lim = 0;
    long long n0 = 0;
    for (int u = set; u < U; u += NSETS) {
      if (ti != cur_ti) {
        lim = tile_rows(ti, n0);
        cur_ti = ti;
      }
      const long long n = n0 + m;
      const bool valid = m < lim;
      const long long b = valid ? n / s.HW : 0, pix = valid ? n % s.HW : 0;
      float* drow = DENSE && prm.dist != nullptr ? prm.dist + b * (long long)s.P * s.HW + pix : nullptr;
      const uint32_t sb = set, pb = set, ph = (u / NSETS) & 1;
      if (q == 0 && lane == 0) TL2(32, u);
      mbar_wait(bar(BAR_SFULL + sb), ph);
      if (q == 0 && lane == 0) TL2(33, u);
      tc_fence_after();
      unsigned char* prow = gbase + pl.p_off(pb) + (uint32_t)m * 128u;
#pragma unroll 1
      for (int hf = 0; hf < 2; ++hf) {
        uint32_t r[32];
        tmem_ld32(tmem + lane_base + TMEM_S0 + sb * CN + hf * 32, r);
        tmem_ld_wait();
        if (q == 0 && lane == 0) TL2(34 + hf, u);
        if (hf == 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(BAR_SEMPTY + sb));
        }
        const int col0 = hf * 32, ncols = ct.n[c] - col0;
        const int p0 = ct.p0[c] + col0;
        float av[32];
        if (is_log) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float d = fmaxf(__uint_as_float(r[j]), 0.0f);
#ifdef SPS_DBG_NOEPI
            av[j] = d * ome;   // debug build only: no MUFU, to see what the pipeline around the epilogue can sustain
#else
            av[j] = lg2_approx(fmaf(ome, rcp_approx(d + eps), 1.0f));
#endif
            if (DENSE) r[j] = __float_as_uint(d);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float d = fmaxf(__uint_as_float(r[j]), 0.0f);
            av[j] = -d;
            if (DENSE) r[j] = __float_as_uint(d);
          }
        }
        if (DENSE) {
          if (drow != nullptr && valid) {
            float* dp = drow + (long long)p0 * s.HW;
            const size_t pitch = (size_t)s.HW;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (j < ncols) *dp = __uint_as_float(r[j]);
              dp += pitch;
            }
          }
          if (prm.act != nullptr) {
            const float sc = is_log ? 0.69314718055994531f : 1.0f;
            float* scr = reinterpret_cast<float*>(gbase + pl.scr0 + warp * ACT_SCRATCH);
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 32; ++j) scr[lane * 33 + j] = av[j] * sc;
            __syncwarp();
            const long long n_first = n - lane;
            float* abase = prm.act + n_first * (long long)s.P + p0 + lane;
            const int rows = lim - q * 32 < 32 ? lim - q * 32 : 32;
            if (lane < ncols) {
              if (rows == 32) {   // the common case without a test per row (it compiled to a branch + reconvergence point each)
                const size_t pitch = (size_t)s.P;
                float* ap = abase;
#pragma unroll
                for (int rr = 0; rr < 32; ++rr) {
                  *ap = scr[rr * 33 + lane];
                  ap += pitch;
                }
              } else {
                for (int rr = 0; rr < rows; ++rr) abase[(long long)rr * s.P] = scr[rr * 33 + lane];
              }
            }
          }
        }
        if (has_w) {
          if (hf == 0) mbar_wait(bar(BAR_PEMPTY + pb), ph ^ 1);
          if (q == 0 && lane == 0 && hf == 0) TL2(36, u);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint32_t w[4], wl[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float x0 = av[g * 8 + 2 * e], x1 = av[g * 8 + 2 * e + 1];
              w[e] = Pack2<X>::pack(x0, x1);
              if (LSPLIT == 3) {
                const float2 h = Pack2<X>::unpack(w[e]);
                wl[e] = Pack2<X>::pack(x0 - h.x, x1 - h.y);
              }
            }
            const uint32_t poff = (((uint32_t)(hf * 4 + g)) ^ ((uint32_t)m & 7u)) << 4;
            *reinterpret_cast<uint4*>(prow + poff) = make_uint4(w[0], w[1], w[2], w[3]);
            if (LSPLIT == 3) *reinterpret_cast<uint4*>(prow + P_BYTES + poff) = make_uint4(wl[0], wl[1], wl[2], wl[3]);
          }
        }
      }
      if (has_w) {
        if (q == 0 && lane == 0) TL2(37, u);
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(BAR_PFULL + pb));
        if (q == 0 && lane == 0) TL2(38, u);
      }
      if (has_w && set == 0) {
        while (next_store < ti - 1) store_logits(next_store++);
        if (next_store == ti - 1 && c >= defer_c) store_logits(next_store++);
      }
      c += NSETS;
      while (c >= NC) { c -= NC; ++ti; }
    }
    if (has_w && set == 0)
      while (next_store < my_tiles) store_logits(next_store++);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == WARP_MMA) {
    tc_fence_after();
    tmem_dealloc(tmem, tmem_cols);
  }
}

// ---------------------------------------------------------------------------------------------
// fp32 NCHW -> 16-bit channels-last re-lay (hi [+ lo]) for the TMA-fed kernels (SPS_FLAG_ZPACK).
// One block = 32 consecutive positions x all C channels: reads are 128-byte warp loads along the position axis, the
// packed (channel pair) words go through a shared-memory tile with an odd pitch, writes are 128-byte warp stores
// along the channel axis.
// ---------------------------------------------------------------------------------------------
template <typename X, bool ZLO>
__global__ void __launch_bounds__(256) z_pack_kernel(const float* __restrict__ z, int C, int HW, long long N,
                                                     uint32_t* __restrict__ out_hi, uint32_t* __restrict__ out_lo) {
  extern __shared__ uint32_t pk_smem[];
  const int C2 = C >> 1, pitch = C2 + 1;
  uint32_t* t_hi = pk_smem;
  uint32_t* t_lo = pk_smem + 32 * pitch;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long n0 = (long long)blockIdx.x * 32, n = n0 + lane;
  const bool live = n < N;
  const long long nn = live ? n : N - 1;
  const long long b = nn / HW, pix = nn % HW;
  const float* src = z + (b * C) * (long long)HW + pix;
  for (int c2 = warp; c2 < C2; c2 += 8 * 4) {
    float v0[4], v1[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int cc = c2 + 8 * u;
      v0[u] = cc < C2 ? __ldg(src + (long long)(2 * cc) * HW) : 0.f;
      v1[u] = cc < C2 ? __ldg(src + (long long)(2 * cc + 1) * HW) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int cc = c2 + 8 * u;
      if (cc < C2) {
        const uint32_t h = Pack2<X>::pack(v0[u], v1[u]);
        t_hi[lane * pitch + cc] = h;
        if (ZLO) {
          const float2 e = Pack2<X>::unpack(h);
          t_lo[lane * pitch + cc] = Pack2<X>::pack(v0[u] - e.x, v1[u] - e.y);
        }
      }
    }
  }
  __syncthreads();
  const int rows = N - n0 < 32 ? (int)(N - n0) : 32;
  for (int r = warp; r < rows; r += 8) {
    for (int w = lane; w < C2; w += 32) {
      out_hi[(n0 + r) * C2 + w] = t_hi[r * pitch + w];
      if (ZLO) out_lo[(n0 + r) * C2 + w] = t_lo[r * pitch + w];
    }
  }
}

}  // namespace tf

#ifdef SPS_TIMELINE
extern "C" int sps_debug_timeline2(long long* host, int count) {
  return (int)cudaMemcpyFromSymbol(host, tf::g_tl2, sizeof(long long) * (size_t)count);
}
#endif

// ---------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------
int build_chunk_table(const HeadShape& s, bool with_last_layer, tc::ChunkTable* ct);
int launch_proto_prep(const HeadShape& s, int math, const tc::ChunkTable& ct, const float* protos, const float* last_w,
                      unsigned char* img, cudaStream_t st);

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int encode_tiled_fn(EncodeTiledFn* out) {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    SPS_CUDA_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
    SPS_REQUIRE(p != nullptr && qres == cudaDriverEntryPointSuccess, SPS_ERR_CUDA, "cuTensorMapEncodeTiled not available in this driver");
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  *out = fn;
  return SPS_OK;
}

int make_tmap_nchw_f32(TmapNchwF32* out, const float* z, long long B, long long C, long long HW, int px, int box_rows) {
  const int box_px = px + 4;
  EncodeTiledFn fn = nullptr;
  const int rc = encode_tiled_fn(&fn);
  if (rc != SPS_OK) return rc;
  SPS_REQUIRE(C % 4 == 0 && (reinterpret_cast<uintptr_t>(z) & 3) == 0 && (box_px * 4) % 16 == 0, SPS_ERR_UNSUPPORTED,
              "fp32 NCHW tensor maps: C=%lld must be a multiple of 4 and the map 4-byte aligned", C);
  for (int r = 0; r < 4; ++r) {
    const uintptr_t row = reinterpret_cast<uintptr_t>(z) + (uintptr_t)r * (uintptr_t)HW * 4u, base = row & ~(uintptr_t)15;
    out->shift[r] = (int)((row - base) / 4);
    const cuuint64_t gdim[3] = {(cuuint64_t)(HW + out->shift[r]), (cuuint64_t)(C / 4), (cuuint64_t)(B > 0 ? B : 1)};
    const cuuint64_t gstride[2] = {(cuuint64_t)HW * 16u, (cuuint64_t)C * (cuuint64_t)HW * 4u};
    const cuuint32_t box[3] = {(cuuint32_t)box_px, (cuuint32_t)box_rows, 1u};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult res = fn(&out->map[r], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, reinterpret_cast<void*>(base), gdim, gstride, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    SPS_REQUIRE(res == CUDA_SUCCESS, SPS_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for an fp32 NCHW map [%lld x %lld x %lld], residue %d",
                (int)res, B, C, HW, r);
  }
  return SPS_OK;
}

int make_tmap_2d_16bit(CUtensorMap* out, const void* base, bool bf16, long long rows, long long cols, int box_rows, int box_cols,
                       int swizzle_bytes) {
  EncodeTiledFn fn = nullptr;
  const int rc0 = encode_tiled_fn(&fn);
  if (rc0 != SPS_OK) return rc0;
  SPS_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, SPS_ERR_BAD_ARG, "16-bit channels-last map: base address must be 16-byte aligned");
  SPS_REQUIRE((cols * 2) % 16 == 0, SPS_ERR_UNSUPPORTED, "16-bit channels-last map: %lld channels give a row pitch that is not a multiple of 16 bytes", cols);
  const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)(rows > 0 ? rows : 1)};
  const cuuint64_t gstride[1] = {(cuuint64_t)cols * 2};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                      : CU_TENSOR_MAP_SWIZZLE_NONE;
  const CUresult r = fn(out, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim,
                        gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SPS_REQUIRE(r == CUDA_SUCCESS, SPS_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for a [%lld x %lld] 16-bit map", (int)r, rows, cols);
  return SPS_OK;
}

static bool math_is_split(int math) { return math == SPS_MATH_F16X3 || math == SPS_MATH_BF16X3; }

// bytes of the 16-bit re-lay of an fp32 map (hi [+ lo]), each part 1 KB aligned
size_t z_pack_bytes(const HeadShape& s, int math) {
  const size_t part = ((size_t)s.N * s.C * 2 + 1023) / 1024 * 1024;
  return (math_is_split(math) ? 2 : 1) * part + 1024;
}

void z_pack_parts(const HeadShape& s, int math, const void* area, const void** hi, const void** lo) {
  const unsigned char* base = reinterpret_cast<const unsigned char*>(((uintptr_t)area + 1023) & ~(uintptr_t)1023);
  const size_t part = ((size_t)s.N * s.C * 2 + 1023) / 1024 * 1024;
  *hi = base;
  *lo = math_is_split(math) ? base + part : nullptr;
}

int launch_z_pack(const HeadShape& s, int math, const float* z, void* out, cudaStream_t st) {
  SPS_REQUIRE(s.C % 2 == 0 && s.C <= 1024, SPS_ERR_UNSUPPORTED, "z re-lay: C=%d must be even and <= 1024", s.C);
  if (s.N == 0) return SPS_OK;
  unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)out + 1023) & ~(uintptr_t)1023);
  const size_t part = ((size_t)s.N * s.C * 2 + 1023) / 1024 * 1024;
  uint32_t* hi = reinterpret_cast<uint32_t*>(base);
  uint32_t* lo = reinterpret_cast<uint32_t*>(base + part);
  const unsigned grid = (unsigned)((s.N + 31) / 32);
  const size_t smem = (size_t)32 * (s.C / 2 + 1) * 4 * (math_is_split(math) ? 2 : 1);
  switch (math) {
    case SPS_MATH_F16:
      SPS_CUDA_TRY(cudaFuncSetAttribute(tf::z_pack_kernel<__half, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      tf::z_pack_kernel<__half, false><<<grid, 256, smem, st>>>(z, s.C, s.HW, s.N, hi, lo);
      break;
    case SPS_MATH_F16X3:
      SPS_CUDA_TRY(cudaFuncSetAttribute(tf::z_pack_kernel<__half, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      tf::z_pack_kernel<__half, true><<<grid, 256, smem, st>>>(z, s.C, s.HW, s.N, hi, lo);
      break;
    case SPS_MATH_BF16:
      SPS_CUDA_TRY(cudaFuncSetAttribute(tf::z_pack_kernel<__nv_bfloat16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      tf::z_pack_kernel<__nv_bfloat16, false><<<grid, 256, smem, st>>>(z, s.C, s.HW, s.N, hi, lo);
      break;
    case SPS_MATH_BF16X3:
      SPS_CUDA_TRY(cudaFuncSetAttribute(tf::z_pack_kernel<__nv_bfloat16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      tf::z_pack_kernel<__nv_bfloat16, true><<<grid, 256, smem, st>>>(z, s.C, s.HW, s.N, hi, lo);
      break;
    default: SPS_REQUIRE(false, SPS_ERR_BAD_ARG, "z re-lay: math=%d is not a tensor-core mode", math);
  }
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

// three epilogue sets when the outputs are only the logits and the third distance accumulator / activation tile fit
template <int NSPLIT, bool ZLO, int LSPLIT> static bool three_sets_fit(int kp, bool dense) {
  if (dense || 3 * tc::CN + 2 * kp > 512) return false;
  const tf::Plan pl = tf::make_plan<NSPLIT, ZLO, 3, LSPLIT>(kp, 0);
  return pl.nst >= 2 && pl.total <= 227 * 1024;
}
template <int NSPLIT, bool ZLO, int LSPLIT> static bool two_sets_fit(int kp, bool dense_act) {
  const tf::Plan pl = tf::make_plan<NSPLIT, ZLO, 2, LSPLIT>(kp, dense_act ? 8 : 0);
  return pl.nst >= 2 && pl.total <= 227 * 1024 && 2 * tc::CN + 2 * kp <= 512;
}

template <typename X, int NSPLIT, bool ZLO, int LSPLIT>
static int launch_fwd_tma_l(const HeadShape& s, const tc::ChunkTable& ct, tf::FwdTmaParams& prm, bool dense, bool dense_act, cudaStream_t st) {
  const int grid = prm.n_tiles < sm_count() ? prm.n_tiles : sm_count();
  if (three_sets_fit<NSPLIT, ZLO, LSPLIT>(ct.kp, dense)) {
    const tf::Plan pl = tf::make_plan<NSPLIT, ZLO, 3, LSPLIT>(ct.kp, 0);
    SPS_CUDA_TRY(cudaFuncSetAttribute(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, false, 3, LSPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total));
    SPS_CUDA_TRY(launch_pdl(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, false, 3, LSPLIT>, dim3(grid), dim3(tf::n_threads(3)), (size_t)pl.total, st, prm));
  } else {
    const tf::Plan pl = tf::make_plan<NSPLIT, ZLO, 2, LSPLIT>(ct.kp, dense_act ? 8 : 0);
    SPS_REQUIRE(pl.nst >= 2 && pl.total <= 227 * 1024, SPS_ERR_UNSUPPORTED,
                "TMA-fed head: %d bytes of shared memory needed (K=%d)", pl.total, s.K);
    if (dense) {
      SPS_CUDA_TRY(cudaFuncSetAttribute(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, true, 2, LSPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total));
      SPS_CUDA_TRY(launch_pdl(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, true, 2, LSPLIT>, dim3(grid), dim3(tf::n_threads(2)), (size_t)pl.total, st, prm));
    } else {
      SPS_CUDA_TRY(cudaFuncSetAttribute(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, false, 2, LSPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total));
      SPS_CUDA_TRY(launch_pdl(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, false, 2, LSPLIT>, dim3(grid), dim3(tf::n_threads(2)), (size_t)pl.total, st, prm));
    }
  }
  count_launch();
  return SPS_OK;
}

template <typename X, int NSPLIT, bool ZLO>
static int launch_fwd_tma_t(const HeadShape& s, const tc::ChunkTable& ct, const void* z_hi, const void* z_lo, const float* protos,
                            const float* last_w, float* logits, float* dist, float* act, unsigned char* img, int math,
                            cudaStream_t st) {
  int rc = launch_proto_prep(s, math, ct, protos, last_w, img, st);
  if (rc != SPS_OK) return rc;
  tf::FwdTmaParams prm;
  const bool bf = std::is_same<X, __nv_bfloat16>::value;
  rc = make_tmap_2d_16bit(&prm.tm_hi, z_hi, bf, s.N, s.C, tc::TILE_M, 64, 128);
  if (rc != SPS_OK) return rc;
  rc = make_tmap_2d_16bit(&prm.tm_lo, ZLO ? z_lo : z_hi, bf, s.N, s.C, tc::TILE_M, 64, 128);
  if (rc != SPS_OK) return rc;
  prm.s = s;
  prm.ct = ct;
  prm.img = img;
  prm.logits = logits;
  prm.dist = dist;
  prm.act = act;
  prm.n_tiles = (int)((s.N + tc::TILE_M - 1) / tc::TILE_M);
  const bool dense = dist != nullptr || act != nullptr;
  // a wide last layer whose hi + lo tiles do not fit keeps the 3-term distances and runs the logits GEMM on one term
  if (NSPLIT == 3 && !two_sets_fit<NSPLIT, ZLO, NSPLIT>(ct.kp, act != nullptr))
    return launch_fwd_tma_l<X, NSPLIT, ZLO, 1>(s, ct, prm, dense, act != nullptr, st);
  return launch_fwd_tma_l<X, NSPLIT, ZLO, NSPLIT>(s, ct, prm, dense, act != nullptr, st);
}

// ---- fp32 NCHW maps: TMA staging + operand in TMEM (F32 mode of the kernel) ----
template <int NSPLIT> static bool f32_sets_fit(int nsets, int kp, bool dense_act) {
  if (nsets * tc::CN + 2 * kp + 2 * tf::a_tmem_cols<NSPLIT>() > 512) return false;
  const tf::Plan pl = nsets == 3 ? tf::make_plan<NSPLIT, NSPLIT == 3, 3, NSPLIT, true>(kp, 0)
                                 : tf::make_plan<NSPLIT, NSPLIT == 3, 2, NSPLIT, true>(kp, dense_act ? 8 : 0);
  return pl.nst >= 2 && pl.total <= 227 * 1024;
}

bool head_fwd_f32tma_supported(const HeadShape& s, int math, bool with_last_layer, bool dense_act) {
  if (math == SPS_MATH_FP32 || s.D != 64 || s.C % 4 != 0 || s.K > 192 || s.N >= (1ll << 31) || s.HW >= (1 << 30)) return false;
  int nc = 0;
  for (int sc = 0; sc < s.S; ++sc) nc += (s.scale_lo[sc + 1] - s.scale_lo[sc] + tc::CN - 1) / tc::CN;
  if (nc > tc::MAX_CHUNKS) return false;
  const int kp = with_last_layer ? (s.K + 15) / 16 * 16 : 0;
  return math_is_split(math) ? f32_sets_fit<3>(2, kp, dense_act) : f32_sets_fit<1>(2, kp, dense_act);
}

template <typename X, int NSPLIT>
static int launch_fwd_f32_t(const HeadShape& s, const tc::ChunkTable& ct, const float* z, const float* protos, const float* last_w,
                            float* logits, float* dist, float* act, unsigned char* img, int math, cudaStream_t st) {
  constexpr bool ZLO = NSPLIT == 3;
  int rc = launch_proto_prep(s, math, ct, protos, last_w, img, st);
  if (rc != SPS_OK) return rc;
  tf::FwdTmaParams prm;
  memset(&prm.tm_hi, 0, sizeof(prm.tm_hi));
  memset(&prm.tm_lo, 0, sizeof(prm.tm_lo));
  rc = make_tmap_nchw_f32(&prm.tm_f32, z, s.B, s.C, s.HW, tc::TILE_M, 16);
  if (rc != SPS_OK) return rc;
  prm.s = s;
  prm.ct = ct;
  prm.img = img;
  prm.logits = logits;
  prm.dist = dist;
  prm.act = act;
  prm.tiles_per_map = (s.HW + tc::TILE_M - 1) / tc::TILE_M;
  prm.n_tiles = s.B * prm.tiles_per_map;
  const bool dense = dist != nullptr || act != nullptr;
  const int grid = prm.n_tiles < sm_count() ? prm.n_tiles : sm_count();
  if (!dense && f32_sets_fit<NSPLIT>(3, ct.kp, false)) {
    const tf::Plan pl = tf::make_plan<NSPLIT, ZLO, 3, NSPLIT, true>(ct.kp, 0);
    SPS_CUDA_TRY(cudaFuncSetAttribute(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, false, 3, NSPLIT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total));
    SPS_CUDA_TRY(launch_pdl(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, false, 3, NSPLIT, true>, dim3(grid), dim3(tf::n_threads(3)), (size_t)pl.total, st, prm));
  } else {
    const tf::Plan pl = tf::make_plan<NSPLIT, ZLO, 2, NSPLIT, true>(ct.kp, act != nullptr ? 8 : 0);
    SPS_REQUIRE(pl.nst >= 2 && pl.total <= 227 * 1024, SPS_ERR_UNSUPPORTED, "fp32 TMA-fed head: %d bytes of shared memory needed (K=%d)", pl.total, s.K);
    if (dense) {
      SPS_CUDA_TRY(cudaFuncSetAttribute(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, true, 2, NSPLIT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total));
      SPS_CUDA_TRY(launch_pdl(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, true, 2, NSPLIT, true>, dim3(grid), dim3(tf::n_threads(2)), (size_t)pl.total, st, prm));
    } else {
      SPS_CUDA_TRY(cudaFuncSetAttribute(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, false, 2, NSPLIT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total));
      SPS_CUDA_TRY(launch_pdl(tf::head_fwd_tma_kernel<X, NSPLIT, ZLO, false, 2, NSPLIT, true>, dim3(grid), dim3(tf::n_threads(2)), (size_t)pl.total, st, prm));
    }
  }
  count_launch();
  return SPS_OK;
}

// z: fp32 NCHW.  img: workspace for the packed prototype images (head_fwd_tc_workspace bytes).
int launch_head_fwd_f32tma(const HeadShape& s, int math, const float* z, const float* protos, const float* last_w, float* logits,
                           float* dist, float* act, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (s.N == 0 || s.P == 0) return SPS_OK;
  tc::ChunkTable ct;
  int rc = build_chunk_table(s, last_w != nullptr, &ct);
  if (rc != SPS_OK) return rc;
  const size_t need = (size_t)ct.n_chunks * ct.img_bytes + 1024;
  SPS_REQUIRE(ws != nullptr && ws_bytes >= need, SPS_ERR_WORKSPACE, "head_fwd workspace: need %zu bytes, got %zu", need, ws_bytes);
  unsigned char* img = reinterpret_cast<unsigned char*>(((uintptr_t)ws + 1023) & ~(uintptr_t)1023);
  switch (math) {
    case SPS_MATH_F16: return launch_fwd_f32_t<__half, 1>(s, ct, z, protos, last_w, logits, dist, act, img, math, st);
    case SPS_MATH_F16X3: return launch_fwd_f32_t<__half, 3>(s, ct, z, protos, last_w, logits, dist, act, img, math, st);
    case SPS_MATH_BF16: return launch_fwd_f32_t<__nv_bfloat16, 1>(s, ct, z, protos, last_w, logits, dist, act, img, math, st);
    case SPS_MATH_BF16X3: return launch_fwd_f32_t<__nv_bfloat16, 3>(s, ct, z, protos, last_w, logits, dist, act, img, math, st);
  }
  SPS_REQUIRE(false, SPS_ERR_BAD_ARG, "math=%d is not a tensor-core mode", math);
}

// Shapes the TMA-fed forward covers.  zlo: the map has a lo part (fp32 source re-laid by z_pack_kernel in a split mode).
bool head_fwd_tma_supported(const HeadShape& s, int math, bool with_last_layer, bool dense_act, bool zlo) {
  if (math == SPS_MATH_FP32 || s.D != 64 || s.K > 192 || s.N >= (1ll << 31)) return false;
  int nc = 0;
  for (int sc = 0; sc < s.S; ++sc) nc += (s.scale_lo[sc + 1] - s.scale_lo[sc] + tc::CN - 1) / tc::CN;
  if (nc > tc::MAX_CHUNKS) return false;
  const int kp = with_last_layer ? (s.K + 15) / 16 * 16 : 0;
  if (!math_is_split(math)) return two_sets_fit<1, false, 1>(kp, dense_act);
  if (zlo) return two_sets_fit<3, true, 3>(kp, dense_act) || two_sets_fit<3, true, 1>(kp, dense_act);
  return two_sets_fit<3, false, 3>(kp, dense_act) || two_sets_fit<3, false, 1>(kp, dense_act);
}

// z_hi / z_lo: 16-bit [N][C] arrays (z_lo may be null: the map is exactly 16 bit).  img: 1 KB aligned workspace for the
// packed prototype images (head_fwd_tc_workspace bytes).
int launch_head_fwd_tma(const HeadShape& s, int math, const void* z_hi, const void* z_lo, const float* protos, const float* last_w,
                        float* logits, float* dist, float* act, void* ws, size_t ws_bytes, cudaStream_t st) {
  SPS_REQUIRE(s.D == 64, SPS_ERR_UNSUPPORTED, "TMA-fed head: D=%d, only 64 channels per scale are supported", s.D);
  if (s.N == 0 || s.P == 0) return SPS_OK;
  tc::ChunkTable ct;
  int rc = build_chunk_table(s, last_w != nullptr, &ct);
  if (rc != SPS_OK) return rc;
  const size_t need = (size_t)ct.n_chunks * ct.img_bytes + 1024;
  SPS_REQUIRE(ws != nullptr && ws_bytes >= need, SPS_ERR_WORKSPACE, "head_fwd workspace: need %zu bytes, got %zu", need, ws_bytes);
  unsigned char* img = reinterpret_cast<unsigned char*>(((uintptr_t)ws + 1023) & ~(uintptr_t)1023);
  const bool zlo = z_lo != nullptr;
  switch (math) {
    case SPS_MATH_F16: return launch_fwd_tma_t<__half, 1, false>(s, ct, z_hi, z_lo, protos, last_w, logits, dist, act, img, math, st);
    case SPS_MATH_BF16: return launch_fwd_tma_t<__nv_bfloat16, 1, false>(s, ct, z_hi, z_lo, protos, last_w, logits, dist, act, img, math, st);
    case SPS_MATH_F16X3:
      return zlo ? launch_fwd_tma_t<__half, 3, true>(s, ct, z_hi, z_lo, protos, last_w, logits, dist, act, img, math, st)
                 : launch_fwd_tma_t<__half, 3, false>(s, ct, z_hi, z_lo, protos, last_w, logits, dist, act, img, math, st);
    case SPS_MATH_BF16X3:
      return zlo ? launch_fwd_tma_t<__nv_bfloat16, 3, true>(s, ct, z_hi, z_lo, protos, last_w, logits, dist, act, img, math, st)
                 : launch_fwd_tma_t<__nv_bfloat16, 3, false>(s, ct, z_hi, z_lo, protos, last_w, logits, dist, act, img, math, st);
  }
  SPS_REQUIRE(false, SPS_ERR_BAD_ARG, "math=%d is not a tensor-core mode", math);
}

}  // namespace sps
