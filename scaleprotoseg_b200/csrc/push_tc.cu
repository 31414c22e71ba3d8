// Tensor-core push scan (sm_100a): per map and per prototype, the nearest position among those labelled with the
// prototype's class.  Replaces compute_distances lines 68-91 of segmentation/push_multiscale_optimization.py
// (one-hot, void column dropped, 1e10 mask, min / argmin over the flattened map).
//
// The distance GEMM is the forward's, TRANSPOSED: prototypes are the M rows (A operand, resident for the whole kernel: -2p
// in TMEM, the norm / class-mask K step in shared memory), positions are the N columns (B operand, streamed), so every
// epilogue thread owns one prototype and keeps its running (distance, index) minimum in registers - no cross-thread
// reduction.
//   acc[p, n] = (-2 p_p).z_s(n) + |p_p|^2 + |z_s(n)|^2       (norms in the extra K step, as in head_tc.cu)
//   key[p]    = min over n with label[n] == class(p)+1 of (bits(max(acc,0)) << 32 | n)   (first occurrence wins)
// One CTA owns ONE scale and up to 256 of its prototypes (two M=128 accumulator tiles) and a contiguous range of
// position tiles (128 positions of one map), so each feature value is read from HBM exactly once per 256
// prototypes of its scale (once in total for the shipped configs).  Feature tiles arrive by TMA in both map formats
// (fp32 NCHW: four residue tensor maps into a staging buffer, converted by the loader warps; 16-bit channels-last:
// straight into the operand layout).  In the 3-term modes the tensor pipe is what bounds the scan.
// Results are merged into keys[map][p] with one 64-bit atomicMin per (CTA, map, prototype).
#include "tma.cuh"

namespace sps {
namespace tcp {
using namespace sps::tc;

#ifdef SPS_PUSH_TIMELINE
// Debug build only: clock stamps of CTA 0, row = event kind, column = tile index (profiles/tools/timeline_push.py)
__device__ long long g_tlp[16 * 1024];
#define TLP(row, idx)                                                                                  \
  do {                                                                                                 \
    if (blockIdx.x == 0 && (idx) >= 0 && (idx) < 1024) g_tlp[(row) * 1024 + (idx)] = clock64();          \
  } while (0)
#else
#define TLP(row, idx) do { } while (0)
#endif

constexpr int PGROUP_CHUNKS = 4;   // 256 prototype rows per CTA
constexpr int MAX_PGROUPS = 64;
constexpr int NTHREADS = 576;      // 8 epilogue warps, 8 loader warps, 1 MMA warp, 1 resident-operand warp
constexpr int WARP_LOAD0 = 8, WARP_MMA = 16, WARP_RES = 17;

struct PGroupTable {
  int n_groups;
  int c0[MAX_PGROUPS];
  short nch[MAX_PGROUPS];
  short scale[MAX_PGROUPS];
  short cta0[MAX_PGROUPS];
  short ncta[MAX_PGROUPS];
};

struct PushParams {
  CUtensorMap tm_z;                // ZM_TMA16: the 16-bit channels-last maps as one [B*HW][C] array
  TmapNchwF32 tm_f32;              // ZM_LDG (fp32 NCHW maps): four tensor maps, one per channel residue mod 4 (tma.cuh)
  HeadShape s;
  ChunkTable ct;
  PGroupTable gt;
  const float* z;
  const unsigned char* img;        // forward chunk images (Bhi | Blo | Bn), kp = 0
  const int16_t* labels;           // [B, HW]
  const int32_t* proto_class;      // [P]
  unsigned long long* keys;        // [B, P], pre-filled with the "absent" key
  int tiles_per_map;
  long long n_tiles;               // B * tiles_per_map
};

// ZLO: the streamed feature tile has a lo part (fp32 source in a split mode; a 16-bit map is exact in one term)
template <int NSPLIT, bool ZLO, bool STAGED> struct PLay {
  static constexpr int RES_N = TILE_M * 32;                                               // norm K step of one 128-row prototype tile
  static constexpr int STAGE = (ZLO ? 2 : 1) * TILE_M * 128 + TILE_M * 32;                // z hi [lo] | norm
  static constexpr int SOFF_LO = TILE_M * 128, SOFF_N = (ZLO ? 2 : 1) * TILE_M * 128;
  static constexpr int SM_RES = 0;
  static constexpr int SM_Z = 2 * RES_N;
  // fp32 NCHW source (STAGED): a tile arrives as four TMA boxes {128 + 4 positions, D/4 channels} (channels r, r+4, ... of the
  // scale, one box per residue r, tma.cuh) in one of NSTG staging buffers; the loader warps convert from there.  Nothing
  // goes through L1 or registers on the way in, so the bytes in flight are bounded by the staging buffers (4 x 32 KB), not
  // by the load queue of the loader warps (LDG-fed, the scan stayed at 2.6 TB/s).  Loader group g owns buffers 2g and 2g+1
  // (tile i: buffer (i&1)*2 + ((i>>1)&1), use number i>>2), so every barrier has ONE waiting group going through its
  // phases in order (boxes of different tiles may land out of order: with a shared ring a parity wait could take phase k
  // for phase k+2)
  static constexpr int STG_PX = TILE_M + 4, STG_BYTES = 64 * STG_PX * 4, NSTG = STAGED ? 4 : 0;   // box rows of 132 floats (tma.cuh)
  static constexpr int SM_FREE = 227 * 1024 - 2048 - 4096 - SM_Z - NSTG * STG_BYTES;
  static constexpr int NSTAGE = SM_FREE / STAGE >= 6 ? 6 : SM_FREE / STAGE;
  // labels of the tile in flight: a ring of LAB_RING >= NSTAGE + 2 slots needs no barrier of its own (the loader of
  // tile i has waited for the MMA of tile i-NSTAGE, which waited for the epilogue of tile i-NSTAGE-1)
  static constexpr int LAB_RING = 8;
  static constexpr int SM_STG = SM_Z + NSTAGE * STAGE;
  static constexpr int SM_LAB = SM_STG + NSTG * STG_BYTES;
  static constexpr int SM_BAR = SM_LAB + LAB_RING * TILE_M * 4;
  static constexpr int SM_TOTAL = SM_BAR + 512 + 1024;
};
// TMEM columns: a ring of three 128-column accumulators (unit u = tile * n_mt + prototype tile, slot u % 3) and the
// resident prototype operand (-2p, hi | lo part: 32 columns each = 64 channels as 16-bit pairs) of prototype tile t at
// TM_A + t*64.  With the A operand in TMEM an M128 x N128 x K16 MMA reads 4 KB of shared memory instead of 8 (measured,
// profiles/tools/mma_ts_probe.cu: same 65 clk), which leaves half of the shared-memory bandwidth to the feature loaders.
// Each slot is used by the two epilogue sets in turn; no wait can alias because a set reaches unit u only after it has
// seen the commit of unit u-2, and commits are ordered (so unit u-3, the slot's previous user, has committed too).
constexpr int D_SLOTS = 3;
constexpr uint32_t TM_A = 384;
enum { PB_ZFULL = 0, PB_ZEMPTY = 8, PB_DFULL = 16, PB_DEMPTY = 20, PB_RES = 24, PB_RES2 = 25, PB_ATMEM = 26, PB_TFULL = 27, PB_SFULL = 36, PB_SEMPTY = 40, PB_COUNT = 44 };

__device__ __forceinline__ unsigned long long absent_key64() { return ((unsigned long long)0x501502F9u) << 32; }  // 1e10f

// MASKED: the class mask rides in the norm K step of the GEMM (labels <= MASK_MAX_LABEL, see ZStage::store_norm_mask)
// and the epilogue is a plain running minimum; otherwise the epilogue compares labels element by element.
template <typename X, int NSPLIT, bool MASKED, int ZMODE>
__global__ void __launch_bounds__(NTHREADS, 1) push_scan_tc_kernel(const __grid_constant__ PushParams prm) {
  constexpr bool ZLO = NSPLIT == 3 && ZMODE == ZM_LDG;
  using L = PLay<NSPLIT, ZLO, ZMODE == ZM_LDG>;
  extern __shared__ unsigned char smem_raw[];
  const HeadShape& s = prm.s;
  const ChunkTable& ct = prm.ct;
  const PGroupTable& gt = prm.gt;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* gbase = smem_raw + (base - smem_u32(smem_raw));
  auto bar = [&](int i) { return base + L::SM_BAR + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + L::SM_BAR + 8 * PB_COUNT);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

#ifdef SPS_PUSH_TIMELINE
  long long cta_t0;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(cta_t0));
#endif
  int g = 0;
  while (g + 1 < gt.n_groups && (int)blockIdx.x >= gt.cta0[g + 1]) ++g;
  const int cta_local = (int)blockIdx.x - gt.cta0[g], ncta = gt.ncta[g];
  const int nch = gt.nch[g], c0 = gt.c0[g], sc = gt.scale[g];
  const int n_mt = (nch + 1) / 2;                       // M tiles (128 rows) in use: 1 or 2
  const long long t_lo = prm.n_tiles * cta_local / ncta, t_hi = prm.n_tiles * (cta_local + 1) / ncta;
  const int my_tiles = (int)(t_hi - t_lo);

  // zero the resident operand area (rows of missing chunks must read as zeros), then barrier init
  for (int i = threadIdx.x; i < 2 * L::RES_N / 16; i += NTHREADS) reinterpret_cast<uint4*>(gbase + L::SM_RES)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) {
    for (int i = 0; i < L::NSTAGE; ++i) {
      mbar_init(bar(PB_ZFULL + i), 4);     // the 4 warps of one loader group
      mbar_init(bar(PB_ZEMPTY + i), 1);    // tcgen05.commit
      mbar_init(bar(PB_TFULL + i), 1);     // TMA mode: arrive.expect_tx + the tile bytes
    }
    for (int i = 0; i < D_SLOTS; ++i) {
      mbar_init(bar(PB_DFULL + i), 1);     // tcgen05.commit
      mbar_init(bar(PB_DEMPTY + i), 4);    // the 4 epilogue warps of the prototype tile that used the slot
    }
    mbar_init(bar(PB_RES), 1);
    mbar_init(bar(PB_RES2), 1);
    mbar_init(bar(PB_ATMEM), 8);
    for (int i = 0; i < 4; ++i) {
      mbar_init(bar(PB_SFULL + i), 1);     // staging buffer: arrive.expect_tx + the bytes of its 64 bulk copies
      mbar_init(bar(PB_SEMPTY + i), 4);    // the 4 warps of the loader group that converted it
    }           // the 8 epilogue warps have written the prototype operand into TMEM
    mbar_fence_init();
  }
  fence_proxy_async();
  if (warp == WARP_MMA) tmem_alloc(smem_u32((const void*)tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == WARP_RES) {
    if (lane == 0) {
      // the norm K step of the prototype tiles stays in shared memory (it also carries the class mask); -2p goes to TMEM
      mbar_arrive_expect_tx(bar(PB_RES), (uint32_t)nch * CN * 32);
      for (int j = 0; j < nch; ++j) {
        const unsigned char* src = prm.img + (size_t)(c0 + j) * ct.img_bytes;
        bulk_g2s(base + L::SM_RES + (j >> 1) * L::RES_N + (uint32_t)(j & 1) * CN * 32, src + OFF_BN, CN * 32, bar(PB_RES));
      }
    }
    if (MASKED) {
      // prototype side of the class mask: a 1 in the slot of each digit of the prototype's label value
      mbar_wait(bar(PB_RES), 0);
      for (int r = lane; r < 2 * TILE_M; r += 32) {
        const int jr = r >> 6, jj = r & 63;
        if (jr < nch && jj < ct.n[c0 + jr]) {
          const int cls1 = __ldg(prm.proto_class + ct.p0[c0 + jr] + jj) + 1;
          ZStage<X, NSPLIT>::write_proto_mask(gbase + L::SM_RES + (r >> 7) * L::RES_N, r & 127, cls1);
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(PB_RES2));
    }
    if (ZMODE == ZM_LDG && lane == 0) {
      // fp32 NCHW maps: this warp is the TMA producer, four boxes per tile
      const int rows = s.D / 4;
      for (int i = 0; i < my_tiles; ++i) {
        const int slot = (i & 1) * 2 + ((i >> 1) & 1);
        mbar_wait(bar(PB_SEMPTY + slot), ((i >> 2) & 1) ^ 1);
        const long long t = t_lo + i;
        const int b = (int)(t / prm.tiles_per_map), pix0 = (int)(t % prm.tiles_per_map) * TILE_M;
        TLP(4, i);
        mbar_arrive_expect_tx(bar(PB_SFULL + slot), (uint32_t)(4 * rows * L::STG_PX * 4));
#pragma unroll
        for (int r = 0; r < 4; ++r)
          tma_load_3d(base + L::SM_STG + slot * L::STG_BYTES + (uint32_t)(r * rows) * (L::STG_PX * 4), &prm.tm_f32.map[r], pix0, sc * rows, b,
                      bar(PB_SFULL + slot));
      }
    }
    if (ZMODE != ZM_LDG && lane == 0) {
      // 16-bit channels-last maps: this warp is the TMA producer. Each tile is ONE box {64 channels of this scale, 128
      // positions} landing in the MMA's own layout; it runs as far ahead as the stage ring allows (its waits block nobody:
      // when the loader groups issued their own look-ahead, the wait for a future stage sat in front of the tile at hand)
      for (int i = 0; i < my_tiles; ++i) {
        const int st = i % L::NSTAGE;
        mbar_wait(bar(PB_ZEMPTY + st), ((i / L::NSTAGE) & 1) ^ 1);
        const long long t = t_lo + i;
        const long long row0 = (t / prm.tiles_per_map) * s.HW + (t % prm.tiles_per_map) * TILE_M;
        TLP(4, i);
        mbar_arrive_expect_tx(bar(PB_TFULL + st), TILE_M * 128u);
        tma_load_2d(base + L::SM_Z + st * L::STAGE, &prm.tm_z, sc * s.D, (int)row0, bar(PB_TFULL + st));
      }
    }
  } else if (warp == WARP_MMA) {
    // ================================ MMA issuer ================================
    constexpr uint32_t ID = idesc_f16(Fmt<X>::id, TILE_M, TILE_M);
    mbar_wait(bar(PB_RES), 0);
    if (MASKED) mbar_wait(bar(PB_RES2), 0);
    mbar_wait(bar(PB_ATMEM), 0);
    tc_fence_after();
    for (int i = 0; i < my_tiles; ++i) {
      const int st = i % L::NSTAGE;
      if (lane == 0) TLP(0, i);
      mbar_wait(bar(PB_ZFULL + st), (i / L::NSTAGE) & 1);
      if (lane == 0) TLP(1, i);
      tc_fence_after();
      const uint32_t zt = base + L::SM_Z + st * L::STAGE;
      for (int t = 0; t < n_mt; ++t) {
        const int u = i * n_mt + t, slot = u % D_SLOTS;
        mbar_wait(bar(PB_DEMPTY + slot), ((u / D_SLOTS) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d = tmem + slot * 128, ah = tmem + TM_A + t * 64, al = ah + 32;
        mma_f16(d, desc_noswz(base + L::SM_RES + t * L::RES_N, NORM_LBO, NORM_SBO), desc_noswz(zt + L::SOFF_N, NORM_LBO, NORM_SBO), ID, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_f16_ts(d, ah + 8 * k, desc_sw128(zt + 32 * k), ID, 1);
        if (NSPLIT == 3) {
          if (ZLO) {
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_f16_ts(d, ah + 8 * k, desc_sw128(zt + L::SOFF_LO + 32 * k), ID, 1);
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_f16_ts(d, al + 8 * k, desc_sw128(zt + 32 * k), ID, 1);
        }
        mma_commit(bar(PB_DFULL + slot));
      }
      mma_commit(bar(PB_ZEMPTY + st));
      if (lane == 0) TLP(3, i);
      __syncwarp();
    }
  } else if (warp >= WARP_LOAD0 && warp < WARP_MMA) {
    // ================================ feature loaders (two groups of 4 warps, alternate tiles) ================================
    // Software pipeline per group: the feature loads of the group's next tile are issued right after the current one
    // has been handed over and fly during the wait for its stage.
    const int lt = threadIdx.x - WARP_LOAD0 * 32, grp = lt >> 7, m = lt & 127;
    if (ZMODE == ZM_LDG) {
      ZStage<X, NSPLIT> zst;
      const int rows = s.D / 4;
      for (int i = grp; i < my_tiles; i += 2) {
        const long long t = t_lo + i;
        const long long b = t / prm.tiles_per_map;
        const int pix = (int)(t % prm.tiles_per_map) * TILE_M + m;
        const int lab = pix < s.HW ? (int)__ldg(prm.labels + b * s.HW + pix) : 0;
        const int slot = (i & 1) * 2 + ((i >> 1) & 1), st = i % L::NSTAGE;
        if (m == 0) TLP(5, i);
        mbar_wait(bar(PB_SFULL + slot), (i >> 2) & 1);
        if (m == 0) TLP(6, i);
        // channel c = 4j + r of this thread's position: box r, row j, element m + shift_r
        const float* stg = reinterpret_cast<const float*>(gbase + L::SM_STG + slot * L::STG_BYTES) + m;
        const int sh0 = prm.tm_f32.shift[0], sh1 = prm.tm_f32.shift[1], sh2 = prm.tm_f32.shift[2], sh3 = prm.tm_f32.shift[3];
        mbar_wait(bar(PB_ZEMPTY + st), ((i / L::NSTAGE) & 1) ^ 1);
        unsigned char* a0 = gbase + L::SM_Z + st * L::STAGE;
        float zn = 0.f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
#pragma unroll
          for (int k = 0; k < 32; ++k) {
            const int c = h * 32 + k;
            const int sh = (c & 3) == 0 ? sh0 : (c & 3) == 1 ? sh1 : (c & 3) == 2 ? sh2 : sh3;
            zst.v[c] = c < s.D ? stg[((c & 3) * rows + (c >> 2)) * L::STG_PX + sh] : 0.0f;
          }
          if (h == 1) {                                   // every value of the tile is in registers: the buffer can be refilled
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(PB_SEMPTY + slot));
          }
          zst.convert_half(h, m, a0, a0 + L::SOFF_LO, zn);
        }
        if (MASKED) {
          ZStage<X, NSPLIT>::store_norm_mask(m, a0 + L::SOFF_N, zn, lab);
        } else {
          ZStage<X, NSPLIT>::store_norm(m, a0 + L::SOFF_N, zn);
          reinterpret_cast<int*>(gbase + L::SM_LAB)[(i % L::LAB_RING) * TILE_M + m] = lab;
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(PB_ZFULL + st));
        if (m == 0) TLP(7, i);
      }
    } else {
      // 16-bit channels-last maps (tiles brought by the producer warp): the thread's job per row is the label, the squared
      // norm (read back from the landed tile) and the norm / class-mask K step.
      for (int i = grp; i < my_tiles; i += 2) {
        const long long t = t_lo + i;
        const long long b = t / prm.tiles_per_map;
        const int pix = (int)(t % prm.tiles_per_map) * TILE_M + m;
        const int lab = pix < s.HW ? (int)__ldg(prm.labels + b * s.HW + pix) : 0;
        const int st = i % L::NSTAGE;
        if (m == 0) TLP(5, i);
        mbar_wait(bar(PB_TFULL + st), (i / L::NSTAGE) & 1);
        if (m == 0) TLP(6, i);
        unsigned char* a0 = gbase + L::SM_Z + st * L::STAGE;
        const float zn = tile_row_norm<X, false>(a0, a0, m);
        if (MASKED) {
          ZStage<X, NSPLIT>::store_norm_mask(m, a0 + L::SOFF_N, zn, lab);
        } else {
          ZStage<X, NSPLIT>::store_norm(m, a0 + L::SOFF_N, zn);
          reinterpret_cast<int*>(gbase + L::SM_LAB)[(i % L::LAB_RING) * TILE_M + m] = lab;
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(PB_ZFULL + st));
        if (m == 0) TLP(7, i);
      }
    }
  }
  else if (warp < WARP_LOAD0) {
    // ================================ epilogue: one prototype per thread, running masked minimum ================================
    const int mt = warp >> 2, q = warp & 3;
    const int r = mt * 128 + q * 32 + lane;              // prototype row inside the CTA's group
    const uint32_t lane_base = ((uint32_t)(q * 32)) << 16;
    int p = -1, cls = -1;
    if ((r >> 6) < nch && (r & 63) < ct.n[c0 + (r >> 6)]) {
      p = ct.p0[c0 + (r >> 6)] + (r & 63);
      cls = __ldg(prm.proto_class + p) + 1;              // labels: 0 = void, c + 1 = class c
    }
    const bool live_warp = mt < n_mt;
    // the resident operand: row r of -2p (hi, lo) from the pre-pass image (K-major, 128-byte swizzle) into TMEM, lane =
    // row, 32-bit column j = channels 2j, 2j+1 (rows of missing chunks: zeros)
    {
      const int j = r >> 6, jr = r & 63;
#pragma unroll 1
      for (int part = 0; part < (NSPLIT == 3 ? 2 : 1); ++part) {
        uint32_t v[32];
        if (j < nch) {
          const uint4* src = reinterpret_cast<const uint4*>(prm.img + (size_t)(c0 + j) * ct.img_bytes + (part ? OFF_BLO : OFF_BHI) + jr * 128);
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const uint4 w = __ldg(src + (c ^ (jr & 7)));
            v[4 * c] = w.x; v[4 * c + 1] = w.y; v[4 * c + 2] = w.z; v[4 * c + 3] = w.w;
          }
        } else {
#pragma unroll
          for (int c = 0; c < 32; ++c) v[c] = 0u;
        }
        tmem_st32(tmem + lane_base + TM_A + mt * 64 + part * 32, v);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(PB_ATMEM));
    }
    float best = 3.0e38f;
    unsigned bidx = 0;
    long long cur_map = -1;
    auto flush = [&]() {
      // MASKED: every differing label digit added MASK_BIG, so a minimum that large means "no position of this class"
      if (p >= 0 && cur_map >= 0 && best < (MASKED ? 0.5f * MASK_BIG : 1.0e38f))
        atomicMin(prm.keys + cur_map * s.P + p, ((unsigned long long)__float_as_uint(best) << 32) | bidx);
    };
    for (int i = 0; live_warp && i < my_tiles; ++i) {
      const long long t = t_lo + i;
      const long long b = t / prm.tiles_per_map;
      const unsigned pix0 = (unsigned)(t % prm.tiles_per_map) * TILE_M;
      if (b != cur_map) {
        flush();
        cur_map = b;
        best = 3.0e38f;
        bidx = 0;
      }
      const int u = i * n_mt + mt, slot = u % D_SLOTS;
      if (threadIdx.x == 0) TLP(8, i);
      mbar_wait(bar(PB_DFULL + slot), (u / D_SLOTS) & 1);
      if (threadIdx.x == 0) TLP(9, i);
      tc_fence_after();
      const uint32_t tsrc = tmem + lane_base + slot * 128;
#ifdef SPS_PUSH_DBG_NOEPI
      if (true) {
      } else
#endif
      if (MASKED) {
        // the load of block cb+1 flies while block cb is reduced (two register buffers); four partial minima per block
        // instead of one 16-deep chain
        uint32_t acc[2][32];
        tmem_ld32(tsrc, acc[0]);
#pragma unroll
        for (int cb = 0; cb < 4; ++cb) {
          tmem_ld_wait();
          if (cb < 3) tmem_ld32(tsrc + (cb + 1) * 32, acc[(cb + 1) & 1]);
          const uint32_t* a = acc[cb & 1];
          float m4[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) m4[k] = __uint_as_float(a[k]);
#pragma unroll
          for (int jj = 4; jj < 32; ++jj) m4[jj & 3] = fminf(m4[jj & 3], __uint_as_float(a[jj]));
          const float mn = fmaxf(fminf(fminf(m4[0], m4[1]), fminf(m4[2], m4[3])), 0.0f);   // relu commutes with the minimum
          if (mn < best) {                                 // strict <: an earlier block keeps a tie
            best = mn;
            int jf = 31;                                   // first column whose clamped value is the minimum
#pragma unroll
            for (int jj = 30; jj >= 0; --jj)
              if (__uint_as_float(a[jj]) <= mn) jf = jj;
            bidx = pix0 + (unsigned)(cb * 32 + jf);
          }
        }
      } else {
        const int4* labs = reinterpret_cast<const int4*>(gbase + L::SM_LAB + (i % L::LAB_RING) * TILE_M * 4);
#pragma unroll 1
        for (int cb = 0; cb < 4; ++cb) {
          uint32_t acc[32];
          tmem_ld32(tsrc + cb * 32, acc);
          tmem_ld_wait();
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const int4 l4 = labs[cb * 8 + j4];
            const int lv[4] = {l4.x, l4.y, l4.z, l4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float d = fmaxf(__uint_as_float(acc[j4 * 4 + e]), 0.0f);
              if (lv[e] == cls && d < best) {            // strict <: the first occurrence keeps a tie
                best = d;
                bidx = pix0 + (unsigned)(cb * 32 + j4 * 4 + e);
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(PB_DEMPTY + slot));
      if (threadIdx.x == 0) TLP(10, i);
    }
    flush();
  }

  tc_fence_before();
  __syncthreads();
#ifdef SPS_PUSH_TIMELINE
  if (threadIdx.x == 0 && blockIdx.x < 512) {   // rows 12/13: start / end of every CTA on the global timer (ns)
    unsigned long long gt;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt));
    g_tlp[13 * 1024 + blockIdx.x] = (long long)gt;
    g_tlp[12 * 1024 + blockIdx.x] = cta_t0;
  }
#endif
  if (warp == WARP_MMA) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

}  // namespace tcp

#ifdef SPS_PUSH_TIMELINE
extern "C" int sps_debug_push_timeline(long long* host, int count) {
  return (int)cudaMemcpyFromSymbol(host, tcp::g_tlp, sizeof(long long) * (size_t)count);
}
#endif

// ---------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------
int build_chunk_table(const HeadShape& s, bool with_last_layer, tc::ChunkTable* ct);
int launch_proto_prep(const HeadShape& s, int math, const tc::ChunkTable& ct, const float* protos, const float* last_w,
                      unsigned char* img, cudaStream_t st);

static int build_pgroups(const tc::ChunkTable& ct, long long n_tiles, int n_sm, tcp::PGroupTable* gt) {
  int ng = 0;
  for (int c = 0; c < ct.n_chunks;) {
    int e = c + 1;
    while (e < ct.n_chunks && ct.scale[e] == ct.scale[c] && e - c < tcp::PGROUP_CHUNKS) ++e;
    SPS_REQUIRE(ng < tcp::MAX_PGROUPS, SPS_ERR_UNSUPPORTED, "tensor-core push: more than %d prototype groups", tcp::MAX_PGROUPS);
    gt->c0[ng] = c;
    gt->nch[ng] = (short)(e - c);
    gt->scale[ng] = ct.scale[c];
    ++ng;
    c = e;
  }
  gt->n_groups = ng;
  SPS_REQUIRE(ng <= n_sm, SPS_ERR_UNSUPPORTED, "tensor-core push: %d prototype groups exceed %d SMs", ng, n_sm);
  int next = 0;
  for (int g = 0; g < ng; ++g) {
    int n = n_sm / ng + (g < n_sm % ng ? 1 : 0);
    if ((long long)n > n_tiles) n = (int)n_tiles;
    if (n < 1) n = 1;
    gt->cta0[g] = (short)next;
    gt->ncta[g] = (short)n;
    next += n;
  }
  return next;
}

template <typename X, int NSPLIT, bool MASKED, int ZMODE>
static int launch_push_tc_t(const HeadShape& s, const tc::ChunkTable& ct, const void* z, const int16_t* labels,
                            const int32_t* proto_class, unsigned long long* keys, const unsigned char* img, cudaStream_t st) {
  tcp::PushParams prm;
  prm.s = s;
  prm.ct = ct;
  prm.tiles_per_map = (s.HW + tc::TILE_M - 1) / tc::TILE_M;
  prm.n_tiles = (long long)s.B * prm.tiles_per_map;
  const int grid = build_pgroups(ct, prm.n_tiles, sm_count(), &prm.gt);
  if (grid <= 0) return SPS_ERR_UNSUPPORTED;
  prm.z = reinterpret_cast<const float*>(z);
  if (ZMODE == tc::ZM_LDG) {
    const int rc = make_tmap_nchw_f32(&prm.tm_f32, prm.z, s.B, s.C, s.HW, tc::TILE_M, s.D / 4);
    if (rc != SPS_OK) return rc;
  }
  if (ZMODE != tc::ZM_LDG) {
    const int rc = make_tmap_2d_16bit(&prm.tm_z, z, std::is_same<X, __nv_bfloat16>::value, s.N, s.C, tc::TILE_M, 64, 128);
    if (rc != SPS_OK) return rc;
  }
  prm.img = img;
  prm.labels = labels;
  prm.proto_class = proto_class;
  prm.keys = keys;
  using L = tcp::PLay<NSPLIT, NSPLIT == 3 && ZMODE == tc::ZM_LDG, ZMODE == tc::ZM_LDG>;
  constexpr int SMEM = L::SM_TOTAL;
  static_assert(sizeof(tcp::PushParams) <= 4096, "kernel parameter block");
  static_assert(SMEM <= 227 * 1024 && L::NSTAGE >= 2 && L::NSTAGE + 2 <= L::LAB_RING, "push kernel shared memory plan");
  SPS_CUDA_TRY(cudaFuncSetAttribute(tcp::push_scan_tc_kernel<X, NSPLIT, MASKED, ZMODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  tcp::push_scan_tc_kernel<X, NSPLIT, MASKED, ZMODE><<<grid, tcp::NTHREADS, SMEM, st>>>(prm);
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

// keys must already hold the "absent" key (launch_push_fill).  z16: the maps are 16-bit channels-last (TMA-fed).
int launch_push_scan_tc(const HeadShape& s, int math, const void* z, bool z16, const float* protos, const int16_t* labels,
                        const int32_t* proto_class, unsigned long long* keys, void* ws, size_t ws_bytes, cudaStream_t st) {
  SPS_REQUIRE(s.D <= 64 && s.D % 8 == 0, SPS_ERR_UNSUPPORTED, "tensor-core push: D=%d must be a multiple of 8, <= 64", s.D);
  SPS_REQUIRE(!z16 || (s.D == 64 && s.N < (1ll << 31)), SPS_ERR_UNSUPPORTED, "TMA-fed push: D=%d (64 only), N=%lld (< 2^31)", s.D, s.N);
  tc::ChunkTable ct;
  int rc = build_chunk_table(s, false, &ct);
  if (rc != SPS_OK) return rc;
  const size_t need = (size_t)ct.n_chunks * ct.img_bytes + 1024;
  SPS_REQUIRE(ws != nullptr && ws_bytes >= need, SPS_ERR_WORKSPACE, "push_scan workspace: need %zu bytes, got %zu", need, ws_bytes);
  unsigned char* img = reinterpret_cast<unsigned char*>(((uintptr_t)ws + 1023) & ~(uintptr_t)1023);
  rc = launch_proto_prep(s, math, ct, protos, nullptr, img, st);
  if (rc != SPS_OK) return rc;
  // label values are class + 1 <= K; the in-GEMM class mask covers label values up to MASK_MAX_LABEL
  const bool masked = s.K <= tc::MASK_MAX_LABEL;
#define SPS_PUSH_CASE(X, NS)                                                                                              \
  if (z16)                                                                                                                \
    return masked ? launch_push_tc_t<X, NS, true, tc::ZM_TMA16>(s, ct, z, labels, proto_class, keys, img, st)             \
                  : launch_push_tc_t<X, NS, false, tc::ZM_TMA16>(s, ct, z, labels, proto_class, keys, img, st);           \
  return masked ? launch_push_tc_t<X, NS, true, tc::ZM_LDG>(s, ct, z, labels, proto_class, keys, img, st)                 \
                : launch_push_tc_t<X, NS, false, tc::ZM_LDG>(s, ct, z, labels, proto_class, keys, img, st)
  switch (math) {
    case SPS_MATH_F16: SPS_PUSH_CASE(__half, 1);
    case SPS_MATH_F16X3: SPS_PUSH_CASE(__half, 3);
    case SPS_MATH_BF16: SPS_PUSH_CASE(__nv_bfloat16, 1);
    case SPS_MATH_BF16X3: SPS_PUSH_CASE(__nv_bfloat16, 3);
  }
#undef SPS_PUSH_CASE
  SPS_REQUIRE(false, SPS_ERR_BAD_ARG, "math=%d is not a tensor-core mode", math);
}

}  // namespace sps
