// Tensor-core push scan (sm_100a): per map and per prototype, the nearest position among those labelled with the
// prototype's class.  Replaces compute_distances lines 68-91 of segmentation/push_multiscale_optimization.py
// (one-hot, void column dropped, 1e10 mask, min / argmin over the flattened map).
//
// The distance GEMM is the forward's, TRANSPOSED: prototypes are the M rows (A operand, resident in shared
// memory for the whole kernel), positions are the N columns (B operand, streamed), so every epilogue thread owns
// one prototype and keeps its running (distance, index) minimum in registers — no cross-thread reduction.
//   acc[p, n] = (-2 p_p).z_s(n) + |p_p|^2 + |z_s(n)|^2       (norms in the extra K step, as in head_tc.cu)
//   key[p]    = min over n with label[n] == class(p)+1 of (bits(max(acc,0)) << 32 | n)   (first occurrence wins)
// One CTA owns ONE scale and up to 256 of its prototypes (two M=128 accumulator tiles) and a contiguous range of
// position tiles (128 positions of one map), so each feature value is read from HBM exactly once per 256
// prototypes of its scale (once in total for the shipped configs).  The scan is HBM-bound: the feature loaders
// (8 warps, two tiles in flight) are the critical resource; MMA and the compare/select epilogue hide under them.
// Results are merged into keys[map][p] with one 64-bit atomicMin per (CTA, map, prototype).
#include "tma.cuh"

namespace sps {
namespace tcp {
using namespace sps::tc;

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
template <int NSPLIT, bool ZLO> struct PLay {
  static constexpr int RES_T = (NSPLIT == 3 ? 2 : 1) * TILE_M * 128 + TILE_M * 32;        // one 128-row prototype tile
  static constexpr int ROFF_LO = TILE_M * 128, ROFF_N = (NSPLIT == 3 ? 2 : 1) * TILE_M * 128;
  static constexpr int STAGE = (ZLO ? 2 : 1) * TILE_M * 128 + TILE_M * 32;                // z hi [lo] | norm
  static constexpr int SOFF_LO = TILE_M * 128, SOFF_N = (ZLO ? 2 : 1) * TILE_M * 128;
  static constexpr int SM_RES = 0;
  static constexpr int SM_Z = 2 * RES_T;
  static constexpr int NSTAGE = (227 * 1024 - 2048 - 4096 - SM_Z) / STAGE >= 6 ? 6 : (227 * 1024 - 2048 - 4096 - SM_Z) / STAGE;
  // labels of the tile in flight: a ring of LAB_RING >= NSTAGE + 2 slots needs no barrier of its own (the loader of
  // tile i has waited for the MMA of tile i-NSTAGE, which waited for the epilogue of tile i-NSTAGE-2)
  static constexpr int LAB_RING = 8;
  static constexpr int SM_LAB = SM_Z + NSTAGE * STAGE;
  static constexpr int SM_BAR = SM_LAB + LAB_RING * TILE_M * 4;
  static constexpr int SM_TOTAL = SM_BAR + 256 + 1024;
};
enum { PB_ZFULL = 0, PB_ZEMPTY = 8, PB_DFULL = 16, PB_DEMPTY = 18, PB_RES = 20, PB_RES2 = 21, PB_TFULL = 22, PB_COUNT = 30 };

__device__ __forceinline__ unsigned long long absent_key64() { return ((unsigned long long)0x501502F9u) << 32; }  // 1e10f

// MASKED: the class mask rides in the norm K step of the GEMM (labels <= MASK_MAX_LABEL, see ZStage::store_norm_mask)
// and the epilogue is a plain running minimum; otherwise the epilogue compares labels element by element.
template <typename X, int NSPLIT, bool MASKED, int ZMODE>
__global__ void __launch_bounds__(NTHREADS, 1) push_scan_tc_kernel(const __grid_constant__ PushParams prm) {
  constexpr bool ZLO = NSPLIT == 3 && ZMODE == ZM_LDG;
  using L = PLay<NSPLIT, ZLO>;
  extern __shared__ unsigned char smem_raw[];
  const HeadShape& s = prm.s;
  const ChunkTable& ct = prm.ct;
  const PGroupTable& gt = prm.gt;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* gbase = smem_raw + (base - smem_u32(smem_raw));
  auto bar = [&](int i) { return base + L::SM_BAR + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + L::SM_BAR + 8 * PB_COUNT);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  int g = 0;
  while (g + 1 < gt.n_groups && (int)blockIdx.x >= gt.cta0[g + 1]) ++g;
  const int cta_local = (int)blockIdx.x - gt.cta0[g], ncta = gt.ncta[g];
  const int nch = gt.nch[g], c0 = gt.c0[g], sc = gt.scale[g];
  const int n_mt = (nch + 1) / 2;                       // M tiles (128 rows) in use: 1 or 2
  const long long t_lo = prm.n_tiles * cta_local / ncta, t_hi = prm.n_tiles * (cta_local + 1) / ncta;
  const int my_tiles = (int)(t_hi - t_lo);

  // zero the resident operand area (rows of missing chunks must read as zeros), then barrier init
  for (int i = threadIdx.x; i < 2 * L::RES_T / 16; i += NTHREADS) reinterpret_cast<uint4*>(gbase + L::SM_RES)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) {
    for (int i = 0; i < L::NSTAGE; ++i) {
      mbar_init(bar(PB_ZFULL + i), 4);     // the 4 warps of one loader group
      mbar_init(bar(PB_ZEMPTY + i), 1);    // tcgen05.commit
      mbar_init(bar(PB_TFULL + i), 1);     // TMA mode: arrive.expect_tx + the tile bytes
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(PB_DFULL + i), 1);
      mbar_init(bar(PB_DEMPTY + i), 8);
    }
    mbar_init(bar(PB_RES), 1);
    mbar_init(bar(PB_RES2), 1);
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
      const uint32_t per = (NSPLIT == 3 ? 2 : 1) * CN * 128 + CN * 32;
      mbar_arrive_expect_tx(bar(PB_RES), (uint32_t)nch * per);
      for (int j = 0; j < nch; ++j) {
        const unsigned char* src = prm.img + (size_t)(c0 + j) * ct.img_bytes;
        const uint32_t dst = base + L::SM_RES + (j >> 1) * L::RES_T, half = (uint32_t)(j & 1);
        bulk_g2s(dst + half * CN * 128, src + OFF_BHI, CN * 128, bar(PB_RES));
        if (NSPLIT == 3) bulk_g2s(dst + L::ROFF_LO + half * CN * 128, src + OFF_BLO, CN * 128, bar(PB_RES));
        bulk_g2s(dst + L::ROFF_N + half * CN * 32, src + OFF_BN, CN * 32, bar(PB_RES));
      }
    }
    if (MASKED) {
      // prototype side of the class mask: a 1 in the slot of each digit of the prototype's label value
      mbar_wait(bar(PB_RES), 0);
      for (int r = lane; r < 2 * TILE_M; r += 32) {
        const int jr = r >> 6, jj = r & 63;
        if (jr < nch && jj < ct.n[c0 + jr]) {
          const int cls1 = __ldg(prm.proto_class + ct.p0[c0 + jr] + jj) + 1;
          ZStage<X, NSPLIT>::write_proto_mask(gbase + L::SM_RES + (r >> 7) * L::RES_T + L::ROFF_N, r & 127, cls1);
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(PB_RES2));
    }
  } else if (warp == WARP_MMA) {
    // ================================ MMA issuer ================================
    constexpr uint32_t ID = idesc_f16(Fmt<X>::id, TILE_M, TILE_M);
    mbar_wait(bar(PB_RES), 0);
    if (MASKED) mbar_wait(bar(PB_RES2), 0);
    for (int i = 0; i < my_tiles; ++i) {
      const int st = i % L::NSTAGE, db = i & 1;
      mbar_wait(bar(PB_ZFULL + st), (i / L::NSTAGE) & 1);
      mbar_wait(bar(PB_DEMPTY + db), ((i >> 1) & 1) ^ 1);
      tc_fence_after();
      {
        const uint32_t zt = base + L::SM_Z + st * L::STAGE;
        for (int t = 0; t < n_mt; ++t) {
          const uint32_t pt = base + L::SM_RES + t * L::RES_T, d = tmem + db * 256 + t * 128;
          mma_f16(d, desc_noswz(pt + L::ROFF_N, NORM_LBO, NORM_SBO), desc_noswz(zt + L::SOFF_N, NORM_LBO, NORM_SBO), ID, 0);
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(pt + 32 * k), desc_sw128(zt + 32 * k), ID, 1);
          if (NSPLIT == 3) {
            if (ZLO) {
#pragma unroll
              for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(pt + 32 * k), desc_sw128(zt + L::SOFF_LO + 32 * k), ID, 1);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(pt + L::ROFF_LO + 32 * k), desc_sw128(zt + 32 * k), ID, 1);
          }
        }
        mma_commit(bar(PB_DFULL + db));
        mma_commit(bar(PB_ZEMPTY + st));
      }
      __syncwarp();
    }
  } else if (warp >= WARP_LOAD0 && warp < WARP_MMA) {
    // ================================ feature loaders (two groups of 4 warps, alternate tiles) ================================
    // Software pipeline per group: the feature loads of the group's next tile are issued right after the current one
    // has been handed over and fly during the wait for its stage.
    const int lt = threadIdx.x - WARP_LOAD0 * 32, grp = lt >> 7, m = lt & 127;
    const uint32_t hw = (uint32_t)s.HW;
    auto row_ptr = [&](int i, long long& b, int& pix) -> const float* {
      const long long t = t_lo + i;
      b = t / prm.tiles_per_map;
      pix = (int)(t % prm.tiles_per_map) * TILE_M + m;
      const int px = pix < s.HW ? pix : s.HW - 1;  // rows past the end of the map read a valid row; their label is void
      return prm.z + (b * s.C + (long long)sc * s.D) * s.HW + px;
    };
    if (ZMODE == ZM_LDG) {
    ZStage<X, NSPLIT> zst;
    long long b = 0, nb = 0;
    int pix = 0, npix = 0;
    const float* cur = nullptr;
    if (grp < my_tiles) {
      cur = row_ptr(grp, b, pix);
      zst.issue_half(0, cur, hw, s.D);
      zst.issue_half(1, cur, hw, s.D);
    }
    for (int i = grp; i < my_tiles; i += 2) {
      const bool have_next = i + 2 < my_tiles;
      const float* nxt = have_next ? row_ptr(i + 2, nb, npix) : cur;
      const int lab = pix < s.HW ? (int)__ldg(prm.labels + b * s.HW + pix) : 0;
      const int st = i % L::NSTAGE;
      mbar_wait(bar(PB_ZEMPTY + st), ((i / L::NSTAGE) & 1) ^ 1);
      unsigned char* a0 = gbase + L::SM_Z + st * L::STAGE;
      float zn = 0.f;
      zst.convert_half(0, m, a0, a0 + L::SOFF_LO, zn);
      zst.convert_half(1, m, a0, a0 + L::SOFF_LO, zn);
      if (MASKED) {
        ZStage<X, NSPLIT>::store_norm_mask(m, a0 + L::SOFF_N, zn, lab);
      } else {
        ZStage<X, NSPLIT>::store_norm(m, a0 + L::SOFF_N, zn);
        reinterpret_cast<int*>(gbase + L::SM_LAB)[(i % L::LAB_RING) * TILE_M + m] = lab;
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(PB_ZFULL + st));
      // the next tile's loads go out AFTER the arrive: mbarrier.arrive has release semantics and would wait for them
      if (have_next) {
        zst.issue_half(0, nxt, hw, s.D);
        zst.issue_half(1, nxt, hw, s.D);
      }
      cur = nxt;
      b = nb;
      pix = npix;
    }
    } else {
      // 16-bit channels-last maps: each tile is ONE TMA box {64 channels of this scale, 128 positions} landing in the
      // MMA's own layout; every loader group keeps three tiles in flight.  The thread's job per row: label, squared norm
      // (read back from the landed tile) and the norm / class-mask K step.
      int next_issue = grp;
      auto issue_upto = [&](int upto) {
        if (m != 0) return;
        while (next_issue <= upto && next_issue < my_tiles) {
          const int st = next_issue % L::NSTAGE;
          mbar_wait(bar(PB_ZEMPTY + st), ((next_issue / L::NSTAGE) & 1) ^ 1);
          const long long t = t_lo + next_issue;
          const long long row0 = (t / prm.tiles_per_map) * s.HW + (t % prm.tiles_per_map) * TILE_M;
          mbar_arrive_expect_tx(bar(PB_TFULL + st), TILE_M * 128u);
          tma_load_2d(base + L::SM_Z + st * L::STAGE, &prm.tm_z, sc * s.D, (int)row0, bar(PB_TFULL + st));
          next_issue += 2;
        }
      };
      issue_upto(grp + 2);
      for (int i = grp; i < my_tiles; i += 2) {
        issue_upto(i + 4);
        const long long t = t_lo + i;
        const long long b = t / prm.tiles_per_map;
        const int pix = (int)(t % prm.tiles_per_map) * TILE_M + m;
        const int lab = pix < s.HW ? (int)__ldg(prm.labels + b * s.HW + pix) : 0;
        const int st = i % L::NSTAGE;
        mbar_wait(bar(PB_TFULL + st), (i / L::NSTAGE) & 1);
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
    float best = 3.0e38f;
    unsigned bidx = 0;
    long long cur_map = -1;
    auto flush = [&]() {
      // MASKED: every differing label digit added MASK_BIG, so a minimum that large means "no position of this class"
      if (p >= 0 && cur_map >= 0 && best < (MASKED ? 0.5f * MASK_BIG : 1.0e38f))
        atomicMin(prm.keys + cur_map * s.P + p, ((unsigned long long)__float_as_uint(best) << 32) | bidx);
    };
    for (int i = 0; i < my_tiles; ++i) {
      const long long t = t_lo + i;
      const long long b = t / prm.tiles_per_map;
      const unsigned pix0 = (unsigned)(t % prm.tiles_per_map) * TILE_M;
      const int db = i & 1;
      if (b != cur_map) {
        flush();
        cur_map = b;
        best = 3.0e38f;
        bidx = 0;
      }
      mbar_wait(bar(PB_DFULL + db), (i >> 1) & 1);
      tc_fence_after();
      if (live_warp && MASKED) {
#pragma unroll 1
        for (int cb = 0; cb < 4; ++cb) {
          uint32_t acc[32];
          tmem_ld32(tmem + lane_base + db * 256 + mt * 128 + cb * 32, acc);
          tmem_ld_wait();
          float mn = __uint_as_float(acc[0]);
#pragma unroll
          for (int j = 1; j < 32; ++j) mn = fminf(mn, __uint_as_float(acc[j]));
          mn = fmaxf(mn, 0.0f);                            // relu commutes with the minimum
          if (mn < best) {                                 // strict <: an earlier batch keeps a tie
            best = mn;
            int jf = 31;                                   // first column whose clamped value is the minimum
#pragma unroll
            for (int j = 30; j >= 0; --j)
              if (__uint_as_float(acc[j]) <= mn) jf = j;
            bidx = pix0 + (unsigned)(cb * 32 + jf);
          }
        }
      } else if (live_warp) {
        const int4* labs = reinterpret_cast<const int4*>(gbase + L::SM_LAB + (i % L::LAB_RING) * TILE_M * 4);
#pragma unroll 1
        for (int cb = 0; cb < 4; ++cb) {
          uint32_t acc[32];
          tmem_ld32(tmem + lane_base + db * 256 + mt * 128 + cb * 32, acc);
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
      if (lane == 0) mbar_arrive(bar(PB_DEMPTY + db));
    }
    flush();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == WARP_MMA) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

}  // namespace tcp

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
  if (ZMODE != tc::ZM_LDG) {
    const int rc = make_tmap_2d_16bit(&prm.tm_z, z, std::is_same<X, __nv_bfloat16>::value, s.N, s.C, tc::TILE_M, 64, 128);
    if (rc != SPS_OK) return rc;
  }
  prm.img = img;
  prm.labels = labels;
  prm.proto_class = proto_class;
  prm.keys = keys;
  using L = tcp::PLay<NSPLIT, NSPLIT == 3 && ZMODE == tc::ZM_LDG>;
  constexpr int SMEM = L::SM_TOTAL;
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
