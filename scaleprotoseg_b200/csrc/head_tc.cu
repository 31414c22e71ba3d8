// Tensor-core prototype head (sm_100a): fused forward.
//
//   distance GEMM  (tcgen05.mma, fp16/bf16 operands, fp32 accumulation in TMEM)
//       acc[n,p] = sum_k z_s(n)[k] * (-2 p_p[k])  +  |p_p|^2 * 1  +  1 * |z_s(n)|^2
//     the two squared norms ride in one extra K step (each split in three 16-bit terms), so the
//     accumulator already holds |z|^2 - 2 z.p + |p|^2 (segmentation/model/model_multiscale.py:267-279).
//   epilogue       (tcgen05.ld -> registers)   d = max(acc, 0);  a = log((d+1)/(d+eps))  (:324-330)
//                  optional dense stores of d (NCHW) and a (N x P); a is packed to 16 bit into shared memory
//   logits GEMM    (tcgen05.mma)               logits[n,k] += a[n,p] * W[k,p]          (:243-244)
//     accumulated in TMEM over every scale and chunk of the tile, so in fused mode no (N x P) map is
//     ever written to HBM.
//
// Work decomposition: a tile is 128 consecutive positions (rows n = (b*H + h)*W + w, the order of
// permute(0,2,3,1).reshape(-1,P), :369-370); the prototypes of each scale (ragged ranges allowed, :400-432)
// are cut into chunks of 64.  One persistent CTA per SM walks tiles; inside a tile it walks the chunks.
// Warp roles: warps 0-7 = epilogue (two per TMEM lane quarter; warps 0-3 take the even units, warps 4-7 the odd
// ones, so the two warps of an SM sub-partition run out of phase), warps 8-11 = feature-tile
// loaders (NCHW fp32 -> 16-bit K-major swizzled tiles + row norms), warp 12 = prototype-chunk producer (bulk
// async copies of the packed images written by the pre-pass), warp 13 = distance-GEMM issuer, warp 14 =
// logits-GEMM issuer (one lane of each issues; tcgen05 work of one thread executes in issue order, so two
// issuers keep the two dependency chains from blocking each other).
#include "tc_common.cuh"

namespace sps {
namespace tc {

#ifdef SPS_TIMELINE
// Debug build only: per-role clock stamps of CTA 0 (role r, event e) -> g_tl[r * 1024 + e].
__device__ long long g_tl[8 * 1024];
#define TL(role, ev)                                                                    \
  do {                                                                                  \
    if (blockIdx.x == 0 && (ev) < 1024) g_tl[(role) * 1024 + (ev)] = clock64();          \
  } while (0)
// every CTA: global-timer stamps (ns) of kernel entry / epilogue loop end / exit -> roles 5, 6, 7 at [512 + blockIdx.x]
__device__ __forceinline__ long long gtimer() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define TL_CTA(role) do { if (blockIdx.x < 512) g_tl[(role) * 1024 + 512 + blockIdx.x] = gtimer(); } while (0)
#else
#define TL(role, ev) do { } while (0)
#define TL_CTA(role) do { } while (0)
#endif

// ---------------------------------------------------------------------------------------------
// Pre-pass: packed prototype / last-layer images, one per chunk (layout in tc_common.cuh)
// ---------------------------------------------------------------------------------------------
template <typename X, int NSPLIT>
__global__ void __launch_bounds__(256) proto_prep_kernel(const __grid_constant__ ChunkTable ct, int D, int P, int K,
                                                          int act_kind, const float* __restrict__ protos,
                                                          const float* __restrict__ last_w,
                                                          unsigned char* __restrict__ img_base) {
  pdl_launch_dependents();  // the main kernel may start its prologue and feature loads; its producer warp waits for us
  const int c = blockIdx.x;
  const int p0 = ct.p0[c], n = ct.n[c];
  unsigned char* img = img_base + (size_t)c * ct.img_bytes;
  __shared__ float pn[CN];
  const int t = threadIdx.x;
  if (t < CN) pn[t] = 0.f;
  __syncthreads();
  // operand rows: 4 threads per row, 16 k each
  {
    const int j = t >> 2, kq = (t & 3) * 16;
    float part = pack_row16<X, NSPLIT>(protos + (size_t)(p0 + (j < n ? j : 0)) * D, j < n, D, -2.0f, img + OFF_BHI, img + OFF_BLO, j, kq);
    part += __shfl_xor_sync(0xffffffffu, part, 1);
    part += __shfl_xor_sync(0xffffffffu, part, 2);
    if ((t & 3) == 0) pn[j] = part;
  }
  __syncthreads();
  if (t < CN) store_proto_norm<X>(img + OFF_BN, t, pn[t], t < n);
  if (last_w != nullptr) {
    const float sc = act_kind == SPS_ACT_LOG ? 0.69314718055994531f : 1.0f;  // the epilogue produces log2
    // batches of 8 loads per thread, then the stores: one load -> store round per iteration made this loop a chain of
    // kp/4 dependent L2 round trips (most of the pre-pass time)
    const int total = ct.kp * CN;
    for (int i0 = t; i0 < total; i0 += 8 * blockDim.x) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int idx = i0 + u * blockDim.x, k = idx / CN, j = idx % CN;
        v[u] = (idx < total && k < K && j < n) ? sc * __ldg(last_w + (size_t)k * P + p0 + j) : 0.0f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int idx = i0 + u * blockDim.x, k = idx / CN, j = idx % CN;
        if (idx >= total) break;
        const X hi = Fmt<X>::from(v[u]);
        *reinterpret_cast<uint16_t*>(img + OFF_W2 + sw128_off(k, j)) = bits16(hi);
        *reinterpret_cast<uint16_t*>(img + OFF_W2 + ct.kp * 128 + sw128_off(k, j)) =
            bits16(Fmt<X>::from(NSPLIT == 3 ? v[u] - Fmt<X>::to(hi) : 0.0f));
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Forward kernel
// ---------------------------------------------------------------------------------------------
struct FwdParams {
  HeadShape s;
  ChunkTable ct;
  const float* z;
  const unsigned char* img;
  float* logits;
  float* dist;
  float* act;
  int n_tiles;
};

constexpr int NTHREADS = 512;
constexpr int WARP_EPI0 = 0, WARP_LOAD0 = 8, WARP_PROD = 12, WARP_MMA = 13, WARP_MMA2 = 14, WARP_PRODW = 15;
constexpr int MAX_NST = 8;                   // prototype-chunk ring depth is chosen at launch (2..MAX_NST)
constexpr int A_HI = TILE_M * 128;           // 16384
constexpr int A_NORM = TILE_M * 32;          // 4096
constexpr int P_BYTES = TILE_M * 128;        // 16384
constexpr int TMEM_S0 = 0, TMEM_L0 = 2 * CN; // S double buffer at columns [0,128), logits after it

template <int NSPLIT> __host__ __device__ constexpr int a_buf_bytes() { return (NSPLIT == 3 ? 2 : 1) * A_HI + A_NORM; }

// shared memory plan: [A x2][P x2][B ring x nst][W ring x NSTW][barriers]; offsets are computed, not tabulated, so
// that a run-time buffer index never turns into a local-memory lookup.  The prototype operands (hi [lo] | norm) and the
// last-layer tile of a chunk travel in SEPARATE rings: a B slot is free again as soon as the distance GEMM of its unit
// has completed, a W slot only after the logits GEMM a unit later, so the B ring runs several units ahead.
constexpr int MAX_NSTW = 4;   // last-layer tile ring: as deep as it is cheap (small K), the chunk producer runs that far ahead
struct SmemPlan {
  int a_stride, p0, b0, b_stride, nst, nstw, w0, w_stride, scr0, bar_off, total;
  __host__ __device__ int a_off(int i) const { return i * a_stride; }
  int p_stride;
  __host__ __device__ int p_off(int i) const { return p0 + i * p_stride; }
  __host__ __device__ int b_off(int i) const { return b0 + i * b_stride; }
  __host__ __device__ int w_off(int i) const { return w0 + i * w_stride; }
};
constexpr int ACT_SCRATCH = 32 * 33 * 4;   // per epilogue warp: a 32 x 32 fp32 block (padded) for the transposed activation store
template <int NSPLIT> __host__ __device__ inline SmemPlan smem_plan(int kp, bool dense_act) {
  SmemPlan pl;
  pl.a_stride = a_buf_bytes<NSPLIT>();
  pl.p0 = 2 * pl.a_stride;
  pl.p_stride = (NSPLIT == 3 ? 2 : 1) * P_BYTES;   // a tile (hi) [+ lo tile]
  pl.b0 = pl.p0 + 2 * pl.p_stride;
  pl.b_stride = ((NSPLIT == 3 ? 2 : 1) * CN * 128 + CN * 32 + 1023) / 1024 * 1024;
  pl.w_stride = ((NSPLIT == 3 ? 2 : 1) * kp * 128 + 1023) / 1024 * 1024;
  pl.nstw = pl.w_stride <= 8 * 1024 ? MAX_NSTW : 2;
  const int w_total = kp > 0 ? pl.nstw * pl.w_stride : 0;
  const int scr_total = dense_act ? 8 * ACT_SCRATCH : 0;
  int nst = (227 * 1024 - 512 - 1024 - pl.b0 - w_total - scr_total) / pl.b_stride;
  pl.nst = nst > MAX_NST ? MAX_NST : nst;
  pl.w0 = pl.b0 + (pl.nst > 0 ? pl.nst : 0) * pl.b_stride;
  pl.scr0 = pl.w0 + w_total;
  pl.bar_off = pl.scr0 + scr_total;
  pl.total = pl.bar_off + 512 + 1024;  // barriers + slack to align the base to 1024
  return pl;
}

// barrier indices
enum { BAR_AFULL = 0, BAR_AEMPTY = 2, BAR_BFULL = 4, BAR_BEMPTY = 4 + MAX_NST, BAR_SFULL = 4 + 2 * MAX_NST,
       BAR_SEMPTY = 6 + 2 * MAX_NST, BAR_PFULL = 8 + 2 * MAX_NST, BAR_PEMPTY = 10 + 2 * MAX_NST, BAR_LFULL = 12 + 2 * MAX_NST,
       BAR_LEMPTY = 14 + 2 * MAX_NST, BAR_WFULL = 16 + 2 * MAX_NST, BAR_WEMPTY = 20 + 2 * MAX_NST, BAR_COUNT = 24 + 2 * MAX_NST };

// DENSE: the distance and / or activation maps leave the SM as well.  A compile-time switch: as a run-time test the
// compiler merged both paths and the fused epilogue carried the (predicated-off) address arithmetic of the dense stores.
template <typename X, int NSPLIT, bool DENSE>
__global__ void __launch_bounds__(NTHREADS, 1) head_fwd_tc_kernel(const __grid_constant__ FwdParams prm) {
  extern __shared__ unsigned char smem_raw[];
  const HeadShape& s = prm.s;
  const ChunkTable& ct = prm.ct;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* gbase = smem_raw + (base - smem_u32(smem_raw));
  const SmemPlan pl = smem_plan<NSPLIT>(ct.kp, prm.act != nullptr);
  const uint32_t bar0 = base + pl.bar_off;
  auto bar = [&](int i) { return bar0 + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + pl.bar_off + 8 * BAR_COUNT);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int NC = ct.n_chunks, KP = ct.kp, NST = pl.nst, NSTW = pl.nstw;
  const bool has_w = KP > 0;
  uint32_t tmem_cols = 32;
  while ((int)tmem_cols < TMEM_L0 + 2 * KP) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    TL_CTA(5);
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(BAR_AFULL + i), 4);    // one arrive per loader warp
      mbar_init(bar(BAR_AEMPTY + i), 1);   // tcgen05.commit
      mbar_init(bar(BAR_SFULL + i), 1);    // tcgen05.commit
      mbar_init(bar(BAR_SEMPTY + i), 4);   // one arrive per epilogue warp of the set that owns the buffer
      mbar_init(bar(BAR_PFULL + i), 4);
      mbar_init(bar(BAR_PEMPTY + i), 1);
      mbar_init(bar(BAR_LFULL + i), 1);
      mbar_init(bar(BAR_LEMPTY + i), 4);   // epilogue warps 0..3 read the logits
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

  if (warp == WARP_MMA) {
    // ================================ distance-GEMM issuer ================================
    // The whole warp walks the loop (waits are warp-uniform); one lane issues.  The two GEMMs have their own
    // issuer warps: with a single issuer the distance GEMM of unit i+1 queued behind the wait for the
    // epilogue of unit i-1, and the epilogue starved (timeline in profiles/).
    constexpr uint32_t ID1 = idesc_f16(Fmt<X>::id, TILE_M, CN);
    const int U = my_tiles * NC;
    uint32_t acount = 0, st = 0, bph = 0;
    int c = 0;
    for (int i = 0; i < U; ++i) {
      const bool first_of_scale = c == 0 || ct.scale[c] != ct.scale[c - 1];
      const bool last_of_scale = c == NC - 1 || ct.scale[c + 1] != ct.scale[c];
      const uint32_t ab = acount & 1, sb = i & 1;
      if (lane == 0) TL(5, i);
      if (first_of_scale) mbar_wait(bar(BAR_AFULL + ab), (acount >> 1) & 1);
      if (lane == 0) TL(6, i);
      mbar_wait(bar(BAR_BFULL + st), bph);
      if (lane == 0) TL(4, i);
      mbar_wait(bar(BAR_SEMPTY + sb), ((i >> 1) & 1) ^ 1);
      if (lane == 0) TL(7, i);
      tc_fence_after();
      {
        const uint32_t a_hi = base + pl.a_off(ab), a_lo = a_hi + A_HI;
        const uint32_t a_n = a_hi + (NSPLIT == 3 ? 2 : 1) * A_HI;
        const uint32_t b_hi = base + pl.b_off(st) + OFF_BHI, b_lo = base + pl.b_off(st) + OFF_BLO;
        const uint32_t b_n = base + pl.b_off(st) + OFF_BN;
        const uint32_t d = tmem + TMEM_S0 + sb * CN;
        mma_f16(d, desc_noswz(a_n, NORM_LBO, NORM_SBO), desc_noswz(b_n, NORM_LBO, NORM_SBO), ID1, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(a_hi + 32 * k), desc_sw128(b_hi + 32 * k), ID1, 1);
        if (NSPLIT == 3) {
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(a_hi + 32 * k), desc_sw128(b_lo + 32 * k), ID1, 1);
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(a_lo + 32 * k), desc_sw128(b_hi + 32 * k), ID1, 1);
        }
        mma_commit(bar(BAR_SFULL + sb));
        if (last_of_scale) mma_commit(bar(BAR_AEMPTY + ab));
        mma_commit(bar(BAR_BEMPTY + st));   // the B slot is free once this unit's distance GEMM has read it
        if (lane == 0) TL(0, i);
      }
      __syncwarp();
      if (last_of_scale) ++acount;
      if (++st == (uint32_t)NST) { st = 0; bph ^= 1; }
      if (++c == NC) c = 0;
    }
  } else if (warp == WARP_MMA2) {
    // ================================ logits-GEMM issuer ================================
    if (has_w) {
      const uint32_t ID2 = idesc_f16(Fmt<X>::id, TILE_M, KP);
      const int U = my_tiles * NC;
      int c = 0, tl = 0;
      uint32_t sw = 0, wph = 0;
      for (int u = 0; u < U; ++u) {
        const uint32_t pb = u & 1, lb = tl & 1;
        if (c == 0) mbar_wait(bar(BAR_LEMPTY + lb), ((tl >> 1) & 1) ^ 1);
        mbar_wait(bar(BAR_WFULL + sw), wph);
        mbar_wait(bar(BAR_PFULL + pb), (u >> 1) & 1);
        tc_fence_after();
        {
          if (lane == 0) TL(1, u);
          const uint32_t pa = base + pl.p_off(pb), w2 = base + pl.w_off(sw);
          const uint32_t d = tmem + TMEM_L0 + lb * KP;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            mma_f16(d, desc_sw128(pa + 32 * k), desc_sw128(w2 + 32 * k), ID2, (c > 0 || k > 0) ? 1 : 0);
          if (NSPLIT == 3) {  // a_hi.W_lo + a_lo.W_hi
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(pa + 32 * k), desc_sw128(w2 + KP * 128 + 32 * k), ID2, 1);
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_f16(d, desc_sw128(pa + P_BYTES + 32 * k), desc_sw128(w2 + 32 * k), ID2, 1);
          }
          mma_commit(bar(BAR_PEMPTY + pb));
          mma_commit(bar(BAR_WEMPTY + sw));
          if (c == NC - 1) mma_commit(bar(BAR_LFULL + lb));
          if (lane == 0) TL(2, u);
        }
        __syncwarp();
        if (++sw == (uint32_t)NSTW) { sw = 0; wph ^= 1; }
        if (++c == NC) { c = 0; ++tl; }
      }
    }
  } else if (warp == WARP_PROD) {
    // ================================ prototype-chunk producer ================================
    // One bulk copy per unit (hi | norm [| lo] are contiguous in the packed image) and running counters; the last-layer
    // tiles have a producer warp of their own (see head_fwd_tma.cu: one warp doing both delivered every chunk just in time).
    {
      pdl_wait();  // the packed images come from the pre-pass launched just before this kernel
      const int U = my_tiles * NC;
      const uint32_t bytes_b = (NSPLIT == 3 ? 2 : 1) * CN * 128 + CN * 32;
      const unsigned char* img = prm.img;
      const int img_bytes = ct.img_bytes;
      int c = 0, st = 0;
      uint32_t ph = 1;
      const unsigned char* src = img;
      for (int i = 0; i < U; ++i) {
        mbar_wait(bar(BAR_BEMPTY + st), ph);
        if (lane == 0) {
          mbar_arrive_expect_tx(bar(BAR_BFULL + st), bytes_b);
          bulk_g2s(base + pl.b_off(st), src, bytes_b, bar(BAR_BFULL + st));
        }
        __syncwarp();
        src += img_bytes;
        if (++c == NC) { c = 0; src = img; }
        if (++st == NST) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == WARP_PRODW) {
    // ================================ last-layer tile producer ================================
    if (has_w) {
      pdl_wait();
      const int U = my_tiles * NC;
      const uint32_t bytes_w = (NSPLIT == 3 ? 2 : 1) * KP * 128;
      const unsigned char* img = prm.img + OFF_W2;
      const int img_bytes = ct.img_bytes;
      int c = 0, sw = 0;
      uint32_t ph = 1;
      const unsigned char* src = img;
      for (int i = 0; i < U; ++i) {
        mbar_wait(bar(BAR_WEMPTY + sw), ph);
        if (lane == 0) {
          mbar_arrive_expect_tx(bar(BAR_WFULL + sw), bytes_w);
          bulk_g2s(base + pl.w_off(sw), src, bytes_w, bar(BAR_WFULL + sw));
        }
        __syncwarp();
        src += img_bytes;
        if (++c == NC) { c = 0; src = img; }
        if (++sw == NSTW) { sw = 0; ph ^= 1; }
      }
    }
  } else if (warp >= WARP_LOAD0 && warp < WARP_PROD) {
    // ================================ feature-tile loaders ================================
    // Work list: (tile, scale) pairs in the order the distance-GEMM issuer consumes them.  Software pipeline: the
    // loads of pair i+1 are issued right after pair i has been handed over and fly during the wait for its buffer.
    const int m = threadIdx.x - WARP_LOAD0 * 32;  // row of the tile
    const uint32_t hw = (uint32_t)s.HW;
    auto row_ptr = [&](int ti, int c) -> const float* {
      long long n = ((long long)blockIdx.x + (long long)ti * gridDim.x) * TILE_M + m;
      if (n >= s.N) n = s.N - 1;  // rows past the end read a valid row; their results are never stored
      const long long b = n / s.HW, pix = n % s.HW;
      return prm.z + (b * (long long)s.C + (long long)ct.scale[c] * s.D) * s.HW + pix;
    };
    auto advance = [&](int& ti, int& c) {  // first chunk of the next scale, wrapping to the next tile
      const int sc = ct.scale[c];
      do { ++c; } while (c < NC && ct.scale[c] == sc);
      if (c == NC) { c = 0; ++ti; }
    };
    ZStage<X, NSPLIT> zst;
    int ti = 0, c = 0;
    uint32_t acount = 0;
    const float* cur = nullptr;
    if (my_tiles > 0) {
      cur = row_ptr(0, 0);
      zst.issue_half(0, cur, hw, s.D);
      zst.issue_half(1, cur, hw, s.D);
    }
    while (ti < my_tiles) {
      int nti = ti, nc = c;
      advance(nti, nc);
      const bool have_next = nti < my_tiles;
      const float* nxt = have_next ? row_ptr(nti, nc) : cur;
      const uint32_t ab = acount & 1;
      mbar_wait(bar(BAR_AEMPTY + ab), ((acount >> 1) & 1) ^ 1);
      unsigned char* a_hi = gbase + pl.a_off(ab);
      unsigned char* a_lo = a_hi + A_HI;
      unsigned char* a_n = a_hi + (NSPLIT == 3 ? 2 : 1) * A_HI;
      float zn = 0.f;
      zst.convert_half(0, m, a_hi, a_lo, zn);
      zst.convert_half(1, m, a_hi, a_lo, zn);
      ZStage<X, NSPLIT>::store_norm(m, a_n, zn);
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(BAR_AFULL + ab));
      // the next pair's loads go out AFTER the arrive: mbarrier.arrive has release semantics and would wait for them
      if (have_next) {
        zst.issue_half(0, nxt, hw, s.D);
        zst.issue_half(1, nxt, hw, s.D);
      }
      ++acount;
      ti = nti;
      c = nc;
      cur = nxt;
    }
  }
  else if (warp < WARP_LOAD0) {
    // ================================ epilogue ================================
    const int ew = warp - WARP_EPI0, q = ew & 3, half = ew >> 2;  // q == warp % 4: the TMEM lane quarter this warp may read
    const int m = q * 32 + lane;
    const uint32_t lane_base = ((uint32_t)(q * 32)) << 16;
    const float eps = s.eps, ome = 1.0f - s.eps;
    const bool is_log = s.act_kind == SPS_ACT_LOG;
    // Logits of tile `tj` (TMEM -> HBM).  Called one unit into the NEXT tile so that the epilogue never
    // sits waiting for the last logits GEMM of a tile (the logits accumulator is double buffered).
    auto store_logits = [&](int tj) {
      const long long nn = ((long long)blockIdx.x + (long long)tj * gridDim.x) * TILE_M + m;
      const uint32_t lb = tj & 1;
      mbar_wait(bar(BAR_LFULL + lb), (tj >> 1) & 1);
      tc_fence_after();
      for (int k0 = 0; k0 < KP; k0 += 16) {
        uint32_t lr[16];
        tmem_ld16(tmem + lane_base + TMEM_L0 + lb * KP + k0, lr);
        tmem_ld_wait();
        if (nn < s.N) {
#pragma unroll
          for (int k = 0; k < 16; ++k)
            if (k0 + k < s.K) prm.logits[nn * s.K + k0 + k] = __uint_as_float(lr[k]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(BAR_LEMPTY + lb));
    };
    // The two warps of a lane quarter (ew and ew+4 share an SM sub-partition) take ALTERNATE units: set 0 (warps
    // 0-3) the even ones, set 1 (warps 4-7) the odd ones, each walking both 32-column halves of its unit.  In
    // lock-step both warps sat in the MUFU phase together (XU pipe saturated) and then in the pack / store / wait
    // phase together (XU idle); out of phase one warp's MUFU work hides under the other's packing and waiting.
    const int set = half;
    const int defer_c = NC > 1 ? 1 : 0;
    int next_store = 0;   // set 0: next tile whose logits go to HBM
    const int U = my_tiles * NC;
    int ti = 0, c = 0;
    long long n = (long long)blockIdx.x * TILE_M + m;
    if (set == 1) { c = 1; if (c >= NC) { c = 0; ti = 1; n += (long long)gridDim.x * TILE_M; } }
    for (int u = set; u < U; u += 2) {
      const bool valid = n < s.N;
      const long long b = valid ? n / s.HW : 0, pix = valid ? n % s.HW : 0;
      float* drow = DENSE && prm.dist != nullptr ? prm.dist + b * (long long)s.P * s.HW + pix : nullptr;
      const uint32_t sb = u & 1, pb = u & 1, ph = (u >> 1) & 1;
      if (lane == 0 && q == 0) TL(3, 8 * u + 0);
      mbar_wait(bar(BAR_SFULL + sb), ph);
      if (lane == 0 && q == 0) TL(3, 8 * u + 1);
      tc_fence_after();
      unsigned char* prow = gbase + pl.p_off(pb) + (uint32_t)m * 128u;
#pragma unroll 1
      for (int hf = 0; hf < 2; ++hf) {
        uint32_t r[32];
        tmem_ld32(tmem + lane_base + TMEM_S0 + sb * CN + hf * 32, r);
        tmem_ld_wait();
        if (lane == 0 && q == 0) TL(3, 8 * u + 2 + 2 * hf);
        if (hf == 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(BAR_SEMPTY + sb));
        }
        const int col0 = hf * 32, ncols = ct.n[c] - col0;  // live columns of this half: [0, ncols)
        const int p0 = ct.p0[c] + col0;
        float av[32];  // log2 of the similarity ratio ('log'), or -d ('linear')
        // activation kind resolved per half-unit, not per element (a run-time select evaluates both formulas)
        if (is_log) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float d = fmaxf(__uint_as_float(r[j]), 0.0f);
            av[j] = lg2_approx(fmaf(ome, rcp_approx(d + eps), 1.0f));
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
            // (N x P) activation map: transpose the warp's 32 x 32 block through shared memory so that every store
            // instruction writes 128 contiguous bytes of one row (a thread's own 32 values belong to ONE row)
            const float sc = is_log ? 0.69314718055994531f : 1.0f;
            float* scr = reinterpret_cast<float*>(gbase + pl.scr0 + ew * ACT_SCRATCH);
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 32; ++j) scr[lane * 33 + j] = av[j] * sc;
            __syncwarp();
            const long long n_first = n - lane;                    // row of lane 0 (the warp's rows are consecutive)
            float* abase = prm.act + n_first * (long long)s.P + p0 + lane;
            const int rows = s.N - n_first < 32 ? (int)(s.N - n_first) : 32;
            if (lane < ncols) {
#pragma unroll 8
              for (int r = 0; r < 32; ++r)
                if (r < rows) abase[(long long)r * s.P] = scr[r * 33 + lane];
            }
          }
        }
        if (has_w) {
          if (hf == 0) mbar_wait(bar(BAR_PEMPTY + pb), ph ^ 1);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint32_t w[4], wl[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float x0 = av[g * 8 + 2 * e], x1 = av[g * 8 + 2 * e + 1];
              w[e] = Pack2<X>::pack(x0, x1);
              if (NSPLIT == 3) {
                const float2 h = Pack2<X>::unpack(w[e]);
                wl[e] = Pack2<X>::pack(x0 - h.x, x1 - h.y);
              }
            }
            const uint32_t poff = (((uint32_t)(hf * 4 + g)) ^ ((uint32_t)m & 7u)) << 4;
            *reinterpret_cast<uint4*>(prow + poff) = make_uint4(w[0], w[1], w[2], w[3]);
            if (NSPLIT == 3) *reinterpret_cast<uint4*>(prow + P_BYTES + poff) = make_uint4(wl[0], wl[1], wl[2], wl[3]);
          }
        }
        if (lane == 0 && q == 0) TL(3, 8 * u + 3 + 2 * hf);
      }
      if (has_w) {
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(BAR_PFULL + pb));
      }
      if (lane == 0 && q == 0) TL(3, 8 * u + 7);
      if (has_w && set == 0) {
        while (next_store < ti - 1) store_logits(next_store++);
        if (next_store == ti - 1 && c >= defer_c) store_logits(next_store++);
      }
      // advance two units
      c += 2;
      while (c >= NC) { c -= NC; ++ti; n += (long long)gridDim.x * TILE_M; }
    }
    if (threadIdx.x == 0) TL_CTA(6);
    if (has_w && set == 0)
      while (next_store < my_tiles) store_logits(next_store++);
  }

  // ================================ teardown ================================
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) TL_CTA(7);
  if (warp == WARP_MMA) {
    tc_fence_after();
    tmem_dealloc(tmem, tmem_cols);
  }
}

}  // namespace tc

#ifdef SPS_TIMELINE
extern "C" int sps_debug_timeline(long long* host, int count) {
  return (int)cudaMemcpyFromSymbol(host, tc::g_tl, sizeof(long long) * (size_t)count);
}
#endif

// ---------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------
int build_chunk_table(const HeadShape& s, bool with_last_layer, tc::ChunkTable* ct) {
  int nc = 0;
  for (int sc = 0; sc < s.S; ++sc) {
    for (int p0 = s.scale_lo[sc]; p0 < s.scale_lo[sc + 1]; p0 += tc::CN) {
      SPS_REQUIRE(nc < tc::MAX_CHUNKS, SPS_ERR_UNSUPPORTED, "tensor-core head: more than %d prototype chunks (P=%d)",
                  tc::MAX_CHUNKS, s.P);
      ct->p0[nc] = p0;
      ct->n[nc] = (short)(s.scale_lo[sc + 1] - p0 < tc::CN ? s.scale_lo[sc + 1] - p0 : tc::CN);
      ct->scale[nc] = (short)sc;
      ++nc;
    }
  }
  ct->n_chunks = nc;
  ct->kp = with_last_layer ? (s.K + 15) / 16 * 16 : 0;
  ct->img_bytes = tc::chunk_img_bytes(ct->kp);
  return SPS_OK;
}

int launch_proto_prep(const HeadShape& s, int math, const tc::ChunkTable& ct, const float* protos, const float* last_w,
                      unsigned char* img, cudaStream_t st) {
  switch (math) {
    case SPS_MATH_F16: tc::proto_prep_kernel<__half, 1><<<ct.n_chunks, 256, 0, st>>>(ct, s.D, s.P, s.K, s.act_kind, protos, last_w, img); break;
    case SPS_MATH_F16X3: tc::proto_prep_kernel<__half, 3><<<ct.n_chunks, 256, 0, st>>>(ct, s.D, s.P, s.K, s.act_kind, protos, last_w, img); break;
    case SPS_MATH_BF16: tc::proto_prep_kernel<__nv_bfloat16, 1><<<ct.n_chunks, 256, 0, st>>>(ct, s.D, s.P, s.K, s.act_kind, protos, last_w, img); break;
    case SPS_MATH_BF16X3: tc::proto_prep_kernel<__nv_bfloat16, 3><<<ct.n_chunks, 256, 0, st>>>(ct, s.D, s.P, s.K, s.act_kind, protos, last_w, img); break;
    default: SPS_REQUIRE(false, SPS_ERR_BAD_ARG, "math=%d is not a tensor-core mode", math);
  }
  count_launch();
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

size_t head_fwd_tc_workspace(const HeadShape& s) {
  tc::ChunkTable ct;
  if (build_chunk_table(s, true, &ct) != SPS_OK) return 0;
  return (size_t)ct.n_chunks * ct.img_bytes + 1024;
}

template <typename X, int NSPLIT>
static int launch_fwd_tc_t(const HeadShape& s, const tc::ChunkTable& ct, const float* z, const float* protos,
                           const float* last_w, float* logits, float* dist, float* act, unsigned char* img,
                           cudaStream_t st) {
  tc::proto_prep_kernel<X, NSPLIT><<<ct.n_chunks, 256, 0, st>>>(ct, s.D, s.P, s.K, s.act_kind, protos, last_w, img);
  SPS_CUDA_TRY(cudaGetLastError());
  count_launch();
  tc::FwdParams prm;
  prm.s = s;
  prm.ct = ct;
  prm.z = z;
  prm.img = img;
  prm.logits = logits;
  prm.dist = dist;
  prm.act = act;
  prm.n_tiles = (int)((s.N + tc::TILE_M - 1) / tc::TILE_M);
  const tc::SmemPlan pl = tc::smem_plan<NSPLIT>(ct.kp, act != nullptr);
  SPS_REQUIRE(pl.nst >= 2 && pl.total <= 227 * 1024, SPS_ERR_UNSUPPORTED, "tensor-core head: %d bytes of shared memory needed (K=%d)",
              pl.total, s.K);
  const int grid = prm.n_tiles < sm_count() ? prm.n_tiles : sm_count();
  if (dist != nullptr || act != nullptr) {
    SPS_CUDA_TRY(cudaFuncSetAttribute(tc::head_fwd_tc_kernel<X, NSPLIT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total));
    SPS_CUDA_TRY(launch_pdl(tc::head_fwd_tc_kernel<X, NSPLIT, true>, dim3(grid), dim3(tc::NTHREADS), (size_t)pl.total, st, prm));
  } else {
    SPS_CUDA_TRY(cudaFuncSetAttribute(tc::head_fwd_tc_kernel<X, NSPLIT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total));
    SPS_CUDA_TRY(launch_pdl(tc::head_fwd_tc_kernel<X, NSPLIT, false>, dim3(grid), dim3(tc::NTHREADS), (size_t)pl.total, st, prm));
  }
  count_launch();
  return SPS_OK;
}

// Shapes the tcgen05 forward covers in this math mode; api.cu routes everything else to the exact fp32 kernels
// (e.g. K = 171 with 3-term products: the hi + lo last-layer tiles leave room for a single ring stage).
bool head_fwd_tc_supported(const HeadShape& s, int math, bool with_last_layer, bool dense_act) {
  if (math == SPS_MATH_FP32 || s.D > 64 || s.D % 8 != 0 || s.K > 192) return false;
  int nc = 0;
  for (int sc = 0; sc < s.S; ++sc) nc += (s.scale_lo[sc + 1] - s.scale_lo[sc] + tc::CN - 1) / tc::CN;
  if (nc > tc::MAX_CHUNKS) return false;
  const int kp = with_last_layer ? (s.K + 15) / 16 * 16 : 0;
  const bool split = math == SPS_MATH_F16X3 || math == SPS_MATH_BF16X3;
  const tc::SmemPlan pl = split ? tc::smem_plan<3>(kp, dense_act) : tc::smem_plan<1>(kp, dense_act);
  return pl.nst >= 2 && pl.total <= 227 * 1024;
}

int launch_head_fwd_tc(const HeadShape& s, int math, const float* z, const float* protos, const float* last_w,
                       float* logits, float* dist, float* act, void* ws, size_t ws_bytes, cudaStream_t st) {
  SPS_REQUIRE(s.D <= 64 && s.D % 8 == 0, SPS_ERR_UNSUPPORTED, "tensor-core head: D=%d must be a multiple of 8, <= 64", s.D);
  SPS_REQUIRE(s.K <= 192, SPS_ERR_UNSUPPORTED, "tensor-core head: K=%d > 192 classes", s.K);
  if (s.N == 0 || s.P == 0) return SPS_OK;
  tc::ChunkTable ct;
  int rc = build_chunk_table(s, last_w != nullptr, &ct);
  if (rc != SPS_OK) return rc;
  const size_t need = (size_t)ct.n_chunks * ct.img_bytes + 1024;
  SPS_REQUIRE(ws != nullptr && ws_bytes >= need, SPS_ERR_WORKSPACE, "head_fwd workspace: need %zu bytes, got %zu", need,
              ws_bytes);
  unsigned char* img = reinterpret_cast<unsigned char*>(((uintptr_t)ws + 1023) & ~(uintptr_t)1023);
  switch (math) {
    case SPS_MATH_F16: return launch_fwd_tc_t<__half, 1>(s, ct, z, protos, last_w, logits, dist, act, img, st);
    case SPS_MATH_F16X3: return launch_fwd_tc_t<__half, 3>(s, ct, z, protos, last_w, logits, dist, act, img, st);
    case SPS_MATH_BF16: return launch_fwd_tc_t<__nv_bfloat16, 1>(s, ct, z, protos, last_w, logits, dist, act, img, st);
    case SPS_MATH_BF16X3: return launch_fwd_tc_t<__nv_bfloat16, 3>(s, ct, z, protos, last_w, logits, dist, act, img, st);
  }
  SPS_REQUIRE(false, SPS_ERR_BAD_ARG, "math=%d is not a tensor-core mode", math);
}

}  // namespace sps
