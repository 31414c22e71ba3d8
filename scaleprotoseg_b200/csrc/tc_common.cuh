// sm_100a primitives used by the tensor-core kernels: mbarrier, bulk async copy, tcgen05 (alloc /
// mma / commit / ld), shared-memory matrix descriptors and the packed operand layouts shared by the
// prototype pre-pass and the main kernels.
#pragma once
#include <cuda_fp16.h>
#include <cuda_bf16.h>

#include "common.cuh"

namespace sps {
namespace tc {

// ---------------------------------------------------------------------------------------------
// Addresses / barriers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Non-blocking probe.  mbarrier.try_wait is NOT one: it suspends the thread for a system-dependent time limit (measured here:
// ~3 000 clk) before it reports "not yet", so an opportunistic "is the slot free?" written with try_wait stalls its caller.
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// generic-proxy writes (st.shared) -> visible to the async proxy (tcgen05.mma / bulk copies)
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// 1-D bulk copy global -> shared, completion counted on an mbarrier (UBLKCP). 16 B aligned, size % 16 == 0.
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---------------------------------------------------------------------------------------------
// tcgen05
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]^T, kind::f16 (fp16 or bf16 operands, fp32 accumulate). One thread issues.
// tcgen05.mma / tcgen05.commit are called by EVERY lane of a converged issuer warp on warp-uniform operands; the
// instruction itself is predicated on elect.sync (the same lane every time for the full mask, so the commit tracks
// the MMAs).  Issuing under `if (lane == 0)` instead makes the compiler wrap each MMA in an elect / branch loop that
// costs ~80 clk per instruction (measured, scratch/mma_rate.cu) against 28-50 clk for the shapes used here.
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// The same with the A operand in TMEM (written there by tcgen05.st: lane = row, 32-bit column j = K elements 2j, 2j+1; one
// K = 16 step is 8 columns).  An M128 x N64 x K16 MMA then takes 32.8 clk instead of 48.9, and any shape reads half the
// shared memory (profiles/tools/mma_ts_probe.cu).
__device__ __forceinline__ void mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0],"
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16,"
      " %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
        "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
        "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0],"
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
               "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// Arrives on the mbarrier when every tcgen05 op issued so far by the elected lane has completed.
// (implies tcgen05.fence::before_thread_sync)
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 32 lanes x 32 consecutive fp32 columns: thread t of the warp gets row (lane base + t).
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32"
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15,"
      " %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32"
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// ---------------------------------------------------------------------------------------------
// Descriptors (bit layout: CUTLASS cute/arch/mma_sm100_desc.hpp; PTX ISA "tcgen05 matrix descriptor")
// ---------------------------------------------------------------------------------------------
// K-major operand tile, 128-byte swizzle: rows of 128 B, 8-row groups of 1024 B (SBO), base 1024-B aligned.
__device__ __forceinline__ uint64_t desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) /*LBO (unused)*/ |
         ((uint64_t)(1024 >> 4) << 32) /*SBO*/ | ((uint64_t)1 << 46) /*version*/ | ((uint64_t)2 << 61) /*SWIZZLE_128B*/;
}
// K-major operand, no swizzle, one 16-element K step: core matrices of 8 rows x 16 B; the two K halves are
// `lbo` bytes apart, 8-row groups `sbo` bytes apart.
__device__ __forceinline__ uint64_t desc_noswz(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
         ((uint64_t)1 << 46);
}
// kind::f16 instruction descriptor: fp32 accumulate, A and B K-major, dense.
enum : uint32_t { FMT_F16 = 0, FMT_BF16 = 1 };
__host__ __device__ constexpr uint32_t idesc_f16(uint32_t fmt, int M, int N) {
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// MN-major operand, 128-byte swizzle (rows = K index, 64 MN elements per 128-B row; 8 K rows = 1024 B (SBO);
// further blocks of 64 MN elements `lbo` bytes apart).  The SAME bytes as a K-major SW128 tile whose rows are the
// K index of this view: the swizzle is a function of the byte address only.  Advance 16 K = +2048 B.
__device__ __forceinline__ uint64_t desc_mn_sw128(uint32_t smem_addr, uint32_t lbo) {
  return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// 64-byte swizzle: rows of 64 B (32 elements), 8 rows = 512 B.  K-major (rows = MN index, advance 16 K = +32 B)
// and MN-major (rows = K index, 32 MN elements per row, advance 16 K = +1024 B) views of the same bytes.
__device__ __forceinline__ uint64_t desc_k_sw64(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(512 >> 4) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)4 << 61);
}
__device__ __forceinline__ uint64_t desc_mn_sw64(uint32_t smem_addr, uint32_t lbo) {
  return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(512 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)4 << 61);
}
__host__ __device__ __forceinline__ uint32_t sw64_off(int r, int k) {  // row r (64 B), element k < 32
  return (uint32_t)r * 64u + ((((uint32_t)k >> 3) ^ (((uint32_t)r >> 1) & 3u)) << 4) + (((uint32_t)k & 7u) << 1);
}
// instruction descriptor with explicit majors (0 = K-major, 1 = MN-major)
__host__ __device__ constexpr uint32_t idesc_f16_major(uint32_t fmt, int M, int N, uint32_t a_mn, uint32_t b_mn) {
  return idesc_f16(fmt, M, N) | (a_mn << 15) | (b_mn << 16);
}

__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2_approx(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// byte offset of element (row r, k) inside a SW128 K-major tile of 2-byte elements with 64 k per row
__host__ __device__ __forceinline__ uint32_t sw128_off(int r, int k) {
  return (uint32_t)r * 128u + ((((uint32_t)k >> 3) ^ ((uint32_t)r & 7u)) << 4) + (((uint32_t)k & 7u) << 1);
}
// byte offset of element (row r, k<16) inside the no-swizzle 16-wide "norm" tile (LBO 128, SBO 256)
__host__ __device__ __forceinline__ uint32_t norm_off(int r, int k) {
  return (uint32_t)(r & 7) * 16u + (uint32_t)(r >> 3) * 256u + (uint32_t)(k >> 3) * 128u + (uint32_t)(k & 7) * 2u;
}
constexpr uint32_t NORM_LBO = 128, NORM_SBO = 256;

// ---------------------------------------------------------------------------------------------
// 16-bit operand formats.  X = __half or __nv_bfloat16.
// ---------------------------------------------------------------------------------------------
template <typename X> struct Fmt;
template <> struct Fmt<__half> {
  static constexpr uint32_t id = FMT_F16;
  static __device__ __forceinline__ __half from(float v) { return __float2half_rn(v); }
  static __device__ __forceinline__ float to(__half v) { return __half2float(v); }
};
template <> struct Fmt<__nv_bfloat16> {
  static constexpr uint32_t id = FMT_BF16;
  static __device__ __forceinline__ __nv_bfloat16 from(float v) { return __float2bfloat16_rn(v); }
  static __device__ __forceinline__ float to(__nv_bfloat16 v) { return __bfloat162float(v); }
};
template <typename X> __device__ __forceinline__ uint16_t bits16(X v) { return *reinterpret_cast<uint16_t*>(&v); }

// Two fp32 -> one packed 32-bit word of 16-bit values (lo = a, hi = b) with ONE F2FP instruction (the scalar
// conversions go through the XU pipe, which the log-similarity epilogue already saturates), and back.
template <typename X> struct Pack2;
template <> struct Pack2<__half> {
  static __device__ __forceinline__ uint32_t pack(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  static __device__ __forceinline__ float2 unpack(uint32_t w) { return __half22float2(*reinterpret_cast<__half2*>(&w)); }
};
template <> struct Pack2<__nv_bfloat16> {
  static __device__ __forceinline__ uint32_t pack(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  static __device__ __forceinline__ float2 unpack(uint32_t w) {
    return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xFFFF0000u));
  }
};

// v ~= h + m + l with three 16-bit terms (used for the squared norms that ride in the extra K step)
template <typename X> __device__ __forceinline__ void split3(float v, X& h, X& m, X& l) {
  h = Fmt<X>::from(v);
  float r = v - Fmt<X>::to(h);
  m = Fmt<X>::from(r);
  r -= Fmt<X>::to(m);
  l = Fmt<X>::from(r);
}

// ---------------------------------------------------------------------------------------------
// Feature-row staging shared by the forward, backward and push loaders.  One thread owns one position (row m
// of a 128-row tile) and the 64 channels of one scale: 64 coalesced LDGs (32 consecutive positions per warp and
// channel), then the 16-bit K-major SW128 tile(s) and the squared norm of the operand the MMA really sees.
// The 64 values are staged in two halves of 32 so that the loads of the NEXT row can be issued while the current
// half is still being converted (software pipeline across tiles; see the loader loops).
// ---------------------------------------------------------------------------------------------
constexpr float MASK_BIG = 32768.0f;   // added per differing label digit; exact in fp16 and bf16
constexpr int MASK_MAX_LABEL = 35;      // 3 * 3 * 4 - 1
template <typename X, int NSPLIT> struct ZStage {
  float v[64];
  // zs: channel 0 of this row (a VALID address: callers clamp out-of-range rows); hw: channel pitch in elements
  __device__ __forceinline__ void issue_half(int h, const float* __restrict__ zs, uint32_t hw, int D) {
    if (D == 64) {
#pragma unroll
      for (int k = 0; k < 32; ++k) v[h * 32 + k] = __ldg(zs + (size_t)((uint32_t)(h * 32 + k) * hw));
    } else {
#pragma unroll
      for (int k = 0; k < 32; ++k) v[h * 32 + k] = (h * 32 + k < D) ? __ldg(zs + (size_t)((uint32_t)(h * 32 + k) * hw)) : 0.0f;
    }
  }
  // converts channels [32h, 32h+32) of row m into a_hi (and a_lo); accumulates the squared norm
  // NACC partial sums for the norm (1, 2 or 4): a single FFMA chain serialises the row when the caller has registers
  // to spare (backward: 128 -> 122 us with 4); the push loaders do not (4 sums: 220 -> 284 us per 16 maps)
  template <int NACC = 1>
  __device__ __forceinline__ void convert_half(int h, int m, unsigned char* a_hi, unsigned char* a_lo, float& zn) {
    float zq[NACC];
    zq[0] = zn;
#pragma unroll
    for (int i = 1; i < NACC; ++i) zq[i] = 0.f;
#pragma unroll
    for (int g4 = 0; g4 < 4; ++g4) {
      const int g = h * 4 + g4;
      uint32_t hw_[4], lw_[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float v0 = v[g * 8 + 2 * q], v1 = v[g * 8 + 2 * q + 1];
        hw_[q] = Pack2<X>::pack(v0, v1);
        float2 e = Pack2<X>::unpack(hw_[q]);
        if (NSPLIT == 3) {
          lw_[q] = Pack2<X>::pack(v0 - e.x, v1 - e.y);
          const float2 el = Pack2<X>::unpack(lw_[q]);
          e.x += el.x;
          e.y += el.y;
        }
        zq[q % NACC] = fmaf(e.x, e.x, zq[q % NACC]);
        zq[q % NACC] = fmaf(e.y, e.y, zq[q % NACC]);
      }
      const uint32_t off = (uint32_t)m * 128u + ((((uint32_t)g) ^ ((uint32_t)m & 7u)) << 4);
      *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(hw_[0], hw_[1], hw_[2], hw_[3]);
      if (NSPLIT == 3) *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(lw_[0], lw_[1], lw_[2], lw_[3]);
    }
    if (NACC == 4) zn = (zq[0] + zq[1]) + (zq[2] + zq[3]);
    else if (NACC == 2) zn = zq[0] + zq[1];
    else zn = zq[0];
  }
  // the extra K step of row m: [1, 1, 1, zn_hi, zn_mid, zn_lo, 0 ...]
  static __device__ __forceinline__ void store_norm(int m, unsigned char* a_n, float zn) {
    X h, mm, l;
    split3<X>(zn, h, mm, l);
    const uint32_t one = bits16(Fmt<X>::from(1.0f));
    *reinterpret_cast<uint4*>(a_n + norm_off(m, 0)) = make_uint4(one | (one << 16), one | ((uint32_t)bits16(h) << 16),
                                                                 (uint32_t)bits16(mm) | ((uint32_t)bits16(l) << 16), 0u);
    *reinterpret_cast<uint4*>(a_n + norm_off(m, 8)) = make_uint4(0u, 0u, 0u, 0u);
  }
  // Push scan: the same K step also carries the class mask in its 10 free slots (k = 6..15).  The label value
  // L = class + 1 (0 = void) is written in three digits (radix 3, 3, 4; L <= 35); the prototype side holds a 1 in the
  // slot of each of ITS digits (write_proto_mask), the position side holds MASK_BIG in every slot that is NOT one of
  // its digits.  The products add MASK_BIG once per differing digit, and every single product is exactly 0 for a
  // matching pair, so the accumulator of a matching (prototype, position) pair is untouched.
  static __device__ __forceinline__ void store_norm_mask(int m, unsigned char* a_n, float zn, int L) {
    X h, mm, l;
    split3<X>(zn, h, mm, l);
    const uint32_t one = bits16(Fmt<X>::from(1.0f)), big = bits16(Fmt<X>::from(MASK_BIG));
    if (L < 0 || L > MASK_MAX_LABEL) L = 0;
    const uint32_t d0 = (uint32_t)L % 3u, d1 = ((uint32_t)L / 3u) % 3u, d2 = (uint32_t)L / 9u;
    const uint32_t mk = 0x3FFu & ~((1u << d0) | (8u << d1) | (64u << d2));
    auto sl = [&](int i) -> uint32_t { return ((mk >> i) & 1u) ? big : 0u; };
    *reinterpret_cast<uint4*>(a_n + norm_off(m, 0)) =
        make_uint4(one | (one << 16), one | ((uint32_t)bits16(h) << 16), (uint32_t)bits16(mm) | ((uint32_t)bits16(l) << 16),
                   sl(0) | (sl(1) << 16));
    *reinterpret_cast<uint4*>(a_n + norm_off(m, 8)) =
        make_uint4(sl(2) | (sl(3) << 16), sl(4) | (sl(5) << 16), sl(6) | (sl(7) << 16), sl(8) | (sl(9) << 16));
  }
  // prototype side of the class mask: row r of a no-swizzle norm tile, label value L = class + 1 (1..35)
  static __device__ __forceinline__ void write_proto_mask(unsigned char* b_n, int r, int L) {
    const uint16_t one = bits16(Fmt<X>::from(1.0f));
    const int d0 = L % 3, d1 = (L / 3) % 3, d2 = L / 9;
    *reinterpret_cast<uint16_t*>(b_n + norm_off(r, 6 + d0)) = one;
    *reinterpret_cast<uint16_t*>(b_n + norm_off(r, 9 + d1)) = one;
    *reinterpret_cast<uint16_t*>(b_n + norm_off(r, 12 + d2)) = one;
  }
};

// Pre-pass helper: 16 consecutive channels [kq, kq+16) of operand row j, value scale * src[k] (zero where !live or
// k >= D), into the SW128 K-major tiles t_hi / t_lo as 16-byte pieces (2-byte stores cost one L2 sector write each and
// made the 12-block pre-pass take 5 us).  Returns sum of (0.5 * p_eff)^2 over the 16 channels, p_eff = the value the
// MMA really sees (hi + lo).
template <typename X, int NSPLIT>
__device__ __forceinline__ float pack_row16(const float* __restrict__ src, bool live, int D, float scale, unsigned char* t_hi,
                                            unsigned char* t_lo, int j, int kq) {
  float v[16];
#pragma unroll
  for (int kk = 0; kk < 16; ++kk) v[kk] = (live && kq + kk < D) ? scale * __ldg(src + kq + kk) : 0.0f;
  float part = 0.f;
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    uint32_t hw[4], lw[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float v0 = v[g * 8 + 2 * q], v1 = v[g * 8 + 2 * q + 1];
      hw[q] = Pack2<X>::pack(v0, v1);
      float2 e = Pack2<X>::unpack(hw[q]);
      lw[q] = 0u;
      if (NSPLIT == 3) {
        lw[q] = Pack2<X>::pack(v0 - e.x, v1 - e.y);
        const float2 el = Pack2<X>::unpack(lw[q]);
        e.x += el.x;
        e.y += el.y;
      }
      part = fmaf(0.5f * e.x, 0.5f * e.x, part);
      part = fmaf(0.5f * e.y, 0.5f * e.y, part);
    }
    const uint32_t off = sw128_off(j, kq + 8 * g);
    *reinterpret_cast<uint4*>(t_hi + off) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
    *reinterpret_cast<uint4*>(t_lo + off) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
  }
  return part;
}
// norm K step of prototype row t: [|p|^2 hi, mid, lo, 1, 1, 1, 0 ...] (zeros for a dead row), two 16-byte pieces
template <typename X>
__device__ __forceinline__ void store_proto_norm(unsigned char* t_n, int t, float pn, bool live) {
  X h, m, l;
  split3<X>(pn, h, m, l);
  const uint32_t one = bits16(Fmt<X>::from(1.0f));
  uint4 a = make_uint4(0u, 0u, 0u, 0u);
  if (live) a = make_uint4((uint32_t)bits16(h) | ((uint32_t)bits16(m) << 16), (uint32_t)bits16(l) | (one << 16), one | (one << 16), 0u);
  *reinterpret_cast<uint4*>(t_n + norm_off(t, 0)) = a;
  *reinterpret_cast<uint4*>(t_n + norm_off(t, 8)) = make_uint4(0u, 0u, 0u, 0u);
}

// ---------------------------------------------------------------------------------------------
// Prototype chunks.  The prototypes of every scale are cut into chunks of <= CN rows; the pre-pass writes
// one packed image per chunk that the main kernels bulk-copy straight into shared memory:
//   [Bhi  CN x 128 B  SW128]  rows j: -2 * p_(p0+j)            (hi part)
//   [Bn   CN x  32 B  noswz]  k=0..2: |p|^2 split in three; k=3..5: 1; rest 0
//   [Blo  CN x 128 B  SW128]  lo part (zeros in the single-term modes)
//   (hi | norm | lo in this order: what a kernel needs per chunk is ONE contiguous bulk copy, 10 KB single-term, 18 KB split)
//   [W2   KP x 128 B  SW128]  rows k: ln2 * W[k, p0+j]  ('log'; W itself for 'linear'); only with a last layer
//   [W2lo KP x 128 B  SW128]  lo part of W2 (used by the split modes)
// ---------------------------------------------------------------------------------------------
constexpr int CN = 64;          // prototypes per chunk (MMA N of the distance GEMM, K of the logits GEMM)
constexpr int MAX_CHUNKS = 160;
constexpr int TILE_M = 128;     // positions per tile (MMA M)

struct ChunkTable {
  int n_chunks;
  int kp;                       // classes padded to a multiple of 16 (0: no last layer)
  int img_bytes;                // bytes of one packed chunk image
  int p0[MAX_CHUNKS];
  short n[MAX_CHUNKS];
  short scale[MAX_CHUNKS];
};

__host__ __device__ inline int chunk_img_bytes(int kp) { return 2 * CN * 128 + CN * 32 + 2 * kp * 128; }
constexpr int OFF_BHI = 0, OFF_BN = CN * 128, OFF_BLO = CN * 128 + CN * 32, OFF_W2 = 2 * CN * 128 + CN * 32;

}  // namespace tc
}  // namespace sps
