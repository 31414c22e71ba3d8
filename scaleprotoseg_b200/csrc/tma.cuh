// TMA (cp.async.bulk.tensor) wrappers and the host-side tensor-map encoder used by the 16-bit channels-last kernels.
//
// A 16-bit NHWC feature map is the 2-D row-major array [N positions][C channels] (row pitch C * 2 bytes, a multiple of
// 16 — a legal tensor map, unlike the fp32 NCHW map whose 4 * H * W pitch is not 16-byte aligned for 65x65 or 129x257).
// A box of {64 channels, 128 positions} with the 128-byte swizzle lands in shared memory in exactly the K-major
// SWIZZLE_128B layout the tcgen05 descriptors of tc_common.cuh describe: no conversion work on the SM.
#pragma once
#include <cuda.h>

#include "tc_common.cuh"

namespace sps {
namespace tc {

// 2-D tile global -> shared, completion counted on an mbarrier (UTMALDG).  c0 = element index along the contiguous
// dimension (channel), c1 = row (position).  Rows / columns outside the tensor read as zeros.
__device__ __forceinline__ void tma_load_2d(uint32_t dst_smem, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst_smem),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
// 3-D box global -> shared (UTMALDG), completion on an mbarrier
__device__ __forceinline__ void tma_load_3d(uint32_t dst_smem, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst_smem),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// 2-D tile shared -> global (UTMASTG), bulk-group completion.  Elements outside the tensor are not written.
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, int c0, int c1, uint32_t src_smem) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(reinterpret_cast<uint64_t>(map)),
               "r"(src_smem), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// waits until the shared-memory source of all but the `N` most recent bulk groups has been read
template <int N> __device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void tma_store_wait() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

// ZMODE of the kernels that can take either source: where the feature tile comes from.
//   ZM_LDG    fp32 NCHW map, converted by loader warps (128-byte warp loads along the position axis)
//   ZM_TMA16  16-bit channels-last map (the user's fp16 / bf16 tensor): TMA box {64 ch, 128 pos} -> SW128 tile; dz is
//             written in the same 16-bit channels-last layout
//   ZM_TMAHL  hi / lo 16-bit re-lay of an fp32 NCHW map (z_pack_kernel): TMA for both parts; dz stays fp32 NCHW
enum { ZM_LDG = 0, ZM_TMA16 = 1, ZM_TMAHL = 2 };

// squared norm of row r of a landed K-major SW128 tile (hi [+ lo]): the value the MMA really sees
template <typename X, bool ZLO>
__device__ __forceinline__ float tile_row_norm(const unsigned char* a_hi, const unsigned char* a_lo, int r) {
  float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#pragma unroll
  for (int g = 0; g < 8; ++g) {
    const uint32_t off = (uint32_t)r * 128u + ((((uint32_t)g) ^ ((uint32_t)r & 7u)) << 4);
    const uint4 h = *reinterpret_cast<const uint4*>(a_hi + off);
    const uint32_t hw[4] = {h.x, h.y, h.z, h.w};
    uint32_t lw[4] = {0u, 0u, 0u, 0u};
    if (ZLO) {
      const uint4 l = *reinterpret_cast<const uint4*>(a_lo + off);
      lw[0] = l.x, lw[1] = l.y, lw[2] = l.z, lw[3] = l.w;
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float2 v = Pack2<X>::unpack(hw[e]);
      if (ZLO) {
        const float2 vl = Pack2<X>::unpack(lw[e]);
        v.x += vl.x;
        v.y += vl.y;
      }
      if (e & 1) {
        q2 = fmaf(v.x, v.x, q2);
        q3 = fmaf(v.y, v.y, q3);
      } else {
        q0 = fmaf(v.x, v.x, q0);
        q1 = fmaf(v.y, v.y, q1);
      }
    }
  }
  return (q0 + q1) + (q2 + q3);
}

}  // namespace tc

// Host: tensor map over a row-major [rows][cols] array of 16-bit elements, box = {box_cols, box_rows}.
// swizzle_bytes: 128 (box_cols * 2 == 128), 64 (box_cols * 2 == 64) or 0.  Returns an SpsStatus.
int make_tmap_2d_16bit(CUtensorMap* out, const void* base, bool bf16, long long rows, long long cols, int box_rows,
                       int box_cols, int swizzle_bytes);

// fp32 NCHW maps by TMA.  A channel row of H*W floats starts at any multiple of 4 bytes when H*W is odd, and a tensor map
// wants 16-byte strides - but every FOURTH channel row is 16*H*W bytes further on.  So the map is described four times,
// once per channel residue r = c mod 4: base = the address of channel r rounded DOWN to 16 bytes, dimensions
// {H*W + shift_r, C/4, B} with strides {16*H*W, 4*C*H*W} bytes, where shift_r = the floats between the rounded base and
// the real start of the row.  The start of a box must be 16-byte aligned as well (an x coordinate that is not a multiple
// of 4 floats faults: profiles/tools/tma_nchw_probe2.cu), so a box is taken at x = pixel (a multiple of 4) with 4 extra
// floats: {px + 4, rows, 1}; pixel p of channel r + 4j is then element p + shift_r of box row j, zeros past the end of
// the row.  Needs C % 4 == 0 and a 4-byte aligned map.
struct TmapNchwF32 {
  CUtensorMap map[4];
  int shift[4];
};
int make_tmap_nchw_f32(TmapNchwF32* out, const float* z, long long B, long long C, long long HW, int px, int box_rows);


}  // namespace sps
