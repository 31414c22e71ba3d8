// Shared device/host helpers for libscaleproto_sm100.so (sm_100a only).
#pragma once
#include <utility>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/scaleproto_b200.h"

namespace sps {

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define SPS_CUDA_TRY(expr)                                   \
  do {                                                       \
    cudaError_t e__ = (expr);                                \
    if (e__ != cudaSuccess) return sps::cuda_fail(e__, #expr); \
  } while (0)

#define SPS_REQUIRE(cond, status, ...) \
  do {                                 \
    if (!(cond)) {                     \
      sps::set_error(__VA_ARGS__);     \
      return (status);                 \
    }                                  \
  } while (0)

// Kernel-side copy of the shape (passed by value).
struct HeadShape {
  int B, HW, S, D, C, P, K;
  long long N;  // B*HW
  int scale_lo[SPS_MAX_SCALES + 1];
  int act_kind;
  float eps;
  int z_dtype;   // SPS_Z_*
  int z_layout;  // SPS_Z_NCHW / SPS_Z_NHWC
  int flags;     // SPS_FLAG_*
};

inline HeadShape make_shape(const SpsHeadDesc* d) {
  HeadShape s;
  s.B = d->B;
  s.HW = d->H * d->W;
  s.S = d->S;
  s.D = d->D;
  s.C = d->S * d->D;
  s.P = d->P;
  s.K = d->K;
  s.N = (long long)d->B * s.HW;
  for (int i = 0; i <= SPS_MAX_SCALES; ++i) s.scale_lo[i] = (i <= d->S) ? d->scale_lo[i] : d->P;
  s.act_kind = d->act_kind;
  s.eps = d->eps;
  s.z_dtype = d->z_dtype;
  s.z_layout = d->z_layout;
  s.flags = d->flags;
  return s;
}

int validate_desc(const SpsHeadDesc* d);
int sm_count();
void count_launch(int n = 1);
void note_fp32_reroute();

// Programmatic dependent launch: the kernel may start while the previous kernel of the stream is still running (that
// kernel calls pdl_launch_dependents() early); every thread that reads what the previous kernel wrote calls
// pdl_wait() first, which blocks until that kernel has completed and its writes are visible.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// distance -> similarity (model_multiscale.py:324-330) and its derivative.
__device__ __forceinline__ float act_of(float d, int kind, float eps) {
  return kind == SPS_ACT_LOG ? __logf((d + 1.0f) / (d + eps)) : -d;
}
__device__ __forceinline__ float act_exact(float d, int kind, float eps) {
  return kind == SPS_ACT_LOG ? logf((d + 1.0f) / (d + eps)) : -d;
}
__device__ __forceinline__ float dact_dd(float d, int kind, float eps) {
  // d/dd log((d+1)/(d+eps)) = 1/(d+1) - 1/(d+eps) = -(1-eps)/((d+1)(d+eps))
  return kind == SPS_ACT_LOG ? -(1.0f - eps) / ((d + 1.0f) * (d + eps)) : -1.0f;
}

// launchers implemented in the .cu files
int launch_head_fwd_simt(const HeadShape& s, const float* z, const float* protos, const float* last_w,
                         float* logits, float* dist, float* act, cudaStream_t st);
size_t head_fwd_tc_workspace(const HeadShape& s);
bool head_fwd_tc_supported(const HeadShape& s, int math, bool with_last_layer, bool dense_act);
int launch_head_fwd_tc(const HeadShape& s, int math, const float* z, const float* protos, const float* last_w,
                       float* logits, float* dist, float* act, void* ws, size_t ws_bytes, cudaStream_t st);
// TMA-fed kernels (16-bit channels-last maps, or the hi / lo re-lay of an fp32 map)
bool head_fwd_tma_supported(const HeadShape& s, int math, bool with_last_layer, bool dense_act, bool zlo);
int launch_head_fwd_tma(const HeadShape& s, int math, const void* z_hi, const void* z_lo, const float* protos,
                        const float* last_w, float* logits, float* dist, float* act, void* ws, size_t ws_bytes,
                        cudaStream_t st);
size_t z_pack_bytes(const HeadShape& s, int math);
int launch_z_pack(const HeadShape& s, int math, const float* z, void* out, cudaStream_t st);
// hi / lo parts inside a re-lay buffer of z_pack_bytes (lo = nullptr in the single-term modes)
void z_pack_parts(const HeadShape& s, int math, const void* area, const void** hi, const void** lo);
bool head_bwd_tc_supported(const HeadShape& s, int math, const float* last_w, const float* dlogits, const float* ddist,
                           const float* dact);
size_t head_bwd_tc_workspace(const HeadShape& s);
// zmode: 0 = z is the fp32 NCHW map; 1 = z is a 16-bit channels-last map (dz likewise); 2 = z / z_lo are the hi / lo
// re-lay of an fp32 map (dz fp32 NCHW)
int launch_head_bwd_tc(const HeadShape& s, int math, int zmode, const void* z, const void* z_lo, const float* protos,
                       const float* last_w, const float* dlogits, const float* ddist, void* dz, float* dprotos,
                       float* dlast_w, void* ws, size_t ws_bytes, cudaStream_t st);
// one-pass tcgen05 backward for more than 32 classes (tile-stationary dz kernel + chunk-stationary dp / dW kernel)
bool head_bwd_wide_supported(const HeadShape& s, int math, const float* last_w, const float* dlogits, const float* ddist,
                             const float* dact);
size_t head_bwd_wide_workspace(const HeadShape& s);
int launch_head_bwd_wide(const HeadShape& s, int math, const void* z16, bool dz16, const float* protos, const float* last_w,
                         const float* dlogits, void* dz, float* dprotos, float* dlast_w, void* ws, size_t ws_bytes, cudaStream_t st);
int launch_head_bwd_simt(const HeadShape& s, int flags, const float* z, const float* protos, const float* last_w,
                         const float* dlogits, const float* ddist, const float* dact, float* dz, float* dprotos,
                         float* dlast_w, void* ws, size_t ws_bytes, cudaStream_t st);
size_t head_bwd_simt_workspace(const HeadShape& s, int flags);
int launch_group_fwd(const HeadShape& s, const SpsGroupDesc& g, const float* act, float* groups, float* logits,
                     cudaStream_t st);
int launch_group_bwd(const HeadShape& s, const SpsGroupDesc& g, const float* act, const float* groups_in, const float* dlogits,
                     const float* dgroups, float* dact, float* dG, float* dWg, float* ws, cudaStream_t st);
int launch_push_scan_simt(const HeadShape& s, const float* z, const float* protos, const int16_t* labels,
                          const int32_t* proto_class, unsigned long long* keys, cudaStream_t st);
int launch_push_fill(unsigned long long* keys, long long count, cudaStream_t st);
bool head_fwd_f32tma_supported(const HeadShape& s, int math, bool with_last_layer, bool dense_act);
int launch_head_fwd_f32tma(const HeadShape& s, int math, const float* z, const float* protos, const float* last_w, float* logits,
                           float* dist, float* act, void* ws, size_t ws_bytes, cudaStream_t st);
int launch_push_scan_tc(const HeadShape& s, int math, const void* z, bool z16, const float* protos, const int16_t* labels,
                        const int32_t* proto_class, unsigned long long* keys, void* ws, size_t ws_bytes, cudaStream_t st);
int launch_push_reduce(const unsigned long long* keys, int n_maps, int P, uint32_t map_base,
                       unsigned long long* best, cudaStream_t st);
int launch_push_gather(const HeadShape& s, const void* z, const unsigned long long* keys,
                       const unsigned long long* best, int n_maps, uint32_t map_base, const int32_t* proto_scale,
                       float* out, cudaStream_t st);

size_t kld_workspace(const HeadShape& s, int Q);
int launch_kld_fwd(const HeadShape& s, const SpsKldDesc& k, const float* dist, const int16_t* labels, void* ws,
                   float* loss, cudaStream_t st);
int launch_kld_bwd(const HeadShape& s, const SpsKldDesc& k, const float* dist, void* ws, const float* dloss,
                   float* ddist, cudaStream_t st);

size_t kld_feat_workspace(const HeadShape& s, const SpsKldDesc& k);
int launch_kld_feat_fwd(const HeadShape& s, const SpsKldDesc& k, const float* z, const float* protos,
                        const int16_t* labels, void* ws, float* loss, cudaStream_t st);
int launch_kld_feat_bwd(const HeadShape& s, const SpsKldDesc& k, const float* z, const float* protos,
                        const int16_t* labels, void* ws, const float* dloss, float* dz, float* dprotos,
                        cudaStream_t st);

int launch_push_topk(const unsigned long long* keys, int n_maps, int P, uint32_t map_base, int k,
                     unsigned long long* top_key, int32_t* top_pix, cudaStream_t st);
int launch_upsample_argext(const float* maps, int B, int C, int h, int w, int H, int W, int argmin, int32_t* out_idx,
                           float* out_val, cudaStream_t st);

}  // namespace sps
