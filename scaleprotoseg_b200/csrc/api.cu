// extern "C" surface of libscaleproto_sm100.so: argument checks, math-mode routing, error text.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "common.cuh"

namespace sps {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return SPS_ERR_CUDA;
}

static std::atomic<unsigned long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }
// calls that asked for a tensor-core math mode but were served by the exact fp32 CUDA-core kernels because the shape
// does not fit the tcgen05 kernels: reported (sps_fp32_reroute_count), never hidden
static std::atomic<unsigned long long> g_reroutes{0};
void note_fp32_reroute() { g_reroutes.fetch_add(1ull, std::memory_order_relaxed); }

int sm_count() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      cached = 148;
  }
  return cached;
}

int validate_desc(const SpsHeadDesc* d) {
  SPS_REQUIRE(d != nullptr, SPS_ERR_BAD_ARG, "desc is NULL");
  SPS_REQUIRE(d->B >= 0 && d->H >= 0 && d->W >= 0, SPS_ERR_BAD_ARG, "negative map shape B=%d H=%d W=%d", d->B, d->H, d->W);
  SPS_REQUIRE(d->S >= 1 && d->S <= SPS_MAX_SCALES, SPS_ERR_BAD_ARG, "S=%d outside [1,%d]", d->S, SPS_MAX_SCALES);
  SPS_REQUIRE(d->D >= 1, SPS_ERR_BAD_ARG, "D=%d", d->D);
  SPS_REQUIRE(d->P >= 0 && d->K >= 1, SPS_ERR_BAD_ARG, "P=%d K=%d", d->P, d->K);
  SPS_REQUIRE(d->scale_lo[0] == 0 && d->scale_lo[d->S] == d->P, SPS_ERR_BAD_ARG,
              "scale_lo must start at 0 and end at P (got %d..%d, P=%d)", d->scale_lo[0], d->scale_lo[d->S], d->P);
  for (int s = 0; s < d->S; ++s)
    SPS_REQUIRE(d->scale_lo[s] <= d->scale_lo[s + 1], SPS_ERR_BAD_ARG, "scale_lo not monotone at scale %d", s);
  SPS_REQUIRE(d->act_kind == SPS_ACT_LOG || d->act_kind == SPS_ACT_LINEAR, SPS_ERR_BAD_ARG, "act_kind=%d", d->act_kind);
  SPS_REQUIRE(d->math >= SPS_MATH_FP32 && d->math <= SPS_MATH_F16X3, SPS_ERR_BAD_ARG, "math=%d", d->math);
  SPS_REQUIRE((long long)d->H * d->W < (1LL << 31), SPS_ERR_UNSUPPORTED, "H*W too large");
  if (d->z_dtype == SPS_Z_F32) {
    SPS_REQUIRE(d->z_layout == SPS_Z_NCHW, SPS_ERR_UNSUPPORTED, "fp32 feature maps must be NCHW (z_layout=%d)", d->z_layout);
  } else {
    SPS_REQUIRE(d->z_dtype == SPS_Z_F16 || d->z_dtype == SPS_Z_BF16, SPS_ERR_BAD_ARG, "z_dtype=%d", d->z_dtype);
    SPS_REQUIRE(d->z_layout == SPS_Z_NHWC, SPS_ERR_UNSUPPORTED, "16-bit feature maps must be channels-last (z_layout=%d)", d->z_layout);
    const bool f16 = d->z_dtype == SPS_Z_F16;
    SPS_REQUIRE(f16 ? (d->math == SPS_MATH_F16 || d->math == SPS_MATH_F16X3) : (d->math == SPS_MATH_BF16 || d->math == SPS_MATH_BF16X3),
                SPS_ERR_UNSUPPORTED, "math=%d does not match the %s feature maps (fp16 maps: F16 / F16X3, bf16 maps: BF16 / BF16X3)", d->math,
                f16 ? "fp16" : "bf16");
    SPS_REQUIRE(d->D == 64, SPS_ERR_UNSUPPORTED, "16-bit channels-last feature maps: D=%d, only 64 channels per scale are supported", d->D);
  }
  return SPS_OK;
}

}  // namespace sps

using namespace sps;

extern "C" {

int sps_version(void) { return SPS_ABI_VERSION; }

const char* sps_last_error(void) { return g_err; }

uint64_t sps_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

uint64_t sps_fp32_reroute_count(void) { return g_reroutes.load(std::memory_order_relaxed); }

// SPS_NO_F32TMA=1 in the environment keeps fp32 maps on the loader-fed forward kernel (A/B measurements only)
static bool getenv_no_f32tma() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("SPS_NO_F32TMA");
    v = (e != nullptr && e[0] == '1') ? 1 : 0;
  }
  return v == 1;
}
static bool use_zpack(const SpsHeadDesc* d) {
  return d->z_dtype == SPS_Z_F32 && d->math != SPS_MATH_FP32 && (d->flags & (SPS_FLAG_ZPACK | SPS_FLAG_ZRELAYED)) != 0 && d->D == 64;
}
static bool is_split(int math) { return math == SPS_MATH_F16X3 || math == SPS_MATH_BF16X3; }
static size_t align1k(size_t v) { return (v + 1023) / 1024 * 1024; }

static const int WIDE_K = 32;   // more classes than this: the one-pass wide backward (head_bwd_wide.cu); the forward may
                                // need the re-lay route as well (3-term operands + a wide last layer do not fit the loader kernel)
static int single_term(int math) { return (math == SPS_MATH_BF16 || math == SPS_MATH_BF16X3) ? SPS_MATH_BF16 : SPS_MATH_F16; }

size_t sps_head_fwd_workspace_bytes(const SpsHeadDesc* desc) {
  if (validate_desc(desc) != SPS_OK || desc->math == SPS_MATH_FP32) return 0;
  const HeadShape s = make_shape(desc);
  size_t n = align1k(head_fwd_tc_workspace(s));
  if (desc->z_dtype == SPS_Z_F32 && !(desc->flags & SPS_FLAG_ZRELAYED) && desc->D == 64 && (use_zpack(desc) || s.K > WIDE_K))
    n += z_pack_bytes(s, desc->math);
  return n;
}

size_t sps_workspace_bytes(const SpsHeadDesc* desc) {
  if (validate_desc(desc) != SPS_OK) return 0;
  HeadShape s = make_shape(desc);
  size_t a = desc->z_dtype == SPS_Z_F32 ? head_bwd_simt_workspace(s, desc->flags) : 0;
  size_t b = desc->math != SPS_MATH_FP32 ? head_bwd_tc_workspace(s) : 0;
  size_t w = (desc->math != SPS_MATH_FP32 && s.K > WIDE_K && s.D == 64) ? head_bwd_wide_workspace(s) : 0;
  size_t n = align1k(a > b ? (a > w ? a : w) : (b > w ? b : w));
  if (desc->z_dtype == SPS_Z_F32 && !(desc->flags & SPS_FLAG_ZRELAYED)) {
    size_t r = use_zpack(desc) ? z_pack_bytes(s, desc->math) : 0;
    if (w > 0) {   // fp32 maps and a wide last layer: the map is re-laid (one 16-bit term) for the TMA-fed kernels
      const size_t r1 = z_pack_bytes(s, single_term(desc->math));
      r = r > r1 ? r : r1;
    }
    n += r;
  }
  return n;
}

size_t sps_z_relay_bytes(const SpsHeadDesc* desc) {
  if (validate_desc(desc) != SPS_OK || desc->math == SPS_MATH_FP32 || desc->z_dtype != SPS_Z_F32) return 0;
  return z_pack_bytes(make_shape(desc), desc->math);
}

int sps_z_relay(const SpsHeadDesc* desc, const float* z, void* out, size_t out_bytes, void* stream) {
  int rc = validate_desc(desc);
  if (rc != SPS_OK) return rc;
  SPS_REQUIRE(desc->z_dtype == SPS_Z_F32 && desc->math != SPS_MATH_FP32, SPS_ERR_BAD_ARG,
              "z_relay: needs an fp32 NCHW map and a tensor-core math mode");
  HeadShape s = make_shape(desc);
  if (s.N == 0) return SPS_OK;
  SPS_REQUIRE(z != nullptr && out != nullptr && out_bytes >= z_pack_bytes(s, desc->math), SPS_ERR_WORKSPACE,
              "z_relay: need %zu bytes, got %zu", z_pack_bytes(s, desc->math), out_bytes);
  return launch_z_pack(s, desc->math, z, out, reinterpret_cast<cudaStream_t>(stream));
}

int sps_head_fwd(const SpsHeadDesc* desc, const void* z, const float* protos, const float* last_w, float* logits,
                 float* dist, float* act, void* workspace, size_t workspace_bytes, void* stream) {
  int rc = validate_desc(desc);
  if (rc != SPS_OK) return rc;
  HeadShape s = make_shape(desc);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (s.N == 0 || s.P == 0) {
    // nothing to contract over: the logits of an empty prototype set are zeros (torch: (N,0) @ (0,K))
    if (logits != nullptr && s.N > 0) SPS_CUDA_TRY(cudaMemsetAsync(logits, 0, sizeof(float) * (size_t)s.N * s.K, st));
    return SPS_OK;
  }
  SPS_REQUIRE(z != nullptr && protos != nullptr, SPS_ERR_BAD_ARG, "z/protos NULL");
  SPS_REQUIRE((last_w == nullptr) == (logits == nullptr), SPS_ERR_BAD_ARG, "last_w and logits go together");
  SPS_REQUIRE(last_w != nullptr || dist != nullptr || act != nullptr, SPS_ERR_BAD_ARG, "no output requested");
  if (desc->z_dtype != SPS_Z_F32) {   // 16-bit channels-last map: TMA-fed kernel or nothing (no silent re-route)
    SPS_REQUIRE(head_fwd_tma_supported(s, desc->math, last_w != nullptr, act != nullptr, false), SPS_ERR_UNSUPPORTED,
                "16-bit channels-last forward: this shape does not fit the tensor-core kernel (K=%d, P=%d)", s.K, s.P);
    return launch_head_fwd_tma(s, desc->math, z, nullptr, protos, last_w, logits, dist, act, workspace, workspace_bytes, st);
  }
  const float* zf = reinterpret_cast<const float*>(z);
  const bool tma_ok = desc->math != SPS_MATH_FP32 && head_fwd_tma_supported(s, desc->math, last_w != nullptr, act != nullptr, is_split(desc->math));
  const bool ldg_ok = desc->math != SPS_MATH_FP32 && head_fwd_tc_supported(s, desc->math, last_w != nullptr, act != nullptr);
  // re-lay route: asked for (SPS_FLAG_ZPACK / ZRELAYED), or the loader-fed kernel does not fit this shape (3-term operands
  // with a wide last layer) while the TMA-fed one does
  if (tma_ok && (use_zpack(desc) || (!ldg_ok && s.K > WIDE_K))) {
    const size_t img_bytes = align1k(head_fwd_tc_workspace(s));
    const void *hi, *lo;
    if (desc->flags & SPS_FLAG_ZRELAYED) {
      z_pack_parts(s, desc->math, z, &hi, &lo);
    } else {
      SPS_REQUIRE(workspace != nullptr && workspace_bytes >= img_bytes + z_pack_bytes(s, desc->math), SPS_ERR_WORKSPACE,
                  "head_fwd workspace: need %zu bytes, got %zu", img_bytes + z_pack_bytes(s, desc->math), workspace_bytes);
      void* area = reinterpret_cast<unsigned char*>(workspace) + img_bytes;
      rc = launch_z_pack(s, desc->math, zf, area, st);
      if (rc != SPS_OK) return rc;
      z_pack_parts(s, desc->math, area, &hi, &lo);
    }
    return launch_head_fwd_tma(s, desc->math, hi, lo, protos, last_w, logits, dist, act, workspace, img_bytes, st);
  }
  SPS_REQUIRE(!(desc->flags & SPS_FLAG_ZRELAYED), SPS_ERR_UNSUPPORTED, "SPS_FLAG_ZRELAYED: this shape has no TMA-fed forward");
  // fp32 NCHW maps by TMA (four residue tensor maps), converted on the SM into a TMEM operand: the default where it fits
  if (desc->math != SPS_MATH_FP32 && !getenv_no_f32tma() && head_fwd_f32tma_supported(s, desc->math, last_w != nullptr, act != nullptr))
    return launch_head_fwd_f32tma(s, desc->math, zf, protos, last_w, logits, dist, act, workspace, workspace_bytes, st);
  if (desc->math == SPS_MATH_FP32 || !head_fwd_tc_supported(s, desc->math, last_w != nullptr, act != nullptr)) {
    if (desc->math != SPS_MATH_FP32) note_fp32_reroute();
    return launch_head_fwd_simt(s, zf, protos, last_w, logits, dist, act, st);
  }
  return launch_head_fwd_tc(s, desc->math, zf, protos, last_w, logits, dist, act, workspace, workspace_bytes, st);
}

// ---- allocation-free forward (no counterpart in the reference; see the header) ---------------------------------------
// The scratch of the forward (the packed prototype / last-layer chunk images the pre-pass writes and the main kernel
// reads) lives in this file-scope device array instead of a caller workspace: every GPU that loads the library holds
// SPS_STATIC_SCRATCH_BYTES of it until exit. In host code `g_fwd_scratch` is only a handle for the runtime, not a device
// address: the address is asked for once per device, with that device current, and cached.
static __device__ __align__(256) unsigned char g_fwd_scratch[SPS_STATIC_SCRATCH_BYTES];

static const int MAX_DEVICES = 64;
static void* g_fwd_scratch_addr[MAX_DEVICES] = {};
static std::mutex g_fwd_scratch_mutex;

static int fwd_scratch_address(void** out) {
  int dev = -1;
  SPS_CUDA_TRY(cudaGetDevice(&dev));
  SPS_REQUIRE(dev >= 0 && dev < MAX_DEVICES, SPS_ERR_UNSUPPORTED, "device ordinal %d outside [0,%d)", dev, MAX_DEVICES);
  std::lock_guard<std::mutex> lock(g_fwd_scratch_mutex);
  if (g_fwd_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    SPS_CUDA_TRY(cudaGetSymbolAddress(&p, g_fwd_scratch));   // the copy of the array on the CURRENT device
    SPS_REQUIRE(p != nullptr, SPS_ERR_CUDA, "cudaGetSymbolAddress returned NULL for the static scratch on device %d", dev);
    g_fwd_scratch_addr[dev] = p;
  }
  *out = g_fwd_scratch_addr[dev];
  return SPS_OK;
}

int sps_head_fwd_static(const SpsHeadDesc* desc, const void* z, const float* protos, const float* last_w, float* logits,
                        float* dist, float* act, void* stream) {
  int rc = validate_desc(desc);
  if (rc != SPS_OK) return rc;
  const size_t need = sps_head_fwd_workspace_bytes(desc);
  SPS_REQUIRE(need <= (size_t)SPS_STATIC_SCRATCH_BYTES, SPS_ERR_WORKSPACE,
              "head_fwd_static: this shape needs %zu bytes of scratch, the static array holds %zu (B=%d P=%d K=%d%s); "
              "use sps_head_fwd with a caller workspace",
              need, (size_t)SPS_STATIC_SCRATCH_BYTES, desc->B, desc->P, desc->K,
              (desc->z_dtype == SPS_Z_F32 && need > (size_t)(1u << 20)) ? ", includes the 16-bit re-lay of the fp32 map" : "");
  void* scratch = nullptr;
  if (need > 0) {
    rc = fwd_scratch_address(&scratch);
    if (rc != SPS_OK) return rc;
  }
  // nothing in the scratch has to start at zero: the pre-pass rewrites every byte the main kernel reads, so what the
  // previous call left behind is never seen
  return sps_head_fwd(desc, z, protos, last_w, logits, dist, act, scratch, need > 0 ? (size_t)SPS_STATIC_SCRATCH_BYTES : 0, stream);
}

int sps_head_bwd(const SpsHeadDesc* desc, const void* z, const float* protos, const float* last_w,
                 const float* dlogits, const float* ddist, const float* dact, void* dz, float* dprotos,
                 float* dlast_w, void* workspace, size_t workspace_bytes, void* stream) {
  int rc = validate_desc(desc);
  if (rc != SPS_OK) return rc;
  HeadShape s = make_shape(desc);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool z16 = desc->z_dtype != SPS_Z_F32;
  if (s.N == 0 || s.P == 0 || (dlogits == nullptr && ddist == nullptr && dact == nullptr)) {
    // empty batch / no prototypes / no upstream gradient: every requested gradient is zero (never uninitialised memory)
    if (dz != nullptr && s.N > 0) SPS_CUDA_TRY(cudaMemsetAsync(dz, 0, (z16 ? 2 : 4) * (size_t)s.N * s.C, st));
    if (dprotos != nullptr && s.P > 0) SPS_CUDA_TRY(cudaMemsetAsync(dprotos, 0, sizeof(float) * (size_t)s.P * s.D, st));
    if (dlast_w != nullptr && s.P > 0) SPS_CUDA_TRY(cudaMemsetAsync(dlast_w, 0, sizeof(float) * (size_t)s.K * s.P, st));
    return SPS_OK;
  }
  SPS_REQUIRE(z != nullptr && protos != nullptr, SPS_ERR_BAD_ARG, "z/protos NULL");
  SPS_REQUIRE(dlogits == nullptr || last_w != nullptr, SPS_ERR_BAD_ARG, "dlogits given without last_w");
  // tensor-core math mode with a last layer (upstream gradient = dlogits, optionally + ddist from the KLD loss):
  // tcgen05 backward; a gradient through the activation map (dact), no dlogits, or SPS_MATH_FP32: exact-operand
  // CUDA-core kernels
  const bool tc_ok = head_bwd_tc_supported(s, desc->math, last_w, dlogits, ddist, dact);
  if (s.K > WIDE_K && head_bwd_wide_supported(s, desc->math, last_w, dlogits, ddist, dact)) {
    // more than 32 classes: one pass over the distances (tile-stationary dz kernel + chunk-stationary dp / dW kernel), TMA-fed
    if (z16)
      return launch_head_bwd_wide(s, desc->math, z, true, protos, last_w, dlogits, dz, dprotos, dlast_w, workspace, workspace_bytes, st);
    const size_t base_bytes = align1k(head_bwd_wide_workspace(s));
    const void *hi, *lo;
    if (desc->flags & SPS_FLAG_ZRELAYED) {
      z_pack_parts(s, desc->math, z, &hi, &lo);      // the hi part of any re-lay is the map rounded to one 16-bit term
    } else {
      const int m1 = single_term(desc->math);
      SPS_REQUIRE(workspace != nullptr && workspace_bytes >= base_bytes + z_pack_bytes(s, m1), SPS_ERR_WORKSPACE,
                  "head_bwd workspace: need %zu bytes, got %zu", base_bytes + z_pack_bytes(s, m1), workspace_bytes);
      void* area = reinterpret_cast<unsigned char*>(workspace) + base_bytes;
      rc = launch_z_pack(s, m1, reinterpret_cast<const float*>(z), area, st);
      if (rc != SPS_OK) return rc;
      z_pack_parts(s, m1, area, &hi, &lo);
    }
    return launch_head_bwd_wide(s, desc->math, hi, false, protos, last_w, dlogits, dz, dprotos, dlast_w, workspace, base_bytes, st);
  }
  if (z16) {
    SPS_REQUIRE(tc_ok, SPS_ERR_UNSUPPORTED,
                "16-bit channels-last backward needs last_w + dlogits (no gradient through the activation map) and K <= 256");
    return launch_head_bwd_tc(s, desc->math, 1, z, nullptr, protos, last_w, dlogits, ddist, dz, dprotos, dlast_w, workspace,
                              workspace_bytes, st);
  }
  const float* zf = reinterpret_cast<const float*>(z);
  if (tc_ok && use_zpack(desc)) {
    const size_t base_bytes = align1k(head_bwd_tc_workspace(s));
    const void *hi, *lo;
    if (desc->flags & SPS_FLAG_ZRELAYED) {
      z_pack_parts(s, desc->math, z, &hi, &lo);
    } else {
      SPS_REQUIRE(workspace != nullptr && workspace_bytes >= base_bytes + z_pack_bytes(s, desc->math), SPS_ERR_WORKSPACE,
                  "head_bwd workspace: need %zu bytes, got %zu", base_bytes + z_pack_bytes(s, desc->math), workspace_bytes);
      void* area = reinterpret_cast<unsigned char*>(workspace) + base_bytes;
      rc = launch_z_pack(s, desc->math, zf, area, st);
      if (rc != SPS_OK) return rc;
      z_pack_parts(s, desc->math, area, &hi, &lo);
    }
    return launch_head_bwd_tc(s, desc->math, 2, hi, lo, protos, last_w, dlogits, ddist, dz, dprotos, dlast_w, workspace, base_bytes, st);
  }
  SPS_REQUIRE(!(desc->flags & SPS_FLAG_ZRELAYED), SPS_ERR_UNSUPPORTED, "SPS_FLAG_ZRELAYED: this call has no TMA-fed backward");
  if (tc_ok)
    return launch_head_bwd_tc(s, desc->math, 0, zf, nullptr, protos, last_w, dlogits, ddist, dz, dprotos, dlast_w, workspace,
                              workspace_bytes, st);
  if (desc->math != SPS_MATH_FP32 && dact == nullptr && dlogits != nullptr) note_fp32_reroute();
  return launch_head_bwd_simt(s, desc->flags, zf, protos, last_w, dlogits, ddist, dact, reinterpret_cast<float*>(dz), dprotos, dlast_w,
                              workspace, workspace_bytes, st);
}

static int check_group(const SpsGroupDesc* g) {
  SPS_REQUIRE(g != nullptr, SPS_ERR_BAD_ARG, "group desc NULL");
  SPS_REQUIRE(g->Kc >= 0 && g->J >= 1 && g->J <= 8, SPS_ERR_UNSUPPORTED, "Kc=%d J=%d (J must be 1..8)", g->Kc, g->J);
  SPS_REQUIRE(g->proto_slot && g->proto_col && g->class_off && g->members && g->G && g->Wg, SPS_ERR_BAD_ARG,
              "group desc has NULL tables");
  return SPS_OK;
}

int sps_group_fwd(const SpsHeadDesc* desc, const SpsGroupDesc* g, const float* act, float* groups, float* logits,
                  void* stream) {
  int rc = validate_desc(desc);
  if (rc != SPS_OK) return rc;
  rc = check_group(g);
  if (rc != SPS_OK) return rc;
  SPS_REQUIRE(act != nullptr, SPS_ERR_BAD_ARG, "act NULL");
  HeadShape s = make_shape(desc);
  return launch_group_fwd(s, *g, act, groups, logits, reinterpret_cast<cudaStream_t>(stream));
}

size_t sps_group_workspace_bytes(const SpsHeadDesc* desc, const SpsGroupDesc* g) {
  if (validate_desc(desc) != SPS_OK || g == nullptr) return 0;
  return sizeof(float) * 2 * (size_t)desc->B * desc->H * desc->W * (size_t)g->J * g->Kc;
}

int sps_group_bwd(const SpsHeadDesc* desc, const SpsGroupDesc* g, const float* act, const float* groups, const float* dlogits,
                  const float* dgroups, float* dact, float* dG, float* dWg, void* workspace, size_t workspace_bytes,
                  void* stream) {
  int rc = validate_desc(desc);
  if (rc != SPS_OK) return rc;
  rc = check_group(g);
  if (rc != SPS_OK) return rc;
  SPS_REQUIRE(act != nullptr, SPS_ERR_BAD_ARG, "act NULL");
  size_t need = sps_group_workspace_bytes(desc, g);
  SPS_REQUIRE(need == 0 || (workspace != nullptr && workspace_bytes >= need), SPS_ERR_WORKSPACE,
              "group_bwd workspace: need %zu bytes, got %zu", need, workspace_bytes);
  HeadShape s = make_shape(desc);
  return launch_group_bwd(s, *g, act, groups, dlogits, dgroups, dact, dG, dWg, reinterpret_cast<float*>(workspace),
                          reinterpret_cast<cudaStream_t>(stream));
}

static int check_kld(const SpsKldDesc* k) {
  SPS_REQUIRE(k != nullptr, SPS_ERR_BAD_ARG, "kld desc NULL");
  SPS_REQUIRE(k->G >= 0 && k->Q >= 0, SPS_ERR_BAD_ARG, "kld desc G=%d Q=%d", k->G, k->Q);
  SPS_REQUIRE(k->proto_class && k->proto_group && k->proto_local && k->group_moff && k->group_poff && k->members,
              SPS_ERR_BAD_ARG, "kld desc has NULL tables");
  SPS_REQUIRE(k->Q == 0 || (k->pair_a && k->pair_b), SPS_ERR_BAD_ARG, "kld desc has NULL pair tables");
  SPS_REQUIRE(k->mmax >= 1 && k->mmax <= 32, SPS_ERR_UNSUPPORTED, "kld: largest (class, scale) group %d outside [1,32]",
              k->mmax);
  return SPS_OK;
}

size_t sps_kld_workspace_bytes(const SpsHeadDesc* desc, const SpsKldDesc* k) {
  if (validate_desc(desc) != SPS_OK || k == nullptr || k->Q < 0) return 0;
  return kld_workspace(make_shape(desc), k->Q);
}

int sps_kld_fwd(const SpsHeadDesc* desc, const SpsKldDesc* k, const float* dist, const int16_t* labels,
                void* workspace, size_t workspace_bytes, float* loss, void* stream) {
  int rc = validate_desc(desc);
  if (rc != SPS_OK) return rc;
  rc = check_kld(k);
  if (rc != SPS_OK) return rc;
  SPS_REQUIRE(loss != nullptr, SPS_ERR_BAD_ARG, "kld_fwd: loss NULL");
  HeadShape s = make_shape(desc);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (s.N == 0 || s.P == 0) {
    SPS_CUDA_TRY(cudaMemsetAsync(loss, 0, 2 * sizeof(float), st));
    return SPS_OK;
  }
  SPS_REQUIRE(dist && labels, SPS_ERR_BAD_ARG, "kld_fwd: dist/labels NULL");
  size_t need = kld_workspace(s, k->Q);
  SPS_REQUIRE(workspace != nullptr && workspace_bytes >= need, SPS_ERR_WORKSPACE,
              "kld workspace: need %zu bytes, got %zu", need, workspace_bytes);
  return launch_kld_fwd(s, *k, dist, labels, workspace, loss, st);
}

int sps_kld_bwd(const SpsHeadDesc* desc, const SpsKldDesc* k, const float* dist, void* workspace,
                size_t workspace_bytes, const float* dloss, float* ddist, void* stream) {
  int rc = validate_desc(desc);
  if (rc != SPS_OK) return rc;
  rc = check_kld(k);
  if (rc != SPS_OK) return rc;
  HeadShape s = make_shape(desc);
  if (s.N == 0 || s.P == 0) return SPS_OK;
  SPS_REQUIRE(dist && dloss && ddist, SPS_ERR_BAD_ARG, "kld_bwd: NULL argument");
  size_t need = kld_workspace(s, k->Q);
  SPS_REQUIRE(workspace != nullptr && workspace_bytes >= need, SPS_ERR_WORKSPACE,
              "kld workspace: need %zu bytes, got %zu", need, workspace_bytes);
  return launch_kld_bwd(s, *k, dist, workspace, dloss, ddist, reinterpret_cast<cudaStream_t>(stream));
}

static int check_kld_feat(const SpsHeadDesc* desc, const SpsKldDesc* k) {
  int rc = validate_desc(desc);
  if (rc != SPS_OK) return rc;
  rc = check_kld(k);
  if (rc != SPS_OK) return rc;
  SPS_REQUIRE(desc->D <= 64, SPS_ERR_UNSUPPORTED, "kld feature mode: prototype depth D=%d > 64", desc->D);
  SPS_REQUIRE(k->G == desc->K * desc->S, SPS_ERR_BAD_ARG, "kld feature mode: G=%d must be K*S=%d", k->G,
              desc->K * desc->S);
  return SPS_OK;
}

size_t sps_kld_feat_workspace_bytes(const SpsHeadDesc* desc, const SpsKldDesc* k) {
  if (validate_desc(desc) != SPS_OK || k == nullptr || k->Q < 0 || k->mmax < 1) return 0;
  return kld_feat_workspace(make_shape(desc), *k);
}

int sps_kld_feat_fwd(const SpsHeadDesc* desc, const SpsKldDesc* k, const float* z, const float* protos,
                     const int16_t* labels, void* workspace, size_t workspace_bytes, float* loss, void* stream) {
  int rc = check_kld_feat(desc, k);
  if (rc != SPS_OK) return rc;
  SPS_REQUIRE(loss != nullptr, SPS_ERR_BAD_ARG, "kld_feat_fwd: loss NULL");
  HeadShape s = make_shape(desc);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (s.N == 0 || s.P == 0) {
    SPS_CUDA_TRY(cudaMemsetAsync(loss, 0, 2 * sizeof(float), st));
    return SPS_OK;
  }
  SPS_REQUIRE(z && protos && labels, SPS_ERR_BAD_ARG, "kld_feat_fwd: NULL argument");
  size_t need = kld_feat_workspace(s, *k);
  SPS_REQUIRE(workspace != nullptr && workspace_bytes >= need, SPS_ERR_WORKSPACE,
              "kld feature workspace: need %zu bytes, got %zu", need, workspace_bytes);
  return launch_kld_feat_fwd(s, *k, z, protos, labels, workspace, loss, st);
}

int sps_kld_feat_bwd(const SpsHeadDesc* desc, const SpsKldDesc* k, const float* z, const float* protos,
                     const int16_t* labels, void* workspace, size_t workspace_bytes, const float* dloss, float* dz,
                     float* dprotos, void* stream) {
  int rc = check_kld_feat(desc, k);
  if (rc != SPS_OK) return rc;
  HeadShape s = make_shape(desc);
  if (s.N == 0 || s.P == 0) return SPS_OK;
  SPS_REQUIRE(z && protos && labels && dloss, SPS_ERR_BAD_ARG, "kld_feat_bwd: NULL argument");
  size_t need = kld_feat_workspace(s, *k);
  SPS_REQUIRE(workspace != nullptr && workspace_bytes >= need, SPS_ERR_WORKSPACE,
              "kld feature workspace: need %zu bytes, got %zu", need, workspace_bytes);
  return launch_kld_feat_bwd(s, *k, z, protos, labels, workspace, dloss, dz, dprotos,
                             reinterpret_cast<cudaStream_t>(stream));
}

int sps_upsample_argext(const float* maps, int32_t B, int32_t C, int32_t h, int32_t w, int32_t H, int32_t W,
                        int32_t argmin, int32_t* out_idx, float* out_val, void* stream) {
  SPS_REQUIRE(B >= 0 && C >= 1 && h >= 1 && w >= 1 && H >= 0 && W >= 0, SPS_ERR_BAD_ARG,
              "upsample_argext: bad shape B=%d C=%d h=%d w=%d H=%d W=%d", B, C, h, w, H, W);
  SPS_REQUIRE(B <= 65535 && (H + 3) / 4 <= 65535, SPS_ERR_UNSUPPORTED, "upsample_argext: B=%d or H=%d too large", B, H);
  SPS_REQUIRE(maps != nullptr && out_idx != nullptr, SPS_ERR_BAD_ARG, "upsample_argext: NULL argument");
  return launch_upsample_argext(maps, B, C, h, w, H, W, argmin, out_idx, out_val, reinterpret_cast<cudaStream_t>(stream));
}

int sps_push_scan(const SpsHeadDesc* desc, const void* z, const float* protos, const int16_t* labels,
                  const int32_t* proto_class, uint64_t* keys, void* workspace, size_t workspace_bytes, void* stream) {
  int rc = validate_desc(desc);
  if (rc != SPS_OK) return rc;
  SPS_REQUIRE(z && protos && labels && proto_class && keys, SPS_ERR_BAD_ARG, "push_scan: NULL argument");
  HeadShape s = make_shape(desc);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool z16 = desc->z_dtype != SPS_Z_F32;
  if (desc->math == SPS_MATH_FP32)
    return launch_push_scan_simt(s, reinterpret_cast<const float*>(z), protos, labels, proto_class,
                                 reinterpret_cast<unsigned long long*>(keys), st);
  rc = launch_push_fill(reinterpret_cast<unsigned long long*>(keys), (long long)s.B * s.P, st);
  if (rc != SPS_OK || s.B == 0 || s.P == 0 || s.HW == 0) return rc;
  return launch_push_scan_tc(s, desc->math, z, z16, protos, labels, proto_class, reinterpret_cast<unsigned long long*>(keys),
                             workspace, workspace_bytes, st);
}

int sps_push_reduce(const uint64_t* keys, int32_t n_maps, int32_t P, uint32_t map_base, uint64_t* best,
                    void* stream) {
  SPS_REQUIRE(keys && best && n_maps >= 0 && P >= 0, SPS_ERR_BAD_ARG, "push_reduce: bad argument");
  return launch_push_reduce(reinterpret_cast<const unsigned long long*>(keys), n_maps, P, map_base,
                            reinterpret_cast<unsigned long long*>(best), reinterpret_cast<cudaStream_t>(stream));
}

int sps_push_topk(const uint64_t* keys, int32_t n_maps, int32_t P, uint32_t map_base, int32_t k, uint64_t* top_key,
                  int32_t* top_pix, void* stream) {
  SPS_REQUIRE(keys && top_key && top_pix && n_maps >= 0 && P >= 0, SPS_ERR_BAD_ARG, "push_topk: bad argument");
  return launch_push_topk(reinterpret_cast<const unsigned long long*>(keys), n_maps, P, map_base, k,
                          reinterpret_cast<unsigned long long*>(top_key), top_pix,
                          reinterpret_cast<cudaStream_t>(stream));
}

int sps_push_gather(const SpsHeadDesc* desc, const void* z, const uint64_t* keys, const uint64_t* best,
                    int32_t n_maps, uint32_t map_base, const int32_t* proto_scale, float* out, void* stream) {
  int rc = validate_desc(desc);
  if (rc != SPS_OK) return rc;
  SPS_REQUIRE(z && keys && best && proto_scale && out, SPS_ERR_BAD_ARG, "push_gather: NULL argument");
  HeadShape s = make_shape(desc);
  return launch_push_gather(s, z, reinterpret_cast<const unsigned long long*>(keys),
                            reinterpret_cast<const unsigned long long*>(best), n_maps, map_base, proto_scale, out,
                            reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
