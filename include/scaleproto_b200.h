/*
 * scaleproto_b200.h — C ABI of libscaleproto_sm100.so
 *
 * B200-native (sm_100a) replacement for the prototype head and the push scan of
 * eceo-epfl/ScaleProtoSeg.  The reference has no native/FFI layer: its "interface" for this path is
 * the method set of the two PPNetMultiScale classes and three functions of
 * segmentation/push_multiscale_optimization.py.  Every entry point below names the reference code
 * it replaces (paths relative to the reference root).  INTEGRATION.md shows the ctypes binding and
 * the method patches a maintainer would add on the reference side.
 *
 * Conventions
 *  - plain pointers and sizes only; every pointer is a DEVICE pointer unless marked "host";
 *  - the caller owns all memory; the library never allocates or frees device memory;
 *  - all work is enqueued on `stream` (a cudaStream_t passed as void*); no implicit syncs;
 *  - return value 0 = OK, otherwise an SpsStatus; text via sps_last_error() (thread local);
 *  - feature maps are fp32 NCHW exactly as the reference holds them, z[b][s*D + c][h][w] (z_dtype = SPS_Z_F32,
 *    z_layout = SPS_Z_NCHW, the default), or 16-bit channels-last, z[b][h][w][s*D + c] (SPS_Z_F16 / SPS_Z_BF16 with
 *    SPS_Z_NHWC: what the stock backbone emits under channels_last + autocast); the 16-bit layout is a legal 2-D tensor
 *    map ([N][C], 512-byte rows) and is fed to the tensor cores by TMA with no conversion work;
 *  - N = B*H*W positions, row n = (b*H + h)*W + w (the order of permute(0,2,3,1).reshape(-1,P),
 *    segmentation/model/model_multiscale.py:369-370).
 */
#ifndef SCALEPROTO_B200_H
#define SCALEPROTO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPS_ABI_VERSION 2
#define SPS_MAX_SCALES 8

typedef enum {
  SPS_OK = 0,
  SPS_ERR_BAD_ARG = 1,      /* null pointer / inconsistent shape */
  SPS_ERR_UNSUPPORTED = 2,  /* shape outside what the kernels cover (message says which) */
  SPS_ERR_WORKSPACE = 3,    /* workspace too small */
  SPS_ERR_CUDA = 4          /* launch / runtime failure, message carries cudaGetErrorString */
} SpsStatus;

/* distance_2_similarity kinds, segmentation/model/model_multiscale.py:324-330 */
#define SPS_ACT_LOG 0    /* log((d+1)/(d+eps)) */
#define SPS_ACT_LINEAR 1 /* -d */

/* Arithmetic of the z.p contraction.  fp32 accumulation everywhere. */
#define SPS_MATH_FP32 0   /* fp32 FFMA on CUDA cores: exact-operand mode (gradient parity, pushed prototypes) */
#define SPS_MATH_BF16X3 1 /* tcgen05 kind::f16, bf16 operands split hi+lo, 3 MMAs per product (16-bit mantissa) */
#define SPS_MATH_BF16 2   /* tcgen05 kind::f16, bf16 operands rounded to nearest (8-bit mantissa) */
#define SPS_MATH_F16 3    /* tcgen05 kind::f16, fp16 operands rounded to nearest (11-bit mantissa; |z|,|p| < 6e4) */
#define SPS_MATH_F16X3 4  /* tcgen05 kind::f16, fp16 operands split hi+lo, 3 MMAs per product (22-bit mantissa) */

/* flags */
#define SPS_FLAG_DETERMINISTIC 1 /* backward reductions over N in a fixed order (no fp32 atomics) */
#define SPS_FLAG_ZPACK 2         /* fp32 NCHW input: re-lay the map as 16-bit channels-last tiles in a pre-pass (inside the
                                    workspace) and run the TMA-fed kernels; tensor-core math modes only */
#define SPS_FLAG_ZRELAYED 4      /* the z pointer of sps_head_fwd / sps_head_bwd is not the fp32 NCHW map itself but its
                                    re-lay, written earlier by sps_z_relay (one re-lay shared by the forward and the
                                    backward of a step); the descriptor still describes the fp32 map, dz stays fp32 NCHW */

/* element type and layout of the feature maps z (and of dz, which always mirrors z) */
#define SPS_Z_F32 0
#define SPS_Z_F16 1   /* requires SPS_MATH_F16 or SPS_MATH_F16X3 */
#define SPS_Z_BF16 2  /* requires SPS_MATH_BF16 or SPS_MATH_BF16X3 */
#define SPS_Z_NCHW 0  /* z[b][c][h][w]  (SPS_Z_F32 only) */
#define SPS_Z_NHWC 1  /* z[b][h][w][c]  (SPS_Z_F16 / SPS_Z_BF16 only; D = 64, 16-byte aligned base) */

/* Shape of one head invocation.  Mirrors the state the reference keeps on the module
 * (segmentation/model/model_multiscale.py:104-149): prototype_vectors (P,D,1,1), num_scales,
 * scale_num_prototypes (possibly ragged after prune_prototypes, :400-432), epsilon. */
typedef struct SpsHeadDesc {
  int32_t B, H, W;                      /* feature map batch and size */
  int32_t S, D;                         /* num_scales, channels per scale; C = S*D */
  int32_t P, K;                         /* prototypes, classes */
  int32_t scale_lo[SPS_MAX_SCALES + 1]; /* scale s owns prototype rows [scale_lo[s], scale_lo[s+1]) */
  int32_t act_kind;                     /* SPS_ACT_* */
  float eps;                            /* 1e-4 in the reference */
  int32_t math;                         /* SPS_MATH_* */
  int32_t flags;                        /* SPS_FLAG_* */
  int32_t z_dtype;                      /* SPS_Z_F32 (default 0) | SPS_Z_F16 | SPS_Z_BF16 */
  int32_t z_layout;                     /* SPS_Z_NCHW (default 0) | SPS_Z_NHWC */
} SpsHeadDesc;

/* Stage-2 grouping head (segmentation/model/model_multiscale_group.py:249-308).
 * Kc = number of classes that own >= 1 prototype; class k' (0..Kc-1, ascending class id) has
 * n_k' = class_off[k'+1]-class_off[k'] prototypes and a (J x n_k') projection.  The projections are
 * packed back to back: G + J*class_off[k'] is the row-major (J, n_k') weight of class k'. */
typedef struct SpsGroupDesc {
  int32_t Kc;               /* classes with prototypes */
  int32_t J;                /* num_groups (3) */
  const int32_t* proto_slot; /* [P] k' of prototype p */
  const int32_t* proto_col;  /* [P] column of p inside its class (rank among the class's prototypes) */
  const int32_t* class_off;  /* [Kc+1] prefix sum of n_k' */
  const int32_t* members;    /* [P] prototype ids grouped by class: members[class_off[k'] + q] (ascending p) */
  const float* G;            /* [J*P] packed projections */
  const float* Wg;           /* [K x J*Kc] last_layer_group.weight, row major */
} SpsGroupDesc;

/* Tables of the KLD loss between same-class prototypes (segmentation/model/loss.py:53-146).
 * A group is one (class, scale) with its prototypes in ascending order; groups are numbered class-major.
 * Pairs are the unordered prototype pairs (a < b) inside a group, numbered group by group, a-major. */
typedef struct SpsKldDesc {
  int32_t G;                  /* groups = K * S (empty groups allowed) */
  int32_t Q;                  /* pairs over all groups */
  int32_t mmax;               /* largest group size, 1..32 (also: rows per scale of the compact block, feature mode) */
  int32_t reserved;           /* 0 */
  const int32_t* proto_class; /* [P] class of prototype p (argmax of prototype_class_identity) */
  const int32_t* proto_group; /* [P] group of p */
  const int32_t* proto_local; /* [P] rank of p inside its group */
  const int32_t* group_moff;  /* [G+1] prefix sum of group sizes (into members) */
  const int32_t* group_poff;  /* [G+1] prefix sum of pairs per group (into pair_a / pair_b) */
  const int32_t* members;     /* [P] prototype ids ordered by group */
  const int32_t* pair_a;      /* [Q] first prototype of pair q */
  const int32_t* pair_b;      /* [Q] second prototype of pair q */
} SpsKldDesc;

int sps_version(void);
const char* sps_last_error(void);
/* Number of kernels this library has launched in this process (bench.py reports it as gpu_launches). */
uint64_t sps_launch_count(void);

/* Calls that asked for a tensor-core math mode and were served by the exact fp32 CUDA-core kernels instead, because the
 * shape does not fit the tcgen05 kernels (bench.py prints it; 0 on every BASELINE shape). */
uint64_t sps_fp32_reroute_count(void);

/* 16-bit channels-last re-lay of an fp32 NCHW map for the TMA-fed kernels: [N][C] hi part, then (3-term math modes) the
 * [N][C] lo part.  out: >= sps_z_relay_bytes(desc) bytes.  Pass `out` as z with SPS_FLAG_ZRELAYED. */
size_t sps_z_relay_bytes(const SpsHeadDesc* desc);
int sps_z_relay(const SpsHeadDesc* desc, const float* z, void* out, size_t out_bytes, void* stream);

/* Bytes of scratch sps_head_bwd needs for this shape (0 is possible). */
size_t sps_workspace_bytes(const SpsHeadDesc* desc);
/* Bytes of scratch sps_head_fwd / sps_push_scan need (packed 16-bit prototype images of the tensor-core
 * modes; 0 for SPS_MATH_FP32). */
size_t sps_head_fwd_workspace_bytes(const SpsHeadDesc* desc);

/* Prototype head forward.
 * Replaces PPNetMultiScale._scale_l2_convolution + _l2_convolution + distance_2_similarity +
 * run_last_layer as chained by forward_from_conv_features
 * (segmentation/model/model_multiscale.py:256-317, 324-330, 243-244, 340-388).
 *   z        [B, S*D, H, W] fp32, or [B, H, W, S*D] fp16 / bf16 (desc->z_dtype, z_layout)
 *   protos   [P, D] fp32
 *   last_w   [K, P] fp32 or NULL (NULL: no logits, e.g. prototype_distances / push_forward / stage 2)
 *   logits   [N, K] fp32 or NULL (required iff last_w != NULL)
 *   dist     [B, P, H, W] fp32 or NULL   (NULL = fused mode, the map is never written to HBM)
 *   act      [N, P] fp32 or NULL
 *   workspace >= sps_head_fwd_workspace_bytes(desc) bytes (may be NULL when that is 0) */
int sps_head_fwd(const SpsHeadDesc* desc, const void* z, const float* protos, const float* last_w,
                 float* logits, float* dist, float* act, void* workspace, size_t workspace_bytes, void* stream);

/* Prototype head forward that allocates nothing and takes no workspace: for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * NO COUNTERPART IN THE REFERENCE.
 * Same arguments, outputs, routing and error codes as sps_head_fwd; the scratch the forward needs (the packed prototype /
 * last-layer images, sps_head_fwd_workspace_bytes(desc)) lives in a static device array of SPS_STATIC_SCRATCH_BYTES
 * inside the library (held by every GPU that loads the library until exit). The call itself uses no cudaMalloc /
 * cudaFree / allocating library routine.
 *   - shapes whose scratch exceeds SPS_STATIC_SCRATCH_BYTES are rejected with SPS_ERR_WORKSPACE and a message naming
 *     both sizes (Cityscapes shape: 0.3 MB; COCO-Stuff shape with 16-bit maps: 6.5 MB; an fp32 map that needs the 16-bit
 *     re-lay, e.g. K > 32 with fp32 maps, does not fit: use sps_head_fwd);
 *   - every call on a device shares the one array: calls that overlap on the GPU (two streams, two threads, two graphs
 *     replayed concurrently) RACE; serialise them on one stream. Contents left by a call are never read by the next
 *     (the pre-pass rewrites all of it), so nothing needs clearing between calls;
 *   - the first call on each device asks the runtime for the array's address (cudaGetSymbolAddress, with that device
 *     current): make it once outside graph capture. */
#define SPS_STATIC_SCRATCH_BYTES (8u << 20)
int sps_head_fwd_static(const SpsHeadDesc* desc, const void* z, const float* protos, const float* last_w,
                        float* logits, float* dist, float* act, void* stream);

/* Prototype head backward (what autograd derives from the lines above; driven by manual_backward,
 * segmentation/model/module_multiscale.py:292).  d and a are recomputed from z and protos.
 *   dlogits [N,K] or NULL, ddist [B,P,H,W] or NULL (KLD loss back-props straight into distances,
 *   module_multiscale.py:242), dact [N,P] or NULL.
 *   dz (same element type and layout as z) or NULL, dprotos [P,D] or NULL, dlast_w [K,P] or NULL: OVERWRITTEN.
 *   workspace: >= sps_workspace_bytes(desc) bytes, contents undefined on entry and exit.
 * Routing: a tensor-core math mode with last_w, dlogits (and optionally ddist) runs the tcgen05 backward (more
 * than 32 classes: the one-pass wide backward, K <= 192, D == 64, no ddist; with ddist they run as slabs of 32); a gradient through act, a call without
 * last_w/dlogits, K > 256, or SPS_MATH_FP32 runs the exact-operand CUDA-core kernels. */
int sps_head_bwd(const SpsHeadDesc* desc, const void* z, const float* protos, const float* last_w,
                 const float* dlogits, const float* ddist, const float* dact, void* dz, float* dprotos,
                 float* dlast_w, void* workspace, size_t workspace_bytes, void* stream);

/* Stage-2 group projection + last_layer_group.
 * Replaces compute_group + run_last_layer (segmentation/model/model_multiscale_group.py:283-308).
 *   act [N,P] -> groups [N, J*Kc] (exp of the projections; may be NULL) and logits [N,K] (may be NULL). */
int sps_group_fwd(const SpsHeadDesc* desc, const SpsGroupDesc* g, const float* act, float* groups,
                  float* logits, void* stream);

/* Scratch bytes sps_group_bwd needs (2 * N * J*Kc floats). */
size_t sps_group_workspace_bytes(const SpsHeadDesc* desc, const SpsGroupDesc* g);

/* Backward of sps_group_fwd.  dlogits [N,K] or NULL, dgroups [N,J*Kc] or NULL.
 *   groups [N,J*Kc] or NULL: the group activations sps_group_fwd returned.  When given, the (N x P) activation map is read
 *   exactly once in the whole backward (by the dG reduction); NULL recomputes them from act.
 *   dact [N,P] or NULL (overwritten), dG [J*P] or NULL, dWg [K x J*Kc] or NULL (overwritten). */
int sps_group_bwd(const SpsHeadDesc* desc, const SpsGroupDesc* g, const float* act, const float* groups, const float* dlogits,
                  const float* dgroups, float* dact, float* dG, float* dWg, void* workspace,
                  size_t workspace_bytes, void* stream);

/* KLD loss between the same-class prototypes of every scale.  Replaces KLDLoss.forward
 * (segmentation/model/loss.py:66-146: the image x class x scale x pair loop of masked_select / log_softmax /
 * kl_div calls) called from segmentation/model/module_multiscale.py:242.
 *   dist    [B, P, H, W] fp32 distances (the head's dense output)
 *   labels  [B, H, W] int16, 0 = void, c+1 = class c (the loss subtracts 1, loss.py:77)
 *   loss    [2] fp32: mean of exp(-kld) over the contributing pairs (0 when none), and their number
 *   workspace >= sps_kld_workspace_bytes bytes; sps_kld_bwd reads what sps_kld_fwd left in it. */
size_t sps_kld_workspace_bytes(const SpsHeadDesc* desc, const SpsKldDesc* k);
int sps_kld_fwd(const SpsHeadDesc* desc, const SpsKldDesc* k, const float* dist, const int16_t* labels,
                void* workspace, size_t workspace_bytes, float* loss, void* stream);
/* Gradient of the loss into the distance map.  dloss [1] fp32 (device) upstream gradient of the scalar;
 * ddist [B, P, H, W] OVERWRITTEN (zero wherever the pixel's class differs from the prototype's). */
int sps_kld_bwd(const SpsHeadDesc* desc, const SpsKldDesc* k, const float* dist, void* workspace,
                size_t workspace_bytes, const float* dloss, float* ddist, void* stream);

/* The same loss straight from the feature maps ("feature mode"): the distances the loss needs are only those between a
 * pixel and the prototypes of the pixel's own class (1/K of the map), so they are computed here from z and the
 * prototypes into a compact block inside the workspace and the (B,P,H,W) map is never needed — the head can run in
 * fused mode (dist = NULL) during stage-1 training.  desc carries the head's S, D and scale_lo; every group must
 * lie inside one scale (it does by construction); D <= 64.
 *   z [B,S*D,H,W], protos [P,D], labels as above; loss [2] as sps_kld_fwd.
 *   sps_kld_feat_bwd: dz [B,S*D,H,W] and dprotos [P,D] (either may be NULL) are OVERWRITTEN with the loss's own
 *   gradient (the caller adds them to the head's); workspace as left by sps_kld_feat_fwd. */
size_t sps_kld_feat_workspace_bytes(const SpsHeadDesc* desc, const SpsKldDesc* k);
int sps_kld_feat_fwd(const SpsHeadDesc* desc, const SpsKldDesc* k, const float* z, const float* protos,
                     const int16_t* labels, void* workspace, size_t workspace_bytes, float* loss, void* stream);
int sps_kld_feat_bwd(const SpsHeadDesc* desc, const SpsKldDesc* k, const float* z, const float* protos,
                     const int16_t* labels, void* workspace, size_t workspace_bytes, const float* dloss, float* dz,
                     float* dprotos, void* stream);

/* Evaluation epilogue: bilinear upsampling (align_corners = False) of a low-resolution map to the annotation size and
 * argmax (logits -> predicted class) or argmin (distances -> nearest prototype) over the channel axis, without
 * materialising the upsampled map.  Replaces F.interpolate + torch.argmax / torch.argmin of
 * segmentation/eval_valid_multiscale.py:229-234.
 *   maps    [B, C, h, w] fp32 (logits permuted to channels-first as :221 does, or the head's distances)
 *   out_idx [B, H, W] int32 winning channel (first one on ties);  out_val [B, H, W] fp32 its value, or NULL */
int sps_upsample_argext(const float* maps, int32_t B, int32_t C, int32_t h, int32_t w, int32_t H, int32_t W,
                        int32_t argmin, int32_t* out_idx, float* out_val, void* stream);

/* Push scan: per map and per prototype, the nearest position among those labelled with the
 * prototype's class.  Replaces compute_distances lines 68-91 of
 * segmentation/push_multiscale_optimization.py (one-hot, void column drop, 1e10 mask, min/argmin).
 *   labels      [B, H, W] int16, already resized to the map (PIL NEAREST stays on the host,
 *               segmentation/data/dataset.py:22-30); 0 = void, c+1 = class c
 *   proto_class [P] int32 class of each prototype (argmax of prototype_class_identity)
 *   keys        [B, P] uint64, OVERWRITTEN: (float_bits(min masked distance) << 32) | flat index h*W+w
 *               of its first occurrence; (bits(1e10f) << 32) | 0 when the class is absent.
 *   workspace   >= sps_head_fwd_workspace_bytes(desc) bytes (NULL / 0 for SPS_MATH_FP32) */
int sps_push_scan(const SpsHeadDesc* desc, const void* z, const float* protos, const int16_t* labels,
                  const int32_t* proto_class, uint64_t* keys, void* workspace, size_t workspace_bytes, void* stream);

/* Dataset arg-min over maps (push_multiscale_optimization.py:135-137, argmin(dim=0): first map wins).
 *   keys [n_maps, P] from sps_push_scan; map_base = global index of keys row 0 (shard offset).
 *   best [P] uint64, MERGED by min: (float_bits(d) << 32) | global map index.  Pre-fill with
 *   0xFFFFFFFFFFFFFFFF (or a previous shard's result); one NCCL min all-reduce combines ranks. */
int sps_push_reduce(const uint64_t* keys, int32_t n_maps, int32_t P, uint32_t map_base, uint64_t* best,
                    void* stream);

/* k nearest maps per prototype: the generalisation of sps_push_reduce that find_k_nearest_patches_to_prototypes needs
 * (find_nearest.py:89-175 keeps one heap of size k per prototype and feeds it every image's closest patch; the
 * nearest_img / nearest_proto analyses use the same lists).  Scan with labels = 1 at every non-void pixel and
 * proto_class = 0 to reproduce its void-only mask (find_nearest.py:138), or with the class labels for class-pure lists.
 *   keys [n_maps, P] from sps_push_scan; k in [1,16];
 *   top_key [P, k] uint64 ascending, MERGED: (float_bits(d) << 32) | global map index; pre-fill with 0xFF..FF
 *   top_pix [P, k] int32 flat position h*W+w of each entry inside its map; pre-fill with 0. */
int sps_push_topk(const uint64_t* keys, int32_t n_maps, int32_t P, uint32_t map_base, int32_t k, uint64_t* top_key,
                  int32_t* top_pix, void* stream);

/* Gather of the winning feature vectors (push_multiscale_optimization.py:182-188).
 *   z [n_maps,S*D,H,W]; best [P] (global map index in the low word); keys [n_maps,P]; only prototypes
 *   whose winning map lies in [map_base, map_base+n_maps) are written.
 *   proto_scale [P] int32 scale whose channel block is read (the reference uses p // (P/S), :161,:167).
 *   out [P, D] fp32. */
int sps_push_gather(const SpsHeadDesc* desc, const void* z, const uint64_t* keys, const uint64_t* best,
                    int32_t n_maps, uint32_t map_base, const int32_t* proto_scale, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SCALEPROTO_B200_H */
