// Fused KLD loss between same-class prototypes (segmentation/model/loss.py:53-146).
//
// The reference walks image x class x scale x prototype-pair in Python, masking the distance map and calling
// log_softmax / kl_div per pair (about 6 small kernels and one host sync per pair).  Here the labels of an image are
// bucketed by class once (pixel index lists), and everything after that touches only the (pixel, prototype) pairs
// whose classes agree — 1/K of the distance map:
//   kld_hist     per image: pixels per class and their exclusive offsets
//   kld_bucket   per (class, image): ordered list of the class's pixels (ballot + prefix: the order is the pixel
//                order, so every later sum has a fixed order)
//   kld_lse      one warp per (image, prototype): log-sum-exp of the prototype's distances over its class's pixels
//   kld_pairs    one warp per (image, pair): KL(a||b) and KL(b||a) over the class's pixels, term = exp(-kld)
//   kld_finish   one CTA: mean of the terms in a fixed order
//   kld_bwd      one warp per (image, prototype): gradient into the distance map at the class's pixels
// All sums are fp32 in a fixed order; exp/log are the full-precision library versions (the work is tiny).
#include "common.cuh"

namespace sps {

namespace {

constexpr int KLD_THREADS = 256;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// cnt[b][c], off[b][c] for c in [0, K): pixels of image b labelled c+1 (0 = void and labels above K are dropped).
__global__ void kld_hist_kernel(const int16_t* __restrict__ labels, int HW, int K, int* __restrict__ cnt,
                                int* __restrict__ off) {
  extern __shared__ int hist[];  // K
  const int b = blockIdx.x;
  for (int c = threadIdx.x; c < K; c += blockDim.x) hist[c] = 0;
  __syncthreads();
  const int16_t* lab = labels + (size_t)b * HW;
  for (int x = threadIdx.x; x < HW; x += blockDim.x) {
    int c = (int)lab[x] - 1;
    if (c >= 0 && c < K) atomicAdd(&hist[c], 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int c = 0; c < K; ++c) {
      cnt[b * K + c] = hist[c];
      off[b * K + c] = run;
      run += hist[c];
    }
  }
}

// idx[b*HW + off[b][c] + r] = r-th pixel (ascending) of image b labelled c+1.
__global__ void kld_bucket_kernel(const int16_t* __restrict__ labels, int HW, int K, const int* __restrict__ cnt,
                                  const int* __restrict__ off, int* __restrict__ idx) {
  __shared__ int warp_cnt[KLD_THREADS / 32];
  __shared__ int base;
  const int c = blockIdx.x, b = blockIdx.y;
  if (cnt[b * K + c] == 0) return;
  const int16_t* lab = labels + (size_t)b * HW;
  int* out = idx + (size_t)b * HW + off[b * K + c];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) base = 0;
  __syncthreads();
  for (int x0 = 0; x0 < HW; x0 += KLD_THREADS) {
    const int x = x0 + threadIdx.x;
    const bool hit = x < HW && (int)lab[x] == c + 1;
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    if (lane == 0) warp_cnt[warp] = __popc(m);
    __syncthreads();
    int before = base;
    for (int w = 0; w < warp; ++w) before += warp_cnt[w];
    if (hit) out[before + __popc(m & ((1u << lane) - 1u))] = x;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < KLD_THREADS / 32; ++w) t += warp_cnt[w];
      base += t;
    }
    __syncthreads();
  }
}

// Row of prototype p inside one image's distance block.  Dense map: the prototype itself (rows = P).  Compact map
// (feature mode): row (scale, rank inside the (class, scale) group) of rows = S * mmax, where the entry at pixel x
// belongs to the group of x's own class — the only prototypes the loss ever pairs with that pixel.
struct RowMap {
  int compact, rows, S, mmax;
  const int* proto_group;
  const int* proto_local;
  __device__ __forceinline__ int operator()(int p) const {
    return compact ? (proto_group[p] % S) * mmax + proto_local[p] : p;
  }
};

// lse[b*P + p] = log sum_x exp(d[b,p,x]) over the pixels of p's class (loss.py:112-115 log_softmax denominator).
__global__ void kld_lse_kernel(const float* __restrict__ dist, RowMap rm, int B, int P, int HW, int K,
                               const int* __restrict__ proto_class, const int* __restrict__ cnt,
                               const int* __restrict__ off, const int* __restrict__ idx, float* __restrict__ lse) {
  const long long wid = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (wid >= (long long)B * P) return;
  const int b = (int)(wid / P), p = (int)(wid % P);
  const int c = proto_class[p];
  const int n = (c >= 0 && c < K) ? cnt[b * K + c] : 0;
  if (n == 0) {
    if (lane == 0) lse[wid] = 0.0f;
    return;
  }
  const int* ix = idx + (size_t)b * HW + off[b * K + c];
  const float* row = dist + ((size_t)b * rm.rows + rm(p)) * HW;
  float mx = -INFINITY;
  for (int i = lane; i < n; i += 32) mx = fmaxf(mx, row[ix[i]]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float s = 0.0f;
  for (int i = lane; i < n; i += 32) s += expf(row[ix[i]] - mx);
  s = warp_sum(s);
  if (lane == 0) lse[wid] = mx + logf(s);
}

// One warp per (image, pair q): kl[b*Q+q] = (KL(a||b), KL(b||a)), term[b*Q+q] = exp(-(sum)/2), or -1 when the pair
// does not contribute (class absent from the image or fewer than two pixels, loss.py:117-135).
__global__ void kld_pairs_kernel(const float* __restrict__ dist, RowMap rm, int B, int P, int HW, int K, int Q,
                                 const int* __restrict__ pair_a, const int* __restrict__ pair_b,
                                 const int* __restrict__ proto_class, const int* __restrict__ cnt,
                                 const int* __restrict__ off, const int* __restrict__ idx,
                                 const float* __restrict__ lse, float2* __restrict__ kl, float* __restrict__ term) {
  const long long wid = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (wid >= (long long)B * Q) return;
  const int b = (int)(wid / Q), q = (int)(wid % Q);
  const int pa = pair_a[q], pb = pair_b[q];
  const int c = proto_class[pa];
  const int n = (c >= 0 && c < K) ? cnt[b * K + c] : 0;
  if (n < 2) {
    if (lane == 0) {
      kl[wid] = make_float2(0.0f, 0.0f);
      term[wid] = -1.0f;  // marks "no contribution" (a real term is exp(-kld) >= 0)
    }
    return;
  }
  const int* ix = idx + (size_t)b * HW + off[b * K + c];
  const float* ra = dist + ((size_t)b * rm.rows + rm(pa)) * HW;
  const float* rb = dist + ((size_t)b * rm.rows + rm(pb)) * HW;
  const float la = lse[(size_t)b * P + pa], lb = lse[(size_t)b * P + pb];
  float kab = 0.0f, kba = 0.0f;
  for (int i = lane; i < n; i += 32) {
    const int x = ix[i];
    const float a = ra[x] - la, bb = rb[x] - lb;
    const float diff = a - bb;
    kab += expf(a) * diff;
    kba -= expf(bb) * diff;
  }
  kab = warp_sum(kab);
  kba = warp_sum(kba);
  if (lane == 0) {
    kl[wid] = make_float2(kab, kba);
    term[wid] = expf(-0.5f * (kab + kba));
  }
}

// out[0] = mean of the contributing terms (0 when none, loss.py:142-146), out[1] = their number.
__global__ void kld_finish_kernel(const float* __restrict__ term, long long total, float* __restrict__ out) {
  __shared__ float ssum[32];
  __shared__ float scnt[32];
  float s = 0.0f, m = 0.0f;
  for (long long i = threadIdx.x; i < total; i += blockDim.x) {
    const float t = term[i];
    if (t >= 0.0f) {
      s += t;
      m += 1.0f;
    }
  }
  s = warp_sum(s);
  m = warp_sum(m);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    ssum[warp] = s;
    scnt[warp] = m;
  }
  __syncthreads();
  if (warp == 0) {
    const int nw = blockDim.x >> 5;
    s = lane < nw ? ssum[lane] : 0.0f;
    m = lane < nw ? scnt[lane] : 0.0f;
    s = warp_sum(s);
    m = warp_sum(m);
    if (lane == 0) {
      out[0] = m > 0.0f ? s / m : 0.0f;
      out[1] = m;
    }
  }
}

// Gradient into the distance block.  One CTA per (group, image); thread = one pixel of the group's class.  With
// l = log-softmax over the class's pixels and p = e^l, for every member a of the group
//   ddist[b,a,y] = dloss * sum_{o != a} (-term_ao / M) * 1/2 * [ p_a(y) * (l_a(y) - l_o(y) + 1 - KL(a||o)) - p_o(y) ]
// (derivative of exp(-kld) through the two log_softmax; the rest of ddist was zeroed by the caller).  The pair weights
// and directional KLs of the group sit in shared memory (m x m), the pixel's m log-probabilities too, so every
// distance is read once and every exp taken once.
constexpr int BWD_THREADS = 128;

__global__ void __launch_bounds__(BWD_THREADS)
kld_bwd_kernel(const float* __restrict__ dist, RowMap rm, int B, int P, int HW, int K, int S, int Q,
               const int* __restrict__ group_moff, const int* __restrict__ group_poff, const int* __restrict__ members,
               const int* __restrict__ cnt, const int* __restrict__ off, const int* __restrict__ idx,
               const float* __restrict__ lse, const float2* __restrict__ kl, const float* __restrict__ term,
               const float* __restrict__ fin, const float* __restrict__ dloss, float* __restrict__ ddist) {
  extern __shared__ float sm[];
  const int g = blockIdx.x, b = blockIdx.y, c = g / S;
  const int m0 = group_moff[g], m = group_moff[g + 1] - m0;
  const int n = cnt[b * K + c];
  const float M = fin[1];
  if (m < 2 || n < 2 || M <= 0.0f) return;
  float* wt = sm;                         // [m][m] weight of pair (a,o): -term/M * dloss/2, 0 on the diagonal
  float* kd = wt + m * m;                 // [m][m] KL(a || o)
  float* ls = kd + m * m;                 // [m] log-sum-exp of member a
  int* rowof = reinterpret_cast<int*>(ls + m);   // [m] row of member a in the distance block
  float* lt = reinterpret_cast<float*>(rowof + m);   // [m][BWD_THREADS] log-probabilities of this thread's pixel
  float* pt = lt + m * BWD_THREADS;                  // [m][BWD_THREADS] probabilities
  const int q0 = group_poff[g];
  const float scale = -0.5f * dloss[0] / M;
  for (int e = threadIdx.x; e < m * m; e += BWD_THREADS) {
    const int a = e / m, o = e % m;
    float w = 0.0f, k = 0.0f;
    if (a != o) {
      const int lo = a < o ? a : o, hi = a < o ? o : a;
      const size_t q = (size_t)b * Q + q0 + lo * (2 * m - lo - 1) / 2 + (hi - lo - 1);
      const float2 k2 = kl[q];
      w = scale * term[q];
      k = a < o ? k2.x : k2.y;
    }
    wt[e] = w;
    kd[e] = k;
  }
  for (int a = threadIdx.x; a < m; a += BWD_THREADS) {
    const int p = members[m0 + a];
    ls[a] = lse[(size_t)b * P + p];
    rowof[a] = rm(p);
  }
  __syncthreads();
  const int* ix = idx + (size_t)b * HW + off[b * K + c];
  const float* dblk = dist + (size_t)b * rm.rows * HW;
  float* gblk = ddist + (size_t)b * rm.rows * HW;
  const int t = threadIdx.x;
  for (int i = t; i < n; i += BWD_THREADS) {
    const int x = ix[i];
    for (int a = 0; a < m; ++a) {
      const float l = dblk[(size_t)rowof[a] * HW + x] - ls[a];
      lt[a * BWD_THREADS + t] = l;
      pt[a * BWD_THREADS + t] = expf(l);
    }
    for (int a = 0; a < m; ++a) {
      const float l_a = lt[a * BWD_THREADS + t], p_a = pt[a * BWD_THREADS + t];
      float acc = 0.0f;
      for (int o = 0; o < m; ++o)   // the diagonal carries weight 0
        acc = fmaf(wt[a * m + o], fmaf(p_a, l_a - lt[o * BWD_THREADS + t] + 1.0f - kd[a * m + o], -pt[o * BWD_THREADS + t]), acc);
      gblk[(size_t)rowof[a] * HW + x] = acc;
    }
  }
}

inline size_t kld_bwd_smem(int mmax) {
  return sizeof(float) * ((size_t)2 * mmax * mmax + 2 * mmax + (size_t)2 * mmax * BWD_THREADS);
}

// ---- feature mode: the class-matched distances straight from z, no (B,P,H,W) map --------------------------------
constexpr int FEAT_THREADS = 128;

// dc[b][s*mmax + j][x] = |z_s(x) - p_j|^2 for the j-th prototype of group (class(x), s)  (the value
// _l2_convolution gives, model_multiscale.py:256-281, in its difference form: never negative, so no clamp).
// One thread per (pixel, scale, image); the pixel's D <= 64 channels sit in registers (coalesced loads along x) and
// every prototype row is read once, 16 bytes at a time (lanes of one class share the address).
constexpr int FEAT_D = 64;

__device__ __forceinline__ float4 proto_quad(const float* pv, int c4, int D, bool vec) {
  if (vec) return reinterpret_cast<const float4*>(pv)[c4];
  float4 r;
  r.x = 4 * c4 + 0 < D ? pv[4 * c4 + 0] : 0.0f;
  r.y = 4 * c4 + 1 < D ? pv[4 * c4 + 1] : 0.0f;
  r.z = 4 * c4 + 2 < D ? pv[4 * c4 + 2] : 0.0f;
  r.w = 4 * c4 + 3 < D ? pv[4 * c4 + 3] : 0.0f;
  return r;
}

__global__ void __launch_bounds__(FEAT_THREADS)
kld_feat_dist_kernel(const float* __restrict__ z, const float* __restrict__ protos, const int16_t* __restrict__ labels,
                     int HW, int K, int S, int D, int mmax, const int* __restrict__ group_moff,
                     const int* __restrict__ members, float* __restrict__ dc) {
  const int x = blockIdx.x * FEAT_THREADS + threadIdx.x, s = blockIdx.y, b = blockIdx.z;
  if (x >= HW) return;
  const int c = (int)labels[(size_t)b * HW + x] - 1;
  if (c < 0 || c >= K) return;
  const int g = c * S + s, m0 = group_moff[g], m = group_moff[g + 1] - m0;
  const float* zp = z + ((size_t)b * S * D + (size_t)s * D) * HW + x;
  float zr[FEAT_D];
#pragma unroll
  for (int ch = 0; ch < FEAT_D; ++ch) zr[ch] = ch < D ? zp[(size_t)ch * HW] : 0.0f;
  float* out = dc + ((size_t)b * S * mmax + (size_t)s * mmax) * HW + x;
  const bool vec = (D & 3) == 0;
  const int nq = (D + 3) >> 2;
  // (unrolling this loop by two members takes 143 registers and is slower: 74 -> 83 us)
  for (int j = 0; j < m; ++j) {
    const float* pv = protos + (size_t)members[m0 + j] * D;
    float a0 = 0.0f, a1 = 0.0f;
#pragma unroll
    for (int c4 = 0; c4 < FEAT_D / 4; ++c4) {
      if (c4 < nq) {
        const float4 p = proto_quad(pv, c4, D, vec);   // beyond D both z and p are 0
        const float d0 = zr[4 * c4 + 0] - p.x, d1 = zr[4 * c4 + 1] - p.y;
        const float d2 = zr[4 * c4 + 2] - p.z, d3 = zr[4 * c4 + 3] - p.w;
        a0 = fmaf(d0, d0, a0);
        a1 = fmaf(d1, d1, a1);
        a0 = fmaf(d2, d2, a0);
        a1 = fmaf(d3, d3, a1);
      }
    }
    out[(size_t)j * HW] = a0 + a1;
  }
}

// dz[b, s*D + ch, x] = 2 * sum_j g_j(x) * (z - p_j)[ch] over the group of x's class; 0 for void pixels.  One thread
// per (pixel, 16-channel slice, scale, image): 16 partial sums -sum_j g_j p_j[ch] in registers (a 64-register tile
// per thread left 20 warps per SM waiting on dependent loads: 154 us; the slices quadruple the warps in flight).
constexpr int DZ_SLICE = 16;

__global__ void __launch_bounds__(FEAT_THREADS)
kld_feat_dz_kernel(const float* __restrict__ z, const float* __restrict__ protos, const int16_t* __restrict__ labels,
                   int HW, int K, int S, int D, int mmax, const int* __restrict__ group_moff,
                   const int* __restrict__ members, const float* __restrict__ gc, float* __restrict__ dz) {
  const int nsl = (D + DZ_SLICE - 1) / DZ_SLICE;
  const int x = blockIdx.x * FEAT_THREADS + threadIdx.x, s = blockIdx.y / nsl, c0 = (blockIdx.y % nsl) * DZ_SLICE;
  const int b = blockIdx.z;
  if (x >= HW) return;
  const int c = (int)labels[(size_t)b * HW + x] - 1;
  float* dzp = dz + ((size_t)b * S * D + (size_t)s * D) * HW + x;
  int m0 = 0, m = 0;
  if (c >= 0 && c < K) {
    m0 = group_moff[c * S + s];
    m = group_moff[c * S + s + 1] - m0;
  }
  // the slice's own z values first: their DRAM latency overlaps the member loop (read one by one next to the stores
  // they were 60 % of the kernel's stall samples)
  const float* zp = z + ((size_t)b * S * D + (size_t)s * D) * HW + x;
  float zr[DZ_SLICE];
#pragma unroll
  for (int i = 0; i < DZ_SLICE; ++i) zr[i] = (m > 0 && c0 + i < D) ? zp[(size_t)(c0 + i) * HW] : 0.0f;
  float acc[DZ_SLICE];
#pragma unroll
  for (int i = 0; i < DZ_SLICE; ++i) acc[i] = 0.0f;
  float gsum = 0.0f;
  const float* gp = gc + ((size_t)b * S * mmax + (size_t)s * mmax) * HW + x;
  const bool vec = (D & 3) == 0;
  for (int j0 = 0; j0 < m; j0 += 2) {   // two members per round: both gradients and all eight prototype quads in flight
    const bool two = j0 + 1 < m;
    const float g0 = gp[(size_t)j0 * HW];
    const float g1 = two ? gp[(size_t)(j0 + 1) * HW] : 0.0f;
    const float* pv0 = protos + (size_t)members[m0 + j0] * D;
    const float* pv1 = protos + (size_t)members[m0 + (two ? j0 + 1 : j0)] * D;
    float4 q0[DZ_SLICE / 4], q1[DZ_SLICE / 4];
#pragma unroll
    for (int i4 = 0; i4 < DZ_SLICE / 4; ++i4) {
      const bool in = c0 + 4 * i4 < D;
      q0[i4] = in ? proto_quad(pv0, c0 / 4 + i4, D, vec) : make_float4(0.f, 0.f, 0.f, 0.f);
      q1[i4] = in ? proto_quad(pv1, c0 / 4 + i4, D, vec) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    gsum += g0 + g1;
#pragma unroll
    for (int i4 = 0; i4 < DZ_SLICE / 4; ++i4) {
      acc[4 * i4 + 0] = fmaf(-g1, q1[i4].x, fmaf(-g0, q0[i4].x, acc[4 * i4 + 0]));
      acc[4 * i4 + 1] = fmaf(-g1, q1[i4].y, fmaf(-g0, q0[i4].y, acc[4 * i4 + 1]));
      acc[4 * i4 + 2] = fmaf(-g1, q1[i4].z, fmaf(-g0, q0[i4].z, acc[4 * i4 + 2]));
      acc[4 * i4 + 3] = fmaf(-g1, q1[i4].w, fmaf(-g0, q0[i4].w, acc[4 * i4 + 3]));
    }
  }
#pragma unroll
  for (int i = 0; i < DZ_SLICE; ++i) {
    const int ch = c0 + i;
    if (ch < D) dzp[(size_t)ch * HW] = m > 0 ? 2.0f * fmaf(gsum, zr[i], acc[i]) : 0.0f;
  }
}

// part[b][p][ch] = 2 * sum_x g_p(x) * (p - z_s(x))[ch] over the pixels of p's class in image b.  One CTA per
// (group, image): tiles of 32 bucketed pixels staged in shared memory (z tile D x 32, gradient tile m x 32).  Thread
// (ch = t mod 64, jq = t / 64) owns channel ch of members jq, jq+4, ...: one z read from shared memory feeds all of
// them, and p * sum g - sum g z needs one FMA per (pixel, member) instead of a subtract and an FMA.
constexpr int DP_THREADS = 256;
constexpr int DP_JQ = DP_THREADS / FEAT_D;   // 4 member lanes
constexpr int DP_MAXE = 8;                   // members per thread: covers groups of up to 32

// E = members per thread = ceil(mmax / 4), a compile-time count: with a fixed 8 every pixel issued 8 predicated
// LDS + FFMA + FADD triples, 5 of them idle at the reference's 10 prototypes per class and scale
template <int E>
__global__ void __launch_bounds__(DP_THREADS)
kld_feat_dp_kernel(const float* __restrict__ z, const float* __restrict__ protos, int HW, int K, int S, int D, int P,
                   int mmax, const int* __restrict__ group_moff, const int* __restrict__ members,
                   const int* __restrict__ cnt, const int* __restrict__ off, const int* __restrict__ idx,
                   const float* __restrict__ gc, float* __restrict__ part) {
  extern __shared__ float sm[];  // zt[D][33], gt[mmax][33]
  const int g = blockIdx.x, b = blockIdx.y;
  const int c = g / S, s = g % S;
  const int m0 = group_moff[g], m = group_moff[g + 1] - m0;
  if (m == 0) return;
  float* zt = sm;
  float* gt = sm + (size_t)D * 33;
  const int n = cnt[b * K + c];
  const int* ix = idx + (size_t)b * HW + off[b * K + c];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = DP_THREADS >> 5;
  const int ch = threadIdx.x % FEAT_D, jq = threadIdx.x / FEAT_D;
  float sgz[E], sg[E];
#pragma unroll
  for (int e = 0; e < E; ++e) sgz[e] = sg[e] = 0.0f;
  const float* zb = z + ((size_t)b * S * D + (size_t)s * D) * HW;
  const float* gb = gc + ((size_t)b * S * mmax + (size_t)s * mmax) * HW;
  for (int i0 = 0; i0 < n; i0 += 32) {
    const int i = i0 + lane;
    const int x = i < n ? ix[i] : -1;
    for (int cc = warp; cc < D; cc += nw) zt[cc * 33 + lane] = x >= 0 ? zb[(size_t)cc * HW + x] : 0.0f;
    for (int j = warp; j < m; j += nw) gt[j * 33 + lane] = x >= 0 ? gb[(size_t)j * HW + x] : 0.0f;
    __syncthreads();
    if (ch < D) {
#pragma unroll 4
      for (int t = 0; t < 32; ++t) {
        const float zv = zt[ch * 33 + t];
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const int j = jq + DP_JQ * e;
          if (j < m) {
            const float gv = gt[j * 33 + t];
            sgz[e] = fmaf(gv, zv, sgz[e]);
            sg[e] += gv;
          }
        }
      }
    }
    __syncthreads();
  }
  if (ch < D) {
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int j = jq + DP_JQ * e;
      if (j < m) {
        const int p = members[m0 + j];
        part[((size_t)b * P + p) * D + ch] = 2.0f * (protos[(size_t)p * D + ch] * sg[e] - sgz[e]);
      }
    }
  }
}

// dprotos[p][ch] = sum_b part[b][p][ch] (fixed order); prototypes outside every group get 0.
__global__ void kld_feat_dp_reduce_kernel(const float* __restrict__ part, int B, long long PD,
                                          const int* __restrict__ proto_class, int D, float* __restrict__ dprotos) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= PD) return;
  float a = 0.0f;
  if (proto_class[e / D] >= 0)
    for (int b = 0; b < B; ++b) a += part[(size_t)b * PD + e];
  dprotos[e] = a;
}

inline size_t align16(size_t v) { return (v + 15) & ~(size_t)15; }

struct KldCarve {
  int *cnt, *off, *idx;
  float *lse, *term, *fin;
  float2* kl;
  float *dc, *gc, *part;  // feature mode only
  size_t bytes;
};

KldCarve carve(const HeadShape& s, int Q, void* ws, int feat_rows = 0) {
  char* base = reinterpret_cast<char*>(ws);
  size_t o = 0;
  KldCarve k;
  auto take = [&](size_t bytes) {
    char* p = base ? base + o : nullptr;
    o += align16(bytes);
    return p;
  };
  k.fin = reinterpret_cast<float*>(take(4 * sizeof(float)));
  k.cnt = reinterpret_cast<int*>(take(sizeof(int) * (size_t)s.B * s.K));
  k.off = reinterpret_cast<int*>(take(sizeof(int) * (size_t)s.B * s.K));
  k.idx = reinterpret_cast<int*>(take(sizeof(int) * (size_t)s.N));
  k.lse = reinterpret_cast<float*>(take(sizeof(float) * (size_t)s.B * s.P));
  k.term = reinterpret_cast<float*>(take(sizeof(float) * (size_t)s.B * Q));
  k.kl = reinterpret_cast<float2*>(take(sizeof(float2) * (size_t)s.B * Q));
  k.dc = k.gc = k.part = nullptr;
  if (feat_rows > 0) {
    k.dc = reinterpret_cast<float*>(take(sizeof(float) * (size_t)s.B * feat_rows * s.HW));
    k.gc = reinterpret_cast<float*>(take(sizeof(float) * (size_t)s.B * feat_rows * s.HW));
    k.part = reinterpret_cast<float*>(take(sizeof(float) * (size_t)s.B * s.P * s.D));
  }
  k.bytes = o;
  return k;
}

inline unsigned warps_grid(long long warps) { return (unsigned)((warps * 32 + KLD_THREADS - 1) / KLD_THREADS); }

}  // namespace

size_t kld_workspace(const HeadShape& s, int Q) { return carve(s, Q, nullptr).bytes; }

namespace {

RowMap dense_rows(const HeadShape& s) { return RowMap{0, s.P, 1, 1, nullptr, nullptr}; }
RowMap compact_rows(const HeadShape& s, const SpsKldDesc& k) {
  return RowMap{1, s.S * k.mmax, s.S, k.mmax, k.proto_group, k.proto_local};
}

void launch_buckets(const HeadShape& s, const int16_t* labels, const KldCarve& w, cudaStream_t st) {
  kld_hist_kernel<<<s.B, KLD_THREADS, sizeof(int) * s.K, st>>>(labels, s.HW, s.K, w.cnt, w.off);
  kld_bucket_kernel<<<dim3(s.K, s.B), KLD_THREADS, 0, st>>>(labels, s.HW, s.K, w.cnt, w.off, w.idx);
  count_launch(2);
}

// lse -> pairs -> mean over a dense or compact distance block; copies (loss, count) out
int launch_stats(const HeadShape& s, const SpsKldDesc& k, const float* dist, RowMap rm, const KldCarve& w, float* loss,
                 cudaStream_t st) {
  kld_lse_kernel<<<warps_grid((long long)s.B * s.P), KLD_THREADS, 0, st>>>(dist, rm, s.B, s.P, s.HW, s.K,
                                                                          k.proto_class, w.cnt, w.off, w.idx, w.lse);
  int launched = 1;
  if (k.Q > 0) {
    kld_pairs_kernel<<<warps_grid((long long)s.B * k.Q), KLD_THREADS, 0, st>>>(dist, rm, s.B, s.P, s.HW, s.K, k.Q,
                                                                              k.pair_a, k.pair_b, k.proto_class, w.cnt,
                                                                              w.off, w.idx, w.lse, w.kl, w.term);
    ++launched;
  }
  kld_finish_kernel<<<1, 1024, 0, st>>>(w.term, (long long)s.B * k.Q, w.fin);
  count_launch(launched + 1);
  SPS_CUDA_TRY(cudaMemcpyAsync(loss, w.fin, 2 * sizeof(float), cudaMemcpyDeviceToDevice, st));
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

}  // namespace

int launch_kld_fwd(const HeadShape& s, const SpsKldDesc& k, const float* dist, const int16_t* labels, void* ws,
                   float* loss, cudaStream_t st) {
  KldCarve w = carve(s, k.Q, ws);
  launch_buckets(s, labels, w, st);
  return launch_stats(s, k, dist, dense_rows(s), w, loss, st);
}

int launch_kld_bwd(const HeadShape& s, const SpsKldDesc& k, const float* dist, void* ws, const float* dloss,
                   float* ddist, cudaStream_t st) {
  KldCarve w = carve(s, k.Q, ws);
  SPS_CUDA_TRY(cudaMemsetAsync(ddist, 0, sizeof(float) * (size_t)s.N * s.P, st));
  if (k.Q > 0) {
    kld_bwd_kernel<<<dim3(k.G, s.B), BWD_THREADS, kld_bwd_smem(k.mmax), st>>>(
        dist, dense_rows(s), s.B, s.P, s.HW, s.K, k.G / s.K, k.Q, k.group_moff, k.group_poff, k.members, w.cnt, w.off,
        w.idx, w.lse, w.kl, w.term, w.fin, dloss, ddist);
    count_launch(1);
  }
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

size_t kld_feat_workspace(const HeadShape& s, const SpsKldDesc& k) { return carve(s, k.Q, nullptr, s.S * k.mmax).bytes; }

int launch_kld_feat_fwd(const HeadShape& s, const SpsKldDesc& k, const float* z, const float* protos,
                        const int16_t* labels, void* ws, float* loss, cudaStream_t st) {
  KldCarve w = carve(s, k.Q, ws, s.S * k.mmax);
  launch_buckets(s, labels, w, st);
  dim3 grid((s.HW + FEAT_THREADS - 1) / FEAT_THREADS, s.S, s.B);
  kld_feat_dist_kernel<<<grid, FEAT_THREADS, 0, st>>>(z, protos, labels, s.HW, s.K, s.S, s.D, k.mmax, k.group_moff,
                                                      k.members, w.dc);
  count_launch(1);
  return launch_stats(s, k, w.dc, compact_rows(s, k), w, loss, st);
}

int launch_kld_feat_bwd(const HeadShape& s, const SpsKldDesc& k, const float* z, const float* protos,
                        const int16_t* labels, void* ws, const float* dloss, float* dz, float* dprotos,
                        cudaStream_t st) {
  const int rows = s.S * k.mmax;
  KldCarve w = carve(s, k.Q, ws, rows);
  SPS_CUDA_TRY(cudaMemsetAsync(w.gc, 0, sizeof(float) * (size_t)s.B * rows * s.HW, st));
  if (k.Q > 0) {
    kld_bwd_kernel<<<dim3(k.G, s.B), BWD_THREADS, kld_bwd_smem(k.mmax), st>>>(
        w.dc, compact_rows(s, k), s.B, s.P, s.HW, s.K, s.S, k.Q, k.group_moff, k.group_poff, k.members, w.cnt, w.off,
        w.idx, w.lse, w.kl, w.term, w.fin, dloss, w.gc);
    count_launch(1);
  }
  if (dz != nullptr) {
    dim3 grid((s.HW + FEAT_THREADS - 1) / FEAT_THREADS, s.S * ((s.D + DZ_SLICE - 1) / DZ_SLICE), s.B);
    kld_feat_dz_kernel<<<grid, FEAT_THREADS, 0, st>>>(z, protos, labels, s.HW, s.K, s.S, s.D, k.mmax, k.group_moff,
                                                      k.members, w.gc, dz);
    count_launch(1);
  }
  if (dprotos != nullptr) {
    SPS_CUDA_TRY(cudaMemsetAsync(w.part, 0, sizeof(float) * (size_t)s.B * s.P * s.D, st));
    SPS_REQUIRE(k.mmax <= DP_JQ * DP_MAXE && s.D <= FEAT_D, SPS_ERR_UNSUPPORTED,
                "kld feature mode: group size %d > %d or depth %d > %d", k.mmax, DP_JQ * DP_MAXE, s.D, FEAT_D);
    const size_t smem = sizeof(float) * 33 * (size_t)(s.D + k.mmax);
    const dim3 grid(k.G, s.B);
#define SPS_DP_LAUNCH(E)                                                                                              \
  kld_feat_dp_kernel<E><<<grid, DP_THREADS, smem, st>>>(z, protos, s.HW, s.K, s.S, s.D, s.P, k.mmax, k.group_moff, \
                                                        k.members, w.cnt, w.off, w.idx, w.gc, w.part)
    switch ((k.mmax + DP_JQ - 1) / DP_JQ) {
      case 1: SPS_DP_LAUNCH(1); break;
      case 2: SPS_DP_LAUNCH(2); break;
      case 3: SPS_DP_LAUNCH(3); break;
      case 4: SPS_DP_LAUNCH(4); break;
      case 5: SPS_DP_LAUNCH(5); break;
      case 6: SPS_DP_LAUNCH(6); break;
      case 7: SPS_DP_LAUNCH(7); break;
      default: SPS_DP_LAUNCH(8); break;
    }
#undef SPS_DP_LAUNCH
    const long long PD = (long long)s.P * s.D;
    kld_feat_dp_reduce_kernel<<<(unsigned)((PD + 255) / 256), 256, 0, st>>>(w.part, s.B, PD, k.proto_class, s.D, dprotos);
    count_launch(2);
  }
  SPS_CUDA_TRY(cudaGetLastError());
  return SPS_OK;
}

}  // namespace sps
