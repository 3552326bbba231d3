// csi_nuts.cuh (included at the end of csi_api.cu) -- the No-U-Turn sampler around the potential kernel, with the
// tree logic on the device.
//
// The reference drives its models with numpyro's NUTS (scripts/model_star.py:132-155; numpyro/infer/hmc_util.py is
// third-party): multinomial NUTS, iterative sub-tree building with checkpointed U-turn tests
// (_iterative_build_subtree / _leaf_idx_to_ckpt_idxs / _is_iterative_turning), biased progressive sampling between
// doublings (_combine_tree).  Here one leapfrog is three stream-ordered launches
//
//   nuts_pre   : [lazy copy of the previous leaf into the sub-tree proposal]  r -= eps/2 g ;  z += eps M^-1 r
//   potential  : g = dU/dz(z), U                       (csi_tc.cu / csi_ffma.cu; the 6 sums exchanged in-kernel)
//   nuts_post  : r -= eps/2 g ; kinetic energy ; r_sum ; checkpoints ; U-turn inner products against the
//                checkpoints ; grid reduce (+ exchange over NVLink when sharded) ; leaf weight, multinomial
//                accept, divergence and U-turn decisions by the finalising block
//
// applied IN PLACE to the end of the trajectory that is being extended, and a doubling of depth d is ONE CUDA graph
// of 2^d such triples + nuts_merge (sub-tree into tree, outer U-turn test).  All decisions are taken on the device
// from a counter-based hash (the same stream on every rank); a sub-tree that turns or diverges sets a stop word and
// the launches queued behind it return at once (Launch::stop).  The host synchronises once per doubling, to learn
// whether to queue the next one.  Every sum is a fixed-order grid reduce: bit-reproducible runs.
//
// Vectors are float[len], len = n_hyper + n_sites * n_local in csi_layout order.  Sharded: the hyper entries are
// replicated (only rank 0 adds them to the inner products), the star entries are summed over the ranks through the
// peer exchange of csi_reduce.cuh; the momentum of element i depends on its GLOBAL index only, so a run does not
// depend on how the stars are sharded (up to the rounding of the reductions).
#pragma once
#include "csi_internal.h"

namespace csi {
namespace nuts {

constexpr int kBlock = 256;
constexpr int kMaxDepth = 12;
constexpr int kNV = 1 + 2 * kMaxDepth;  // kinetic energy + two inner products per checkpoint
static_assert(kNV < kPxSlot, "exchange slot");

// ---- counter-based random numbers (host and device; the Python mirror is nuts.hash_u64) -------------------------
__host__ __device__ inline unsigned long long mix64(unsigned long long x) {
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return x;
}
enum Kind : unsigned long long { K_MOMENTUM = 1, K_DIRECTION = 2, K_LEAF = 3, K_MERGE = 4 };
__host__ __device__ inline unsigned long long hash_u64(unsigned long long seed, unsigned long long draw,
                                                       unsigned long long kind, unsigned long long idx) {
  unsigned long long h = mix64(seed + 0x9E3779B97F4A7C15ull);
  h = mix64(h ^ (draw + 0x9E3779B97F4A7C15ull));
  h = mix64(h ^ (kind * 0xD1B54A32D192ED03ull));
  return mix64(h ^ (idx + 0x9E3779B97F4A7C15ull));
}
__host__ __device__ inline double hash_uniform(unsigned long long seed, unsigned long long draw,
                                               unsigned long long kind, unsigned long long idx) {
  return (double)(hash_u64(seed, draw, kind, idx) >> 11) * (1.0 / 9007199254740992.0);  // [0, 1)
}

struct Scalars {  // device copy + pinned host mirror
  double h0, u_prop, u_sub, u_end, kin_end, logw_tree, logw_sub, sum_accept, eps;
  long long n_prop;
  unsigned long long seed, draw;
  int depth, stop, done, turning, diverging, pending_take, sub_turning, pad;
};

struct Vecs {  // everything a launch needs (passed by value)
  long long len, n_local, n_total, offset;
  int nh, add_hyper;  // add_hyper: this rank adds the (replicated) hyper entries to the sums
  float *z, *g;       // current state = tree proposal (caller's buffers)
  const float *im;    // inverse mass (caller's buffer)
  float *zl, *rl, *gl, *zr, *rr, *gr, *z_sub, *g_sub, *r_sum, *r_sum_sub;
  float *ck_r[kMaxDepth], *ck_rs[kMaxDepth];
  Scalars *sc;
  double *partials;       // [grid][kNV]
  unsigned int *counter;
  const double *out6;     // the potential's sums (U at the new leaf)
  PeerX px;
  int max_depth;
  double max_de;
};

__device__ __forceinline__ double block_sum(double v, double *sh) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x < kBlock / 32) t = sh[threadIdx.x];
  if (warp == 0) t = warp_sum(t);
  return t;  // valid on thread 0
}

// vals[0..nv) of this block (thread 0 holds them in shared memory) -> fixed-order sum over the blocks -> sum over
// the ranks; returns true in the block that holds the totals in tot[] (the last one to arrive).
__device__ __forceinline__ bool grid_reduce(const Vecs &V, const double *vals_sh, int nv, double *tot) {
  __shared__ bool is_last;
  const int tid = threadIdx.x;
  __syncthreads();
  if (tid < nv) V.partials[(size_t)blockIdx.x * kNV + tid] = vals_sh[tid];
  __threadfence();
  __syncthreads();
  if (tid == 0) is_last = atomicAdd(V.counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return false;
  __threadfence();
  if (tid < kNV) {
    double s = 0.0;
    if (tid < nv) {
      const volatile double *p = V.partials;
      for (unsigned int b = 0; b < gridDim.x; ++b) s += p[(size_t)b * kNV + tid];
    }
    tot[tid] = s;
  }
  __syncthreads();
  if (V.px.world > 1) peer_allreduce<kNV>(V.px, tot);
  if (tid == 0) *V.counter = 0u;
  __syncthreads();
  return true;
}

__device__ __forceinline__ double logaddexp(double a, double b) {
  const double m = fmax(a, b);
  if (m == -INFINITY) return m;
  return m + log(exp(a - m) + exp(b - m));
}

// global index of element i (momentum stream): hyper entries first, then [site][global star]
__device__ __forceinline__ unsigned long long global_index(const Vecs &V, long long i) {
  if (i < V.nh) return (unsigned long long)i;
  const long long j = i - V.nh, s = j / V.n_local, k = j - s * V.n_local;
  return (unsigned long long)(V.nh + s * V.n_total + V.offset + k);
}

#define CSI_NUTS_LOOP(i) \
  for (long long i = (long long)blockIdx.x * kBlock + threadIdx.x; i < V.len; i += (long long)gridDim.x * kBlock)

// ---- start of a transition: momentum, both ends = the current state, initial energy -----------------------------
__global__ void __launch_bounds__(kBlock) nuts_begin_kernel(const __grid_constant__ Vecs V, unsigned long long seed,
                                                            unsigned long long draw, double eps, double u_cur,
                                                            int use_u_arg) {
  __shared__ double sh[kBlock / 32], vals[kNV], tot[kNV];
  double k = 0.0;
  CSI_NUTS_LOOP(i) {
    const unsigned long long x = hash_u64(seed, draw, K_MOMENTUM, global_index(V, i));
    const float u1 = ((float)(x >> 40) + 0.5f) * (1.0f / 16777216.0f);             // (0, 1), 24 bits
    const float u2 = ((float)((x >> 8) & 0xffffffull) + 0.5f) * (1.0f / 16777216.0f);
    const float im = V.im[i];
    const float r = sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2) * rsqrtf(im);      // r ~ N(0, M), M = 1 / im
    const float z = V.z[i], g = V.g[i];
    V.zl[i] = z; V.zr[i] = z; V.gl[i] = g; V.gr[i] = g; V.rl[i] = r; V.rr[i] = r; V.r_sum[i] = r;
    if (i >= V.nh || V.add_hyper) k += (double)im * (double)r * (double)r;
  }
  const double t = block_sum(k, sh);
  if (threadIdx.x == 0) vals[0] = t;
  if (!grid_reduce(V, vals, 1, tot)) return;
  if (threadIdx.x == 0) {
    Scalars &s = *V.sc;
    if (use_u_arg) s.u_prop = u_cur;
    s.h0 = s.u_prop + 0.5 * tot[0];
    s.logw_tree = 0.0; s.logw_sub = -INFINITY; s.sum_accept = 0.0; s.n_prop = 0; s.eps = eps;
    s.seed = seed; s.draw = draw;
    s.depth = 0; s.stop = 0; s.done = 0; s.turning = 0; s.diverging = 0; s.pending_take = 0; s.sub_turning = 0;
    s.u_sub = s.u_prop; s.u_end = s.u_prop; s.kin_end = 0.5 * tot[0];
  }
}

// ---- first half of a leapfrog on the end (z, r, g) ----------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) nuts_pre_kernel(const __grid_constant__ Vecs V, float *z, float *r,
                                                          const float *g, float sign) {
  const Scalars &s = *V.sc;
  if (*reinterpret_cast<const volatile int *>(&s.stop) != 0) return;
  const bool take = s.pending_take != 0;  // the previous leaf was accepted as the sub-tree proposal: keep it
  const float eps = sign * (float)s.eps, he = 0.5f * eps;
  CSI_NUTS_LOOP(i) {
    const float zi = z[i], gi = g[i];
    if (take) { V.z_sub[i] = zi; V.g_sub[i] = gi; }
    const float rh = r[i] - he * gi;
    r[i] = rh;
    z[i] = zi + eps * (V.im[i] * rh);
  }
}

// ---- second half + tree bookkeeping of leaf `leaf` of the sub-tree --------------------------------------------------
__global__ void __launch_bounds__(kBlock) nuts_post_kernel(const __grid_constant__ Vecs V, float *r, const float *g,
                                                           float sign, int leaf, int idx_min, int idx_max, int depth) {
  __shared__ double sh[kBlock / 32], vals[kNV], tot[kNV];
  Scalars &s = *V.sc;
  if (*reinterpret_cast<const volatile int *>(&s.stop) != 0) return;
  const float he = 0.5f * sign * (float)s.eps;
  const bool even = (leaf & 1) == 0;
  float *ckr = V.ck_r[even ? idx_max : 0], *ckrs = V.ck_rs[even ? idx_max : 0];
  double k = 0.0;
  CSI_NUTS_LOOP(i) {
    const float rv = r[i] - he * g[i];
    r[i] = rv;
    const float rs = (leaf == 0 ? 0.0f : V.r_sum_sub[i]) + rv;
    V.r_sum_sub[i] = rs;
    if (even) { ckr[i] = rv; ckrs[i] = rs; }
    if (i >= V.nh || V.add_hyper) k += (double)V.im[i] * (double)rv * (double)rv;
  }
  double t = block_sum(k, sh);
  if (threadIdx.x == 0) vals[0] = t;
  // U-turn tests of the sub-trees that end at this (odd) leaf: checkpoints idx_max .. idx_min
  // (numpyro hmc_util._is_iterative_turning); each thread re-reads the r / r_sum it has just written
  int nv = 1;
  for (int c = idx_max; c >= idx_min; --c) {
    const float *cr = V.ck_r[c], *crs = V.ck_rs[c];
    double a = 0.0, b = 0.0;
    CSI_NUTS_LOOP(i) {
      if (i >= V.nh || V.add_hyper) {
        const float rl = cr[i], rr = r[i];
        const float sub = V.r_sum_sub[i] - crs[i] + rl;
        const double m = (double)V.im[i], rc = (double)(sub - 0.5f * (rl + rr));
        a += m * (double)rl * rc;
        b += m * (double)rr * rc;
      }
    }
    t = block_sum(a, sh);
    if (threadIdx.x == 0) vals[nv] = t;
    t = block_sum(b, sh);
    if (threadIdx.x == 0) vals[nv + 1] = t;
    nv += 2;
  }
  if (!grid_reduce(V, vals, nv, tot)) return;
  if (threadIdx.x == 0) {
    const double u1 = V.out6[0], kin = 0.5 * tot[0];
    double de = (u1 + kin) - s.h0;
    if (!(de == de) || isinf(de)) de = INFINITY;           // NaN / inf energy: zero weight, divergence
    const double w = -de;
    const bool diverging = de > V.max_de;
    s.sum_accept += de < INFINITY ? fmin(1.0, exp(-de)) : 0.0;
    s.n_prop += 1;
    bool take;
    if (leaf == 0) {
      take = true;
      s.logw_sub = w;
    } else {  // uniform progressive sampling inside the sub-tree: P(new leaf) = w / (w_sub + w)
      const double p = w == -INFINITY ? 0.0 : 1.0 / (1.0 + exp(s.logw_sub - w));
      take = hash_uniform(s.seed, s.draw, K_LEAF, ((unsigned long long)depth << 32) | (unsigned)leaf) < p;
      s.logw_sub = logaddexp(s.logw_sub, w);
    }
    s.pending_take = take ? 1 : 0;
    if (take) s.u_sub = u1;
    s.u_end = u1; s.kin_end = kin;
    bool turning = false;
    for (int j = 1; j < nv; j += 2) turning = turning || tot[j] <= 0.0 || tot[j + 1] <= 0.0;
    if (diverging) s.diverging = 1;
    if (turning) s.sub_turning = 1;
    if (diverging || turning) { __threadfence(); s.stop = 1; }
  }
}

// ---- end of a doubling: sub-tree into tree (biased progressive sampling), outer U-turn test ------------------------
__global__ void __launch_bounds__(kBlock) nuts_merge_kernel(const __grid_constant__ Vecs V, const float *z_end,
                                                            const float *g_end, int depth) {
  __shared__ double sh[kBlock / 32], vals[kNV], tot[kNV];
  Scalars &s = *V.sc;
  if (s.done != 0) return;  // a doubling queued behind the end of the tree
  const bool valid = s.stop == 0;
  bool take = false;
  if (valid) {
    const double p = fmin(1.0, exp(s.logw_sub - s.logw_tree));
    take = hash_uniform(s.seed, s.draw, K_MERGE, (unsigned long long)depth) < p;
  }
  const bool pending = s.pending_take != 0;
  const float *src_z = pending ? z_end : V.z_sub, *src_g = pending ? g_end : V.g_sub;
  double a = 0.0, b = 0.0;
  if (valid) {
    CSI_NUTS_LOOP(i) {
      if (take) { V.z[i] = src_z[i]; V.g[i] = src_g[i]; }
      const float rs = V.r_sum[i] + V.r_sum_sub[i];
      V.r_sum[i] = rs;
      if (i >= V.nh || V.add_hyper) {
        const float rl = V.rl[i], rr = V.rr[i];
        const double m = (double)V.im[i], rc = (double)(rs - 0.5f * (rl + rr));
        a += m * (double)rl * rc;
        b += m * (double)rr * rc;
      }
    }
  }
  double t = block_sum(a, sh);
  if (threadIdx.x == 0) vals[0] = t;
  t = block_sum(b, sh);
  if (threadIdx.x == 0) vals[1] = t;
  if (!grid_reduce(V, vals, 2, tot)) return;
  if (threadIdx.x == 0) {
    if (!valid) {  // the sub-tree turned or diverged: no proposal is taken from it, the tree ends here
      if (s.sub_turning) s.turning = 1;
      s.depth = depth + 1;
      s.done = 1;
    } else {
      if (take) s.u_prop = s.u_sub;
      s.logw_tree = logaddexp(s.logw_tree, s.logw_sub);
      s.depth = depth + 1;
      const bool turning = tot[0] <= 0.0 || tot[1] <= 0.0;
      if (turning) s.turning = 1;
      if (turning || s.depth >= V.max_depth) s.done = 1;
      s.pending_take = 0;
    }
  }
}

// ---- deterministic versions of the single-step vector kernels (csi_leapfrog_pre / _post, csi_uturn) ---------------
__global__ void __launch_bounds__(kBlock) leapfrog_pre_kernel(long long n, const float *__restrict__ z,
                                                              const float *__restrict__ r, const float *__restrict__ g,
                                                              const float *__restrict__ im, float eps,
                                                              float *__restrict__ z1, float *__restrict__ r_half) {
  const float he = 0.5f * eps;
  for (long long i = (long long)blockIdx.x * kBlock + threadIdx.x; i < n; i += (long long)gridDim.x * kBlock) {
    const float rh = r[i] - he * g[i];
    r_half[i] = rh;
    z1[i] = z[i] + eps * (im[i] * rh);
  }
}

// vals -> partials -> fixed-order sum by the last block into out[0..nv) (no exchange: these entry points return the
// LOCAL star sums and the hyper sums apart, as documented in csi.h)
__device__ __forceinline__ void reduce_to(const double *vals_sh, int nv, double *partials, unsigned int *counter,
                                          double *out) {
  __shared__ bool is_last;
  const int tid = threadIdx.x;
  __syncthreads();
  if (tid < nv) partials[(size_t)blockIdx.x * kNV + tid] = vals_sh[tid];
  __threadfence();
  __syncthreads();
  if (tid == 0) is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (tid < nv) {
    double s = 0.0;
    const volatile double *p = partials;
    for (unsigned int b = 0; b < gridDim.x; ++b) s += p[(size_t)b * kNV + tid];
    out[tid] = s;
  }
  if (tid == 0) *counter = 0u;
}

__global__ void __launch_bounds__(kBlock) leapfrog_post_kernel(long long n, int nh, const float *__restrict__ r_half,
                                                               const float *__restrict__ grad,
                                                               const double *__restrict__ out6,
                                                               const float *__restrict__ im, float eps,
                                                               float *__restrict__ r1, float *__restrict__ g1,
                                                               double *kin, double *partials, unsigned int *counter) {
  __shared__ double sh[kBlock / 32], vals[2];
  const float he = 0.5f * eps;
  double ks = 0.0, kh = 0.0;
  for (long long i = (long long)blockIdx.x * kBlock + threadIdx.x; i < n; i += (long long)gridDim.x * kBlock) {
    const float gi = i < nh ? (float)out6[1 + i] : grad[i];  // hyper slots: the reduced sums, not the local partials
    const float rv = r_half[i] - he * gi;
    g1[i] = gi;
    r1[i] = rv;
    const double e = (double)im[i] * (double)rv * (double)rv;
    if (i < nh) kh += e; else ks += e;
  }
  double t = block_sum(ks, sh);
  if (threadIdx.x == 0) vals[0] = t;
  t = block_sum(kh, sh);
  if (threadIdx.x == 0) vals[1] = t;
  reduce_to(vals, 2, partials, counter, kin);
}

__global__ void __launch_bounds__(kBlock) uturn_kernel(long long n, int nh, const float *__restrict__ rl,
                                                       const float *__restrict__ rr, const float *__restrict__ rs,
                                                       const float *__restrict__ im, double *dots, double *partials,
                                                       unsigned int *counter) {
  __shared__ double sh[kBlock / 32], vals[4];
  double a = 0.0, b = 0.0, ah = 0.0, bh = 0.0;
  for (long long i = (long long)blockIdx.x * kBlock + threadIdx.x; i < n; i += (long long)gridDim.x * kBlock) {
    const float l = rl[i], r = rr[i];
    const float rc = rs[i] - 0.5f * (l + r);
    const double m = (double)im[i], c = (double)rc;
    if (i < nh) { ah += m * l * c; bh += m * r * c; } else { a += m * l * c; b += m * r * c; }
  }
  double t = block_sum(a, sh);
  if (threadIdx.x == 0) vals[0] = t;
  t = block_sum(b, sh);
  if (threadIdx.x == 0) vals[1] = t;
  t = block_sum(ah, sh);
  if (threadIdx.x == 0) vals[2] = t;
  t = block_sum(bh, sh);
  if (threadIdx.x == 0) vals[3] = t;
  reduce_to(vals, 4, partials, counter, dots);
}

static int grid_for(const csi_handle *h, long long n) {
  const long long want = (n + kBlock - 1) / kBlock;
  return (int)std::max<long long>(1, std::min<long long>(want, (long long)h->sm_count * 4));
}

}  // namespace nuts
}  // namespace csi

// ---------------------------------------------------------------------------------------------------------------
// single-step entry points (kept from round 1; sums are now fixed-order grid reductions, no atomics on doubles)
// ---------------------------------------------------------------------------------------------------------------
static int vec_scratch(csi_handle *h) {  // partial sums of the vector kernels: [4 * SMs][kNV] doubles + a ticket
  if (h->d_vec_partials) return CSI_OK;
  CK(cudaMalloc(&h->d_vec_partials, (size_t)h->sm_count * 4 * nuts::kNV * sizeof(double)));
  CK(cudaMalloc(&h->d_vec_counter, sizeof(unsigned int)));
  CK(cudaMemset(h->d_vec_counter, 0, sizeof(unsigned int)));
  h->owned.push_back(h->d_vec_partials);
  h->owned.push_back(h->d_vec_counter);
  return CSI_OK;
}

extern "C" int csi_leapfrog_pre(csi_handle *h, const float *z, const float *r, const float *g, const float *inv_mass,
                                float eps, float *z1, float *r_half, double *kin2, void *stream) {
  if (!h || !z || !r || !g || !inv_mass || !z1 || !r_half || !kin2) return fail(CSI_EINVAL, "null argument");
  if (h->n_local < 1) return fail(CSI_ESTATE, "csi_set_data has not been called");
  if (int rcd = check_device(h)) return rcd;
  const long long n = h->stage_len;
  nuts::leapfrog_pre_kernel<<<nuts::grid_for(h, n), nuts::kBlock, 0, (cudaStream_t)stream>>>(n, z, r, g, inv_mass, eps,
                                                                                            z1, r_half);
  CK(cudaGetLastError());
  return CSI_OK;
}

extern "C" int csi_leapfrog_post(csi_handle *h, const float *r_half, const float *grad, const double *out6,
                                 const float *inv_mass, float eps, float *r1, float *g1, double *kin2, void *stream) {
  if (!h || !r_half || !grad || !out6 || !inv_mass || !r1 || !g1 || !kin2) return fail(CSI_EINVAL, "null argument");
  if (h->n_local < 1) return fail(CSI_ESTATE, "csi_set_data has not been called");
  if (int rcd = check_device(h)) return rcd;
  if (int rcs = vec_scratch(h)) return rcs;
  const long long n = h->stage_len;
  nuts::leapfrog_post_kernel<<<nuts::grid_for(h, n), nuts::kBlock, 0, (cudaStream_t)stream>>>(
      n, h->n_hyper, r_half, grad, out6, inv_mass, eps, r1, g1, kin2, h->d_vec_partials, h->d_vec_counter);
  CK(cudaGetLastError());
  return CSI_OK;
}

extern "C" int csi_uturn(csi_handle *h, const float *r_left, const float *r_right, const float *r_sum,
                         const float *inv_mass, double *dots4, void *stream) {
  if (!h || !r_left || !r_right || !r_sum || !inv_mass || !dots4) return fail(CSI_EINVAL, "null argument");
  if (h->n_local < 1) return fail(CSI_ESTATE, "csi_set_data has not been called");
  if (int rcd = check_device(h)) return rcd;
  if (int rcs = vec_scratch(h)) return rcs;
  const long long n = h->stage_len;
  nuts::uturn_kernel<<<nuts::grid_for(h, n), nuts::kBlock, 0, (cudaStream_t)stream>>>(
      n, h->n_hyper, r_left, r_right, r_sum, inv_mass, dots4, h->d_vec_partials, h->d_vec_counter);
  CK(cudaGetLastError());
  return CSI_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// the sampler object
// ---------------------------------------------------------------------------------------------------------------
struct csi_nuts {
  csi_handle *h = nullptr;
  csi::nuts::Vecs V{};
  std::vector<void *> owned;
  csi::nuts::Scalars *h_sc = nullptr;  // pinned mirror
  double *d_out = nullptr;
  cudaStream_t stream = nullptr;
  cudaGraphExec_t exec[csi::nuts::kMaxDepth][2] = {{nullptr}};
  unsigned long long seed = 0;
  int grid = 1;
  bool use_graphs = true;
  long long launches = 0;  // kernel launches queued by the last transition (graph nodes counted)
};

static int nuts_alloc(csi_nuts *s, size_t bytes, void **out) {
  CK(cudaMalloc(out, std::max<size_t>(bytes, 16)));
  s->owned.push_back(*out);
  return CSI_OK;
}

extern "C" int csi_nuts_create(csi_handle *h, float *z_state, float *g_state, const float *inv_mass,
                               int32_t max_tree_depth, double max_delta_energy, uint64_t seed, int64_t star_offset,
                               int64_t n_total, csi_nuts **out) {
  if (!h || !z_state || !g_state || !inv_mass || !out) return fail(CSI_EINVAL, "null argument");
  *out = nullptr;
  if (h->n_local < 1) return fail(CSI_ESTATE, "csi_set_data has not been called");
  if (max_tree_depth < 1 || max_tree_depth > csi::nuts::kMaxDepth) return fail(CSI_EINVAL, "max_tree_depth must be 1..12");
  if (h->world > 1 && h->px.world != h->world)
    return fail(CSI_ESTATE, "a sharded sampler needs csi_set_peer_exchange: its inner products use the in-kernel exchange");
  if (star_offset < 0 || n_total < star_offset + h->n_local) return fail(CSI_EINVAL, "star_offset / n_total");
  CK(cudaSetDevice(h->device));
  csi_nuts *s = new csi_nuts();
  s->h = h; s->seed = seed;
  using namespace csi::nuts;
  Vecs &V = s->V;
  V.len = h->stage_len; V.n_local = h->n_local; V.n_total = n_total; V.offset = star_offset;
  V.nh = h->n_hyper; V.add_hyper = h->rank == 0 ? 1 : 0;
  V.z = z_state; V.g = g_state; V.im = inv_mass;
  V.px = h->px; V.max_depth = max_tree_depth; V.max_de = max_delta_energy;
  int rc = CSI_OK;
  const size_t vb = (size_t)V.len * sizeof(float);
  float **vecs[] = {&V.zl, &V.rl, &V.gl, &V.zr, &V.rr, &V.gr, &V.z_sub, &V.g_sub, &V.r_sum, &V.r_sum_sub};
  for (float **p : vecs)
    if (rc == CSI_OK) rc = nuts_alloc(s, vb, (void **)p);
  for (int d = 0; d < kMaxDepth; ++d) {  // unused checkpoint slots alias slot 0 (never addressed)
    if (d < max_tree_depth) {
      if (rc == CSI_OK) rc = nuts_alloc(s, vb, (void **)&V.ck_r[d]);
      if (rc == CSI_OK) rc = nuts_alloc(s, vb, (void **)&V.ck_rs[d]);
    } else {
      V.ck_r[d] = V.ck_r[0]; V.ck_rs[d] = V.ck_rs[0];
    }
  }
  s->grid = grid_for(h, V.len);
  if (rc == CSI_OK) rc = nuts_alloc(s, sizeof(Scalars), (void **)&V.sc);
  if (rc == CSI_OK) rc = nuts_alloc(s, (size_t)s->grid * kNV * sizeof(double), (void **)&V.partials);
  if (rc == CSI_OK) rc = nuts_alloc(s, sizeof(unsigned int), (void **)&V.counter);
  if (rc == CSI_OK) rc = nuts_alloc(s, 8 * sizeof(double), (void **)&s->d_out);
  if (rc == CSI_OK) {
    V.out6 = s->d_out;
    if (cudaMemset(V.sc, 0, sizeof(Scalars)) != cudaSuccess || cudaMemset(V.counter, 0, sizeof(unsigned int)) != cudaSuccess ||
        cudaMallocHost(&s->h_sc, sizeof(Scalars)) != cudaSuccess ||
        cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess)
      rc = fail(CSI_ECUDA, "sampler set-up");
  }
  s->use_graphs = !(getenv("CSI_NUTS_NO_GRAPH") && atoi(getenv("CSI_NUTS_NO_GRAPH")));
  if (rc != CSI_OK) { std::string keep = g_err; csi_nuts_destroy(s); g_err = keep; return rc; }
  memset(s->h_sc, 0, sizeof(Scalars));
  *out = s;
  return CSI_OK;
}

extern "C" void csi_nuts_destroy(csi_nuts *s) {
  if (!s) return;
  cudaSetDevice(s->h->device);
  if (s->stream) cudaStreamSynchronize(s->stream);
  for (auto &row : s->exec)
    for (cudaGraphExec_t e : row)
      if (e) cudaGraphExecDestroy(e);
  for (void *p : s->owned) cudaFree(p);
  if (s->h_sc) cudaFreeHost(s->h_sc);
  if (s->stream) cudaStreamDestroy(s->stream);
  delete s;
}

// number of non-zero bits above the last one / number of trailing ones (numpyro hmc_util._leaf_idx_to_ckpt_idxs)
static void leaf_to_ckpt(int n, int *idx_min, int *idx_max) {
  int imax = 0;
  for (int m = n >> 1; m > 0; m >>= 1) imax += m & 1;
  int trailing = 0;
  for (int m = n; m & 1; m >>= 1) ++trailing;
  *idx_max = imax;
  *idx_min = imax - trailing + 1;
}

// queue the three launches of leaf `leaf` of a sub-tree of depth `depth` growing to the right / left
static int nuts_queue_leaf(csi_nuts *s, int depth, int leaf, bool right) {
  using namespace csi::nuts;
  const Vecs &V = s->V;
  float *z = right ? V.zr : V.zl, *r = right ? V.rr : V.rl, *g = right ? V.gr : V.gl;
  const float sign = right ? 1.0f : -1.0f;
  nuts_pre_kernel<<<s->grid, kBlock, 0, s->stream>>>(V, z, r, g, sign);
  CK(cudaGetLastError());
  csi_handle *h = s->h;
  h->stop_flag = &V.sc->stop;
  int rc = csi_potential(h, z, s->d_out, g, s->stream);
  h->stop_flag = nullptr;
  if (rc != CSI_OK) return rc;
  int idx_min, idx_max;
  leaf_to_ckpt(leaf, &idx_min, &idx_max);
  nuts_post_kernel<<<s->grid, kBlock, 0, s->stream>>>(V, r, g, sign, leaf, idx_min, idx_max, depth);
  CK(cudaGetLastError());
  s->launches += 3;
  return CSI_OK;
}

static int nuts_queue_doubling(csi_nuts *s, int depth, bool right) {
  using namespace csi::nuts;
  for (int leaf = 0; leaf < (1 << depth); ++leaf)
    if (int rc = nuts_queue_leaf(s, depth, leaf, right)) return rc;
  const Vecs &V = s->V;
  nuts_merge_kernel<<<s->grid, kBlock, 0, s->stream>>>(V, right ? V.zr : V.zl, right ? V.gr : V.gl, depth);
  CK(cudaGetLastError());
  s->launches += 1;
  return CSI_OK;
}

static int nuts_run_doubling(csi_nuts *s, int depth, bool right) {
  if (!s->use_graphs) return nuts_queue_doubling(s, depth, right);
  cudaGraphExec_t &e = s->exec[depth][right ? 1 : 0];
  if (!e) {  // captured once per (depth, direction): the step size and every decision live in device memory
    cudaGraph_t g = nullptr;
    CK(cudaStreamBeginCapture(s->stream, cudaStreamCaptureModeThreadLocal));
    const long long before = s->launches;
    int rc = nuts_queue_doubling(s, depth, right);
    s->launches = before;
    cudaError_t ce = cudaStreamEndCapture(s->stream, &g);
    if (rc != CSI_OK) { if (g) cudaGraphDestroy(g); return rc; }
    if (ce != cudaSuccess) return fail(CSI_ECUDA, std::string("graph capture: ") + cudaGetErrorString(ce));
    ce = cudaGraphInstantiate(&e, g, 0);
    cudaGraphDestroy(g);
    if (ce != cudaSuccess) { e = nullptr; return fail(CSI_ECUDA, std::string("graph instantiate: ") + cudaGetErrorString(ce)); }
  }
  CK(cudaGraphLaunch(e, s->stream));
  s->launches += 3LL * (1 << depth) + 1;
  return CSI_OK;
}

static int nuts_begin(csi_nuts *s, double eps, uint64_t draw, const double *u_cur) {
  using namespace csi::nuts;
  nuts_begin_kernel<<<s->grid, kBlock, 0, s->stream>>>(s->V, s->seed, draw, eps, u_cur ? *u_cur : 0.0, u_cur ? 1 : 0);
  CK(cudaGetLastError());
  s->launches += 1;
  return CSI_OK;
}

static int nuts_fetch(csi_nuts *s) {
  CK(cudaMemcpyAsync(s->h_sc, s->V.sc, sizeof(csi::nuts::Scalars), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return CSI_OK;
}

extern "C" int csi_nuts_init_state(csi_nuts *s, double *u_out) {
  if (!s) return fail(CSI_EINVAL, "null argument");
  CK(cudaSetDevice(s->h->device));
  int rc = csi_potential(s->h, s->V.z, s->d_out, s->V.g, s->stream);
  if (rc != CSI_OK) return rc;
  double out[6];
  CK(cudaMemcpyAsync(out, s->d_out, sizeof(out), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  CK(cudaMemcpy(&s->V.sc->u_prop, &out[0], sizeof(double), cudaMemcpyHostToDevice));
  s->h_sc->u_prop = out[0];
  if (u_out) *u_out = out[0];
  return CSI_OK;
}

extern "C" int csi_nuts_transition(csi_nuts *s, double step_size, uint64_t draw, csi_nuts_stats *st) {
  if (!s || !st) return fail(CSI_EINVAL, "null argument");
  if (!(step_size > 0)) return fail(CSI_EINVAL, "step_size must be > 0");
  CK(cudaSetDevice(s->h->device));
  using namespace csi::nuts;
  s->launches = 0;
  int rc = nuts_begin(s, step_size, draw, nullptr);
  if (rc != CSI_OK) return rc;
  int syncs = 0;
  for (int depth = 0; depth < s->V.max_depth; ++depth) {
    const bool right = (hash_u64(s->seed, draw, K_DIRECTION, (unsigned long long)depth) >> 63) != 0;
    if ((rc = nuts_run_doubling(s, depth, right)) != CSI_OK) break;
    if ((rc = nuts_fetch(s)) != CSI_OK) break;
    ++syncs;
    if (s->h_sc->done) break;
  }
  if (rc != CSI_OK) { cudaStreamSynchronize(s->stream); return rc; }
  const Scalars &c = *s->h_sc;
  st->potential = c.u_prop;
  st->accept_prob = c.n_prop > 0 ? c.sum_accept / (double)c.n_prop : 0.0;
  st->energy0 = c.h0;
  st->num_leapfrog = c.n_prop;
  st->depth = c.depth;
  st->diverging = c.diverging;
  st->turning = c.turning;
  st->host_syncs = syncs;
  st->launches = s->launches;
  return CSI_OK;
}

// one leapfrog with a fresh momentum from the current state, which is left untouched: energy change H1 - H0
// (find_heuristic_step_size of numpyro's NUTS, scripts/model_star.py:20)
extern "C" int csi_nuts_probe_step(csi_nuts *s, double step_size, uint64_t draw, double *delta_energy) {
  if (!s || !delta_energy) return fail(CSI_EINVAL, "null argument");
  CK(cudaSetDevice(s->h->device));
  s->launches = 0;
  int rc = nuts_begin(s, step_size, draw, nullptr);
  if (rc == CSI_OK) rc = nuts_queue_leaf(s, 0, 0, true);
  if (rc == CSI_OK) rc = nuts_fetch(s);
  if (rc != CSI_OK) { cudaStreamSynchronize(s->stream); return rc; }
  *delta_energy = -s->h_sc->logw_sub;
  return CSI_OK;
}

// test / debugging access to the internal vectors: 0 z_left, 1 r_left, 2 g_left, 3 z_right, 4 r_right, 5 g_right,
// 6 r_sum.  csi_nuts_begin_only runs the first launch of a transition (momentum draw) and stops.
extern "C" int csi_nuts_begin_only(csi_nuts *s, double step_size, uint64_t draw) {
  if (!s) return fail(CSI_EINVAL, "null argument");
  CK(cudaSetDevice(s->h->device));
  int rc = nuts_begin(s, step_size, draw, nullptr);
  if (rc == CSI_OK) rc = nuts_fetch(s);
  return rc;
}
extern "C" int csi_nuts_peek(csi_nuts *s, int32_t which, float *out_host) {
  if (!s || !out_host) return fail(CSI_EINVAL, "null argument");
  const csi::nuts::Vecs &V = s->V;
  const float *src[] = {V.zl, V.rl, V.gl, V.zr, V.rr, V.gr, V.r_sum};
  if (which < 0 || which > 6) return fail(CSI_EINVAL, "which");
  CK(cudaSetDevice(s->h->device));
  CK(cudaStreamSynchronize(s->stream));
  CK(cudaMemcpy(out_host, src[which], (size_t)V.len * sizeof(float), cudaMemcpyDeviceToHost));
  return CSI_OK;
}

// the decision stream, callable without a device: tests pin the Python mirror (nuts.hash_u64) against it
extern "C" uint64_t csi_nuts_hash(uint64_t seed, uint64_t draw, uint64_t kind, uint64_t idx) {
  return csi::nuts::hash_u64(seed, draw, kind, idx);
}
