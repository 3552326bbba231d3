// csi_nuts.cuh (included at the end of csi_api.cu) -- the vector arithmetic of one NUTS leapfrog / U-turn check around the potential kernel.
//
// numpyro's velocity-Verlet (hmc_util.velocity_verlet) and the generalised U-turn test (hmc_util._is_turning) are
// a handful of elementwise updates and inner products on vectors as long as the star axis; issued one framework
// op at a time they cost more than the potential itself (tens of launches and two host synchronisations per
// leapfrog).  Here they are three memory-bound kernels over the flat [hyper | local stars] vector:
//
//   leapfrog_pre  : r_half = r - eps/2 g ;  z1 = z + eps M^-1 r_half                    (28 B/element read+write x ..)
//   leapfrog_post : g1 = gradient (hyper slots from the reduced 6 sums) ;  r1 = r_half - eps/2 g1 ;
//                   kin[0] += sum_stars M^-1 r1^2 ,  kin[1] += sum_hyper M^-1 r1^2      (fp64 accumulation)
//   uturn         : rc = r_sum - (r_left + r_right)/2 ;  dots[0..1] += star parts of (M^-1 r_left).rc, (M^-1 r_right).rc
//                   dots[2..3] += the hyper parts
//
// The star parts are the ones a sharded run allreduces; the hyper entries are replicated on every rank.
// HBM-bound: 148 x 4 blocks of 256 threads, grid-stride, one atomicAdd(double) per block and sum.
#pragma once
#include "csi_internal.h"

namespace csi {
namespace nuts {

constexpr int kBlock = 256;

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

__global__ void __launch_bounds__(kBlock) leapfrog_pre_kernel(long long n, const float *__restrict__ z,
                                                              const float *__restrict__ r, const float *__restrict__ g,
                                                              const float *__restrict__ im, float eps,
                                                              float *__restrict__ z1, float *__restrict__ r_half,
                                                              double *kin) {
  if (blockIdx.x == 0 && threadIdx.x < 2) kin[threadIdx.x] = 0.0;  // the sums leapfrog_post accumulates into
  const float he = 0.5f * eps;
  for (long long i = (long long)blockIdx.x * kBlock + threadIdx.x; i < n; i += (long long)gridDim.x * kBlock) {
    const float rh = r[i] - he * g[i];
    r_half[i] = rh;
    z1[i] = z[i] + eps * (im[i] * rh);
  }
}

__global__ void __launch_bounds__(kBlock) leapfrog_post_kernel(long long n, int nh, const float *__restrict__ r_half,
                                                               const float *__restrict__ grad,
                                                               const double *__restrict__ out6,
                                                               const float *__restrict__ im, float eps,
                                                               float *__restrict__ r1, float *__restrict__ g1,
                                                               double *kin) {
  __shared__ double sh[kBlock / 32];
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
  const double ts = block_sum(ks, sh);
  if (threadIdx.x == 0 && ts != 0.0) atomicAdd(&kin[0], ts);
  if (blockIdx.x == 0) {  // the first nh elements belong to block 0
    const double th = block_sum(kh, sh);
    if (threadIdx.x == 0) kin[1] = th;
  }
}

__global__ void __launch_bounds__(kBlock) uturn_kernel(long long n, int nh, const float *__restrict__ rl,
                                                       const float *__restrict__ rr, const float *__restrict__ rs,
                                                       const float *__restrict__ im, double *dots) {
  __shared__ double sh[kBlock / 32];
  double a = 0.0, b = 0.0, ah = 0.0, bh = 0.0;
  for (long long i = (long long)blockIdx.x * kBlock + threadIdx.x; i < n; i += (long long)gridDim.x * kBlock) {
    const float l = rl[i], r = rr[i];
    const float rc = rs[i] - 0.5f * (l + r);
    const double m = (double)im[i], c = (double)rc;
    if (i < nh) { ah += m * l * c; bh += m * r * c; } else { a += m * l * c; b += m * r * c; }
  }
  const double ta = block_sum(a, sh);
  if (threadIdx.x == 0 && ta != 0.0) atomicAdd(&dots[0], ta);
  const double tb = block_sum(b, sh);
  if (threadIdx.x == 0 && tb != 0.0) atomicAdd(&dots[1], tb);
  if (blockIdx.x == 0) {
    const double tah = block_sum(ah, sh), tbh = block_sum(bh, sh);
    if (threadIdx.x == 0) { dots[2] = tah; dots[3] = tbh; }
  }
}

static int grid_for(const csi_handle *h, long long n) {
  const long long want = (n + kBlock - 1) / kBlock;
  return (int)std::max<long long>(1, std::min<long long>(want, (long long)h->sm_count * 4));
}

}  // namespace nuts
}  // namespace csi


extern "C" int csi_leapfrog_pre(csi_handle *h, const float *z, const float *r, const float *g, const float *inv_mass,
                                float eps, float *z1, float *r_half, double *kin2, void *stream) {
  if (!h || !z || !r || !g || !inv_mass || !z1 || !r_half || !kin2) return fail(CSI_EINVAL, "null argument");
  if (h->n_local < 1) return fail(CSI_ESTATE, "csi_set_data has not been called");
  const long long n = h->stage_len;
  nuts::leapfrog_pre_kernel<<<nuts::grid_for(h, n), nuts::kBlock, 0, (cudaStream_t)stream>>>(n, z, r, g, inv_mass, eps,
                                                                                            z1, r_half, kin2);
  CK(cudaGetLastError());
  return CSI_OK;
}

extern "C" int csi_leapfrog_post(csi_handle *h, const float *r_half, const float *grad, const double *out6,
                                 const float *inv_mass, float eps, float *r1, float *g1, double *kin2, void *stream) {
  if (!h || !r_half || !grad || !out6 || !inv_mass || !r1 || !g1 || !kin2) return fail(CSI_EINVAL, "null argument");
  if (h->n_local < 1) return fail(CSI_ESTATE, "csi_set_data has not been called");
  const long long n = h->stage_len;
  nuts::leapfrog_post_kernel<<<nuts::grid_for(h, n), nuts::kBlock, 0, (cudaStream_t)stream>>>(
      n, h->n_hyper, r_half, grad, out6, inv_mass, eps, r1, g1, kin2);
  CK(cudaGetLastError());
  return CSI_OK;
}

extern "C" int csi_uturn(csi_handle *h, const float *r_left, const float *r_right, const float *r_sum,
                         const float *inv_mass, double *dots4, void *stream) {
  if (!h || !r_left || !r_right || !r_sum || !inv_mass || !dots4) return fail(CSI_EINVAL, "null argument");
  if (h->n_local < 1) return fail(CSI_ESTATE, "csi_set_data has not been called");
  const long long n = h->stage_len;
  CK(cudaMemsetAsync(dots4, 0, 4 * sizeof(double), (cudaStream_t)stream));
  nuts::uturn_kernel<<<nuts::grid_for(h, n), nuts::kBlock, 0, (cudaStream_t)stream>>>(n, h->n_hyper, r_left, r_right,
                                                                                     r_sum, inv_mass, dots4);
  CK(cudaGetLastError());
  return CSI_OK;
}
