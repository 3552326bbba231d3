// csi_reduce.cuh -- block reduce of the (U, hyper-gradient) partial sums and the single-pass grid reduce.
#pragma once
#include "csi_model.cuh"

namespace csi {

// -------------------------------------------------------------------------------------------------
// block reduce of the 6 partial sums + single-pass grid reduce (last block finalises, fixed order)
// -------------------------------------------------------------------------------------------------
__device__ __noinline__ static void block_reduce_finalize(const DevModel &m, const HyperVals &hv, HyperAcc acc, double *red_smem,
                                      double *partials, unsigned int *counter, double *out6, float *grad_hyper) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = (blockDim.x + 31) >> 5;
  double v[6] = {acc.lp, acc.g[0], acc.g[1], acc.g[2], acc.g[3], acc.g[4]};
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    double s = warp_sum(v[j]);
    if (lane == 0) red_smem[warp * 6 + j] = s;
  }
  __syncthreads();
  __shared__ bool is_last;
  if (tid == 0) {
    for (int j = 0; j < 6; ++j) {
      double s = 0.0;
      for (int w = 0; w < nwarp; ++w) s += red_smem[w * 6 + j];
      partials[(size_t)blockIdx.x * 8 + j] = s;
    }
    __threadfence();
    unsigned int t = atomicAdd(counter, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    // fixed-order sum over blocks: deterministic for a given grid size
    if (tid < 6) {
      double s = 0.0;
      const volatile double *p = partials;
      for (unsigned int b = 0; b < gridDim.x; ++b) s += p[(size_t)b * 8 + tid];
      red_smem[tid] = s;
    }
    __syncthreads();
    if (tid == 0) {
      const bool hier = (m.prior_mode == P_HIER);
      const double hon = (hier && m.hyper_on) ? 1.0 : 0.0;
      out6[0] = -(red_smem[0] + hon * (double)hv.prior_lp);
      double chain[5] = {1.0, (double)hv.ddydz, -0.5 * (double)hv.sigma_y, (double)hv.dmu_a, -0.5 * (double)hv.sigma_a};
      for (int j = 0; j < 5; ++j) {
        double g = hier ? -(hon * (double)hv.gprior[j] + chain[j] * red_smem[1 + j]) : 0.0;
        out6[1 + j] = g;
        if (grad_hyper != nullptr && hier) grad_hyper[j] = (float)g;
      }
      *counter = 0u;  // re-arm for the next launch
    }
  }
}


}  // namespace csi
