// csi_reduce.cuh -- block reduce of the (U, hyper-gradient) partial sums and the single-pass grid reduce.
#pragma once
#include "csi_model.cuh"

namespace csi {

// Star-sharded evaluation: the allreduce of the 6 sums fused into the finalising block.  Every rank owns an
// exchange buffer double[2][world][8] (epoch parity x source rank x {6 values, spare, flag}) that all ranks can
// write over NVLink peer memory (csi_set_peer_exchange).  The finalising block stores its 6 sums + the epoch flag
// into slot [epoch & 1][my rank] of EVERY rank's buffer, then waits for the world's flags in its own buffer and
// adds the contributions in rank order (bit-identical totals on all ranks).  Double buffering by epoch parity is
// enough: a rank can only be one evaluation ahead of the slowest one (it needs everybody's flag to finish).
constexpr int kMaxPeers = 8;
struct PeerX {
  double *bufs[kMaxPeers];
  int world, rank;
  unsigned long long epoch;  // 1, 2, ...: number of this evaluation (same on all ranks)
};

// -------------------------------------------------------------------------------------------------
// block reduce of the 6 partial sums + single-pass grid reduce (last block finalises, fixed order)
// -------------------------------------------------------------------------------------------------
__device__ __noinline__ static void block_reduce_finalize(const DevModel &m, const HyperVals &hv, HyperAcc acc, double *red_smem,
                                      double *partials, unsigned int *counter, double *out6, float *grad_hyper,
                                      const PeerX &px) {
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
    __shared__ double tot[6];
    if (tid == 0) {
      const bool hier = (m.prior_mode == P_HIER);
      const double hon = (hier && m.hyper_on) ? 1.0 : 0.0;
      tot[0] = -(red_smem[0] + hon * (double)hv.prior_lp);
      double chain[5] = {1.0, (double)hv.ddydz, -0.5 * (double)hv.sigma_y, (double)hv.dmu_a, -0.5 * (double)hv.sigma_a};
      for (int j = 0; j < 5; ++j)
        tot[1 + j] = hier ? -(hon * (double)hv.gprior[j] + chain[j] * red_smem[1 + j]) : 0.0;
    }
    __syncthreads();
    if (px.world > 1) {
      // ---- fused allreduce over NVLink peer memory: thread p talks to rank p ----------------------------------
      __shared__ double got[kMaxPeers][6];
      const int par = (int)(px.epoch & 1ull);
      if (tid < px.world) {
        volatile double *dst = px.bufs[tid] + ((size_t)par * px.world + px.rank) * 8;
        for (int j = 0; j < 6; ++j) dst[j] = tot[j];
        __threadfence_system();
        *reinterpret_cast<volatile unsigned long long *>(dst + 7) = px.epoch;
        volatile double *src = px.bufs[px.rank] + ((size_t)par * px.world + tid) * 8;
        const volatile unsigned long long *flag = reinterpret_cast<volatile unsigned long long *>(src + 7);
        long long spins = 0;
        while (*flag != px.epoch && ++spins < (1ll << 27)) {
        }
        __threadfence_system();
        const bool ok = *flag == px.epoch;
        for (int j = 0; j < 6; ++j) got[tid][j] = ok ? src[j] : nan("");  // a missing peer poisons the result
      }
      __syncthreads();
      if (tid == 0)
        for (int j = 0; j < 6; ++j) {
          double t = 0.0;
          for (int r = 0; r < px.world; ++r) t += got[r][j];
          tot[j] = t;
        }
      __syncthreads();
    }
    if (tid == 0) {
      const bool hier = (m.prior_mode == P_HIER);
      for (int j = 0; j < 6; ++j) out6[j] = tot[j];
      for (int j = 0; j < 5; ++j)
        if (grad_hyper != nullptr && hier) grad_hyper[j] = (float)tot[1 + j];
      *counter = 0u;  // re-arm for the next launch
    }
  }
}


}  // namespace csi
