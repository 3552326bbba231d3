// csi_reduce.cuh -- block reduce of the (U, hyper-gradient) partial sums, the single-pass grid reduce and the
// allreduce over NVLink peer memory that the finalising block performs itself.
#pragma once
#include "csi_model.cuh"

namespace csi {

// Star-sharded evaluation: the allreduce of a handful of doubles fused into the finalising block of a kernel.
// Every rank owns an exchange buffer double[2][world][kPxSlot] (epoch parity x source rank x {values, flag}) that
// all ranks can write over NVLink peer memory (csi_set_peer_exchange).  The finalising block stores its sums + the
// epoch flag into slot [epoch & 1][my rank] of EVERY rank's buffer, then waits for the world's flags in its own
// buffer and adds the contributions in rank order (bit-identical totals on all ranks).  Double buffering by epoch
// parity is enough: a rank can only be one exchange ahead of the slowest one (it needs everybody's flag to finish).
// The epoch lives in device memory and is advanced by the exchange itself, so a launch replayed from a CUDA graph,
// or skipped on every rank by the same device-side decision (Launch::stop), keeps the ranks in step.
constexpr int kMaxPeers = 8;
constexpr int kPxSlot = 32;  // doubles per slot: up to 31 values, the flag in the last one
struct PeerX {
  double *bufs[kMaxPeers];
  int world, rank;
  unsigned long long *epoch;  // device word of this rank: exchanges completed so far (the same count on all ranks)
  unsigned long long timeout_ns;  // how long a block waits for a peer before it poisons the result with NaN
};

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// vals[0..NV) (shared memory) <- sum over the ranks, in rank order.  Called by ALL threads of one block
// (blockDim.x >= world); a peer that does not answer within px.timeout_ns poisons the result with NaN.
template <int NV>
__device__ __forceinline__ void peer_allreduce(const PeerX &px, double *vals) {
  static_assert(NV < kPxSlot, "exchange slot");
  __shared__ double got[kMaxPeers][NV];
  const int tid = threadIdx.x;
  __syncthreads();
  const unsigned long long epoch = *reinterpret_cast<volatile unsigned long long *>(px.epoch) + 1ull;
  const int par = (int)(epoch & 1ull);
  if (tid < px.world) {
    volatile double *dst = px.bufs[tid] + ((size_t)par * px.world + px.rank) * kPxSlot;
    for (int j = 0; j < NV; ++j) dst[j] = vals[j];
    __threadfence_system();
    *reinterpret_cast<volatile unsigned long long *>(dst + kPxSlot - 1) = epoch;
    volatile double *src = px.bufs[px.rank] + ((size_t)par * px.world + tid) * kPxSlot;
    const volatile unsigned long long *flag = reinterpret_cast<volatile unsigned long long *>(src + kPxSlot - 1);
    const unsigned long long t0 = global_ns();
    while (*flag != epoch && global_ns() - t0 < px.timeout_ns) {
    }
    __threadfence_system();
    const bool ok = *flag == epoch;
    for (int j = 0; j < NV; ++j) got[tid][j] = ok ? src[j] : nan("");
  }
  __syncthreads();
  if (tid < NV) {
    double t = 0.0;
    for (int r = 0; r < px.world; ++r) t += got[r][tid];
    vals[tid] = t;
  }
  if (tid == 0) *px.epoch = epoch;
  __syncthreads();
}

// -------------------------------------------------------------------------------------------------
// block reduce of the 6 partial sums + single-pass grid reduce (last block finalises, fixed order)
// -------------------------------------------------------------------------------------------------
// per-warp partial sums of one thread's accumulators -> red_smem[warp * 6 + j] (lane 0 writes)
__device__ __forceinline__ void warp_partials_to_smem(const HyperAcc &acc, double *red_smem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double v[6] = {acc.lp, acc.g[0], acc.g[1], acc.g[2], acc.g[3], acc.g[4]};
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    double s = warp_sum(v[j]);
    if (lane == 0) red_smem[warp * 6 + j] = s;
  }
}

__device__ __noinline__ static void block_reduce_finalize_tail(const DevModel &m, const HyperVals &hv, int nwarp, double *red_smem,
                                      double *partials, unsigned int *counter, double *out6, float *grad_hyper,
                                      const PeerX &px);

__device__ __forceinline__ void block_reduce_finalize(const DevModel &m, const HyperVals &hv, HyperAcc acc, double *red_smem,
                                      double *partials, unsigned int *counter, double *out6, float *grad_hyper,
                                      const PeerX &px) {
  warp_partials_to_smem(acc, red_smem);
  block_reduce_finalize_tail(m, hv, (blockDim.x + 31) >> 5, red_smem, partials, counter, out6, grad_hyper, px);
}

// red_smem[w * 6 + j], w < nwarp: the warps' partial sums (written before the call; the barrier below orders them)
__device__ __noinline__ static void block_reduce_finalize_tail(const DevModel &m, const HyperVals &hv, int nwarp, double *red_smem,
                                      double *partials, unsigned int *counter, double *out6, float *grad_hyper,
                                      const PeerX &px) {
  const int tid = threadIdx.x;
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
    if (px.world > 1) peer_allreduce<6>(px, tot);  // fused allreduce over NVLink peer memory
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
