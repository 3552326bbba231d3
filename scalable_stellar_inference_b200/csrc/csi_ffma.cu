// csi_ffma.cu -- CUDA-core (fp32 FFMA) engine of the potential + the standalone operator kernels
// (emulator / BC-net forward and VJP, BC table interpolation and VJP, predict).
//
// The FFMA engine evaluates exactly the same per-star phases as the tcgen05 engine (csi_tc.cu) but runs the
// MLP contractions on the FP32 pipe with plain fmaf accumulation in k-order.  It is the in-repo fp32 GPU
// reference for the tensor-core engine, the engine for tiny problems, and the only engine that supports the
// BC-emulator net inside the potential in this round.
#include <cooperative_groups.h>

#include "csi_internal.h"
#include "csi_reduce.cuh"

namespace csi {

__device__ __forceinline__ void load_hyper(const DevModel &m, const StarIO &io, HyperVals *hv_smem) {
  if (threadIdx.x == 0) {
    HyperVals h{};
    if (m.prior_mode == P_HIER) hyper_eval(m, io.z_hyper, h);
    *hv_smem = h;
  }
  __syncthreads();
}

// -------------------------------------------------------------------------------------------------
// potential, FFMA engine: 128 stars per block, one thread per star, thread-private smem columns
// -------------------------------------------------------------------------------------------------
struct FfmaSmem {
  float bufA[128 * kRows];
  float bufB[128 * kRows];
  uint32_t mask[6 * 4 * kRows];     // stellar emulator: 6 hidden layers x 128 bits
  uint32_t bcmask[6 * 2 * kRows];   // BC emulator: 6 hidden layers x 64 bits
  double red[4 * 6];
  HyperVals hv;
};

template <int BC, bool PREDICT>
__global__ void __launch_bounds__(kRows) potential_ffma_kernel(const __grid_constant__ Launch L) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FfmaSmem &S = *reinterpret_cast<FfmaSmem *>(smem_raw);
  if (L.stop != nullptr && *reinterpret_cast<const volatile int *>(L.stop) != 0) return;  // skipped leapfrog (csi_nuts.cuh)
  const DevModel &m = L.model;
  const StarIO &io = L.io;
  load_hyper(m, io, &S.hv);
  const HyperVals hv = S.hv;
  const int tid = threadIdx.x;
  const long long i = (long long)blockIdx.x * kRows + tid;
  HyperAcc acc{};
  if (i < io.n) {
    StarState st;
    star_pre(m, hv, io, i, st, acc);
    float o[4], g_o[4], g_x[8];
    float xin[8] = {st.x[0], st.x[1], st.x[2], st.x[3], st.x[4], 0.f, 0.f, 0.f};
    net_forward_ffma(L.emu, xin, o, S.bufA, S.bufB, S.mask, tid);
    float *mean_out = nullptr;
    float mean_local[kMaxMean];
    if (PREDICT) mean_out = mean_local;
    StarObs so;
    star_load_obs(m, io, i, so);
    if (BC == 0) { BcNone bc; star_mid(m, io, i, st, o, g_o, bc, so, mean_out); }
    if (BC == 1) { BcGrid bc{L.grid}; star_mid(m, io, i, st, o, g_o, bc, so, mean_out); }
    if (BC == 2) { BcNetFfma bc{L.bcnet, S.bufA, S.bufB, S.bcmask, tid}; star_mid(m, io, i, st, o, g_o, bc, so, mean_out); }
    if (PREDICT) {
      if (L.mean_out) {
#pragma unroll
        for (int k = 0; k < kMaxMean; ++k) L.mean_out[i * kMaxMean + k] = mean_local[k];
      }
      if (L.params_out) {
        float *p = L.params_out + i * 8;
        p[0] = st.x[0]; p[1] = st.x[1]; p[2] = st.x[2]; p[3] = st.x[3]; p[4] = st.x[4];
        p[5] = st.log_mass; p[6] = st.tau; p[7] = st.dist;
      }
    } else {
      net_vjp_ffma(L.emu, g_o, g_x, S.bufA, S.bufB, S.mask, tid);
      star_post(m, io, i, st, g_x);
      acc.lp += (double)st.lp;
    }
  }
  if (!PREDICT) block_reduce_finalize(m, hv, acc, S.red, L.partials, L.counter, L.out6, L.grad_hyper, L.px);
}

template <int BC, bool PREDICT>
static int launch_ffma_t(const Launch &L, cudaStream_t st, int *grid_out, int *smem_out) {
  const int smem = (int)sizeof(FfmaSmem);
  auto kern = potential_ffma_kernel<BC, PREDICT>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return -(int)e;
  const int grid = (int)((L.io.n + kRows - 1) / kRows);
  kern<<<grid, kRows, smem, st>>>(L);
  e = cudaGetLastError();
  if (e != cudaSuccess) return -(int)e;
  if (grid_out) *grid_out = grid;
  if (smem_out) *smem_out = smem;
  return 1;
}

int launch_potential_ffma(const Launch &L, cudaStream_t st, int *grid_out, int *smem_out) {
  switch (L.model.bands > 0 ? L.model.bc_mode : 0) {
    case 0: return launch_ffma_t<0, false>(L, st, grid_out, smem_out);
    case 1: return launch_ffma_t<1, false>(L, st, grid_out, smem_out);
    default: return launch_ffma_t<2, false>(L, st, grid_out, smem_out);
  }
}

int launch_predict_ffma(const Launch &L, cudaStream_t st) {
  switch (L.model.bands > 0 ? L.model.bc_mode : 0) {
    case 0: return launch_ffma_t<0, true>(L, st, nullptr, nullptr);
    case 1: return launch_ffma_t<1, true>(L, st, nullptr, nullptr);
    default: return launch_ffma_t<2, true>(L, st, nullptr, nullptr);
  }
}

int max_partial_blocks(int sm_count, long long n) {
  long long b = (n + kRows - 1) / kRows;
  return (int)(b > sm_count ? b : sm_count);
}

// -------------------------------------------------------------------------------------------------
// standalone operators: Emulator.model / BolometricCorrections.model forward + VJP
// -------------------------------------------------------------------------------------------------
struct NetSmem {
  float bufA[128 * kRows];
  float bufB[128 * kRows];
  uint32_t mask[6 * 4 * kRows];
};

template <bool VJP>
__global__ void __launch_bounds__(kRows) net_kernel(const __grid_constant__ NetDev net, const float *__restrict__ x,
                                                    const float *__restrict__ gy, float *__restrict__ out, long long n) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  NetSmem &S = *reinterpret_cast<NetSmem *>(smem_raw);
  const int tid = threadIdx.x;
  const long long i = (long long)blockIdx.x * kRows + tid;
  if (i >= n) return;
  float xi[8], y[4];
#pragma unroll
  for (int k = 0; k < 8; ++k) xi[k] = (k < net.n_in) ? x[i * net.n_in + k] : 0.0f;
  net_forward_ffma(net, xi, y, S.bufA, S.bufB, S.mask, tid);
  if (!VJP) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (j < net.n_out) out[i * net.n_out + j] = y[j];
  } else {
    float g[4], gx[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) g[j] = (j < net.n_out) ? gy[i * net.n_out + j] : 0.0f;
    net_vjp_ffma(net, g, gx, S.bufA, S.bufB, S.mask, tid);
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (k < net.n_in) out[i * net.n_in + k] = gx[k];
  }
}

template <bool VJP>
static int launch_net_t(const NetDev &net, const float *x, const float *gy, float *out, long long n, cudaStream_t st) {
  if (n <= 0) return 0;
  const int smem = (int)sizeof(NetSmem);
  auto kern = net_kernel<VJP>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return -(int)e;
  const int grid = (int)((n + kRows - 1) / kRows);
  kern<<<grid, kRows, smem, st>>>(net, x, gy, out, n);
  e = cudaGetLastError();
  return e == cudaSuccess ? 1 : -(int)e;
}

int launch_net_forward(const NetDev &net, const float *x, float *y, long long n, cudaStream_t st) {
  return launch_net_t<false>(net, x, nullptr, y, n, st);
}
int launch_net_vjp(const NetDev &net, const float *x, const float *gy, float *gx, long long n, cudaStream_t st) {
  return launch_net_t<true>(net, x, gy, gx, n, st);
}

// -------------------------------------------------------------------------------------------------
// BC table interpolation (K2): 16 float4 corner gathers per star, Jacobian formed with the value
// -------------------------------------------------------------------------------------------------
template <bool VJP>
__global__ void __launch_bounds__(256) grid_kernel(const __grid_constant__ GridDev gd, int bands,
                                                   const float4 *__restrict__ x, const float *__restrict__ gy,
                                                   float *__restrict__ out, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 xv = __ldg(x + i);
  float xi[4] = {xv.x, xv.y, xv.z, xv.w}, val[4], jac[4][4];
  grid_eval(gd, xi, val, jac);
  if (!VJP) {
    for (int b = 0; b < bands; ++b) out[i * bands + b] = val[b];
  } else {
    float g[4] = {0.f, 0.f, 0.f, 0.f};
    for (int b = 0; b < bands; ++b) g[b] = gy[i * bands + b];
    float4 r;
    r.x = g[0] * jac[0][0] + g[1] * jac[1][0] + g[2] * jac[2][0] + g[3] * jac[3][0];
    r.y = g[0] * jac[0][1] + g[1] * jac[1][1] + g[2] * jac[2][1] + g[3] * jac[3][1];
    r.z = g[0] * jac[0][2] + g[1] * jac[1][2] + g[2] * jac[2][2] + g[3] * jac[3][2];
    r.w = g[0] * jac[0][3] + g[1] * jac[1][3] + g[2] * jac[2][3] + g[3] * jac[3][3];
    reinterpret_cast<float4 *>(out)[i] = r;
  }
}

int launch_grid_forward(const GridDev &g, int bands, const float *x, float *y, long long n, cudaStream_t st) {
  if (n <= 0) return 0;
  const int grid = (int)((n + 255) / 256);
  grid_kernel<false><<<grid, 256, 0, st>>>(g, bands, reinterpret_cast<const float4 *>(x), nullptr, y, n);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 1 : -(int)e;
}
int launch_grid_vjp(const GridDev &g, int bands, const float *x, const float *gy, float *gx, long long n,
                    cudaStream_t st) {
  if (n <= 0) return 0;
  const int grid = (int)((n + 255) / 256);
  grid_kernel<true><<<grid, 256, 0, st>>>(g, bands, reinterpret_cast<const float4 *>(x), gy, gx, n);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 1 : -(int)e;
}

}  // namespace csi
