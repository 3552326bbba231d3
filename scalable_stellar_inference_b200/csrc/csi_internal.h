// csi_internal.h -- host-side handle and the launch entry points of the two MLP engines.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/csi.h"
#include "csi_net.cuh"
#include "csi_reduce.cuh"

namespace csi {

// hidden-layer weight images for the tcgen05 engine (see csi_tc.cu for the layout)
struct TcWeights {
  const uint8_t *img;      // [10][kStepBytes]: streamed images, fwd layers 1..5 then transposed layers 5..1
  const uint8_t *resident; // images of the four narrow stages (layer 0, 6, 6^T, 0^T), kept in shared memory
  const float *small;      // packed fp32 scalings / gains
  int n_steps;
  long long *trace;        // CSI_TC_DEBUG=8: clock64 trace of one CTA (csi_tc_trace); selects the TRACE instantiation
  int debug;               // 8 + (first traced iteration << 8) + (traced CTA << 16); read by the TRACE instantiation only
  float act_scale, w_scale;  // power-of-two operand scalings (see csi_tc.cu)
  int half_k;              // hidden width <= 64 (BC emulator, zero-padded images): N = 64 MMAs over the first K-half only
};

struct Launch {           // everything a potential launch needs
  DevModel model;
  StarIO io;
  NetDev emu, bcnet;
  GridDev grid;
  TcWeights tc;
  double *partials;       // [grid][8]
  unsigned int *counter;
  double *out6;
  float *grad_hyper;      // may be null
  float *mean_out;        // predict mode only: [n][kMaxMean]
  float *params_out;      // predict mode only: [n][8]
  const float *net_x, *net_gy;  // network-operator modes: x [n][net_nin], gy [n][net_nout]
  float *net_y, *net_gx;        // y [n][net_nout], gx [n][net_nin]
  int net_nin, net_nout;        // 5, 4 (stellar emulator) / 4, 3 (BC emulator)
  PeerX px;               // fused allreduce over peer memory (world <= 1: off)
  int sm_count;
  // Early exit (device-side NUTS, csi_nuts.cuh): when *stop != 0 the launch returns at once without touching
  // anything -- a stopped sub-tree skips the leapfrogs queued behind it.  null = always run.
  const int *stop;
  int grid_limit;         // upper bound on the number of CTAs (0 = one per SM): several chains share one GPU
};

// returns the number of kernel launches issued (>0) or a negative cudaError
int launch_potential_ffma(const Launch &L, cudaStream_t st, int *grid_out, int *smem_out);
int launch_potential_tc(const Launch &L, cudaStream_t st, int *grid_out, int *smem_out);
int launch_predict_ffma(const Launch &L, cudaStream_t st);
int launch_predict_tc(const Launch &L, cudaStream_t st);
int launch_net_tc(const Launch &L, bool vjp, cudaStream_t st);
int launch_net_forward(const NetDev &net, const float *x, float *y, long long n, cudaStream_t st);
int launch_net_vjp(const NetDev &net, const float *x, const float *gy, float *gx, long long n, cudaStream_t st);
int launch_grid_forward(const GridDev &g, int bands, const float *x, float *y, long long n, cudaStream_t st);
int launch_grid_vjp(const GridDev &g, int bands, const float *x, const float *gy, float *gx, long long n,
                    cudaStream_t st);
int max_partial_blocks(int sm_count, long long n);

// host-side preparation of the tcgen05 weight images (csi_tc.cu)
size_t tc_image_bytes(const csi_net &net);
size_t tc_resident_bytes(const csi_net &net);
size_t tc_small_floats(const csi_net &net);
void tc_build_images(const csi_net &net, uint8_t *img_host, uint8_t *resident_host, float *small_host,
                     float *act_scale, float *w_scale);
bool tc_supported(const csi_net &net);   // any 7-layer net up to 5 -> 128 x 6 -> 4 (narrower ones are zero-padded)
bool tc_half_k(const csi_net &net);     // hidden width <= 64: TcWeights::half_k

}  // namespace csi

struct csi_handle {
  int device = 0, sm_count = 0, engine = 0;
  int level = 0, prior = 0;
  int n_hyper = 0, n_sites = 0;
  int slot_of_site[csi::kSlots] = {0};
  const char *names[CSI_MAX_HYPER + CSI_MAX_STAR_SITES] = {nullptr};
  int rank = 0, world = 1;
  csi::DevModel model{};
  csi::NetDev emu{}, bcnet{};
  csi::GridDev grid{};
  csi::TcWeights tc{}, tc_bc{};  // tensor-core operand images: stellar emulator / BC emulator (operator only)
  std::vector<void *> owned;  // device allocations released in csi_destroy
  // data
  long long n_local = 0;
  float *d_obs = nullptr, *d_diag = nullptr, *d_mhloc = nullptr, *d_mhscale = nullptr, *d_mhlogz = nullptr;
  bool has_diag = false, has_mh = false;
  double *d_partials = nullptr;
  unsigned int *d_counter = nullptr;
  int partial_cap = 0;
  // host-variant staging
  float *h_z = nullptr, *h_grad = nullptr, *d_z = nullptr, *d_grad = nullptr;
  double *h_out = nullptr, *d_out = nullptr;
  long long stage_len = 0;
  cudaStream_t stream = nullptr;
  csi::PeerX px{};          // csi_set_peer_exchange
  unsigned long long *d_epoch = nullptr;  // exchange epoch, advanced on the device by every exchange
  const int *stop_flag = nullptr;         // Launch::stop of the next csi_potential (set by the NUTS driver)
  int grid_limit = 0;
  double *d_vec_partials = nullptr;         // scratch of the NUTS vector kernels (csi_nuts.cuh)
  unsigned int *d_vec_counter = nullptr;
  long long last_launches = 0;
  int last_grid = 0, last_smem = 0;
};
