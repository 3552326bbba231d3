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
  long long *trace;        // CSI_TC_DEBUG & 8: clock64 trace of CTA 0 (csi_tc_trace)
  int debug;               // CSI_TC_DEBUG experiments (timing only; results are wrong when non-zero)
  float act_scale, w_scale;  // power-of-two operand scalings (see csi_tc.cu)
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
  const float *net_x, *net_gy;  // emulator-operator modes: x [n][5], gy [n][4]
  float *net_y, *net_gx;        // y [n][4], gx [n][5]
  PeerX px;               // fused allreduce over peer memory (world <= 1: off)
  int sm_count;
  // Gated sites (csi_potential_host): the copy engine fills io.z[] block by block while the kernel runs and
  // publishes z_base + (stars copied so far) in *z_ready after each block; star i may be staged once
  // *z_ready - z_base > i.  null = the sites are all there at launch.
  const unsigned long long *z_ready;
  unsigned long long z_base;
  // hybrid: stars below z_direct_end are read from the caller's page-locked buffer itself (= io.z[] + z_host_delta
  // floats, no gate), so the kernel starts at once while the copy engine works on the blocks behind them
  long long z_direct_end, z_host_delta;
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
bool tc_supported(const csi_net &net);

}  // namespace csi

constexpr int kGateMaxBlocks = 64;  // star blocks of the gated host-buffer copy

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
  csi::TcWeights tc{};
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
  // gated host-buffer path: copy stream, progress word (device) and the values the copy engine writes into it
  cudaStream_t copy_stream = nullptr;
  unsigned long long *d_zflag = nullptr, *h_zflag_vals = nullptr;
  unsigned long long zgate_base = 0;
  bool zgate_on = false;
  long long zgate_direct_end = 0, zgate_host_delta = 0;
  csi::PeerX px{};          // csi_set_peer_exchange
  long long last_launches = 0;
  int last_grid = 0, last_smem = 0;
};
