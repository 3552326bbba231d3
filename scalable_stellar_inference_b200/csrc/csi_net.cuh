// csi_net.cuh -- device-side description of a dense ReLU network and the CUDA-core (FFMA) layer routines.
//
// y = out_loc + out_scale * MLP((x - in_loc) / in_scale), hidden layers ReLU
// (src/celestify/_jax/emulator.py:44-50, src/celestify/_jax/bolometric_corrections.py:17-25).
//
// FFMA engine: one thread owns one row (star); its activations live in a thread-private column of a
// shared-memory buffer act[k * kRows + tid] (conflict-free, no block synchronisation needed); weights
// are read with warp-uniform float4 loads (L1-resident).  The ReLU masks are kept as bit words so that
// the input-VJP never re-reads activations.
#pragma once
#include "csi_model.cuh"

namespace csi {

constexpr int kMaxLayers = 8;
constexpr int kRows = 128;  // rows (stars) per block in the FFMA engine

struct NetDev {
  int n_layers, n_in, n_hidden, n_out;
  int kdim[kMaxLayers], ndim[kMaxLayers];  // true in/out width per layer
  int np[kMaxLayers];                      // out width padded to a multiple of 4 (row stride of W)
  int kp[kMaxLayers];                      // in width padded to a multiple of 4 (row stride of Wt)
  const float *W[kMaxLayers];              // [kdim][np]   kernel, zero-padded columns
  const float *Wt[kMaxLayers];             // [ndim][kp]   transposed kernel, zero-padded columns
  const float *b[kMaxLayers];              // [np]
  float in_loc[8], in_scale[8], out_loc[4], out_scale[4];
};

// out[j] = (bias ? b[j] : 0) + sum_k in[k] * W[k][j],  j < np (np % 4 == 0); thread-private smem columns.
// MODE 0: plain store.  MODE 1: ReLU + record mask bits (mask words at mask[(j/32) * kRows + tid]).
// MODE 2: multiply by a previously recorded mask (VJP through ReLU).
template <int MODE>
CSI_DEV void dense_ffma(const float *__restrict__ in, float *__restrict__ out, const float *__restrict__ W,
                        const float *__restrict__ bias, int K, int np, uint32_t *mask, int tid) {
  for (int j0 = 0; j0 < np; j0 += 16) {
    const int nq = min(4, (np - j0) >> 2);
    float acc[16];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
      if (bias != nullptr && q < nq) bv = __ldg(reinterpret_cast<const float4 *>(bias + j0) + q);
      acc[4 * q] = bv.x; acc[4 * q + 1] = bv.y; acc[4 * q + 2] = bv.z; acc[4 * q + 3] = bv.w;
    }
    for (int k = 0; k < K; ++k) {
      const float a = in[k * kRows + tid];
      const float4 *wr = reinterpret_cast<const float4 *>(W + (size_t)k * np + j0);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (q < nq) {
          float4 w = __ldg(wr + q);
          acc[4 * q] = fmaf(a, w.x, acc[4 * q]);
          acc[4 * q + 1] = fmaf(a, w.y, acc[4 * q + 1]);
          acc[4 * q + 2] = fmaf(a, w.z, acc[4 * q + 2]);
          acc[4 * q + 3] = fmaf(a, w.w, acc[4 * q + 3]);
        }
      }
    }
    uint32_t bits = 0;
    uint32_t mword = 0;
    if (MODE == 2) mword = mask[(j0 >> 5) * kRows + tid] >> (j0 & 31);
#pragma unroll
    for (int jj = 0; jj < 16; ++jj) {
      if (jj < 4 * nq) {
        float v = acc[jj];
        if (MODE == 1) { bool p = v > 0.0f; bits |= (p ? 1u : 0u) << jj; v = p ? v : 0.0f; }
        if (MODE == 2) v = ((mword >> jj) & 1u) ? v : 0.0f;
        out[(j0 + jj) * kRows + tid] = v;
      }
    }
    if (MODE == 1) {
      uint32_t *mw = mask + (j0 >> 5) * kRows + tid;
      uint32_t old = (j0 & 31) ? *mw : 0u;
      *mw = old | (bits << (j0 & 31));
    }
  }
}

// words of mask storage per hidden layer
CSI_DEV int mask_words(const NetDev &net) { return (net.n_hidden + 31) >> 5; }

// forward: x[8] (registers, first n_in used) -> y[4]; bufA/bufB thread-private smem columns; masks recorded.
CSI_DEV void net_forward_ffma(const NetDev &net, const float *x, float *y, float *bufA, float *bufB,
                              uint32_t *mask, int tid) {
  // fixed-trip, guarded loops: x / y stay in registers (a runtime-bounded loop over a caller's local array
  // made nvcc 12.9 drop the stores to y in the forward-only instantiation)
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if (k < net.n_in) bufA[k * kRows + tid] = (x[k] - net.in_loc[k]) / net.in_scale[k];
  float *a = bufA, *b = bufB;
  const int mw = mask_words(net);
  const int L = net.n_layers;
  for (int l = 0; l < L - 1; ++l) {
    dense_ffma<1>(a, b, net.W[l], net.b[l], net.kdim[l], net.np[l], mask + (size_t)l * mw * kRows, tid);
    float *t = a; a = b; b = t;
  }
  dense_ffma<0>(a, b, net.W[L - 1], net.b[L - 1], net.kdim[L - 1], net.np[L - 1], nullptr, tid);
#pragma unroll
  for (int j = 0; j < 4; ++j) y[j] = (j < net.n_out) ? net.out_loc[j] + net.out_scale[j] * b[j * kRows + tid] : 0.0f;
}

// input-VJP: g_y[4] -> g_x[8], using the masks recorded by net_forward_ffma
CSI_DEV void net_vjp_ffma(const NetDev &net, const float *g_y, float *g_x, float *bufA, float *bufB,
                          uint32_t *mask, int tid) {
  const int L = net.n_layers;
  const int mw = mask_words(net);
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (j < net.n_out) bufA[j * kRows + tid] = g_y[j] * net.out_scale[j];
  float *a = bufA, *b = bufB;
  for (int l = L - 1; l >= 1; --l) {
    // g_in[k] = sum_j g_out[j] W[k][j]  -> "K" = ndim[l], out width = kp[l]; then ReLU mask of layer l-1
    dense_ffma<2>(a, b, net.Wt[l], nullptr, net.ndim[l], net.kp[l], mask + (size_t)(l - 1) * mw * kRows, tid);
    float *t = a; a = b; b = t;
  }
  dense_ffma<0>(a, b, net.Wt[0], nullptr, net.ndim[0], net.kp[0], nullptr, tid);
#pragma unroll
  for (int k = 0; k < 8; ++k) g_x[k] = (k < net.n_in) ? b[k * kRows + tid] / net.in_scale[k] : 0.0f;
}

// BC emulator as a BC source for star_mid (FFMA engine): forward then VJP on the same buffers
struct BcNetFfma {
  const NetDev &net;
  float *bufA, *bufB;
  uint32_t *mask;
  int tid;
  CSI_DEV void forward(const float in[4], float val[4]) {
    float y[4] = {0.f, 0.f, 0.f, 0.f};
    const float x8[8] = {in[0], in[1], in[2], in[3], 0.f, 0.f, 0.f, 0.f};
    net_forward_ffma(net, x8, y, bufA, bufB, mask, tid);
#pragma unroll
    for (int b = 0; b < 4; ++b) val[b] = y[b];
  }
  CSI_DEV void vjp(const float gb[4], float gin[4]) {
    float gx[8];
    net_vjp_ffma(net, gb, gx, bufA, bufB, mask, tid);
#pragma unroll
    for (int d = 0; d < 4; ++d) gin[d] = gx[d];
  }
};

}  // namespace csi
