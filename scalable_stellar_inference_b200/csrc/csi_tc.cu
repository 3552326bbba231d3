// csi_tc.cu -- tcgen05 engine: the whole potential (priors -> emulator MLP fwd -> likelihood -> MLP input-VJP
// -> gradient stores -> reduction) as ONE persistent kernel for sm_100a.
//
// Tile = 128 stars = the 128 TMEM lanes; a thread of an "epilogue" warpgroup owns one star end to end.
// The five hidden 128x128 layers (forward) and their transposes (VJP) run on the tensor cores:
//
//   D[128 x 128] (TMEM, fp32)  =  A[128 x 128] (TMEM, fp16)  x  B[128 x 128] (smem, fp16, K-major)
//
// with every fp32 operand split into an fp16 (hi, lo) pair and three MMA passes
//   A_lo*B_hi + A_hi*B_lo + A_hi*B_hi          (fp32 accumulate in TMEM)
// which keeps ~23 significant bits per operand -- the parity bar is the reference's fp32 CPU path, which
// single-pass TF32/BF16/FP16 misses by 3 orders of magnitude (DESIGN.md "precision").  Operands are scaled
// by powers of two into the fp16 range (activations: fixed scale; cotangents: per-row scale), so the
// splitting is exact up to the dropped lo*lo term.
//
// Activations never leave the SM: the epilogue reads D with tcgen05.ld, applies bias / ReLU / mask,
// splits, and writes the next layer's A operand straight back to TMEM with tcgen05.st.  ReLU masks are
// 6 x 128 bits per star held in registers.  The first (5->128) and last (128->4) layers, the priors, the
// BC table gather and the likelihood run on the CUDA cores of the same thread.
//
// Weights: the 5 hidden layers x {forward, transposed} x {hi, lo} fp16 images (640 KB) do not fit in shared
// memory, so a loader warp streams the 64 KB image of the current step from L2 with bulk async copies into a
// 2-deep ring; each image is used by the TWO tiles a CTA keeps in flight (one warpgroup each), so that the
// MMAs of one tile overlap the epilogue of the other.
//
// Warp roles (320 threads): warps 0-3 epilogue of tile A, warps 4-7 epilogue of tile B, warp 8 MMA issuer
// (one elected lane) + TMEM allocator, warp 9 weight loader.
#include <cuda_fp16.h>

#include <cmath>
#include <cstring>

#include "csi_internal.h"
#include "csi_ptx.cuh"
#include "csi_reduce.cuh"

namespace csi {
namespace tc {

constexpr int kTile = 128;
constexpr int kHid = 128;
constexpr int kSteps = 10;                      // 5 forward + 5 transposed hidden layers
constexpr uint32_t kImgBytes = 32768;           // one fp16 128x128 operand image
constexpr uint32_t kStepBytes = 2 * kImgBytes;  // hi + lo
constexpr uint32_t kKStepBytes = 4096;          // 128 (N) x 16 (K) fp16
constexpr uint32_t kLbo = 128, kSbo = 256;
constexpr int kThreads = 320;
constexpr uint32_t kTmemCols = 512;
constexpr float kActScale = 512.0f;             // 2^9: activations up to ~127 stay finite in fp16

// offsets (floats) into the small-parameter blob
constexpr int oW0 = 0;                  // [5][128]
constexpr int oW0t = oW0 + 5 * 128;     // [128][8]
constexpr int oB = oW0t + 128 * 8;      // [6][128]
constexpr int oW6 = oB + 6 * 128;       // [128][4]
constexpr int oB6 = oW6 + 128 * 4;      // [4]
constexpr int oInLoc = oB6 + 4;         // [8]
constexpr int oInScale = oInLoc + 8;    // [8]
constexpr int oOutLoc = oInScale + 8;   // [4]
constexpr int oOutScale = oOutLoc + 4;  // [4]
constexpr int oW6Max = oOutScale + 4;   // [4] max_j |W6[j][c]|
constexpr int kSmallFloats = oW6Max + 4;

struct __align__(128) Smem {
  uint8_t wring[2][kStepBytes];  // weight ring: [buf][hi image | lo image]
  float small[kSmallFloats];
  uint32_t mask[24][256];        // ReLU masks: [hidden layer * 4 + word][epilogue thread]
  uint64_t w_full[2], w_empty[2], a_ready[2], d_ready[2];
  uint32_t tmem_base;
  double red[10 * 6];
  HyperVals hv;
};

// byte offset of element (n, k) inside one operand image (no-swizzle K-major core-matrix layout)
__host__ __device__ constexpr uint32_t img_offset(int n, int k) {
  return (uint32_t)(k >> 4) * kKStepBytes + (uint32_t)(n >> 3) * kSbo + (uint32_t)((k >> 3) & 1) * kLbo +
         (uint32_t)(n & 7) * 16 + (uint32_t)(k & 7) * 2;
}

// fp32 -> (hi, lo) fp16 pair: hi = rn16(v), lo = rn16(v - hi) (v - hi is exact in fp32); round-to-nearest on
// both halves keeps 23 significant bits and leaves the dropped lo*lo term without a preferred sign
__device__ __forceinline__ void split2(float v0, float v1, uint32_t &hi, uint32_t &lo) {
  __half2 hh = __floats2half2_rn(v0, v1);
  float2 hf = __half22float2(hh);
  __half2 ll = __floats2half2_rn(v0 - hf.x, v1 - hf.y);
  hi = *reinterpret_cast<uint32_t *>(&hh);
  lo = *reinterpret_cast<uint32_t *>(&ll);
}

// write 32 consecutive K-elements (chunk c) of this thread's row into the A operand region
__device__ __forceinline__ void store_a_chunk(uint32_t a_base, int c, const float (&v)[32]) {
  uint32_t hi[16], lo[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) split2(v[2 * q], v[2 * q + 1], hi[q], lo[q]);
  ptx::tmem_st16(a_base + 16 * c, hi);
  ptx::tmem_st16(a_base + 64 + 16 * c, lo);
}

// PREDICT = forward only (5 steps): writes predict's means / parameters instead of gradients (csi_predict).
template <int BC, bool PREDICT>
__global__ void __launch_bounds__(kThreads, 1) potential_tc_kernel(const __grid_constant__ Launch L) {
  constexpr int kNSteps = PREDICT ? 5 : kSteps;
  extern __shared__ unsigned char smem_raw[];
  Smem &S = *reinterpret_cast<Smem *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const DevModel &m = L.model;
  const StarIO &io = L.io;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long long n = io.n;
  const long long n_tiles = (n + kTile - 1) / kTile;
  const long long n_pairs = (n_tiles + 1) / 2;
  long long my_pairs = 0;
  if ((long long)blockIdx.x < n_pairs) my_pairs = (n_pairs - blockIdx.x + gridDim.x - 1) / gridDim.x;

  // ---- one-time setup ----------------------------------------------------------------------------
  for (int i = tid; i < kSmallFloats; i += kThreads) S.small[i] = __ldg(L.tc.small + i);
  if (tid == 0) {
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(&S.w_full[b], 1);
      ptx::mbar_init(&S.w_empty[b], 1);
      ptx::mbar_init(&S.a_ready[b], 128);  // every epilogue thread of the tile arrives
      ptx::mbar_init(&S.d_ready[b], 1);
    }
    ptx::fence_barrier_init();
    HyperVals h{};
    if (m.prior_mode == P_HIER) hyper_eval(m, io.z_hyper, h);
    S.hv = h;
  }
  if (warp == 8) ptx::tmem_alloc<kTmemCols>(&S.tmem_base);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = S.tmem_base;
  const HyperVals hv = S.hv;
  HyperAcc acc{};

  if (warp == 9) {
    // ===== weight loader: stream the 64 KB image of every step into the 2-deep ring ====================
    if (lane == 0) {
      const long long total = my_pairs * kNSteps;
      for (long long it = 0; it < total; ++it) {
        const int buf = (int)(it & 1);
        const int step = (int)(it % kNSteps);
        ptx::mbar_wait(&S.w_empty[buf], (uint32_t)(((it >> 1) & 1) ^ 1));
        ptx::mbar_arrive_expect_tx(&S.w_full[buf], kStepBytes);
        const uint8_t *src = L.tc.img + (size_t)step * kStepBytes;
#pragma unroll
        for (int q = 0; q < 4; ++q)
          ptx::bulk_g2s(&S.wring[buf][q * 16384], src + q * 16384, 16384, &S.w_full[buf]);
      }
    }
    __syncwarp();
  } else if (warp == 8) {
    // ===== MMA issuer ===============================================================================
    const uint32_t idesc = ptx::make_idesc_f16(128, 128);
    const long long total = my_pairs * kNSteps;
    for (long long it = 0; it < total; ++it) {
      const int buf = (int)(it & 1);
      ptx::mbar_wait(&S.w_full[buf], (uint32_t)((it >> 1) & 1));
      const uint32_t w_hi = ptx::smem_u32(&S.wring[buf][0]);
      const uint32_t w_lo = w_hi + kImgBytes;
      for (int t = 0; t < 2; ++t) {
        ptx::mbar_wait(&S.a_ready[t], (uint32_t)(it & 1));
        ptx::tc_fence_after();
        if (lane == 0) {
          const uint32_t d_t = tmem + (uint32_t)t * 256u;
          const uint32_t a_hi = d_t + 128u, a_lo = d_t + 192u;
          uint32_t accumulate = 0;
#pragma unroll
          for (int pass = 0; pass < 3; ++pass) {
            const uint32_t a_sel = pass == 0 ? a_lo : a_hi;
            const uint32_t w_sel = pass == 1 ? w_lo : w_hi;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint64_t bdesc = ptx::make_smem_desc(w_sel + (uint32_t)j * kKStepBytes, kLbo, kSbo);
              ptx::mma_f16_ts(d_t, a_sel + 8u * (uint32_t)j, bdesc, idesc, accumulate);
              accumulate = 1;
            }
          }
          ptx::mma_commit(&S.d_ready[t]);
        }
        __syncwarp();
      }
      if (lane == 0) ptx::mma_commit(&S.w_empty[buf]);
      __syncwarp();
    }
  } else {
    // ===== epilogue warpgroups: one thread = one star ===================================================
    const int t = warp >> 2;                   // tile slot within the pair
    const int row = (warp & 3) * 32 + lane;    // TMEM lane == row of the tile
    const uint32_t lane_base = ((uint32_t)((warp & 3) * 32)) << 16;
    const uint32_t d_base = tmem + lane_base + (uint32_t)t * 256u;
    const uint32_t a_base = d_base + 128u;
    const float *sm = S.small;
    const float inv_fwd = 1.0f / (kActScale * L.tc.w_scale);
    const float inv_wscale = 1.0f / L.tc.w_scale;
    long long it = 0;
    for (long long p = 0; p < my_pairs; ++p) {
      const long long pair = (long long)blockIdx.x + p * gridDim.x;
      const long long i = (pair * 2 + t) * kTile + row;
      const bool valid = i < n;
      StarState st;
      float x[5];
      if (valid) {
        star_pre(m, hv, io, i, st, acc);
#pragma unroll
        for (int k = 0; k < 5; ++k) x[k] = (st.x[k] - sm[oInLoc + k]) / sm[oInScale + k];
      } else {
#pragma unroll
        for (int k = 0; k < 5; ++k) x[k] = 0.0f;
      }
      uint32_t *mask = &S.mask[0][tid];  // word w of hidden layer h at mask[(h * 4 + w) * 256]
      // ---- layer 0 (5 -> 128) on the CUDA cores, ReLU, mask, split, A operand of step 0 ----------------
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        float v[32];
        uint32_t bits = 0;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const int j = 32 * c + 4 * q;
          float4 a = *reinterpret_cast<const float4 *>(sm + oB + j);
#pragma unroll
          for (int k = 0; k < 5; ++k) {
            const float4 w = *reinterpret_cast<const float4 *>(sm + oW0 + k * 128 + j);
            a.x = fmaf(x[k], w.x, a.x); a.y = fmaf(x[k], w.y, a.y);
            a.z = fmaf(x[k], w.z, a.z); a.w = fmaf(x[k], w.w, a.w);
          }
          v[4 * q] = a.x; v[4 * q + 1] = a.y; v[4 * q + 2] = a.z; v[4 * q + 3] = a.w;
        }
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const bool on = v[e] > 0.0f;
          bits |= (on ? 1u : 0u) << e;
          v[e] = on ? v[e] * kActScale : 0.0f;
        }
        mask[(0 * 4 + c) * 256] = bits;
        store_a_chunk(a_base, c, v);
      }
      ptx::tmem_wait_st();
      ptx::tc_fence_before();
      ptx::mbar_arrive(&S.a_ready[t]);

      float g_o[4] = {0.f, 0.f, 0.f, 0.f};
      float un = 1.0f;  // cotangent un-scaling: true value = D * un
#pragma unroll 1
      for (int step = 0; step < kNSteps; ++step, ++it) {
        ptx::mbar_wait(&S.d_ready[t], (uint32_t)(it & 1));
        ptx::tc_fence_after();
        if (step < 4) {
          // ---- hidden layer step+1 pre-activations -> ReLU -> next A operand -----------------------------
          const float *bias = sm + oB + (step + 1) * 128;
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t r[32];
            ptx::tmem_ld32(d_base + 32 * c, r);
            ptx::tmem_wait_ld();
            float v[32];
            uint32_t bits = 0;
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const float z = fmaf(__uint_as_float(r[e]), inv_fwd, bias[32 * c + e]);
              const bool on = z > 0.0f;
              bits |= (on ? 1u : 0u) << e;
              v[e] = on ? z * kActScale : 0.0f;
            }
            mask[((step + 1) * 4 + c) * 256] = bits;
            store_a_chunk(a_base, c, v);
          }
        } else if (step == 4) {
          // ---- last hidden layer -> output layer (128 -> 4) -> likelihood -> cotangent of hidden 5 --------
          const float *bias = sm + oB + 5 * 128;
          float y[4] = {sm[oB6], sm[oB6 + 1], sm[oB6 + 2], sm[oB6 + 3]};
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t r[32];
            ptx::tmem_ld32(d_base + 32 * c, r);
            ptx::tmem_wait_ld();
            uint32_t bits = 0;
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const float z = fmaf(__uint_as_float(r[e]), inv_fwd, bias[32 * c + e]);
              const bool on = z > 0.0f;
              bits |= (on ? 1u : 0u) << e;
              const float a = on ? z : 0.0f;
              const float4 w = *reinterpret_cast<const float4 *>(sm + oW6 + (32 * c + e) * 4);
              y[0] = fmaf(a, w.x, y[0]); y[1] = fmaf(a, w.y, y[1]);
              y[2] = fmaf(a, w.z, y[2]); y[3] = fmaf(a, w.w, y[3]);
            }
            mask[(5 * 4 + c) * 256] = bits;
          }
          float g_y[4] = {0.f, 0.f, 0.f, 0.f};
          if (valid) {
            float o[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) o[c] = sm[oOutLoc + c] + sm[oOutScale + c] * y[c];
            float mean_local[kMaxMean];
            float *mo = PREDICT ? mean_local : nullptr;
            if (BC == 0) { BcNone bc; star_mid(m, io, i, st, o, g_o, bc, mo); }
            if (BC == 1) { BcGrid bc{L.grid}; star_mid(m, io, i, st, o, g_o, bc, mo); }
#pragma unroll
            for (int c = 0; c < 4; ++c) g_y[c] = g_o[c] * sm[oOutScale + c];
            if (PREDICT) {
              if (L.mean_out) {
#pragma unroll
                for (int k = 0; k < kMaxMean; ++k) L.mean_out[i * kMaxMean + k] = mean_local[k];
              }
              if (L.params_out) {
                float4 *p = reinterpret_cast<float4 *>(L.params_out + i * 8);
                p[0] = make_float4(st.x[0], st.x[1], st.x[2], st.x[3]);
                p[1] = make_float4(st.x[4], st.log_mass, st.tau, st.dist);
              }
            }
          }
          if (!PREDICT) {
          // per-row power-of-two scale from a bound on |g_y W6^T|
          float bound = fabsf(g_y[0]) * sm[oW6Max] + fabsf(g_y[1]) * sm[oW6Max + 1] + fabsf(g_y[2]) * sm[oW6Max + 2] +
                        fabsf(g_y[3]) * sm[oW6Max + 3];
          float f = 1.0f;
          if (bound > 0.0f && bound < 3.0e38f) {
            int e = (int)((__float_as_uint(bound) >> 23) & 0xffu) - 127;
            e = max(-100, min(100, e));
            f = __uint_as_float((uint32_t)(9 - e + 127) << 23);
          }
          un = 1.0f / f * inv_wscale;  // after the MMA: D = sum (g f)(w s_w)  ->  g_in = D / (f s_w)
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            float v[32];
            const uint32_t bits = mask[(5 * 4 + c) * 256];
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const float4 w = *reinterpret_cast<const float4 *>(sm + oW6 + (32 * c + e) * 4);
              const float g = (g_y[0] * w.x + g_y[1] * w.y + g_y[2] * w.z + g_y[3] * w.w) * f;
              v[e] = ((bits >> e) & 1u) ? g : 0.0f;
            }
            store_a_chunk(a_base, c, v);
          }
          }  // !PREDICT
        } else if (step < 9) {
          // ---- cotangent of hidden layer h = 9 - step: ReLU mask, re-scale, next A operand ----------------
          const int h = 9 - step;
          float mx = 0.0f;
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t r[32];
            ptx::tmem_ld32(d_base + 32 * c, r);
            ptx::tmem_wait_ld();
#pragma unroll
            for (int e = 0; e < 32; ++e) mx = fmaxf(mx, fabsf(__uint_as_float(r[e])));
          }
          float f = 1.0f;
          if (mx > 0.0f && mx < 3.0e38f) {
            int e = (int)((__float_as_uint(mx) >> 23) & 0xffu) - 127;
            e = max(-100, min(100, e));
            f = __uint_as_float((uint32_t)(9 - e + 127) << 23);
          }
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t r[32];
            ptx::tmem_ld32(d_base + 32 * c, r);
            ptx::tmem_wait_ld();
            float v[32];
            const uint32_t bits = mask[(h * 4 + c) * 256];
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] = ((bits >> e) & 1u) ? __uint_as_float(r[e]) * f : 0.0f;
            store_a_chunk(a_base, c, v);
          }
          un = un / f * inv_wscale;
        } else {
          // ---- cotangent of hidden layer 0 -> layer-0 VJP (128 -> 5) -> gradient stores -----------------
          float gh[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t r[32];
            ptx::tmem_ld32(d_base + 32 * c, r);
            ptx::tmem_wait_ld();
            const uint32_t bits = mask[(0 * 4 + c) * 256];
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const float g = ((bits >> e) & 1u) ? __uint_as_float(r[e]) : 0.0f;
              const float4 w0 = *reinterpret_cast<const float4 *>(sm + oW0t + (32 * c + e) * 8);
              const float w4 = sm[oW0t + (32 * c + e) * 8 + 4];
              gh[0] = fmaf(g, w0.x, gh[0]); gh[1] = fmaf(g, w0.y, gh[1]); gh[2] = fmaf(g, w0.z, gh[2]);
              gh[3] = fmaf(g, w0.w, gh[3]); gh[4] = fmaf(g, w4, gh[4]);
            }
          }
          if (valid) {
            float g_x[5];
#pragma unroll
            for (int k = 0; k < 5; ++k) g_x[k] = gh[k] * un / sm[oInScale + k];
            star_post(m, io, i, st, g_x);
            acc.lp += (double)st.lp;
          }
        }
        if (step < kNSteps - 1) {
          ptx::tmem_wait_st();
          ptx::tc_fence_before();
          ptx::mbar_arrive(&S.a_ready[t]);
        } else {
          // all reads of D are complete (wait::ld above); order them before the next pair's MMAs
          ptx::tc_fence_before();
        }
      }
    }
  }

  // ---- teardown + reduction (all 320 threads) -----------------------------------------------------------
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 8) ptx::tmem_dealloc<kTmemCols>(tmem);
  if (!PREDICT) block_reduce_finalize(m, hv, acc, S.red, L.partials, L.counter, L.out6, L.grad_hyper);
}

}  // namespace tc

// -----------------------------------------------------------------------------------------------------
// host side
// -----------------------------------------------------------------------------------------------------
bool tc_supported(const csi_net &net) {
  return net.n_layers == 7 && net.n_in == 5 && net.n_hidden == 128 && net.n_out == 4;
}
size_t tc_image_bytes(const csi_net &) { return (size_t)tc::kSteps * tc::kStepBytes; }
size_t tc_small_floats(const csi_net &) { return tc::kSmallFloats; }

static uint16_t half_bits(float v) {
  __half h = __float2half_rn(v);
  uint16_t b;
  std::memcpy(&b, &h, 2);
  return b;
}

void tc_build_images(const csi_net &net, uint8_t *img, float *small, float *act_scale, float *w_scale) {
  using namespace tc;
  // weight scale: largest power of two with max|W| * s <= 32768 over the hidden layers
  float wmax = 0.f;
  for (int l = 1; l <= 5; ++l)
    for (int i = 0; i < kHid * kHid; ++i) wmax = std::fmax(wmax, std::fabs(net.weights[l][i]));
  float s = 1.0f;
  if (wmax > 0.f) s = std::exp2(std::floor(std::log2(32768.0f / wmax)));
  *w_scale = s;
  *act_scale = kActScale;
  std::memset(img, 0, (size_t)kSteps * kStepBytes);
  for (int step = 0; step < kSteps; ++step) {
    const bool fwd = step < 5;
    const int l = fwd ? step + 1 : 10 - step;  // bwd: layers 5,4,3,2,1
    const float *W = net.weights[l];           // [in][out]
    uint8_t *hi = img + (size_t)step * kStepBytes, *lo = hi + kImgBytes;
    for (int nn = 0; nn < kHid; ++nn)
      for (int k = 0; k < kHid; ++k) {
        // forward: B[n = out][k = in] = W[k][n];  transposed: B[n = in][k = out] = W[n][k]
        const float w = (fwd ? W[(size_t)k * kHid + nn] : W[(size_t)nn * kHid + k]) * s;
        const uint16_t hb = half_bits(w);
        __half hh;
        std::memcpy(&hh, &hb, 2);
        const uint16_t lb = half_bits(w - __half2float(hh));
        const uint32_t off = img_offset(nn, k);
        std::memcpy(hi + off, &hb, 2);
        std::memcpy(lo + off, &lb, 2);
      }
  }
  std::memset(small, 0, sizeof(float) * kSmallFloats);
  for (int k = 0; k < 5; ++k)
    for (int j = 0; j < 128; ++j) {
      small[oW0 + k * 128 + j] = net.weights[0][k * 128 + j];
      small[oW0t + j * 8 + k] = net.weights[0][k * 128 + j];
    }
  for (int l = 0; l < 6; ++l)
    for (int j = 0; j < 128; ++j) small[oB + l * 128 + j] = net.bias[l][j];
  for (int j = 0; j < 128; ++j)
    for (int c = 0; c < 4; ++c) {
      small[oW6 + j * 4 + c] = net.weights[6][j * 4 + c];
      small[oW6Max + c] = std::fmax(small[oW6Max + c], std::fabs(net.weights[6][j * 4 + c]));
    }
  for (int c = 0; c < 4; ++c) {
    small[oB6 + c] = net.bias[6][c];
    small[oOutLoc + c] = (float)net.out_loc[c];
    small[oOutScale + c] = (float)net.out_scale[c];
  }
  for (int k = 0; k < 5; ++k) {
    small[oInLoc + k] = (float)net.in_loc[k];
    small[oInScale + k] = (float)net.in_scale[k];
  }
}

template <int BC, bool PREDICT>
static int launch_tc_t(const Launch &L, cudaStream_t st, int *grid_out, int *smem_out) {
  const int smem = (int)sizeof(tc::Smem) + 1024;
  auto kern = tc::potential_tc_kernel<BC, PREDICT>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return -(int)e;
  const long long n_tiles = (L.io.n + tc::kTile - 1) / tc::kTile;
  const long long n_pairs = (n_tiles + 1) / 2;
  const int grid = (int)std::min<long long>(n_pairs, L.sm_count);
  kern<<<grid, tc::kThreads, smem, st>>>(L);
  e = cudaGetLastError();
  if (e != cudaSuccess) return -(int)e;
  if (grid_out) *grid_out = grid;
  if (smem_out) *smem_out = smem;
  return 1;
}

int launch_potential_tc(const Launch &L, cudaStream_t st, int *grid_out, int *smem_out) {
  if (L.model.bands > 0 && L.model.bc_mode == 1) return launch_tc_t<1, false>(L, st, grid_out, smem_out);
  return launch_tc_t<0, false>(L, st, grid_out, smem_out);
}

int launch_predict_tc(const Launch &L, cudaStream_t st) {
  if (L.model.bands > 0 && L.model.bc_mode == 1) return launch_tc_t<1, true>(L, st, nullptr, nullptr);
  return launch_tc_t<0, true>(L, st, nullptr, nullptr);
}

}  // namespace csi
