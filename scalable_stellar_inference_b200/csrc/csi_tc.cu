// csi_tc.cu -- tcgen05 engine: the whole potential (priors -> emulator MLP fwd -> likelihood -> MLP input-VJP
// -> gradient stores -> reduction) as ONE persistent kernel for sm_100a.
//
// Tile = 128 stars = the 128 TMEM lanes; a thread of an "epilogue" warpgroup owns one star end to end.
// ALL SEVEN layers of the emulator and ALL SEVEN transposed layers of its input-VJP run on the tensor cores:
//
//   D[128 x N] (TMEM, fp32)  =  A[128 x K] (TMEM, fp16)  x  B[N x K] (smem, fp16, K-major)
//
// with every fp32 operand split into an fp16 (hi, lo) pair and three products per contraction
//   A_lo*B_hi + A_hi*B_lo + A_hi*B_hi          (fp32 accumulate in TMEM, small terms first)
// which keeps ~23 significant bits per operand -- the parity bar is the reference's fp32 CPU path, which
// single-pass TF32/BF16/FP16 misses by 3 orders of magnitude (DESIGN.md "precision").  Operands are scaled
// by powers of two into the fp16 range (activations: fixed scale; cotangents: per-row scale), so the
// splitting is exact up to the dropped lo*lo term.  Forward biases are folded into the contraction (constant
// A entries x a 3-term fp16 split of the bias), so the epilogue never adds them.
//
// Stages of one tile (MMA -> epilogue), 14 per evaluation:
//    0      layer 0   (5 -> 128):   K = 32  [x_hi | x_lo | x_hi | c c c] x [W0_hi; W0_hi; W0_lo; bias split]
//    1..5   layers 1..5 (128 -> 128): 24 MMAs + 1 bias MMA, weights streamed (ring)
//    6      layer 6   (128 -> 4):   N = 16, 24 MMAs + bias; epilogue = likelihood + cotangent of the outputs
//    7      layer 6^T (4 -> 128):   K = 16  [g_hi | g_lo | g_hi | 0] x [W6_hi; W6_hi; W6_lo; 0]
//    8..12  layers 5..1 transposed, weights streamed (ring)
//    13     layer 0^T (128 -> 5):   N = 16, 24 MMAs; epilogue = chain rule to the sites + gradient stores
// Activations never leave the SM: the epilogue reads D with tcgen05.ld, applies ReLU / mask, splits, and
// writes the next stage's A operand straight back to TMEM with tcgen05.st.  ReLU masks are the SIGN bits of
// the pre-activations, 6 x 128 bits per star, gathered with one funnel shift per element and applied in the
// VJP with byte-permute sign replication on the packed fp16 pairs.  The priors, the BC table gather and the
// likelihood run on the CUDA cores of the same thread.
//
// Weights: the 5 hidden layers x {forward, transposed} x {hi, lo} fp16 images (+ bias tiles, 680 KB) do not
// fit in shared memory, so a loader warp streams the 68 KB image of the current stage from L2 with bulk async
// copies into a 2-deep ring; each image is used by the TWO tiles a CTA keeps in flight (one warpgroup each),
// so that the MMAs of one tile overlap the epilogue of the other.  The images of the four narrow stages
// (28.5 KB) stay resident in shared memory.
//
// Warp roles (320 threads): warps 0-3 epilogue of tile A, warps 4-7 epilogue of tile B, warp 8 MMA issuer
// (one elected lane) + TMEM allocator, warp 9 weight loader.
#include <cuda_fp16.h>

#include <cmath>
#include <cstring>
#include <vector>

#include "csi_internal.h"
#include "csi_ptx.cuh"
#include "csi_reduce.cuh"

namespace csi {
namespace tc {

constexpr int kTile = 128;
constexpr int kHid = 128;
constexpr int kStages = 14;                     // MMA stages per evaluation of a tile (7 in predict mode)
constexpr int kRingImages = 10;                 // 5 forward + 5 transposed hidden layers
constexpr uint32_t kImgBytes = 32768;           // one fp16 128x128 operand image
constexpr uint32_t kBiasBytes = 4096;           // one fp16 128 (N) x 16 (K) bias tile
constexpr uint32_t kStepBytes = 2 * kImgBytes + kBiasBytes;  // hi + lo + bias tile
constexpr uint32_t kKStepBytes = 4096;          // 128 (N) x 16 (K) fp16
constexpr uint32_t kKStep16 = 512;              // 16 (N) x 16 (K) fp16
constexpr uint32_t kLbo = 128, kSbo = 256;
constexpr uint32_t kHalfBytes = 16384;         // one K-half (4 K-steps) of an fp16 128x128 operand image
constexpr int kThreads = 384;  // 8 epilogue warps, 2 MMA issuers (one per tile), 2 weight loaders (one per tile)
constexpr uint32_t kTmemCols = 512;
constexpr int kGridPtsMax = 512;                // BC table axes kept in shared memory when they fit
constexpr float kActScale = 512.0f;             // 2^9: activations up to ~127 stay finite in fp16

// resident images (bytes into the resident blob)
constexpr uint32_t rB0 = 0;                     // layer 0:  N = 128, K = 32
constexpr uint32_t rB6 = rB0 + 2 * kKStepBytes; // layer 6:  N = 16, K = 128, hi then lo
constexpr uint32_t rB6Bias = rB6 + 2 * 8 * kKStep16;  // N = 16, K = 16
constexpr uint32_t rB7 = rB6Bias + kKStep16;    // layer 6^T: N = 128, K = 16
constexpr uint32_t rB13 = rB7 + kKStepBytes;    // layer 0^T: N = 16, K = 128, hi then lo
constexpr uint32_t kResidentBytes = rB13 + 2 * 8 * kKStep16;

// offsets (floats) into the small-parameter blob
constexpr int oInLoc = 0;               // [8]
constexpr int oInScale = oInLoc + 8;    // [8]
constexpr int oOutLoc = oInScale + 8;   // [4]
constexpr int oOutScale = oOutLoc + 4;  // [4]
constexpr int oGain = oOutScale + 4;    // [8] s_w * (largest row L1 norm) of the B image of stages 7..12
constexpr int oBiasA = oGain + 8;       // [1] constant A entry of the bias products (+3 pad)
constexpr int oInInv = oBiasA + 4;      // [8] 1 / in_scale (fp64 reciprocal, rounded once)
constexpr int kSmallFloats = oInInv + 8;

struct __align__(128) Smem {
  uint8_t wring[2][2][2 * kHalfBytes];  // weight slots: [tile][K-half][hi half | lo half]
  uint8_t wbias[2][kBiasBytes];         // bias tile of the tile's current forward stage (travels with K-half 0)
  uint8_t resident[kResidentBytes];
  uint8_t a_const[kBiasBytes];   // constant A tile of the bias MMA: columns k = 0..2 hold oBiasA, rest 0
  float small[kSmallFloats];
  uint32_t mask[24][256];        // ReLU sign words: [hidden layer * 4 + chunk][epilogue thread]
  float stage[kStageRows][kStageStride];  // per-thread input staging column (star_take_z / star_take_obs)
  StageList sites_list, obs_list;         // which arrays go into the column (built once, uniform copy loops)
  float grid_pts[kGridPtsMax];   // BC table axes + their lookup tables (copied once; the search then stays on-chip)
  unsigned short grid_lut[4][kLutBins];
  uint64_t w_full[2][2], w_empty[2][2], a_ready[2][2], d_ready[2], start_b;  // a_ready: one barrier per K-half
  uint32_t tmem_base;
  double red[(kThreads / 32) * 6];
  HyperVals hv;
};

static_assert(sizeof(Smem) + 1024 <= 232448, "shared memory of one CTA");

// byte offset of element (n, k) inside an operand image with 128 rows (no-swizzle K-major core matrices)
__host__ __device__ constexpr uint32_t img_offset(int n, int k) {
  return (uint32_t)(k >> 4) * kKStepBytes + (uint32_t)(n >> 3) * kSbo + (uint32_t)((k >> 3) & 1) * kLbo +
         (uint32_t)(n & 7) * 16 + (uint32_t)(k & 7) * 2;
}
// same for an image with 16 rows
__host__ __device__ constexpr uint32_t img_offset16(int n, int k) {
  return (uint32_t)(k >> 4) * kKStep16 + (uint32_t)(n >> 3) * kSbo + (uint32_t)((k >> 3) & 1) * kLbo +
         (uint32_t)(n & 7) * 16 + (uint32_t)(k & 7) * 2;
}

// fp32 -> (hi, lo) fp16 pair: hi = rn16(v), lo = rn16(v - hi); v - hi is formed exactly by the mixed-precision
// FMA (fma.rn.f32.f16: hi * -1 + v, SASS FHFMA).  Round-to-nearest on both halves keeps 23 significant bits
// and leaves the dropped lo*lo term without a preferred sign.
__device__ __forceinline__ void split2(float v0, float v1, uint32_t &hi, uint32_t &lo) {
  __half2 hh = __floats2half2_rn(v0, v1);
  hi = *reinterpret_cast<uint32_t *>(&hh);
  float l0, l1;
  asm("{\n\t.reg .f16 h0, h1, m;\n\tmov.b32 {h0, h1}, %2;\n\tmov.b16 m, 0xBC00;\n\t"
      "fma.rn.f32.f16 %0, h0, m, %3;\n\tfma.rn.f32.f16 %1, h1, m, %4;\n\t}"
      : "=f"(l0), "=f"(l1)
      : "r"(hi), "f"(v0), "f"(v1));
  __half2 ll = __floats2half2_rn(l0, l1);
  lo = *reinterpret_cast<uint32_t *>(&ll);
}

// (a, b) *= s as one packed FP32x2 multiply (SASS FMUL2)
__device__ __forceinline__ void mul2(float &a, float &b, float s) {
  asm("{\n\t.reg .b64 v, sc, o;\n\tmov.b64 v, {%0, %1};\n\tmov.b64 sc, {%2, %2};\n\tmul.rn.f32x2 o, v, sc;\n\t"
      "mov.b64 {%0, %1}, o;\n\t}"
      : "+f"(a), "+f"(b)
      : "f"(s));
}

// byte permute with sign replication (selector nibbles with bit 3 set replicate the msb of the chosen byte)
template <uint32_t kSel>
__device__ __forceinline__ uint32_t prmt_sign(uint32_t a) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(0u), "n"(kSel));
  return d;
}

// Element e = 4 s + q of a 32-element chunk is shifted into the sign word at step t = 8 q + s, i.e. ends at bit
// 31 - 8 q - s: after `word << s` the four elements of group s sit on the four byte msbs, which is what the
// sign-replicating byte permute of the VJP needs.  bit = 1 means pre-activation < 0 (unit inactive).
__device__ __forceinline__ uint32_t sign_word(const uint32_t (&r)[32]) {
  uint32_t bits = 0;
#pragma unroll
  for (int t = 0; t < 32; ++t) bits = __funnelshift_l(r[4 * (t & 7) + (t >> 3)], bits, 1);
  return bits;
}

// forward chunk: D (scaled pre-activations) -> ReLU -> activation scale -> (hi, lo) -> A operand columns
__device__ __forceinline__ void fwd_chunk(const uint32_t (&r)[32], float scale, uint32_t a_base, int c) {
  uint32_t hi[16], lo[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) {
    float v0 = fmaxf(__uint_as_float(r[2 * q]), 0.0f);
    float v1 = fmaxf(__uint_as_float(r[2 * q + 1]), 0.0f);
    mul2(v0, v1, scale);
    split2(v0, v1, hi[q], lo[q]);
  }
  ptx::tmem_st16(a_base + 32 * c, hi);       // in place over the D chunk just consumed:
  ptx::tmem_st16(a_base + 32 * c + 16, lo);  // columns [32c, 32c+16) = hi halves, [32c+16, 32c+32) = lo halves
}

// VJP chunk: D (cotangent before the ReLU mask) -> row scale -> (hi, lo) -> mask on the packed pairs -> A operand
__device__ __forceinline__ void bwd_chunk(const uint32_t (&r)[32], float f, uint32_t signs, uint32_t a_base, int c) {
  uint32_t hi[16], lo[16];
#pragma unroll
  for (int s = 0; s < 8; ++s) {
    const uint32_t w = signs << s;
    const uint32_t ma = prmt_sign<0xAABBu>(w);  // elements 4s (low half), 4s+1 (high half)
    const uint32_t mb = prmt_sign<0x8899u>(w);  // elements 4s+2, 4s+3
    uint32_t h0, l0, h1, l1;
    float a0 = __uint_as_float(r[4 * s]), a1 = __uint_as_float(r[4 * s + 1]);
    float a2 = __uint_as_float(r[4 * s + 2]), a3 = __uint_as_float(r[4 * s + 3]);
    mul2(a0, a1, f);
    mul2(a2, a3, f);
    split2(a0, a1, h0, l0);
    split2(a2, a3, h1, l1);
    hi[2 * s] = h0 & ~ma; lo[2 * s] = l0 & ~ma;
    hi[2 * s + 1] = h1 & ~mb; lo[2 * s + 1] = l1 & ~mb;
  }
  ptx::tmem_st16(a_base + 32 * c, hi);       // in place over the D chunk just consumed:
  ptx::tmem_st16(a_base + 32 * c + 16, lo);  // columns [32c, 32c+16) = hi halves, [32c+16, 32c+32) = lo halves
}

// power of two f with f * x in [2^9, 2^10) (1 when x is 0 / non-finite)
__device__ __forceinline__ float pow2_scale(float x) {
  float f = 1.0f;
  if (x > 0.0f && x < 3.0e38f) {
    int e = (int)((__float_as_uint(x) >> 23) & 0xffu) - 127;
    e = max(-100, min(100, e));
    f = __uint_as_float((uint32_t)(9 - e + 127) << 23);
  }
  return f;
}

// pack two fp16 bit patterns (low, high)
__device__ __forceinline__ uint32_t pack16(uint32_t lo16, uint32_t hi16) { return (lo16 & 0xffffu) | (hi16 << 16); }

// D chunks (4 x 32 columns) through a two-register-set pipeline: the tcgen05.ld of the next chunk is in flight
// while the current one is processed.  f(r, c) consumes chunk c (and may overwrite ITS columns: the A operand of
// the next stage is written in place, chunk by chunk).
template <class F>
__device__ __forceinline__ void for_chunks(uint32_t d_base, int ncc, F &&f) {
  uint32_t r0[32], r1[32];
  ptx::tmem_ld32(d_base, r0);
  ptx::tmem_wait_ld_dep(r0);
#pragma unroll 1
  for (int cc = 0; cc < ncc; ++cc) {
    ptx::tmem_ld32(d_base + 64 * cc + 32, r1);
    f(r0, 2 * cc);
    ptx::tmem_wait_ld_dep(r1);
    if (cc + 1 < ncc) ptx::tmem_ld32(d_base + 64, r0);
    f(r1, 2 * cc + 1);
    if (cc + 1 < ncc) ptx::tmem_wait_ld_dep(r0);
  }
}

// One K-HALF (4 K-steps = 64 contraction indices) of a K = 128 stage, issued as ONE straight-line run of 12 (13)
// tcgen05.mma: A_lo*B_hi, A_hi*B_lo, (bias), A_hi*B_hi -- small terms first within the half.  Measured on B200
// (profiles/r1_ubench_issue_patterns.txt): every break in the MMA instruction stream (a loop back-edge, a poll)
// costs ~250 cycles of tensor-pipe bubble, and per-thread (non-uniform) operands make issue slower than
// execution; so a run is as long as the pipelining allows, everything is "uniform base + immediate", and each
// tile has its own issuer warp so that one tile's bubble is filled by the other tile's MMAs.
// The A operand sits in place in the previous stage's D region: K-step kk has its hi halves at column
// 32 (kk >> 1) + 8 (kk & 1) and its lo halves 16 columns further.  blo_hi / blo_lo: LOW descriptor words of the
// B_hi / B_lo images at the half's first K-step (address field = bits 0..13 in 16-byte units: offsets add
// without carry); KSTEP16 = K-step stride in 16-byte units.
template <int HH, uint32_t KSTEP16>
__device__ __forceinline__ void issue_half(uint32_t d, uint32_t a_reg, uint32_t blo_hi, uint32_t blo_lo, uint32_t dhi,
                                           uint32_t idesc, bool bias, uint64_t a_const_desc, uint64_t bias_desc) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t a_lo = a_reg + 32u * (uint32_t)((4 * HH + j) >> 1) + 8u * (uint32_t)(j & 1) + 16u;
    ptx::mma_f16_ts(d, a_lo, ptx::desc_from_words(blo_hi + (uint32_t)j * KSTEP16, dhi), idesc, (HH | j) ? 1u : 0u);
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t a_hi = a_reg + 32u * (uint32_t)((4 * HH + j) >> 1) + 8u * (uint32_t)(j & 1);
    ptx::mma_f16_ts(d, a_hi, ptx::desc_from_words(blo_lo + (uint32_t)j * KSTEP16, dhi), idesc, 1u);
  }
  if (bias) ptx::mma_f16_ss(d, a_const_desc, bias_desc, idesc, 1);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t a_hi = a_reg + 32u * (uint32_t)((4 * HH + j) >> 1) + 8u * (uint32_t)(j & 1);
    ptx::mma_f16_ts(d, a_hi, ptx::desc_from_words(blo_hi + (uint32_t)j * KSTEP16, dhi), idesc, 1u);
  }
}

// clock64 trace (CSI_TC_DEBUG=8): a recorder appends (tag, clock) pairs to its 1024-entry region.  The marks cost
// ~3.5 % of the kernel even when idle (a predicated store pair and a counter per mark), so they -- and every use of
// TcWeights::debug -- exist only in the TRACE instantiation, which launches on a handle created under CSI_TC_DEBUG select.
template <bool ON>
struct Tracer {
  long long *p;
  int n;
  __device__ __forceinline__ void rec(int tag) {
    if constexpr (ON) {
      if (p != nullptr && n < 510) { p[2 * n] = tag; p[2 * n + 1] = clock64(); ++n; }
    }
  }
};

__device__ __forceinline__ bool ring_stage(int s) { return (s >= 1 && s <= 5) || (s >= 8 && s <= 12); }

// MODE 0: the potential.  MODE 1 (predict): forward only (stages 0..6), writes predict's means / parameters instead
// of gradients (csi_predict).  MODE 2 / 3: the emulator operator on its own (csi_net_forward / csi_net_vjp, plug
// point Emulator(backend)): x[n,5] -> y[n,4], and (x, gy) -> gx[n,5]; no priors, likelihood or reduction.
enum Mode { M_POTENTIAL = 0, M_PREDICT = 1, M_NET_FWD = 2, M_NET_VJP = 3 };
template <int BC, int MODE, bool TRACE = false>
__global__ void __launch_bounds__(kThreads, 1) potential_tc_kernel(const __grid_constant__ Launch L) {
  constexpr bool PREDICT = MODE == M_PREDICT, NET = MODE == M_NET_FWD || MODE == M_NET_VJP;
  constexpr bool FWD_ONLY = MODE == M_PREDICT || MODE == M_NET_FWD;
  constexpr int kNStages = FWD_ONLY ? 7 : kStages;
  constexpr int kNRing = FWD_ONLY ? 5 : kRingImages;
  extern __shared__ unsigned char smem_raw[];
  // align inside the shared window (keeps the pointer provably in the shared address space: LDS/STS, not LD/ST)
  Smem &S = *reinterpret_cast<Smem *>(smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u));
  if (L.stop != nullptr && *reinterpret_cast<const volatile int *>(L.stop) != 0) return;  // skipped leapfrog (csi_nuts.cuh)
  const DevModel &m = L.model;
  const StarIO &io = L.io;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long long n = io.n;
  const long long n_tiles = (n + kTile - 1) / kTile;
  // Work assignment at TILE granularity: tile k belongs to CTA k % grid; a CTA's tiles alternate between its two
  // tile slots (slot 0 takes the 1st, 3rd, ... -- one more than slot 1 when the count is odd).  Each slot has its
  // own epilogue warpgroup, MMA issuer and weight loader, so the slots simply run different iteration counts.
  long long my_tiles = 0;
  if ((long long)blockIdx.x < n_tiles) my_tiles = (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
  auto slot_iters = [&](int slot) { return (my_tiles + 1 - slot) / 2; };

  // CTA life-cycle marks of the traced CTA: 1 entry, 2 set-up done, 3 all tiles done, 4 reduction done
  Tracer<TRACE> ctr{nullptr, 0};
  if constexpr (TRACE)
    if (L.tc.trace && (int)blockIdx.x == (L.tc.debug >> 16) && tid == 0) ctr.p = L.tc.trace + 5 * 1024;
  ctr.rec(1);
  // ---- one-time setup ----------------------------------------------------------------------------
  // Warp 11 initialises the barriers and evaluates the hyper-parameter terms (a ~3 k-cycle dependent chain behind
  // a cold global load) while warps 0..10 copy the resident operands, so the two run side by side.
  constexpr int kCopyThreads = kThreads - 32;
  GridDev gl = L.grid;
  const bool grid_in_smem = BC == 1 && gl.shape[0] + gl.shape[1] + gl.shape[2] + gl.shape[3] <= kGridPtsMax;
  const uint32_t ca16 = (uint32_t)__half_as_ushort(__float2half_rn(__ldg(L.tc.small + oBiasA)));
  if (tid < kCopyThreads) {
    for (int i = tid; i < kSmallFloats; i += kCopyThreads) S.small[i] = __ldg(L.tc.small + i);
    for (int i = tid; i < (int)(kResidentBytes / 16); i += kCopyThreads)
      reinterpret_cast<uint4 *>(S.resident)[i] = __ldg(reinterpret_cast<const uint4 *>(L.tc.resident) + i);
    if (tid < 128) {
      // constant A tile of the bias MMA, K-major core-matrix layout: row tid, k = 0..2 -> c_A, k = 3..15 -> 0
      uint4 *row = reinterpret_cast<uint4 *>(S.a_const + (tid >> 3) * kSbo + (tid & 7) * 16);
      row[0] = make_uint4(ca16 | (ca16 << 16), ca16, 0u, 0u);
      row[kLbo / 16] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (grid_in_smem) {
      int off = 0;
#pragma unroll
      for (int d = 0; d < 4; ++d) {
        for (int i = tid; i < gl.shape[d]; i += kCopyThreads) S.grid_pts[off + i] = L.grid.pts[d][i];
        for (int i = tid; i < kLutBins; i += kCopyThreads) S.grid_lut[d][i] = L.grid.lut[d][i];
        off += gl.shape[d];
      }
    }
    if (!NET && tid == 32) { stage_list_sites(io, S.sites_list); stage_list_obs(m, io, S.obs_list); }
  } else if (tid == kCopyThreads) {
    for (int b = 0; b < 2; ++b) {
      for (int hh = 0; hh < 2; ++hh) {
        ptx::mbar_init(&S.w_full[b][hh], 1);
        ptx::mbar_init(&S.w_empty[b][hh], 1);  // one commit: the MMAs of this tile that read the slot
      }
      for (int c = 0; c < 2; ++c) ptx::mbar_init(&S.a_ready[b][c], 4);  // one elected lane per epilogue warp
      ptx::mbar_init(&S.d_ready[b], 1);
    }
    ptx::mbar_init(&S.start_b, 1);
    ptx::fence_barrier_init();
    HyperVals h{};
    if (m.prior_mode == P_HIER) hyper_eval(m, io.z_hyper, h);
    S.hv = h;
  }
  if (grid_in_smem) {
    int off = 0;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      gl.pts[d] = S.grid_pts + off;
      gl.lut[d] = S.grid_lut[d];
      off += gl.shape[d];
    }
  }
  ptx::fence_proxy_async_smem();  // generic-proxy writes above are read by tcgen05.mma (async proxy)
  if (warp == 8) ptx::tmem_alloc<kTmemCols>(&S.tmem_base);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = S.tmem_base;
  const HyperVals hv = S.hv;
  HyperAcc acc{};
  ctr.rec(2);

  if (warp >= 10) {
    // ===== weight loaders: warp 10 + t streams the images of tile slot t, one K-half per slot =============
    // The two tiles of a CTA run out of phase (see the start gate below), so each has its own stream: slot h of
    // tile t holds K-steps 4h..4h+3 of the (hi, lo) images of the tile's current 128 x 128 stage and is refilled
    // with the next stage's half as soon as the MMAs that read it have completed (tcgen05.commit -> w_empty).
    const int t = warp - 10;
    if (lane == 0) {
      const long long total = slot_iters(t) * kNRing * 2;
      for (long long it = 0; it < total; ++it) {
        const int hh = (int)(it & 1);
        const long long q = it >> 1;           // ring stage counter of this tile
        const int r = (int)(q % kNRing);
        const bool with_bias = hh == 0 && r < 5;  // only forward images carry a bias tile
        ptx::mbar_wait(&S.w_empty[t][hh], (uint32_t)((q & 1) ^ 1));
        ptx::mbar_arrive_expect_tx(&S.w_full[t][hh], 2 * kHalfBytes + (with_bias ? kBiasBytes : 0u));
        const uint8_t *src = L.tc.img + (size_t)r * kStepBytes + (size_t)hh * kHalfBytes;
        ptx::bulk_g2s(&S.wring[t][hh][0], src, kHalfBytes, &S.w_full[t][hh]);
        ptx::bulk_g2s(&S.wring[t][hh][kHalfBytes], src + kImgBytes, kHalfBytes, &S.w_full[t][hh]);
        if (with_bias)
          ptx::bulk_g2s(&S.wbias[t][0], L.tc.img + (size_t)r * kStepBytes + 2 * kImgBytes, kBiasBytes, &S.w_full[t][hh]);
      }
    }
    __syncwarp();
  } else if (warp >= 8) {
    // ===== MMA issuers: warp 8 + t issues the tensor work of tile slot t =================================
    // Everything here is kept PROVABLY warp-uniform (barrier tests through a vote, the TMEM base through a warp
    // reduction, all addresses derived from parameters, constants and loop counters): ptxas then holds the
    // tcgen05.mma operands in uniform registers instead of wrapping every UTCHMMA in an elect / R2UR / branch loop.
    const int t = warp - 8;
    const uint32_t full = 0xffffffffu;
    const uint32_t tmem_u = __reduce_or_sync(full, tmem) + (uint32_t)t * 256u;
    // hidden width <= 64 (BC emulator on zero-padded images): N = 64 and only the first K-half is contracted, so the
    // second half of A (never written: the epilogue converts chunks 0, 1 only) is never read
    const bool hk = NET && L.tc.half_k != 0;
    const uint32_t idesc128 = hk ? ptx::make_idesc_f16(128, 64) : ptx::make_idesc_f16(128, 128);
    const uint32_t idesc16 = ptx::make_idesc_f16(128, 16);
    const uint64_t a_const_desc = ptx::make_smem_desc(ptx::smem_u32(S.a_const), kLbo, kSbo);
    const uint32_t res = ptx::smem_u32(S.resident);
    // descriptor words, computed once: low word = address field | leading-byte offset, high word = stride | version
    auto lo_word = [](uint32_t addr) { return ((addr & 0x3FFFFu) >> 4) | ((kLbo >> 4) << 16); };
    const uint32_t dhi = (kSbo >> 4) | (1u << 14);
    const uint32_t ring_lo = lo_word(ptx::smem_u32(&S.wring[t][0][0]));
    const uint64_t bias_desc = ptx::make_smem_desc(ptx::smem_u32(&S.wbias[t][0]), kLbo, kSbo);
    const uint32_t r6_lo = lo_word(res + rB6), r13_lo = lo_word(res + rB13);
    const uint64_t r6_bias_desc = ptx::make_smem_desc(res + rB6Bias, kLbo, kSbo);
    const uint64_t r0_desc0 = ptx::make_smem_desc(res + rB0, kLbo, kSbo);
    const uint64_t r0_desc1 = ptx::make_smem_desc(res + rB0 + kKStepBytes, kLbo, kSbo);
    const uint64_t r7_desc = ptx::make_smem_desc(res + rB7, kLbo, kSbo);
    constexpr uint32_t kSlot16 = 2 * kHalfBytes / 16, kLoOff16 = kHalfBytes / 16, kK128 = kKStepBytes / 16, kK16 = kKStep16 / 16;
    auto wait_bar = [&](uint64_t *bar, uint32_t parity) {
      while (!__all_sync(full, ptx::mbar_try_wait(bar, parity))) {
      }
    };
    Tracer<TRACE> tr{nullptr, 0};
    if constexpr (TRACE)
      if (L.tc.trace && blockIdx.x == 0 && lane == 0 && t == 0) tr.p = L.tc.trace + 4 * 1024;
    constexpr int kGateStage = FWD_ONLY ? 2 : 4;  // tile B starts when tile A's stage kGateStage is issued
    uint32_t k = 0, q = 0;            // stage counter of the tile's whole life (region parity); streamed-stage counter
    uint32_t apar0 = 0u, apar1 = 0u;  // parities of the next phases of a_ready[t][0 / 1]
    const long long my_iters = slot_iters(t);
    for (long long p = 0; p < my_iters; ++p) {
#pragma unroll 1
      for (int s = 0; s < kNStages; ++s, ++k) {
        const bool ring = ring_stage(s);
        const uint32_t d_t = tmem_u + (k & 1u) * 128u;             // D of this stage
        const uint32_t a_reg = tmem_u + ((k & 1u) ^ 1u) * 128u;    // A: in place in the previous stage's D region
        wait_bar(&S.a_ready[t][0], apar0);
        apar0 ^= 1u;
        tr.rec(1000 + 100 * t + s);
        if (s == 0 || s == 7) {
          if (ptx::elect_one()) {
            if (s == 0) {
              ptx::mma_f16_ts(d_t, a_reg, r0_desc0, idesc128, 0);
              ptx::mma_f16_ts(d_t, a_reg + 8u, r0_desc1, idesc128, 1);
            } else {
              ptx::mma_f16_ts(d_t, a_reg, r7_desc, idesc128, 0);
            }
            ptx::mma_commit(&S.d_ready[t]);
          }
        } else {
          // K = 128: the A operand arrives in two K-halves; the MMAs of half 0 run under the epilogue of half 1
          if (ring) wait_bar(&S.w_full[t][0], q & 1u);
          if (ptx::elect_one()) {
            if (ring) {
              issue_half<0, kK128>(d_t, a_reg, ring_lo, ring_lo + kLoOff16, dhi, idesc128, s <= 5, a_const_desc, bias_desc);
              ptx::mma_commit(&S.w_empty[t][0]);  // slot 0 and the bias tile are refillable once these complete
            } else if (s == 6) {
              issue_half<0, kK16>(d_t, a_reg, r6_lo, r6_lo + 8 * kK16, dhi, idesc16, true, a_const_desc, r6_bias_desc);
            } else {
              issue_half<0, kK16>(d_t, a_reg, r13_lo, r13_lo + 8 * kK16, dhi, idesc16, false, a_const_desc, r6_bias_desc);
            }
            if (t == 0 && p == 0 && s == kGateStage) ptx::mbar_arrive(&S.start_b);
          }
          __syncwarp();
          wait_bar(&S.a_ready[t][1], apar1);
          apar1 ^= 1u;
          if (ring) wait_bar(&S.w_full[t][1], q & 1u);
          if (ptx::elect_one()) {
            if (ring) {
              if (!hk)
                issue_half<1, kK128>(d_t, a_reg, ring_lo + kSlot16, ring_lo + kSlot16 + kLoOff16, dhi, idesc128, false,
                                     a_const_desc, bias_desc);
              ptx::mma_commit(&S.w_empty[t][1]);
            } else if (hk) {
            } else if (s == 6) {
              issue_half<1, kK16>(d_t, a_reg, r6_lo + 4 * kK16, r6_lo + 12 * kK16, dhi, idesc16, false, a_const_desc, r6_bias_desc);
            } else {
              issue_half<1, kK16>(d_t, a_reg, r13_lo + 4 * kK16, r13_lo + 12 * kK16, dhi, idesc16, false, a_const_desc, r6_bias_desc);
            }
            ptx::mma_commit(&S.d_ready[t]);
          }
          if (ring) ++q;
        }
        __syncwarp();
        tr.rec(2000 + 100 * t + s);
      }
    }
  } else {
    // ===== epilogue warpgroups: one thread = one star ===================================================
    const int t = warp >> 2;                   // tile slot within the pair
    const int row = (warp & 3) * 32 + lane;    // TMEM lane == row of the tile
    const uint32_t lane_base = ((uint32_t)((warp & 3) * 32)) << 16;
    // TMEM of a tile: two 128-column regions.  Stage k (counted over the tile's whole life, `it`) accumulates D
    // into region k & 1; its epilogue converts D into the next stage's A operand IN PLACE, so stage k + 1 reads A
    // from region k & 1 and accumulates into the other one.  Because A appears chunk by chunk, the MMAs of the
    // next stage start on K-chunk c while the epilogue is still converting chunk c + 1.
    const uint32_t t_base = tmem + lane_base + (uint32_t)t * 256u;
    const float *sm = S.small;
    const float inv_fwd = 1.0f / (kActScale * L.tc.w_scale);
    const float inv_wscale = 1.0f / L.tc.w_scale;
    uint32_t *mask = &S.mask[0][tid];  // sign word of chunk c of hidden layer h at mask[(h * 4 + c) * 256]
    // chunk pairs per 128-wide epilogue: a 64-wide network (operator modes only) has columns 0..63 only
    const int ncc = (NET && L.tc.half_k) ? 1 : 2;
    // hand the A operand to the MMA warp: all tcgen05.st of this warp done, then one elected arrive
    // (one barrier per K-half: a half's next phase belongs to the next stage, which cannot begin before the MMA
    // warp has consumed this one, so the producer can never run two phases ahead of the consumer's parity)
    auto publish_a = [&](int c) {
      ptx::tmem_wait_st();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&S.a_ready[t][c]);
    };
    Tracer<TRACE> tr{nullptr, 0};
    long long *tr_base = nullptr;  // trace of CTA (debug >> 16), starting at this slot's iteration (debug >> 8) & 255
    if constexpr (TRACE)
      if (L.tc.trace && (int)blockIdx.x == (L.tc.debug >> 16) && lane == 0 && (warp == 0 || warp == 3 || warp == 4 || warp == 7))
        tr_base = L.tc.trace + (warp == 0 ? 0 : warp == 3 ? 1 : warp == 4 ? 2 : 3) * 1024;
    long long it = 0;
    // Start gate: tile B begins a quarter of an evaluation chain after tile A, so that the long CUDA-core phases
    // of one tile (priors; table gather + likelihood; gradient stores) run under the MMA stages of the other.
    const long long my_iters = slot_iters(t);
    // Input staging: a thread copies its star's inputs into its own column of S.stage with 4-byte asynchronous
    // copies, most of a tile before it needs them -- the sites of tile p + 1 once star_mid of tile p has taken the
    // observations out of the column, the observations of tile p once its priors have taken the sites.  Nothing
    // but the thread itself touches the column, so cp.async.wait_all is the only synchronisation.
    float *col = &S.stage[0][tid];
    auto stage = [&](const StageList &l, long long ii) {
      if (NET || ii >= n) return;
      const int cnt = l.n;
#pragma unroll 4
      for (int j = 0; j < cnt; ++j) ptx::cp_async4(col + l.row[j] * kStageStride, l.src[j] + ii);
    };
    auto stage_sites = [&](long long ii) { stage(S.sites_list, ii); };
    auto stage_obs = [&](long long ii) { stage(S.obs_list, ii); };
    if (my_iters > 0) stage_sites(((long long)blockIdx.x + t * gridDim.x) * kTile + row);
    if (t == 1 && my_iters > 0) ptx::mbar_wait(&S.start_b, 0u);
    for (long long p = 0; p < my_iters; ++p) {
      const long long tile = (long long)blockIdx.x + (2 * p + t) * gridDim.x;
      const long long i = tile * kTile + row;
      const bool valid = i < n;
      if constexpr (TRACE)
        if (tr_base && p >= ((L.tc.debug >> 8) & 255)) tr.p = tr_base;
      const long long i_next = p + 1 < my_iters ? i + 2LL * kTile * gridDim.x : n;  // same row, this slot's next tile
      StarState st;
      float x[5];
      if (valid && NET) {
#pragma unroll
        for (int k = 0; k < 5; ++k)
          x[k] = ((k < L.net_nin ? L.net_x[i * L.net_nin + k] : 0.0f) - sm[oInLoc + k]) * (sm[oInInv + k] * kActScale);
      } else if (valid) {
        tr.rec(501);
        StarIn in;
        ptx::cp_async_wait_all();
        star_take_z(m, io, col, in);
        if (TRACE && tr.p) {  // trace only: stall until the loads have landed so that their latency shows up on its own
          uint32_t x_ = __float_as_uint(in.mloc) ^ __float_as_uint(in.mscale) ^ __float_as_uint(in.mlogz);
#pragma unroll
          for (int sl = 0; sl < kSlots; ++sl) x_ ^= __float_as_uint(in.zv[sl]);
          if (x_ == 0xdeadbeefu) tr.rec(999);
          tr.rec(502);
        }
        star_pre(m, hv, in, st, acc);
#pragma unroll
        for (int k = 0; k < 5; ++k) x[k] = (st.x[k] - sm[oInLoc + k]) * (sm[oInInv + k] * kActScale);
        if (TRACE && tr.p) { if (x[0] + x[1] + x[2] + x[3] + x[4] + st.lp == 1.2345e30f) tr.rec(999); tr.rec(503); }
      } else {
#pragma unroll
        for (int k = 0; k < 5; ++k) x[k] = 0.0f;
      }
      // ---- A operand of stage 0: [x_hi(5) | x_lo(5) | x_hi(5) | c_A c_A c_A | 0 ...], K = 32 -> 16 columns ----
      {
        uint32_t h01, l01, h23, l23, h4, l4;
        split2(x[0], x[1], h01, l01);
        split2(x[2], x[3], h23, l23);
        split2(x[4], 0.0f, h4, l4);
        uint32_t a0[16];
        a0[0] = h01; a0[1] = h23; a0[2] = pack16(h4, l01); a0[3] = pack16(l01 >> 16, l23);
        a0[4] = pack16(l23 >> 16, l4); a0[5] = h01; a0[6] = h23; a0[7] = pack16(h4, ca16);
        a0[8] = ca16 | (ca16 << 16);
#pragma unroll
        for (int q = 9; q < 16; ++q) a0[q] = 0u;
        ptx::tmem_st16(t_base + (uint32_t)(((it & 1) ^ 1) * 128), a0);  // A of stage `it` lives in the other region
      }
      tr.rec(504);
      publish_a(0);
      tr.rec(505);
      stage_obs(i);  // the column is free: the priors hold the sites in registers

      float un = 1.0f;    // cotangent un-scaling: true value = D * un
      float amax = 0.0f;  // max |A operand| of the current VJP stage (after scaling)
#pragma unroll 1
      for (int s = 0; s < kNStages; ++s, ++it) {
        tr.rec(100 + s);
        ptx::mbar_wait(&S.d_ready[t], (uint32_t)(it & 1));
        ptx::tc_fence_after();
        const uint32_t d_base = t_base + (uint32_t)((it & 1) * 128);
        const uint32_t a_base = d_base;
        tr.rec(200 + s);
        if (s <= 5) {
          // ---- pre-activations of hidden layer s (bias already in D) -> sign word, ReLU -> next A operand ----
          uint32_t *mk = mask + s * 4 * 256;
          for_chunks(d_base, ncc, [&](const uint32_t (&r)[32], int c) {
            mk[c * 256] = sign_word(r);
            fwd_chunk(r, inv_wscale, a_base, c);
            if (c & 1) publish_a(c >> 1);  // K-half c >> 1 of the next stage's A operand is complete
          });
          if (ncc == 1) publish_a(1);  // 64-wide network: the second K-half does not exist (protocol only)
        } else if (s == 6) {
          // ---- emulator outputs -> predict + likelihood -> cotangent of the outputs -> A operand of stage 7 ----
          uint32_t r8[8];
          ptx::tmem_ld8(d_base, r8);
          ptx::tmem_wait_ld_dep8(r8);
          float g_y[4] = {0.f, 0.f, 0.f, 0.f};
          if (valid) {
            float o[4], g_o[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) o[c] = sm[oOutLoc + c] + sm[oOutScale + c] * (__uint_as_float(r8[c]) * inv_fwd);
            float mean_local[kMaxMean];
            float *mo = PREDICT ? mean_local : nullptr;
            if (NET) {
              if (MODE == M_NET_FWD) {
#pragma unroll
                for (int c = 0; c < 4; ++c)
                  if (c < L.net_nout) L.net_y[i * L.net_nout + c] = o[c];
              } else {
#pragma unroll
                for (int c = 0; c < 4; ++c) g_o[c] = c < L.net_nout ? L.net_gy[i * L.net_nout + c] : 0.0f;
              }
            } else {
            StarObs so;
            tr.rec(601);
            ptx::cp_async_wait_all();
            tr.rec(611);
            star_take_obs(m, io, col, so);
            if (TRACE && tr.p) {
              uint32_t x_ = 0;
#pragma unroll
              for (int k = 0; k < kMaxMean; ++k) x_ ^= __float_as_uint(so.ob[k]) ^ __float_as_uint(so.dg[k]);
              if (x_ == 0xdeadbeefu) tr.rec(999);
              tr.rec(602);
            }
            if (BC == 0) { BcNone bc; star_mid(m, io, i, st, o, g_o, bc, so, mo); }
            else if (BC == 1) { BcGrid bc{gl}; star_mid(m, io, i, st, o, g_o, bc, so, mo); }
            stage_sites(i_next);  // the observations have been consumed: the column takes the next tile's sites
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) g_y[c] = g_o[c] * sm[oOutScale + c];
            if (PREDICT) {
              if (L.mean_out) {
#pragma unroll
                for (int k = 0; k < kMaxMean; ++k) L.mean_out[i * kMaxMean + k] = mean_local[k];
              }
              if (L.params_out) {
                float4 *pp = reinterpret_cast<float4 *>(L.params_out + i * 8);
                pp[0] = make_float4(st.x[0], st.x[1], st.x[2], st.x[3]);
                pp[1] = make_float4(st.x[4], st.log_mass, st.tau, st.dist);
              }
            }
          }
          if (!FWD_ONLY) {
            // [g_hi(4) | g_lo(4) | g_hi(4) | 0(4)] with a per-row power-of-two scale, K = 16 -> 8 columns
            const float gmax = fmaxf(fmaxf(fabsf(g_y[0]), fabsf(g_y[1])), fmaxf(fabsf(g_y[2]), fabsf(g_y[3])));
            const float f = pow2_scale(gmax);
            un = 1.0f / f * inv_wscale;  // after the MMA: D = sum (g f)(w s_w)  ->  g_in = D / (f s_w)
            amax = gmax * f;
            uint32_t h01, l01, h23, l23;
            split2(g_y[0] * f, g_y[1] * f, h01, l01);
            split2(g_y[2] * f, g_y[3] * f, h23, l23);
            const uint32_t ag[8] = {h01, h23, l01, l23, h01, h23, 0u, 0u};
            ptx::tmem_st8(a_base, ag);
          }
        } else if (s <= 12) {
          // ---- cotangent of hidden layer h = 12 - s: re-scale, ReLU mask, next A operand --------------------
          // |D| <= amax * (s_w * G) with G = the largest row L1 norm of this stage's B image, so the scale is
          // known BEFORE D is read (one pass); the exact row maximum is gathered for the next stage.
          const int h = 12 - s;
          const float f = pow2_scale(amax * sm[oGain + (s - 7)]) * 32.0f;  // f * |D| < 2^15
          float mx = 0.0f;
          const uint32_t *mk = mask + h * 4 * 256;
          for_chunks(d_base, ncc, [&](const uint32_t (&r)[32], int c) {
#pragma unroll
            for (int e = 0; e < 32; e += 2)
              mx = fmaxf(mx, fmaxf(fabsf(__uint_as_float(r[e])), fabsf(__uint_as_float(r[e + 1]))));
            bwd_chunk(r, f, mk[c * 256], a_base, c);
            if (c & 1) publish_a(c >> 1);
          });
          if (ncc == 1) publish_a(1);
          amax = mx * f;
          un = un / f * inv_wscale;
        } else {
          // ---- cotangent of the (normalised) emulator inputs -> chain rule to the sites -> gradient stores ------
          uint32_t r8[8];
          ptx::tmem_ld8(d_base, r8);
          ptx::tmem_wait_ld_dep8(r8);
          if (valid) {
            float g_x[5];
#pragma unroll
            for (int k = 0; k < 5; ++k) g_x[k] = __uint_as_float(r8[k]) * un * sm[oInInv + k];
            if (NET) {
#pragma unroll
              for (int k = 0; k < 5; ++k)
                if (k < L.net_nin) L.net_gx[i * L.net_nin + k] = g_x[k];
            } else {
              star_post(m, io, i, st, g_x);
              acc.lp += (double)st.lp;
            }
          }
        }
        tr.rec(300 + s);
        if (s == 6 && !FWD_ONLY) {
          publish_a(0);  // the narrow A operand of stage 7 (the 128-wide stages publish chunk by chunk above)
          tr.rec(400 + s);
        } else if (s == kNStages - 1) {
          // all reads of D are complete (wait::ld above); order them before the next pair's MMAs
          ptx::tc_fence_before();
        }
      }
    }
  }

  // ---- teardown + reduction (all threads) -----------------------------------------------------------
  ptx::tc_fence_before();
  __syncthreads();
  ctr.rec(3);
  if (warp == 8) ptx::tmem_dealloc<kTmemCols>(tmem);
  if (MODE == M_POTENTIAL) block_reduce_finalize(m, hv, acc, S.red, L.partials, L.counter, L.out6, L.grad_hyper, L.px);
  ctr.rec(4);
}

}  // namespace tc

// -----------------------------------------------------------------------------------------------------
// host side
// -----------------------------------------------------------------------------------------------------
static uint16_t half_bits(float v) {
  __half h = __float2half_rn(v);
  uint16_t b;
  std::memcpy(&b, &h, 2);
  return b;
}
static float half_value(uint16_t b) {
  __half h;
  std::memcpy(static_cast<void *>(&h), &b, 2);
  return __half2float(h);
}
static void put16(uint8_t *base, uint32_t off, uint16_t v) { std::memcpy(base + off, &v, 2); }
// (hi, lo) fp16 split of a host value, round-to-nearest on both halves
static void split_host(float v, uint16_t *hi, uint16_t *lo) {
  *hi = half_bits(v);
  *lo = half_bits(v - half_value(*hi));
}

// operand scalings: weights by s_w = the largest power of two with max|W| s_w <= 32768 (all layers); the bias
// products use a constant A entry c_A = the smallest power of two with act_scale * s_w * max|b| / c_A <= 32768
static void tc_scales(const csi_net &net, float *w_scale, float *bias_a) {
  float wmax = 0.f, bmax = 0.f;
  for (int l = 0; l < 7; ++l) {
    const int K = l == 0 ? 5 : tc::kHid, N = l == 6 ? 4 : tc::kHid;
    for (int i = 0; i < K * N; ++i) wmax = std::fmax(wmax, std::fabs(net.weights[l][i]));
    for (int i = 0; i < N; ++i) bmax = std::fmax(bmax, std::fabs(net.bias[l][i]));
  }
  float s = 1.0f;
  if (wmax > 0.f) s = std::exp2(std::floor(std::log2(32768.0f / wmax)));
  float ca = 1.0f;
  const float need = tc::kActScale * s * bmax / 32768.0f;
  if (need > 1.0f) ca = std::exp2(std::ceil(std::log2(need)));
  *w_scale = s;
  *bias_a = ca;
}

// A network narrower than the kernel's 5 -> 128 x 6 -> 4 shape (the BC emulator: 4 -> 64 x 6 -> 3) runs on the same
// operand layout zero-padded: absent inputs get weight 0 (and loc 0 / scale 1), absent hidden units weight 0 and bias
// 0 (ReLU(0) = 0 contributes nothing, its cotangent is multiplied by zero weights), absent outputs weight 0.
struct PaddedNet {
  std::vector<float> w[7], b[7];
  double in_loc[5], in_scale[5], out_loc[4], out_scale[4];
  csi_net net{};
};
static bool pad_net(const csi_net &src, PaddedNet &p) {
  if (!(src.n_layers == 7 && src.n_in >= 1 && src.n_in <= 5 && src.n_hidden >= 8 && src.n_hidden <= tc::kHid &&
        src.n_out >= 1 && src.n_out <= 4))
    return false;
  for (int l = 0; l < 7; ++l) {
    const int K = l == 0 ? 5 : tc::kHid, N = l == 6 ? 4 : tc::kHid;
    const int k0 = l == 0 ? src.n_in : src.n_hidden, n0 = l == 6 ? src.n_out : src.n_hidden;
    p.w[l].assign((size_t)K * N, 0.0f);
    p.b[l].assign(N, 0.0f);
    for (int k = 0; k < k0; ++k)
      for (int j = 0; j < n0; ++j) p.w[l][(size_t)k * N + j] = src.weights[l][(size_t)k * n0 + j];
    for (int j = 0; j < n0; ++j) p.b[l][j] = src.bias[l][j];
    p.net.weights[l] = p.w[l].data();
    p.net.bias[l] = p.b[l].data();
  }
  for (int k = 0; k < 5; ++k) { p.in_loc[k] = k < src.n_in ? src.in_loc[k] : 0.0; p.in_scale[k] = k < src.n_in ? src.in_scale[k] : 1.0; }
  for (int j = 0; j < 4; ++j) { p.out_loc[j] = j < src.n_out ? src.out_loc[j] : 0.0; p.out_scale[j] = j < src.n_out ? src.out_scale[j] : 1.0; }
  p.net.n_layers = 7; p.net.n_in = 5; p.net.n_hidden = tc::kHid; p.net.n_out = 4;
  p.net.in_loc = p.in_loc; p.net.in_scale = p.in_scale; p.net.out_loc = p.out_loc; p.net.out_scale = p.out_scale;
  return true;
}

bool tc_supported(const csi_net &src) {
  PaddedNet p;
  if (!pad_net(src, p)) return false;
  float s, ca;
  tc_scales(p.net, &s, &ca);
  return std::isfinite(s) && ca <= 32768.0f;  // the bias tile must fit the fp16 range
}
bool tc_half_k(const csi_net &net) { return net.n_hidden <= 64; }
size_t tc_image_bytes(const csi_net &) { return (size_t)tc::kRingImages * tc::kStepBytes; }
size_t tc_resident_bytes(const csi_net &) { return tc::kResidentBytes; }
size_t tc_small_floats(const csi_net &) { return tc::kSmallFloats; }

void tc_build_images(const csi_net &src, uint8_t *img, uint8_t *res, float *small, float *act_scale, float *w_scale) {
  using namespace tc;
  PaddedNet padded;
  pad_net(src, padded);
  const csi_net &net = padded.net;
  float s, ca;
  tc_scales(net, &s, &ca);
  *w_scale = s;
  *act_scale = kActScale;
  const float bias_mul = kActScale * s / ca;
  const double slack = 1.0 + 1.0 / 256.0;  // covers the rounding of the split operands in the a-priori bounds
  std::memset(img, 0, (size_t)kRingImages * kStepBytes);
  std::memset(res, 0, kResidentBytes);
  std::memset(small, 0, sizeof(float) * kSmallFloats);
  auto put_bias = [&](uint8_t *tile, uint32_t off0, uint32_t off1, uint32_t off2, float b) {
    const uint16_t b0 = half_bits(b);
    const float r1 = b - half_value(b0);
    const uint16_t b1 = half_bits(r1);
    const uint16_t b2 = half_bits(r1 - half_value(b1));
    put16(tile, off0, b0); put16(tile, off1, b1); put16(tile, off2, b2);
  };
  // ---- streamed images: forward layers 1..5, then transposed layers 5..1 ----------------------------------
  for (int r = 0; r < kRingImages; ++r) {
    const bool fwd = r < 5;
    const int l = fwd ? r + 1 : 10 - r;
    const float *W = net.weights[l];  // [in][out]
    uint8_t *hi = img + (size_t)r * kStepBytes, *lo = hi + kImgBytes, *bt = lo + kImgBytes;
    double gain = 0.0;
    for (int nn = 0; nn < kHid; ++nn) {
      double l1 = 0.0;
      for (int k = 0; k < kHid; ++k) {
        // forward: B[n = out][k = in] = W[k][n];  transposed: B[n = in][k = out] = W[n][k]
        const float w = (fwd ? W[(size_t)k * kHid + nn] : W[(size_t)nn * kHid + k]) * s;
        uint16_t hb, lb;
        split_host(w, &hb, &lb);
        put16(hi, img_offset(nn, k), hb);
        put16(lo, img_offset(nn, k), lb);
        l1 += std::fabs((double)w);
      }
      gain = std::fmax(gain, l1);
      if (fwd) put_bias(bt, img_offset(nn, 0), img_offset(nn, 1), img_offset(nn, 2), net.bias[l][nn] * bias_mul);
    }
    if (!fwd) small[oGain + 1 + (r - 5)] = (float)(gain * slack);  // stages 8..12
  }
  // ---- stage 0: layer 0, K = 32: [W0_hi (5) | W0_hi (5) | W0_lo (5) | bias split (3) | 0] per output unit ----
  for (int j = 0; j < kHid; ++j) {
    for (int k = 0; k < 5; ++k) {
      uint16_t hb, lb;
      split_host(net.weights[0][k * kHid + j] * s, &hb, &lb);
      put16(res + rB0, img_offset(j, k), hb);
      put16(res + rB0, img_offset(j, 5 + k), hb);
      put16(res + rB0, img_offset(j, 10 + k), lb);
    }
    put_bias(res + rB0, img_offset(j, 15), img_offset(j, 16), img_offset(j, 17), net.bias[0][j] * bias_mul);
  }
  // ---- stage 6: layer 6, N = 16 (4 used), K = 128, + bias tile; stage 7: its transpose, K = 16 -----------------
  double gain7 = 0.0;
  for (int j = 0; j < kHid; ++j) {
    double l1 = 0.0;
    for (int c = 0; c < 4; ++c) {
      uint16_t hb, lb;
      split_host(net.weights[6][j * 4 + c] * s, &hb, &lb);
      put16(res + rB6, img_offset16(c, j), hb);
      put16(res + rB6 + 8 * kKStep16, img_offset16(c, j), lb);
      put16(res + rB7, img_offset(j, c), hb);
      put16(res + rB7, img_offset(j, 4 + c), hb);
      put16(res + rB7, img_offset(j, 8 + c), lb);
      l1 += std::fabs((double)net.weights[6][j * 4 + c] * s);
    }
    gain7 = std::fmax(gain7, l1);
  }
  small[oGain + 0] = (float)(gain7 * slack);  // stage 7
  for (int c = 0; c < 4; ++c)
    put_bias(res + rB6Bias, img_offset16(c, 0), img_offset16(c, 1), img_offset16(c, 2), net.bias[6][c] * bias_mul);
  // ---- stage 13: layer 0 transposed, N = 16 (5 used), K = 128 ---------------------------------------------
  for (int k = 0; k < 5; ++k)
    for (int j = 0; j < kHid; ++j) {
      uint16_t hb, lb;
      split_host(net.weights[0][k * kHid + j] * s, &hb, &lb);
      put16(res + rB13, img_offset16(k, j), hb);
      put16(res + rB13 + 8 * kKStep16, img_offset16(k, j), lb);
    }
  small[oBiasA] = ca;
  for (int c = 0; c < 4; ++c) {
    small[oOutLoc + c] = (float)net.out_loc[c];
    small[oOutScale + c] = (float)net.out_scale[c];
  }
  for (int k = 0; k < 5; ++k) {
    small[oInLoc + k] = (float)net.in_loc[k];
    small[oInScale + k] = (float)net.in_scale[k];
    small[oInInv + k] = (float)(1.0 / net.in_scale[k]);
  }
}

template <int BC, int MODE>
static int launch_tc_t(const Launch &L, cudaStream_t st, int *grid_out, int *smem_out) {
  const int smem = (int)sizeof(tc::Smem) + 1024;
  auto kern = tc::potential_tc_kernel<BC, MODE>;
  if (BC == 1 && MODE == tc::M_POTENTIAL && L.tc.trace) kern = tc::potential_tc_kernel<BC == 1 ? 1 : 0, tc::M_POTENTIAL, BC == 1 && MODE == tc::M_POTENTIAL>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return -(int)e;
  const long long n_tiles = (L.io.n + tc::kTile - 1) / tc::kTile;
  const int ctas = L.grid_limit > 0 ? std::min(L.grid_limit, L.sm_count) : L.sm_count;
  const int grid = (int)std::min<long long>(n_tiles, ctas);
  kern<<<grid, tc::kThreads, smem, st>>>(L);
  e = cudaGetLastError();
  if (e != cudaSuccess) return -(int)e;
  if (grid_out) *grid_out = grid;
  if (smem_out) *smem_out = smem;
  return 1;
}

int launch_potential_tc(const Launch &L, cudaStream_t st, int *grid_out, int *smem_out) {
  if (L.model.bands > 0 && L.model.bc_mode == 1) return launch_tc_t<1, tc::M_POTENTIAL>(L, st, grid_out, smem_out);
  return launch_tc_t<0, tc::M_POTENTIAL>(L, st, grid_out, smem_out);
}

int launch_predict_tc(const Launch &L, cudaStream_t st) {
  if (L.model.bands > 0 && L.model.bc_mode == 1) return launch_tc_t<1, tc::M_PREDICT>(L, st, nullptr, nullptr);
  return launch_tc_t<0, tc::M_PREDICT>(L, st, nullptr, nullptr);
}

// the stellar emulator on its own on the tensor-core engine: L.io.n rows, L.net_* pointers set
int launch_net_tc(const Launch &L, bool vjp, cudaStream_t st) {
  return vjp ? launch_tc_t<0, tc::M_NET_VJP>(L, st, nullptr, nullptr) : launch_tc_t<0, tc::M_NET_FWD>(L, st, nullptr, nullptr);
}

}  // namespace csi
