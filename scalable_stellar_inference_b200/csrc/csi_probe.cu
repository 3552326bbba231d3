// csi_probe.cu -- standalone hardware probe for the tcgen05 building blocks libcsi relies on.
//
//   csi_probe <variant>
//     0: A from TMEM (tcgen05.st, packed fp16x2), B from smem, descriptor LBO=K-dir(128) SBO=N-dir(256)
//     1: same, descriptor fields swapped
//     2: A and B from smem, descriptor as in 0
//     3: A and B from smem, descriptor as in 1
//     4: TMEM st/ld round trip only (lane/column mapping of the 32x32b shape)
//
// Computes D[128x128] = A[128x128] * B[128x128]^T (fp16 inputs, fp32 accumulate) as 8 K=16 MMAs and compares
// with a host reference.  Exit code 0 = match.  Each variant is a separate process so that a faulting
// variant cannot poison the others.
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdlib.h>

#include <cmath>
#include <cstring>
#include <vector>

#include "csi_ptx.cuh"

using namespace csi;

constexpr uint32_t kKStep = 4096;

static uint32_t img_off(int n, int k) {
  return (uint32_t)(k >> 4) * kKStep + (uint32_t)(n >> 3) * 256 + (uint32_t)((k >> 3) & 1) * 128 + (uint32_t)(n & 7) * 16 +
         (uint32_t)(k & 7) * 2;
}

struct __align__(128) PSmem {
  uint8_t a_img[32768];
  uint8_t b_img[32768];
  uint64_t bar;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(128) probe_kernel(const uint8_t *a_img, const uint8_t *b_img, const __half *a_rows,
                                                    float *d_out, uint32_t *rt_out, int variant, uint32_t lbo,
                                                    uint32_t sbo) {
  extern __shared__ unsigned char raw[];
  PSmem &S = *reinterpret_cast<PSmem *>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 32768 / 16; i += 128) {
    reinterpret_cast<uint4 *>(S.a_img)[i] = reinterpret_cast<const uint4 *>(a_img)[i];
    reinterpret_cast<uint4 *>(S.b_img)[i] = reinterpret_cast<const uint4 *>(b_img)[i];
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (tid == 0) {
    ptx::mbar_init(&S.bar, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 0) ptx::tmem_alloc<256>(&S.tmem_base);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = S.tmem_base;
  const uint32_t lane_base = ((uint32_t)(warp * 32)) << 16;
  if (tid == 0) rt_out[128 * 16] = tmem;

  if (variant == 4) {
    uint32_t w[16], r[32];
    for (int j = 0; j < 16; ++j) w[j] = (uint32_t)(tid * 1000 + j);
    ptx::tmem_st16(tmem + lane_base + 5, w);
    ptx::tmem_wait_st();
    ptx::tmem_ld32(tmem + lane_base, r);
    ptx::tmem_wait_ld();
    for (int j = 0; j < 16; ++j) rt_out[tid * 16 + j] = r[5 + j];
  } else {
    if (variant < 2) {
      // A operand: row tid, 128 fp16 = 64 packed columns at TMEM columns [128, 192)
      const uint32_t *row = reinterpret_cast<const uint32_t *>(a_rows + (size_t)tid * 128);
      for (int c = 0; c < 4; ++c) {
        uint32_t w[16];
        for (int j = 0; j < 16; ++j) w[j] = row[16 * c + j];
        ptx::tmem_st16(tmem + lane_base + 128 + 16 * c, w);
      }
      ptx::tmem_wait_st();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    if (tid == 0) {
      const uint32_t idesc = ptx::make_idesc_f16(128, 128);
      const uint32_t a_s = ptx::smem_u32(S.a_img), b_s = ptx::smem_u32(S.b_img);
      for (int j = 0; j < 8; ++j) {
        const uint64_t bd = ptx::make_smem_desc(b_s + j * kKStep, lbo, sbo);
        if (variant < 2) {
          ptx::mma_f16_ts(tmem, tmem + 128 + 8 * j, bd, idesc, j > 0);
        } else {
          const uint64_t ad = ptx::make_smem_desc(a_s + j * kKStep, lbo, sbo);
          ptx::mma_f16_ss(tmem, ad, bd, idesc, j > 0);
        }
      }
      ptx::mma_commit(&S.bar);
    }
    ptx::mbar_wait(&S.bar, 0);
    ptx::tc_fence_after();
    for (int c = 0; c < 4; ++c) {
      uint32_t r[32];
      ptx::tmem_ld32(tmem + lane_base + 32 * c, r);
      ptx::tmem_wait_ld();
      for (int j = 0; j < 32; ++j) d_out[(size_t)tid * 128 + 32 * c + j] = __uint_as_float(r[j]);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc<256>(tmem);
}

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e = (x);                                                           \
    if (e != cudaSuccess) {                                                        \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      return 2;                                                                    \
    }                                                                              \
  } while (0)

int main(int argc, char **argv) {
  const int variant = argc > 1 ? atoi(argv[1]) : 0;
  std::vector<__half> A(128 * 128), B(128 * 128);
  std::vector<float> Af(128 * 128), Bf(128 * 128);
  srand(1234);
  for (int i = 0; i < 128 * 128; ++i) {
    float a = (float)((rand() % 2001) - 1000) / 512.0f, b = (float)((rand() % 2001) - 1000) / 1024.0f;
    A[i] = __float2half_rn(a); B[i] = __float2half_rn(b);
    Af[i] = __half2float(A[i]); Bf[i] = __half2float(B[i]);
  }
  std::vector<uint8_t> a_img(32768), b_img(32768);
  for (int n = 0; n < 128; ++n)
    for (int k = 0; k < 128; ++k) {
      memcpy(&a_img[img_off(n, k)], &A[n * 128 + k], 2);
      memcpy(&b_img[img_off(n, k)], &B[n * 128 + k], 2);
    }
  uint8_t *d_a, *d_b; __half *d_rows; float *d_d; uint32_t *d_rt;
  CK(cudaMalloc(&d_a, 32768)); CK(cudaMalloc(&d_b, 32768)); CK(cudaMalloc(&d_rows, 128 * 128 * 2));
  CK(cudaMalloc(&d_d, 128 * 128 * 4)); CK(cudaMalloc(&d_rt, (128 * 16 + 1) * 4));
  CK(cudaMemcpy(d_a, a_img.data(), 32768, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_b, b_img.data(), 32768, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_rows, A.data(), 128 * 128 * 2, cudaMemcpyHostToDevice));
  CK(cudaMemset(d_d, 0, 128 * 128 * 4));
  CK(cudaMemset(d_rt, 0, (128 * 16 + 1) * 4));
  const uint32_t lbo = (variant & 1) ? 256 : 128, sbo = (variant & 1) ? 128 : 256;
  const int smem = (int)sizeof(PSmem) + 1024;
  CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  probe_kernel<<<1, 128, smem>>>(d_a, d_b, d_rows, d_d, d_rt, variant, lbo, sbo);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  std::vector<uint32_t> rt(128 * 16 + 1);
  CK(cudaMemcpy(rt.data(), d_rt, rt.size() * 4, cudaMemcpyDeviceToHost));
  printf("variant %d: tmem base 0x%08x\n", variant, rt[128 * 16]);
  if (variant == 4) {
    int bad = 0;
    for (int t = 0; t < 128; ++t)
      for (int j = 0; j < 16; ++j) bad += rt[t * 16 + j] != (uint32_t)(t * 1000 + j);
    printf("variant 4: tmem st/ld round trip mismatches = %d (row0: %u %u %u, row 77: %u %u)\n", bad, rt[0], rt[1], rt[2],
           rt[77 * 16], rt[77 * 16 + 1]);
    return bad ? 1 : 0;
  }
  std::vector<float> D(128 * 128);
  CK(cudaMemcpy(D.data(), d_d, D.size() * 4, cudaMemcpyDeviceToHost));
  double maxerr = 0, maxref = 0;
  for (int mrow = 0; mrow < 128; ++mrow)
    for (int n = 0; n < 128; ++n) {
      double s = 0;
      for (int k = 0; k < 128; ++k) s += (double)Af[mrow * 128 + k] * (double)Bf[n * 128 + k];
      maxerr = std::fmax(maxerr, std::fabs(s - D[mrow * 128 + n]));
      maxref = std::fmax(maxref, std::fabs(s));
    }
  printf("variant %d: max |D - ref| = %.6g (max |ref| = %.4g)  D[0][0..3] = %g %g %g %g\n", variant, maxerr, maxref, D[0],
         D[1], D[2], D[3]);
  return maxerr < 1e-2 * maxref ? 0 : 1;
}
