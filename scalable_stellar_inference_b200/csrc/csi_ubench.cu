// csi_ubench.cu -- micro-benchmarks of the tcgen05 building blocks on one SM (cycle counts via clock64):
//   MMA throughput for the shapes the engine uses (A from TMEM, B from smem), dependent vs independent
//   accumulators, commit -> mbarrier wake-up latency, tcgen05.ld / tcgen05.st round trips.
// usage: csi_ubench            (prints one line per measurement)
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdlib.h>

#include "csi_ptx.cuh"

using namespace csi;

struct __align__(128) USmem {
  uint8_t b_img[65536];
  uint64_t bar, bar2, bar3, bar4, bar5;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(160) ubench_kernel(long long *out) {
  extern __shared__ unsigned char raw[];
  USmem &S = *reinterpret_cast<USmem *>(raw + ((1024u - (ptx::smem_u32(raw) & 1023u)) & 1023u));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 65536 / 16; i += blockDim.x) reinterpret_cast<uint4 *>(S.b_img)[i] = make_uint4(0, 0, 0, 0);
  ptx::fence_proxy_async_smem();
  if (tid == 0) { ptx::mbar_init(&S.bar, 1); ptx::mbar_init(&S.bar2, 1); ptx::fence_barrier_init(); }
  if (warp == 4) ptx::tmem_alloc<512>(&S.tmem_base);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = S.tmem_base;
  const uint32_t b_s = ptx::smem_u32(S.b_img);
  uint32_t par = 0;
  // ---- MMA throughput: n_mma MMAs of (N, one or two accumulators), issued by one thread -------------------
  const int shapesN[3] = {128, 64, 16};
  for (int sh = 0; sh < 3; ++sh)
    for (int two = 0; two < 2; ++two) {
      const int N = shapesN[sh];
      const uint32_t idesc = ptx::make_idesc_f16(128, N);
      const int n_mma = 200;
      __syncthreads();
      if (tid == 128) {
        long long t0 = clock64();
        for (int j = 0; j < n_mma; ++j) {
          const uint32_t d = tmem + ((two && (j & 1)) ? 256u : 0u);
          ptx::mma_f16_ts(d, tmem + 128u + 8u * (j & 7), ptx::make_smem_desc(b_s + (j & 7) * 4096, 128, 256), idesc, j > 1);
        }
        long long t1 = clock64();
        ptx::mma_commit(&S.bar);
        ptx::mbar_wait(&S.bar, par);
        long long t2 = clock64();
        out[(sh * 2 + two) * 2] = t1 - t0;
        out[(sh * 2 + two) * 2 + 1] = t2 - t0;
      }
      par ^= 1;
      __syncthreads();
    }
  // ---- SS-mode (A from smem) N = 128 ------------------------------------------------------------------
  {
    const uint32_t idesc = ptx::make_idesc_f16(128, 128);
    __syncthreads();
    if (tid == 128) {
      long long t0 = clock64();
      for (int j = 0; j < 200; ++j)
        ptx::mma_f16_ss(tmem, ptx::make_smem_desc(b_s + 32768 + (j & 7) * 4096, 128, 256),
                        ptx::make_smem_desc(b_s + (j & 7) * 4096, 128, 256), idesc, j > 0);
      ptx::mma_commit(&S.bar);
      ptx::mbar_wait(&S.bar, par);
      out[12] = clock64() - t0;
    }
    par ^= 1;
    __syncthreads();
  }
  // ---- hand-off latency: 1 MMA + commit -> waiter on another warp sees it ---------------------------------
  {
    const uint32_t idesc = ptx::make_idesc_f16(128, 128);
    __shared__ long long t_issue, t_seen;
    __syncthreads();
    if (tid == 128) {
      t_issue = clock64();
      ptx::mma_f16_ts(tmem, tmem + 128u, ptx::make_smem_desc(b_s, 128, 256), idesc, 0);
      ptx::mma_commit(&S.bar2);
    }
    if (warp == 0) {
      ptx::mbar_wait(&S.bar2, 0);
      if (lane == 0) t_seen = clock64();
    }
    __syncthreads();
    if (tid == 0) out[13] = t_seen - t_issue;
  }
  // ---- tcgen05.ld x32 + wait (dependent use), idle tensor pipe vs MMAs streaming from another warp -------------
  for (int busy = 0; busy < 2; ++busy) {
    __syncthreads();
    if (tid == 128 && busy) {
      const uint32_t idesc = ptx::make_idesc_f16(128, 128);
      for (int j = 0; j < 120; ++j)
        ptx::mma_f16_ts(tmem + 256u, tmem + 384u + 8u * (j & 7), ptx::make_smem_desc(b_s + (j & 7) * 4096, 128, 256), idesc, j > 0);
      ptx::mma_commit(&S.bar);
    }
    if (warp < 4) {
      const uint32_t lane_base = ((uint32_t)(warp * 32)) << 16;
      uint32_t r[32], acc = 0;
      if (busy) __nanosleep(1000);
      long long t0 = clock64();
      for (int it = 0; it < 8; ++it) {
        ptx::tmem_ld32(tmem + lane_base + 32 * (it & 3), r);
        ptx::tmem_wait_ld_dep(r);
#pragma unroll
        for (int q = 0; q < 32; ++q) acc += r[q];
        asm volatile("" ::"r"(acc) : "memory");
      }
      long long t1 = clock64();
      uint32_t h[16];
#pragma unroll
      for (int q = 0; q < 16; ++q) h[q] = acc + q;
      for (int it = 0; it < 8; ++it) {
        ptx::tmem_st16(tmem + lane_base + 128 + 16 * (it & 3), h);
        ptx::tmem_st16(tmem + lane_base + 192 + 16 * (it & 3), h);
        ptx::tmem_wait_st();
      }
      long long t2 = clock64();
      if (lane == 0 && warp == 0) {
        out[14 + busy * 3] = (t1 - t0) / 8;
        out[15 + busy * 3] = (t2 - t1) / 8;
        out[16 + busy * 3] = acc;
      }
    }
    if (tid == 128 && busy) { ptx::mbar_wait(&S.bar, par); par ^= 1; }
    __syncthreads();
  }
  // ---- issue-pattern variants: 32 batches of 6 MMAs (immediate operands), whole warp 4 in the loop -----------
  //  0: elected lane issues, __syncwarp            1: + vote(mbarrier.test_wait) on a completed barrier before each batch
  //  2: variant 1 + clock64/global store per batch  3: variant 0 + tcgen05.commit (unwaited barrier) after each batch
  //  4: batches of 24 MMAs x 8 (same total)         5: variant 1 with a failing poll of a second barrier as well
  //  6: single thread (tid 128 only), 6-MMA batches, no warp-level ops
  {
    const uint32_t idesc = ptx::make_idesc_f16(128, 128);
    const uint32_t lo0 = ((b_s & 0x3FFFFu) >> 4) | ((128u >> 4) << 16), dhi = (256u >> 4) | (1u << 14);
    if (tid == 0) { ptx::mbar_init(&S.bar3, 1); ptx::mbar_init(&S.bar4, 1); ptx::mbar_init(&S.bar5, 1); ptx::fence_barrier_init(); }
    __syncthreads();
    if (tid == 0) ptx::mbar_arrive(&S.bar3);  // bar3: phase 0 complete -> test_wait(parity 0) succeeds; bar4 never completes
    __syncthreads();
    for (int v = 0; v < 7; ++v) {
      __syncthreads();
      if (warp == 4 && (v != 6 || lane == 0)) {
        long long t0 = clock64();
        const int nb = (v == 4) ? 8 : 32;
        for (int bt = 0; bt < nb; ++bt) {
          bool ok = true;
          if (v == 1 || v == 2 || v == 5) ok = __all_sync(0xffffffffu, ptx::mbar_test_wait(&S.bar3, 0u));
          if (v == 5) ok = ok && !__all_sync(0xffffffffu, ptx::mbar_test_wait(&S.bar4, 0u));
          if (!ok) continue;
          if (v == 6 || ptx::elect_one()) {
            if (v == 4) {
#pragma unroll
              for (int j = 0; j < 24; ++j)
                ptx::mma_f16_ts(tmem, tmem + 128u + 8u * (j & 7), ptx::desc_from_words(lo0 + 256u * (j & 7), dhi), idesc, 1u);
            } else {
#pragma unroll
              for (int j = 0; j < 6; ++j)
                ptx::mma_f16_ts(tmem, tmem + 128u + 8u * j, ptx::desc_from_words(lo0 + 256u * j, dhi), idesc, 1u);
            }
            if (v == 3) ptx::mma_commit(&S.bar5);
            if (v == 2) { out[40 + (bt & 3)] = clock64(); }
          }
          if (v != 6) __syncwarp();
        }
        if (lane == 0) {
          ptx::mma_commit(&S.bar);
          ptx::mbar_wait(&S.bar, par);
          out[30 + v] = clock64() - t0;
        }
      }
      par ^= 1;
      __syncthreads();
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 4) ptx::tmem_dealloc<512>(tmem);
}

int main() {
  long long *d, h[48] = {0};
  cudaMalloc(&d, sizeof(h));
  cudaMemset(d, 0, sizeof(h));
  const int smem = (int)sizeof(USmem) + 1024;
  cudaFuncSetAttribute(ubench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int rep = 0; rep < 2; ++rep) {
    ubench_kernel<<<1, 160, smem>>>(d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
  }
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  const char *names[3] = {"N=128", "N=64", "N=16"};
  for (int sh = 0; sh < 3; ++sh)
    for (int two = 0; two < 2; ++two)
      printf("mma TS M=128 %s K=16 x200, %s accumulator(s): issue %lld cyc total, complete %lld cyc total = %.1f cyc/MMA\n",
             names[sh], two ? "two alternating" : "one", h[(sh * 2 + two) * 2], h[(sh * 2 + two) * 2 + 1],
             h[(sh * 2 + two) * 2 + 1] / 200.0);
  printf("mma SS M=128 N=128 x200: %.1f cyc/MMA\n", h[12] / 200.0);
  printf("1 MMA + commit -> waiting warp resumes: %lld cyc\n", h[13]);
  for (int busy = 0; busy < 2; ++busy)
    printf("4 warps, tensor pipe %s: tcgen05.ld x32 + wait + use %lld cyc, 2 x tcgen05.st x16 + wait %lld cyc\n",
           busy ? "busy (MMA N=128 stream)" : "idle", h[14 + busy * 3], h[15 + busy * 3]);
  for (int v = 0; v < 7; ++v) printf("issue-pattern variant %d: %.1f cyc/MMA\n", v, h[30 + v] / 192.0);
  return 0;
}
