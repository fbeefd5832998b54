// Microbenchmark: cycles per tcgen05.mma (kind::f16, M = 128, K = 16) as a function of N, operand source of A
// (shared memory descriptor vs TMEM) and accumulator dependence.  One CTA per SM, one issuing thread.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/umma_probe scripts/umma_probe.cu
// Output: one line per configuration: N, mode, clk per MMA (median over CTAs), % of the 128xNx16 ideal (N/2 clk).
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ bool elect_one_sync() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred P1;\n\telect.sync _|P1, 0xffffffff;\n\tselp.u32 %0, 1, 0, P1;\n\t}" : "=r"(pred));
  return pred != 0;
}

__global__ void __launch_bounds__(128, 1) probe(int N, int mode, int iters, int sbo_bytes, long long* out) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar2[2];
  __shared__ uint32_t tmem_slot;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  // A: up to 24 KB region (16 groups of 8 rows, group pitch sbo_bytes <= 1536); B: 256 rows x 128 B after it
  const uint32_t a_addr = base, b_addr = base + 32768;
  for (uint32_t i = threadIdx.x; i < (32768 + 32768) / 4; i += blockDim.x) {
    uint32_t h = (i * 2654435761u) ^ (blockIdx.x * 40503u);
    // two bf16 values in [-1, 1): sign + exponent 0x3f00.. keeps them finite
    uint32_t v = (0x3f00u | (h & 0x80ffu)) | ((0x3f00u | ((h >> 16) & 0x80ffu)) << 16);
    reinterpret_cast<uint32_t*>(raw + (base - smem_u32(raw)))[i] = v;
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar2[0])) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar2[1])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  const int wid = threadIdx.x >> 5;
  if (wid == 0 || (wid == 1 && (mode & 8))) {
    uint64_t& bar = bar2[wid];
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t hi_a = (uint64_t)((uint32_t)(sbo_bytes >> 4) | (1u << 14) | (2u << 29)) << 32;
    const uint64_t hi_b = (uint64_t)(64u | (1u << 14) | (2u << 29)) << 32;
    const uint32_t lo_a = (a_addr >> 4) | (1u << 16), lo_b = (b_addr >> 4) | (1u << 16);
    const uint32_t a_tmem = tmem + 448 - ((mode & 4) ? 64 : 0);  // TS mode: A operand columns (128 lanes x 8 columns per K = 16), past the accumulators
    const bool ts = (mode & 1) != 0, alt = (mode & 2) != 0, cp = (mode & 4) != 0;
    const long long t0 = clock64();
    if (elect_one_sync()) {
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t d = tmem + (uint32_t)(wid * 224) + ((alt && (k & 1)) ? (uint32_t)N : 0u);
        const uint64_t db = hi_b | (uint64_t)(lo_b + 2 * k);
        if (cp) {  // refill the A slice this MMA reads: 128 rows x 32 B (one K = 16 step) from shared memory to TMEM
          const uint64_t dsrc = hi_a | (uint64_t)(lo_a + 2 * k);
          asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(a_tmem + 8 * k + ((it & 1) ? 32u : 0u)), "l"(dsrc) : "memory");
        }
        if (ts) {
          asm volatile(
              "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t"
              "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %4, p;\n\t}" ::"r"(d), "r"(a_tmem + 8 * k + ((cp && (it & 1)) ? 32u : 0u)), "l"(db),
              "r"(1u), "r"(idesc)
              : "memory");
        } else {
          const uint64_t da = hi_a | (uint64_t)(lo_a + 2 * k);
          asm volatile(
              "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t"
              "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %4, p;\n\t}" ::"r"(d), "l"(da), "l"(db), "r"(1u), "r"(idesc)
              : "memory");
        }
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    __syncwarp();
    uint32_t ok = 0;
    while (!ok) {
      asm volatile(
          "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(ok)
          : "r"(smem_u32(&bar)), "r"(0u)
          : "memory");
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = (mode & 8) ? (t1 - t0) / 2 : (t1 - t0);  // dual issue: per MMA of the pair
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
  }
}

int main() {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  long long* d_out;
  cudaMalloc(&d_out, sizeof(long long) * sms);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 70 * 1024);
  const int iters = 2000;
  const int Ns[] = {16, 32, 64, 96, 128, 192, 256};
  const char* names[] = {"SS", "TS", "SS alt-acc", "TS alt-acc", "-", "TS+cp", "-", "TS+cp alt", "SS 2warps", "TS 2warps"};
  for (int sbo : {1024, 1280}) {
    for (int mode = 0; mode < 10; ++mode) {
      if (mode == 4 || mode == 6 || (sbo == 1280 && mode != 0 && mode != 5)) continue;
      for (int N : Ns) {
        if ((mode & 2) && 2 * N > 384) continue;
        if ((mode & 4) && N > 192) continue;
        if ((mode & 8) && N > 192) continue;
        for (int rep = 0; rep < 2; ++rep) probe<<<sms, 128, 70 * 1024>>>(N, mode, iters, sbo, d_out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("N=%d mode=%d: %s\n", N, mode, cudaGetErrorString(e)); return 1; }
        std::vector<long long> h(sms);
        cudaMemcpy(h.data(), d_out, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
        std::sort(h.begin(), h.end());
        const double per = (double)h[sms / 2] / (iters * 4.0);
        printf("SBO=%4d  %-10s N=%3d  %7.1f clk/MMA  (ideal %5.1f, %5.1f %%)   min %.1f max %.1f\n", sbo, names[mode], N, per,
               N / 2.0, 100.0 * (N / 2.0) / per, (double)h[0] / (iters * 4.0), (double)h[sms - 1] / (iters * 4.0));
      }
    }
  }
  return 0;
}
