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

__global__ void __launch_bounds__(128, 1) probe(int N, int mode, int iters, int sbo_bytes, long long* out, int M = 128,
                                                int sbo_b_bytes = 1024, int b_shift_rows = 0, int a_tiles = 4, int hammer = 0) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar2[2];
  __shared__ uint32_t tmem_slot;
  __shared__ volatile int stop_flag;
  if (threadIdx.x == 0) stop_flag = 0;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  // A: up to 24 KB region (16 groups of 8 rows, group pitch sbo_bytes <= 1536); B: 256 rows x 128 B after it
  const uint32_t a_addr = base, b_addr = base + 32768 + 128u * (uint32_t)b_shift_rows;
  for (uint32_t i = threadIdx.x; i < (32768 + 49152) / 4; i += blockDim.x) {
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
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (((uint32_t)M >> 4) << 24);
    const uint64_t hi_a = (uint64_t)((uint32_t)(sbo_bytes >> 4) | (1u << 14) | (2u << 29)) << 32;
    const uint64_t hi_b = (uint64_t)((uint32_t)(sbo_b_bytes >> 4) | (1u << 14) | (2u << 29)) << 32;
    const uint32_t lo_a = (a_addr >> 4) | (1u << 16), lo_b = (b_addr >> 4) | (1u << 16);
    const uint32_t a_tmem = tmem + 448 - ((mode & 4) ? 64 : 0);  // TS mode: A operand columns (128 lanes x 8 columns per K = 16), past the accumulators
    const bool ts = (mode & 1) != 0, alt = (mode & 2) != 0, cp = (mode & 4) != 0;
    const long long t0 = clock64();
    uint32_t a_cur = 0;
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
              "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %4, p;\n\t}" ::"r"(d), "r"(a_tiles > 4 ? tmem + 256u + 8u * a_cur : a_tmem + 8 * k + ((cp && (it & 1)) ? 32u : 0u)), "l"(db),
              "r"(1u), "r"(idesc)
              : "memory");
        } else {
          const uint64_t da = hi_a | (uint64_t)(lo_a + 2 * k);
          asm volatile(
              "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t"
              "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %4, p;\n\t}" ::"r"(d), "l"(da), "l"(db), "r"(1u), "r"(idesc)
              : "memory");
        }
        if (++a_cur == (uint32_t)a_tiles) a_cur = 0;
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    __syncwarp();
    uint32_t ok = 0;
    const long long t_wait = clock64();
    while (!ok) {
      if (clock64() - t_wait > 4000000000ll) __trap();  // ~2 s: a protocol error must not hang the GPU
      asm volatile(
          "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(ok)
          : "r"(smem_u32(&bar)), "r"(0u)
          : "memory");
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = (mode & 8) ? (t1 - t0) / 2 : (t1 - t0);  // dual issue: per MMA of the pair
    if (threadIdx.x == 0) stop_flag = 1;
  } else if (hammer) {
    // r2t: what slows queued MMAs inside the conv kernel?  Warps 1-3 load one resource while warp 0 issues:
    // 1 = ld.shared.v4, 2 = st.shared.v4 (A region: the TS MMAs read only B), 3 = tcgen05.ld of the accumulator columns,
    // 4 = ALU only (issue slots)
    uint32_t acc = threadIdx.x;
    const uint32_t my = a_addr + (threadIdx.x - 32) * 16;
    while (!stop_flag) {
#pragma unroll 1
      for (int r = 0; r < 16; ++r) {
        if (hammer == 1) {
          uint32_t x0, x1, x2, x3;
          asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x0), "=r"(x1), "=r"(x2), "=r"(x3) : "r"(my + (r & 7) * 1536));
          acc += x0 ^ x1 ^ x2 ^ x3;
        } else if (hammer == 2) {
          asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(my + (r & 7) * 1536), "r"(acc) : "memory");
        } else if (hammer == 3) {
          uint32_t v[8];
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                       : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                       : "r"(tmem + ((uint32_t)(wid * 32) << 16) + (uint32_t)(r * 8)));
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          acc += v[0] ^ v[7];
        } else {
#pragma unroll
          for (int u = 0; u < 16; ++u) acc = acc * 3u + 1u;
        }
      }
    }
    if (acc == 0x12345678u) out[0] = acc;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
  }
}

// ---- cta_group::2: a CTA pair (cluster of 2) issues ONE MMA of M = 256 (128 rows of A and 128 lanes of D per CTA), B's N
// rows split N/2 per CTA.  Only the leader (cluster rank 0) issues; the commit is multicast to both CTAs' barriers.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) probe2(int N, int mode, int iters, long long* out) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const uint32_t a_addr = base, b_addr = base + 32768;
  for (uint32_t i = threadIdx.x; i < (32768 + 32768) / 4; i += blockDim.x) {
    uint32_t h = (i * 2654435761u) ^ (blockIdx.x * 40503u);
    uint32_t v = (0x3f00u | (h & 0x80ffu)) | ((0x3f00u | ((h >> 16) & 0x80ffu)) << 16);
    reinterpret_cast<uint32_t*>(raw + (base - smem_u32(raw)))[i] = v;
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x < 32) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((256u >> 4) << 24);
    const uint64_t hi = (uint64_t)(64u | (1u << 14) | (2u << 29)) << 32;
    const uint32_t lo_a = (a_addr >> 4) | (1u << 16), lo_b = (b_addr >> 4) | (1u << 16);
    const uint32_t a_tmem = tmem + 448;
    const bool ts = (mode & 1) != 0;
    const long long t0 = clock64();
    if (rank == 0 && elect_one_sync()) {
#pragma unroll 1
      for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t db = hi | (uint64_t)(lo_b + 2 * k);
          if (ts) {
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t"
                "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %4, p;\n\t}" ::"r"(tmem), "r"(a_tmem + 8 * k), "l"(db), "r"(1u),
                "r"(idesc)
                : "memory");
          } else {
            const uint64_t da = hi | (uint64_t)(lo_a + 2 * k);
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t"
                "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %4, p;\n\t}" ::"r"(tmem), "l"(da), "l"(db), "r"(1u), "r"(idesc)
                : "memory");
          }
        }
      }
      asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                       smem_u32(&bar)),
                   "h"((uint16_t)3)
                   : "memory");
    }
    __syncwarp();
    uint32_t ok = 0;
    const long long t_wait = clock64();
    while (!ok) {
      if (clock64() - t_wait > 4000000000ll) __trap();  // ~2 s: a protocol error must not hang the GPU
      asm volatile(
          "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(ok)
          : "r"(smem_u32(&bar)), "r"(0u)
          : "memory");
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0 && rank == 0) out[blockIdx.x / 2] = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
  }
}

int main() {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  long long* d_out;
  cudaMalloc(&d_out, sizeof(long long) * sms);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 90 * 1024);
  cudaFuncSetAttribute(probe2, cudaFuncAttributeMaxDynamicSharedMemorySize, 90 * 1024);
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
        for (int rep = 0; rep < 2; ++rep) probe<<<sms, 128, 90 * 1024>>>(N, mode, iters, sbo, d_out);
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
  // ---- round 2: the questions behind the conv re-mapping ----------------------------------------
  auto run1 = [&](const char* label, int N, int mode, int M, int sbo_b) {
    for (int rep = 0; rep < 2; ++rep) probe<<<sms, 128, 90 * 1024>>>(N, mode, iters, 1024, d_out, M, sbo_b);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s N=%d: %s\n", label, N, cudaGetErrorString(e)); exit(1); }
    std::vector<long long> h(sms);
    cudaMemcpy(h.data(), d_out, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    std::sort(h.begin(), h.end());
    const double per = (double)h[sms / 2] / (iters * 4.0), ideal = (double)M * N / 256.0;
    printf("%-28s M=%3d N=%3d  %7.1f clk/MMA  (ideal %5.1f, %5.1f %%)\n", label, M, N, per, ideal, 100.0 * ideal / per);
  };
  for (int N : {64, 128, 256}) run1("M=64 SS", N, 0, 64, 1024);
  for (int N : {64, 128, 256}) run1("M=64 TS", N, 1, 64, 1024);
  for (int N : {80, 144, 160, 176, 208, 240}) run1("TS odd N", N, 1, 128, 1024);
  for (int N : {128, 144, 256}) run1("TS, B groups 1280 B apart", N, 1, 128, 1280);
  for (int N : {128, 144, 256}) run1("SS, B groups 1280 B apart", N, 0, 128, 1280);
  // r2r: is the TS rate kept when the B start is shifted by whole rows inside a swizzle atom (the conv's tap shifts) and
  // when every MMA reads a different A tile (24 tiles = the conv's filter)?
  auto run2 = [&](const char* label, int N, int shift, int tiles) {
    for (int rep = 0; rep < 2; ++rep) probe<<<sms, 128, 90 * 1024>>>(N, 1, iters, 1024, d_out, 128, 1280, shift, tiles);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", label, cudaGetErrorString(e)); exit(1); }
    std::vector<long long> h(sms);
    cudaMemcpy(h.data(), d_out, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    std::sort(h.begin(), h.end());
    printf("%-28s N=%3d B shift %2d rows, %2d A tiles  %7.1f clk/MMA\n", label, N, shift, tiles, (double)h[sms / 2] / (iters * 4.0));
  };
  for (int shift : {0, 1, 2, 4, 8, 20, 21, 22}) run2("TS shifted B", 128, shift, 4);
  for (int hm : {0, 1, 2, 3, 4}) {
    for (int rep = 0; rep < 2; ++rep) probe<<<sms, 128, 90 * 1024>>>(128, 1, iters, 1024, d_out, 128, 1280, 1, 24, hm);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("hammer %d: %s\n", hm, cudaGetErrorString(e)); exit(1); }
    std::vector<long long> h(sms);
    cudaMemcpy(h.data(), d_out, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    std::sort(h.begin(), h.end());
    printf("TS N=128 with 3 warps hammering %-22s %7.1f clk/MMA\n",
           hm == 0 ? "nothing" : hm == 1 ? "ld.shared.v4" : hm == 2 ? "st.shared.v4" : hm == 3 ? "tcgen05.ld (acc cols)" : "ALU",
           (double)h[sms / 2] / (iters * 4.0));
  }
  for (int tiles : {5, 6, 8, 12, 24, 32}) run2("TS many A tiles", 128, 1, tiles);
  for (int tiles : {8, 24}) run2("TS many A tiles", 64, 1, tiles);
  for (int mode = 0; mode < 2; ++mode)
    for (int N : {64, 128, 256}) {
      for (int rep = 0; rep < 2; ++rep) probe2<<<sms / 2 * 2, 128, 90 * 1024>>>(N, mode, iters, d_out);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("cta_group::2 mode %d N=%d: %s\n", mode, N, cudaGetErrorString(e)); return 1; }
      const int pairs = sms / 2;
      std::vector<long long> h(pairs);
      cudaMemcpy(h.data(), d_out, sizeof(long long) * pairs, cudaMemcpyDeviceToHost);
      std::sort(h.begin(), h.end());
      const double per = (double)h[pairs / 2] / (iters * 4.0);
      printf("cta_group::2 %s             M=256 N=%3d  %7.1f clk/MMA per pair (ideal %5.1f per SM, %5.1f %%)\n", mode ? "TS" : "SS", N, per,
             N / 2.0, 100.0 * (N / 2.0) / per);
    }
  return 0;
}
