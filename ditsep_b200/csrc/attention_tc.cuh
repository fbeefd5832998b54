// Single-head self-attention on the 5th-generation tensor cores (tcgen05 / TMEM), bf16, C = 128 or 256 channels.
// o = softmax(q k^T / sqrt(C)) v over the S = H*W tokens of one image (AttnBlockpp core, layerspp.py:88-92).
// Included by attention.cu; the CUDA-core kernel there stays for fp32, C = 64 and tiny images.
//
// One CTA = 128 queries of one image (TMEM lane = query).  Keys are walked in tiles of kKeys (128 at C = 128, 64 at
// C = 256) TWICE: pass 0 only takes the row maxima of the logits, pass 1 recomputes the logits, exponentiates against
// the final maximum and accumulates P V -- no online rescaling of the output accumulator (the tensor work is 0.2 % of
// the step; the second Q K^T costs less than a correction pass over TMEM would).
//   S = Q K^T : A = Q tile (128 x C, K-major, TMA), B = K tile (kKeys x C, K-major, TMA)        -> TMEM cols [0, kKeys)
//   O += P V  : A = P (128 x kKeys bf16, K-major, written by the softmax threads with the 128-byte swizzle),
//               B = V tile AS IT LIES IN MEMORY (kKeys rows of C channels): the MN-major operand form -- N (channels)
//               contiguous, 64-channel atoms LBO apart, 8-key groups SBO = 1024 bytes apart         -> TMEM cols [256, 256 + C)
// Warps 0-7: softmax + epilogue (thread = query row of lane quadrant w & 3, column half w >> 2 of the logits and of the
// output; row maxima / sums of the halves meet in shared memory); warp 8: TMA producer and MMA issuer (one elected lane).
// Everything is single-buffered and handed over with mbarriers: the kernel is latency-bound by construction (10 key
// tiles at S = 1280) and runs in tens of microseconds where the CUDA-core kernel needs hundreds.
#pragma once

#include "tc_ptx.cuh"

namespace ditsep {

constexpr int kAtQ = 128;          // queries per CTA
constexpr int kAtThreads = 288;    // 8 softmax / epilogue warps (lane quadrant x column half) + 1 producer / MMA warp
constexpr int kAtOCol = 256;       // first TMEM column of the output accumulator

__device__ __forceinline__ float ex2_approx(float x) {  // 2^x, one MUFU; -inf -> 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int C, int kKeys>
__global__ void __launch_bounds__(kAtThreads, 1) attention_tc_kernel(const __grid_constant__ CUtensorMap mapQ,
                                                                     const __grid_constant__ CUtensorMap mapK,
                                                                     const __grid_constant__ CUtensorMap mapV,
                                                                     int S, bf16* __restrict__ o, int ld_o, float scale) {
  constexpr int kCh = C / 64;               // 64-channel chunks (one 128-byte swizzle row each)
  constexpr int kPc = kKeys / 64;           // 64-key chunks of P
  constexpr uint32_t kQBytes = kCh * kAtQ * 128, kKVBytes = kCh * kKeys * 128, kPBytes = kPc * kAtQ * 128;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_q, bar_kv[2], bar_s_full, bar_s_free, bar_p_ready, bar_pv_done;
  __shared__ uint32_t tmem_slot;
  __shared__ float part[2][kAtQ];  // row maxima, then row sums, of the two column halves

  griddep_launch_dependents();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // Q | K[2] | V[2] | P: the next tile's K (and V) land while the current one is multiplied and exponentiated
  const uint32_t q_s = base, k_s0 = q_s + kQBytes, v_s0 = k_s0 + 2 * kKVBytes, p_s = v_s0 + 2 * kKVBytes;
  uint8_t* const p_gen = smem_raw + (base - smem_u32(smem_raw)) + kQBytes + 4 * kKVBytes;
  const int b = blockIdx.y, q0 = blockIdx.x * kAtQ;
  const int nkt = (S + kKeys - 1) / kKeys;

  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar_q), 1);
    mbar_init(smem_u32(&bar_kv[0]), 1);
    mbar_init(smem_u32(&bar_kv[1]), 1);
    mbar_init(smem_u32(&bar_s_full), 1);
    mbar_init(smem_u32(&bar_s_free), 8);
    mbar_init(smem_u32(&bar_p_ready), 8);
    mbar_init(smem_u32(&bar_pv_done), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) tmem_alloc(smem_u32(&tmem_slot), 512u);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  griddep_wait();  // q, k, v are the predecessor's output

  if (warp == 8) {
    if (lane == 0) {
      // instruction descriptors: D = f32, A = B = bf16; M = 128; QK: N = kKeys, both K-major; PV: N = C, B MN-major
      const uint32_t idesc_qk = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kKeys >> 3) << 17) | ((uint32_t)(kAtQ >> 4) << 24);
      const uint32_t idesc_pv = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | ((uint32_t)(C >> 3) << 17) |
                                ((uint32_t)(kAtQ >> 4) << 24);
      constexpr uint32_t k_hi = 64u | (1u << 14) | (2u << 29);  // K-major SW128: 8-row groups 1024 bytes apart
      constexpr uint32_t lbo1 = 1u << 16;
      constexpr uint32_t v_lbo = (uint32_t)((kKeys * 128) >> 4) << 16;  // MN-major: 64-channel atoms one chunk apart
      mbar_expect_tx(smem_u32(&bar_q), kQBytes);
      for (int c = 0; c < kCh; ++c) tma_load_3d(q_s + (uint32_t)c * (kAtQ * 128), &mapQ, smem_u32(&bar_q), c * 64, q0, b);
      mbar_wait(smem_u32(&bar_q), 0);
      auto load_tile = [&](const uint32_t t) {  // tile t of the 2 * nkt (pass 0: K only; pass 1: K and V) -> buffer t & 1
        const bool p1 = t >= (uint32_t)nkt;
        const int kt = (int)(p1 ? t - (uint32_t)nkt : t);
        const uint32_t bar = smem_u32(&bar_kv[t & 1u]), off = (t & 1u) * kKVBytes;
        mbar_expect_tx(bar, p1 ? 2 * kKVBytes : kKVBytes);
        for (int c = 0; c < kCh; ++c) {
          tma_load_3d(k_s0 + off + (uint32_t)c * (kKeys * 128), &mapK, bar, c * 64, kt * kKeys, b);
          if (p1) tma_load_3d(v_s0 + off + (uint32_t)c * (kKeys * 128), &mapV, bar, c * 64, kt * kKeys, b);
        }
      };
      load_tile(0);
      uint32_t n = 0;  // tiles issued so far (both passes): the phase of every per-tile barrier
      for (int pass = 0; pass < 2; ++pass) {
        for (int kt = 0; kt < nkt; ++kt, ++n) {
          // the previous tile's consumers are done with S, K (pass 0: logits read) and with V, P (pass 1: P V complete):
          // its buffer may take the NEXT tile, and S may be overwritten
          if (n > 0) {
            if (pass == 0 || kt == 0) mbar_wait(smem_u32(&bar_s_free), (n - 1) & 1u);
            else mbar_wait(smem_u32(&bar_pv_done), (uint32_t)(kt - 1) & 1u);
          }
          if (n + 1 < 2u * (uint32_t)nkt) load_tile(n + 1);
          const uint32_t k_s = k_s0 + (n & 1u) * kKVBytes, v_s = v_s0 + (n & 1u) * kKVBytes;
          mbar_wait(smem_u32(&bar_kv[n & 1u]), (n >> 1) & 1u);
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < kCh; ++c) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const uint64_t da = ((uint64_t)k_hi << 32) | (uint64_t)((((q_s + (uint32_t)c * (kAtQ * 128)) >> 4) | lbo1) + 2 * kk);
              const uint64_t db = ((uint64_t)k_hi << 32) | (uint64_t)((((k_s + (uint32_t)c * (kKeys * 128)) >> 4) | lbo1) + 2 * kk);
              umma_bf16(tmem, da, db, idesc_qk, (c | kk) != 0 ? 1u : 0u);
            }
          }
          umma_commit(smem_u32(&bar_s_full));
          if (pass) {
            mbar_wait(smem_u32(&bar_p_ready), (uint32_t)kt & 1u);
            tc_fence_after();
#pragma unroll
            for (int kk = 0; kk < kKeys / 16; ++kk) {  // 16 keys per MMA: two 8-key groups of V, 32 bytes along P's rows
              const uint64_t da = ((uint64_t)k_hi << 32) |
                                  (uint64_t)((((p_s + (uint32_t)(kk >> 2) * (kAtQ * 128)) >> 4) | lbo1) + 2 * (kk & 3));
              const uint64_t db = ((uint64_t)k_hi << 32) | (uint64_t)(((v_s + (uint32_t)kk * 2048u) >> 4) | v_lbo);
              umma_bf16(tmem + kAtOCol, da, db, idesc_pv, (kt | kk) != 0 ? 1u : 0u);
            }
            umma_commit(smem_u32(&bar_pv_done));
          }
        }
      }
    }
  } else {
    // warp w: TMEM lane quadrant w & 3 (query rows), column half w >> 2 of every logits tile and of the output
    const int qd = warp & 3, hf = warp >> 2;
    const int r = qd * 32 + lane;  // query row = TMEM lane
    const uint32_t lane_addr = tmem + ((uint32_t)(qd * 32) << 16);
    constexpr int kHalf = kKeys / 2;
    const float sc = scale * 1.4426950408889634f;  // exp(x * scale) = exp2(x * sc)
    float m = -INFINITY, l = 0.f;
    uint32_t n = 0;
    for (int pass = 0; pass < 2; ++pass) {
      for (int kt = 0; kt < nkt; ++kt, ++n) {
        mbar_wait_sleep(smem_u32(&bar_s_full), n & 1u);
        tc_fence_after();
        const int kvalid = min(kKeys, S - kt * kKeys);  // keys of this tile that exist (TMA zero-fills the rest)
#pragma unroll 1
        for (int c0 = hf * kHalf; c0 < (hf + 1) * kHalf; c0 += 32) {
          uint32_t v0[16], v1[16];
          tmem_ld16_nowait(lane_addr + (uint32_t)c0, v0);
          tmem_ld16_nowait(lane_addr + (uint32_t)c0 + 16u, v1);
          tmem_ld_wait();
          if (pass == 0) {
            if (c0 + 32 <= kvalid) {
#pragma unroll
              for (int j = 0; j < 16; ++j) m = fmaxf(m, fmaxf(__uint_as_float(v0[j]), __uint_as_float(v1[j])));
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                if (c0 + j < kvalid) m = fmaxf(m, __uint_as_float(v0[j]));
                if (c0 + 16 + j < kvalid) m = fmaxf(m, __uint_as_float(v1[j]));
              }
            }
          } else {
            // p = exp2((s - m) sc); P row r, keys c0 .. c0+31 = four 16-byte units, unit u at physical unit u ^ (r & 7)
            uint32_t pk[16];
            const float ms = m * sc;
            if (c0 + 32 > kvalid) {  // ragged last tile: keys that do not exist get weight 0 (-inf logit)
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                if (c0 + j >= kvalid) v0[j] = 0xff800000u;
                if (c0 + 16 + j >= kvalid) v1[j] = 0xff800000u;
              }
            }
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
              const float a0 = ex2_approx(fmaf(__uint_as_float(v0[j]), sc, -ms));
              const float a1 = ex2_approx(fmaf(__uint_as_float(v0[j + 1]), sc, -ms));
              const float b0 = ex2_approx(fmaf(__uint_as_float(v1[j]), sc, -ms));
              const float b1 = ex2_approx(fmaf(__uint_as_float(v1[j + 1]), sc, -ms));
              const __nv_bfloat162 pa = __floats2bfloat162_rn(a0, a1), pb = __floats2bfloat162_rn(b0, b1);
              // the row sum is taken over the bf16-rounded weights the tensor core multiplies with
              l += (__low2float(pa) + __high2float(pa)) + (__low2float(pb) + __high2float(pb));
              pk[j >> 1] = *reinterpret_cast<const uint32_t*>(&pa);
              pk[8 + (j >> 1)] = *reinterpret_cast<const uint32_t*>(&pb);
            }
            uint8_t* row = p_gen + (c0 >> 6) * (kAtQ * 128) + r * 128;
            const int u0 = (c0 & 63) >> 3;
#pragma unroll
            for (int u = 0; u < 4; ++u)
              *reinterpret_cast<uint4*>(row + (((u0 + u) ^ (r & 7)) << 4)) =
                  make_uint4(pk[4 * u], pk[4 * u + 1], pk[4 * u + 2], pk[4 * u + 3]);
          }
        }
        tc_fence_before();
        if (pass == 0) {
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&bar_s_free));
          if (kt == nkt - 1) {  // the row maximum over both column halves
            part[hf][r] = m;
            named_bar_sync(1, 256);
            m = fmaxf(part[0][r], part[1][r]);
            named_bar_sync(1, 256);  // part[] is reused for the row sums
          }
        } else {
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // P: generic-proxy writes -> tensor-core reads
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&bar_p_ready));  // also: this warp is done reading S
        }
      }
    }
    // epilogue: O / l -> bf16, one row of C channels per thread
    mbar_wait_sleep(smem_u32(&bar_pv_done), (uint32_t)(nkt - 1) & 1u);
    tc_fence_after();
    part[hf][r] = l;
    named_bar_sync(1, 256);
    const float inv = 1.0f / (part[0][r] + part[1][r]);
    const int qi = q0 + r;
    bf16* dst = o + ((int64_t)b * S + qi) * ld_o;
#pragma unroll 1
    for (int c0 = hf * (C / 2); c0 < (hf + 1) * (C / 2); c0 += 16) {
      uint32_t v[16];
      tmem_ld16(lane_addr + (uint32_t)(kAtOCol + c0), v);
      if (qi < S) {
        uint32_t pk[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[2 * j]) * inv, __uint_as_float(v[2 * j + 1]) * inv);
          pk[j] = *reinterpret_cast<const uint32_t*>(&h);
        }
        *reinterpret_cast<uint4*>(dst + c0) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        *reinterpret_cast<uint4*>(dst + c0 + 8) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem, 512u);
  }
}

}  // namespace ditsep
