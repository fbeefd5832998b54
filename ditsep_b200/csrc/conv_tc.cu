// Implicit-GEMM convolution on the 5th-generation tensor cores (tcgen05 / TMEM), TMA-fed.
//
//   out[b,y,x,co] = out_scale * ( sum_{tap,ci} in[b,y+dy,x+dx,ci] * w[co,tap,ci]
//                                 + bias[co] + bias_b[b,co] + residual[b,y,x,co] )
// 3x3 (pad 1) or 1x1, stride 1, NHWC bf16 activations, bf16 filters [CoutPad][tap][Cin],
// fp32 accumulation in TMEM.  Reference semantics: nn.Conv2d through ddpm_conv3x3 /
// ddpm_conv1x1 (ncsnpp_utils/layers.py:112-156), NIN (layers.py:678-689), with the
// res-block tail (layerspp.py:314-328) fused as the epilogue.
//
// GEMM view: M = 128 output pixels per CTA (a bw x bh x bb box of the image batch),
// N = BN <= 128 output channels, K = taps * Cin walked in blocks of 64 channels.
//  * A operand: NO im2col.  The activation tensor is described to TMA as a 4-D tensor
//    (C, W, H, B); the K-block of tap (dy,dx) is ONE cp.async.bulk.tensor box
//    (64 ch, bw, bh, bb) at coordinates (c0, x0+dx, y0+dy, b0).  Out-of-image pixels are
//    zero-filled by the TMA unit (that is the conv's zero padding), and the box lands in
//    shared memory as 128 rows of 128 bytes with the 128-byte swizzle = the canonical
//    K-major UMMA operand layout.
//  * B operand: the filter as a 2-D tensor (K, CoutPad), box (64, BN), same swizzle.
//  * warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread tcgen05.mma issuer,
//    warps 2-5 = epilogue (tcgen05.ld 32 lanes x 16 columns -> bias / temb / residual /
//    scale -> global).  kStages-deep full/empty mbarrier ring between producer and MMA;
//    tcgen05.commit releases a stage when the MMAs that read it have retired.
//  * several CTAs are resident per SM (smem ring 72-128 KB, <= 128 TMEM columns each), so one
//    CTA's epilogue overlaps another's main loop.
#include <cuda.h>

#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace ditsep {

constexpr int kTileM = 128;
constexpr int kBlockK = 64;  // bf16 elements = 128 bytes = one swizzle row
constexpr int kABytes = kTileM * kBlockK * 2;
constexpr int kMaxStages = 6;
constexpr int kThreads = 192;

struct ConvTcParams {
  int B, H, W, Cin, Cout, taps, ksize;
  int bw, bh, bb;          // pixel box of one M tile (bw*bh*bb == 128), all powers of two
  int tiles_x, tiles_y, tiles_b;
  int BN, n_tiles, tmem_cols, stages;
  const float* bias;
  const float* bias_b;
  int ld_bias_b;
  const void* residual;
  int res_dtype, ld_res;
  float out_scale;
  void* out;
  int out_dtype, ld_out;
  float* stats;            // nullable: per-(b, channel) {sum, sum of squares} of the OUTPUT, [B][ld_stats][2]
  int ld_stats;
  // halo mode (3x3): 8 x 16 pixel tiles, persistent CTAs, optional resident filter
  int num_tiles, chunks, w_resident, a_stages, w_stages, base_off_mode;
};

// ---- PTX wrappers ------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must not hang the GPU; after ~seconds it traps (sticky error).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 28)) __trap();
  }
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128-byte-swizzled operand tile: rows of 128 bytes, 8-row groups 1024 bytes apart.
// (cute::UMMA::SmemDescriptor: start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46), version=1 [46,48),
//  layout SWIZZLE_128B=2 [61,64).)
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFF) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

// ---- shared epilogue -----------------------------------------------------------------------
// One thread = one output pixel (TMEM lane), 16 consecutive output channels per call.
// v: raw fp32 accumulators.  Applies bias / per-image bias / residual / scale, stores, and (if
// p.stats) accumulates the GroupNorm statistics of the stored tensor: a 4-step halving exchange
// leaves each lane pair with one channel's sum over the warp's 32 pixels, then one red.add per
// channel per warp.  `uniform_b`: all lanes of the warp belong to the same image.
__device__ __forceinline__ void epilogue16(const ConvTcParams& p, const uint32_t (&v)[16], bool valid, int pb,
                                           int64_t pix, int co0, bool uniform_b, int lane) {
  const int nvalid = min(16, p.Cout - co0);  // > 0 guaranteed by the caller
  float f[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    float a = __uint_as_float(v[j]);
    if (j < nvalid) {
      if (p.bias) a += __ldg(p.bias + co0 + j);
      if (p.bias_b && valid) a += __ldg(p.bias_b + (int64_t)pb * p.ld_bias_b + co0 + j);
    }
    f[j] = a;
  }
  if (valid) {
    if (p.residual) {
      if (p.res_dtype == DITSEP_BF16) {
        const bf16* rp = reinterpret_cast<const bf16*>(p.residual) + pix * p.ld_res + co0;
        if (nvalid == 16 && ((reinterpret_cast<uintptr_t>(rp) & 15) == 0)) {
          float t[8];
          Pack8<bf16>::load(rp, t);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] += t[j];
          Pack8<bf16>::load(rp + 8, t);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[8 + j] += t[j];
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (j < nvalid) f[j] += __bfloat162float(rp[j]);
        }
      } else {
        const float* rp = reinterpret_cast<const float*>(p.residual) + pix * p.ld_res + co0;
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < nvalid) f[j] += rp[j];
      }
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) f[j] *= p.out_scale;
    if (p.out_dtype == DITSEP_BF16) {
      bf16* op = reinterpret_cast<bf16*>(p.out) + pix * p.ld_out + co0;
      if (nvalid == 16 && ((reinterpret_cast<uintptr_t>(op) & 15) == 0)) {
        float t[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) t[j] = f[j];
        Pack8<bf16>::store(op, t);
#pragma unroll
        for (int j = 0; j < 8; ++j) t[j] = f[8 + j];
        Pack8<bf16>::store(op + 8, t);
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < nvalid) op[j] = __float2bfloat16_rn(f[j]);
      }
      if (p.stats) {  // statistics of what the consumer will read: the bf16-rounded values
#pragma unroll
        for (int j = 0; j < 16; ++j) f[j] = __bfloat162float(__float2bfloat16_rn(f[j]));
      }
    } else {
      float* op = reinterpret_cast<float*>(p.out) + pix * p.ld_out + co0;
      if (nvalid == 16 && ((reinterpret_cast<uintptr_t>(op) & 15) == 0)) {
#pragma unroll
        for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(op + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < nvalid) op[j] = f[j];
      }
    }
  }
  if (p.stats == nullptr) return;
  if (!uniform_b) {  // deep levels: a warp spans several images -> plain per-thread reds (few pixels)
    if (valid) {
      float* sp = p.stats + ((int64_t)pb * p.ld_stats + co0) * 2;
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (j < nvalid) {
          atomicAdd(sp + 2 * j, f[j]);
          atomicAdd(sp + 2 * j + 1, f[j] * f[j]);
        }
    }
    return;
  }
  float s[16], q[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    s[j] = valid ? f[j] : 0.f;
    q[j] = s[j] * s[j];
  }
  // halving exchange: after the steps with offsets 16, 8, 4, 2 lane l holds channel
  // c = 8*b4 + 4*b3 + 2*b2 + b1 (bits of l) summed over the 16 lanes that share those bits' complement
#pragma unroll
  for (int step = 0; step < 4; ++step) {
    const int off = 16 >> step;       // lane offset
    const int half = 8 >> step;       // values kept
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (j < half) {
        const float send_s = upper ? s[j] : s[j + half];
        const float send_q = upper ? q[j] : q[j + half];
        const float keep_s = upper ? s[j + half] : s[j];
        const float keep_q = upper ? q[j + half] : q[j];
        s[j] = keep_s + __shfl_xor_sync(0xffffffffu, send_s, off);
        q[j] = keep_q + __shfl_xor_sync(0xffffffffu, send_q, off);
      }
    }
  }
  s[0] += __shfl_xor_sync(0xffffffffu, s[0], 1);
  q[0] += __shfl_xor_sync(0xffffffffu, q[0], 1);
  const int c = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
  if ((lane & 1) == 0 && c < nvalid) {
    float* sp = p.stats + ((int64_t)pb * p.ld_stats + co0 + c) * 2;
    atomicAdd(sp, s[0]);
    atomicAdd(sp + 1, q[0]);
  }
}

// ---- kernel 1: one TMA box per (tap, 64-channel block) --------------------------------------
__global__ void __launch_bounds__(kThreads) conv_tc_kernel(const __grid_constant__ CUtensorMap mapA,
                                                           const __grid_constant__ CUtensorMap mapW,
                                                           const ConvTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[kMaxStages];
  __shared__ __align__(8) uint64_t bar_empty[kMaxStages];
  __shared__ __align__(8) uint64_t bar_accum;
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1024-byte aligned operand ring: [stage][A 16 KB | W BN*128 B]
  const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t w_bytes = (uint32_t)p.BN * 128u;
  const uint32_t stage_bytes = kABytes + w_bytes;  // multiple of 1024 (BN multiple of 16 -> 2048)

  // tile coordinates
  int mt = blockIdx.x;
  const int tx = mt % p.tiles_x;
  mt /= p.tiles_x;
  const int ty = mt % p.tiles_y;
  const int tb = mt / p.tiles_y;
  const int x0 = tx * p.bw, y0 = ty * p.bh, b0 = tb * p.bb;
  const int n0 = blockIdx.y * p.BN;
  const int kc_per_tap = p.Cin / kBlockK;
  const int num_kb = p.taps * kc_per_tap;
  const int pad = p.ksize / 2;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    mbar_init(smem_u32(&bar_accum), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_slot), (uint32_t)p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = tmem_base_slot;

  if (warp == 0) {
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapW) : "memory");
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % p.stages;
        const uint32_t ph = (uint32_t)(kb / p.stages) & 1u;
        mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1u);
        const int tap = kb / kc_per_tap, kc = kb - tap * kc_per_tap;
        const int dy = tap / p.ksize - pad, dx = tap % p.ksize - pad;
        const uint32_t full = smem_u32(&bar_full[s]);
        const uint32_t dstA = ring + (uint32_t)s * stage_bytes;
        mbar_expect_tx(full, stage_bytes);
        tma_load_4d(dstA, &mapA, full, kc * kBlockK, x0 + dx, y0 + dy, b0);
        tma_load_2d(dstA + kABytes, &mapW, full, tap * p.Cin + kc * kBlockK, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // instruction descriptor (cute::UMMA::InstrDescriptor): D=f32 [4,6)=1, A=bf16 [7,10)=1,
      // B=bf16 [10,13)=1, both K-major, N>>3 at [17,23), M>>4 at [24,29)
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.BN >> 3) << 17) |
                             ((uint32_t)(kTileM >> 4) << 24);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % p.stages;
        const uint32_t ph = (uint32_t)(kb / p.stages) & 1u;
        mbar_wait(smem_u32(&bar_full[s]), ph);
        tc_fence_after();
        const uint32_t a_addr = ring + (uint32_t)s * stage_bytes;
        const uint64_t da = make_sw128_desc(a_addr), db = make_sw128_desc(a_addr + kABytes);
#pragma unroll
        for (int k = 0; k < kBlockK / 16; ++k)  // +32 bytes along K per UMMA_K = 16
          umma_bf16(tmem_acc, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
        umma_commit(smem_u32(&bar_empty[s]));
      }
      umma_commit(smem_u32(&bar_accum));
    }
  } else {
    // ---- epilogue: warp q owns TMEM lanes [32q, 32q+32) = tile rows (pixels)
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const int xo = r % p.bw, yo = (r / p.bw) % p.bh, bo = r / (p.bw * p.bh);
    const int px = x0 + xo, py = y0 + yo, pb = b0 + bo;
    const bool valid = px < p.W && py < p.H && pb < p.B;
    const int64_t pix = ((int64_t)pb * p.H + py) * p.W + px;
    mbar_wait(smem_u32(&bar_accum), 0);
    tc_fence_after();
    const uint32_t lane_base = tmem_acc + ((uint32_t)(q * 32) << 16);
    for (int c0 = 0; c0 < p.BN; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(lane_base + (uint32_t)c0, v);
      if (n0 + c0 >= p.Cout) continue;
      epilogue16(p, v, valid, pb, pix, n0 + c0, p.bb == 1, lane);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_acc, (uint32_t)p.tmem_cols);
  }
}

// ---- kernel 2: 3x3, halo tiles, persistent CTAs ----------------------------------------------
// M tile = 8 (x) by 16 (y) output pixels of one image.  Per 64-channel block ONE TMA box
// (64 ch, 10, 18) brings the tile plus its one-pixel halo: 180 rows of 128 bytes, swizzled.
// The A operand of tap (dy,dx) is the SAME shared-memory tile read through a descriptor whose
// start is shifted by ((1+dy)*10 + (1+dx)) rows and whose 8-row groups are 10 rows (1280 B)
// apart -- every input element is fetched from L2 once per tile instead of nine times.
// CTAs are persistent over M tiles (static stride); the filter slice of the CTA's N tile stays
// resident in shared memory when it fits, else it streams through a small ring.  Two TMEM
// accumulators alternate so the epilogue of tile i overlaps the MMAs of tile i+1.
constexpr int kHaloW = 10, kHaloH = 18;
constexpr int kHaloBytes = 23 * 1024;  // 180 rows x 128 B = 23040, padded to a 1024 multiple
constexpr int kHaloThreads = 256;
constexpr int kMaxAStages = 4, kMaxWStages = 4;

__global__ void __launch_bounds__(kHaloThreads, 1) conv_tc_halo_kernel(const __grid_constant__ CUtensorMap mapA,
                                                                       const __grid_constant__ CUtensorMap mapW,
                                                                       const ConvTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_a_full[kMaxAStages], bar_a_empty[kMaxAStages];
  __shared__ __align__(8) uint64_t bar_w_full[kMaxWStages], bar_w_empty[kMaxWStages];
  __shared__ __align__(8) uint64_t bar_w_res, bar_acc_full[2], bar_acc_empty[2];
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t w_tile_bytes = (uint32_t)p.BN * 128u;
  const int n_wtiles = p.w_resident ? p.chunks * 9 : p.w_stages;
  const uint32_t a_ring = base + (uint32_t)n_wtiles * w_tile_bytes;  // stays 1024-aligned (BN % 16 == 0)
  const int n0 = blockIdx.y * p.BN;
  const int my_tiles = (p.num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kMaxAStages; ++s) {
      mbar_init(smem_u32(&bar_a_full[s]), 1);
      mbar_init(smem_u32(&bar_a_empty[s]), 1);
    }
    for (int s = 0; s < kMaxWStages; ++s) {
      mbar_init(smem_u32(&bar_w_full[s]), 1);
      mbar_init(smem_u32(&bar_w_empty[s]), 1);
    }
    mbar_init(smem_u32(&bar_w_res), 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&bar_acc_full[s]), 1);
      mbar_init(smem_u32(&bar_acc_empty[s]), 4);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) tmem_alloc(smem_u32(&tmem_base_slot), (uint32_t)p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = tmem_base_slot;

  if (warp == 0) {
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapW) : "memory");
      if (p.w_resident) {
        const uint32_t bar = smem_u32(&bar_w_res);
        mbar_expect_tx(bar, (uint32_t)(p.chunks * 9) * w_tile_bytes);
        for (int ch = 0; ch < p.chunks; ++ch)
          for (int tap = 0; tap < 9; ++tap)
            tma_load_2d(base + (uint32_t)(ch * 9 + tap) * w_tile_bytes, &mapW, bar, tap * p.Cin + ch * kBlockK, n0);
      }
      uint32_t ia = 0, iw = 0;  // running A-block / W-tile counters (ring position + phase)
      for (int it = 0; it < my_tiles; ++it) {
        int mt = (int)blockIdx.x + it * (int)gridDim.x;
        const int tx = mt % p.tiles_x;
        mt /= p.tiles_x;
        const int ty = mt % p.tiles_y;
        const int b = mt / p.tiles_y;
        for (int ch = 0; ch < p.chunks; ++ch, ++ia) {
          const int sa = (int)(ia % (uint32_t)p.a_stages);
          mbar_wait(smem_u32(&bar_a_empty[sa]), ((ia / (uint32_t)p.a_stages) & 1u) ^ 1u);
          const uint32_t full = smem_u32(&bar_a_full[sa]);
          mbar_expect_tx(full, (uint32_t)(kHaloW * kHaloH * 128));
          tma_load_4d(a_ring + (uint32_t)sa * kHaloBytes, &mapA, full, ch * kBlockK, tx * 8 - 1, ty * 16 - 1, b);
          if (!p.w_resident) {
            for (int tap = 0; tap < 9; ++tap, ++iw) {
              const int sw = (int)(iw % (uint32_t)p.w_stages);
              mbar_wait(smem_u32(&bar_w_empty[sw]), ((iw / (uint32_t)p.w_stages) & 1u) ^ 1u);
              const uint32_t wfull = smem_u32(&bar_w_full[sw]);
              mbar_expect_tx(wfull, w_tile_bytes);
              tma_load_2d(base + (uint32_t)sw * w_tile_bytes, &mapW, wfull, tap * p.Cin + ch * kBlockK, n0);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.BN >> 3) << 17) |
                             ((uint32_t)(kTileM >> 4) << 24);
      // A: rows 128 B apart inside a group of 8 output pixels (one tile row), groups 10 rows apart
      const uint64_t a_hi = (1ull << 16) | ((uint64_t)(kHaloW * 128 >> 4) << 32) | (1ull << 46) | (2ull << 61);
      if (p.w_resident) {
        mbar_wait(smem_u32(&bar_w_res), 0);
        tc_fence_after();
      }
      uint32_t ia = 0, iw = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int ab = it & 1;
        mbar_wait(smem_u32(&bar_acc_empty[ab]), (((uint32_t)it >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_acc + (uint32_t)(ab * p.BN);
        for (int ch = 0; ch < p.chunks; ++ch, ++ia) {
          const int sa = (int)(ia % (uint32_t)p.a_stages);
          mbar_wait(smem_u32(&bar_a_full[sa]), (ia / (uint32_t)p.a_stages) & 1u);
          tc_fence_after();
          const uint32_t a_buf = a_ring + (uint32_t)sa * kHaloBytes;
          for (int tap = 0; tap < 9; ++tap) {
            uint32_t w_addr;
            int sw = 0;
            if (p.w_resident) {
              w_addr = base + (uint32_t)(ch * 9 + tap) * w_tile_bytes;
            } else {
              sw = (int)(iw % (uint32_t)p.w_stages);
              mbar_wait(smem_u32(&bar_w_full[sw]), (iw / (uint32_t)p.w_stages) & 1u);
              tc_fence_after();
              w_addr = base + (uint32_t)sw * w_tile_bytes;
              ++iw;
            }
            const int dy = tap / 3, dx = tap % 3;  // already offset by the halo origin (-1, -1)
            const uint32_t a_start = a_buf + (uint32_t)((dy * kHaloW + dx) * 128);
            uint64_t da = a_hi | (uint64_t)((a_start >> 4) & 0x3FFF);
            if (p.base_off_mode == 1) da |= (uint64_t)((a_start >> 7) & 7u) << 49;
            const uint64_t db = make_sw128_desc(w_addr);
#pragma unroll
            for (int k = 0; k < kBlockK / 16; ++k)
              umma_bf16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (ch | tap | k) != 0);
            if (!p.w_resident) umma_commit(smem_u32(&bar_w_empty[sw]));
          }
          umma_commit(smem_u32(&bar_a_empty[sa]));
        }
        umma_commit(smem_u32(&bar_acc_full[ab]));
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const int xo = r & 7, yo = r >> 3;
    for (int it = 0; it < my_tiles; ++it) {
      int mt = (int)blockIdx.x + it * (int)gridDim.x;
      const int tx = mt % p.tiles_x;
      mt /= p.tiles_x;
      const int ty = mt % p.tiles_y;
      const int pb = mt / p.tiles_y;
      const int px = tx * 8 + xo, py = ty * 16 + yo;
      const bool valid = px < p.W && py < p.H;
      const int64_t pix = ((int64_t)pb * p.H + py) * p.W + px;
      const int ab = it & 1;
      mbar_wait(smem_u32(&bar_acc_full[ab]), ((uint32_t)it >> 1) & 1u);
      tc_fence_after();
      const uint32_t lane_base = tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * p.BN);
      for (int c0 = 0; c0 < p.BN; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(lane_base + (uint32_t)c0, v);
        if (c0 + 16 >= p.BN) {  // last TMEM read of this tile: hand the accumulator back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bar_acc_empty[ab])) : "memory");
        }
        if (n0 + c0 >= p.Cout) continue;
        epilogue16(p, v, valid, pb, pix, n0 + c0, true, lane);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_acc, (uint32_t)p.tmem_cols);
  }
}

// ---- host side ------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult qres;
  // resolved at run time so the library has no link-time dependency on libcuda
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(sym);
  return fn;
}

static int pow2_ceil(int v) {
  int p = 1;
  while (p < v) p *= 2;
  return p;
}

}  // namespace ditsep

using namespace ditsep;

struct ditsep_conv_plan {
  CUtensorMap mapA, mapW;
  ConvTcParams p;
  size_t smem;
  int halo;      // 1 = conv_tc_halo_kernel
  int grid_x;
};

extern "C" int ditsep_conv_tc_plan_create(ditsep_conv_plan** plan, const void* in, int B, int H, int W, int Cin,
                                          int ld_in, const void* w_bf16, int Cout, int ksize) {
  DITSEP_REQUIRE(plan && in && w_bf16, "conv_tc_plan_create: null pointer");
  DITSEP_REQUIRE(ksize == 1 || ksize == 3, "conv_tc: ksize=%d (1 or 3)", ksize);
  DITSEP_REQUIRE(B > 0 && H > 0 && W > 0, "conv_tc: empty problem");
  DITSEP_REQUIRE(Cin > 0 && Cin % kBlockK == 0, "conv_tc: Cin=%d must be a multiple of %d", Cin, kBlockK);
  DITSEP_REQUIRE(Cout > 0, "conv_tc: Cout=%d", Cout);
  DITSEP_REQUIRE(ld_in >= Cin && ld_in % 8 == 0, "conv_tc: ld_in=%d must be >= Cin and a multiple of 8", ld_in);
  DITSEP_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(w_bf16) & 15) == 0,
                 "conv_tc: operands must be 16-byte aligned");
  EncodeTiledFn encode = get_encode_fn();
  DITSEP_REQUIRE(encode != nullptr, "conv_tc: cuTensorMapEncodeTiled not available from the driver");

  ConvTcParams p{};
  p.B = B; p.H = H; p.W = W; p.Cin = Cin; p.Cout = Cout; p.ksize = ksize; p.taps = ksize * ksize;
  // pixel box: search power-of-two widths for the best tile utilisation (ties -> widest rows)
  double best = -1.0;
  for (int bw = 128; bw >= 1; bw /= 2) {
    if (bw > pow2_ceil(W)) continue;
    int bh = pow2_ceil(H);
    if (bh > kTileM / bw) bh = kTileM / bw;
    int bb = kTileM / (bw * bh);
    int64_t tiles = (int64_t)ceil_div(W, bw) * ceil_div(H, bh) * ceil_div(B, bb);
    double util = (double)B * H * W / ((double)tiles * kTileM);
    if (util > best + 1e-9) {
      best = util;
      p.bw = bw; p.bh = bh; p.bb = bb;
    }
  }
  p.tiles_x = ceil_div(W, p.bw);
  p.tiles_y = ceil_div(H, p.bh);
  p.tiles_b = ceil_div(B, p.bb);
  const int cout_pad = (Cout + 15) / 16 * 16;
  p.BN = cout_pad <= 128 ? cout_pad : 128;
  if (cout_pad > 128 && cout_pad % 128 != 0) {
    // keep every N tile full: fall back to 64-wide tiles (Cout multiple of 64) or reject
    DITSEP_REQUIRE(cout_pad % 64 == 0, "conv_tc: Cout=%d > 128 must be a multiple of 64", Cout);
    p.BN = 64;
  }
  p.n_tiles = ceil_div(cout_pad, p.BN);
  p.tmem_cols = pow2_ceil(p.BN) < 32 ? 32 : pow2_ceil(p.BN);
  const size_t stage_bytes = (size_t)kABytes + (size_t)p.BN * 128;
  p.stages = p.BN > 64 ? 4 : 3;
  const int num_kb = p.taps * (Cin / kBlockK);
  if (p.stages > num_kb) p.stages = num_kb;

  // ---- halo mode: 3x3 on images of at least one 8 x 16 tile (DITSEP_CONV_MODE=tap disables it)
  const char* mode_env = getenv("DITSEP_CONV_MODE");
  int halo = (ksize == 3 && W >= 8 && H >= 16 && !(mode_env && strcmp(mode_env, "tap") == 0)) ? 1 : 0;
  size_t smem = 1024 + (size_t)p.stages * stage_bytes;
  int grid_x = 0;
  if (halo) {
    const char* bo = getenv("DITSEP_TC_BASEOFF");
    p.base_off_mode = bo ? atoi(bo) : 0;
    p.chunks = Cin / kBlockK;
    p.tiles_x = ceil_div(W, 8);
    p.tiles_y = ceil_div(H, 16);
    p.tiles_b = B;
    p.num_tiles = p.tiles_x * p.tiles_y * B;
    p.bw = 8; p.bh = 16; p.bb = 1;
    if (p.BN > 64 && cout_pad % 64 == 0) p.BN = 64;  // more M tiles per resident filter byte
    p.n_tiles = ceil_div(cout_pad, p.BN);
    p.tmem_cols = pow2_ceil(2 * p.BN) < 32 ? 32 : pow2_ceil(2 * p.BN);
    const size_t w_tile = (size_t)p.BN * 128;
    const size_t budget = 225 * 1024 - 1024;
    const size_t w_all = (size_t)p.chunks * 9 * w_tile;
    p.a_stages = p.chunks >= 2 ? 3 : 3;
    if (w_all + (size_t)2 * kHaloBytes <= budget) {
      p.w_resident = 1;
      p.w_stages = 0;
      while (p.a_stages > 2 && w_all + (size_t)p.a_stages * kHaloBytes > budget) --p.a_stages;
      smem = 1024 + w_all + (size_t)p.a_stages * kHaloBytes;
    } else {
      p.w_resident = 0;
      p.w_stages = 4;
      p.a_stages = 4;
      smem = 1024 + (size_t)p.w_stages * w_tile + (size_t)p.a_stages * kHaloBytes;
    }
    int dev = 0, sms = kNumSMs;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    grid_x = sms / p.n_tiles;
    if (grid_x < 1) grid_x = 1;
    if (grid_x > p.num_tiles) grid_x = p.num_tiles;
  }

  ditsep_conv_plan* pl = new ditsep_conv_plan();
  pl->p = p;
  pl->smem = smem;
  pl->halo = halo;
  pl->grid_x = grid_x;
  {
    cuuint64_t gdim[4] = {(cuuint64_t)Cin, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t gstr[3] = {(cuuint64_t)ld_in * 2, (cuuint64_t)W * ld_in * 2, (cuuint64_t)H * W * ld_in * 2};
    cuuint32_t box[4] = {(cuuint32_t)kBlockK, (cuuint32_t)p.bw, (cuuint32_t)p.bh, (cuuint32_t)p.bb};
    if (halo) { box[1] = kHaloW; box[2] = kHaloH; box[3] = 1; }
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = encode(&pl->mapA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(in), gdim, gstr, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      delete pl;
      DITSEP_REQUIRE(false, "conv_tc: cuTensorMapEncodeTiled(activations) failed with %d", (int)r);
    }
  }
  {
    const int K = p.taps * Cin;
    cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)cout_pad};
    cuuint64_t gstr[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {(cuuint32_t)kBlockK, (cuuint32_t)p.BN};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = encode(&pl->mapW, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w_bf16), gdim, gstr, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      delete pl;
      DITSEP_REQUIRE(false, "conv_tc: cuTensorMapEncodeTiled(filter) failed with %d", (int)r);
    }
  }
  static bool attr_set = false;
  if (!attr_set) {
    DITSEP_CUDA(cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     1024 + kMaxStages * (kABytes + 128 * 128)));
    DITSEP_CUDA(cudaFuncSetAttribute(conv_tc_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
    attr_set = true;
  }
  *plan = pl;
  return 0;
}

extern "C" void ditsep_conv_tc_plan_destroy(ditsep_conv_plan* plan) { delete plan; }

extern "C" int ditsep_conv_tc_run(const ditsep_conv_plan* plan, const float* bias, const float* bias_b, int ld_bias_b,
                                  const void* residual, int res_dtype, int ld_res, float out_scale, void* out,
                                  int out_dtype, int ld_out, float* stats, int ld_stats, void* stream) {
  DITSEP_REQUIRE(plan && out, "conv_tc_run: null pointer");
  DITSEP_REQUIRE(out_dtype == DITSEP_F32 || out_dtype == DITSEP_BF16, "conv_tc_run: unknown out dtype %d", out_dtype);
  DITSEP_REQUIRE(!residual || res_dtype == DITSEP_F32 || res_dtype == DITSEP_BF16, "conv_tc_run: unknown residual dtype");
  DITSEP_REQUIRE(ld_out >= plan->p.Cout, "conv_tc_run: ld_out=%d < Cout=%d", ld_out, plan->p.Cout);
  DITSEP_REQUIRE(!stats || ld_stats >= plan->p.Cout, "conv_tc_run: ld_stats=%d < Cout=%d", ld_stats, plan->p.Cout);
  ConvTcParams p = plan->p;
  p.bias = bias; p.bias_b = bias_b; p.ld_bias_b = ld_bias_b;
  p.residual = residual; p.res_dtype = res_dtype; p.ld_res = ld_res;
  p.out_scale = out_scale; p.out = out; p.out_dtype = out_dtype; p.ld_out = ld_out;
  p.stats = stats; p.ld_stats = ld_stats;
  if (plan->halo) {
    dim3 grid((unsigned)plan->grid_x, (unsigned)p.n_tiles);
    conv_tc_halo_kernel<<<grid, kHaloThreads, plan->smem, (cudaStream_t)stream>>>(plan->mapA, plan->mapW, p);
    return check_launch("conv_tc_halo_kernel");
  }
  dim3 grid((unsigned)((int64_t)p.tiles_x * p.tiles_y * p.tiles_b), (unsigned)p.n_tiles);
  conv_tc_kernel<<<grid, kThreads, plan->smem, (cudaStream_t)stream>>>(plan->mapA, plan->mapW, p);
  return check_launch("conv_tc_kernel");
}
