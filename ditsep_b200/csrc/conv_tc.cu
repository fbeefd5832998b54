// Implicit-GEMM convolution on the 5th-generation tensor cores (tcgen05 / TMEM), TMA-fed.
//
//   out[b,y,x,co] = out_scale * ( sum_{tap,ci} in[b,y+dy,x+dx,ci] * w[co,tap,ci]
//                                 + bias[co] + bias_b[b,co] + residual[b,y,x,co] )
// 3x3 (pad 1) or 1x1, stride 1, NHWC bf16 activations, bf16 filters [CoutPad][tap][Cin],
// fp32 accumulation in TMEM.  Reference semantics: nn.Conv2d through ddpm_conv3x3 /
// ddpm_conv1x1 (ncsnpp_utils/layers.py:112-156), NIN (layers.py:678-689), with the
// res-block tail (layerspp.py:314-328) fused as the epilogue.
//
// GEMM view: M = 128 output pixels per tile, N = BN output channels, K = taps * Cin walked in
// blocks of 64 channels.  NO im2col: the activation tensor is described to TMA as a 4-D tensor
// (C, W, H, B); out-of-image pixels are zero-filled by the TMA unit (= the conv's zero padding)
// and a box lands in shared memory as rows of 128 bytes with the 128-byte swizzle, which is the
// canonical K-major UMMA operand layout.  Two kernels:
//
//  * conv_tc_tile_kernel (images of at least 8 x 16 pixels): the M tile is 8 (x) by 16 (y)
//    pixels of one image.  Per 64-channel block ONE TMA box brings the tile plus its halo
//    ((8+2h) x (16+2h) pixels, h = 1 for 3x3, 0 for 1x1); the A operand of tap (dy,dx) is the
//    SAME shared-memory tile read through a UMMA descriptor whose start is shifted by
//    (dy*(8+2h) + dx) rows and whose 8-row groups are (8+2h) rows apart -- every input element
//    is fetched from L2 once per tile instead of nine times.  CTAs are persistent over a
//    CONTIGUOUS range of segments (1-4 consecutive tiles of one image); the filter slice of the
//    CTA's N tile stays resident in shared memory when it fits (else it streams through a ring).
//    Two TMEM accumulators alternate, and 16 epilogue warps (4 per TMEM lane quadrant, 16 output
//    channels each) drain one while the single MMA thread fills the other.  The epilogue also
//    accumulates, in registers, the per-channel sum / sum of squares of what it stores (the
//    GroupNorm statistics of the output tensor, layerspp.py:297,318) and writes them once per
//    (segment, quadrant) -- no atomics, and a summation order that depends on the image shape
//    only, so the statistics (and with them the whole bf16 network) are bit-reproducible and
//    independent of the batch size.
//  * conv_tc_kernel (smaller images, where a 128-pixel tile spans several images): one TMA box
//    per (tap, 64-channel block), one tile per CTA.
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "tc_ptx.cuh"

namespace ditsep {

constexpr int kTileM = 128;
constexpr int kBlockK = 64;  // bf16 elements = 128 bytes = one swizzle row
constexpr int kABytes = kTileM * kBlockK * 2;
constexpr int kMaxStages = 8;
constexpr int kThreads = 192;
constexpr int kTraceTiles = 96, kTraceCtas = 4;

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
  // tile kernel
  int num_tiles, tiles_per_image, chunks, w_resident, a_stages, w_stages;
  int halo, row_pitch, a_rows, a_stage_bytes;
  int xchunks;             // fused shortcut: 64-channel blocks of a second input x accumulated as a 1x1 tap (0 = none)
  long long* stats;        // nullable: fixed-point per-channel sums [B][Cout][4] (common.cuh fx_add), ADDED to
  int seg_len, segs_per_image;
  long long* trace;        // debug (ditsep_debug_conv_trace): [4 CTAs][kTraceTiles][8] clock64 stamps, else nullptr
};

// Programmatic dependent launch: the tile kernel may start while the kernel before it in the stream drains (that
// kernel calls griddep_launch_dependents early).  Everything before griddep_wait() -- barrier init, TMEM allocation,
// tensor-map prefetch, the resident filter preload (static data) -- overlaps that tail; no thread touches a tensor the
// predecessor writes, or writes anything at all to global memory, before it (griddep_wait: common.cuh).
__device__ __forceinline__ void trace_stamp(const ConvTcParams& p, uint32_t it, int slot) {
  if (p.trace && blockIdx.x < kTraceCtas && blockIdx.y == 0 && it < kTraceTiles)
    p.trace[((int64_t)blockIdx.x * kTraceTiles + it) * 16 + slot] = clock64();
}

// ---- shared epilogue -----------------------------------------------------------------------
// One thread = one output pixel (TMEM lane), 16 consecutive output channels.  v: raw fp32
// accumulators.  Applies bias / per-image bias / residual / scale and stores; on return f[] holds
// the values as stored (bf16-rounded for bf16 outputs; 0 for channels >= Cout or invalid pixels).
__device__ __forceinline__ void epilogue16(const ConvTcParams& p, const uint32_t (&v)[16], bool valid, int pb,
                                           int64_t pix, int co0, float (&f)[16]) {
  const int nvalid = min(16, p.Cout - co0);  // > 0 guaranteed by the caller
#pragma unroll
  for (int j = 0; j < 16; ++j) f[j] = __uint_as_float(v[j]);
  if (p.bias) {
    const float* bp = p.bias + co0;
    if (nvalid == 16 && ((reinterpret_cast<uintptr_t>(bp) & 15) == 0)) {
#pragma unroll
      for (int j = 0; j < 16; j += 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(bp + j));
        f[j] += t.x; f[j + 1] += t.y; f[j + 2] += t.z; f[j + 3] += t.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (j < nvalid) f[j] += __ldg(bp + j);
    }
  }
  if (!valid) {
#pragma unroll
    for (int j = 0; j < 16; ++j) f[j] = 0.f;
    return;
  }
  if (p.bias_b) {
    const float* bp = p.bias_b + (int64_t)pb * p.ld_bias_b + co0;
    if (nvalid == 16 && ((reinterpret_cast<uintptr_t>(bp) & 15) == 0)) {
#pragma unroll
      for (int j = 0; j < 16; j += 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(bp + j));
        f[j] += t.x; f[j + 1] += t.y; f[j + 2] += t.z; f[j + 3] += t.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (j < nvalid) f[j] += __ldg(bp + j);
    }
  }
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
  for (int j = 0; j < 16; ++j) f[j] = (j < nvalid) ? f[j] * p.out_scale : 0.f;
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
    // statistics are those of what the consumer will read: the bf16-rounded values
#pragma unroll
    for (int j = 0; j < 16; ++j) f[j] = __bfloat162float(__float2bfloat16_rn(f[j]));
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

// Out-of-line copy for the tile kernel's rare shapes (fewer than 16 real channels, fp32 output): keeps the
// register budget of its hot path for the bf16 fast path.
__device__ __noinline__ void epilogue16_slow(const ConvTcParams& p, const uint32_t* v, bool valid, int pb, int64_t pix,
                                             int co0, float* f) {
  uint32_t vv[16];
  float ff[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) vv[j] = v[j];
  epilogue16(p, vv, valid, pb, pix, co0, ff);
#pragma unroll
  for (int j = 0; j < 16; ++j) f[j] = ff[j];
}

// ---- kernel 1: one TMA box per (tap, 64-channel block) --------------------------------------
__global__ void __launch_bounds__(kThreads) conv_tc_kernel(const __grid_constant__ CUtensorMap mapA,
                                                           const __grid_constant__ CUtensorMap mapW,
                                                           const ConvTcParams p) {
  griddep_launch_dependents();
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
  griddep_wait();  // launched with programmatic stream serialization: the set-up above overlaps the predecessor

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
      float f[16];
      epilogue16(p, v, valid, pb, pix, n0 + c0, f);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_acc, (uint32_t)p.tmem_cols);
  }
}

// ---- kernel 2: 8 x 16 pixel tiles with halo, persistent CTAs over contiguous segment ranges ----
constexpr int kTileW = 8, kTileH = 16;
// Warps: 0 TMA producer, 1 TMEM allocator + MMA issuer (even tiles), 2 MMA issuer (odd tiles), 3 idle; then 16 epilogue
// warps.  20 warps x 96 registers fill the register file.  (Round 1 also carried four warps that applied the INPUT's
// GroupNorm + SiLU to the landed A tiles in shared memory; measured slower than a separate pass and removed --
// DESIGN.md, negative results.)
constexpr int kCtrlWarps = 4;
constexpr int kEpiWarps = 16;   // warp w: TMEM lane quadrant w % 4, output channels [16*((w-4)/4), +16)
constexpr int kTileThreads = (kCtrlWarps + kEpiWarps) * 32;
constexpr int kMaxAStages = 4, kMaxWStages = 4;
constexpr int kTileMaxBN = 64;

// The unit of work is a SEGMENT: seg_len consecutive tiles of one image (segs_per_image per image; both are
// functions of the image shape only).  CTA c walks segments [c*S/G, (c+1)*S/G), S = B * segs_per_image.
// The GroupNorm partial sums are flushed once per segment, so which pixels are summed together -- and in
// which order -- never depends on the batch size or on the grid: batch-invariant, bit-reproducible.
struct TileWalk {
  int seg, seg_end, b, sidx, lt, lt_end, tx, ty;
  template <typename P>
  __device__ __forceinline__ void start_segment(const P& p) {
    b = seg / p.segs_per_image;
    sidx = seg - b * p.segs_per_image;
    lt = sidx * p.seg_len;
    lt_end = min(lt + p.seg_len, p.tiles_per_image);
    ty = lt / p.tiles_x;
    tx = lt - ty * p.tiles_x;
  }
  template <typename P>
  __device__ __forceinline__ void init(const P& p) {
    const int64_t S = (int64_t)p.B * p.segs_per_image;
    seg = (int)((int64_t)blockIdx.x * S / gridDim.x);
    seg_end = (int)((int64_t)(blockIdx.x + 1) * S / gridDim.x);
    if (seg < seg_end) start_segment(p);
  }
  __device__ __forceinline__ bool done() const { return seg >= seg_end; }
  __device__ __forceinline__ bool last_of_segment() const { return lt + 1 == lt_end; }
  template <typename P>
  __device__ __forceinline__ void next(const P& p) {
    if (++lt == lt_end) {
      if (++seg < seg_end) start_segment(p);
    } else if (++tx == p.tiles_x) {
      tx = 0;
      ++ty;
    }
  }
};

template <int KS, bool kRes, bool kFast>
__global__ void __launch_bounds__(kTileThreads, 1) conv_tc_tile_kernel(const __grid_constant__ CUtensorMap mapA,
                                                                       const __grid_constant__ CUtensorMap mapW,
                                                                       const __grid_constant__ CUtensorMap mapX,
                                                                       const __grid_constant__ ConvTcParams p) {
  constexpr int kTaps = KS * KS, kHalo = KS / 2, kRowPitch = kTileW + 2 * kHalo;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_a_full[kMaxAStages], bar_a_empty[kMaxAStages];
  __shared__ __align__(8) uint64_t bar_w_full[kMaxWStages], bar_w_empty[kMaxWStages];
  __shared__ __align__(8) uint64_t bar_w_res, bar_acc_full[2], bar_acc_empty[2], bar_turn[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(16) float comb[2][kTileMaxBN];  // bias[co] + bias_b[b][co] of the image being stored

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t w_tile_bytes = (uint32_t)p.BN * 128u;
  const int n_wtiles = kRes ? p.chunks * kTaps + p.xchunks : p.w_stages * KS;  // streamed: one stage = one filter ROW
  const uint32_t a_ring = (base + (uint32_t)n_wtiles * w_tile_bytes + 1023u) & ~1023u;
  const int n0 = blockIdx.y * p.BN;

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
      mbar_init(smem_u32(&bar_turn[s]), 1);
      mbar_init(smem_u32(&bar_acc_full[s]), 1);
      mbar_init(smem_u32(&bar_acc_empty[s]), kEpiWarps);  // one arrival per epilogue warp
    }
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
      if (kRes) {
        const uint32_t bar = smem_u32(&bar_w_res);
        mbar_expect_tx(bar, (uint32_t)(p.chunks * kTaps + p.xchunks) * w_tile_bytes);
        for (int ch = 0; ch < p.chunks; ++ch)
          for (int tap = 0; tap < kTaps; ++tap)
            tma_load_2d(base + (uint32_t)(ch * kTaps + tap) * w_tile_bytes, &mapW, bar, tap * p.Cin + ch * kBlockK, n0);
        for (int cx = 0; cx < p.xchunks; ++cx)  // shortcut filter: K columns past the taps
          tma_load_2d(base + (uint32_t)(p.chunks * kTaps + cx) * w_tile_bytes, &mapW, bar, kTaps * p.Cin + cx * kBlockK, n0);
        if (p.xchunks) asm volatile("prefetch.tensormap [%0];" ::"l"(&mapX) : "memory");
      }
      griddep_wait();  // the activations are the predecessor's output
      int sa = 0, sw = 0;
      uint32_t pa = 1, pw = 1;  // "empty" barriers start free: the first pass waits on the opposite parity
      TileWalk tw;
      uint32_t itp = 0;
      const bool xfirst = kRes && p.xchunks > 0 && p.a_stages >= p.xchunks + 2;  // unit order of the MMA warp
      auto load_shortcut = [&](const TileWalk& t) {  // shortcut input: the tile's own 8 x 16 pixels, no halo
        for (int cx = 0; cx < p.xchunks; ++cx) {
          mbar_wait_sleep(smem_u32(&bar_a_empty[sa]), pa);
          const uint32_t full = smem_u32(&bar_a_full[sa]);
          mbar_expect_tx(full, (uint32_t)(kTileM * 128));
          tma_load_4d(a_ring + (uint32_t)(sa * p.a_stage_bytes), &mapX, full, cx * kBlockK, t.tx * kTileW, t.ty * kTileH, t.b);
          if (++sa == p.a_stages) { sa = 0; pa ^= 1u; }
        }
      };
      for (tw.init(p); !tw.done(); tw.next(p), ++itp) {
        if (xfirst) load_shortcut(tw);
        for (int ch = 0; ch < p.chunks; ++ch) {
          mbar_wait_sleep(smem_u32(&bar_a_empty[sa]), pa);
          if (ch == 0) trace_stamp(p, itp, 6);
          const uint32_t full = smem_u32(&bar_a_full[sa]);
          mbar_expect_tx(full, (uint32_t)(p.a_rows * 128));
          tma_load_4d(a_ring + (uint32_t)(sa * p.a_stage_bytes), &mapA, full, ch * kBlockK, tw.tx * kTileW - kHalo,
                      tw.ty * kTileH - kHalo, tw.b);
          if (++sa == p.a_stages) { sa = 0; pa ^= 1u; }
          if (!kRes) {
            for (int row = 0; row < KS; ++row) {  // a filter row (KS taps) per stage: one barrier round trip per 4*KS MMAs
              mbar_wait(smem_u32(&bar_w_empty[sw]), pw);
              const uint32_t wfull = smem_u32(&bar_w_full[sw]);
              mbar_expect_tx(wfull, (uint32_t)KS * w_tile_bytes);
              for (int col = 0; col < KS; ++col)
                tma_load_2d(base + (uint32_t)(sw * KS + col) * w_tile_bytes, &mapW, wfull,
                            (row * KS + col) * p.Cin + ch * kBlockK, n0);
              if (++sw == p.w_stages) { sw = 0; pw ^= 1u; }
            }
          }
        }
        if (!xfirst) load_shortcut(tw);
      }
    }
  } else if (warp == 1 || (kRes && warp == 2)) {
    // All 32 lanes walk the loop (converged, so every operand lives in uniform registers); one elected lane
    // issues.  Everything that can
    // be a compile-time constant is one (tap offsets, K steps), ring positions are wrap-around counters, and
    // descriptors are built from 32-bit low words (shared-memory addresses >> 4 fit 14 bits, so adding row / K
    // offsets never carries into the next field).
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.BN >> 3) << 17) |
                           ((uint32_t)(kTileM >> 4) << 24);
    // A: rows 128 B apart inside a group of 8 output pixels (one tile row), groups kRowPitch rows apart.
    // The swizzle is keyed on absolute shared-memory address bits, so a row-shifted start needs no base
    // offset (measured: tests/test_gpu_kernels.py::test_conv_tc_shifted_descriptor_views).
    constexpr uint32_t a_hi = (uint32_t)((kRowPitch * 128) >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t b_hi = 64u | (1u << 14) | (2u << 29);
    constexpr uint32_t a_hi_x = (uint32_t)((kTileW * 128) >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t lbo = 1u << 16;
    const uint32_t w_step = w_tile_bytes >> 4;
    if (kRes) {
      mbar_wait(smem_u32(&bar_w_res), 0);
      tc_fence_after();
    }
    int sa = 0, sw = 0;
    uint32_t pa = 0, pw = 0, it = 0;
    TileWalk tw;
    if constexpr (kRes) {
      // Resident filter.  The work of this CTA is a flat sequence of UNITS: per tile, the shortcut chunks (one 1x1 tap
      // of 4 MMAs each) and the main K chunks (KS * KS taps of 4 MMAs each), each unit owning one stage of the A ring.
      // Measured (r1k/r1l clock64 traces, scripts/umma_probe.cu): a 128 x 64 x 16 MMA takes 75 cycles in SS mode, the
      // tensor pipe's queue is only ~4 MMAs deep, and every dependent instruction of an issuing warp costs 4-6 cycles,
      // so the barrier probes + fence + loop bookkeeping between two units (300-700 cycles) drain the queue whatever
      // their position in the loop (two software-pipelined variants were measured and were no faster).  Hence TWO
      // issuing warps: warp 1 takes the even tiles of this CTA, warp 2 the odd ones (accumulator = tile parity), and
      // one warp's bookkeeping runs under the other's MMAs.  MMAs of different threads are unordered, which is fine:
      // they touch different accumulators and different stages; each warp commits its own stages and accumulator.
      // The A ring stays a single in-order queue: a warp probes its tile's stages only once the other warp has passed
      // the probe of ITS last stage (bar_turn hand-over) -- a parity wait is only meaningful for the very next phase
      // of a barrier, and a warp probing a whole tile ahead would alias phases.
      constexpr int kRow16 = 128 >> 4;
      const uint32_t my = (uint32_t)(warp - 1);
      int n_my = 0;
      for (tw.init(p); !tw.done(); tw.next(p)) ++n_my;
      const int nc = p.chunks, nx = p.xchunks, nsteps = nc + nx;
      const bool xfirst = nx > 0 && p.a_stages >= nx + 2;  // the producer follows the same order
      const uint32_t w_base = (base >> 4) | lbo;
      for (it = 0; it < (uint32_t)n_my; ++it) {
        if ((it & 1u) != my) {  // the other warp's tile: step over its stages
          for (int i = 0; i < nsteps; ++i)
            if (++sa == p.a_stages) { sa = 0; pa ^= 1u; }
          continue;
        }
        const uint32_t ab = it & 1u;
        mbar_wait(smem_u32(&bar_acc_empty[ab]), ((it >> 1) & 1u) ^ 1u);
        if (it != 0) mbar_wait(smem_u32(&bar_turn[my]), (my ? (it >> 1) : ((it >> 1) - 1u)) & 1u);
        if (lane == 0) trace_stamp(p, it, 0);
        const uint32_t tmem_d = tmem_acc + ab * (uint32_t)p.BN;
        for (int pos = 0; pos < nsteps; ++pos) {
          const bool is_x = xfirst ? pos < nx : pos >= nc;  // fused shortcut: a 1x1 tap over the tile's own pixels of x
          const bool last_st = pos == nsteps - 1;
          const int widx = is_x ? nc * kTaps + (xfirst ? pos : pos - nc) : (xfirst ? pos - nx : pos) * kTaps;
          const uint32_t w_lo = w_base + (uint32_t)widx * w_step;
          mbar_wait(smem_u32(&bar_a_full[sa]), pa);
          if (last_st && lane == 0) mbar_arrive(smem_u32(&bar_turn[my ^ 1u]));  // the ring is the other warp's now
          tc_fence_after();
          if (pos == 0 && lane == 0) trace_stamp(p, it, 1);
          const uint32_t a_lo = ((a_ring + (uint32_t)(sa * p.a_stage_bytes)) >> 4) | lbo;
          if (elect_one_sync()) {
            if (!is_x) {
              // The row loop stays ROLLED: fully unrolled, ptxas hoists all descriptor words to the front and spills
              // uniform registers.  Rolled, a tap is 4 MMAs plus a handful of uniform adds.
              uint32_t a_cur = a_lo, w_cur = w_lo, acc = (uint32_t)(pos != 0);
#pragma unroll 1
              for (int row = 0; row < KS; ++row) {
#pragma unroll
                for (int col = 0; col < KS; ++col) {
#pragma unroll
                  for (int k = 0; k < kBlockK / 16; ++k) {
                    const uint64_t da = ((uint64_t)a_hi << 32) | (uint64_t)(a_cur + col * kRow16 + 2 * k);
                    const uint64_t db = ((uint64_t)b_hi << 32) | (uint64_t)(w_cur + 2 * k);
                    umma_bf16(tmem_d, da, db, idesc, (col | k) != 0 ? 1u : acc);
                  }
                  w_cur += w_step;
                }
                acc = 1u;
                a_cur += (uint32_t)(kRowPitch * kRow16);
              }
            } else {
#pragma unroll
              for (int k = 0; k < kBlockK / 16; ++k) {  // 8 x 16 box without halo: 8-row groups 1024 bytes apart
                const uint64_t da = ((uint64_t)a_hi_x << 32) | (uint64_t)(a_lo + 2 * k);
                const uint64_t db = ((uint64_t)b_hi << 32) | (uint64_t)(w_lo + 2 * k);
                umma_bf16(tmem_d, da, db, idesc, k != 0 ? 1u : (uint32_t)(pos != 0));
              }
            }
            umma_commit(smem_u32(&bar_a_empty[sa]));
            if (last_st) umma_commit(smem_u32(&bar_acc_full[ab]));
          }
          __syncwarp();
          if (++sa == p.a_stages) { sa = 0; pa ^= 1u; }
        }
        if (lane == 0) trace_stamp(p, it, 2);
      }
    } else if (warp == 1) {
      for (tw.init(p); !tw.done(); tw.next(p), ++it) {
        const uint32_t ab = it & 1u;
        mbar_wait(smem_u32(&bar_acc_empty[ab]), ((it >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_acc + ab * (uint32_t)p.BN;
        uint32_t w_lo;
        for (int ch = 0; ch < p.chunks; ++ch) {
          mbar_wait(smem_u32(&bar_a_full[sa]), pa);
          tc_fence_after();
          const uint32_t a_lo = ((a_ring + (uint32_t)(sa * p.a_stage_bytes)) >> 4) | lbo;
          constexpr int kRow16 = 128 >> 4;
          // Streamed filter: the per-tap version waited on a barrier every 4 MMAs (~150 cycles of probe + fence per
          // ~300 cycles of tensor work: the streamed layers ran at 550 TFLOP/s against 800-1000 resident, r1l); a
          // stage is now a whole filter row.
#pragma unroll
          for (int row = 0; row < KS; ++row) {
            mbar_wait(smem_u32(&bar_w_full[sw]), pw);
            tc_fence_after();
            w_lo = ((base + (uint32_t)(sw * KS) * w_tile_bytes) >> 4) | lbo;
            const uint32_t a_row = a_lo + (uint32_t)(row * kRowPitch * kRow16);
            if (elect_one_sync()) {
#pragma unroll
              for (int col = 0; col < KS; ++col) {
#pragma unroll
                for (int k = 0; k < kBlockK / 16; ++k) {
                  const uint64_t da = ((uint64_t)a_hi << 32) | (uint64_t)(a_row + col * kRow16 + 2 * k);
                  const uint64_t db = ((uint64_t)b_hi << 32) | (uint64_t)(w_lo + (uint32_t)col * w_step + 2 * k);
                  umma_bf16(tmem_d, da, db, idesc, (row | col | k) != 0 ? 1u : (uint32_t)(ch != 0));
                }
              }
              umma_commit(smem_u32(&bar_w_empty[sw]));
              if (row == KS - 1) {
                umma_commit(smem_u32(&bar_a_empty[sa]));
                if (ch == p.chunks - 1) umma_commit(smem_u32(&bar_acc_full[ab]));
              }
            }
            __syncwarp();
            if (++sw == p.w_stages) { sw = 0; pw ^= 1u; }
          }
          if (++sa == p.a_stages) { sa = 0; pa ^= 1u; }
        }
      }
    }
  } else if (warp >= kCtrlWarps) {
    griddep_wait();  // bias_b, residual, out and the statistics buffer are shared with earlier kernels
    const int e = warp - kCtrlWarps;
    const int q = warp & 3;            // TMEM lane quadrant this warp may read
    // N tiles of 16 columns (the 6-channel pyramid outputs) would leave twelve of the sixteen warps idle and the
    // other four serialised on global-memory latency: there the four warp groups take the tiles in turn instead.
    const bool spread = p.BN <= 16 && p.stats == nullptr;
    const uint32_t grp = (uint32_t)(e >> 2);
    const int c0 = spread ? 0 : (e >> 2) * 16;  // its 16 output channels inside the N tile
    const bool cols_any = c0 < p.BN;
    const int nvalid = min(16, p.Cout - (n0 + c0));
    const int r = q * 32 + lane;
    const int xo = r & 7, yo = r >> 3;
    const int etid = e * 32 + lane;    // 0..511
    // kFast (host-checked): Cout a multiple of 16, bf16 output (and residual) whose 32-byte channel runs are
    // 16-byte aligned.  The general path lives in a separate instantiation so that it costs the hot one no registers.
    constexpr bool fast = kFast;
    const uint32_t tmem_lane = tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
    float s[16], sq[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) s[j] = sq[j] = 0.f;
    int cur_b = -1;
    uint32_t it = 0, nimg = 0;
    const float* cb = comb[0];
    TileWalk tw;
    for (tw.init(p); !tw.done(); tw.next(p), ++it) {
      const int pb = tw.b;
      if (pb != cur_b) {  // same tile sequence in every epilogue warp: uniform across the 16 warps
        cur_b = pb;
        float* cw = comb[nimg & 1u];
        if (etid < p.BN) {
          float a = 0.f;
          if (n0 + etid < p.Cout) {
            if (p.bias) a += __ldg(p.bias + n0 + etid);
            if (p.bias_b) a += __ldg(p.bias_b + (int64_t)pb * p.ld_bias_b + n0 + etid);
          }
          cw[etid] = kFast ? a * p.out_scale : a;  // fast path: out = fma(acc, scale, comb) (+ fma(res, scale, .))
        }
        // one barrier per image: the buffer being overwritten was last read two images ago
        named_bar_sync(1, kEpiWarps * 32);
        cb = cw;
        ++nimg;
      }
      const bool has_cols = cols_any && (!spread || (it & 3u) == grp);
      const bool active = has_cols && nvalid > 0;
      const int px = tw.tx * kTileW + xo, py = tw.ty * kTileH + yo;
      const bool valid = px < p.W && py < p.H;
      const int64_t pix = (int64_t)((pb * p.H + py) * p.W + px);  // B*H*W < 2^31 pixels
      uint4 r0 = make_uint4(0, 0, 0, 0), r1 = r0;
      if (fast && active && valid && p.residual) {  // issued before the accumulator wait: latency overlaps the MMAs
        const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(p.residual) + pix * p.ld_res + n0 + c0);
        r0 = rp[0];
        r1 = rp[1];
      }
      const uint32_t ab = it & 1u;
      mbar_wait_sleep(smem_u32(&bar_acc_full[ab]), (it >> 1) & 1u);
      tc_fence_after();
      if (e == 0 && lane == 0) trace_stamp(p, it, 3);
      if constexpr (fast) {
        // two passes of 8 channels keep the live set small: v[8] -> 4 packed bf16 pairs -> one 16-byte store
        uint32_t va[8], vb[8];
        if (has_cols) {
          tmem_ld8_nowait(tmem_lane + ab * (uint32_t)p.BN, va);
          tmem_ld8_nowait(tmem_lane + ab * (uint32_t)p.BN + 8u, vb);
          tmem_ld_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bar_acc_empty[ab]));  // accumulator handed back
        if (e == 0 && lane == 0) trace_stamp(p, it, 4);
        if (active) {
          uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(p.out) + pix * p.ld_out + n0 + c0);
          const float osc = p.out_scale;
          // out = acc * scale + (bias + bias_b) * scale [+ residual * scale]: one FMA per element (two with a
          // residual).  The statistics are taken from the fp32 values before the bf16 rounding (the rounding error
          // is zero-mean and 2^-9 relative: 1e-5 on a variance) -- the r1e profile showed this epilogue, not the
          // tensor pipe, pacing the 64-channel layers (6400 warp instructions per tile against 1152 MMA cycles).
          auto half_tile = [&](auto has_res, const int h) {
            const uint32_t (&v)[8] = h ? vb : va;
            const uint4 rq = h ? r1 : r0;
            const uint32_t rr[4] = {rq.x, rq.y, rq.z, rq.w};
            const float4 t0 = *reinterpret_cast<const float4*>(cb + c0 + 8 * h);
            const float4 t1 = *reinterpret_cast<const float4*>(cb + c0 + 8 * h + 4);
            const float cbv[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
            float f[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = fmaf(__uint_as_float(v[j]), osc, cbv[j]);
            if constexpr (decltype(has_res)::value) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                f[2 * j] = fmaf(__uint_as_float(rr[j] << 16), osc, f[2 * j]);
                f[2 * j + 1] = fmaf(__uint_as_float(rr[j] & 0xffff0000u), osc, f[2 * j + 1]);
              }
            }
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const __nv_bfloat162 hh = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
              o[j] = *reinterpret_cast<const uint32_t*>(&hh);
            }
            if (valid) {
              op[h] = make_uint4(o[0], o[1], o[2], o[3]);
              if (p.stats) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  s[8 * h + j] += f[j];
                  sq[8 * h + j] = fmaf(f[j], f[j], sq[8 * h + j]);
                }
              }
            }
          };
          if (p.residual) {
            half_tile(std::true_type{}, 0);
            half_tile(std::true_type{}, 1);
          } else {
            half_tile(std::false_type{}, 0);
            half_tile(std::false_type{}, 1);
          }
        }
      } else {
        uint32_t v[16];
        if (has_cols) tmem_ld16(tmem_lane + ab * (uint32_t)p.BN, v);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bar_acc_empty[ab]));  // accumulator handed back
        if (active) {
          float f[16];
          epilogue16(p, v, valid, pb, pix, n0 + c0, f);
          if (p.stats) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              s[j] += f[j];
              sq[j] = fmaf(f[j], f[j], sq[j]);
            }
          }
        }
      }
      if (e == 0 && lane == 0) trace_stamp(p, it, 5);
      if (p.stats && tw.last_of_segment()) {
        // warp-level transpose-reduce: after offsets 16, 8, 4, 2 lane l holds channel
        // c = 8*b4 + 4*b3 + 2*b2 + b1 (bits of l) summed over 16 lanes; offset 1 completes it
#pragma unroll
        for (int step = 0; step < 4; ++step) {
          const int off = 16 >> step, half = 8 >> step;
          const bool upper = (lane & off) != 0;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (j < half) {
              const float send_s = upper ? s[j] : s[j + half], send_q = upper ? sq[j] : sq[j + half];
              const float keep_s = upper ? s[j + half] : s[j], keep_q = upper ? sq[j + half] : sq[j];
              s[j] = keep_s + __shfl_xor_sync(0xffffffffu, send_s, off);
              sq[j] = keep_q + __shfl_xor_sync(0xffffffffu, send_q, off);
            }
          }
        }
        s[0] += __shfl_xor_sync(0xffffffffu, s[0], 1);
        sq[0] += __shfl_xor_sync(0xffffffffu, sq[0], 1);
        const int c = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
        // Order-independent fixed-point accumulation (fx_add): integer atomics commute, so the per-image sums -- and
        // every bf16 rounding downstream -- are bit-reproducible and independent of the batch size and of which CTA
        // handled which segment; what is added here (one fp32 partial per segment, quadrant and channel) is a function
        // of the image shape only.
        if ((lane & 1) == 0 && has_cols && n0 + c0 + c < p.Cout) {
          long long* dst = p.stats + ((int64_t)pb * p.Cout + n0 + c0 + c) * 4;
          fx_add(dst, s[0]);
          fx_add(dst + 2, sq[0]);
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) s[j] = sq[j] = 0.f;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem_acc, (uint32_t)p.tmem_cols);
  }
}


}  // namespace ditsep
#include "conv_ts.cuh"
namespace ditsep {

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
  CUtensorMap mapA, mapW, mapX;  // mapX: the fused shortcut's input (a copy of mapA when there is none)
  ConvTcParams p;
  size_t smem;
  int tile;      // 1 = conv_tc_tile_kernel
  int grid_x;
  int ts;        // 1 = conv_ts_kernel (64 -> 64 3x3, filter in tensor memory) when the call's epilogue allows it
  ConvTsParams tp;
  size_t ts_smem;
  int ts_grid;
};

extern "C" int ditsep_conv_tc_plan_create(ditsep_conv_plan** plan, const void* in, int B, int H, int W, int Cin,
                                          int ld_in, const void* w_bf16, int Cout, int ksize) {
  return ditsep_conv_tc_plan_create_sc(plan, in, B, H, W, Cin, ld_in, w_bf16, Cout, ksize, nullptr, 0, 0);
}

extern "C" int ditsep_conv_tc_plan_create_sc(ditsep_conv_plan** plan, const void* in, int B, int H, int W, int Cin,
                                             int ld_in, const void* w_bf16, int Cout, int ksize, const void* x_sc, int Cx,
                                             int ld_x) {
  DITSEP_REQUIRE(plan && in && w_bf16, "conv_tc_plan_create: null pointer");
  DITSEP_REQUIRE(!x_sc || (Cx > 0 && Cx % kBlockK == 0 && ld_x >= Cx && ld_x % 8 == 0 &&
                           (reinterpret_cast<uintptr_t>(x_sc) & 15) == 0),
                 "conv_tc: fused shortcut needs Cx %% 64 == 0, ld_x %% 8 == 0, 16-byte alignment (Cx=%d ld_x=%d)", Cx, ld_x);
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

  // ---- tile kernel: images of at least one 8 x 16 tile, N tiles of at most 64 channels
  // (DITSEP_CONV_MODE=tap forces the per-tap kernel: debug aid)
  const char* mode_env = getenv("DITSEP_CONV_MODE");
  int tile = (W >= kTileW && H >= kTileH && (cout_pad <= kTileMaxBN || cout_pad % kTileMaxBN == 0) &&
              !(mode_env && strcmp(mode_env, "tap") == 0)) ? 1 : 0;
  size_t smem = 1024 + (size_t)p.stages * stage_bytes;
  if (!tile) {
    // Small images give few M tiles (4 x 4 at batch 32: four), and a CTA that streams the whole filter through one
    // SM's load path is latency-bound: narrow the N tile until about half the SMs have a CTA, and use the shared
    // memory that frees for a deeper ring.
    const int64_t m_tiles = (int64_t)p.tiles_x * p.tiles_y * p.tiles_b;
    while (p.BN > 32 && p.BN % 32 == 0 && cout_pad % (p.BN / 2) == 0 && m_tiles * ceil_div(cout_pad, p.BN) < 74) p.BN /= 2;
    p.n_tiles = ceil_div(cout_pad, p.BN);
    p.tmem_cols = pow2_ceil(p.BN) < 32 ? 32 : pow2_ceil(p.BN);
    const size_t sb = (size_t)kABytes + (size_t)p.BN * 128;
    p.stages = (int)((size_t)(192 * 1024) / sb);
    if (p.stages > kMaxStages) p.stages = kMaxStages;
    if (p.stages > num_kb) p.stages = num_kb;
    smem = 1024 + (size_t)p.stages * sb;
  }
  int grid_x = 0;
  DITSEP_REQUIRE(!x_sc || tile, "conv_tc: fused shortcut needs the tile kernel (image >= 8 x 16)");
  if (tile) {
    p.halo = ksize / 2;
    p.row_pitch = kTileW + 2 * p.halo;
    p.a_rows = p.row_pitch * (kTileH + 2 * p.halo);
    p.a_stage_bytes = (p.a_rows * 128 + 1023) / 1024 * 1024;
    p.chunks = Cin / kBlockK;
    p.tiles_x = ceil_div(W, kTileW);
    p.tiles_y = ceil_div(H, kTileH);
    p.tiles_b = B;
    p.tiles_per_image = p.tiles_x * p.tiles_y;
    p.num_tiles = p.tiles_per_image * B;
    p.bw = kTileW; p.bh = kTileH; p.bb = 1;
    if (cout_pad > kTileMaxBN) p.BN = kTileMaxBN;
    p.n_tiles = ceil_div(cout_pad, p.BN);
    p.tmem_cols = pow2_ceil(2 * p.BN) < 32 ? 32 : pow2_ceil(2 * p.BN);
    const size_t w_tile = (size_t)p.BN * 128;
    const size_t budget = 222 * 1024;  // dynamic; static (barriers + reduction scratch) is ~2.3 KB
    p.xchunks = x_sc ? Cx / kBlockK : 0;
    const size_t w_all = ((size_t)p.chunks * p.taps + p.xchunks) * w_tile;
    if (w_all + 2 * (size_t)p.a_stage_bytes + 2048 <= budget) {
      p.w_resident = 1;
      p.w_stages = 0;
      p.a_stages = kMaxAStages;
      while (p.a_stages > 2 && w_all + (size_t)p.a_stages * p.a_stage_bytes + 2048 > budget) --p.a_stages;
      smem = 2048 + w_all + (size_t)p.a_stages * p.a_stage_bytes;
    } else {
      DITSEP_REQUIRE(!x_sc, "conv_tc: fused shortcut needs the filter resident in shared memory (%zu bytes do not fit)", w_all);
      p.w_resident = 0;
      p.w_stages = ksize == 3 ? 3 : kMaxWStages;  // a stage holds one filter row (ksize taps)
      p.a_stages = kMaxAStages;
      smem = 2048 + (size_t)p.w_stages * ksize * w_tile + (size_t)p.a_stages * p.a_stage_bytes;
    }
    int dev = 0, sms = kNumSMs;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // segments: the unit of work AND of the statistics partials; a function of the image shape only
    p.seg_len = p.tiles_per_image >= 512 ? 4 : (p.tiles_per_image >= 128 ? 2 : 1);
    p.segs_per_image = ceil_div(p.tiles_per_image, p.seg_len);
    const int64_t n_segs = (int64_t)B * p.segs_per_image;
    grid_x = sms / p.n_tiles;
    if (grid_x < 1) grid_x = 1;
    if (grid_x > n_segs) grid_x = (int)n_segs;
  }

  // ---- filter-in-tensor-memory kernel (conv_ts.cuh): 3x3, 64 -> 64 channels, shortcut of 0 / 64 / 128 channels,
  // images of at least 64 x 64 (below that the 15-of-16 row tiling wastes too much); DITSEP_CONV_TS=0 turns it off
  ConvTsParams tp{};
  int ts = 0, ts_grid = 0;
  size_t ts_smem = 0;
  {
    const char* ts_env = getenv("DITSEP_CONV_TS");
    const int cx = x_sc ? Cx : 0;
    if (tile && ksize == 3 && Cin == 64 && Cout == 64 && (cx == 0 || cx == 64 || cx == 128) && W % kTileW == 0 &&
        H >= 64 && W >= 64 && !(ts_env && strcmp(ts_env, "0") == 0)) {
      ts = 1;
      tp.B = B; tp.H = H; tp.W = W; tp.xchunks = cx / kBlockK;
      tp.tiles_x = W / kTileW;
      tp.tiles_y = ceil_div(H, kTsRows);
      tp.tiles_per_image = tp.tiles_x * tp.tiles_y;
      tp.seg_len = tp.tiles_per_image >= 512 ? 3 : (tp.tiles_per_image >= 128 ? 2 : 1);
      tp.segs_per_image = ceil_div(tp.tiles_per_image, tp.seg_len);
      tp.stages = 5;
      tp.w = reinterpret_cast<const bf16*>(w_bf16);
      tp.ldw = 9 * 64 + cx;
      ts_smem = 1024 + (size_t)tp.stages * kTsStageBytes + kTsExchBytes;
      int dev = 0, sms = kNumSMs;
      if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      const int64_t n_t = (int64_t)B * tp.tiles_per_image;
      ts_grid = (int)(sms < n_t ? sms : n_t);
    }
  }

  ditsep_conv_plan* pl = new ditsep_conv_plan();
  pl->ts = ts; pl->tp = tp; pl->ts_smem = ts_smem; pl->ts_grid = ts_grid;
  pl->p = p;
  pl->smem = smem;
  pl->tile = tile;
  pl->grid_x = grid_x;
  {
    cuuint64_t gdim[4] = {(cuuint64_t)Cin, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t gstr[3] = {(cuuint64_t)ld_in * 2, (cuuint64_t)W * ld_in * 2, (cuuint64_t)H * W * ld_in * 2};
    cuuint32_t box[4] = {(cuuint32_t)kBlockK, (cuuint32_t)p.bw, (cuuint32_t)p.bh, (cuuint32_t)p.bb};
    if (tile) { box[1] = (cuuint32_t)p.row_pitch; box[2] = (cuuint32_t)(kTileH + 2 * p.halo); box[3] = 1; }
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
    const int K = p.taps * Cin + (x_sc ? Cx : 0);  // the shortcut filter rides behind the taps along K
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
  pl->mapX = pl->mapA;
  if (x_sc) {
    cuuint64_t gdim[4] = {(cuuint64_t)Cx, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t gstr[3] = {(cuuint64_t)ld_x * 2, (cuuint64_t)W * ld_x * 2, (cuuint64_t)H * W * ld_x * 2};
    cuuint32_t box[4] = {(cuuint32_t)kBlockK, (cuuint32_t)kTileW, (cuuint32_t)kTileH, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = encode(&pl->mapX, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x_sc), gdim, gstr, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      delete pl;
      DITSEP_REQUIRE(false, "conv_tc: cuTensorMapEncodeTiled(shortcut input) failed with %d", (int)r);
    }
  }
  // cudaFuncAttributeMaxDynamicSharedMemorySize is a PER-DEVICE attribute: one flag per device ordinal
  static std::atomic<bool> attr_set[64];
  int attr_dev = 0;
  DITSEP_CUDA(cudaGetDevice(&attr_dev));
  const bool attr_known = attr_dev >= 0 && attr_dev < 64;
  if (!attr_known || !attr_set[attr_dev].load(std::memory_order_acquire)) {
    DITSEP_CUDA(cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     1024 + 192 * 1024));
#define DITSEP_TILE_ATTR(KS, RES, FAST) \
  DITSEP_CUDA(cudaFuncSetAttribute(conv_tc_tile_kernel<KS, RES, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024))
    DITSEP_TILE_ATTR(1, true, true); DITSEP_TILE_ATTR(1, true, false); DITSEP_TILE_ATTR(1, false, true);
    DITSEP_TILE_ATTR(1, false, false); DITSEP_TILE_ATTR(3, true, true); DITSEP_TILE_ATTR(3, true, false);
    DITSEP_TILE_ATTR(3, false, true); DITSEP_TILE_ATTR(3, false, false);
#undef DITSEP_TILE_ATTR
    DITSEP_CUDA(cudaFuncSetAttribute(conv_ts_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
    DITSEP_CUDA(cudaFuncSetAttribute(conv_ts_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
    DITSEP_CUDA(cudaFuncSetAttribute(conv_ts_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
    DITSEP_CUDA(cudaFuncSetAttribute(conv_ts_kernel<192>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
    if (attr_known) attr_set[attr_dev].store(true, std::memory_order_release);
  }
  *plan = pl;
  return 0;
}

static long long* g_conv_trace = nullptr;
extern "C" void ditsep_debug_conv_trace(void* buf) { g_conv_trace = reinterpret_cast<long long*>(buf); }

extern "C" void ditsep_conv_tc_plan_destroy(ditsep_conv_plan* plan) { delete plan; }

extern "C" int ditsep_conv_tc_has_stats(const ditsep_conv_plan* plan) { return plan && plan->tile ? 1 : 0; }

extern "C" int ditsep_conv_tc_run(const ditsep_conv_plan* plan, const float* bias, const float* bias_b, int ld_bias_b,
                                  const void* residual, int res_dtype, int ld_res, float out_scale, void* out,
                                  int out_dtype, int ld_out, long long* stats, void* stream) {
  DITSEP_REQUIRE(plan && out, "conv_tc_run: null pointer");
  DITSEP_REQUIRE(out_dtype == DITSEP_F32 || out_dtype == DITSEP_BF16, "conv_tc_run: unknown out dtype %d", out_dtype);
  DITSEP_REQUIRE(!residual || res_dtype == DITSEP_F32 || res_dtype == DITSEP_BF16, "conv_tc_run: unknown residual dtype");
  DITSEP_REQUIRE(ld_out >= plan->p.Cout, "conv_tc_run: ld_out=%d < Cout=%d", ld_out, plan->p.Cout);
  DITSEP_REQUIRE(!stats || plan->tile, "conv_tc_run: this plan (image smaller than one 8x16 tile) cannot produce statistics");
  ConvTcParams p = plan->p;
  p.bias = bias; p.bias_b = bias_b; p.ld_bias_b = ld_bias_b;
  p.residual = residual; p.res_dtype = res_dtype; p.ld_res = ld_res;
  p.out_scale = out_scale; p.out = out; p.out_dtype = out_dtype; p.ld_out = ld_out;
  p.stats = stats;
  p.trace = g_conv_trace;
  if (plan->tile) {
    dim3 grid((unsigned)plan->grid_x, (unsigned)p.n_tiles);
    cudaStream_t st = (cudaStream_t)stream;
    const bool fast = p.Cout % 16 == 0 && out_dtype == DITSEP_BF16 && (ld_out & 7) == 0 &&
                      (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
                      (!residual || (res_dtype == DITSEP_BF16 && (ld_res & 7) == 0 &&
                                     (reinterpret_cast<uintptr_t>(residual) & 15) == 0));
    // launched with the programmatic-stream-serialization attribute (see griddep_wait); DITSEP_PDL=0 turns it off
    const bool pdl = pdl_enabled();
    cudaLaunchAttribute lattr[1];
    lattr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    lattr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t lcfg = {};
    lcfg.gridDim = grid;
    lcfg.dynamicSmemBytes = plan->smem;
    lcfg.stream = st;
    lcfg.attrs = lattr;
    lcfg.numAttrs = pdl ? 1 : 0;
    if (plan->ts && fast && !residual) {
      ConvTsParams tp = plan->tp;
      tp.bias = bias; tp.bias_b = bias_b; tp.ld_bias_b = ld_bias_b;
      tp.out_scale = out_scale; tp.out = reinterpret_cast<bf16*>(out); tp.ld_out = ld_out; tp.stats = stats;
      tp.trace = g_conv_trace;
      lcfg.gridDim = dim3((unsigned)plan->ts_grid);
      lcfg.blockDim = dim3(kTsThreads);
      lcfg.dynamicSmemBytes = plan->ts_smem;
      if (ld_out == 64) DITSEP_CUDA(cudaLaunchKernelEx(&lcfg, conv_ts_kernel<64>, plan->mapA, plan->mapX, tp));
      else if (ld_out == 128) DITSEP_CUDA(cudaLaunchKernelEx(&lcfg, conv_ts_kernel<128>, plan->mapA, plan->mapX, tp));
      else if (ld_out == 192) DITSEP_CUDA(cudaLaunchKernelEx(&lcfg, conv_ts_kernel<192>, plan->mapA, plan->mapX, tp));
      else DITSEP_CUDA(cudaLaunchKernelEx(&lcfg, conv_ts_kernel<0>, plan->mapA, plan->mapX, tp));
      return check_launch("conv_ts_kernel");
    }
#define DITSEP_TILE_LAUNCH(KS, RES, FAST)                                                                             \
  do {                                                                                                                \
    lcfg.blockDim = dim3(kTileThreads);                                                                               \
    DITSEP_CUDA(cudaLaunchKernelEx(&lcfg, conv_tc_tile_kernel<KS, RES, FAST>, plan->mapA, plan->mapW, plan->mapX, p)); \
  } while (0)
    if (p.ksize == 3) {
      if (p.w_resident) { if (fast) DITSEP_TILE_LAUNCH(3, true, true); else DITSEP_TILE_LAUNCH(3, true, false); }
      else { if (fast) DITSEP_TILE_LAUNCH(3, false, true); else DITSEP_TILE_LAUNCH(3, false, false); }
    } else {
      if (p.w_resident) { if (fast) DITSEP_TILE_LAUNCH(1, true, true); else DITSEP_TILE_LAUNCH(1, true, false); }
      else { if (fast) DITSEP_TILE_LAUNCH(1, false, true); else DITSEP_TILE_LAUNCH(1, false, false); }
    }
#undef DITSEP_TILE_LAUNCH
    return check_launch("conv_tc_tile_kernel");
  }
  dim3 grid((unsigned)((int64_t)p.tiles_x * p.tiles_y * p.tiles_b), (unsigned)p.n_tiles);
  DITSEP_CUDA(launch_pdl(conv_tc_kernel, grid, dim3(kThreads), plan->smem, (cudaStream_t)stream, plan->mapA, plan->mapW, p));
  return check_launch("conv_tc_kernel");
}
