// 3x3 convolution 64 -> 64 channels with the FILTER as the tensor-memory A operand (tcgen05.mma "TS" form) and the
// PIXELS on the N axis.  Included by conv_tc.cu (shares its TileWalk, tensor maps and plan).
//
// Why (profiles/r2a_umma_probe.txt): with pixels on M = 128 and Cout = 64 on N, both operands in shared memory, one
// 128 x 64 x 16 MMA takes 43 + N/2 = 75 cycles against an ideal 32 -- the floor of conv_tc_tile_kernel on the
// 64-channel layers (81 % of the network's FLOPs).  With A in tensor memory the cost is 10 + N/2.  The filter is
// static per layer, so it is written to tensor memory once per CTA and the pixel tile becomes the B operand
// (N = 128 pixels: 74 cycles per 128 x 128 x 16).
//
// M = 128 rows are filled by stacking TWO TAPS of the 64 output channels:
//   rows  0..63  (half a): W[co][(ky,   kx)]      rows 64..127 (half b): W[co][(ky+1, kx)]
// Both halves multiply the SAME pixel tile, so half b's column of input row r belongs to the output row ABOVE the one
// half a's column belongs to: with columns ordered (row group g, x offset j), out(g, j) = Da(g, j) + Db(g + 1, j) --
// a fixed column offset of one 8-pixel group, added in the epilogue.  An N tile of 16 groups therefore yields 15
// output rows.  Per 16-channel K slice: 3 pair MMAs (ky = 0 | 1, kx = 0..2) + 3 single MMAs (ky = 2, half b zero)
// = 6 MMAs for 9 taps; the fused 1x1 shortcut (conv_tc.cu, xchunks) adds one single MMA per K slice of x.
//
// Pixel tile (B operand): the same TMA box as conv_tc_tile_kernel's A tile -- (8 + 2) x (16 + 2) pixels x 64 channels,
// 128-byte swizzled rows, 8-row groups 10 rows apart, start shifted by (dy * 10 + dx) rows per tap.
//
// Tensor memory (all 512 columns): [0, 256) two 128-column fp32 accumulators; [256, 448) the filter: 24 A tiles of 8
// columns (4 K slices x 6 MMAs); [448, 512) up to 8 A tiles of the shortcut filter.
//
// Epilogue: channels sit on TMEM LANES (thread = channel), pixels on columns.  Two groups of 8 warps take the tiles in
// turn (group = accumulator = tile parity); a warp (lane quadrant q, column half kh) covers 64 columns in two passes of
// 4 row groups and hands the accumulator back as soon as both passes are in registers.  Quadrants 0/1 hold half a, 2/3
// half b of the same channels; the partner warps (q, kh) and (q + 2, kh) exchange half of their columns through shared
// memory (64-thread named barriers) and each finishes two of a pass's four output rows: scale + bias in packed fp32x2
// arithmetic, bf16 store (64 contiguous bytes per warp and pixel), GroupNorm statistics in registers (no cross-lane
// reduction: a thread IS a channel) -- one fp32 partial per thread and TILE, converted to fixed point at once and
// carried as integers, so the per-image totals are bit-reproducible and independent of the batch size and of how
// the tiles are dealt to CTAs; they reach global memory when the CTA moves to another image.
//
// Measured (profiles/r2w_conv_ts.txt): 1655 cycles per tile of 120 output pixels = 24 MMAs x 69 cycles, i.e. the
// kernel runs at the tensor pipe's TS rate; under sustained load the SM clock drops to ~1.3-1.6 GHz (power cap).
#pragma once

namespace ditsep {

constexpr int kTsRows = 15;        // output rows per tile (16 row groups on N, one lost to the half-b offset)
constexpr int kTsN = 128;
constexpr int kTsABase = 256;      // first TMEM column of the filter tiles
constexpr int kTsAX = 448;         // first TMEM column of the shortcut filter tiles
constexpr int kTsMaxStages = 6;
constexpr int kTsStageBytes = 23 * 1024;  // (8 + 2) x (16 + 2) rows of 128 bytes = 23040, rounded to 1024
constexpr int kTsExchRow = 20;     // floats per channel row of an exchange buffer: 16 pixels + 4 (conflict-free float4 access)
constexpr int kTsExchBytes = 8 * 2 * 2 * 32 * kTsExchRow * 4;  // [pair][tile parity][direction][32 channels][20] fp32
constexpr int kTsThreads = (kCtrlWarps + kEpiWarps) * 32;

struct ConvTsParams {
  int B, H, W, xchunks;
  int tiles_x, tiles_y, tiles_per_image, seg_len, segs_per_image;
  int stages;
  const bf16* w;           // [64][ldw]: taps (ky * 3 + kx) * 64 + ci, then the shortcut filter
  int ldw;
  const float* bias;
  const float* bias_b;
  int ld_bias_b;
  float out_scale;
  bf16* out;
  int ld_out;
  long long* stats;
  long long* trace;      // debug (ditsep_debug_conv_trace): [4 CTAs][kTraceTiles][16] clock64 stamps, else nullptr
};

// Packed fp32 pairs in a 64-bit register (sm_100: add / fma on two fp32 lanes per instruction).
__device__ __forceinline__ uint64_t f2_pack(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}

// CTA c walks the tiles [c*T/G, (c+1)*T/G) of the B * tiles_per_image tiles (x fastest, then y, then image): divisions
// once at the start, increments afterwards.  The statistics are accumulated as fixed-point integers per TILE (below),
// so no unit coarser than a tile has to be kept together for batch invariance.
struct TsWalk {
  int left, b, tx, ty;
  template <typename P>
  __device__ __forceinline__ void init(const P& p) {
    const int64_t T = (int64_t)p.B * p.tiles_per_image;
    const int t0 = (int)((int64_t)blockIdx.x * T / gridDim.x);
    left = (int)((int64_t)(blockIdx.x + 1) * T / gridDim.x) - t0;
    b = t0 / p.tiles_per_image;
    const int r = t0 - b * p.tiles_per_image;
    ty = r / p.tiles_x;
    tx = r - ty * p.tiles_x;
  }
  __device__ __forceinline__ bool done() const { return left <= 0; }
  template <typename P>
  __device__ __forceinline__ void next(const P& p) {
    --left;
    if (++tx == p.tiles_x) {
      tx = 0;
      if (++ty == p.tiles_y) { ty = 0; ++b; }
    }
  }
};

// Per-tile clock64 stamps (scripts/trace_conv.py) only in builds with -DDITSEP_TS_TRACE: each call site costs the
// 16 epilogue warps three instructions per tile otherwise.
__device__ __forceinline__ void ts_stamp(const ConvTsParams& p, uint32_t it, int slot) {
#ifdef DITSEP_TS_TRACE
  if (p.trace && blockIdx.x < kTraceCtas && it < kTraceTiles)
    p.trace[((int64_t)blockIdx.x * kTraceTiles + it) * 16 + slot] = clock64();
#endif
}

template <int kLd>  // kLd: the output's pixel stride in elements as a compile-time constant (64, 128, 192: dense tensors
                    // and the U-Net's concat buffers), or 0 = read it from the parameters (one 64-bit add per store)
__global__ void __launch_bounds__(kTsThreads, 1) conv_ts_kernel(const __grid_constant__ CUtensorMap mapA,
                                                                const __grid_constant__ CUtensorMap mapX,
                                                                const __grid_constant__ ConvTsParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[kTsMaxStages], bar_empty[kTsMaxStages];
  __shared__ __align__(8) uint64_t bar_acc_full[2], bar_acc_empty[2], bar_turn[2];
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef DITSEP_TS_TRACE
  if (threadIdx.x == 0 && p.trace && blockIdx.x < kTraceCtas) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    p.trace[(int64_t)blockIdx.x * kTraceTiles * 16 + 10] = clock64();
    p.trace[(int64_t)blockIdx.x * kTraceTiles * 16 + 12] = (long long)gt;
  }
#endif
  const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
  float* const exch = reinterpret_cast<float*>(smem_raw + (ring - smem_u32(smem_raw)) + (size_t)p.stages * kTsStageBytes);
  const int nx = p.xchunks, nsteps = nx + 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTsMaxStages; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&bar_turn[s]), 1);
      mbar_init(smem_u32(&bar_acc_full[s]), 1);
      mbar_init(smem_u32(&bar_acc_empty[s]), kEpiWarps / 2);  // one arrival per warp of the accumulator's group
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_slot), 512u);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_slot;

  // ---- the filter goes to tensor memory once (static data: runs before griddep_wait, under the predecessor's tail).
  // Warp (q, k) writes TMEM lanes [32q, 32q + 32) of the A tiles u = k, k + 4, ...; lane row r = 32q + lane is output
  // channel r & 63 of half r >> 6.  One A tile = 16 input channels = 32 bytes per row = 8 columns.
  if (warp >= kCtrlWarps) {
    const int q = warp & 3, k = (warp - kCtrlWarps) >> 2;
    const int half = q >> 1, co = (q & 1) * 32 + lane;
    const bf16* wrow = p.w + (int64_t)co * p.ldw;
    const uint32_t lane_addr = tmem + ((uint32_t)(q * 32) << 16);
    const int n_units = 24 + 4 * nx;
    for (int u = k; u < n_units; u += 4) {
      int koff = -1;  // element offset of the 16 channels inside the filter row; -1 = zero rows
      uint32_t col;
      if (u < 24) {
        const int s = u / 6, m = u - s * 6;
        const int tap = m < 3 ? (half == 0 ? m : 3 + m) : (half == 0 ? 6 + (m - 3) : -1);
        if (tap >= 0) koff = tap * 64 + 16 * s;
        col = (uint32_t)(kTsABase + u * 8);
      } else {
        const int v = u - 24;  // cx * 4 + s
        if (half == 0) koff = 576 + 16 * v;
        col = (uint32_t)(kTsAX + v * 8);
      }
      uint32_t r[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
      if (koff >= 0) {
        const uint4 lo = __ldg(reinterpret_cast<const uint4*>(wrow + koff));
        const uint4 hi = __ldg(reinterpret_cast<const uint4*>(wrow + koff + 8));
        r[0] = lo.x; r[1] = lo.y; r[2] = lo.z; r[3] = lo.w;
        r[4] = hi.x; r[5] = hi.y; r[6] = hi.z; r[7] = hi.w;
      }
      tmem_st8(lane_addr + col, r);
    }
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (warp == 0) {
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
      if (nx) asm volatile("prefetch.tensormap [%0];" ::"l"(&mapX) : "memory");
      griddep_wait();  // the activations are the predecessor's output
      int sa = 0;
      uint32_t pa = 1;  // "empty" barriers start free
      TsWalk tw;
      uint32_t itp = 0;
      for (tw.init(p); !tw.done(); tw.next(p), ++itp) {
        const int x0 = tw.tx * kTileW, y0 = tw.ty * kTsRows;
        for (int pos = 0; pos < nsteps; ++pos) {
          mbar_wait_sleep(smem_u32(&bar_empty[sa]), pa);
          if (pos == nsteps - 1) ts_stamp(p, itp, 6);
          const uint32_t full = smem_u32(&bar_full[sa]);
          const uint32_t dst = ring + (uint32_t)(sa * kTsStageBytes);
          if (pos < nx) {  // shortcut input: the tile's own 8 x 16 pixels of x, no halo
            mbar_expect_tx(full, (uint32_t)(kTileM * 128));
            tma_load_4d(dst, &mapX, full, pos * kBlockK, x0, y0, tw.b);
          } else {
            mbar_expect_tx(full, (uint32_t)((kTileW + 2) * (kTileH + 2) * 128));
            tma_load_4d(dst, &mapA, full, 0, x0 - 1, y0 - 1, tw.b);
          }
          if (++sa == p.stages) { sa = 0; pa ^= 1u; }
        }
      }
    }
  }
  else if (warp == 1 || warp == 2) {
    // Two issuing warps, as in conv_tc_tile_kernel: warp 1 takes the even tiles of this CTA, warp 2 the odd ones
    // (accumulator = tile parity); one warp's barrier bookkeeping runs under the other's MMAs.  The operand ring is a
    // single in-order queue handed over through bar_turn.
    const uint32_t my = (uint32_t)(warp - 1);
    // instruction descriptor: D = f32, A = B = bf16, K-major, M = 128; N is set per tile (last row of tiles: fewer groups)
    const uint32_t idesc0 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kTileM >> 4) << 24);
    constexpr uint32_t b_hi = (uint32_t)(((kTileW + 2) * 128) >> 4) | (1u << 14) | (2u << 29);  // groups 10 rows apart
    constexpr uint32_t b_hi_x = (uint32_t)((kTileW * 128) >> 4) | (1u << 14) | (2u << 29);       // shortcut box: 8 rows
    constexpr uint32_t lbo = 1u << 16;
    constexpr uint32_t kRow16 = 128 >> 4;
    int sa = 0;
    uint32_t pa = 0, it = 0;
    TsWalk tw;
    for (tw.init(p); !tw.done(); tw.next(p), ++it) {
      if ((it & 1u) != my) {  // the other warp's tile: step over its stages
        for (int i = 0; i < nsteps; ++i)
          if (++sa == p.stages) { sa = 0; pa ^= 1u; }
        continue;
      }
      const uint32_t ab = it & 1u;
      mbar_wait(smem_u32(&bar_acc_empty[ab]), ((it >> 1) & 1u) ^ 1u);
      if (it != 0) mbar_wait(smem_u32(&bar_turn[my]), (my ? (it >> 1) : ((it >> 1) - 1u)) & 1u);
      if (lane == 0) ts_stamp(p, it, 0);
      const uint32_t tmem_d = tmem + ab * (uint32_t)kTsN;
      // row groups this tile needs: its valid output rows + 1 (half b reads one group further), rounded up to even
      const int rows = min(kTsRows, p.H - tw.ty * kTsRows);
      const uint32_t n_cols = (uint32_t)(((rows + 2) >> 1) << 4);  // 16 .. 128
      const uint32_t idesc = idesc0 | ((n_cols >> 3) << 17);
      for (int pos = 0; pos < nsteps; ++pos) {
        const bool last_st = pos == nsteps - 1;
        mbar_wait(smem_u32(&bar_full[sa]), pa);
        if (last_st && lane == 0) mbar_arrive(smem_u32(&bar_turn[my ^ 1u]));  // the ring is the other warp's now
        tc_fence_after();
        if (last_st && lane == 0) ts_stamp(p, it, 1);
        const uint32_t b_lo = ((ring + (uint32_t)(sa * kTsStageBytes)) >> 4) | lbo;
        if (elect_one_sync()) {
          if (pos >= nx) {
            // pair MMAs (m < 3): taps (0, m) | (1, m), pixel tile shifted by (dy = 0, dx = m);
            // single MMAs: tap (2, m - 3), shifted by (dy = 2, dx = m - 3)
#pragma unroll
            for (int s = 0; s < 4; ++s) {
#pragma unroll
              for (int m = 0; m < 6; ++m) {
                const uint32_t shift = (uint32_t)(m < 3 ? m : 2 * (kTileW + 2) + (m - 3));
                const uint64_t db = ((uint64_t)b_hi << 32) | (uint64_t)(b_lo + shift * kRow16 + 2 * s);
                umma_bf16_ts(tmem_d, tmem + (uint32_t)(kTsABase + (s * 6 + m) * 8), db, idesc,
                             (s | m) != 0 ? 1u : (uint32_t)(pos != 0));
              }
            }
          } else {
#pragma unroll
            for (int s = 0; s < 4; ++s) {
              const uint64_t db = ((uint64_t)b_hi_x << 32) | (uint64_t)(b_lo + 2 * s);
              umma_bf16_ts(tmem_d, tmem + (uint32_t)(kTsAX + (pos * 4 + s) * 8), db, idesc,
                           s != 0 ? 1u : (uint32_t)(pos != 0));
            }
          }
          umma_commit(smem_u32(&bar_empty[sa]));
          if (last_st) umma_commit(smem_u32(&bar_acc_full[ab]));
        }
        __syncwarp();
        if (++sa == p.stages) { sa = 0; pa ^= 1u; }
      }
      if (lane == 0) ts_stamp(p, it, 2);
    }
  }
  else if (warp >= kCtrlWarps) {
    griddep_wait();  // bias_b, out and the statistics buffer are shared with earlier kernels
    // Two GROUPS of 8 warps (4 lane quadrants x 2 column halves) take the tiles in turn (group = accumulator = tile
    // parity): the groups run out of phase, so one group's TMEM-load / barrier latencies sit under the other's
    // arithmetic, and the per-tile bookkeeping is paid by 8 warps instead of 16.  A warp covers its 64 columns in two
    // passes of 32 (4 row groups each) and hands the accumulator back after the second pass's loads.
    const int e = warp - kCtrlWarps;
    const int q = warp & 3, grp = e >> 3, kh = (e >> 2) & 1;
    const bool is_b = q >= 2;
    const int qa = q & 1, co = qa * 32 + lane;  // this thread's output channel
    const int pair = grp * 4 + kh * 2 + qa;     // partner warps (qa, kh) and (qa + 2, kh) share named barrier 1 + pair
    float* const ex = exch + pair * (2 * 2 * 32 * kTsExchRow) + lane * kTsExchRow;
    const uint32_t lane_addr = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)grp * (uint32_t)kTsN;
    const int etid = e * 32 + lane;
    const float osc = p.out_scale;
    // byte stride between pixels of the output: a compile-time constant for dense tensors (immediate store offsets)
    const int64_t pix_bytes = kLd ? (int64_t)kLd * 2 : (int64_t)p.ld_out * 2;
    float cb = 0.f;
    uint64_t cb2 = 0ull;
    // statistics of this thread's channel: fixed-point {integer, fraction * 2^32} of sum and sum of squares (fx_add's
    // format), accumulated per TILE from an fp32 partial that depends on the tile only -> order-independent, batch-invariant
    long long fx_s = 0, fx_sf = 0, fx_q = 0, fx_qf = 0;
    const uint64_t osc2 = f2_pack(osc, osc);
    int cur_b = -1;
    uint32_t it = 0;
    // the image's sums get this thread's fixed-point totals when the CTA moves to the next image and at the end
    auto flush_stats = [&](const int b) {
      if (p.stats) {
        unsigned long long* d = reinterpret_cast<unsigned long long*>(p.stats + ((int64_t)b * 64 + co) * 4);
        atomicAdd(d, (unsigned long long)fx_s);
        atomicAdd(d + 1, (unsigned long long)fx_sf);
        atomicAdd(d + 2, (unsigned long long)fx_q);
        atomicAdd(d + 3, (unsigned long long)fx_qf);
        fx_s = fx_sf = fx_q = fx_qf = 0;
      }
    };
    TsWalk tw;
    for (tw.init(p); !tw.done(); tw.next(p), ++it) {
      if ((it & 1u) != (uint32_t)grp) continue;  // the other group's tile
      const int pb = tw.b;
      if (etid == 0) ts_stamp(p, it, 9);
      uint64_t s2 = 0ull, sq2 = 0ull;  // this tile's partial sums, packed fp32 pairs (even | odd pixels of a row)
      if (pb != cur_b) {
        if (cur_b >= 0) flush_stats(cur_b);
        cur_b = pb;
        float a = 0.f;
        if (p.bias) a += __ldg(p.bias + co);
        if (p.bias_b) a += __ldg(p.bias_b + (int64_t)pb * p.ld_bias_b + co);
        cb = a * osc;  // out = fma(acc, scale, (bias + bias_b) * scale)
        cb2 = f2_pack(cb, cb);
      }
      mbar_wait_sleep(smem_u32(&bar_acc_full[grp]), (it >> 1) & 1u);
      tc_fence_after();
      if (etid == 0) ts_stamp(p, it, 3);
      uint8_t* const otile = reinterpret_cast<uint8_t*>(p.out + ((int64_t)(pb * p.H + tw.ty * kTsRows) * p.W + tw.tx * kTileW) * p.ld_out + co);
      const int64_t row_bytes = (int64_t)p.W * p.ld_out * 2;
      const int rows_valid = min(kTsRows, p.H - tw.ty * kTsRows);
      // Both passes' accumulator columns are read before any output is produced, so that the accumulator goes back to
      // the MMA warps ~300 cycles after it filled (r2v trace: holding it through the first pass cost 800 cycles per
      // tile on the MMA side).  Of a pass's 32 values a thread keeps the 16 it finishes; the other 16 go to the partner.
      auto ld_pass = [&](const int ps, uint32_t (&v0)[8], uint32_t (&v1)[8], uint32_t (&v2)[8], uint32_t (&v3)[8]) {
        const int g0 = 4 * (2 * kh + ps) + (is_b ? 1 : 0);  // half b reads one row group further: out(g) = Da(g) + Db(g + 1)
        const uint32_t col = lane_addr + (uint32_t)(g0 * 8);
        tmem_ld8_nowait(col, v0);
        tmem_ld8_nowait(col + 8u, v1);
        tmem_ld8_nowait(col + 16u, v2);
        if (g0 + 3 < 16) {  // half b of the last column group: row group 16 does not exist (output row 15 is not produced)
          tmem_ld8_nowait(col + 24u, v3);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) v3[j] = 0u;
        }
        tmem_ld_wait();
      };
      // exchange buffer = [channel][16 pixels + 4]: a thread moves its 16 values as four float4, conflict-free
      auto send = [&](const int ps, const uint32_t (&lo)[8], const uint32_t (&hi)[8]) {
        float4* const exw = reinterpret_cast<float4*>(ex + (ps * 2 + (is_b ? 1 : 0)) * (32 * kTsExchRow));
        exw[0] = make_float4(__uint_as_float(lo[0]), __uint_as_float(lo[1]), __uint_as_float(lo[2]), __uint_as_float(lo[3]));
        exw[1] = make_float4(__uint_as_float(lo[4]), __uint_as_float(lo[5]), __uint_as_float(lo[6]), __uint_as_float(lo[7]));
        exw[2] = make_float4(__uint_as_float(hi[0]), __uint_as_float(hi[1]), __uint_as_float(hi[2]), __uint_as_float(hi[3]));
        exw[3] = make_float4(__uint_as_float(hi[4]), __uint_as_float(hi[5]), __uint_as_float(hi[6]), __uint_as_float(hi[7]));
      };
      // Packed fp32x2 arithmetic (add.rn.f32x2 / fma.rn.f32x2: two IEEE operations per instruction, same bits as the
      // scalar forms): the SM is issue-bound here (profiles/r2s).
      auto finish = [&](const int ps, const uint32_t (&v)[8], const int i) {
        const int g = 4 * (2 * kh + ps) + (is_b ? 2 : 0) + i;  // half a finishes rows 0,1 of the pass, half b rows 2,3
        if (g < rows_valid) {
          const float4* const exr = reinterpret_cast<const float4*>(ex + (ps * 2 + (is_b ? 0 : 1)) * (32 * kTsExchRow));
          const float4 e0 = exr[2 * i], e1 = exr[2 * i + 1];
          const uint64_t ee[4] = {f2_pack(e0.x, e0.y), f2_pack(e0.z, e0.w), f2_pack(e1.x, e1.y), f2_pack(e1.z, e1.w)};
          uint8_t* op = otile + g * row_bytes;
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const uint64_t a = f2_add(f2_pack(__uint_as_float(v[2 * jj]), __uint_as_float(v[2 * jj + 1])), ee[jj]);
            const uint64_t f = f2_fma(a, osc2, cb2);
            s2 = f2_add(s2, f);
            sq2 = f2_fma(f, f, sq2);
            float f0, f1;
            f2_unpack(f, f0, f1);
            const __nv_bfloat162 h = __floats2bfloat162_rn(f0, f1);  // .x (low half) = f0
            const uint32_t hb = *reinterpret_cast<const uint32_t*>(&h);
            *reinterpret_cast<uint16_t*>(op + (2 * jj) * pix_bytes) = (uint16_t)hb;
            *reinterpret_cast<uint16_t*>(op + (2 * jj + 1) * pix_bytes) = (uint16_t)(hb >> 16);
          }
        }
      };
      auto tile_body = [&](auto isb) {
        constexpr bool kB = decltype(isb)::value;
        uint32_t v0[8], v1[8], v2[8], v3[8], w0[8], w1[8], w2[8], w3[8];
        ld_pass(0, v0, v1, v2, v3);
        if (kB) send(0, v0, v1); else send(0, v2, v3);
        ld_pass(1, w0, w1, w2, w3);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bar_acc_empty[grp]));  // accumulator handed back
        if (etid == 0) ts_stamp(p, it, 4);
        if (kB) send(1, w0, w1); else send(1, w2, w3);
        named_bar_sync(1 + pair, 64);
        if (etid == 0) ts_stamp(p, it, 7);
        if (kB) { finish(0, v2, 0); finish(0, v3, 1); finish(1, w2, 0); finish(1, w3, 1); }
        else { finish(0, v0, 0); finish(0, v1, 1); finish(1, w0, 0); finish(1, w1, 1); }
        named_bar_sync(1 + pair, 64);  // the partner has read this tile's buffers: they may be rewritten
      };
      if (is_b) tile_body(std::true_type{}); else tile_body(std::false_type{});
      if (etid == 0) ts_stamp(p, it, 5);
      if (p.stats) {
        float sa, sb, qa2, qb2;
        f2_unpack(s2, sa, sb);
        f2_unpack(sq2, qa2, qb2);
        const float ts = sa + sb, tq = qa2 + qb2;
        const float fs = floorf(ts), fq = floorf(tq);
        fx_s += (long long)fs;
        fx_sf += (long long)((ts - fs) * 4294967296.0f);
        fx_q += (long long)fq;
        fx_qf += (long long)((tq - fq) * 4294967296.0f);
      }
    }
    if (cur_b >= 0) flush_stats(cur_b);
  }
  tc_fence_before();
  __syncthreads();
#ifdef DITSEP_TS_TRACE
  if (threadIdx.x == 0 && p.trace && blockIdx.x < kTraceCtas) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    p.trace[(int64_t)blockIdx.x * kTraceTiles * 16 + 11] = clock64();
    p.trace[(int64_t)blockIdx.x * kTraceTiles * 16 + 13] = (long long)gt;
  }
#endif
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem, 512u);
  }
}

}  // namespace ditsep
