// HBM-bound NHWC kernels around the convolutions: GroupNorm statistics, the fused
// normalise + affine + SiLU apply, FIR x2 resampling (optionally of both the raw tensor and
// its normalised+activated version in one pass), the time-embedding MLPs and the 1x1
// output layer with the 1/t scaling.
//
// Reference semantics: nn.GroupNorm(min(C/4,32), C, eps=1e-6) + SiLU (layerspp.py:269-271,
// 283-285,297,318), upsample_2d / downsample_2d (up_or_down_sampling.py:206-273 ->
// upfirdn2d, op/upfirdn2d_kernel.cu:107-207), temb (ncsnpp.py:325-346, layerspp.py:44-46,
// 316-317), output scaling + output_layer (ncsnpp.py:473-477).
#include <mutex>
#include <type_traits>

#include "common.cuh"

// fir_down is latency-bound at 4 warps per scheduler (r1p ncu: long-scoreboard stall 4.8 per issue, 22 % occupancy at
// 128 registers): 4 CTAs per SM with ~430 bytes of spills measured 251 us against 303 (2 CTAs, no spills) at 256x256x64.
#ifndef DITSEP_FIR_MINB
#define DITSEP_FIR_MINB 4
#endif
namespace ditsep {

// ------------------------------------------------------------------- GroupNorm statistics
// Explicit statistics pass (for tensors whose producer left none): grid (nsplit, B); block = cv * ppb threads (cv = C/8
// channel packs, ppb pixels in flight).  Each thread keeps sum / sum-of-squares of its 8 channels in registers over its
// pixels; the CTA combines them in shared memory and ADDS one fp32 partial per channel to the fixed-point sums
// [B][C][4] (fx_add).  nsplit is a function of the image shape only, so what is added never depends on the batch.
template <typename T>
__global__ void __launch_bounds__(256) gn_sums_kernel(const T* __restrict__ x, int64_t HW, int C, int ld, int cv, int ppb,
                                                      long long* __restrict__ sums) {
  griddep_launch_dependents();
  griddep_wait();
  extern __shared__ float red[];  // [ppb][C][2]
  const int b = blockIdx.y, split = blockIdx.x, nsplit = gridDim.x;
  const int v = threadIdx.x % cv, pp = threadIdx.x / cv;
  const int64_t per = (HW + nsplit - 1) / nsplit;
  const int64_t p0 = (int64_t)split * per, p1 = min(HW, p0 + per);
  float s[8], q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = q[i] = 0.f;
  if (pp < ppb) {
    const T* base = x + (int64_t)b * HW * ld + v * 8;
    for (int64_t p = p0 + pp; p < p1; p += ppb) {
      float f[8];
      Pack8<T>::load(base + p * ld, f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s[i] += f[i];
        q[i] += f[i] * f[i];
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      red[((pp * C) + v * 8 + i) * 2] = s[i];
      red[((pp * C) + v * 8 + i) * 2 + 1] = q[i];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = 0.f, bq = 0.f;
    for (int j = 0; j < ppb; ++j) {
      a += red[(j * C + c) * 2];
      bq += red[(j * C + c) * 2 + 1];
    }
    long long* dst = sums + ((int64_t)b * C + c) * 4;
    fx_add(dst, a);
    fx_add(dst + 2, bq);
  }
}

// Per-(b, c) scale / shift of one GroupNorm consumer from the fixed-point sums of up to 4 producers laid end to end
// along the channel axis (torch.cat([h, skip], dim=1) feeding GroupNorm_0, ncsnpp.py:413 + layerspp.py:297).
// One image per call; executed by a whole CTA (any block size).  The arithmetic from the exact integer sums on is
// double precision in a fixed order: bit-reproducible.  grp: shared scratch of >= 2 * G doubles.
// Every channel's four words are fetched by a thread of its own first (ONE L2 round trip for the whole image; the
// first version walked a group's channels serially: cpg dependent round trips, ~5 us per image), then a thread per
// group adds its channels in ascending order.
constexpr int kGnMaxC = 512;

template <bool kToShared>
__device__ __forceinline__ void gn_job_image(const ditsep_gn_job& job, int b, double* chan, double* grp, float* ss_out) {
  const int C = job.C, G = job.G, cpg = C / G;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    int cc = c, k = 0;
    while (k < job.n_src - 1 && cc >= job.src_C[k]) {
      cc -= job.src_C[k];
      ++k;
    }
    const longlong2* sp = reinterpret_cast<const longlong2*>(job.src_sums[k] + ((int64_t)b * job.src_C[k] + cc) * 4);
    const longlong2 s = sp[0], q = sp[1];
    const long long sv[2] = {s.x, s.y}, qv[2] = {q.x, q.y};
    chan[2 * c] = fx_value(sv);
    chan[2 * c + 1] = fx_value(qv);
  }
  __syncthreads();
  for (int g = threadIdx.x; g < G; g += blockDim.x) {
    double a = 0.0, q = 0.0;
    for (int i = 0; i < cpg; ++i) {
      a += chan[2 * (g * cpg + i)];
      q += chan[2 * (g * cpg + i) + 1];
    }
    const double n = (double)job.HW * cpg, mean = a / n;
    double var = q / n - mean * mean;
    if (var < 0.0) var = 0.0;
    grp[2 * g] = mean;
    grp[2 * g + 1] = 1.0 / sqrt(var + (double)job.eps);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int g = c / cpg;
    const double sc = grp[2 * g + 1] * (double)job.gamma[c];
    const float fs = (float)sc, fh = (float)((double)job.beta[c] - grp[2 * g] * sc);
    if (kToShared) {
      ss_out[c] = fs;
      ss_out[C + c] = fh;
    } else {
      job.scale_shift[((int64_t)b * 2) * C + c] = fs;
      job.scale_shift[((int64_t)b * 2 + 1) * C + c] = fh;
    }
  }
  __syncthreads();
}

// grid B, block 128: a few hundred bytes of sums per image (the round-1 finalize read ~8 MB of per-CTA partials).
__global__ void __launch_bounds__(128) gn_finalize_kernel(const ditsep_gn_job job) {
  griddep_launch_dependents();
  griddep_wait();
  __shared__ double chan[2 * kGnMaxC];
  __shared__ double grp[2 * 64];
  gn_job_image<false>(job, blockIdx.x, chan, grp, nullptr);
}

// ------------------------------------------------------------------- normalise + SiLU
// grid (blocks per image, B); block = cv * ppb threads (cv = C/8 channel packs, ppb pixels in flight).  A thread
// keeps ONE channel pack for its whole life, so the per-(b,c) scale / shift sit in registers and the loop body is
// load 16 B (bf16) -> 8 fma + SiLU -> store 16 B, four pixels in flight per thread.
// kJob: the per-(b, c) scale / shift are formed by the CTA itself from the producers' fixed-point sums (the gn_finalize
// arithmetic, ~1 us per CTA) instead of being read from a buffer a separate launch wrote: 51 of the 63 gn_finalize
// launches of a score step disappear.
template <typename T, bool kJob>
__global__ void __launch_bounds__(256) gn_apply_kernel(const T* __restrict__ x, int64_t HW, int C, int ld_x,
                                                       const float* __restrict__ scale_shift, const ditsep_gn_job job,
                                                       int silu, T* __restrict__ y, int ld_y, int cv, int ppb) {
  griddep_launch_dependents();
  griddep_wait();
  const int b = blockIdx.y;
  const int v = threadIdx.x % cv, pp = threadIdx.x / cv;
  float sc[8], sh[8];
  if (kJob) {
    __shared__ double chan[2 * kGnMaxC];
    __shared__ double grp[2 * 64];
    __shared__ __align__(16) float ss_s[2 * kGnMaxC];
    gn_job_image<true>(job, b, chan, grp, ss_s);
    if (pp >= ppb) return;
    Pack8<float>::load(ss_s + v * 8, sc);
    Pack8<float>::load(ss_s + C + v * 8, sh);
  } else {
    if (pp >= ppb) return;
    Pack8<float>::load(scale_shift + ((int64_t)b * 2) * C + v * 8, sc);
    Pack8<float>::load(scale_shift + ((int64_t)b * 2 + 1) * C + v * 8, sh);
  }
  const T* xb = x + (int64_t)b * HW * ld_x + v * 8;
  T* yb = y + (int64_t)b * HW * ld_y + v * 8;
  const int64_t stride = (int64_t)gridDim.x * ppb;
  int64_t p = (int64_t)blockIdx.x * ppb + pp;
  for (; p + 3 * stride < HW; p += 4 * stride) {
    float f[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u) Pack8<T>::load(xb + (p + u * stride) * ld_x, f[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float t = fmaf(f[u][i], sc[i], sh[i]);
        f[u][i] = silu ? silu_f(t) : t;
      }
      Pack8<T>::store(yb + (p + u * stride) * ld_y, f[u]);
    }
  }
  for (; p < HW; p += stride) {
    float f[8];
    Pack8<T>::load(xb + p * ld_x, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float t = fmaf(f[i], sc[i], sh[i]);
      f[i] = silu ? silu_f(t) : t;
    }
    Pack8<T>::store(yb + p * ld_y, f);
  }
}

// ------------------------------------------------------------------- small images: statistics + normalise + SiLU in ONE launch
// At the 16x16 .. 4x4 levels an image is a few tens of KB (L1 / L2 resident) and the three-launch sequence
// partial sums -> finalize -> apply is pure launch latency (~100 tiny dependent launches per score step).  One CTA per
// image: pass 1 accumulates per-channel sums (thread = one 8-channel pack x a strided set of pixels, fp32), the CTA
// reduces them in shared memory (double from there on, as gn_finalize does), forms the per-channel scale / shift,
// pass 2 re-reads the image and applies them.  y may alias x.
template <typename T>
__global__ void __launch_bounds__(256) gn_small_kernel(const T* __restrict__ x, int64_t HW, int C, int ld_x, int G,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       float eps, int silu, T* __restrict__ y, int ld_y) {
  griddep_launch_dependents();
  griddep_wait();
  __shared__ float red[2][256 * 8];     // [sum | sum of squares][pixel lane][channel] (ppb * C = 2048 floats each)
  __shared__ double chan[2][256];
  __shared__ double grp[2][64];
  __shared__ float ss[2][256];
  const int b = blockIdx.x, cv = C / 8, ppb = 256 / cv;
  const int v = threadIdx.x % cv, pp = threadIdx.x / cv;
  const T* xb = x + (int64_t)b * HW * ld_x + v * 8;
  T* yb = y + (int64_t)b * HW * ld_y + v * 8;
  float s[8], q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = q[i] = 0.f;
  if (pp < ppb) {
    for (int64_t p = pp; p < HW; p += ppb) {
      float f[8];
      Pack8<T>::load(xb + p * ld_x, f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s[i] += f[i];
        q[i] = fmaf(f[i], f[i], q[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      red[0][pp * C + v * 8 + i] = s[i];
      red[1][pp * C + v * 8 + i] = q[i];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    double a = 0.0, d = 0.0;
    for (int j = 0; j < ppb; ++j) {
      a += (double)red[0][j * C + c];
      d += (double)red[1][j * C + c];
    }
    chan[0][c] = a;
    chan[1][c] = d;
  }
  __syncthreads();
  const int cpg = C / G;
  for (int g = threadIdx.x; g < G; g += blockDim.x) {
    double a = 0.0, d = 0.0;
    for (int i = 0; i < cpg; ++i) {
      a += chan[0][g * cpg + i];
      d += chan[1][g * cpg + i];
    }
    const double n = (double)HW * cpg, mean = a / n;
    double var = d / n - mean * mean;
    if (var < 0.0) var = 0.0;
    grp[0][g] = mean;
    grp[1][g] = 1.0 / sqrt(var + (double)eps);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int g = c / cpg;
    const double sc = grp[1][g] * (double)gamma[c];
    ss[0][c] = (float)sc;
    ss[1][c] = (float)((double)beta[c] - grp[0][g] * sc);
  }
  __syncthreads();
  if (pp < ppb) {
    float sc[8], sh[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      sc[i] = ss[0][v * 8 + i];
      sh[i] = ss[1][v * 8 + i];
    }
    for (int64_t p = pp; p < HW; p += ppb) {
      float f[8];
      Pack8<T>::load(xb + p * ld_x, f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float t = fmaf(f[i], sc[i], sh[i]);
        f[i] = silu ? silu_f(t) : t;
      }
      Pack8<T>::store(yb + p * ld_y, f);
    }
  }
}

// ------------------------------------------------------------------- FIR x2 resampling
// Separable [1,3,3,1]: down = taps (1,3,3,1)/8 per axis at input 2o-1..2o+2 (zero outside);
// up = polyphase: out[2i] = (x[i-1] + 3 x[i])/4, out[2i+1] = (3 x[i] + x[i+1])/4 per axis.
// If scale_shift != nullptr a second output gets FIR(SiLU(x*scale+shift)).
//
// Column-strip kernels: a thread owns one 8-channel pack of one output column (down) / input column (up)
// and walks a segment of rows.  Each input row is filtered horizontally ONCE per thread and its result is
// added into the (at most two) output rows it contributes to, so SiLU is evaluated 2x (down) / 3x (up) per
// input element instead of 4x / 9x for a one-thread-per-output-pixel stencil (that version was issue-bound at
// 9 % of HBM peak, r1e ncu); neighbouring threads re-read the overlapping pixels from L1.  Channel packs are
// the fastest thread index: a warp reads whole 128-byte pixel rows.
template <typename T, bool kDual>
__device__ __forceinline__ void fir_load_px(const T* __restrict__ p, bool ok, const float (&sc)[8], const float (&sh)[8],
                                            float (&f)[8], float (&h)[8]) {
  if (ok) {
    Pack8<T>::load(p, f);
    if (kDual) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        // bf16 tensors: one MUFU per element instead of two (the strip kernels are MUFU- and issue-bound, r1g ncu);
        // the 5e-4 * |x|/2 error of tanh.approx is below the bf16 rounding of the result.  fp32 tensors stay exact.
        const float t = fmaf(f[i], sc[i], sh[i]);
        h[i] = std::is_same<T, bf16>::value ? silu_tanh_f(t) : silu_f(t);
      }
    }
  } else {  // zero padding applies to the activated tensor too (the FIR pads AFTER GroupNorm + SiLU)
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = h[i] = 0.f;
  }
}

template <typename T, bool kDual>
__device__ __forceinline__ void fir_down_hrow(const T* __restrict__ xb, int iy, int H, int W, int ld_x, int ix0,
                                              const float (&sc)[8], const float (&sh)[8], float (&hx)[8], float (&hh)[8]) {
  // horizontal pass of input row iy at one output column: taps (1,3,3,1)/8 at ix0 .. ix0+3
#pragma unroll
  for (int i = 0; i < 8; ++i) hx[i] = hh[i] = 0.f;
  if (iy < 0 || iy >= H) return;
  const T* row = xb + (int64_t)iy * W * ld_x;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const int ix = ix0 + c;
    const float w = (c == 0 || c == 3) ? 0.125f : 0.375f;
    float f[8], g[8];
    fir_load_px<T, kDual>(row + (int64_t)ix * ld_x, ix >= 0 && ix < W, sc, sh, f, g);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      hx[i] = fmaf(w, f[i], hx[i]);
      if (kDual) hh[i] = fmaf(w, g[i], hh[i]);
    }
  }
}

// r1g ncu of the first strip version: 146 registers -> one CTA (8 warps) per SM, 12 % occupancy, latency-bound at
// 11 % of HBM peak with 46 instructions per input element.  Now: two CTAs per SM (launch bound), and the row loop
// advances TWO input rows per trip so the vertical taps are compile-time constants and the accumulator hand-over
// (row oy+1 becomes row oy) is a register rename instead of a select + copy.
template <typename T, bool kDual>
__global__ void __launch_bounds__(256, DITSEP_FIR_MINB) fir_down_kernel(const T* __restrict__ x, int H, int W, int C, int ld_x,
                                                          const float* __restrict__ scale_shift, T* __restrict__ out_x,
                                                          int ld_ox, T* __restrict__ out_h, int ld_oh, int seg_rows,
                                                          int n_seg, int64_t total) {
  griddep_launch_dependents();
  griddep_wait();
  const int cv = C / 8, Ho = H / 2, Wo = W / 2;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int v = (int)(idx % cv);
  int64_t r = idx / cv;
  const int ox = (int)(r % Wo);
  r /= Wo;
  const int seg = (int)(r % n_seg);
  const int b = (int)(r / n_seg);
  const int oy0 = seg * seg_rows, oy1 = min(oy0 + seg_rows, Ho);
  float sc[8], sh[8];
  if (kDual) {
    Pack8<float>::load(scale_shift + ((int64_t)b * 2) * C + v * 8, sc);
    Pack8<float>::load(scale_shift + ((int64_t)b * 2 + 1) * C + v * 8, sh);
  }
  const T* xb = x + (int64_t)b * H * W * ld_x + v * 8;
  const int ix0 = 2 * ox - 1;
  // output row oy = 1/8 r(2oy-1) + 3/8 r(2oy) + 3/8 r(2oy+1) + 1/8 r(2oy+2), r() = horizontally filtered input row.
  // carry = 1/8 r(2oy-1) + 3/8 r(2oy): the part of row oy that the previous trip already saw.
  float cx[8], chh[8];
  {
    float ax[8], ah[8], bx[8], bh[8];
    fir_down_hrow<T, kDual>(xb, 2 * oy0 - 1, H, W, ld_x, ix0, sc, sh, ax, ah);
    fir_down_hrow<T, kDual>(xb, 2 * oy0, H, W, ld_x, ix0, sc, sh, bx, bh);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      cx[i] = fmaf(0.375f, bx[i], 0.125f * ax[i]);
      chh[i] = fmaf(0.375f, bh[i], 0.125f * ah[i]);
    }
  }
  for (int oy = oy0; oy < oy1; ++oy) {
    float ax[8], ah[8], bx[8], bh[8];  // rows 2oy+1 and 2oy+2
    fir_down_hrow<T, kDual>(xb, 2 * oy + 1, H, W, ld_x, ix0, sc, sh, ax, ah);
    fir_down_hrow<T, kDual>(xb, 2 * oy + 2, H, W, ld_x, ix0, sc, sh, bx, bh);
    float ox_[8], oh_[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      ox_[i] = fmaf(0.125f, bx[i], fmaf(0.375f, ax[i], cx[i]));
      cx[i] = fmaf(0.375f, bx[i], 0.125f * ax[i]);  // rows 2oy+1, 2oy+2 are taps 0, 1 of output row oy+1
      if (kDual) {
        oh_[i] = fmaf(0.125f, bh[i], fmaf(0.375f, ah[i], chh[i]));
        chh[i] = fmaf(0.375f, bh[i], 0.125f * ah[i]);
      }
    }
    const int64_t opix = ((int64_t)b * Ho + oy) * Wo + ox;
    if (out_x) Pack8<T>::store(out_x + opix * ld_ox + v * 8, ox_);
    if (kDual) Pack8<T>::store(out_h + opix * ld_oh + v * 8, oh_);
  }
}

// (r1o, measured and removed: a "pair" version in which a thread owns input pixels (2m, 2m+1), activates each element
// once and exchanges the two neighbour pixels through shared memory -- half the SiLU evaluations of the strip kernel
// above -- was correct but SLOWER, 340 vs 304 us at 256x256x64: one __syncthreads per input row serialises the load
// latency.  The strip kernel's limit is therefore not its redundant arithmetic; see DESIGN.md.)
template <typename T, bool kDual>
__global__ void __launch_bounds__(256) fir_up_kernel(const T* __restrict__ x, int H, int W, int C, int ld_x,
                                                     const float* __restrict__ scale_shift, T* __restrict__ out_x,
                                                     int ld_ox, T* __restrict__ out_h, int ld_oh, int seg_rows, int n_seg,
                                                     int64_t total) {
  griddep_launch_dependents();
  griddep_wait();
  const int cv = C / 8, Ho = 2 * H, Wo = 2 * W;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int v = (int)(idx % cv);
  int64_t r = idx / cv;
  const int ix = (int)(r % W);
  r /= W;
  const int seg = (int)(r % n_seg);
  const int b = (int)(r / n_seg);
  const int r0 = seg * seg_rows, r1 = min(r0 + seg_rows, H);
  float sc[8], sh[8];
  if (kDual) {
    Pack8<float>::load(scale_shift + ((int64_t)b * 2) * C + v * 8, sc);
    Pack8<float>::load(scale_shift + ((int64_t)b * 2 + 1) * C + v * 8, sh);
  }
  const T* xb = x + (int64_t)b * H * W * ld_x + v * 8;
  // horizontally filtered previous input row: [even output column 2ix | odd output column 2ix+1]
  float pxe[8], pxo[8], phe[8], pho[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) pxe[i] = pxo[i] = phe[i] = pho[i] = 0.f;
  for (int iy = r0 - 1; iy <= r1; ++iy) {
    float cxe[8], cxo[8], che[8], cho[8];
    if (iy >= 0 && iy < H) {
      const T* row = xb + (int64_t)iy * W * ld_x;
      float f[3][8], g[3][8];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const int xx = ix + c - 1;
        fir_load_px<T, kDual>(row + (int64_t)xx * ld_x, xx >= 0 && xx < W, sc, sh, f[c], g[c]);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        cxe[i] = fmaf(0.75f, f[1][i], 0.25f * f[0][i]);
        cxo[i] = fmaf(0.75f, f[1][i], 0.25f * f[2][i]);
        if (kDual) {
          che[i] = fmaf(0.75f, g[1][i], 0.25f * g[0][i]);
          cho[i] = fmaf(0.75f, g[1][i], 0.25f * g[2][i]);
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) cxe[i] = cxo[i] = che[i] = cho[i] = 0.f;
    }
    // on arrival of filtered row iy: output row 2(iy-1)+1 = (3 prev + cur)/4, output row 2iy = (prev + 3 cur)/4
    float o0[8], o1[8];
    if (iy - 1 >= r0) {
      const int64_t opix = ((int64_t)b * Ho + 2 * (iy - 1) + 1) * Wo + 2 * ix;
      if (out_x) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          o0[i] = fmaf(0.75f, pxe[i], 0.25f * cxe[i]);
          o1[i] = fmaf(0.75f, pxo[i], 0.25f * cxo[i]);
        }
        Pack8<T>::store(out_x + opix * ld_ox + v * 8, o0);
        Pack8<T>::store(out_x + (opix + 1) * ld_ox + v * 8, o1);
      }
      if (kDual) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          o0[i] = fmaf(0.75f, phe[i], 0.25f * che[i]);
          o1[i] = fmaf(0.75f, pho[i], 0.25f * cho[i]);
        }
        Pack8<T>::store(out_h + opix * ld_oh + v * 8, o0);
        Pack8<T>::store(out_h + (opix + 1) * ld_oh + v * 8, o1);
      }
    }
    if (iy >= r0 && iy < r1) {
      const int64_t opix = ((int64_t)b * Ho + 2 * iy) * Wo + 2 * ix;
      if (out_x) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          o0[i] = fmaf(0.75f, cxe[i], 0.25f * pxe[i]);
          o1[i] = fmaf(0.75f, cxo[i], 0.25f * pxo[i]);
        }
        Pack8<T>::store(out_x + opix * ld_ox + v * 8, o0);
        Pack8<T>::store(out_x + (opix + 1) * ld_ox + v * 8, o1);
      }
      if (kDual) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          o0[i] = fmaf(0.75f, che[i], 0.25f * phe[i]);
          o1[i] = fmaf(0.75f, cho[i], 0.25f * pho[i]);
        }
        Pack8<T>::store(out_h + opix * ld_oh + v * 8, o0);
        Pack8<T>::store(out_h + (opix + 1) * ld_oh + v * 8, o1);
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      pxe[i] = cxe[i];
      pxo[i] = cxo[i];
      phe[i] = che[i];
      pho[i] = cho[i];
    }
  }
}

// Narrow tensors (the 2(n+1)-channel input pyramid and the progressive output pyramid):
// scalar channel loop, optional accumulate-into (pyramid = up(pyramid) + conv(...)).
template <typename T>
__global__ void __launch_bounds__(256) fir_narrow_kernel(const T* __restrict__ x, int H, int W, int C, int ld_x, int up,
                                                         T* __restrict__ out, int ld_o, int64_t total) {
  const int Ho = up ? 2 * H : H / 2, Wo = up ? 2 * W : W / 2;
  const float kd[4] = {0.125f, 0.375f, 0.375f, 0.125f};
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % C);
    int64_t r = idx / C;
    const int ox = (int)(r % Wo);
    r /= Wo;
    const int oy = (int)(r % Ho);
    const int b = (int)(r / Ho);
    float acc = 0.f;
    if (!up) {
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const int iy = 2 * oy + a - 1;
        if (iy < 0 || iy >= H) continue;
#pragma unroll
        for (int d = 0; d < 4; ++d) {
          const int ix = 2 * ox + d - 1;
          if (ix < 0 || ix >= W) continue;
          acc = fmaf(kd[a] * kd[d], to_f<T>(x[(((int64_t)b * H + iy) * W + ix) * ld_x + c]), acc);
        }
      }
    } else {
      const int iy = oy >> 1, ix = ox >> 1;
      const int y2 = (oy & 1) ? iy + 1 : iy - 1, x2 = (ox & 1) ? ix + 1 : ix - 1;
      const int ys[2] = {iy, y2}, xs[2] = {ix, x2};
      const float ws[2] = {0.75f, 0.25f};
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        if (ys[a] < 0 || ys[a] >= H) continue;
#pragma unroll
        for (int d = 0; d < 2; ++d) {
          if (xs[d] < 0 || xs[d] >= W) continue;
          acc = fmaf(ws[a] * ws[d], to_f<T>(x[(((int64_t)b * H + ys[a]) * W + xs[d]) * ld_x + c]), acc);
        }
      }
    }
    out[(((int64_t)b * Ho + oy) * Wo + ox) * ld_o + c] = from_f<T>(acc);
  }
}

// Narrow fp32 tensors with an even channel count (the 6-channel pyramids): one thread per OUTPUT PIXEL, all channels
// as float2 loads, so the index arithmetic (three 64-bit divisions per element in the scalar kernel above: 60-73 us
// per launch at batch 32 for 12-50 MB of data, r1x) is paid once per pixel.
template <int C>
__global__ void __launch_bounds__(256) fir_narrow_px_kernel(const float* __restrict__ x, int H, int W, int ld_x, int up,
                                                            float* __restrict__ out, int ld_o, int64_t npix_out) {
  griddep_launch_dependents();
  griddep_wait();
  static_assert(C % 2 == 0, "float2 path");
  const int Ho = up ? 2 * H : H / 2, Wo = up ? 2 * W : W / 2;
  const float kd[4] = {0.125f, 0.375f, 0.375f, 0.125f};
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < npix_out;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int ox = (int)(idx % Wo);
    int64_t r = idx / Wo;
    const int oy = (int)(r % Ho);
    const int b = (int)(r / Ho);
    float acc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = 0.f;
    const float* xb = x + (int64_t)b * H * W * ld_x;
    if (!up) {
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const int iy = 2 * oy + a - 1;
        if (iy < 0 || iy >= H) continue;
#pragma unroll
        for (int d = 0; d < 4; ++d) {
          const int ix = 2 * ox + d - 1;
          if (ix < 0 || ix >= W) continue;
          const float wgt = kd[a] * kd[d];
          const float2* px = reinterpret_cast<const float2*>(xb + ((int64_t)iy * W + ix) * ld_x);
#pragma unroll
          for (int c = 0; c < C / 2; ++c) {
            const float2 v = __ldg(px + c);
            acc[2 * c] = fmaf(wgt, v.x, acc[2 * c]);
            acc[2 * c + 1] = fmaf(wgt, v.y, acc[2 * c + 1]);
          }
        }
      }
    } else {
      const int iy = oy >> 1, ix = ox >> 1;
      const int ys[2] = {iy, (oy & 1) ? iy + 1 : iy - 1}, xs[2] = {ix, (ox & 1) ? ix + 1 : ix - 1};
      const float ws[2] = {0.75f, 0.25f};
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        if (ys[a] < 0 || ys[a] >= H) continue;
#pragma unroll
        for (int d = 0; d < 2; ++d) {
          if (xs[d] < 0 || xs[d] >= W) continue;
          const float wgt = ws[a] * ws[d];
          const float2* px = reinterpret_cast<const float2*>(xb + ((int64_t)ys[a] * W + xs[d]) * ld_x);
#pragma unroll
          for (int c = 0; c < C / 2; ++c) {
            const float2 v = __ldg(px + c);
            acc[2 * c] = fmaf(wgt, v.x, acc[2 * c]);
            acc[2 * c + 1] = fmaf(wgt, v.y, acc[2 * c + 1]);
          }
        }
      }
    }
    float2* po = reinterpret_cast<float2*>(out + idx * ld_o);
#pragma unroll
    for (int c = 0; c < C / 2; ++c) po[c] = make_float2(acc[2 * c], acc[2 * c + 1]);
  }
}

// ------------------------------------------------------------------- hi/lo split of a narrow fp32 tensor
// The network's first layer has 2(n+1) = 6 input channels: too narrow for the tensor-core kernel (K blocks of 64)
// and, on CUDA cores, 1.06 ms per score step at batch 32 (r1g ncu).  This kernel widens it to a 64-channel bf16
// tensor the tcgen05 kernel accepts WITHOUT giving up fp32 input accuracy: channels [0,C) = hi = bf16(x),
// [C,2C) = lo = bf16(x - hi), [2C,3C) = hi again, the rest zero; against filter channels [wh | wh | wl]
// (w = wh + wl) the tensor core computes xh*wh + xl*wh + xh*wl = x*w to ~2^-16.
// One thread = one pixel: C fp32 loads, 3C bf16 values in ceil(3C/8) 16-byte stores.  Channels >= 8*ceil(3C/8) are
// NOT written: the caller zeroes the buffer once (engine: torch.zeros at plan build) and they stay zero.
template <int C>
__global__ void __launch_bounds__(256) split_hilo_kernel(const float* __restrict__ x, int64_t npix, bf16* __restrict__ out,
                                                         int ld) {
  griddep_launch_dependents();
  constexpr int kVec = (3 * C + 7) / 8;
  for (int64_t pix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += (int64_t)gridDim.x * blockDim.x) {
    float xv[C], f[kVec * 8];
#pragma unroll
    for (int c = 0; c < C; ++c) xv[c] = __ldg(x + pix * C + c);
#pragma unroll
    for (int j = 0; j < kVec * 8; ++j) f[j] = 0.f;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const float hi = __bfloat162float(__float2bfloat16_rn(xv[c]));
      f[c] = hi;
      f[C + c] = xv[c] - hi;
      f[2 * C + c] = hi;
    }
#pragma unroll
    for (int v = 0; v < kVec; ++v) {
      float t[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) t[i] = f[v * 8 + i];
      Pack8<bf16>::store(out + pix * ld + v * 8, t);
    }
  }
}

// ------------------------------------------------------------------- time embedding
// grid B, block 1024.  smem: feat[2nf] | hid[4nf].  A warp computes four output rows at a time so that their
// weight loads are in flight together (the layer is a chain of global-load latencies, not of flops); the
// summation order inside a row is lane-strided then warp_sum, whatever the row grouping.
template <int ROWS>
__device__ __forceinline__ void temb_rows(const float* __restrict__ w, const float* __restrict__ x, int K, int o0,
                                          int ostride, int n_out, int lane, float (&acc)[ROWS]) {
#pragma unroll
  for (int r = 0; r < ROWS; ++r) acc[r] = 0.f;
  for (int k = lane; k < K; k += 32) {
    const float xv = x[k];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      const int o = o0 + r * ostride;
      if (o < n_out) acc[r] = fmaf(w[(int64_t)o * K + k], xv, acc[r]);
    }
  }
#pragma unroll
  for (int r = 0; r < ROWS; ++r) acc[r] = warp_sum(acc[r]);
}

__global__ void __launch_bounds__(1024) temb_mlp_kernel(const float* __restrict__ t, const float* __restrict__ gfp_w,
                                                        const float* __restrict__ w1, const float* __restrict__ b1,
                                                        const float* __restrict__ w2, const float* __restrict__ b2,
                                                        float* __restrict__ act_temb, int nf) {
  extern __shared__ float sm[];
  float* feat = sm;
  float* hid = sm + 2 * nf;
  const int b = blockIdx.x;
  const float lt = logf(t[b]);
  for (int i = threadIdx.x; i < nf; i += blockDim.x) {
    // x_proj = log(t) * W * 2 * pi  (layerspp.py:45), evaluated like torch: ((lt*W)*2)*pi in fp32
    float xp = lt * gfp_w[i] * 2.0f * 3.14159265358979323846f;
    float s, c;
    sincosf(xp, &s, &c);
    feat[i] = s;
    feat[nf + i] = c;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  constexpr int ROWS = 4;
  float acc[ROWS];
  for (int o0 = warp; o0 < 4 * nf; o0 += nw * ROWS) {
    temb_rows<ROWS>(w1, feat, 2 * nf, o0, nw, 4 * nf, lane, acc);
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < ROWS; ++r) {
        const int o = o0 + r * nw;
        if (o < 4 * nf) {
          const float h = acc[r] + b1[o];
          hid[o] = h / (1.0f + expf(-h));
        }
      }
    }
  }
  __syncthreads();
  for (int o0 = warp; o0 < 4 * nf; o0 += nw * ROWS) {
    temb_rows<ROWS>(w2, hid, 4 * nf, o0, nw, 4 * nf, lane, acc);
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < ROWS; ++r) {
        const int o = o0 + r * nw;
        if (o < 4 * nf) {
          const float h = acc[r] + b2[o];
          act_temb[(int64_t)b * 4 * nf + o] = h / (1.0f + expf(-h));
        }
      }
    }
  }
}

// One warp per output row r (all res-blocks' Dense_0 rows concatenated).  The CTA stages `bchunk` rows of the
// activated embedding in shared memory and every warp keeps its weight row in registers (K <= 512), so the
// per-mixture work is shared-memory loads only.
__global__ void __launch_bounds__(256) temb_dense_kernel(const float* __restrict__ act_temb,
                                                         const float* __restrict__ dense_w,
                                                         const float* __restrict__ dense_b, float* __restrict__ out,
                                                         int B, int K, int R, int bchunk) {
  extern __shared__ float sa[];  // [bchunk][K]
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const bool live = row < R;
  constexpr int KR = 16;
  float wv[KR];
#pragma unroll
  for (int j = 0; j < KR; ++j) wv[j] = (live && lane + 32 * j < K) ? dense_w[(int64_t)row * K + lane + 32 * j] : 0.f;
  const float bias = live ? dense_b[row] : 0.f;
  for (int b0 = 0; b0 < B; b0 += bchunk) {
    const int nb = min(bchunk, B - b0);
    __syncthreads();
    for (int i = threadIdx.x; i < nb * K; i += blockDim.x) sa[i] = act_temb[(int64_t)b0 * K + i];
    __syncthreads();
    if (!live) continue;
    for (int b = 0; b < nb; ++b) {
      const float* a = sa + b * K;
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < KR; ++j)
        if (lane + 32 * j < K) acc = fmaf(wv[j], a[lane + 32 * j], acc);
      acc = warp_sum(acc);
      if (lane == 0) out[(int64_t)(b0 + b) * R + row] = acc + bias;
    }
  }
}

// ------------------------------------------------------------------- output layer
// out[b,p,co] = sum_ci w[co,ci] * (x[b,p,ci] / t[b]) + bias[co]   (ncsnpp.py:473-477)
template <typename T>
__global__ void __launch_bounds__(256) output_layer_kernel(const T* __restrict__ x, int64_t HW, int Cin, int ld_x,
                                                           const float* __restrict__ t, int scale_by_sigma,
                                                           const float* __restrict__ w, const float* __restrict__ bias,
                                                           int Cout, float* __restrict__ out, int64_t total) {
  __shared__ float ws[16 * 16 + 16];
  for (int i = threadIdx.x; i < Cout * Cin; i += blockDim.x) ws[i] = w[i];
  for (int i = threadIdx.x; i < Cout; i += blockDim.x) ws[256 + i] = bias[i];
  __syncthreads();
  for (int64_t pix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; pix < total;
       pix += (int64_t)gridDim.x * blockDim.x) {
    const int b = (int)(pix / HW);
    const float inv = scale_by_sigma ? 1.0f / t[b] : 1.0f;
    float v[16];
    for (int ci = 0; ci < Cin; ++ci) v[ci] = to_f<T>(x[pix * ld_x + ci]) * inv;
    for (int co = 0; co < Cout; ++co) {
      float acc = ws[256 + co];
      for (int ci = 0; ci < Cin; ++ci) acc = fmaf(ws[co * Cin + ci], v[ci], acc);
      out[pix * Cout + co] = acc;
    }
  }
}

static int grid_for(int64_t total, int threads = 256) {
  int64_t g = (total + threads - 1) / threads;
  int64_t cap = (int64_t)kNumSMs * 16;
  return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

}  // namespace ditsep

using namespace ditsep;

// pixel splits of the explicit statistics pass: a function of the image shape only (batch invariance)
static int gn_nsplit(int64_t HW, int C) {
  const int ppb = 256 / (C / 8) > 0 ? 256 / (C / 8) : 1;
  int64_t n = HW / ((int64_t)ppb * 8);
  if (n < 1) n = 1;
  if (n > 64) n = 64;
  return (int)n;
}

extern "C" int ditsep_gn_sums(const void* x, int dtype, int B, int64_t HW, int C, int ld, long long* sums, void* stream) {
  DITSEP_REQUIRE(x && sums, "gn_sums: null pointer");
  DITSEP_REQUIRE(B > 0 && HW > 0 && C > 0 && C % 8 == 0 && C <= 2048 && ld >= C, "gn_sums: bad sizes B=%d HW=%lld C=%d ld=%d",
                 B, (long long)HW, C, ld);
  const int cv = C / 8;
  int ppb = 256 / cv;
  if (ppb < 1) ppb = 1;
  DITSEP_REQUIRE(cv * ppb <= 256 && cv <= 256, "gn_sums: C=%d too wide", C);
  const int nsplit = gn_nsplit(HW, C);
  dim3 grid(nsplit, B), block(cv * ppb);
  const size_t smem = (size_t)ppb * C * 2 * sizeof(float);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == DITSEP_F32)
    DITSEP_CUDA(launch_pdl<2>(gn_sums_kernel<float>, grid, block, smem, st, (const float*)x, HW, C, ld, cv, ppb, sums));
  else if (dtype == DITSEP_BF16)
    DITSEP_CUDA(launch_pdl<2>(gn_sums_kernel<bf16>, grid, block, smem, st, (const bf16*)x, HW, C, ld, cv, ppb, sums));
  else
    DITSEP_REQUIRE(false, "gn_sums: unknown dtype %d", dtype);
  return check_launch("gn_sums_kernel");
}

static int check_gn_job(const ditsep_gn_job* j, const char* who, bool need_out = true) {
  DITSEP_REQUIRE(j != nullptr, "%s: null job", who);
  DITSEP_REQUIRE(j->n_src >= 1 && j->n_src <= 4, "%s: 1..4 statistics sources (got %d)", who, j->n_src);
  int ctot = 0;
  for (int i = 0; i < j->n_src; ++i) {
    DITSEP_REQUIRE(j->src_sums[i] != nullptr && j->src_C[i] > 0, "%s: bad statistics source %d", who, i);
    ctot += j->src_C[i];
  }
  DITSEP_REQUIRE(ctot == j->C, "%s: sources cover %d channels, GroupNorm has %d", who, ctot, j->C);
  DITSEP_REQUIRE(j->G > 0 && j->G <= 64 && j->C % j->G == 0, "%s: C=%d not divisible into G=%d groups (G <= 64)", who, j->C,
                 j->G);
  DITSEP_REQUIRE(j->C <= kGnMaxC, "%s: C=%d exceeds %d", who, j->C, kGnMaxC);
  DITSEP_REQUIRE(j->HW > 0 && j->gamma && j->beta && (j->scale_shift || !need_out), "%s: null gamma / beta / scale_shift", who);
  return 0;
}

extern "C" int ditsep_gn_finalize(const ditsep_gn_job* host_job, int B, void* stream) {
  if (int rc = check_gn_job(host_job, "gn_finalize")) return rc;
  DITSEP_REQUIRE(B > 0, "gn_finalize: B=%d", B);
  DITSEP_CUDA(launch_pdl(gn_finalize_kernel, dim3(B), dim3(128), 0, (cudaStream_t)stream, *host_job));
  return check_launch("gn_finalize_kernel");
}

extern "C" int ditsep_gn_small(const void* x, int dtype, int B, int64_t HW, int C, int ld_x, int G, const float* gamma,
                               const float* beta, float eps, int silu, void* y, int ld_y, void* stream) {
  DITSEP_REQUIRE(x && y && gamma && beta, "gn_small: null pointer");
  DITSEP_REQUIRE(C % 8 == 0 && ld_x % 8 == 0 && ld_y % 8 == 0 && C <= 256 && 256 % (C / 8) == 0,
                 "gn_small: C=%d must be a multiple of 8 dividing 2048, strides multiples of 8", C);
  DITSEP_REQUIRE(G > 0 && G <= 64 && C % G == 0, "gn_small: G=%d must divide C=%d (G <= 64)", G, C);
  DITSEP_REQUIRE(B > 0 && HW > 0 && HW <= 4096, "gn_small: meant for small images (HW=%lld)", (long long)HW);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == DITSEP_F32)
    DITSEP_CUDA(launch_pdl(gn_small_kernel<float>, dim3(B), dim3(256), 0, st, (const float*)x, HW, C, ld_x, G, gamma, beta, eps,
                           silu, (float*)y, ld_y));
  else if (dtype == DITSEP_BF16)
    DITSEP_CUDA(launch_pdl(gn_small_kernel<bf16>, dim3(B), dim3(256), 0, st, (const bf16*)x, HW, C, ld_x, G, gamma, beta, eps,
                           silu, (bf16*)y, ld_y));
  else
    DITSEP_REQUIRE(false, "gn_small: unknown dtype %d", dtype);
  return check_launch("gn_small_kernel");
}

static int launch_gn_apply(const void* x, int dtype, int B, int64_t HW, int C, int ld_x, const float* scale_shift,
                           const ditsep_gn_job* job, int silu, void* y, int ld_y, void* stream) {
  DITSEP_REQUIRE(C % 8 == 0 && ld_x % 8 == 0 && ld_y % 8 == 0, "gn_apply: C/ld must be multiples of 8");
  const int cv = C / 8;
  DITSEP_REQUIRE(cv <= 256, "gn_apply: C=%d too wide", C);
  const int ppb = 256 / cv;
  // enough blocks to fill the machine a few times over, but at least 4 pixels per thread where the image allows
  int64_t gx = (HW + (int64_t)ppb * 4 - 1) / ((int64_t)ppb * 4);
  const int64_t cap = ((int64_t)kNumSMs * 16 + B - 1) / B;
  if (gx > cap) gx = cap;
  if (gx < 1) gx = 1;
  dim3 grid((unsigned)gx, (unsigned)B);
  cudaStream_t st = (cudaStream_t)stream;
  const ditsep_gn_job none = {};
#define DITSEP_GN_APPLY(T)                                                                                              \
  do {                                                                                                                  \
    if (job)                                                                                                            \
      DITSEP_CUDA(launch_pdl(gn_apply_kernel<T, true>, grid, dim3(256), 0, st, (const T*)x, HW, C, ld_x, scale_shift, *job, \
                             silu, (T*)y, ld_y, cv, ppb));                                                              \
    else                                                                                                                \
      DITSEP_CUDA(launch_pdl(gn_apply_kernel<T, false>, grid, dim3(256), 0, st, (const T*)x, HW, C, ld_x, scale_shift, none, \
                             silu, (T*)y, ld_y, cv, ppb));                                                              \
  } while (0)
  if (dtype == DITSEP_F32)
    DITSEP_GN_APPLY(float);
  else if (dtype == DITSEP_BF16)
    DITSEP_GN_APPLY(bf16);
  else
    DITSEP_REQUIRE(false, "gn_apply: unknown dtype %d", dtype);
#undef DITSEP_GN_APPLY
  return check_launch("gn_apply_kernel");
}

extern "C" int ditsep_gn_apply(const void* x, int dtype, int B, int64_t HW, int C, int ld_x, const float* scale_shift,
                               int silu, void* y, int ld_y, void* stream) {
  DITSEP_REQUIRE(x && y && scale_shift, "gn_apply: null pointer");
  return launch_gn_apply(x, dtype, B, HW, C, ld_x, scale_shift, nullptr, silu, y, ld_y, stream);
}

extern "C" int ditsep_gn_apply_job(const void* x, int dtype, int B, int64_t HW, const ditsep_gn_job* host_job, int ld_x,
                                   int silu, void* y, int ld_y, void* stream) {
  DITSEP_REQUIRE(x && y, "gn_apply_job: null pointer");
  if (int rc = check_gn_job(host_job, "gn_apply_job", false)) return rc;
  DITSEP_REQUIRE(host_job->HW == HW, "gn_apply_job: the job normalises %lld pixels per image, the tensor has %lld",
                 (long long)host_job->HW, (long long)HW);
  return launch_gn_apply(x, dtype, B, HW, host_job->C, ld_x, nullptr, host_job, silu, y, ld_y, stream);
}

template <typename T>
static int fir_dispatch(const T* x, int B, int H, int W, int C, int ld_x, int up, const float* ss, T* out_x, int ld_ox,
                        T* out_h, int ld_oh, cudaStream_t st) {
  if (C % 8 != 0 || ld_x % 8 != 0) {
    DITSEP_REQUIRE(ss == nullptr && out_h == nullptr && out_x != nullptr,
                   "fir_resample: narrow tensors (C=%d) support the plain path only", C);
    const int64_t total = (int64_t)B * (up ? 4 : 1) * H * W / (up ? 1 : 4) * C;
    if constexpr (std::is_same<T, float>::value) {
      if ((C == 6 || C == 4 || C == 8) && ld_x % 2 == 0 && ld_ox % 2 == 0 && (reinterpret_cast<uintptr_t>(x) & 7) == 0 &&
          (reinterpret_cast<uintptr_t>(out_x) & 7) == 0) {
        const int64_t npix = total / C;
        if (C == 6) DITSEP_CUDA(launch_pdl<2>(fir_narrow_px_kernel<6>, dim3(grid_for(npix)), dim3(256), 0, st, x, H, W, ld_x, up, out_x, ld_ox, npix));
        else if (C == 4) DITSEP_CUDA(launch_pdl<2>(fir_narrow_px_kernel<4>, dim3(grid_for(npix)), dim3(256), 0, st, x, H, W, ld_x, up, out_x, ld_ox, npix));
        else DITSEP_CUDA(launch_pdl<2>(fir_narrow_px_kernel<8>, dim3(grid_for(npix)), dim3(256), 0, st, x, H, W, ld_x, up, out_x, ld_ox, npix));
        return check_launch("fir_narrow_px_kernel");
      }
    }
    fir_narrow_kernel<T><<<grid_for(total), 256, 0, st>>>(x, H, W, C, ld_x, up, out_x, ld_ox, total);
    return check_launch("fir_narrow_kernel");
  }
  DITSEP_REQUIRE((out_x == nullptr || ld_ox % 8 == 0) && (out_h == nullptr || ld_oh % 8 == 0),
                 "fir_resample: output strides must be multiples of 8");
  const int cv = C / 8;
  // rows per strip: long strips amortise the 2-row warm-up, but keep >= ~8 threads per SM lane-slot in flight
  const int cols = up ? W : W / 2, rows = up ? H : H / 2;
  int seg_rows = rows;
  while (seg_rows > 4 && (int64_t)B * cols * cv * ceil_div(rows, seg_rows) < (int64_t)kNumSMs * 2048) seg_rows = (seg_rows + 1) / 2;
  const int n_seg = ceil_div(rows, seg_rows);
  const int64_t total = (int64_t)B * n_seg * cols * cv;
  const unsigned grid = (unsigned)((total + 255) / 256);
  if (up) {
    if (ss)
      DITSEP_CUDA(launch_pdl<2>(fir_up_kernel<T, true>, dim3(grid), dim3(256), 0, st, x, H, W, C, ld_x, ss, out_x, ld_ox, out_h, ld_oh, seg_rows, n_seg, total));
    else
      DITSEP_CUDA(launch_pdl<2>(fir_up_kernel<T, false>, dim3(grid), dim3(256), 0, st, x, H, W, C, ld_x, ss, out_x, ld_ox, out_h, ld_oh, seg_rows, n_seg, total));
    return check_launch("fir_up_kernel");
  }
  if (ss)
    DITSEP_CUDA(launch_pdl<2>(fir_down_kernel<T, true>, dim3(grid), dim3(256), 0, st, x, H, W, C, ld_x, ss, out_x, ld_ox, out_h, ld_oh, seg_rows, n_seg, total));
  else
    DITSEP_CUDA(launch_pdl<2>(fir_down_kernel<T, false>, dim3(grid), dim3(256), 0, st, x, H, W, C, ld_x, ss, out_x, ld_ox, out_h, ld_oh, seg_rows, n_seg, total));
  return check_launch("fir_down_kernel");
}

extern "C" int ditsep_fir_resample(const void* x, int dtype, int B, int H, int W, int C, int ld_x, int up,
                                   const float* scale_shift, void* out_x, int ld_ox, void* out_h, int ld_oh,
                                   void* stream) {
  DITSEP_REQUIRE(x && (out_x || out_h), "fir_resample: null pointer");
  DITSEP_REQUIRE(B > 0 && H > 0 && W > 0 && C > 0, "fir_resample: empty tensor");
  DITSEP_REQUIRE(up || (H % 2 == 0 && W % 2 == 0), "fir_resample: down needs even H, W (got %dx%d)", H, W);
  DITSEP_REQUIRE((scale_shift != nullptr) == (out_h != nullptr), "fir_resample: out_h requires scale_shift");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == DITSEP_F32)
    return fir_dispatch<float>((const float*)x, B, H, W, C, ld_x, up, scale_shift, (float*)out_x, ld_ox, (float*)out_h,
                               ld_oh, st);
  if (dtype == DITSEP_BF16)
    return fir_dispatch<bf16>((const bf16*)x, B, H, W, C, ld_x, up, scale_shift, (bf16*)out_x, ld_ox, (bf16*)out_h,
                              ld_oh, st);
  DITSEP_REQUIRE(false, "fir_resample: unknown dtype %d", dtype);
}

extern "C" int ditsep_split_hilo(const float* x, int64_t npix, int C, void* out_bf16, int ld, void* stream) {
  DITSEP_REQUIRE(x && out_bf16, "split_hilo: null pointer");
  DITSEP_REQUIRE(npix > 0 && C > 0 && ld % 8 == 0 && 3 * C <= ld, "split_hilo: need 3*C <= ld, ld %% 8 == 0 (C=%d ld=%d)", C, ld);
  DITSEP_REQUIRE((reinterpret_cast<uintptr_t>(out_bf16) & 15) == 0, "split_hilo: output must be 16-byte aligned");
  DITSEP_REQUIRE(C == 6 || C == 8 || C == 3 || C == 4, "split_hilo: C=%d (built for 3, 4, 6, 8 input channels)", C);
  const int grid = grid_for(npix);
  cudaStream_t st = (cudaStream_t)stream;
  if (C == 6) split_hilo_kernel<6><<<grid, 256, 0, st>>>(x, npix, (bf16*)out_bf16, ld);
  else if (C == 8) split_hilo_kernel<8><<<grid, 256, 0, st>>>(x, npix, (bf16*)out_bf16, ld);
  else if (C == 3) split_hilo_kernel<3><<<grid, 256, 0, st>>>(x, npix, (bf16*)out_bf16, ld);
  else split_hilo_kernel<4><<<grid, 256, 0, st>>>(x, npix, (bf16*)out_bf16, ld);
  return check_launch("split_hilo_kernel");
}

extern "C" int ditsep_temb_mlp(const float* t, const float* gfp_w, const float* w1, const float* b1, const float* w2,
                               const float* b2, float* act_temb, int B, int nf, void* stream) {
  DITSEP_REQUIRE(t && gfp_w && w1 && b1 && w2 && b2 && act_temb, "temb_mlp: null pointer");
  DITSEP_REQUIRE(B > 0 && nf > 0 && nf <= 1024, "temb_mlp: bad B=%d nf=%d", B, nf);
  temb_mlp_kernel<<<B, 1024, (size_t)6 * nf * sizeof(float), (cudaStream_t)stream>>>(t, gfp_w, w1, b1, w2, b2, act_temb,
                                                                                    nf);
  return check_launch("temb_mlp_kernel");
}

extern "C" int ditsep_temb_dense(const float* act_temb, const float* dense_w, const float* dense_b, float* out, int B,
                                 int K, int R, void* stream) {
  DITSEP_REQUIRE(act_temb && dense_w && dense_b && out, "temb_dense: null pointer");
  DITSEP_REQUIRE(B > 0 && K > 0 && K <= 512 && R > 0, "temb_dense: bad sizes (K <= 512)");
  const int bchunk = std::max(1, std::min(B, 12288 / K));  // at most 48 KB of shared memory
  temb_dense_kernel<<<ceil_div((int64_t)R * 32, 256), 256, (size_t)bchunk * K * sizeof(float), (cudaStream_t)stream>>>(
      act_temb, dense_w, dense_b, out, B, K, R, bchunk);
  return check_launch("temb_dense_kernel");
}

extern "C" int ditsep_output_layer(const void* x, int dtype, int B, int64_t HW, int Cin, int ld_x, const float* t,
                                   int scale_by_sigma, const float* w, const float* bias, int Cout, float* out,
                                   void* stream) {
  DITSEP_REQUIRE(x && t && w && bias && out, "output_layer: null pointer");
  DITSEP_REQUIRE(Cin >= 1 && Cin <= 16 && Cout >= 1 && Cout <= 16, "output_layer: Cin=%d Cout=%d (1..16)", Cin, Cout);
  const int64_t total = (int64_t)B * HW;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == DITSEP_F32)
    output_layer_kernel<float><<<grid_for(total), 256, 0, st>>>((const float*)x, HW, Cin, ld_x, t, scale_by_sigma, w,
                                                                bias, Cout, out, total);
  else if (dtype == DITSEP_BF16)
    output_layer_kernel<bf16><<<grid_for(total), 256, 0, st>>>((const bf16*)x, HW, Cin, ld_x, t, scale_by_sigma, w, bias,
                                                               Cout, out, total);
  else
    DITSEP_REQUIRE(false, "output_layer: unknown dtype %d", dtype);
  return check_launch("output_layer_kernel");
}

// ------------------------------------------------------------------- allocation-free GroupNorm statistics
// ditsep_gn_stats_static: gn_sums + gn_finalize with the fixed-point sums in a file-scope device array, for callers
// that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  Every GPU that loads the library holds the array (DITSEP_GN_STATIC_MAX_BC * 32 bytes = 512 KiB) until
// exit.  All calls share it: it is cleared at the head of every call, and calls that overlap on the GPU (different
// streams) race.
static __device__ __align__(256) long long g_gn_static_sums[(size_t)DITSEP_GN_STATIC_MAX_BC * 4];

__global__ void __launch_bounds__(256) gn_static_zero_kernel(int64_t n) {
  // device code names the array directly; host code below uses the address cudaGetSymbolAddress returns
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    g_gn_static_sums[i] = 0;
}

// device address of g_gn_static_sums on the CURRENT device, resolved once per device
static int gn_static_address(long long** out) {
  static std::mutex mu;
  static long long* addr[64] = {};
  int dev = -1;
  DITSEP_CUDA(cudaGetDevice(&dev));
  DITSEP_REQUIRE(dev >= 0 && dev < 64, "gn_stats_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    DITSEP_CUDA(cudaGetSymbolAddress(&p, g_gn_static_sums));
    DITSEP_REQUIRE(p != nullptr, "gn_stats_static: cudaGetSymbolAddress returned null on device %d", dev);
    addr[dev] = static_cast<long long*>(p);
  }
  *out = addr[dev];
  return 0;
}

extern "C" int ditsep_gn_stats_static(const void* x, int dtype, int B, int64_t HW, int C, int ld, int G, const float* gamma,
                                      const float* beta, float eps, float* scale_shift, void* stream) {
  DITSEP_REQUIRE(x && gamma && beta && scale_shift, "gn_stats_static: null pointer");
  DITSEP_REQUIRE(B > 0 && C > 0, "gn_stats_static: bad sizes B=%d C=%d", B, C);
  DITSEP_REQUIRE((int64_t)B * C <= DITSEP_GN_STATIC_MAX_BC,
                 "gn_stats_static: B*C = %lld exceeds the static scratch (DITSEP_GN_STATIC_MAX_BC = %d); use "
                 "ditsep_gn_sums + ditsep_gn_finalize with a caller-owned sums buffer",
                 (long long)B * C, (int)DITSEP_GN_STATIC_MAX_BC);
  long long* sums = nullptr;
  if (int rc = gn_static_address(&sums)) return rc;
  const int64_t n = (int64_t)B * C * 4;
  gn_static_zero_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(n);
  if (int rc = check_launch("gn_static_zero_kernel")) return rc;
  if (int rc = ditsep_gn_sums(x, dtype, B, HW, C, ld, sums, stream)) return rc;
  ditsep_gn_job job = {};
  job.src_sums[0] = sums;
  job.src_C[0] = C;
  job.n_src = 1;
  job.C = C;
  job.G = G;
  job.eps = eps;
  job.HW = HW;
  job.gamma = gamma;
  job.beta = beta;
  job.scale_shift = scale_shift;
  return ditsep_gn_finalize(&job, B, stream);
}
