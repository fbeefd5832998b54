// HBM-bound NHWC kernels around the convolutions: GroupNorm statistics, the fused
// normalise + affine + SiLU apply, FIR x2 resampling (optionally of both the raw tensor and
// its normalised+activated version in one pass), the time-embedding MLPs and the 1x1
// output layer with the 1/t scaling.
//
// Reference semantics: nn.GroupNorm(min(C/4,32), C, eps=1e-6) + SiLU (layerspp.py:269-271,
// 283-285,297,318), upsample_2d / downsample_2d (up_or_down_sampling.py:206-273 ->
// upfirdn2d, op/upfirdn2d_kernel.cu:107-207), temb (ncsnpp.py:325-346, layerspp.py:44-46,
// 316-317), output scaling + output_layer (ncsnpp.py:473-477).
#include "common.cuh"

namespace ditsep {

// ------------------------------------------------------------------- GroupNorm statistics
// grid (nsplit, B); block = cv * ppb threads (cv = C/8 channel packs, ppb pixels in flight).
// Each thread keeps sum / sum-of-squares of its 8 channels in registers over its pixels.
// partial: [B][nsplit][C][2].
template <typename T>
__global__ void __launch_bounds__(256) gn_partial_kernel(const T* __restrict__ x, int64_t HW, int C, int ld, int cv,
                                                         int ppb, float* __restrict__ partial) {
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
    float* dst = partial + (((int64_t)b * nsplit + split) * C + c) * 2;
    dst[0] = a;
    dst[1] = bq;
  }
}

// grid B, block 256: combine partials in double, produce per-(b,c) scale / shift.
__global__ void __launch_bounds__(256) gn_finalize_kernel(const float* __restrict__ partial, int nsplit, int64_t HW, int C,
                                                          int G, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float eps,
                                                          float* __restrict__ scale_shift) {
  extern __shared__ double sh[];  // [C][2] then [G][2]
  double* chs = sh;
  double* grp = sh + 2 * C;
  const int b = blockIdx.x;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    double a = 0.0, q = 0.0;
    for (int s = 0; s < nsplit; ++s) {
      const float* src = partial + (((int64_t)b * nsplit + s) * C + c) * 2;
      a += (double)src[0];
      q += (double)src[1];
    }
    chs[2 * c] = a;
    chs[2 * c + 1] = q;
  }
  __syncthreads();
  const int cpg = C / G;
  for (int g = threadIdx.x; g < G; g += blockDim.x) {
    double a = 0.0, q = 0.0;
    for (int i = 0; i < cpg; ++i) {
      a += chs[2 * (g * cpg + i)];
      q += chs[2 * (g * cpg + i) + 1];
    }
    double n = (double)HW * cpg;
    double mean = a / n;
    double var = q / n - mean * mean;
    if (var < 0.0) var = 0.0;
    grp[2 * g] = mean;
    grp[2 * g + 1] = 1.0 / sqrt(var + (double)eps);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int g = c / cpg;
    double sc = grp[2 * g + 1] * (double)gamma[c];
    scale_shift[((int64_t)b * 2) * C + c] = (float)sc;
    scale_shift[((int64_t)b * 2 + 1) * C + c] = (float)((double)beta[c] - grp[2 * g] * sc);
  }
}

// Per-channel {sum, sum of squares} left by the producers of the normalised tensor (up to 4
// sources laid end to end along the channel axis, see ditsep_stats_src) -> per-(b,c) scale / shift.
// grid B, block 256.  Fixed summation order, double accumulation.
struct StatsSrcs {
  ditsep_stats_src s[4];
  int n;
};

__global__ void __launch_bounds__(512) gn_finalize_srcs_kernel(const StatsSrcs srcs, int64_t HW, int C, int G,
                                                               const float* __restrict__ gamma,
                                                               const float* __restrict__ beta, float eps,
                                                               float* __restrict__ scale_shift) {
  extern __shared__ double sh[];  // [C][2], [G][2], [nj][C][2]
  double* chs = sh;
  double* grp = sh + 2 * C;
  const int b = blockIdx.x;
  // thread -> (channel c, slice j of nj): slice j adds partials j, j+nj, ... ; the nj slice sums of a channel
  // are then added in order.  nj depends on (C, blockDim) only: fixed order, bit-reproducible.
  double* slice = sh + 2 * C + 2 * G;  // [nj][C][2]
  const int nj = max(1, (int)blockDim.x / C);
  for (int idx = threadIdx.x; idx < nj * C; idx += blockDim.x) {
    const int c = idx % C, j = idx / C;
    int cc = c, i = 0;
    while (i < srcs.n - 1 && cc >= srcs.s[i].C) {
      cc -= srcs.s[i].C;
      ++i;
    }
    const ditsep_stats_src& src = srcs.s[i];
    double a = 0.0, q = 0.0;
    const float2* sp = reinterpret_cast<const float2*>(src.ptr) + (int64_t)b * src.n_part * src.ld + cc;
    int k = j;
    for (; k + 7 * nj < src.n_part; k += 8 * nj) {  // 8 independent loads in flight, added in index order
      float2 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = sp[(int64_t)(k + u * nj) * src.ld];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        a += (double)v[u].x;
        q += (double)v[u].y;
      }
    }
    for (; k < src.n_part; k += nj) {
      const float2 v = sp[(int64_t)k * src.ld];
      a += (double)v.x;
      q += (double)v.y;
    }
    slice[(j * C + c) * 2] = a;
    slice[(j * C + c) * 2 + 1] = q;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    double a = 0.0, q = 0.0;
    for (int j = 0; j < nj; ++j) {
      a += slice[(j * C + c) * 2];
      q += slice[(j * C + c) * 2 + 1];
    }
    chs[2 * c] = a;
    chs[2 * c + 1] = q;
  }
  __syncthreads();
  const int cpg = C / G;
  for (int g = threadIdx.x; g < G; g += blockDim.x) {
    double a = 0.0, q = 0.0;
    for (int i = 0; i < cpg; ++i) {
      a += chs[2 * (g * cpg + i)];
      q += chs[2 * (g * cpg + i) + 1];
    }
    const double n = (double)HW * cpg;
    const double mean = a / n;
    double var = q / n - mean * mean;
    if (var < 0.0) var = 0.0;
    grp[2 * g] = mean;
    grp[2 * g + 1] = 1.0 / sqrt(var + (double)eps);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int g = c / cpg;
    const double sc = grp[2 * g + 1] * (double)gamma[c];
    scale_shift[((int64_t)b * 2) * C + c] = (float)sc;
    scale_shift[((int64_t)b * 2 + 1) * C + c] = (float)((double)beta[c] - grp[2 * g] * sc);
  }
}

// ------------------------------------------------------------------- normalise + SiLU
// grid (blocks per image, B); block = cv * ppb threads (cv = C/8 channel packs, ppb pixels in flight).  A thread
// keeps ONE channel pack for its whole life, so the per-(b,c) scale / shift sit in registers and the loop body is
// load 16 B (bf16) -> 8 fma + SiLU -> store 16 B, four pixels in flight per thread.
template <typename T>
__global__ void __launch_bounds__(256) gn_apply_kernel(const T* __restrict__ x, int64_t HW, int C, int ld_x,
                                                       const float* __restrict__ scale_shift, int silu,
                                                       T* __restrict__ y, int ld_y, int cv, int ppb) {
  const int b = blockIdx.y;
  const int v = threadIdx.x % cv, pp = threadIdx.x / cv;
  if (pp >= ppb) return;
  float sc[8], sh[8];
  Pack8<float>::load(scale_shift + ((int64_t)b * 2) * C + v * 8, sc);
  Pack8<float>::load(scale_shift + ((int64_t)b * 2 + 1) * C + v * 8, sh);
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

// ------------------------------------------------------------------- FIR x2 resampling
// Separable [1,3,3,1]: down = taps (1,3,3,1)/8 per axis at input 2o-1..2o+2 (zero outside);
// up = polyphase: out[2i] = (x[i-1] + 3 x[i])/4, out[2i+1] = (3 x[i] + x[i+1])/4 per axis.
// One thread = one channel pack of one OUTPUT pixel (down) or one INPUT pixel -> 2x2 outputs
// (up).  If scale_shift != nullptr a second output gets FIR(SiLU(x*scale+shift)).
template <typename T, bool kDual>
__global__ void __launch_bounds__(256) fir_down_kernel(const T* __restrict__ x, int H, int W, int C, int ld_x,
                                                       const float* __restrict__ scale_shift, T* __restrict__ out_x,
                                                       int ld_ox, T* __restrict__ out_h, int ld_oh, int64_t total) {
  const int cv = C / 8, Ho = H / 2, Wo = W / 2;
  const float kw[4] = {0.125f, 0.375f, 0.375f, 0.125f};
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(idx % cv);
    int64_t r = idx / cv;
    const int ox = (int)(r % Wo);
    r /= Wo;
    const int oy = (int)(r % Ho);
    const int b = (int)(r / Ho);
    float sc[8], sh[8];
    if (kDual) {
      Pack8<float>::load(scale_shift + ((int64_t)b * 2) * C + v * 8, sc);
      Pack8<float>::load(scale_shift + ((int64_t)b * 2 + 1) * C + v * 8, sh);
    }
    float ax[8], ah[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) ax[i] = ah[i] = 0.f;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int iy = 2 * oy + a - 1;
      if (iy < 0 || iy >= H) continue;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int ix = 2 * ox + c - 1;
        if (ix < 0 || ix >= W) continue;
        float f[8];
        Pack8<T>::load(x + (((int64_t)b * H + iy) * W + ix) * ld_x + v * 8, f);
        const float w = kw[a] * kw[c];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          ax[i] = fmaf(w, f[i], ax[i]);
          if (kDual) ah[i] = fmaf(w, silu_f(fmaf(f[i], sc[i], sh[i])), ah[i]);
        }
      }
    }
    const int64_t opix = ((int64_t)b * Ho + oy) * Wo + ox;
    if (out_x) Pack8<T>::store(out_x + opix * ld_ox + v * 8, ax);
    if (kDual) Pack8<T>::store(out_h + opix * ld_oh + v * 8, ah);
  }
}

template <typename T, bool kDual>
__global__ void __launch_bounds__(256) fir_up_kernel(const T* __restrict__ x, int H, int W, int C, int ld_x,
                                                     const float* __restrict__ scale_shift, T* __restrict__ out_x,
                                                     int ld_ox, T* __restrict__ out_h, int ld_oh, int64_t total) {
  const int cv = C / 8, Ho = 2 * H, Wo = 2 * W;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(idx % cv);
    int64_t r = idx / cv;
    const int ix = (int)(r % W);
    r /= W;
    const int iy = (int)(r % H);
    const int b = (int)(r / H);
    float sc[8], sh[8];
    if (kDual) {
      Pack8<float>::load(scale_shift + ((int64_t)b * 2) * C + v * 8, sc);
      Pack8<float>::load(scale_shift + ((int64_t)b * 2 + 1) * C + v * 8, sh);
    }
    // out[2i+dy][2j+dx] = sum over (ry, rx) in the 3x3 neighbourhood with per-axis weights
    // dy=0: rows (i-1: 1/4, i: 3/4) ; dy=1: rows (i: 3/4, i+1: 1/4)
    float ox[4][8], oh[4][8];
#pragma unroll
    for (int o = 0; o < 4; ++o)
#pragma unroll
      for (int i = 0; i < 8; ++i) ox[o][i] = oh[o][i] = 0.f;
#pragma unroll
    for (int ry = -1; ry <= 1; ++ry) {
      const int yy = iy + ry;
      if (yy < 0 || yy >= H) continue;
      const float wy0 = (ry == -1) ? 0.25f : (ry == 0 ? 0.75f : 0.f);
      const float wy1 = (ry == 1) ? 0.25f : (ry == 0 ? 0.75f : 0.f);
#pragma unroll
      for (int rx = -1; rx <= 1; ++rx) {
        const int xx = ix + rx;
        if (xx < 0 || xx >= W) continue;
        const float wx0 = (rx == -1) ? 0.25f : (rx == 0 ? 0.75f : 0.f);
        const float wx1 = (rx == 1) ? 0.25f : (rx == 0 ? 0.75f : 0.f);
        float f[8];
        Pack8<T>::load(x + (((int64_t)b * H + yy) * W + xx) * ld_x + v * 8, f);
        const float w00 = wy0 * wx0, w01 = wy0 * wx1, w10 = wy1 * wx0, w11 = wy1 * wx1;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          ox[0][i] = fmaf(w00, f[i], ox[0][i]);
          ox[1][i] = fmaf(w01, f[i], ox[1][i]);
          ox[2][i] = fmaf(w10, f[i], ox[2][i]);
          ox[3][i] = fmaf(w11, f[i], ox[3][i]);
          if (kDual) {
            const float hh = silu_f(fmaf(f[i], sc[i], sh[i]));
            oh[0][i] = fmaf(w00, hh, oh[0][i]);
            oh[1][i] = fmaf(w01, hh, oh[1][i]);
            oh[2][i] = fmaf(w10, hh, oh[2][i]);
            oh[3][i] = fmaf(w11, hh, oh[3][i]);
          }
        }
      }
    }
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int64_t opix = ((int64_t)b * Ho + 2 * iy + dy) * Wo + 2 * ix + dx;
        if (out_x) Pack8<T>::store(out_x + opix * ld_ox + v * 8, ox[dy * 2 + dx]);
        if (kDual) Pack8<T>::store(out_h + opix * ld_oh + v * 8, oh[dy * 2 + dx]);
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

// ------------------------------------------------------------------- time embedding
// grid B, block 256.  smem: feat[2nf] | hid[4nf]
__global__ void __launch_bounds__(256) temb_mlp_kernel(const float* __restrict__ t, const float* __restrict__ gfp_w,
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
  for (int o = warp; o < 4 * nf; o += nw) {
    const float* wr = w1 + (int64_t)o * 2 * nf;
    float acc = 0.f;
    for (int k = lane; k < 2 * nf; k += 32) acc = fmaf(wr[k], feat[k], acc);
    acc = warp_sum(acc);
    if (lane == 0) {
      float h = acc + b1[o];
      hid[o] = h / (1.0f + expf(-h));
    }
  }
  __syncthreads();
  for (int o = warp; o < 4 * nf; o += nw) {
    const float* wr = w2 + (int64_t)o * 4 * nf;
    float acc = 0.f;
    for (int k = lane; k < 4 * nf; k += 32) acc = fmaf(wr[k], hid[k], acc);
    acc = warp_sum(acc);
    if (lane == 0) {
      float h = acc + b2[o];
      act_temb[(int64_t)b * 4 * nf + o] = h / (1.0f + expf(-h));
    }
  }
}

// One warp per output row r (all res-blocks' Dense_0 rows concatenated), looping over B.
__global__ void __launch_bounds__(256) temb_dense_kernel(const float* __restrict__ act_temb,
                                                         const float* __restrict__ dense_w,
                                                         const float* __restrict__ dense_b, float* __restrict__ out,
                                                         int B, int K, int R) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= R) return;
  const float* wr = dense_w + (int64_t)warp * K;
  for (int b = 0; b < B; ++b) {
    const float* a = act_temb + (int64_t)b * K;
    float acc = 0.f;
    for (int k = lane; k < K; k += 32) acc = fmaf(wr[k], a[k], acc);
    acc = warp_sum(acc);
    if (lane == 0) out[(int64_t)b * R + warp] = acc + dense_b[warp];
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

static int launch_gn_partial(const void* x, int dtype, int B, int64_t HW, int C, int ld, float* partial, int nsplit,
                             cudaStream_t st) {
  DITSEP_REQUIRE(x && partial, "gn partial sums: null pointer");
  DITSEP_REQUIRE(C % 8 == 0 && ld % 8 == 0 && C <= ld, "gn partial sums: C=%d ld=%d must be multiples of 8, C<=ld", C, ld);
  DITSEP_REQUIRE(B > 0 && HW > 0 && nsplit > 0 && nsplit <= HW, "gn partial sums: bad B/HW/nsplit");
  const int cv = C / 8;
  DITSEP_REQUIRE(cv <= 256, "gn partial sums: C=%d too wide", C);
  const int ppb = 256 / cv;
  dim3 grid(nsplit, B);
  size_t smem = (size_t)ppb * C * 2 * sizeof(float);
  if (dtype == DITSEP_F32)
    gn_partial_kernel<float><<<grid, 256, smem, st>>>((const float*)x, HW, C, ld, cv, ppb, partial);
  else if (dtype == DITSEP_BF16)
    gn_partial_kernel<bf16><<<grid, 256, smem, st>>>((const bf16*)x, HW, C, ld, cv, ppb, partial);
  else
    DITSEP_REQUIRE(false, "gn partial sums: unknown dtype %d", dtype);
  return check_launch("gn_partial_kernel");
}

extern "C" int ditsep_gn_partial_sums(const void* x, int dtype, int B, int64_t HW, int C, int ld, float* partial,
                                      int nsplit, void* stream) {
  return launch_gn_partial(x, dtype, B, HW, C, ld, partial, nsplit, (cudaStream_t)stream);
}

extern "C" int ditsep_gn_stats(const void* x, int dtype, int B, int64_t HW, int C, int ld, int G, const float* gamma,
                               const float* beta, float eps, float* partial, int nsplit, float* scale_shift,
                               void* stream) {
  DITSEP_REQUIRE(gamma && beta && scale_shift, "gn_stats: null pointer");
  DITSEP_REQUIRE(G > 0 && C % G == 0, "gn_stats: C=%d not divisible by G=%d", C, G);
  cudaStream_t st = (cudaStream_t)stream;
  if (int rc = launch_gn_partial(x, dtype, B, HW, C, ld, partial, nsplit, st)) return rc;
  gn_finalize_kernel<<<B, 256, (size_t)(2 * C + 2 * G) * sizeof(double), st>>>(partial, nsplit, HW, C, G, gamma, beta,
                                                                                eps, scale_shift);
  return check_launch("gn_finalize_kernel");
}

extern "C" int ditsep_gn_finalize(const ditsep_stats_src* host_srcs, int n_srcs, int B, int64_t HW, int G,
                                  const float* gamma, const float* beta, float eps, float* scale_shift, void* stream) {
  DITSEP_REQUIRE(host_srcs && gamma && beta && scale_shift, "gn_finalize: null pointer");
  DITSEP_REQUIRE(n_srcs >= 1 && n_srcs <= 4, "gn_finalize: n_srcs=%d (1..4)", n_srcs);
  StatsSrcs srcs;
  srcs.n = n_srcs;
  int C = 0;
  for (int i = 0; i < n_srcs; ++i) {
    const ditsep_stats_src& s = host_srcs[i];
    DITSEP_REQUIRE(s.ptr && s.kind == DITSEP_STATS_SPLIT_SUMS, "gn_finalize: source %d has no statistics (kind %d)", i,
                   s.kind);
    DITSEP_REQUIRE(s.C > 0 && s.ld >= s.C && s.n_part > 0, "gn_finalize: source %d: bad C/ld/n_part", i);
    srcs.s[i] = s;
    C += s.C;
  }
  DITSEP_REQUIRE(B > 0 && HW > 0 && G > 0 && C % G == 0, "gn_finalize: bad sizes C=%d G=%d", C, G);
  DITSEP_REQUIRE(C <= 512, "gn_finalize: C=%d too wide", C);
  const int threads = 512, nj = threads / C > 0 ? threads / C : 1;
  gn_finalize_srcs_kernel<<<B, threads, (size_t)(2 * C + 2 * G + 2 * nj * C) * sizeof(double), (cudaStream_t)stream>>>(
      srcs, HW, C, G, gamma, beta, eps, scale_shift);
  return check_launch("gn_finalize_srcs_kernel");
}

extern "C" int ditsep_gn_apply(const void* x, int dtype, int B, int64_t HW, int C, int ld_x, const float* scale_shift,
                               int silu, void* y, int ld_y, void* stream) {
  DITSEP_REQUIRE(x && y && scale_shift, "gn_apply: null pointer");
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
  if (dtype == DITSEP_F32)
    gn_apply_kernel<float><<<grid, 256, 0, st>>>((const float*)x, HW, C, ld_x, scale_shift, silu, (float*)y, ld_y, cv, ppb);
  else if (dtype == DITSEP_BF16)
    gn_apply_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)x, HW, C, ld_x, scale_shift, silu, (bf16*)y, ld_y, cv, ppb);
  else
    DITSEP_REQUIRE(false, "gn_apply: unknown dtype %d", dtype);
  return check_launch("gn_apply_kernel");
}

template <typename T>
static int fir_dispatch(const T* x, int B, int H, int W, int C, int ld_x, int up, const float* ss, T* out_x, int ld_ox,
                        T* out_h, int ld_oh, cudaStream_t st) {
  if (C % 8 != 0 || ld_x % 8 != 0) {
    DITSEP_REQUIRE(ss == nullptr && out_h == nullptr && out_x != nullptr,
                   "fir_resample: narrow tensors (C=%d) support the plain path only", C);
    const int64_t total = (int64_t)B * (up ? 4 : 1) * H * W / (up ? 1 : 4) * C;
    fir_narrow_kernel<T><<<grid_for(total), 256, 0, st>>>(x, H, W, C, ld_x, up, out_x, ld_ox, total);
    return check_launch("fir_narrow_kernel");
  }
  DITSEP_REQUIRE((out_x == nullptr || ld_ox % 8 == 0) && (out_h == nullptr || ld_oh % 8 == 0),
                 "fir_resample: output strides must be multiples of 8");
  const int cv = C / 8;
  if (up) {
    const int64_t total = (int64_t)B * H * W * cv;
    if (ss)
      fir_up_kernel<T, true><<<grid_for(total), 256, 0, st>>>(x, H, W, C, ld_x, ss, out_x, ld_ox, out_h, ld_oh, total);
    else
      fir_up_kernel<T, false><<<grid_for(total), 256, 0, st>>>(x, H, W, C, ld_x, ss, out_x, ld_ox, out_h, ld_oh, total);
    return check_launch("fir_up_kernel");
  }
  const int64_t total = (int64_t)B * (H / 2) * (W / 2) * cv;
  if (ss)
    fir_down_kernel<T, true><<<grid_for(total), 256, 0, st>>>(x, H, W, C, ld_x, ss, out_x, ld_ox, out_h, ld_oh, total);
  else
    fir_down_kernel<T, false><<<grid_for(total), 256, 0, st>>>(x, H, W, C, ld_x, ss, out_x, ld_ox, out_h, ld_oh, total);
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

extern "C" int ditsep_temb_mlp(const float* t, const float* gfp_w, const float* w1, const float* b1, const float* w2,
                               const float* b2, float* act_temb, int B, int nf, void* stream) {
  DITSEP_REQUIRE(t && gfp_w && w1 && b1 && w2 && b2 && act_temb, "temb_mlp: null pointer");
  DITSEP_REQUIRE(B > 0 && nf > 0 && nf <= 1024, "temb_mlp: bad B=%d nf=%d", B, nf);
  temb_mlp_kernel<<<B, 256, (size_t)6 * nf * sizeof(float), (cudaStream_t)stream>>>(t, gfp_w, w1, b1, w2, b2, act_temb,
                                                                                    nf);
  return check_launch("temb_mlp_kernel");
}

extern "C" int ditsep_temb_dense(const float* act_temb, const float* dense_w, const float* dense_b, float* out, int B,
                                 int K, int R, void* stream) {
  DITSEP_REQUIRE(act_temb && dense_w && dense_b && out, "temb_dense: null pointer");
  DITSEP_REQUIRE(B > 0 && K > 0 && R > 0, "temb_dense: bad sizes");
  temb_dense_kernel<<<ceil_div((int64_t)R * 32, 256), 256, 0, (cudaStream_t)stream>>>(act_temb, dense_w, dense_b, out,
                                                                                      B, K, R);
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
