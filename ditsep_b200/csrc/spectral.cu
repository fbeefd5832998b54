// STFT + power-law compression, and expansion + iSTFT, for n_fft = 510 / hop = 128.
//
// n_fft = 510 = 2 * 3 * 5 * 17 is not a power of two, and zero-padding to 512 would move
// the bin frequencies, so this is a true 510-point real DFT: even/odd samples are packed
// into a 255-point complex sequence, transformed with a shared-memory-staged two-pass
// Cooley-Tukey FFT (255 = 15 x 17; each lane owns 8 butterflies' outputs per pass, passes
// exchange through per-warp shared memory with warp-level syncs only), then untangled
// into the 256 one-sided bins.  Twiddles come from one fp32 table of e^{-2 pi i j/510}
// evaluated in double precision.  The window multiply, the centre / right zero padding,
// |X|^e e^{j arg X} * factor, the [Re.., Im..] channel packing, the frame padding to W and
// NCSN++'s "2x-1" are all fused into the same kernel; the inverse kernel fuses the
// expansion, the Hermitian inverse FFT, the synthesis window, overlap-add, the
// window-envelope normalisation and the trim to T samples.
//
// Reference semantics: ScoreModelNCSNpp.pre_process / post_process
// (src/models/diffsep/score_models.py:41-124) over torchaudio Spectrogram /
// InverseSpectrogram (= torch.stft / torch.istft, center=True, pad_mode="constant").
#include "common.cuh"
#include "dft_codelets.cuh"

namespace ditsep {

constexpr int kNfft = 510;
constexpr int kHalf = 255;
constexpr int kHop = 128;
constexpr int kBins = 256;
constexpr int kWarps = 8;

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cconj(float2 a) { return make_float2(a.x, -a.y); }

__device__ __forceinline__ void fill_table(float2* tab) {
  for (int j = threadIdx.x; j < kNfft; j += blockDim.x) {
    double s, c;
    sincospi(-2.0 * (double)j / (double)kNfft, &s, &c);
    tab[j] = make_float2((float)c, (float)s);
  }
}

// 255-point complex DFT of zbuf (in place, via ybuf), executed by one warp.
// kInverse uses conjugated twiddles (no 1/255 scaling).  Index maps: n = 17 n1 + n2,
// k = k1 + 15 k2;  Z[k] = sum_n2 W17^{n2 k2} W255^{n2 k1} sum_n1 z[17 n1 + n2] W15^{n1 k1}.
#ifndef DITSEP_DFT_CODELETS
#define DITSEP_DFT_CODELETS 1
#endif
#if DITSEP_DFT_CODELETS
// Register-resident sub-transforms (csrc/dft_codelets.cuh, generated): lane n2 < 17 runs one straight-line 15-point DFT
// (Good-Thomas 3 x 5, literal twiddles) over its column, lane k1 < 15 one paired 17-point DFT over its row; the only
// table reads left are the 15 inter-pass twiddles per lane.  ~750 warp instructions per transform against ~2600 for
// the table-driven loops below (kept behind -DDITSEP_DFT_CODELETS=0 for A/B timing).  The inverse is the forward
// transform of the input with real and imaginary parts swapped, swapped back on the way out.
template <bool kInverse>
__device__ __forceinline__ void warp_dft255(float2* zbuf, float2* ybuf, const float2* tab, int lane) {
  if (lane < 17) {
    float2 in[15], out[15];
#pragma unroll
    for (int n1 = 0; n1 < 15; ++n1) {
      const float2 v = zbuf[17 * n1 + lane];
      in[n1] = kInverse ? make_float2(v.y, v.x) : v;
    }
    dft15(in, out);
    int tw = 0;
#pragma unroll
    for (int k1 = 0; k1 < 15; ++k1) {
      ybuf[k1 * 17 + lane] = cmul(out[k1], tab[tw]);  // W255^{n2 k1} = tab[2 n2 k1 mod 510]
      tw += 2 * lane;
      if (tw >= kNfft) tw -= kNfft;
    }
  }
  __syncwarp();
  if (lane < 15) {
    float2 in[17], out[17];
#pragma unroll
    for (int n2 = 0; n2 < 17; ++n2) in[n2] = ybuf[lane * 17 + n2];
    dft17(in, out);
#pragma unroll
    for (int k2 = 0; k2 < 17; ++k2) zbuf[lane + 15 * k2] = kInverse ? make_float2(out[k2].y, out[k2].x) : out[k2];
  }
  __syncwarp();
}
#else
template <bool kInverse>
__device__ __forceinline__ void warp_dft255(float2* zbuf, float2* ybuf, const float2* tab, int lane) {
  for (int idx = lane; idx < kHalf; idx += 32) {
    const int k1 = idx / 17, n2 = idx - k1 * 17;
    const int step = (34 * k1) % kNfft;
    int tw = 0;
    float2 acc = make_float2(0.f, 0.f);
#pragma unroll
    for (int n1 = 0; n1 < 15; ++n1) {
      float2 w = tab[tw];
      if (kInverse) w.y = -w.y;
      float2 v = zbuf[17 * n1 + n2];
      acc.x += v.x * w.x - v.y * w.y;
      acc.y += v.x * w.y + v.y * w.x;
      tw += step;
      if (tw >= kNfft) tw -= kNfft;
    }
    float2 w = tab[(2 * n2 * k1) % kNfft];
    if (kInverse) w.y = -w.y;
    ybuf[idx] = cmul(acc, w);  // ybuf[k1 * 17 + n2]
  }
  __syncwarp();
  for (int k = lane; k < kHalf; k += 32) {
    const int k2 = k / 15, k1 = k - k2 * 15;
    const int step = (30 * k2) % kNfft;
    int tw = 0;
    float2 acc = make_float2(0.f, 0.f);
#pragma unroll
    for (int n2 = 0; n2 < 17; ++n2) {
      float2 w = tab[tw];
      if (kInverse) w.y = -w.y;
      float2 v = ybuf[k1 * 17 + n2];
      acc.x += v.x * w.x - v.y * w.y;
      acc.y += v.x * w.y + v.y * w.x;
      tw += step;
      if (tw >= kNfft) tw -= kNfft;
    }
    zbuf[k] = acc;
  }
  __syncwarp();
}
#endif

// ------------------------------------------------------------------------------- forward
constexpr int kFT = 8;                               // frames per block
constexpr int kSigLen = (kFT - 1) * kHop + kNfft;    // 1406 samples feed 8 frames

// grid (W / kFT, B), block 256.  dyn smem: tab | win | sig[NC][kSigLen] | scratch | ostage
template <typename T>
__global__ void __launch_bounds__(256) stft_compress_kernel(const float* __restrict__ xt, const float* __restrict__ mix,
                                                            const float* __restrict__ window, T* __restrict__ feat,
                                                            int n_src, int64_t Tlen, int W, int n_frames,
                                                            float exponent, float factor, int affine,
                                                            float2* __restrict__ mix_cache, int mix_mode) {
  // mix_mode: 0 = no cache; 1 = also write the mixture channel's finished (Re, Im) to mix_cache [B][W][256]; 2 = read
  // it from there instead of transforming the mixture again (it does not change during a sampler run: a third of the
  // transforms of every score evaluation)
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int NC = n_src + 1, C2 = 2 * NC;
  float2* tab = reinterpret_cast<float2*>(smem_raw);
  float* win = reinterpret_cast<float*>(tab + kNfft);
  float* sig = win + kNfft + 2;  // keep 8-byte alignment of what follows
  float2* scratch = reinterpret_cast<float2*>(sig + NC * kSigLen + ((NC * kSigLen) & 1));
  float* ostage = reinterpret_cast<float*>(scratch + kWarps * 2 * kBins);

  const int b = blockIdx.y, f0 = blockIdx.x * kFT;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  fill_table(tab);
  for (int j = threadIdx.x; j < kNfft; j += blockDim.x) win[j] = window[j];
  // signal tile: sample index of sig[.][0] in the un-padded waveform
  const int64_t s0 = (int64_t)f0 * kHop - kHalf;
  for (int idx = threadIdx.x; idx < NC * kSigLen; idx += blockDim.x) {
    int ch = idx / kSigLen, i = idx - ch * kSigLen;
    int64_t j = s0 + i;
    float v = 0.f;
    if (j >= 0 && j < Tlen) v = (ch < n_src) ? xt[((int64_t)b * n_src + ch) * Tlen + j] : mix[(int64_t)b * Tlen + j];
    sig[idx] = v;
  }
  __syncthreads();

  float2* zbuf = scratch + warp * 2 * kBins;
  float2* ybuf = zbuf + kBins;
  for (int item = warp; item < kFT * NC; item += kWarps) {
    const int fl = item % kFT, ch = item / kFT;
    const int f = f0 + fl;
    const bool is_mix = ch == n_src;
    float2* const cache_row = mix_mode ? mix_cache + ((int64_t)b * W + f) * kBins : nullptr;
    if (is_mix && mix_mode == 2) {
      for (int k = lane; k < kBins; k += 32) {
        const float2 v = cache_row[k];
        ostage[k * (kFT * C2 + 1) + fl * C2 + ch] = v.x;
        ostage[k * (kFT * C2 + 1) + fl * C2 + NC + ch] = v.y;
      }
      continue;
    }
    if (f < n_frames) {
      const float* s = sig + ch * kSigLen + fl * kHop;
      for (int m = lane; m < kHalf; m += 32)
        zbuf[m] = make_float2(win[2 * m] * s[2 * m], win[2 * m + 1] * s[2 * m + 1]);
      __syncwarp();
      warp_dft255<false>(zbuf, ybuf, tab, lane);
    }
    for (int k = lane; k < kBins; k += 32) {
      float re = 0.f, im = 0.f;
      if (f < n_frames) {
        float2 zk = zbuf[k == kHalf ? 0 : k];
        float2 zm = cconj(zbuf[k == 0 ? 0 : kHalf - k]);
        float2 e = make_float2(0.5f * (zk.x + zm.x), 0.5f * (zk.y + zm.y));
        float2 d = make_float2(0.5f * (zk.x - zm.x), 0.5f * (zk.y - zm.y));
        float2 o = cmul(tab[k], d);  // X = e - i * T[k] * d
        float2 X = make_float2(e.x + o.y, e.y - o.x);
        if (k == 0 || k == kHalf) X.y = 0.f;
        // |X|^e e^{j arg X} * factor = X * |X|^(e-1) * factor
        float mag2 = X.x * X.x + X.y * X.y;
        float sc = 0.f;
        if (exponent == 0.f) {  // transform="log" (score_models.py:49-51): log1p(|X|) e^{j arg X} * |factor|
          const float mag = sqrtf(mag2);
          sc = mag > 0.f ? fabsf(factor) * log1pf(mag) / mag : 0.f;
        } else {
          if (mag2 > 0.f) sc = (exponent == 0.5f) ? factor * rsqrtf(sqrtf(mag2)) : factor * powf(mag2, 0.5f * (exponent - 1.0f));
          if (exponent == 1.0f) sc = factor;
        }
        re = X.x * sc;
        im = X.y * sc;
      }
      if (affine) {
        re = 2.f * re - 1.f;
        im = 2.f * im - 1.f;
      }
      ostage[k * (kFT * C2 + 1) + fl * C2 + ch] = re;  // +1: odd row stride, conflict-free over k
      ostage[k * (kFT * C2 + 1) + fl * C2 + NC + ch] = im;
      if (is_mix && mix_mode == 1) cache_row[k] = make_float2(re, im);
    }
    __syncwarp();
  }
  __syncthreads();
  const int row = kFT * C2;
  for (int idx = threadIdx.x; idx < kBins * row; idx += blockDim.x) {
    int k = idx / row, r = idx - k * row;
    feat[(((int64_t)b * kBins + k) * W + f0) * C2 + r] = from_f<T>(ostage[k * (row + 1) + r]);
  }
}

// ------------------------------------------------------------------------------- inverse
constexpr int kIT = 16;             // output hop-blocks (frames) per CTA
constexpr int kIF = kIT + 3;        // frames that overlap them

// grid (ceil((T + 255) / (128 * kIT)), B), block 256.
// dyn smem: tab | win | scratch | tbuf[n_src][kIF][510]
__global__ void __launch_bounds__(256) expand_istft_kernel(const float* __restrict__ spec, const float* __restrict__ window,
                                                           float* __restrict__ out, int n_src, int64_t Tlen, int W,
                                                           int n_frames, float exponent, float factor) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tab = reinterpret_cast<float2*>(smem_raw);
  float* win = reinterpret_cast<float*>(tab + kNfft);
  float2* scratch = reinterpret_cast<float2*>(win + kNfft + 2);
  float* tbuf = reinterpret_cast<float*>(scratch + kWarps * 2 * kBins);
  const int C = 2 * n_src;
  const int b = blockIdx.y, f0 = blockIdx.x * kIT;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  fill_table(tab);
  for (int j = threadIdx.x; j < kNfft; j += blockDim.x) win[j] = window[j];
  __syncthreads();

  float2* zbuf = scratch + warp * 2 * kBins;
  float2* xbuf = zbuf + kBins;
  const float inv_factor = 1.0f / fabsf(factor);
  const float pw = exponent > 0.f ? 1.0f / exponent - 1.0f : 0.f;
  for (int item = warp; item < kIF * n_src; item += kWarps) {
    const int src = item % n_src, fl = item / n_src;
    const int f = f0 - 3 + fl;
    float* dst = tbuf + (src * kIF + fl) * kNfft;
    if (f < 0 || f >= n_frames) {
      for (int n = lane; n < kNfft; n += 32) dst[n] = 0.f;
      continue;
    }
    for (int k = lane; k < kBins; k += 32) {
      const float* px = spec + (((int64_t)b * kBins + k) * W + f) * C;
      float re = px[src] * inv_factor, im = px[n_src + src] * inv_factor;
      float mag2 = re * re + im * im;
      float sc = 1.0f;
      if (exponent == 0.f) {  // transform="log" backwards (score_models.py:66-68): (exp(|Y|) - 1) e^{j arg Y}
        const float mag = sqrtf(mag2);
        sc = mag > 0.f ? expm1f(mag) / mag : 0.f;
      } else if (exponent != 1.0f) {
        sc = (mag2 > 0.f) ? ((exponent == 0.5f) ? sqrtf(mag2) : powf(mag2, 0.5f * pw)) : 0.f;
      }
      xbuf[k] = make_float2(re * sc, im * sc);
    }
    __syncwarp();
    // Hermitian pack: Z[k] = E[k] + i O[k], E = (X[k] + conj X[255-k])/2,
    // O = (X[k] - conj X[255-k])/2 * e^{+2 pi i k/510}; DC / Nyquist imaginary parts ignored.
    for (int k = lane; k < kHalf; k += 32) {
      float2 xk = xbuf[k], xm = cconj(xbuf[kHalf - k]);
      if (k == 0) { xk.y = 0.f; xm.y = 0.f; }
      float2 e = make_float2(0.5f * (xk.x + xm.x), 0.5f * (xk.y + xm.y));
      float2 d = make_float2(0.5f * (xk.x - xm.x), 0.5f * (xk.y - xm.y));
      float2 o = cmul(cconj(tab[k]), d);
      zbuf[k] = make_float2(e.x - o.y, e.y + o.x);
    }
    __syncwarp();
    warp_dft255<true>(zbuf, xbuf, tab, lane);
    const float sc255 = 1.0f / 255.0f;
    for (int m = lane; m < kHalf; m += 32) {
      float2 z = zbuf[m];
      dst[2 * m] = z.x * sc255 * win[2 * m];
      dst[2 * m + 1] = z.y * sc255 * win[2 * m + 1];
    }
    __syncwarp();
  }
  __syncthreads();
  // overlap-add + envelope + trim: OLA index p = j + 255
  for (int idx = threadIdx.x; idx < n_src * kIT * kHop; idx += blockDim.x) {
    const int src = idx / (kIT * kHop), pl = idx - src * (kIT * kHop);
    const int64_t p = (int64_t)f0 * kHop + pl;
    const int64_t j = p - kHalf;
    if (j < 0 || j >= Tlen) continue;
    const int fhi = (int)(p / kHop);
    float acc = 0.f, env = 0.f;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      const int f = fhi - d;
      const int off = (int)(p - (int64_t)f * kHop);
      if (f >= 0 && f < n_frames && off < kNfft) {
        acc += tbuf[(src * kIF + (f - f0 + 3)) * kNfft + off];
        env += win[off] * win[off];
      }
    }
    out[((int64_t)b * n_src + src) * Tlen + j] = acc / env;
  }
}

}  // namespace ditsep

using namespace ditsep;

static int launch_stft_compress(const float* xt, const float* mix, const float* window, void* feat, int dtype, int B,
                                int n_src, int64_t T, int W, float exponent, float factor, int affine_2x_minus_1,
                                float* mix_cache, int mix_mode, void* stream) {
  DITSEP_REQUIRE(xt && mix && window && feat, "stft_compress: null pointer");
  DITSEP_REQUIRE(mix_mode >= 0 && mix_mode <= 2 && (mix_mode == 0 || mix_cache != nullptr),
                 "stft_compress: mix_mode=%d (0 none, 1 fill, 2 use) needs a cache buffer", mix_mode);
  DITSEP_REQUIRE((reinterpret_cast<uintptr_t>(mix_cache) & 7) == 0, "stft_compress: mix_cache must be 8-byte aligned");
  DITSEP_REQUIRE(B > 0 && T > 0 && n_src >= 1 && n_src <= 4, "stft_compress: bad sizes B=%d T=%lld n_src=%d", B,
                 (long long)T, n_src);
  const int n_frames = 1 + (int)((T + (kNfft - kHop)) / kHop);
  DITSEP_REQUIRE(W % kFT == 0 && W >= n_frames, "stft_compress: W=%d must be a multiple of %d and >= %d frames", W,
                 kFT, n_frames);
  DITSEP_REQUIRE(exponent >= 0.f, "stft_compress: exponent must be > 0 (power law) or 0 (log1p transform)");
  const int NC = n_src + 1, C2 = 2 * NC;
  size_t smem = kNfft * 8 + (kNfft + 2) * 4 + (size_t)(NC * kSigLen + ((NC * kSigLen) & 1)) * 4 +
                (size_t)kWarps * 2 * kBins * 8 + (size_t)kBins * (kFT * C2 + 1) * 4;
  dim3 grid(W / kFT, B);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == DITSEP_F32) {
    DITSEP_CUDA(cudaFuncSetAttribute(stft_compress_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    stft_compress_kernel<float><<<grid, 256, smem, st>>>(xt, mix, window, (float*)feat, n_src, T, W, n_frames, exponent,
                                                         factor, affine_2x_minus_1, reinterpret_cast<float2*>(mix_cache),
                                                         mix_mode);
  } else if (dtype == DITSEP_BF16) {
    DITSEP_CUDA(cudaFuncSetAttribute(stft_compress_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    stft_compress_kernel<bf16><<<grid, 256, smem, st>>>(xt, mix, window, (bf16*)feat, n_src, T, W, n_frames, exponent,
                                                        factor, affine_2x_minus_1, reinterpret_cast<float2*>(mix_cache),
                                                        mix_mode);
  } else {
    DITSEP_REQUIRE(false, "stft_compress: unknown dtype %d", dtype);
  }
  return check_launch("stft_compress_kernel");
}

extern "C" int ditsep_stft_compress(const float* xt, const float* mix, const float* window, void* feat, int dtype,
                                    int B, int n_src, int64_t T, int W, float exponent, float factor,
                                    int affine_2x_minus_1, void* stream) {
  return launch_stft_compress(xt, mix, window, feat, dtype, B, n_src, T, W, exponent, factor, affine_2x_minus_1, nullptr, 0,
                              stream);
}

extern "C" int ditsep_stft_compress_cached(const float* xt, const float* mix, const float* window, void* feat, int dtype,
                                           int B, int n_src, int64_t T, int W, float exponent, float factor,
                                           int affine_2x_minus_1, float* mix_cache, int mix_mode, void* stream) {
  return launch_stft_compress(xt, mix, window, feat, dtype, B, n_src, T, W, exponent, factor, affine_2x_minus_1, mix_cache,
                              mix_mode, stream);
}

extern "C" int ditsep_expand_istft(const float* spec, const float* window, float* out, int B, int n_src, int64_t T,
                                   int W, float exponent, float factor, void* stream) {
  DITSEP_REQUIRE(spec && window && out, "expand_istft: null pointer");
  DITSEP_REQUIRE(B > 0 && T > 0 && n_src >= 1 && n_src <= 4, "expand_istft: bad sizes");
  const int n_frames = 1 + (int)((T + (kNfft - kHop)) / kHop);
  DITSEP_REQUIRE(W >= n_frames, "expand_istft: W=%d < %d frames", W, n_frames);
  DITSEP_REQUIRE(exponent >= 0.f && factor != 0.f, "expand_istft: bad exponent/factor (exponent 0 = log1p transform)");
  size_t smem = kNfft * 8 + (kNfft + 2) * 4 + (size_t)kWarps * 2 * kBins * 8 + (size_t)n_src * kIF * kNfft * 4;
  dim3 grid(ceil_div(T + kHalf, (int64_t)kHop * kIT), B);
  DITSEP_CUDA(cudaFuncSetAttribute(expand_istft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  expand_istft_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(spec, window, out, n_src, T, W, n_frames, exponent,
                                                                 factor);
  return check_launch("expand_istft_kernel");
}
