// Fused SDE drift/diffusion + predictor / Langevin-corrector updates.
//
// One vectorised, coalesced elementwise kernel per update: every thread owns 4
// consecutive time samples of ALL sources of one mixture (the mixing matrix couples
// sources at the same sample), reads x and the score once, draws (or reads) the noise,
// keeps the per-mixture step coefficients in registers and writes x (and x_mean only
// when asked).  Replaces ~25 ATen kernels + 2 Philox kernels per update in the
// reference.
//
// Closed forms (A = 11^T/n averaging matrix, Pn = I - A; bar = mean over sources):
//   MixSDE  L(t) = sqrt(ev1) A + sqrt(ev2) Pn,   L L = ev1 A + ev2 Pn     (sdes.py:298-322)
//   ald2    x_mean = x + 2 snr^2 (ev1 sbar + ev2 (s - sbar))
//           x      = x_mean + 2 snr (sqrt(ev1) zbar + sqrt(ev2) (z - zbar))   (correctors.py:102-121)
//   rd      x_mean = x + lambda dt (x - xbar) + g^2 dt s ; x = x_mean + g sqrt(dt) z
//                                                          (predictors.py:60-66, sdes.py:94-108,165-173,277-286)
//   OUVE    ald: eps = 2 (snr std)^2, x_mean = x + eps s, x = x_mean + sqrt(2 eps) z  (correctors.py:70-84)
//           rd : x_mean = x - theta (y - x) dt + g^2 dt s ; x = x_mean + g sqrt(dt) z (sdes.py:652-660)
// All are instances of
//   x_mean_i = a0 x_i + a1 xbar + a2 s_i + a3 sbar + a4 y ;  x_i = x_mean_i + b0 z_i + b1 zbar.
#include "common.cuh"

namespace ditsep {

struct Coef {
  float a0, a1, a2, a3, a4, b0, b1;
};

// ---- Philox4x32-10 --------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

// Box-Muller on 4 uniform 32-bit words -> 4 standard normals.
__device__ __forceinline__ float4 box_muller4(uint4 r) {
  const float k = 2.3283064365386963e-10f;  // 2^-32
  float u0 = ((float)r.x + 0.5f) * k, u1 = ((float)r.y + 0.5f) * k;
  float u2 = ((float)r.z + 0.5f) * k, u3 = ((float)r.w + 0.5f) * k;
  // u in (0, 1]: (x + 0.5) * 2^-32 can round to 1.0f, never to 0.  __logf is only specified to an ABSOLUTE error
  // near 1 (2^-21.41 on [0.5, 2]), so it may return a tiny positive value for u = 1 - 2^-24: clamp, or the
  // square root is NaN and poisons the whole mixture.
  float r0 = sqrtf(fmaxf(0.0f, -2.0f * __logf(u0))), r1 = sqrtf(fmaxf(0.0f, -2.0f * __logf(u2)));
  float s0, c0, s1, c1;
  __sincosf(6.283185307179586f * u1, &s0, &c0);
  __sincosf(6.283185307179586f * u3, &s1, &c1);
  return make_float4(r0 * c0, r0 * s0, r1 * c1, r1 * s1);
}

// 4 standard normals for state elements [4q, 4q+4) of draw `draw`.
__device__ __forceinline__ float4 philox_normal4(uint64_t q, uint64_t seed, uint64_t draw) {
  return box_muller4(philox4x32_10(make_uint4((uint32_t)q, (uint32_t)(q >> 32), (uint32_t)draw, (uint32_t)(draw >> 32)),
                                   make_uint2((uint32_t)seed, (uint32_t)(seed >> 32))));
}

// ---- per-mixture coefficients ----------------------------------------------------------
enum Mode { kCorrector = 0, kPredictor = 1, kPrior = 2, kLangevin = 3, kPerturb = 4 };

__device__ __forceinline__ Coef make_coef(const ditsep_sde_params& p, int mode, float t) {
  Coef c = {1.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (mode == kLangevin) {  // t = the batch-wide step size eps: x_mean = x + eps s, x = x_mean + sqrt(2 eps) z
    c.a2 = t;
    c.b0 = sqrtf(2.0f * t);
    return c;
  }
  const float ratio = p.sigma_max / p.sigma_min;
  const float logsig = logf(ratio);
  const float dt = 1.0f / (float)p.N;
  if (p.sde == DITSEP_SDE_MIX) {
    const float lam = p.p0;
    float srp = powf(ratio, 2.0f * t);
    float mult = p.sigma_min * p.sigma_min;
    float ev1 = mult * (srp - 1.0f);
    float ev2 = mult * (srp - expf(-2.0f * lam * t)) / (1.0f + lam / logsig);
    float r1 = sqrtf(ev1), r2 = sqrtf(ev2);
    if (mode == kCorrector) {
      float st = 2.0f * p.snr * p.snr;
      c.a2 = st * ev2;
      c.a3 = st * (ev1 - ev2);
      c.b0 = 2.0f * p.snr * r2;
      c.b1 = 2.0f * p.snr * (r1 - r2);
    } else if (mode == kPredictor) {
      float g = p.sigma_min * powf(ratio, t) * sqrtf(2.0f * logsig);
      float G = g * sqrtf(dt);
      c.a0 = 1.0f + lam * dt;
      c.a1 = -lam * dt;
      c.a2 = G * G;
      c.b0 = G;
    } else if (mode == kPerturb) {  // marginal sample: (A + e^{-lam t} Pn) x0 + L(t) z   (sdes.py:288-326)
      const float e = expf(-lam * t);
      c.a0 = e;
      c.a1 = 1.0f - e;
      c.b0 = r2;
      c.b1 = r1 - r2;
    } else {  // prior: 0.5 y + L(T) z   (t = T = 1 passed by the host)
      c.a0 = 0.f;
      c.a4 = 0.5f;
      c.b0 = r2;
      c.b1 = r1 - r2;
    }
  } else {
    const float th = p.p0;
    float std = sqrtf(p.sigma_min * p.sigma_min * expf(-2.0f * th * t) * (expf(2.0f * (th + logsig) * t) - 1.0f) *
                      logsig / (th + logsig));
    if (mode == kCorrector) {
      float eps = 2.0f * (p.snr * std) * (p.snr * std);
      c.a2 = eps;
      c.b0 = sqrtf(2.0f * eps);
    } else if (mode == kPredictor) {
      float g = p.sigma_min * powf(ratio, t) * sqrtf(2.0f * logsig);
      float G = g * sqrtf(dt);
      c.a0 = 1.0f + th * dt;
      c.a4 = -th * dt;
      c.a2 = G * G;
      c.b0 = G;
    } else if (mode == kPerturb) {  // e^{-theta t} x0 + (1 - e^{-theta t}) y + std(t) z   (sdes.py:662-685)
      const float e = expf(-th * t);
      c.a0 = e;
      c.a4 = 1.0f - e;
      c.b0 = std;
    } else {
      c.a0 = 0.f;
      c.a4 = 1.0f;
      c.b0 = std;
    }
  }
  return c;
}

constexpr int kMaxSrc = 4;

// grid: (ceil(L/4/256), B); block 256.  L % 4 == 0 required by the host wrapper for the
// vector path; a scalar tail handles the remainder.
template <int NS>
__global__ void __launch_bounds__(256) pc_update_kernel(ditsep_sde_params p, int mode, float* __restrict__ x,
                                                        const float* __restrict__ score,
                                                        const float* __restrict__ noise, uint64_t seed,
                                                        uint64_t draw, const float* __restrict__ t,
                                                        const float* __restrict__ y, float* __restrict__ x_mean,
                                                        int64_t L, const int* __restrict__ lengths) {
  const int b = blockIdx.y;
  // ragged batches: row b is a zero-padded signal of lengths[b] <= L samples; its state is kept at exactly zero past
  // its end, so the STFT of the next score evaluation sees the same frames as a batch-1 run of that mixture alone
  const int64_t len = lengths ? (int64_t)lengths[b] : L;
  const float tb = (mode == kPrior) ? 1.0f : (mode == kLangevin ? t[0] : t[b]);
  const Coef c = make_coef(p, mode, tb);
  // PriorMixSDE (sdes.py:453-472,519-531): the diffusion carries a per-mixture, per-sample envelope m[b][l], i.e. the
  // score terms scale with m^2 and the noise terms with m
  const float* __restrict__ env = p.elem_scale ? p.elem_scale + (int64_t)b * L : nullptr;
  const int64_t nvec = (L + 3) / 4;
  const float inv_n = 1.0f / (float)NS;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t l0 = q * 4;
    const bool full = (l0 + 4 <= L) && ((L & 3) == 0);
    float xv[NS][4], sv[NS][4], zv[NS][4], yv[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      const int64_t off = ((int64_t)b * NS + i) * L + l0;
      if (full) {
        if (mode != kPrior) {
          float4 a = *reinterpret_cast<const float4*>(x + off);
          xv[i][0] = a.x; xv[i][1] = a.y; xv[i][2] = a.z; xv[i][3] = a.w;
          if (score) {
            float4 s = __ldg(reinterpret_cast<const float4*>(score + off));
            sv[i][0] = s.x; sv[i][1] = s.y; sv[i][2] = s.z; sv[i][3] = s.w;
          } else {
            sv[i][0] = sv[i][1] = sv[i][2] = sv[i][3] = 0.f;
          }
        }
        if (noise) {
          float4 z = __ldg(reinterpret_cast<const float4*>(noise + off));
          zv[i][0] = z.x; zv[i][1] = z.y; zv[i][2] = z.z; zv[i][3] = z.w;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          bool ok = l0 + j < L;
          xv[i][j] = (ok && mode != kPrior) ? x[off + j] : 0.f;
          sv[i][j] = (ok && mode != kPrior && score) ? score[off + j] : 0.f;
          zv[i][j] = (ok && noise) ? noise[off + j] : 0.f;
        }
      }
      if (mode == kPrior) {
#pragma unroll
        for (int j = 0; j < 4; ++j) xv[i][j] = sv[i][j] = 0.f;
      }
      if (!noise) {
        // element index in the (B, NS, L) state, in units of 4
        float4 z = philox_normal4((uint64_t)(((int64_t)b * NS + i) * nvec + q), seed, draw);
        zv[i][0] = z.x; zv[i][1] = z.y; zv[i][2] = z.z; zv[i][3] = z.w;
      }
    }
    if (y) {
      const int64_t off = (int64_t)b * L + l0;
      if (full) {
        float4 a = __ldg(reinterpret_cast<const float4*>(y + off));
        yv[0] = a.x; yv[1] = a.y; yv[2] = a.z; yv[3] = a.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) yv[j] = (l0 + j < L) ? y[off + j] : 0.f;
      }
    }
    float xm[NS][4], xo[NS][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float xb = 0.f, sb = 0.f, zb = 0.f;
#pragma unroll
      for (int i = 0; i < NS; ++i) {
        xb += xv[i][j];
        sb += sv[i][j];
        zb += zv[i][j];
      }
      xb *= inv_n; sb *= inv_n; zb *= inv_n;
      const bool live = l0 + j < len;
      const float ev = (env && l0 + j < L) ? __ldg(env + l0 + j) : 1.0f;
      const float ev2 = ev * ev;
#pragma unroll
      for (int i = 0; i < NS; ++i) {
        float m = c.a0 * xv[i][j] + c.a1 * xb + ev2 * (c.a2 * sv[i][j] + c.a3 * sb) + c.a4 * yv[j];
        xm[i][j] = live ? m : 0.f;
        xo[i][j] = live ? m + ev * (c.b0 * zv[i][j] + c.b1 * zb) : 0.f;
      }
    }
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      const int64_t off = ((int64_t)b * NS + i) * L + l0;
      if (full) {
        *reinterpret_cast<float4*>(x + off) = make_float4(xo[i][0], xo[i][1], xo[i][2], xo[i][3]);
        if (x_mean) *reinterpret_cast<float4*>(x_mean + off) = make_float4(xm[i][0], xm[i][1], xm[i][2], xm[i][3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (l0 + j < L) {
            x[off + j] = xo[i][j];
            if (x_mean) x_mean[off + j] = xm[i][j];
          }
      }
    }
  }
}

static int launch_pc(const ditsep_sde_params* hp, int mode, float* x, const float* score, const float* noise,
                     uint64_t seed, uint64_t draw, const float* t, const float* y, float* x_mean, int B, int64_t L,
                     const int* lengths, void* stream) {
  DITSEP_REQUIRE(hp != nullptr && x != nullptr, "pc_step: null params or state");
  DITSEP_REQUIRE(hp->n_src >= 1 && hp->n_src <= kMaxSrc, "pc_step: n_src=%d unsupported (1..%d)", hp->n_src, kMaxSrc);
  DITSEP_REQUIRE(mode == kLangevin || hp->sde == DITSEP_SDE_MIX || hp->sde == DITSEP_SDE_OUVE, "pc_step: unknown sde %d", hp->sde);
  DITSEP_REQUIRE(B > 0 && L > 0, "pc_step: empty state B=%d L=%lld", B, (long long)L);
  DITSEP_REQUIRE(mode == kPrior || ((score != nullptr || mode == kPerturb) && t != nullptr), "pc_step: score/t required");
  DITSEP_REQUIRE(mode != kPerturb || hp->elem_scale == nullptr, "pc_perturb: PriorMixSDE's envelope is not supported here");
  DITSEP_REQUIRE(!(hp->sde == DITSEP_SDE_MIX && mode == kPrior && hp->n_src != 2 && hp->elem_scale == nullptr),
                 "MixSDE.prior_sampling supports exactly 2 sources (reference sdes.py:346), got %d", hp->n_src);
  DITSEP_REQUIRE((mode == kCorrector) || (mode == kLangevin) || y != nullptr ||
                     (hp->sde == DITSEP_SDE_MIX && (mode == kPredictor || mode == kPerturb)),
                 "pc_step: y required");
  const int64_t nvec = (L + 3) / 4;
  int gx = (int)((nvec + 255) / 256);
  int maxgx = (kNumSMs * 8 + B - 1) / B;  // ~8 blocks of 256 threads per SM across the batch
  if (gx > maxgx) gx = maxgx;
  if (gx < 1) gx = 1;
  dim3 grid(gx, B), block(256);
  cudaStream_t st = (cudaStream_t)stream;
  const float* yy = ((hp->sde == DITSEP_SDE_MIX && mode != kPrior) || mode == kLangevin) ? nullptr : y;
  switch (hp->n_src) {
    case 1: pc_update_kernel<1><<<grid, block, 0, st>>>(*hp, mode, x, score, noise, seed, draw, t, yy, x_mean, L, lengths); break;
    case 2: pc_update_kernel<2><<<grid, block, 0, st>>>(*hp, mode, x, score, noise, seed, draw, t, yy, x_mean, L, lengths); break;
    case 3: pc_update_kernel<3><<<grid, block, 0, st>>>(*hp, mode, x, score, noise, seed, draw, t, yy, x_mean, L, lengths); break;
    default: pc_update_kernel<4><<<grid, block, 0, st>>>(*hp, mode, x, score, noise, seed, draw, t, yy, x_mean, L, lengths); break;
  }
  return check_launch("pc_update_kernel");
}

__global__ void philox_normal_kernel(float* out, int64_t n, uint64_t seed, uint64_t draw) {
  int64_t nvec = (n + 3) / 4;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += (int64_t)gridDim.x * blockDim.x) {
    float4 z = philox_normal4((uint64_t)q, seed, draw);
    float v[4] = {z.x, z.y, z.z, z.w};
    for (int j = 0; j < 4; ++j)
      if (q * 4 + j < n) out[q * 4 + j] = v[j];
  }
}

__global__ void box_muller_kernel(const uint32_t* bits, float* out, int64_t nvec) {
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += (int64_t)gridDim.x * blockDim.x) {
    const uint4 r = *reinterpret_cast<const uint4*>(bits + 4 * q);
    *reinterpret_cast<float4*>(out + 4 * q) = box_muller4(r);
  }
}

}  // namespace ditsep

using namespace ditsep;

extern "C" int ditsep_pc_corrector(const ditsep_sde_params* host_p, float* x, const float* score, const float* noise,
                                   uint64_t seed, uint64_t draw, const float* t, float* x_mean, int B, int64_t L,
                                   const int* lengths, void* stream) {
  return launch_pc(host_p, kCorrector, x, score, noise, seed, draw, t, nullptr, x_mean, B, L, lengths, stream);
}

extern "C" int ditsep_pc_predictor(const ditsep_sde_params* host_p, float* x, const float* score, const float* noise,
                                   uint64_t seed, uint64_t draw, const float* t, const float* y, float* x_mean, int B,
                                   int64_t L, const int* lengths, void* stream) {
  return launch_pc(host_p, kPredictor, x, score, noise, seed, draw, t, y, x_mean, B, L, lengths, stream);
}

extern "C" int ditsep_pc_prior(const ditsep_sde_params* host_p, float* x, const float* noise, uint64_t seed,
                               uint64_t draw, const float* y, int B, int64_t L, const int* lengths, void* stream) {
  return launch_pc(host_p, kPrior, x, nullptr, noise, seed, draw, nullptr, y, nullptr, B, L, lengths, stream);
}

extern "C" int ditsep_pc_perturb(const ditsep_sde_params* host_p, float* x, const float* noise, uint64_t seed,
                                 uint64_t draw, const float* t, const float* y, float* x_mean, int B, int64_t L,
                                 const int* lengths, void* stream) {
  return launch_pc(host_p, kPerturb, x, nullptr, noise, seed, draw, t, y, x_mean, B, L, lengths, stream);
}

extern "C" int ditsep_pc_langevin(int n_src, float* x, const float* score, const float* noise, const float* step,
                                  float* x_mean, int B, int64_t L, const int* lengths, void* stream) {
  DITSEP_REQUIRE(noise != nullptr && step != nullptr, "pc_langevin: the step size depends on the noise draw: pass both");
  ditsep_sde_params p = {};
  p.sde = DITSEP_SDE_MIX;
  p.n_src = n_src;
  p.sigma_min = 1.f;
  p.sigma_max = 2.f;
  p.N = 1;
  return launch_pc(&p, kLangevin, x, score, noise, 0, 0, step, nullptr, x_mean, B, L, lengths, stream);
}

// ---- PriorMixSDE's energy envelope -------------------------------------------------------------
// out[b][l] = 0.5 sqrt(max(1e-4, (1/k) sum_{j = l - k/2}^{l - k/2 + k - 1} mix[b][j]^2)), zeros outside [0, L)
// (avg_pool1d(mix^2, k, stride 1, padding k // 2) with the padding counted, sdes.py:480-492).  One block per 1024
// outputs: the squares it needs (1024 + k - 1) go to shared memory once, every thread then walks its window.
namespace ditsep {
__global__ void __launch_bounds__(256) energy_envelope_kernel(const float* __restrict__ mix, float* __restrict__ out,
                                                              int64_t L, int k) {
  extern __shared__ float sq[];
  const int b = blockIdx.y;
  const int64_t l0 = (int64_t)blockIdx.x * 1024;
  const int n = 1024 + k - 1;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int64_t j = l0 - k / 2 + i;
    const float v = (j >= 0 && j < L) ? mix[(int64_t)b * L + j] : 0.f;
    sq[i] = v * v;
  }
  __syncthreads();
  const float inv = 1.0f / (float)k;
  for (int o = threadIdx.x; o < 1024; o += blockDim.x) {
    if (l0 + o >= L) break;
    float a = 0.f;
    for (int j = 0; j < k; ++j) a += sq[o + j];
    out[(int64_t)b * L + l0 + o] = 0.5f * sqrtf(fmaxf(a * inv, 1e-4f));
  }
}
}  // namespace ditsep

extern "C" int ditsep_energy_envelope(const float* mix, float* out, int B, int64_t L, int avg_len, void* stream) {
  DITSEP_REQUIRE(mix && out && B > 0 && L > 0, "energy_envelope: empty input");
  DITSEP_REQUIRE(avg_len >= 1 && avg_len <= 4096, "energy_envelope: avg_len=%d (1..4096)", avg_len);
  dim3 grid((unsigned)((L + 1023) / 1024), (unsigned)B);
  const size_t smem = (size_t)(1024 + avg_len - 1) * sizeof(float);
  energy_envelope_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(mix, out, L, avg_len);
  return check_launch("energy_envelope_kernel");
}

// ---- Schroedinger-bridge sampler step ----------------------------------------------------------
// One launch per sampler step (get_sb_sampler, sdes/__init__.py:284-389) with DiffSepOU's EDM preconditioning
// (diffsep.py:813-828) folded in.  coef row of sample b: [w_prev, w_est, w_y, w_z, c_skip, c_out, f_div, c_in_next]:
//   est = c_skip x + c_out (f / f_div)                    (c_skip = 0, c_out = f_div = 1: est = f)
//   x'  = (w_prev x + w_est est) + w_y y | w_z z | both     (the reference's evaluation order, unfused roundings)
//   x_in = c_in_next x',  y_in = c_in_next y               (the NEXT network evaluation's scaled inputs)
namespace ditsep {
__global__ void __launch_bounds__(256) sb_step_kernel(const float* __restrict__ x, const float* __restrict__ f,
                                                      const float* __restrict__ y, const float* __restrict__ noise,
                                                      int use_noise, uint64_t seed, uint64_t draw,
                                                      const float* __restrict__ coef, float* __restrict__ x_out,
                                                      float* __restrict__ x_in, float* __restrict__ y_in,
                                                      float* __restrict__ est_out, int n_src, int64_t L,
                                                      const int* __restrict__ lengths) {
  const int b = blockIdx.y;
  const int64_t len = lengths ? (int64_t)lengths[b] : L;
  const float* cb = coef + (int64_t)b * 8;
  const float w_prev = cb[0], w_est = cb[1], w_y = cb[2], w_z = cb[3];
  const float c_skip = cb[4], c_out = cb[5], f_div = cb[6], c_in = cb[7];
  const int64_t nvec = (L + 3) / 4;
  const bool vec = (L & 3) == 0;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t l0 = q * 4;
    float yv[4] = {0.f, 0.f, 0.f, 0.f};
    if (y) {
      const int64_t off = (int64_t)b * L + l0;
      if (vec) {
        float4 a = __ldg(reinterpret_cast<const float4*>(y + off));
        yv[0] = a.x; yv[1] = a.y; yv[2] = a.z; yv[3] = a.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) yv[j] = (l0 + j < L) ? y[off + j] : 0.f;
      }
      if (y_in) {
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] = (l0 + j < len) ? __fmul_rn(c_in, yv[j]) : 0.f;
        if (vec) *reinterpret_cast<float4*>(y_in + off) = make_float4(o[0], o[1], o[2], o[3]);
        else
          for (int j = 0; j < 4; ++j)
            if (l0 + j < L) y_in[off + j] = o[j];
      }
    }
    for (int i = 0; i < n_src; ++i) {
      const int64_t off = ((int64_t)b * n_src + i) * L + l0;
      float xv[4], fv[4] = {0.f, 0.f, 0.f, 0.f}, zv[4] = {0.f, 0.f, 0.f, 0.f};
      if (vec) {
        float4 a = *reinterpret_cast<const float4*>(x + off);
        xv[0] = a.x; xv[1] = a.y; xv[2] = a.z; xv[3] = a.w;
        if (f) {
          float4 s = __ldg(reinterpret_cast<const float4*>(f + off));
          fv[0] = s.x; fv[1] = s.y; fv[2] = s.z; fv[3] = s.w;
        }
        if (use_noise && noise) {
          float4 z = __ldg(reinterpret_cast<const float4*>(noise + off));
          zv[0] = z.x; zv[1] = z.y; zv[2] = z.z; zv[3] = z.w;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const bool ok = l0 + j < L;
          xv[j] = ok ? x[off + j] : 0.f;
          fv[j] = (ok && f) ? f[off + j] : 0.f;
          zv[j] = (ok && use_noise && noise) ? noise[off + j] : 0.f;
        }
      }
      if (use_noise && !noise) {
        float4 z = philox_normal4((uint64_t)(((int64_t)b * n_src + i) * nvec + q), seed, draw);
        zv[0] = z.x; zv[1] = z.y; zv[2] = z.z; zv[3] = z.w;
      }
      float xo[4], eo[4], uo[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool live = l0 + j < len;
        const float est = f ? __fadd_rn(__fmul_rn(c_skip, xv[j]), __fmul_rn(c_out, __fdiv_rn(fv[j], f_div))) : 0.f;
        float v = __fadd_rn(__fmul_rn(w_prev, xv[j]), __fmul_rn(w_est, est));
        if (use_noise != 1) v = __fadd_rn(v, __fmul_rn(w_y, yv[j]));  // 0: ODE update; 2: marginal sample (mean first, then noise)
        if (use_noise != 0) v = __fadd_rn(v, __fmul_rn(w_z, zv[j]));
        eo[j] = live ? est : 0.f;
        xo[j] = live ? v : 0.f;
        uo[j] = live ? __fmul_rn(c_in, v) : 0.f;
      }
      if (vec) {
        if (x_out) *reinterpret_cast<float4*>(x_out + off) = make_float4(xo[0], xo[1], xo[2], xo[3]);
        if (x_in) *reinterpret_cast<float4*>(x_in + off) = make_float4(uo[0], uo[1], uo[2], uo[3]);
        if (est_out) *reinterpret_cast<float4*>(est_out + off) = make_float4(eo[0], eo[1], eo[2], eo[3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (l0 + j < L) {
            if (x_out) x_out[off + j] = xo[j];
            if (x_in) x_in[off + j] = uo[j];
            if (est_out) est_out[off + j] = eo[j];
          }
      }
    }
  }
}
}  // namespace ditsep

extern "C" int ditsep_sb_step(const float* x, const float* f, const float* y, const float* noise, int use_noise,
                              uint64_t seed, uint64_t draw, const float* coef, float* x_out, float* x_in, float* y_in,
                              float* est_out, int B, int n_src, int64_t L, const int* lengths, void* stream) {
  DITSEP_REQUIRE(x != nullptr && coef != nullptr, "sb_step: null state or coefficient table");
  DITSEP_REQUIRE(B > 0 && L > 0 && n_src >= 1 && n_src <= 64, "sb_step: bad shape B=%d n_src=%d L=%lld", B, n_src, (long long)L);
  DITSEP_REQUIRE(x_out || x_in || y_in || est_out, "sb_step: no output requested");
  DITSEP_REQUIRE(use_noise >= 0 && use_noise <= 2, "sb_step: use_noise=%d (0: y term, 1: noise term, 2: both)", use_noise);
  DITSEP_REQUIRE(use_noise == 1 || y != nullptr, "sb_step: the y term (use_noise = 0 or 2) needs the mixture y");
  DITSEP_REQUIRE(!y_in || y != nullptr, "sb_step: y_in requested without y");
  if ((L & 3) == 0) {
    uintptr_t a = reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(f) | reinterpret_cast<uintptr_t>(y) |
                  reinterpret_cast<uintptr_t>(noise) | reinterpret_cast<uintptr_t>(x_out) |
                  reinterpret_cast<uintptr_t>(x_in) | reinterpret_cast<uintptr_t>(y_in) | reinterpret_cast<uintptr_t>(est_out);
    DITSEP_REQUIRE((a & 15) == 0, "sb_step: 16-byte aligned buffers required when L %% 4 == 0");
  }
  const int64_t nvec = (L + 3) / 4;
  int gx = (int)((nvec + 255) / 256);
  int maxgx = (kNumSMs * 8 + B - 1) / B;
  if (gx > maxgx) gx = maxgx;
  if (gx < 1) gx = 1;
  ditsep::sb_step_kernel<<<dim3(gx, B), 256, 0, (cudaStream_t)stream>>>(x, f, y, noise, use_noise, seed, draw, coef, x_out,
                                                                       x_in, y_in, est_out, n_src, L, lengths);
  return check_launch("sb_step_kernel");
}

// ---- denoising score-matching loss (forward value) ---------------------------------------------
// loss[b] = mean over (source, sample) of (L(t_b) pred + z)^2 with L(t) = sqrt(ev1) A + sqrt(ev2) Pn (MixSDE) or std(t) I
// (OUVESDE): DiffSepModel.compute_score_loss (src/diffsep.py:344-358: mult_std, MSELoss(L_score, -z), mean over the last
// two axes).  z is read from `noise`, or regenerated from the Philox stream that ditsep_pc_perturb consumed (same seed,
// draw and element numbering), so a perturb -> network -> loss pass never stores the noise.  Two launches with a fixed
// reduction order (per-block partial sums in double, then one warp per mixture): bit-reproducible.
namespace ditsep {
constexpr int kLossMaxParts = 256;

template <int NS>
__global__ void __launch_bounds__(256) score_loss_partial_kernel(ditsep_sde_params p, const float* __restrict__ pred,
                                                                 const float* __restrict__ noise, uint64_t seed,
                                                                 uint64_t draw, const float* __restrict__ t,
                                                                 const float* __restrict__ sigma,
                                                                 double* __restrict__ partial, int64_t L) {
  const int b = blockIdx.y;
  Coef c = {1.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (sigma) c.b0 = sigma[b];  // scalar std per mixture given by the caller (SBVESDE: closed form kept on the host)
  else c = make_coef(p, kPerturb, t[b]);
  const int64_t nvec = (L + 3) / 4;
  const bool vec = (L & 3) == 0;
  const float inv_n = 1.0f / (float)NS;
  double acc = 0.0;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t l0 = q * 4;
    float pv[NS][4], zv[NS][4];
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      const int64_t off = ((int64_t)b * NS + i) * L + l0;
      if (vec) {
        float4 a = __ldg(reinterpret_cast<const float4*>(pred + off));
        pv[i][0] = a.x; pv[i][1] = a.y; pv[i][2] = a.z; pv[i][3] = a.w;
        if (noise) {
          float4 z = __ldg(reinterpret_cast<const float4*>(noise + off));
          zv[i][0] = z.x; zv[i][1] = z.y; zv[i][2] = z.z; zv[i][3] = z.w;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const bool ok = l0 + j < L;
          pv[i][j] = ok ? pred[off + j] : 0.f;
          zv[i][j] = (ok && noise) ? noise[off + j] : 0.f;
        }
      }
      if (!noise) {
        float4 z = philox_normal4((uint64_t)(((int64_t)b * NS + i) * nvec + q), seed, draw);
        zv[i][0] = z.x; zv[i][1] = z.y; zv[i][2] = z.z; zv[i][3] = z.w;
      }
    }
    float part = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (l0 + j >= L) continue;
      float pb = 0.f;
#pragma unroll
      for (int i = 0; i < NS; ++i) pb += pv[i][j];
      pb *= inv_n;
#pragma unroll
      for (int i = 0; i < NS; ++i) {
        const float v = c.b0 * pv[i][j] + c.b1 * pb + zv[i][j];
        part += v * v;
      }
    }
    acc += (double)part;
  }
  __shared__ double red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w];
    partial[(int64_t)b * gridDim.x + blockIdx.x] = s;
  }
}

__global__ void __launch_bounds__(32) score_loss_final_kernel(const double* __restrict__ partial, float* __restrict__ loss,
                                                              int parts, double inv_count) {
  const int b = blockIdx.x;
  double s = 0.0;
  for (int i = threadIdx.x; i < parts; i += 32) s += partial[(int64_t)b * parts + i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (threadIdx.x == 0) loss[b] = (float)(s * inv_count);
}
}  // namespace ditsep

static int launch_score_loss(const ditsep_sde_params* hp, const float* pred, const float* noise, uint64_t seed,
                             uint64_t draw, const float* t, const float* sigma, float* loss, double* workspace, int B,
                             int64_t L, void* stream) {
  DITSEP_REQUIRE(hp && pred && (t || sigma) && loss && workspace, "score_loss: null argument");
  DITSEP_REQUIRE(sigma || hp->sde == DITSEP_SDE_MIX || hp->sde == DITSEP_SDE_OUVE, "score_loss: unknown sde %d", hp->sde);
  DITSEP_REQUIRE(hp->n_src >= 1 && hp->n_src <= kMaxSrc, "score_loss: n_src=%d unsupported (1..%d)", hp->n_src, kMaxSrc);
  DITSEP_REQUIRE(hp->elem_scale == nullptr, "score_loss: PriorMixSDE's envelope is not supported here");
  DITSEP_REQUIRE(B > 0 && L > 0, "score_loss: empty input B=%d L=%lld", B, (long long)L);
  const int64_t nvec = (L + 3) / 4;
  int parts = (int)((nvec + 1023) / 1024);  // ~4 vectors per thread
  if (parts > kLossMaxParts) parts = kLossMaxParts;
  if (parts < 1) parts = 1;
  dim3 grid(parts, B);
  cudaStream_t st = (cudaStream_t)stream;
  switch (hp->n_src) {
    case 1: score_loss_partial_kernel<1><<<grid, 256, 0, st>>>(*hp, pred, noise, seed, draw, t, sigma, workspace, L); break;
    case 2: score_loss_partial_kernel<2><<<grid, 256, 0, st>>>(*hp, pred, noise, seed, draw, t, sigma, workspace, L); break;
    case 3: score_loss_partial_kernel<3><<<grid, 256, 0, st>>>(*hp, pred, noise, seed, draw, t, sigma, workspace, L); break;
    default: score_loss_partial_kernel<4><<<grid, 256, 0, st>>>(*hp, pred, noise, seed, draw, t, sigma, workspace, L); break;
  }
  if (int rc = check_launch("score_loss_partial_kernel")) return rc;
  score_loss_final_kernel<<<B, 32, 0, st>>>(workspace, loss, parts, 1.0 / ((double)hp->n_src * (double)L));
  return check_launch("score_loss_final_kernel");
}

extern "C" int ditsep_score_loss(const ditsep_sde_params* hp, const float* pred, const float* noise, uint64_t seed,
                                 uint64_t draw, const float* t, float* loss, double* workspace, int B, int64_t L,
                                 void* stream) {
  DITSEP_REQUIRE(t != nullptr, "score_loss: null time vector");
  return launch_score_loss(hp, pred, noise, seed, draw, t, nullptr, loss, workspace, B, L, stream);
}

extern "C" int ditsep_score_loss_sigma(const float* pred, const float* noise, uint64_t seed, uint64_t draw,
                                       const float* sigma, float* loss, double* workspace, int B, int n_src, int64_t L,
                                       void* stream) {
  DITSEP_REQUIRE(sigma != nullptr, "score_loss_sigma: null sigma");
  ditsep_sde_params p = {};
  p.sde = DITSEP_SDE_OUVE;
  p.n_src = n_src;
  p.sigma_min = 1.f;
  p.sigma_max = 2.f;
  p.N = 1;
  return launch_score_loss(&p, pred, noise, seed, draw, nullptr, sigma, loss, workspace, B, L, stream);
}

extern "C" int ditsep_philox_normal(float* out, int64_t n, uint64_t seed, uint64_t draw, void* stream) {
  DITSEP_REQUIRE(out != nullptr && n > 0, "philox_normal: empty output");
  int64_t nvec = (n + 3) / 4;
  int grid = (int)((nvec + 255) / 256);
  if (grid > kNumSMs * 8) grid = kNumSMs * 8;
  philox_normal_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(out, n, seed, draw);
  return check_launch("philox_normal_kernel");
}

extern "C" int ditsep_box_muller(const uint32_t* bits, float* out, int64_t n, void* stream) {
  DITSEP_REQUIRE(bits && out && n > 0 && n % 4 == 0, "box_muller: n=%lld must be a positive multiple of 4", (long long)n);
  DITSEP_REQUIRE(((reinterpret_cast<uintptr_t>(bits) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
                 "box_muller: 16-byte alignment required");
  const int64_t nvec = n / 4;
  int grid = (int)((nvec + 255) / 256);
  if (grid > kNumSMs * 8) grid = kNumSMs * 8;
  box_muller_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(bits, out, nvec);
  return check_launch("box_muller_kernel");
}
