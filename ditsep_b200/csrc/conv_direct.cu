// CUDA-core implicit-GEMM convolution (3x3 pad 1 / 1x1, stride 1, NHWC), fp32 accumulate.
//
// Role: (1) the fp32 score network (parity target rel-L2 <= 1e-3 needs fp32/tf32
// arithmetic, BASELINE.json north_star) and (2) the narrow layers whose channel counts do
// not fill a tensor-core tile (input conv 6->nf, Combine 1x1 6->C).  The wide bf16 layers go
// through the tcgen05 kernel in conv_tc.cu.  Same fused epilogue as there:
//   out = out_scale * (conv + bias[co] + bias_b[b,co] + residual)
// (res-block tail, layerspp.py:314-328; Combine sum layerspp.py:57-62).
//
// GEMM view: M = B*H*W pixels, N = Cout, K = taps*Cin with k = tap*Cin + ci; weights are
// [Cout][taps][Cin] fp32 (K contiguous).  64x64 block tile, BK = 16, 256 threads, 4x4
// micro-tile per thread, A gathered on the fly (zero outside the image).
#include "common.cuh"

namespace ditsep {

constexpr int BM = 64, BN = 64, BK = 16;

template <typename TI, typename TO, typename TR>
__global__ void __launch_bounds__(256) conv_direct_kernel(const TI* __restrict__ in, int B, int H, int W, int Cin,
                                                          int ld_in, const float* __restrict__ w,
                                                          const float* __restrict__ bias,
                                                          const float* __restrict__ bias_b, int ld_bias_b,
                                                          const TR* __restrict__ residual, int ld_res, float out_scale,
                                                          int Cout, int ksize, TO* __restrict__ out, int ld_out,
                                                          float* __restrict__ stats, int ld_stats) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  __shared__ float red[2][BN];
  const int taps = ksize * ksize, K = taps * Cin, pad = ksize / 2;
  const int64_t M = (int64_t)B * H * W;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int tid = threadIdx.x;
  const int tx = tid % 16, ty = tid / 16;  // tx -> n micro-tile, ty -> m micro-tile

  // A gather assignment: 64 pixels x 16 k = 1024 elements, 4 per thread: k = tid % 16 (channels are
  // contiguous in NHWC, so half-warps read 16 consecutive channels), pixel = tid / 16 + 16 j.
  const int a_k = tid % BK, a_p0 = tid / BK;
  int ab[4], ay[4], ax[4];
  bool a_ok[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int64_t am = m0 + a_p0 + 16 * j;
    a_ok[j] = am < M;
    ab[j] = ay[j] = ax[j] = 0;
    if (a_ok[j]) {
      ax[j] = (int)(am % W);
      int64_t r = am / W;
      ay[j] = (int)(r % H);
      ab[j] = (int)(r / H);
    }
  }
  // B load assignment: 64 cout x 16 k, 4 per thread: k = tid % 16, co = tid / 16 + 16 j
  const int b_k = tid % BK, b_c0 = tid / BK;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += BK) {
    {
      const int k = k0 + a_k;
      const int tap = k / Cin, ci = k - tap * Cin;
      const int dy = tap / ksize - pad, dx = tap % ksize - pad;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float v = 0.f;
        const int yy = ay[j] + dy, xx = ax[j] + dx;
        if (a_ok[j] && k < K && yy >= 0 && yy < H && xx >= 0 && xx < W)
          v = to_f<TI>(in[(((int64_t)ab[j] * H + yy) * W + xx) * ld_in + ci]);
        As[a_k][a_p0 + 16 * j] = v;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = n0 + b_c0 + 16 * j, k = k0 + b_k;
      Bs[b_k][b_c0 + 16 * j] = (co < Cout && k < K) ? w[(int64_t)co * K + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], bb[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bb[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }
  // GroupNorm statistics of the stored output (per image, per channel): block-level reduction in
  // shared memory when the 64-pixel tile lies inside one image, else per-thread reds.
  const int64_t HWl = (int64_t)H * W;
  const int64_t m_last = (m0 + BM - 1 < M ? m0 + BM - 1 : M - 1);
  const bool one_image = stats != nullptr && (m0 / HWl) == (m_last / HWl);
  if (stats) {
    for (int i = tid; i < 2 * BN; i += 256) (&red[0][0])[i] = 0.f;
    __syncthreads();
  }
  float cs[4] = {0.f, 0.f, 0.f, 0.f}, cq[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + ty * 4 + i;
    if (m >= M) continue;
    const int b = (int)(m / HWl);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = n0 + tx * 4 + j;
      if (co >= Cout) continue;
      float v = acc[i][j];
      if (bias) v += bias[co];
      if (bias_b) v += bias_b[(int64_t)b * ld_bias_b + co];
      if (residual) v += to_f<TR>(residual[m * ld_res + co]);
      const TO o = from_f<TO>(v * out_scale);
      out[m * ld_out + co] = o;
      if (stats) {
        const float r = to_f<TO>(o);
        if (one_image) {
          cs[j] += r;
          cq[j] += r * r;
        } else {
          atomicAdd(stats + ((int64_t)b * ld_stats + co) * 2, r);
          atomicAdd(stats + ((int64_t)b * ld_stats + co) * 2 + 1, r * r);
        }
      }
    }
  }
  if (one_image) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      atomicAdd(&red[0][tx * 4 + j], cs[j]);
      atomicAdd(&red[1][tx * 4 + j], cq[j]);
    }
    __syncthreads();
    const int b = (int)(m0 / HWl);
    for (int i = tid; i < 2 * BN; i += 256) {
      const int which = i / BN, c = i % BN;
      if (n0 + c < Cout) atomicAdd(stats + ((int64_t)b * ld_stats + n0 + c) * 2 + which, red[which][c]);
    }
  }
}

// Narrow 1x1 convolution (Combine: Conv_0 of the 2(n+1)-channel input pyramid, layerspp.py:57-62, fused with its sum):
// out[p][co] = out_scale * (sum_ci w[co][ci] x[p][ci] + bias[co] + residual[p][co]).  Pure HBM streaming: a thread owns
// 8 output channels of one pixel (one 16-byte residual load and store); the CIN <= 8 input values of a pixel are read
// by the Cout/8 threads that share it (L1 broadcast) and the filter sits in shared memory.  The GEMM-tiled kernel above
// spent 146 us on the 128x128 level at batch 32 (r1n) where the bytes allow ~25 us.
template <typename TO, int CIN>
__global__ void __launch_bounds__(256) conv1x1_narrow_kernel(const float* __restrict__ in, int64_t npix, int ld_in,
                                                             const float* __restrict__ w, const float* __restrict__ bias,
                                                             const TO* __restrict__ residual, int ld_res, float out_scale,
                                                             int Cout, TO* __restrict__ out, int ld_out) {
  extern __shared__ float wsm[];  // [Cout][CIN] then bias[Cout]
  for (int i = threadIdx.x; i < Cout * CIN; i += blockDim.x) wsm[i] = w[i];
  for (int i = threadIdx.x; i < Cout; i += blockDim.x) wsm[Cout * CIN + i] = bias ? bias[i] : 0.f;
  __syncthreads();
  const int cv = Cout / 8;
  const int64_t total = npix * cv;
  // the grid stride is a multiple of cv (host: Cout / 8 divides 256), so a thread keeps ONE channel pack for its whole
  // life: its 8 x CIN filter values and 8 biases live in registers, the loop body is CIN loads + 8 CIN FMAs
  const int v = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) % cv);
  float wreg[8][CIN], breg[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    breg[j] = wsm[Cout * CIN + v * 8 + j];
#pragma unroll
    for (int c = 0; c < CIN; ++c) wreg[j][c] = wsm[(v * 8 + j) * CIN + c];
  }
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t pix = idx / cv;
    float xin[CIN];
#pragma unroll
    for (int c = 0; c < CIN; ++c) xin[c] = __ldg(in + pix * ld_in + c);
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float a = breg[j];
#pragma unroll
      for (int c = 0; c < CIN; ++c) a = fmaf(wreg[j][c], xin[c], a);
      f[j] = a;
    }
    if (residual) {
      float r[8];
      Pack8<TO>::load(residual + pix * ld_res + v * 8, r);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] += r[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= out_scale;
    Pack8<TO>::store(out + pix * ld_out + v * 8, f);
  }
}

}  // namespace ditsep

using namespace ditsep;

extern "C" int ditsep_conv_direct(const void* in, int dtype, int B, int H, int W, int Cin, int ld_in, const float* w,
                                  const float* bias, const float* bias_b, int ld_bias_b, const void* residual,
                                  int ld_res, float out_scale, int Cout, int ksize, void* out, int out_dtype,
                                  int ld_out, void* stream) {
  DITSEP_REQUIRE(in && w && out, "conv_direct: null pointer");
  DITSEP_REQUIRE(ksize == 1 || ksize == 3, "conv_direct: ksize=%d (1 or 3)", ksize);
  DITSEP_REQUIRE(B > 0 && H > 0 && W > 0 && Cin > 0 && Cout > 0, "conv_direct: empty problem");
  DITSEP_REQUIRE(ld_in >= Cin && ld_out >= Cout, "conv_direct: strides smaller than channel counts");
  const int64_t M = (int64_t)B * H * W;
  dim3 grid(ceil_div(M, BM), ceil_div(Cout, BN));
  cudaStream_t st = (cudaStream_t)stream;
  if (ksize == 1 && dtype == DITSEP_F32 && (Cin == 6 || Cin == 8 || Cin == 4) && !bias_b && Cout % 8 == 0 &&
      ld_out % 8 == 0 && 256 % (Cout / 8) == 0 && (!residual || ld_res % 8 == 0) && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
      (!residual || (reinterpret_cast<uintptr_t>(residual) & 15) == 0) && Cout <= 512) {
    const int64_t total = M * (Cout / 8);
    const unsigned g1 = (unsigned)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    const size_t smem = (size_t)(Cout * Cin + Cout) * sizeof(float);
#define LAUNCH1(TO, CIN)                                                                                              \
  conv1x1_narrow_kernel<TO, CIN><<<g1, 256, smem, st>>>((const float*)in, M, ld_in, w, bias, (const TO*)residual, ld_res, \
                                                        out_scale, Cout, (TO*)out, ld_out)
    if (out_dtype == DITSEP_BF16) {
      if (Cin == 6) LAUNCH1(bf16, 6); else if (Cin == 8) LAUNCH1(bf16, 8); else LAUNCH1(bf16, 4);
    } else if (out_dtype == DITSEP_F32) {
      if (Cin == 6) LAUNCH1(float, 6); else if (Cin == 8) LAUNCH1(float, 8); else LAUNCH1(float, 4);
    } else {
      DITSEP_REQUIRE(false, "conv_direct: unknown out dtype %d", out_dtype);
    }
#undef LAUNCH1
    return check_launch("conv1x1_narrow_kernel");
  }
  // residual shares the OUTPUT dtype (it is an activation tensor of the same network)
#define LAUNCH(TI, TO)                                                                                              \
  conv_direct_kernel<TI, TO, TO><<<grid, 256, 0, st>>>((const TI*)in, B, H, W, Cin, ld_in, w, bias, bias_b, ld_bias_b, \
                                                       (const TO*)residual, ld_res, out_scale, Cout, ksize, (TO*)out,  \
                                                       ld_out, nullptr, 0)
  if (dtype == DITSEP_F32 && out_dtype == DITSEP_F32) LAUNCH(float, float);
  else if (dtype == DITSEP_BF16 && out_dtype == DITSEP_BF16) LAUNCH(bf16, bf16);
  else if (dtype == DITSEP_BF16 && out_dtype == DITSEP_F32) LAUNCH(bf16, float);
  else if (dtype == DITSEP_F32 && out_dtype == DITSEP_BF16) LAUNCH(float, bf16);
  else DITSEP_REQUIRE(false, "conv_direct: unknown dtype %d/%d", dtype, out_dtype);
#undef LAUNCH
  return check_launch("conv_direct_kernel");
}
