// Single-head full self-attention over the H*W tokens of one image (AttnBlockpp core,
// layerspp.py:88-92): o = softmax(q k^T / sqrt(C)) v.  S is 16..1280 tokens, C = 64/128/256.
//
// Flash-style (online softmax, never materialises the S x S logits the reference builds):
// one CTA owns 32 queries of one image (4 per warp, processed together so every K/V
// shared-memory read feeds 4 FMAs); K/V tiles of 64 keys are staged in shared memory as
// fp32.  Logits: lane-per-key dot products (no shuffles in the inner loop); P.V: lane-per-
// channel accumulation with P broadcast from shared memory.  The attention FLOPs are 0.15 %
// of the step, so this stays on the CUDA cores; the q/k/v/out projections
// (NIN = 1x1 conv) run on the tensor-core convolution path.
#include <atomic>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "attention_tc.cuh"

namespace ditsep {

constexpr int kQPerWarp = 4;
constexpr int kAttWarps = 8;
constexpr int kQTile = kQPerWarp * kAttWarps;  // 32 queries per CTA
constexpr int kKTile = 64;

// CPL = C / 32 channels per lane.
template <typename T, int CPL>
__global__ void __launch_bounds__(256) attention_kernel(const T* __restrict__ q, const T* __restrict__ k,
                                                        const T* __restrict__ v, int S, int ld, T* __restrict__ o,
                                                        int ld_o, float scale) {
  griddep_launch_dependents();
  griddep_wait();
  constexpr int C = CPL * 32;
  extern __shared__ __align__(16) float sm[];
  float* Ks = sm;                                   // [kKTile][C + 1]
  float* Vs = Ks + kKTile * (C + 1);                // [kKTile][C]
  float* Qs = Vs + kKTile * C;                      // [kQTile][C]
  float* Ps = Qs + kQTile * C;                      // [kAttWarps][kQPerWarp][kKTile]
  const int b = blockIdx.y, q0 = blockIdx.x * kQTile;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const T* qb = q + (int64_t)b * S * ld;
  const T* kb = k + (int64_t)b * S * ld;
  const T* vb = v + (int64_t)b * S * ld;

  for (int idx = threadIdx.x; idx < kQTile * (C / 8); idx += blockDim.x) {
    const int r = idx / (C / 8), cvi = idx % (C / 8);
    float f[8];
    if (q0 + r < S) {
      Pack8<T>::load(qb + (int64_t)(q0 + r) * ld + cvi * 8, f);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] = 0.f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) Qs[r * C + cvi * 8 + i] = f[i] * scale;
  }

  float m[kQPerWarp], l[kQPerWarp], acc[kQPerWarp][CPL];
#pragma unroll
  for (int i = 0; i < kQPerWarp; ++i) {
    m[i] = -INFINITY;
    l[i] = 0.f;
#pragma unroll
    for (int c = 0; c < CPL; ++c) acc[i][c] = 0.f;
  }
  const float* myq = Qs + warp * kQPerWarp * C;
  float* myp = Ps + warp * kQPerWarp * kKTile;

  for (int kt = 0; kt < S; kt += kKTile) {
    __syncthreads();  // previous tile fully consumed (also orders the Qs fill on the first pass)
    for (int idx = threadIdx.x; idx < kKTile * (C / 8); idx += blockDim.x) {
      const int r = idx / (C / 8), cvi = idx % (C / 8);
      float fk[8], fv[8];
      if (kt + r < S) {
        Pack8<T>::load(kb + (int64_t)(kt + r) * ld + cvi * 8, fk);
        Pack8<T>::load(vb + (int64_t)(kt + r) * ld + cvi * 8, fv);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) fk[i] = fv[i] = 0.f;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        Ks[r * (C + 1) + cvi * 8 + i] = fk[i];
        Vs[r * C + cvi * 8 + i] = fv[i];
      }
    }
    __syncthreads();
    // logits for keys (lane, lane + 32) x 4 queries
    float s[kQPerWarp][2];
#pragma unroll
    for (int i = 0; i < kQPerWarp; ++i) s[i][0] = s[i][1] = 0.f;
    const float* k0p = Ks + lane * (C + 1);
    const float* k1p = Ks + (lane + 32) * (C + 1);
#pragma unroll 8
    for (int c = 0; c < C; ++c) {
      const float ka = k0p[c], kb2 = k1p[c];
#pragma unroll
      for (int i = 0; i < kQPerWarp; ++i) {
        const float qv = myq[i * C + c];
        s[i][0] = fmaf(qv, ka, s[i][0]);
        s[i][1] = fmaf(qv, kb2, s[i][1]);
      }
    }
    const bool ok0 = kt + lane < S, ok1 = kt + lane + 32 < S;
    float corr[kQPerWarp];
#pragma unroll
    for (int i = 0; i < kQPerWarp; ++i) {
      float s0 = ok0 ? s[i][0] : -INFINITY, s1 = ok1 ? s[i][1] : -INFINITY;
      float mx = warp_max(fmaxf(s0, s1));
      float mn = fmaxf(m[i], mx);  // finite: every tile has at least one valid key
      float p0 = ok0 ? __expf(s0 - mn) : 0.f, p1 = ok1 ? __expf(s1 - mn) : 0.f;
      corr[i] = __expf(m[i] - mn);
      l[i] = l[i] * corr[i] + warp_sum(p0 + p1);
      m[i] = mn;
      myp[i * kKTile + lane] = p0;
      myp[i * kKTile + lane + 32] = p1;
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < kQPerWarp; ++i)
#pragma unroll
      for (int c = 0; c < CPL; ++c) acc[i][c] *= corr[i];
    for (int j = 0; j < kKTile; ++j) {
      float vv[CPL];
#pragma unroll
      for (int c = 0; c < CPL; ++c) vv[c] = Vs[j * C + c * 32 + lane];
#pragma unroll
      for (int i = 0; i < kQPerWarp; ++i) {
        const float p = myp[i * kKTile + j];
#pragma unroll
        for (int c = 0; c < CPL; ++c) acc[i][c] = fmaf(p, vv[c], acc[i][c]);
      }
    }
    __syncwarp();
  }
#pragma unroll
  for (int i = 0; i < kQPerWarp; ++i) {
    const int qi = q0 + warp * kQPerWarp + i;
    if (qi >= S) continue;
    const float inv = 1.0f / l[i];
    T* dst = o + ((int64_t)b * S + qi) * ld_o;
#pragma unroll
    for (int c = 0; c < CPL; ++c) dst[c * 32 + lane] = from_f<T>(acc[i][c] * inv);
  }
}

template <typename T, int CPL>
static int launch_attention(const void* q, const void* k, const void* v, int B, int S, int ld, void* o, int ld_o,
                            cudaStream_t st) {
  constexpr int C = CPL * 32;
  size_t smem = (size_t)(kKTile * (C + 1) + kKTile * C + kQTile * C + kAttWarps * kQPerWarp * kKTile) * sizeof(float);
  DITSEP_CUDA(cudaFuncSetAttribute(attention_kernel<T, CPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(ceil_div(S, kQTile), B);
  DITSEP_CUDA(launch_pdl<2>(attention_kernel<T, CPL>, dim3(grid), dim3(256), smem, st, (const T*)q, (const T*)k, (const T*)v, S, ld, (T*)o,
                         ld_o, 1.0f / sqrtf((float)C)));
  return check_launch("attention_kernel");
}

// ---- tensor-core path (attention_tc.cuh): bf16, C = 128 / 256, at least one full key tile ----------------------------
typedef CUresult (*AttEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static AttEncodeFn att_encode_fn() {
  static AttEncodeFn fn = nullptr;
  if (fn) return fn;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<AttEncodeFn>(sym);
  return fn;
}

// (C, S, B) view of one of q / k / v: rows of one image are `ld` elements apart; rows past S are zero-filled by TMA
static int att_map(CUtensorMap* map, const void* p, int B, int S, int C, int ld, int rows) {
  AttEncodeFn enc = att_encode_fn();
  DITSEP_REQUIRE(enc != nullptr, "attention: cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t gdim[3] = {(cuuint64_t)C, (cuuint64_t)S, (cuuint64_t)B};
  cuuint64_t gstr[2] = {(cuuint64_t)ld * 2, (cuuint64_t)S * ld * 2};
  cuuint32_t box[3] = {64u, (cuuint32_t)rows, 1u};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(p), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DITSEP_REQUIRE(r == CUDA_SUCCESS, "attention: cuTensorMapEncodeTiled failed with %d", (int)r);
  return 0;
}

template <int C, int kKeys>
static int launch_attention_tc(const void* q, const void* k, const void* v, int B, int S, int ld, void* o, int ld_o,
                               cudaStream_t st) {
  CUtensorMap mq, mk, mv;
  if (int rc = att_map(&mq, q, B, S, C, ld, kAtQ)) return rc;
  if (int rc = att_map(&mk, k, B, S, C, ld, kKeys)) return rc;
  if (int rc = att_map(&mv, v, B, S, C, ld, kKeys)) return rc;
  const size_t smem = 1024 + (size_t)(C / 64) * kAtQ * 128 + 4 * (size_t)(C / 64) * kKeys * 128 + (size_t)(kKeys / 64) * kAtQ * 128;
  static std::atomic<bool> attr_set[64];
  int dev = 0;
  DITSEP_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev].load(std::memory_order_acquire)) {
    DITSEP_CUDA(cudaFuncSetAttribute(attention_tc_kernel<C, kKeys>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (dev >= 0 && dev < 64) attr_set[dev].store(true, std::memory_order_release);
  }
  dim3 grid(ceil_div(S, kAtQ), B);
  DITSEP_CUDA(launch_pdl<2>(attention_tc_kernel<C, kKeys>, grid, dim3(kAtThreads), smem, st, mq, mk, mv, S, (bf16*)o, ld_o,
                            1.0f / sqrtf((float)C)));
  return check_launch("attention_tc_kernel");
}

}  // namespace ditsep

using namespace ditsep;

extern "C" int ditsep_attention(const void* q, const void* k, const void* v, int dtype, int B, int S, int C, int ld_qkv,
                                void* o, int ld_o, void* stream) {
  DITSEP_REQUIRE(q && k && v && o, "attention: null pointer");
  DITSEP_REQUIRE(B > 0 && S > 0, "attention: empty problem");
  DITSEP_REQUIRE(C == 64 || C == 128 || C == 256, "attention: C=%d unsupported (64, 128, 256)", C);
  DITSEP_REQUIRE(ld_qkv % 8 == 0 && ld_qkv >= C && ld_o >= C, "attention: bad strides");
  cudaStream_t st = (cudaStream_t)stream;
  {
    // tcgen05 kernel: bf16, C = 128 / 256, 16-byte aligned operands, a sequence of at least 64 tokens
    // (DITSEP_ATTN_TC=0 keeps the CUDA-core kernel: debug aid and the comparison arm of the tests)
    const char* env = getenv("DITSEP_ATTN_TC");
    const bool aligned = ((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v) |
                           reinterpret_cast<uintptr_t>(o)) & 15) == 0 && ld_o % 8 == 0;
    if (dtype == DITSEP_BF16 && (C == 128 || C == 256) && S >= 64 && aligned && !(env && strcmp(env, "0") == 0)) {
      if (C == 128) return launch_attention_tc<128, 128>(q, k, v, B, S, ld_qkv, o, ld_o, st);
      return launch_attention_tc<256, 64>(q, k, v, B, S, ld_qkv, o, ld_o, st);
    }
  }
#define CASE(T, CPL) return launch_attention<T, CPL>(q, k, v, B, S, ld_qkv, o, ld_o, st)
  if (dtype == DITSEP_F32) {
    if (C == 64) CASE(float, 2);
    if (C == 128) CASE(float, 4);
    CASE(float, 8);
  } else if (dtype == DITSEP_BF16) {
    if (C == 64) CASE(bf16, 2);
    if (C == 128) CASE(bf16, 4);
    CASE(bf16, 8);
  }
#undef CASE
  DITSEP_REQUIRE(false, "attention: unknown dtype %d", dtype);
}
