// Shared device helpers for the k3m sm_100a kernels: counter-based RNG (dropout / Gumbel),
// warp/block reductions, vector load/store, error plumbing for the C ABI.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

typedef __nv_bfloat16 bf16;
// Activation / weight element type of the row-wise, embedding and loss kernels.  The product library is bf16.  The
// fp32 CHECK build (-DK3M_ACT_F32 -> libk3m_sm100a_f32.so, tests only: north_star's "1e-5 in an fp32 check mode")
// compiles the very same kernel sources with fp32 storage; its GEMM and attention are plain CUDA-core fp32 kernels
// (gemm_f32.cu, attention_f32.cu) behind the same entry points.
#ifdef K3M_ACT_F32
typedef float act_t;
#else
typedef bf16 act_t;
#endif

// ---- error handling across the C ABI (no exceptions) ------------------------------------------
#define K3M_OK 0
#define K3M_ERR_ARG -1
#define K3M_ERR_CUDA -2
#define K3M_ERR_UNSUPPORTED -3

void k3m_set_error(const char* fmt, ...);
int k3m_check_launch(const char* what);

#define K3M_REQUIRE(cond, ...)          \
  do {                                  \
    if (!(cond)) {                      \
      k3m_set_error(__VA_ARGS__);       \
      return K3M_ERR_ARG;               \
    }                                   \
  } while (0)

// ---- activation codes (same values in include/k3m.h and the Python op table) --------------------
enum { ACT_NONE = 0, ACT_GELU = 1, ACT_DGELU = 2, ACT_SIGMOID = 3, ACT_TANH = 4, ACT_DTANH = 5,
       ACT_GELU_GRAD = 6,   // out = gelu(v), aux_out = gelu'(v): the derivative is computed where Phi(v) already is
       ACT_MULAUX = 7 };    // out = v * aux_in (the backward of ACT_GELU_GRAD: a plain multiply)

// ---- dropout / RNG ------------------------------------------------------------------------------
// rng points to two device uint64: {seed, step}; site identifies the dropout application inside
// the step.  thresh == 0 means "no dropout".  Keep element iff its 16 random bits >= thresh.
struct DropSpec {
  const unsigned long long* rng;
  unsigned int site;
  unsigned int thresh;  // round(p * 65536)
  float scale;          // 1 / (1 - p)
};

// 7 rounds: the smallest Crush-resistant round count of Philox4x32 (Salmon et al., SC'11); dropout
// masks and Gumbel draws need decorrelated, reproducible bits, not a cryptographic margin.
constexpr int K3M_PHILOX_ROUNDS = 7;
__device__ __forceinline__ void philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t out[4]) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < K3M_PHILOX_ROUNDS; ++r) {
    uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// random words for counter `idx` of dropout/noise site `site`
__device__ __forceinline__ void rng_words(const unsigned long long* rng, unsigned int site,
                                          unsigned long long idx, uint32_t out[4]) {
  unsigned long long seed = rng[0], step = rng[1];
  philox4x32((uint32_t)idx, (uint32_t)(idx >> 32), site, (uint32_t)step, (uint32_t)seed,
                (uint32_t)(seed >> 32), out);
}

// multiplier (0 or scale) for element with linear index `lin` given the words of counter lin>>3
__device__ __forceinline__ float drop_mult(const DropSpec& d, const uint32_t w[4], unsigned int e) {
  uint32_t u16 = (w[e >> 1] >> (16 * (e & 1))) & 0xFFFFu;
  return u16 >= d.thresh ? d.scale : 0.0f;
}

// scalar convenience (one Philox call per element; used only in low-volume kernels)
__device__ __forceinline__ float drop_mult_at(const DropSpec& d, unsigned long long lin) {
  uint32_t w[4];
  rng_words(d.rng, d.site, lin >> 3, w);
  return drop_mult(d, w, (unsigned int)(lin & 7));
}

// ---- math ---------------------------------------------------------------------------------------
// Phi(x) = 0.5 (1 + erf(x / sqrt 2)) through the Abramowitz-Stegun 7.1.26 rational form of erfc
// (|abs error| <= 1.5e-7 on erf): branch-free, 2 MUFU + ~10 FMA per value instead of erff's two
// divergent polynomial branches.  The epilogues that use it are instruction-issue bound (K = 768
// leaves ~6000 clocks of main loop per 128x256 tile), so this is what the tile time is made of.
// `e` returns exp(-x^2 / 2) for the derivative.
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float gauss_cdf(float x, float& e) {
  const float z = fabsf(x) * 0.70710678118654752f;
  const float t = rcp_approx(fmaf(0.3275911f, z, 1.0f));        // MUFU.RCP, no IEEE fix-up path
  e = ex2_approx(x * x * -0.72134752044448170f);                 // exp(-x^2 / 2) = 2^(-x^2 log2(e) / 2)
  float p = fmaf(0.5f * 1.061405429f, t, 0.5f * -1.453152027f);  // coefficients carry the 0.5 of erfc / 2
  p = fmaf(p, t, 0.5f * 1.421413741f);
  p = fmaf(p, t, 0.5f * -0.284496736f);
  p = fmaf(p, t, 0.5f * 0.254829592f);
  const float h = p * t * e;          // 0.5 * erfc(|x| / sqrt 2)
  return x >= 0.0f ? 1.0f - h : h;
}
__device__ __forceinline__ float gelu_f(float x) {
  float e;
  return x * gauss_cdf(x, e);
}
__device__ __forceinline__ float dgelu_f(float x) {
  float e;
  const float c = gauss_cdf(x, e);
  return fmaf(x * e, 0.3989422804014327f, c);
}
__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + __expf(-x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// block-wide sum, result broadcast to all threads; `red` is >= 32 floats of shared memory
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.0f;
  r = warp_sum(r);
  return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : -3.0e38f;
  r = warp_max(r);
  return r;
}

// ---- typed element access (activations may be bf16 or fp32) ------------------------------------------
__device__ __forceinline__ float ld_f(const bf16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ float ld_f(const float* p) { return *p; }
__device__ __forceinline__ void st_f(bf16* p, float v) { *p = __float2bfloat16(v); }
__device__ __forceinline__ void st_f(float* p, float v) { *p = v; }

// 8 consecutive bf16 <-> 8 floats through one 16-byte access (pointer must be 16 B aligned)
__device__ __forceinline__ void ld8(const bf16* p, float v[8]) {
  uint4 u = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x; v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ void st8(bf16* p, const float v[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ void ld8(const float* p, float v[8]) {
  float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void st8(float* p, const float v[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

// ---- value conversions / 8-element packs in the activation type ----------------------------------------------
__device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ act_t act_from(float v) {
#ifdef K3M_ACT_F32
  return v;
#else
  return __float2bfloat16(v);
#endif
}
// eight consecutive activation values kept as loaded (16 bytes of bf16, 32 bytes of fp32)
#ifdef K3M_ACT_F32
struct Pack8 { float4 a, b; };
__device__ __forceinline__ Pack8 load_pack8(const float* p) {
  Pack8 r;
  r.a = *reinterpret_cast<const float4*>(p);
  r.b = *reinterpret_cast<const float4*>(p + 4);
  return r;
}
__device__ __forceinline__ void unpack8(const Pack8& u, float v[8]) {
  v[0] = u.a.x; v[1] = u.a.y; v[2] = u.a.z; v[3] = u.a.w; v[4] = u.b.x; v[5] = u.b.y; v[6] = u.b.z; v[7] = u.b.w;
}
#else
struct Pack8 { uint4 u; };
__device__ __forceinline__ Pack8 load_pack8(const bf16* p) {
  Pack8 r;
  r.u = *reinterpret_cast<const uint4*>(p);
  return r;
}
__device__ __forceinline__ void unpack8(const Pack8& u, float v[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u.u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x; v[2 * i + 1] = f.y;
  }
}
#endif
