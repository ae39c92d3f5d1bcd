// fp32 CHECK build only (-DK3M_ACT_F32, libk3m_sm100a_f32.so): the GEMM entry points of include/k3m.h with fp32
// operands, fp32 FMA accumulation on the CUDA cores and the same fused epilogue contract as gemm_sm100.cu
//   v = acc (+ bias[n]); activation; dropout(p); v += addend[m,n]; out = accumulate ? out + v : v
// (same dropout counter stream: element (row, col) -> linear index row * round_up(N, 8) + col).  It exists so
// that the step can be compared with the unmodified reference at fp32 tolerances (north_star: 1e-5); it is not a
// product path and not tuned: 64 x 64 x 16 tiles, 4 x 4 outputs per thread.
#ifdef K3M_ACT_F32
#include <stdio.h>

#include "common.cuh"

namespace {

struct F32Gemm {
  const float *a, *b;
  int lda, ldb, layout, M, N, K;
  float* out;
  int ldc, accumulate;
  const float* bias;
  int act;
  const float* aux_in;
  int ld_aux_in;
  float* aux_out;
  int ld_aux_out;
  const float* addend;
  int ld_add;
  DropSpec drop;
  int drop_ld;
};

__device__ __forceinline__ float gelu_exact(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float dgelu_exact(float x) {
  return 0.5f * (1.0f + erff(x * 0.70710678118654752f)) + x * 0.3989422804014327f * expf(-0.5f * x * x);
}

__global__ void __launch_bounds__(256) gemm_f32_kernel(F32Gemm g) {
  __shared__ float As[16][65];
  __shared__ float Bs[16][65];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < g.K; k0 += 16) {
    for (int idx = threadIdx.x; idx < 16 * 64; idx += 256) {
      // A tile element (k, m), B tile element (k, n): the contiguous direction of each layout varies fastest
      int k, m, n;
      if (g.layout == 2) { k = idx >> 6; m = idx & 63; } else { m = idx >> 4; k = idx & 15; }
      float va = 0.f;
      if (m0 + m < g.M && k0 + k < g.K)
        va = g.layout == 2 ? g.a[(size_t)(k0 + k) * g.lda + m0 + m] : g.a[(size_t)(m0 + m) * g.lda + k0 + k];
      As[k][m] = va;
      if (g.layout == 0) { n = idx >> 4; k = idx & 15; } else { k = idx >> 6; n = idx & 63; }
      float vb = 0.f;
      if (n0 + n < g.N && k0 + k < g.K)
        vb = g.layout == 0 ? g.b[(size_t)(n0 + n) * g.ldb + k0 + k] : g.b[(size_t)(k0 + k) * g.ldb + n0 + n];
      Bs[k][n] = vb;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = As[k][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = Bs[k][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int row = m0 + ty * 4 + i;
    if (row >= g.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int col = n0 + tx * 4 + j;
      if (col >= g.N) continue;
      float v = acc[i][j];
      if (g.bias) v += g.bias[col];
      if (g.act == ACT_GELU) {
        if (g.aux_out) g.aux_out[(size_t)row * g.ld_aux_out + col] = v;
        v = gelu_exact(v);
      } else if (g.act == ACT_GELU_GRAD) {
        if (g.aux_out) g.aux_out[(size_t)row * g.ld_aux_out + col] = dgelu_exact(v);
        v = gelu_exact(v);
      } else if (g.act == ACT_MULAUX) {
        v *= g.aux_in[(size_t)row * g.ld_aux_in + col];
      } else if (g.act == ACT_DGELU) {
        v *= dgelu_exact(g.aux_in[(size_t)row * g.ld_aux_in + col]);
      } else if (g.act == ACT_DTANH) {
        const float t = g.aux_in[(size_t)row * g.ld_aux_in + col];
        v *= 1.0f - t * t;
      } else if (g.act == ACT_SIGMOID) {
        v = 1.0f / (1.0f + expf(-v));
      } else if (g.act == ACT_TANH) {
        v = tanhf(v);
      }
      if (g.drop.thresh) v *= drop_mult_at(g.drop, (unsigned long long)row * g.drop_ld + col);
      if (g.addend) v += g.addend[(size_t)row * g.ld_add + col];
      float* o = g.out + (size_t)row * g.ldc + col;
      *o = g.accumulate ? *o + v : v;
    }
  }
}

struct Desc {      // == K3MGemmDesc of include/k3m.h
  const void* a; int lda; const void* b; int ldb; int layout; int M, N, K;
  void* out; int ldc; int out_f32; int accumulate; const void* bias; int act;
  const void* aux_in; int ld_aux_in; void* aux_out; int ld_aux_out;
  const void* addend; int ld_add; int addend_f32; const void* rng; unsigned int drop_site; float drop_p;
};

int run(const Desc& d, cudaStream_t st) {
  K3M_REQUIRE(d.M > 0 && d.N > 0 && d.K > 0, "gemm: empty problem M=%d N=%d K=%d", d.M, d.N, d.K);
  K3M_REQUIRE(d.layout >= 0 && d.layout <= 2, "gemm: layout must be 0 (nt), 1 (nn) or 2 (tn)");
  K3M_REQUIRE(!(d.act == ACT_DGELU || d.act == ACT_DTANH || d.act == ACT_MULAUX) || d.aux_in, "gemm: activation-derivative epilogue needs aux_in");
  F32Gemm g;
  g.a = (const float*)d.a; g.b = (const float*)d.b; g.lda = d.lda; g.ldb = d.ldb; g.layout = d.layout;
  g.M = d.M; g.N = d.N; g.K = d.K;
  g.out = (float*)d.out; g.ldc = d.ldc; g.accumulate = d.accumulate;
  g.bias = (const float*)d.bias; g.act = d.act;
  g.aux_in = (const float*)d.aux_in; g.ld_aux_in = d.ld_aux_in;
  g.aux_out = (float*)d.aux_out; g.ld_aux_out = d.ld_aux_out;
  g.addend = (const float*)d.addend; g.ld_add = d.ld_add;
  g.drop.rng = (const unsigned long long*)d.rng;
  g.drop.site = d.drop_site;
  g.drop.thresh = (d.rng && d.drop_p > 0.f) ? (unsigned int)(d.drop_p * 65536.0f + 0.5f) : 0u;
  g.drop.scale = 1.0f / (1.0f - d.drop_p);
  g.drop_ld = (d.N + 7) / 8 * 8;
  gemm_f32_kernel<<<dim3((d.N + 63) / 64, (d.M + 63) / 64), 256, 0, st>>>(g);
  return k3m_check_launch("k3m_gemm_bf16 (fp32 check build)");
}

}  // namespace

extern "C" const char* k3m_gemm_last_variant(void) { return "fp32 check build: CUDA-core 64x64x16 tiles"; }
extern "C" void k3m_gemm_set_sm_limit(int) {}

extern "C" int k3m_gemm_bf16(const void* a, int lda, const void* b, int ldb, int layout, int M, int N, int K,
                             void* out, int ldc, int out_f32, int accumulate, const void* bias, int act,
                             const void* aux_in, int ld_aux_in, void* aux_out, int ld_aux_out, const void* addend,
                             int ld_add, int addend_f32, const void* rng, unsigned int drop_site, float drop_p,
                             void* stream) {
  Desc d = {a, lda, b, ldb, layout, M, N, K, out, ldc, out_f32, accumulate, bias, act, aux_in, ld_aux_in, aux_out,
            ld_aux_out, addend, ld_add, addend_f32, rng, drop_site, drop_p};
  return run(d, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int k3m_gemm_grouped_bf16(const void* descs, int n, void* stream) {
  const Desc* d = reinterpret_cast<const Desc*>(descs);
  K3M_REQUIRE(n >= 1 && n <= 8, "gemm_grouped: 1..8 problems per launch (got %d)", n);
  for (int i = 0; i < n; ++i)
    if (int rc = run(d[i], reinterpret_cast<cudaStream_t>(stream))) return rc;
  return K3M_OK;
}
#endif  // K3M_ACT_F32
