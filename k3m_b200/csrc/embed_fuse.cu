// Embedding gathers, the 5-d box projection, modal fusion (hard Gumbel selection) and pooling.
// All HBM-bound or tiny; index gathers are exact integer arithmetic.
#include "common.cuh"

namespace {

// out[r] = word[ids[r]] + pos[r % L] + typ[type_ids[r]]     (BertEmbeddings, vilbert_k3m.py:361-376)
__global__ void __launch_bounds__(128) embed_fwd_kernel(const long long* __restrict__ ids, const long long* __restrict__ tids,
                                                        const act_t* __restrict__ word, const act_t* __restrict__ pos,
                                                        const act_t* __restrict__ typ, act_t* __restrict__ out, int M, int L, int H) {
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= M) return;
  const act_t* w = word + (size_t)ids[row] * H;
  const act_t* p = pos + (size_t)(row % L) * H;
  const act_t* t = typ + (size_t)tids[row] * H;
  for (int c = lane * 8; c < H; c += 256) {
    float a[8], b[8], e[8];
    ld8(w + c, a); ld8(p + c, b); ld8(t + c, e);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] += b[j] + e[j];
    st8(out + (size_t)row * H + c, a);
  }
}

// scatter-add of dx rows into the fp32 embedding gradients; word row 0 (padding_idx) is skipped
__global__ void __launch_bounds__(128) embed_bwd_kernel(const act_t* __restrict__ dx, const long long* __restrict__ ids,
                                                        const long long* __restrict__ tids, float* __restrict__ dword,
                                                        float* __restrict__ dpos, float* __restrict__ dtyp, int M, int L, int H) {
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= M) return;
  const long long id = ids[row];
  float* w = dword + (size_t)id * H;
  float* p = dpos + (size_t)(row % L) * H;
  float* t = dtyp + (size_t)tids[row] * H;
  for (int c = lane; c < H; c += 32) {
    const float g = to_f(dx[(size_t)row * H + c]);
    if (id != 0) atomicAdd(w + c, g);
    atomicAdd(p + c, g);
    atomicAdd(t + c, g);
  }
}

// out[m, n] = sum_k loc[m, k] w[n, k] + b[n],  k < 5
__global__ void loc_fwd_kernel(const float* __restrict__ loc, const act_t* __restrict__ w, const act_t* __restrict__ b,
                               act_t* __restrict__ out, int M, int N) {
  const unsigned n = (unsigned)(M * N);   // B * L rows: far below 2^31 elements; 32-bit index math (64-bit div / mod dominated these kernels)
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int m = (int)(i / N), c = (int)(i % N);
    float acc = ld_f(b + c);
#pragma unroll
    for (int k = 0; k < 5; ++k) acc = fmaf(loc[(size_t)m * 5 + k], ld_f(w + (size_t)c * 5 + k), acc);
    out[i] = act_from(acc);
  }
}

// dw[n, k] += sum_m dy[m, n] loc[m, k] ; db[n] += sum_m dy[m, n]       one thread per column n
__global__ void __launch_bounds__(128) loc_bwd_kernel(const act_t* __restrict__ dy, const float* __restrict__ loc,
                                                      float* __restrict__ dw, float* __restrict__ db, int M, int N, int rows_per) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= N) return;
  const int r0 = blockIdx.y * rows_per;
  int r1 = r0 + rows_per;
  if (r1 > M) r1 = M;
  float acc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int m = r0; m < r1; ++m) {
    const float g = to_f(dy[(size_t)m * N + c]);
#pragma unroll
    for (int k = 0; k < 5; ++k) acc[k] = fmaf(g, loc[(size_t)m * 5 + k], acc[k]);
    acc[5] += g;
  }
#pragma unroll
  for (int k = 0; k < 5; ++k) atomicAdd(dw + (size_t)c * 5 + k, acc[k]);
  atomicAdd(db + c, acc[5]);
}

// cat[m] = [relu(a[m]) | relu(b[m]) | relu(c[m])]
__global__ void relu_cat3_kernel(const act_t* a, int lda, const act_t* b, int ldb, const act_t* c, int ldc, act_t* out, int M, int H) {
  const int hv = H / 8;
  const unsigned n = (unsigned)(M * 3 * hv);   // B * L rows: far below 2^31 elements; 32-bit index math (64-bit div / mod dominated these kernels)
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int m = (int)(i / (3 * hv)), r = (int)(i % (3 * hv)), k = r / hv, col = (r % hv) * 8;
    const act_t* src = k == 0 ? a + (size_t)m * lda : (k == 1 ? b + (size_t)m * ldb : c + (size_t)m * ldc);
    float v[8];
    ld8(src + col, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
    st8(out + (size_t)m * 3 * H + k * H + col, v);
  }
}

// Gumbel(0,1) draw for element lin of a noise site: 4 draws per Philox call
__device__ __forceinline__ float gumbel_at(const unsigned long long* rng, unsigned int site, unsigned long long lin) {
  uint32_t w[4];
  rng_words(rng, site, lin >> 2, w);
  // 23 random bits + 1/2 is exact in fp32, so u stays inside [2^-24, 1 - 2^-24] and the noise is
  // finite (with 24 bits, 16777215.5 rounds to 2^24 -> u = 1 -> g = +inf -> NaN in the backward softmax)
  const float u = ((float)(w[lin & 3] >> 9) + 0.5f) * (1.0f / 8388608.0f);
  return -logf(-logf(u));
}

// y_k = softmax_k(sigmoid(gate_k) + g_k), hard pick = first arg-max  (pre_sampling_sequence,
// vilbert_k3m.py:2331-2374; F.gumbel_softmax(hard=True, tau=1)).
// Four consecutive channels per thread: the Gumbel draws of elements lin .. lin + 3 (lin % 4 == 0) are the four words
// of ONE Philox call, which the one-channel-per-thread form recomputed in each of its four threads — the generator was
// two thirds of these ALU-bound kernels' instructions.  Same counters, same words, same arithmetic per element.
__device__ __forceinline__ void ysoft3x4(const float* gates, const float* noise, const unsigned long long* rng,
                                         unsigned int site, int m, int h, int H, float ak[3][4], float y[3][4], int pick[4]) {
  float z[3][4];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const size_t lin = ((size_t)m * 3 + k) * H + h;            // % 4 == 0 (H % 4 == 0, h % 4 == 0)
    const float4 gt = *reinterpret_cast<const float4*>(gates + lin);
    const float gk[4] = {gt.x, gt.y, gt.z, gt.w};
    float g[4];
    if (noise) {
      const float4 nz = *reinterpret_cast<const float4*>(noise + lin);
      g[0] = nz.x; g[1] = nz.y; g[2] = nz.z; g[3] = nz.w;
    } else {
      uint32_t w[4];
      rng_words(rng, site, lin >> 2, w);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float u = ((float)(w[j] >> 9) + 0.5f) * (1.0f / 8388608.0f);     // == gumbel_at(rng, site, lin + j)
        g[j] = -logf(-logf(u));
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      ak[k][j] = sigmoid_f(gk[j]);
      z[k][j] = ak[k][j] + g[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float mx = fmaxf(z[0][j], fmaxf(z[1][j], z[2][j]));
    float e[3], sum = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) { e[k] = __expf(z[k][j] - mx); sum += e[k]; }
#pragma unroll
    for (int k = 0; k < 3; ++k) y[k][j] = e[k] / sum;
    int p = 0;
    if (y[1][j] > y[p][j]) p = 1;
    if (y[2][j] > y[p][j]) p = 2;
    pick[j] = p;
  }
}
__device__ __forceinline__ void ld4(const bf16* p, float v[4]) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
  const float2 a = __bfloat1622float2(h[0]), b = __bfloat1622float2(h[1]);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void st4(bf16* p, const float v[4]) {
  uint2 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
  h[0] = __floats2bfloat162_rn(v[0], v[1]);
  h[1] = __floats2bfloat162_rn(v[2], v[3]);
  *reinterpret_cast<uint2*>(p) = u;
}
__device__ __forceinline__ void ld4(const float* p, float v[4]) {
  const float4 u = *reinterpret_cast<const float4*>(p);
  v[0] = u.x; v[1] = u.y; v[2] = u.z; v[3] = u.w;
}
__device__ __forceinline__ void st4(float* p, const float v[4]) { *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]); }

__global__ void fuse_fwd_kernel(const act_t* __restrict__ cat, const float* __restrict__ gates, const float* __restrict__ noise,
                                const unsigned long long* rng, unsigned int site, act_t* __restrict__ out, int ldo, int M, int H) {
  const unsigned H4 = (unsigned)(H >> 2), n = (unsigned)M * H4;   // 32-bit index math (64-bit div / mod dominated these kernels)
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int m = (int)(i / H4), h = (int)(i % H4) * 4;
    float ak[3][4], y[3][4], c[3][4], o[4];
    int pick[4];
    ysoft3x4(gates, noise, rng, site, m, h, H, ak, y, pick);
#pragma unroll
    for (int k = 0; k < 3; ++k) ld4(cat + (size_t)m * 3 * H + k * H + h, c[k]);
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = pick[j] == 0 ? c[0][j] : (pick[j] == 1 ? c[1][j] : c[2][j]);
    st4(out + (size_t)m * ldo + h, o);
  }
}

// straight-through backward: dgates[m,k,h] and dcat_direct[m,k,h] = dout * onehot
__global__ void fuse_bwd_kernel(const act_t* __restrict__ dout, int ldd, const act_t* __restrict__ cat,
                                const float* __restrict__ gates, const float* __restrict__ noise,
                                const unsigned long long* rng, unsigned int site, act_t* __restrict__ dgates,
                                act_t* __restrict__ dcat, int M, int H) {
  const unsigned H4 = (unsigned)(H >> 2), n = (unsigned)M * H4;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int m = (int)(i / H4), h = (int)(i % H4) * 4;
    float ak[3][4], y[3][4], go[4], c[3][4];
    int pick[4];
    ysoft3x4(gates, noise, rng, site, m, h, H, ak, y, pick);
    ld4(dout + (size_t)m * ldd + h, go);
#pragma unroll
    for (int k = 0; k < 3; ++k) ld4(cat + (size_t)m * 3 * H + k * H + h, c[k]);
    float dg[3][4], dc[3][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float dy[3], dot = 0.f;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        dy[k] = go[j] * c[k][j];
        dot += y[k][j] * dy[k];
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        dg[k][j] = y[k][j] * (dy[k] - dot) * ak[k][j] * (1.0f - ak[k][j]);
        dc[k][j] = k == pick[j] ? go[j] : 0.f;
      }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const size_t o = (size_t)m * 3 * H + k * H + h;
      st4(dgates + o, dg[k]);
      st4(dcat + o, dc[k]);
    }
  }
}

// d{a,b,c}[m,h] += dcat[m, k*H + h] * (cat[m, k*H + h] > 0)
__global__ void relu_split3_kernel(const act_t* __restrict__ dcat, const act_t* __restrict__ cat, act_t* da, int lda, act_t* db,
                                   int ldb, act_t* dc, int ldc, int M, int H) {
  const unsigned n = (unsigned)(M * 3 * H);   // B * L rows: far below 2^31 elements; 32-bit index math (64-bit div / mod dominated these kernels)
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int m = (int)(i / (3 * H)), r = (int)(i % (3 * H)), k = r / H, h = r % H;
    if (to_f(cat[i]) > 0.f) {
      act_t* dst = k == 0 ? da + (size_t)m * lda + h : (k == 1 ? db + (size_t)m * ldb + h : dc + (size_t)m * ldc + h);
      *dst = act_from(to_f(*dst) + to_f(dcat[i]));
    }
  }
}

// eight channels per thread (16-byte accesses); same arithmetic per element as the scalar form above
__global__ void relu_split3_vec_kernel(const act_t* __restrict__ dcat, const act_t* __restrict__ cat, act_t* da, int lda,
                                       act_t* db, int ldb, act_t* dc, int ldc, int M, int H) {
  const unsigned hv = (unsigned)(H >> 3), n = (unsigned)M * 3 * hv;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int m = (int)(i / (3 * hv)), r = (int)(i % (3 * hv)), k = r / (int)hv, h = (r % (int)hv) * 8;
    float c[8], d[8], o[8];
    ld8(cat + (size_t)i * 8, c);
    ld8(dcat + (size_t)i * 8, d);
    act_t* dst = k == 0 ? da + (size_t)m * lda + h : (k == 1 ? db + (size_t)m * ldb + h : dc + (size_t)m * ldc + h);
    ld8(dst, o);
    bool any = false;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      // the scalar form leaves dst untouched where cat <= 0 and re-rounds (dst + dcat) elsewhere
      if (c[j] > 0.f) { o[j] = to_f(act_from(o[j] + d[j])); any = true; }
    }
    if (any) st8(dst, o);
  }
}

// pooled[b, h] = mean_{i=1..L-1} seq[b*L + i, h]       (vilbert_k3m.py:2405-2409)
template <typename TO>
__global__ void pool_fwd_kernel(const act_t* __restrict__ seq, int ld, TO* __restrict__ out, int B, int L, int H) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * H) return;
  const int b = i / H, h = i % H;
  float acc = 0.f;
  for (int t = 1; t < L; ++t) acc += to_f(seq[(size_t)(b * L + t) * ld + h]);
  st_f(out + i, acc / (float)(L - 1));
}

// dseq[b*L + i, h] += dpool[b, h] / (L-1), i >= 1
__global__ void pool_bwd_kernel(const float* __restrict__ dpool, act_t* __restrict__ dseq, int ld, int B, int L, int H) {
  const unsigned n = (unsigned)(B * L * H);   // B * L rows: far below 2^31 elements; 32-bit index math (64-bit div / mod dominated these kernels)
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int h = (int)(i % H);
    const int r = (int)(i / H), b = r / L, t = r % L;
    if (t == 0) continue;
    act_t* p = dseq + (size_t)r * ld + h;
    *p = act_from(to_f(*p) + dpool[(size_t)b * H + h] / (float)(L - 1));
  }
}

// ---- soft fusion (if_pre_sampling = 2, pre_sampling_sequence_soft, vilbert_k3m.py:2300-2329) --------------
// gated = cat * alpha  (cat = relu(cat3), alpha = sigmoid(score GEMM), both act_t [M, 3H], 8 elements per thread)
__global__ void gate_mul_kernel(const act_t* __restrict__ cat, const act_t* __restrict__ alpha, act_t* __restrict__ out,
                                size_t n8) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    float a[8], b[8];
    ld8(cat + i * 8, a);
    ld8(alpha + i * 8, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] *= b[j];
    st8(out + i * 8, a);
  }
}
// dgated -> dcat (direct path) = dgated * alpha ;  dscore = dgated * cat * alpha * (1 - alpha)
__global__ void gate_bwd_kernel(const act_t* __restrict__ dgated, const act_t* __restrict__ cat,
                                const act_t* __restrict__ alpha, act_t* __restrict__ dcat, act_t* __restrict__ dscore,
                                size_t n8) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    float d[8], c[8], a[8], o1[8], o2[8];
    ld8(dgated + i * 8, d);
    ld8(cat + i * 8, c);
    ld8(alpha + i * 8, a);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      o1[j] = d[j] * a[j];
      o2[j] = d[j] * c[j] * a[j] * (1.0f - a[j]);
    }
    st8(dcat + i * 8, o1);
    st8(dscore + i * 8, o2);
  }
}

// ---- labelled-row gather for the masked-LM head --------------------------------------------------------------
// The reference runs the vocabulary decoder on EVERY position and lets CrossEntropyLoss ignore the ~85 % of rows whose
// label is -1 (vilbert_k3m.py:1836-1839, :2817-2826); rows without a label contribute neither loss nor gradient, so
// the head runs on the labelled rows only: out[i] = src[idx[i]] (idx < 0: a zero row of padding), and its input
// gradient goes back with dst[idx[i]] += dsrc[i] (indices are unique: plain read-modify-write).
__global__ void gather_rows_kernel(const act_t* __restrict__ src, int ld, const int* __restrict__ idx,
                                   act_t* __restrict__ out, int n, int H8) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)n * H8; i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / H8), c = (int)(i % H8) * 8;
    float v[8];
    const int s_ = idx[r];
    if (s_ >= 0) {
      ld8(src + (size_t)s_ * ld + c, v);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = 0.f;
    }
    st8(out + (size_t)r * (H8 * 8) + c, v);
  }
}
__global__ void scatter_add_rows_kernel(const act_t* __restrict__ dsrc, const int* __restrict__ idx,
                                        act_t* __restrict__ dst, int ld, int n, int H8) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)n * H8; i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / H8), c = (int)(i % H8) * 8;
    const int s_ = idx[r];
    if (s_ < 0) continue;
    float a[8], b[8];
    ld8(dsrc + (size_t)r * (H8 * 8) + c, a);
    ld8(dst + (size_t)s_ * ld + c, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) b[j] += a[j];
    st8(dst + (size_t)s_ * ld + c, b);
  }
}

inline int grid_for(size_t n, int block) {
  size_t g = (n + block - 1) / block;
  if (g > 148 * 16) g = 148 * 16;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace

#define ST(s) reinterpret_cast<cudaStream_t>(s)

extern "C" int k3m_embed_fwd(const long long* ids, const long long* type_ids, const void* word, const void* pos,
                             const void* typ, void* out, int M, int L, int H, void* stream) {
  K3M_REQUIRE(M > 0 && H % 8 == 0, "embed_fwd: H must be a multiple of 8");
  embed_fwd_kernel<<<(M + 3) / 4, 128, 0, ST(stream)>>>(ids, type_ids, (const act_t*)word, (const act_t*)pos,
                                                        (const act_t*)typ, (act_t*)out, M, L, H);
  return k3m_check_launch("k3m_embed_fwd");
}

extern "C" int k3m_embed_bwd(const void* dx, const long long* ids, const long long* type_ids, float* dword, float* dpos,
                             float* dtyp, int M, int L, int H, void* stream) {
  embed_bwd_kernel<<<(M + 3) / 4, 128, 0, ST(stream)>>>((const act_t*)dx, ids, type_ids, dword, dpos, dtyp, M, L, H);
  return k3m_check_launch("k3m_embed_bwd");
}

extern "C" int k3m_loc_embed_fwd(const float* loc, const void* w, const void* b, void* out, int M, int N, void* stream) {
  loc_fwd_kernel<<<grid_for((size_t)M * N, 256), 256, 0, ST(stream)>>>(loc, (const act_t*)w, (const act_t*)b, (act_t*)out, M, N);
  return k3m_check_launch("k3m_loc_embed_fwd");
}

extern "C" int k3m_loc_embed_bwd(const void* dy, const float* loc, float* dw, float* db, int M, int N, void* stream) {
  int gy = (M + 127) / 128;
  if (gy > 64) gy = 64;
  loc_bwd_kernel<<<dim3((N + 127) / 128, gy), 128, 0, ST(stream)>>>((const act_t*)dy, loc, dw, db, M, N, (M + gy - 1) / gy);
  return k3m_check_launch("k3m_loc_embed_bwd");
}

extern "C" int k3m_relu_cat3(const void* a, int lda, const void* b, int ldb, const void* c, int ldc, void* out, int M,
                             int H, void* stream) {
  K3M_REQUIRE(H % 8 == 0 && lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "relu_cat3: widths must be multiples of 8");
  relu_cat3_kernel<<<grid_for((size_t)M * 3 * H / 8, 256), 256, 0, ST(stream)>>>((const act_t*)a, lda, (const act_t*)b, ldb,
                                                                                 (const act_t*)c, ldc, (act_t*)out, M, H);
  return k3m_check_launch("k3m_relu_cat3");
}

extern "C" int k3m_fuse_select_fwd(const void* cat, const float* gates, const float* noise, const void* rng,
                                   unsigned int site, void* out, int ldo, int M, int H, void* stream) {
  K3M_REQUIRE(noise || rng, "fuse_select: need either a noise tensor or the RNG state");
  K3M_REQUIRE(H % 4 == 0 && ldo % 4 == 0, "fuse_select: H and the output row stride must be multiples of 4 (H=%d ldo=%d)", H, ldo);
  fuse_fwd_kernel<<<grid_for((size_t)M * H / 4, 256), 256, 0, ST(stream)>>>((const act_t*)cat, gates, noise,
                                                                        (const unsigned long long*)rng, site, (act_t*)out, ldo, M, H);
  return k3m_check_launch("k3m_fuse_select_fwd");
}

extern "C" int k3m_fuse_select_bwd(const void* dout, int ldd, const void* cat, const float* gates, const float* noise,
                                   const void* rng, unsigned int site, void* dgates, void* dcat, int M, int H, void* stream) {
  K3M_REQUIRE(noise || rng, "fuse_select: need either a noise tensor or the RNG state");
  K3M_REQUIRE(H % 4 == 0 && ldd % 4 == 0, "fuse_select: H and the gradient row stride must be multiples of 4 (H=%d ldd=%d)", H, ldd);
  fuse_bwd_kernel<<<grid_for((size_t)M * H / 4, 256), 256, 0, ST(stream)>>>(
      (const act_t*)dout, ldd, (const act_t*)cat, gates, noise, (const unsigned long long*)rng, site, (act_t*)dgates, (act_t*)dcat, M, H);
  return k3m_check_launch("k3m_fuse_select_bwd");
}

extern "C" int k3m_gather_rows(const void* src, int ld, const int* idx, void* out, int n, int H, void* stream) {
  K3M_REQUIRE(n > 0 && H > 0 && H % 8 == 0 && ld % 8 == 0, "gather_rows: widths must be multiples of 8");
  gather_rows_kernel<<<grid_for((size_t)n * H / 8, 256), 256, 0, ST(stream)>>>((const act_t*)src, ld, idx, (act_t*)out, n, H / 8);
  return k3m_check_launch("k3m_gather_rows");
}

extern "C" int k3m_scatter_add_rows(const void* dsrc, const int* idx, void* dst, int ld, int n, int H, void* stream) {
  K3M_REQUIRE(n > 0 && H > 0 && H % 8 == 0 && ld % 8 == 0, "scatter_add_rows: widths must be multiples of 8");
  scatter_add_rows_kernel<<<grid_for((size_t)n * H / 8, 256), 256, 0, ST(stream)>>>((const act_t*)dsrc, idx, (act_t*)dst, ld, n, H / 8);
  return k3m_check_launch("k3m_scatter_add_rows");
}

extern "C" int k3m_gate_mul(const void* cat, const void* alpha, void* out, long long n, void* stream) {
  K3M_REQUIRE(n > 0 && n % 8 == 0, "gate_mul: element count must be a positive multiple of 8");
  gate_mul_kernel<<<grid_for((size_t)n / 8, 256), 256, 0, ST(stream)>>>((const act_t*)cat, (const act_t*)alpha, (act_t*)out,
                                                                       (size_t)n / 8);
  return k3m_check_launch("k3m_gate_mul");
}

extern "C" int k3m_gate_bwd(const void* dgated, const void* cat, const void* alpha, void* dcat, void* dscore, long long n,
                            void* stream) {
  K3M_REQUIRE(n > 0 && n % 8 == 0, "gate_bwd: element count must be a positive multiple of 8");
  gate_bwd_kernel<<<grid_for((size_t)n / 8, 256), 256, 0, ST(stream)>>>((const act_t*)dgated, (const act_t*)cat,
                                                                       (const act_t*)alpha, (act_t*)dcat, (act_t*)dscore,
                                                                       (size_t)n / 8);
  return k3m_check_launch("k3m_gate_bwd");
}

extern "C" int k3m_relu_bwd_split3(const void* dcat, const void* cat, void* da, int lda, void* db, int ldb, void* dc,
                                   int ldc, int M, int H, void* stream) {
  auto al = [](const void* q) { return (((uintptr_t)q) & 15) == 0; };
  if (H % 8 == 0 && lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0 && al(dcat) && al(cat) && al(da) && al(db) && al(dc)) {
    relu_split3_vec_kernel<<<grid_for((size_t)M * 3 * H / 8, 256), 256, 0, ST(stream)>>>(
        (const act_t*)dcat, (const act_t*)cat, (act_t*)da, lda, (act_t*)db, ldb, (act_t*)dc, ldc, M, H);
    return k3m_check_launch("k3m_relu_bwd_split3");
  }
  relu_split3_kernel<<<grid_for((size_t)M * 3 * H, 256), 256, 0, ST(stream)>>>((const act_t*)dcat, (const act_t*)cat, (act_t*)da,
                                                                               lda, (act_t*)db, ldb, (act_t*)dc, ldc, M, H);
  return k3m_check_launch("k3m_relu_bwd_split3");
}

extern "C" int k3m_mean_pool_fwd(const void* seq, int ld, void* out, int out_f32, int B, int L, int H, void* stream) {
  K3M_REQUIRE(L >= 2, "mean_pool: need at least two tokens");
  const int g = (B * H + 255) / 256;
  if (out_f32) pool_fwd_kernel<float><<<g, 256, 0, ST(stream)>>>((const act_t*)seq, ld, (float*)out, B, L, H);
  else pool_fwd_kernel<act_t><<<g, 256, 0, ST(stream)>>>((const act_t*)seq, ld, (act_t*)out, B, L, H);
  return k3m_check_launch("k3m_mean_pool_fwd");
}

extern "C" int k3m_mean_pool_bwd(const float* dpool, void* dseq, int ld, int B, int L, int H, void* stream) {
  pool_bwd_kernel<<<grid_for((size_t)B * L * H, 256), 256, 0, ST(stream)>>>(dpool, (act_t*)dseq, ld, B, L, H);
  return k3m_check_launch("k3m_mean_pool_bwd");
}
