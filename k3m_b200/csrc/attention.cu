// Fused masked-softmax attention for the short K3M sequences (36-token titles / PV lists, 37
// regions; up to 256 keys forward, 128x128 backward).  One CTA per (batch, head): K and V of
// the head live in shared memory for the whole CTA, one warp owns one query row, the softmax is
// a warp-shuffle reduction in fp32, dropout comes from the counter RNG (regenerated, never
// stored) and the context is written head-merged ([B*Lq, H*d]) so the next GEMM reads it as is.
// Q, K, V are column slices of the fused projection buffers (row stride ld, column offset).
//
// Reference: BertSelfAttention.forward vilbert_k3m.py:440-464, BertBiAttention.forward :786-824.
#include "common.cuh"

namespace {

constexpr int MAX_NJ = 8;  // keys per lane -> Lk <= 256

struct AttnArgs {
  const bf16* q; int ldq, q_off;
  const bf16* kv; int ldkv, k_off, v_off;
  const float* mask;  // [B, Lk] additive
  int B, H, Lq, Lk, d;
  float scale;
  DropSpec drop;
};

#include "attn_mma.cuh"

__device__ __forceinline__ void load_tile(bf16* dst, int dstride, const bf16* src, int ld, int rows, int d) {
  // rows x d bf16, global row stride ld, 16 B vector loads (d % 8 == 0), smem row stride dstride
  const int vec_per_row = d >> 3;
  for (int idx = threadIdx.x; idx < rows * vec_per_row; idx += blockDim.x) {
    const int r = idx / vec_per_row, c = (idx % vec_per_row) * 8;
    const uint4 u = *reinterpret_cast<const uint4*>(src + (size_t)r * ld + c);
    uint32_t* w = reinterpret_cast<uint32_t*>(dst + (size_t)r * dstride + c);
    w[0] = u.x; w[1] = u.y; w[2] = u.z; w[3] = u.w;
  }
}

__device__ __forceinline__ float dot_row(const float* qs, const bf16* krow, int d) {
  // qs: fp32 [d] (broadcast reads), krow: bf16 row in smem (4-byte aligned)
  const __nv_bfloat162* k2 = reinterpret_cast<const __nv_bfloat162*>(krow);
  float acc = 0.f;
#pragma unroll 8
  for (int c = 0; c < d / 2; ++c) {
    const float2 kf = __bfloat1622float2(k2[c]);
    acc = fmaf(qs[2 * c], kf.x, acc);
    acc = fmaf(qs[2 * c + 1], kf.y, acc);
  }
  return acc;
}

__global__ void __launch_bounds__(128) attn_fwd_kernel(AttnArgs a, bf16* __restrict__ ctx, int ldo, float* __restrict__ lse_out) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int h = blockIdx.x, b = blockIdx.y;
  const int d = a.d, ds = d + 2, Lk = a.Lk, Lq = a.Lq;
  const int lkp = (Lk + 7) & ~7;
  bf16* Ks = reinterpret_cast<bf16*>(smem);
  bf16* Vs = Ks + (size_t)Lk * ds;
  float* qs = reinterpret_cast<float*>(Vs + (size_t)Lk * ds);  // 4*Lk*ds bytes in: 4 B aligned
  float* ps = qs + 4 * d;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  load_tile(Ks, ds, a.kv + (size_t)b * Lk * a.ldkv + a.k_off + h * d, a.ldkv, Lk, d);
  load_tile(Vs, ds, a.kv + (size_t)b * Lk * a.ldkv + a.v_off + h * d, a.ldkv, Lk, d);
  __syncthreads();
  const float* mrow = a.mask + (size_t)b * Lk;
  float* q_w = qs + warp * d;
  float* p_w = ps + warp * lkp;
  for (int i = warp; i < Lq; i += 4) {
    const bf16* qg = a.q + (size_t)(b * Lq + i) * a.ldq + a.q_off + h * d;
    for (int c = lane; c < d; c += 32) q_w[c] = ld_f(qg + c) * a.scale;
    __syncwarp();
    float s[MAX_NJ];
    float mx = -3.0e38f;
#pragma unroll
    for (int t = 0; t < MAX_NJ; ++t) {
      const int j = lane + 32 * t;
      s[t] = -3.0e38f;
      if (j < Lk) {
        s[t] = dot_row(q_w, Ks + (size_t)j * ds, d) + mrow[j];
        mx = fmaxf(mx, s[t]);
      }
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int t = 0; t < MAX_NJ; ++t) {
      const int j = lane + 32 * t;
      if (j < Lk) {
        s[t] = __expf(s[t] - mx);
        sum += s[t];
      }
    }
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
    const size_t prow = (size_t)(b * a.H + h) * Lq + i;
    if (lane == 0) lse_out[prow] = mx + __logf(sum);
#pragma unroll
    for (int t = 0; t < MAX_NJ; ++t) {
      const int j = lane + 32 * t;
      if (j < Lk) {
        float p = s[t] * inv;
        if (a.drop.thresh) p *= drop_mult_at(a.drop, (unsigned long long)prow * lkp + j);
        p_w[j] = p;
      }
    }
    __syncwarp();
    bf16* og = ctx + (size_t)(b * Lq + i) * ldo + h * d;
    for (int c = lane; c < d; c += 32) {
      float o = 0.f;
      for (int j = 0; j < Lk; ++j) o = fmaf(p_w[j], __bfloat162float(Vs[(size_t)j * ds + c]), o);
      og[c] = __float2bfloat16(o);
    }
    __syncwarp();
  }
}

// Backward.  P is recomputed from Q, K and the saved log-sum-exp; D_i = sum_j p_ij dp_ij.
//   dV_j = sum_i (p_ij m_ij) dO_i ;  dp_ij = m_ij (dO_i . V_j) ;  dS_ij = p_ij (dp_ij - D_i)
//   dQ_i = scale * sum_j dS_ij K_j ;  dK_j = scale * sum_i dS_ij Q_i
__global__ void __launch_bounds__(256) attn_bwd_kernel(AttnArgs a, const bf16* __restrict__ dctx, int ldo,
                                                       const float* __restrict__ lse, bf16* __restrict__ dq, int lddq,
                                                       bf16* __restrict__ dkv, int lddkv) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int h = blockIdx.x, b = blockIdx.y;
  const int d = a.d, ds = d + 2, Lk = a.Lk, Lq = a.Lq;
  const int lkp = (Lk + 7) & ~7;
  const int lks = lkp + 2;
  bf16* Qs = reinterpret_cast<bf16*>(smem);
  bf16* dOs = Qs + (size_t)Lq * ds;
  bf16* Ks = dOs + (size_t)Lq * ds;
  bf16* Vs = Ks + (size_t)Lk * ds;
  bf16* PD = Vs + (size_t)Lk * ds;
  bf16* DS = PD + (size_t)Lq * lks;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  load_tile(Qs, ds, a.q + (size_t)b * Lq * a.ldq + a.q_off + h * d, a.ldq, Lq, d);
  load_tile(dOs, ds, dctx + (size_t)b * Lq * ldo + h * d, ldo, Lq, d);
  load_tile(Ks, ds, a.kv + (size_t)b * Lk * a.ldkv + a.k_off + h * d, a.ldkv, Lk, d);
  load_tile(Vs, ds, a.kv + (size_t)b * Lk * a.ldkv + a.v_off + h * d, a.ldkv, Lk, d);
  __syncthreads();
  const float* mrow = a.mask + (size_t)b * Lk;
  for (int i = warp; i < Lq; i += 8) {
    const size_t prow = (size_t)(b * a.H + h) * Lq + i;
    const float l = lse[prow];
    const __nv_bfloat162* q2 = reinterpret_cast<const __nv_bfloat162*>(Qs + (size_t)i * ds);
    const __nv_bfloat162* o2 = reinterpret_cast<const __nv_bfloat162*>(dOs + (size_t)i * ds);
    float p[4], dp[4], mlt[4];
    float D = 0.f;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int j = lane + 32 * t;
      p[t] = dp[t] = 0.f;
      mlt[t] = 1.f;
      if (j < Lk) {
        const __nv_bfloat162* k2 = reinterpret_cast<const __nv_bfloat162*>(Ks + (size_t)j * ds);
        const __nv_bfloat162* v2 = reinterpret_cast<const __nv_bfloat162*>(Vs + (size_t)j * ds);
        float s = 0.f, g = 0.f;
#pragma unroll 8
        for (int c = 0; c < d / 2; ++c) {
          const float2 qf = __bfloat1622float2(q2[c]), kf = __bfloat1622float2(k2[c]);
          const float2 of = __bfloat1622float2(o2[c]), vf = __bfloat1622float2(v2[c]);
          s = fmaf(qf.x, kf.x, s); s = fmaf(qf.y, kf.y, s);
          g = fmaf(of.x, vf.x, g); g = fmaf(of.y, vf.y, g);
        }
        p[t] = __expf(s * a.scale + mrow[j] - l);
        if (a.drop.thresh) mlt[t] = drop_mult_at(a.drop, (unsigned long long)prow * lkp + j);
        dp[t] = g * mlt[t];
        D += p[t] * dp[t];
      }
    }
    D = warp_sum(D);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int j = lane + 32 * t;
      if (j < Lk) {
        PD[(size_t)i * lks + j] = __float2bfloat16(p[t] * mlt[t]);
        DS[(size_t)i * lks + j] = __float2bfloat16(p[t] * (dp[t] - D));
      }
    }
  }
  __syncthreads();
  for (int j = warp; j < Lk; j += 8) {
    bf16* dk = dkv + (size_t)(b * Lk + j) * lddkv + a.k_off + h * d;
    bf16* dv = dkv + (size_t)(b * Lk + j) * lddkv + a.v_off + h * d;
    for (int c = lane; c < d; c += 32) {
      float accv = 0.f, acck = 0.f;
      for (int i = 0; i < Lq; ++i) {
        accv = fmaf(__bfloat162float(PD[(size_t)i * lks + j]), __bfloat162float(dOs[(size_t)i * ds + c]), accv);
        acck = fmaf(__bfloat162float(DS[(size_t)i * lks + j]), __bfloat162float(Qs[(size_t)i * ds + c]), acck);
      }
      dv[c] = __float2bfloat16(accv);
      dk[c] = __float2bfloat16(acck * a.scale);
    }
  }
  for (int i = warp; i < Lq; i += 8) {
    bf16* dqr = dq + (size_t)(b * Lq + i) * lddq + a.q_off + h * d;
    for (int c = lane; c < d; c += 32) {
      float acc = 0.f;
      for (int j = 0; j < Lk; ++j)
        acc = fmaf(__bfloat162float(DS[(size_t)i * lks + j]), __bfloat162float(Ks[(size_t)j * ds + c]), acc);
      dqr[c] = __float2bfloat16(acc * a.scale);
    }
  }
}

int fill_args(AttnArgs* a, const void* q, int ldq, int q_off, const void* kv, int ldkv, int k_off, int v_off,
              const float* mask, int B, int H, int Lq, int Lk, int d, const void* rng, unsigned int site, float p) {
  K3M_REQUIRE(B > 0 && H > 0 && Lq > 0 && Lk > 0, "attn: empty problem");
  K3M_REQUIRE(d % 8 == 0 && d <= 256, "attn: head size must be a multiple of 8 and <= 256 (d=%d)", d);
  K3M_REQUIRE(ldq % 8 == 0 && ldkv % 8 == 0 && q_off % 8 == 0 && k_off % 8 == 0 && v_off % 8 == 0,
              "attn: strides / offsets must be multiples of 8 elements");
  a->q = (const bf16*)q; a->ldq = ldq; a->q_off = q_off;
  a->kv = (const bf16*)kv; a->ldkv = ldkv; a->k_off = k_off; a->v_off = v_off;
  a->mask = mask;
  a->B = B; a->H = H; a->Lq = Lq; a->Lk = Lk; a->d = d;
  a->scale = 1.0f / sqrtf((float)d);
  a->drop.rng = (const unsigned long long*)rng;
  a->drop.site = site;
  a->drop.thresh = (rng && p > 0.f) ? (unsigned int)(p * 65536.0f + 0.5f) : 0u;
  a->drop.scale = 1.0f / (1.0f - p);
  return K3M_OK;
}

// ---- tensor-core path dispatch (d in {16, 64, 96, 128}; up to 128 x 128 tokens per head) ----------------
template <int D, int NKT>
int launch_fwd_mma(const AttnArgs& a, bf16* ctx, int ldo, float* lse, cudaStream_t st) {
  const int nw = (a.Lq + 15) / 16;
  const size_t smem = (size_t)(nw * 16 + 2 * NKT * 8) * (D + 8) * 2 + nw * 16 * 16;    // + dropout keep-bytes
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(attn_fwd_mma<D, NKT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 16 * 3 * (D + 8) * 2 + 8 * 16 * 16);
    configured = true;
  }
  attn_fwd_mma<D, NKT><<<dim3(a.H, a.B), nw * 32, smem, st>>>(a, ctx, ldo, lse);
  return k3m_check_launch("k3m_attn_fwd");
}

template <int D, int NKT>
int launch_bwd_mma(const AttnArgs& a, const bf16* dctx, int ldo, const float* lse, bf16* dq, int lddq, bf16* dkv,
                   int lddkv, cudaStream_t st) {
  const int nw = (a.Lq + 15) / 16;
  const size_t smem = (size_t)2 * (nw * 16 + NKT * 8) * (D + 8) * 2 + (size_t)2 * nw * 16 * (NKT * 8 + 8) * 2 + nw * 16 * 16;
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(attn_bwd_mma<D, NKT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         2 * (128 + NKT * 8) * (D + 8) * 2 + 2 * 128 * (NKT * 8 + 8) * 2 + 8 * 16 * 16);
    configured = true;
  }
  attn_bwd_mma<D, NKT><<<dim3(a.H, a.B), nw * 32, smem, st>>>(a, dctx, ldo, lse, dq, lddq, dkv, lddkv);
  return k3m_check_launch("k3m_attn_bwd");
}

bool mma_ok(const AttnArgs& a) {
  return (a.d == 16 || a.d == 64 || a.d == 96 || a.d == 128) && a.Lq <= 128 && a.Lk <= 128;
}

#define K3M_ATTN_DISPATCH(FN, ...)                                       \
  do {                                                                   \
    const bool small = a.Lk <= 48;                                       \
    switch (a.d) {                                                       \
      case 16: return small ? FN<16, 6>(__VA_ARGS__) : FN<16, 16>(__VA_ARGS__);    \
      case 64: return small ? FN<64, 6>(__VA_ARGS__) : FN<64, 16>(__VA_ARGS__);    \
      case 96: return small ? FN<96, 6>(__VA_ARGS__) : FN<96, 16>(__VA_ARGS__);    \
      default: return small ? FN<128, 6>(__VA_ARGS__) : FN<128, 16>(__VA_ARGS__);  \
    }                                                                    \
  } while (0)

int fwd_mma(const AttnArgs& a, bf16* ctx, int ldo, float* lse, cudaStream_t st) {
  K3M_ATTN_DISPATCH(launch_fwd_mma, a, ctx, ldo, lse, st);
}
int bwd_mma(const AttnArgs& a, const bf16* dctx, int ldo, const float* lse, bf16* dq, int lddq, bf16* dkv, int lddkv,
            cudaStream_t st) {
  K3M_ATTN_DISPATCH(launch_bwd_mma, a, dctx, ldo, lse, dq, lddq, dkv, lddkv, st);
}

// ---- long sequences: flash-style kernels, any Lq / Lk, d in {16, 64, 96, 128} --------------------------
bool long_ok(const AttnArgs& a) { return a.d == 16 || a.d == 64 || a.d == 96 || a.d == 128; }

template <int D>
int launch_fwd_long(const AttnArgs& a, bf16* ctx, int ldo, float* lse, cudaStream_t st) {
  const size_t smem = (size_t)(LONG_BQ + 4 * LONG_BK) * (D + 8) * 2 + 4 * 16 * 8;   // Q + two (K, V) buffers + keep-bytes
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(attn_fwd_long<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = true;
  }
  attn_fwd_long<D><<<dim3(a.H, a.B, (a.Lq + LONG_BQ - 1) / LONG_BQ), 128, smem, st>>>(a, ctx, ldo, lse);
  return k3m_check_launch("k3m_attn_fwd");
}

template <int D>
int launch_bwd_long(const AttnArgs& a, const bf16* dctx, int ldo, float* lse, bf16* dq, int lddq, bf16* dkv,
                    int lddkv, cudaStream_t st) {
  float* dsum = lse + (size_t)a.B * a.H * a.Lq;      // second half of the [2, B*H*Lq] buffer
  const size_t smem_q = (size_t)(2 * LONG_BQ + 4 * LONG_BK) * (D + 8) * 2 + 4 * 16 * 8;   // Q, dO + two (K, V) + keep-bytes
  const size_t smem_k = (size_t)(2 * LONG_BK + 4 * LONG_BQ2) * (D + 8) * 2 + 4 * LONG_BQ2 * sizeof(float) +
                        2 * LONG_BQ2 * 8;                                                  // K, V + two (Q, dO, lse, D, keep-bytes)
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(attn_bwd_long_dq<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_q);
    cudaFuncSetAttribute(attn_bwd_long_dkv<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_k);
    configured = true;
  }
  attn_bwd_long_dq<D><<<dim3(a.H, a.B, (a.Lq + LONG_BQ - 1) / LONG_BQ), 128, smem_q, st>>>(a, dctx, ldo, lse, dsum,
                                                                                           dq, lddq);
  int rc = k3m_check_launch("k3m_attn_bwd(dq)");
  if (rc) return rc;
  attn_bwd_long_dkv<D><<<dim3(a.H, a.B, (a.Lk + LONG_BK - 1) / LONG_BK), 128, smem_k, st>>>(a, dctx, ldo, lse, dsum,
                                                                                            dkv, lddkv);
  return k3m_check_launch("k3m_attn_bwd(dkv)");
}

#define K3M_ATTN_LONG_DISPATCH(FN, ...)                \
  do {                                                 \
    switch (a.d) {                                     \
      case 16: return FN<16>(__VA_ARGS__);             \
      case 64: return FN<64>(__VA_ARGS__);             \
      case 96: return FN<96>(__VA_ARGS__);             \
      default: return FN<128>(__VA_ARGS__);            \
    }                                                  \
  } while (0)

int fwd_long(const AttnArgs& a, bf16* ctx, int ldo, float* lse, cudaStream_t st) {
  K3M_ATTN_LONG_DISPATCH(launch_fwd_long, a, ctx, ldo, lse, st);
}
int bwd_long(const AttnArgs& a, const bf16* dctx, int ldo, float* lse, bf16* dq, int lddq, bf16* dkv, int lddkv,
             cudaStream_t st) {
  K3M_ATTN_LONG_DISPATCH(launch_bwd_long, a, dctx, ldo, lse, dq, lddq, dkv, lddkv, st);
}

}  // namespace

extern "C" int k3m_attn_fwd(const void* q, int ldq, int q_off, const void* kv, int ldkv, int k_off, int v_off,
                            const float* mask, int B, int H, int Lq, int Lk, int d, void* ctx, int ldo, float* lse,
                            const void* rng, unsigned int site, float drop_p, void* stream) {
  AttnArgs a;
  int rc = fill_args(&a, q, ldq, q_off, kv, ldkv, k_off, v_off, mask, B, H, Lq, Lk, d, rng, site, drop_p);
  if (rc) return rc;
  if (mma_ok(a)) return fwd_mma(a, (bf16*)ctx, ldo, lse, reinterpret_cast<cudaStream_t>(stream));
  if (long_ok(a)) return fwd_long(a, (bf16*)ctx, ldo, lse, reinterpret_cast<cudaStream_t>(stream));
  K3M_REQUIRE(Lk <= 32 * MAX_NJ, "attn_fwd: at most %d keys (Lk=%d)", 32 * MAX_NJ, Lk);
  const int lkp = (Lk + 7) & ~7;
  const size_t smem = (size_t)2 * Lk * (d + 2) * 2 + 4 + (size_t)4 * d * 4 + (size_t)4 * lkp * 4;
  K3M_REQUIRE(smem <= 220 * 1024, "attn_fwd: tile does not fit shared memory (%zu B)", smem);
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    configured = 220 * 1024;
  }
  attn_fwd_kernel<<<dim3(H, B), 128, smem, reinterpret_cast<cudaStream_t>(stream)>>>(a, (bf16*)ctx, ldo, lse);
  return k3m_check_launch("k3m_attn_fwd");
}

extern "C" int k3m_attn_bwd(const void* dctx, int ldo, const void* q, int ldq, int q_off, const void* kv, int ldkv,
                            int k_off, int v_off, const float* mask, float* lse, void* dq, int lddq, void* dkv,
                            int lddkv, int B, int H, int Lq, int Lk, int d, const void* rng, unsigned int site,
                            float drop_p, void* stream) {
  AttnArgs a;
  int rc = fill_args(&a, q, ldq, q_off, kv, ldkv, k_off, v_off, mask, B, H, Lq, Lk, d, rng, site, drop_p);
  if (rc) return rc;
  K3M_REQUIRE(lddq % 8 == 0 && lddkv % 8 == 0 && ldo % 8 == 0, "attn_bwd: strides must be multiples of 8");
  if (mma_ok(a))
    return bwd_mma(a, (const bf16*)dctx, ldo, lse, (bf16*)dq, lddq, (bf16*)dkv, lddkv, reinterpret_cast<cudaStream_t>(stream));
  if (long_ok(a))
    return bwd_long(a, (const bf16*)dctx, ldo, lse, (bf16*)dq, lddq, (bf16*)dkv, lddkv, reinterpret_cast<cudaStream_t>(stream));
  K3M_REQUIRE(Lk <= 128 && Lq <= 128, "attn_bwd: at most 128 x 128 tokens per head (Lq=%d Lk=%d)", Lq, Lk);
  K3M_REQUIRE(lddq % 8 == 0 && lddkv % 8 == 0 && ldo % 8 == 0, "attn_bwd: strides must be multiples of 8");
  const int lkp = (Lk + 7) & ~7;
  const size_t smem = (size_t)2 * (Lq + Lk) * (d + 2) * 2 + (size_t)2 * Lq * (lkp + 2) * 2;
  K3M_REQUIRE(smem <= 220 * 1024, "attn_bwd: tile does not fit shared memory (%zu B)", smem);
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    configured = 220 * 1024;
  }
  attn_bwd_kernel<<<dim3(H, B), 256, smem, reinterpret_cast<cudaStream_t>(stream)>>>(a, (const bf16*)dctx, ldo, lse,
                                                                                      (bf16*)dq, lddq, (bf16*)dkv, lddkv);
  return k3m_check_launch("k3m_attn_bwd");
}
