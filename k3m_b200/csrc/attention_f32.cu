// fp32 CHECK build only (-DK3M_ACT_F32, libk3m_sm100a_f32.so): k3m_attn_fwd / k3m_attn_bwd of include/k3m.h with
// fp32 Q / K / V / context and fp32 arithmetic throughout (no rounding of P), any Lq / Lk / d.  Same contract and the
// same dropout counter stream as attention.cu (element (row p, key j) -> p * round_up(Lk, 8) + j, p = (b*H+h)*Lq+i).
// One CTA per (batch, head); a warp owns a query row (forward, dQ) or a key row (dK / dV); every q.k and dO.v
// product is a lane-parallel dot over d with a shuffle reduction.  Tests only — not a product path, not tuned.
//   reference: BertSelfAttention.forward vilbert_k3m.py:440-464, BertBiAttention.forward :786-824
#ifdef K3M_ACT_F32
#include "common.cuh"

namespace {

struct A32 {
  const float* q; int ldq, q_off;
  const float* kv; int ldkv, k_off, v_off;
  const float* mask;
  int B, H, Lq, Lk, d;
  float scale;
  DropSpec drop;
};

__device__ __forceinline__ float dot_d(const float* a, const float* b, int d, int lane) {
  float s = 0.f;
  for (int c = lane; c < d; c += 32) s = fmaf(a[c], b[c], s);
  return warp_sum(s);
}

__global__ void __launch_bounds__(128) attn32_fwd(A32 a, float* __restrict__ ctx, int ldo, float* __restrict__ lse) {
  extern __shared__ float sm[];                 // [4 warps][Lk]
  const int h = blockIdx.x, b = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lkp = (a.Lk + 7) & ~7;
  float* sc = sm + (size_t)warp * a.Lk;
  const float* mrow = a.mask + (size_t)b * a.Lk;
  for (int i = warp; i < a.Lq; i += 4) {
    const float* qr = a.q + (size_t)(b * a.Lq + i) * a.ldq + a.q_off + h * a.d;
    float mx = -3.0e38f;
    for (int j = 0; j < a.Lk; ++j) {
      const float* kr = a.kv + (size_t)(b * a.Lk + j) * a.ldkv + a.k_off + h * a.d;
      const float s = dot_d(qr, kr, a.d, lane) * a.scale + mrow[j];
      if (lane == 0) sc[j] = s;
      mx = fmaxf(mx, s);
    }
    __syncwarp();
    float sum = 0.f;
    for (int j = lane; j < a.Lk; j += 32) sum += expf(sc[j] - mx);
    sum = warp_sum(sum);
    const size_t prow = (size_t)(b * a.H + h) * a.Lq + i;
    if (lane == 0) lse[prow] = mx + logf(sum);
    const float inv = 1.0f / sum;
    for (int j = lane; j < a.Lk; j += 32) {
      float p = expf(sc[j] - mx) * inv;
      if (a.drop.thresh) p *= drop_mult_at(a.drop, (unsigned long long)prow * lkp + j);
      sc[j] = p;
    }
    __syncwarp();
    float* og = ctx + (size_t)(b * a.Lq + i) * ldo + h * a.d;
    for (int c = lane; c < a.d; c += 32) {
      float o = 0.f;
      for (int j = 0; j < a.Lk; ++j)
        o = fmaf(sc[j], a.kv[(size_t)(b * a.Lk + j) * a.ldkv + a.v_off + h * a.d + c], o);
      og[c] = o;
    }
    __syncwarp();
  }
}

// P recomputed from the saved log-sum-exp.  m_ij = dropout multiplier.
//   dp_ij = m_ij (dO_i . V_j);  D_i = sum_j p_ij dp_ij;  dS_ij = p_ij (dp_ij - D_i)
//   dQ_i = scale sum_j dS_ij K_j;   dK_j = scale sum_i dS_ij Q_i;   dV_j = sum_i p_ij m_ij dO_i
// pass 1 (warp per query row): D_i -> lse row 1, dQ_i.   pass 2 (warp per key row): dK_j, dV_j.
__global__ void __launch_bounds__(128) attn32_bwd_q(A32 a, const float* __restrict__ dctx, int ldo, float* __restrict__ lse,
                                                    float* __restrict__ dq, int lddq) {
  extern __shared__ float sm[];                 // [4 warps][2][Lk]: p, dp
  const int h = blockIdx.x, b = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lkp = (a.Lk + 7) & ~7;
  float* ps = sm + (size_t)warp * 2 * a.Lk;
  float* dps = ps + a.Lk;
  const float* mrow = a.mask + (size_t)b * a.Lk;
  const size_t nrows = (size_t)a.B * a.H * a.Lq;
  for (int i = warp; i < a.Lq; i += 4) {
    const size_t prow = (size_t)(b * a.H + h) * a.Lq + i;
    const float* qr = a.q + (size_t)(b * a.Lq + i) * a.ldq + a.q_off + h * a.d;
    const float* dor = dctx + (size_t)(b * a.Lq + i) * ldo + h * a.d;
    const float l = lse[prow];
    float D = 0.f;
    for (int j = 0; j < a.Lk; ++j) {
      const float* kr = a.kv + (size_t)(b * a.Lk + j) * a.ldkv + a.k_off + h * a.d;
      const float* vr = a.kv + (size_t)(b * a.Lk + j) * a.ldkv + a.v_off + h * a.d;
      const float p = expf(dot_d(qr, kr, a.d, lane) * a.scale + mrow[j] - l);
      float dp = dot_d(dor, vr, a.d, lane);
      if (a.drop.thresh) dp *= drop_mult_at(a.drop, (unsigned long long)prow * lkp + j);
      if (lane == 0) { ps[j] = p; dps[j] = dp; }
      D = fmaf(p, dp, D);
    }
    __syncwarp();
    if (lane == 0) lse[nrows + prow] = D;
    float* dqr = dq + (size_t)(b * a.Lq + i) * lddq + a.q_off + h * a.d;
    for (int c = lane; c < a.d; c += 32) {
      float o = 0.f;
      for (int j = 0; j < a.Lk; ++j)
        o = fmaf(ps[j] * (dps[j] - D), a.kv[(size_t)(b * a.Lk + j) * a.ldkv + a.k_off + h * a.d + c], o);
      dqr[c] = o * a.scale;
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(128) attn32_bwd_kv(A32 a, const float* __restrict__ dctx, int ldo, const float* __restrict__ lse,
                                                     float* __restrict__ dkv, int lddkv) {
  extern __shared__ float sm[];                 // [4 warps][2][Lq]: dS, p*m
  const int h = blockIdx.x, b = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lkp = (a.Lk + 7) & ~7;
  float* dss = sm + (size_t)warp * 2 * a.Lq;
  float* pms = dss + a.Lq;
  const size_t nrows = (size_t)a.B * a.H * a.Lq;
  for (int j = warp; j < a.Lk; j += 4) {
    const float* kr = a.kv + (size_t)(b * a.Lk + j) * a.ldkv + a.k_off + h * a.d;
    const float* vr = a.kv + (size_t)(b * a.Lk + j) * a.ldkv + a.v_off + h * a.d;
    const float mj = a.mask[(size_t)b * a.Lk + j];
    for (int i = 0; i < a.Lq; ++i) {
      const size_t prow = (size_t)(b * a.H + h) * a.Lq + i;
      const float* qr = a.q + (size_t)(b * a.Lq + i) * a.ldq + a.q_off + h * a.d;
      const float* dor = dctx + (size_t)(b * a.Lq + i) * ldo + h * a.d;
      const float p = expf(dot_d(qr, kr, a.d, lane) * a.scale + mj - lse[prow]);
      float dp = dot_d(dor, vr, a.d, lane);
      float mlt = 1.0f;
      if (a.drop.thresh) mlt = drop_mult_at(a.drop, (unsigned long long)prow * lkp + j);
      dp *= mlt;
      if (lane == 0) { dss[i] = p * (dp - lse[nrows + prow]); pms[i] = p * mlt; }
    }
    __syncwarp();
    float* dk = dkv + (size_t)(b * a.Lk + j) * lddkv + a.k_off + h * a.d;
    float* dv = dkv + (size_t)(b * a.Lk + j) * lddkv + a.v_off + h * a.d;
    for (int c = lane; c < a.d; c += 32) {
      float ok = 0.f, ov = 0.f;
      for (int i = 0; i < a.Lq; ++i) {
        ok = fmaf(dss[i], a.q[(size_t)(b * a.Lq + i) * a.ldq + a.q_off + h * a.d + c], ok);
        ov = fmaf(pms[i], dctx[(size_t)(b * a.Lq + i) * ldo + h * a.d + c], ov);
      }
      dk[c] = ok * a.scale;
      dv[c] = ov;
    }
    __syncwarp();
  }
}

int fill(A32* a, const void* q, int ldq, int q_off, const void* kv, int ldkv, int k_off, int v_off, const float* mask,
         int B, int H, int Lq, int Lk, int d, const void* rng, unsigned int site, float p) {
  K3M_REQUIRE(B > 0 && H > 0 && Lq > 0 && Lk > 0 && d > 0, "attn: empty problem");
  K3M_REQUIRE(Lq <= 4096 && Lk <= 4096, "attn (fp32 check build): at most 4096 tokens");
  a->q = (const float*)q; a->ldq = ldq; a->q_off = q_off;
  a->kv = (const float*)kv; a->ldkv = ldkv; a->k_off = k_off; a->v_off = v_off;
  a->mask = mask;
  a->B = B; a->H = H; a->Lq = Lq; a->Lk = Lk; a->d = d;
  a->scale = 1.0f / sqrtf((float)d);
  a->drop.rng = (const unsigned long long*)rng;
  a->drop.site = site;
  a->drop.thresh = (rng && p > 0.f) ? (unsigned int)(p * 65536.0f + 0.5f) : 0u;
  a->drop.scale = 1.0f / (1.0f - p);
  return K3M_OK;
}

}  // namespace

extern "C" int k3m_attn_fwd(const void* q, int ldq, int q_off, const void* kv, int ldkv, int k_off, int v_off,
                            const float* mask, int B, int H, int Lq, int Lk, int d, void* ctx, int ldo, float* lse,
                            const void* rng, unsigned int site, float drop_p, void* stream) {
  A32 a;
  if (int rc = fill(&a, q, ldq, q_off, kv, ldkv, k_off, v_off, mask, B, H, Lq, Lk, d, rng, site, drop_p)) return rc;
  attn32_fwd<<<dim3(H, B), 128, (size_t)4 * Lk * sizeof(float), reinterpret_cast<cudaStream_t>(stream)>>>(a, (float*)ctx, ldo, lse);
  return k3m_check_launch("k3m_attn_fwd (fp32 check build)");
}

extern "C" int k3m_attn_bwd(const void* dctx, int ldo, const void* q, int ldq, int q_off, const void* kv, int ldkv,
                            int k_off, int v_off, const float* mask, float* lse, void* dq, int lddq, void* dkv,
                            int lddkv, int B, int H, int Lq, int Lk, int d, const void* rng, unsigned int site,
                            float drop_p, void* stream) {
  A32 a;
  if (int rc = fill(&a, q, ldq, q_off, kv, ldkv, k_off, v_off, mask, B, H, Lq, Lk, d, rng, site, drop_p)) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  attn32_bwd_q<<<dim3(H, B), 128, (size_t)8 * Lk * sizeof(float), st>>>(a, (const float*)dctx, ldo, lse, (float*)dq, lddq);
  attn32_bwd_kv<<<dim3(H, B), 128, (size_t)8 * Lq * sizeof(float), st>>>(a, (const float*)dctx, ldo, lse, (float*)dkv, lddkv);
  return k3m_check_launch("k3m_attn_bwd (fp32 check build)");
}
#endif  // K3M_ACT_F32
