// Tensor-core attention for the short K3M sequences: one CTA per (batch, head), one warp per 16 query
// rows, whole K / V of the head in shared memory, S = Q K^T and O = P V as warp-level bf16 MMAs
// (m16n8k16, fp32 accumulate) with the softmax done on the accumulator registers (quad shuffles).
// Tiles are far below the 128-row tcgen05 atom (L = 36..128, d = 64..128), so the 5th-gen tensor
// core path would spend its time on TMEM allocation and barriers; mma.sync keeps a (b, h) problem
// inside one small CTA and lets ~8 of them share an SM.
//
// Included by attention.cu inside its anonymous namespace (uses AttnArgs / DropSpec).
#pragma once

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void ldsm_x4(uint32_t r[4], const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_addr(p)));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t r[4], const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_addr(p)));
}
__device__ __forceinline__ void mma16816(float c[4], const uint32_t a[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// rows x D bf16 tile -> shared memory with row stride DS; rows >= valid are zero-filled
template <int D, int DS>
__device__ __forceinline__ void load_tile_z(bf16* dst, const bf16* src, int ld, int valid, int rows) {
  constexpr int VPR = D / 8;
  for (int idx = threadIdx.x; idx < rows * VPR; idx += blockDim.x) {
    const int r = idx / VPR, c = (idx % VPR) * 8;
    uint4 u = make_uint4(0u, 0u, 0u, 0u);
    if (r < valid) u = *reinterpret_cast<const uint4*>(src + (size_t)r * ld + c);
    *reinterpret_cast<uint4*>(dst + (size_t)r * DS + c) = u;
  }
}

// asynchronous global -> shared copies (LDGSTS): the next block streams in while the tensor cores work on
// the current one; src-size 0 zero-fills rows beyond the sequence
__device__ __forceinline__ void cp_async16(void* dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_addr(dst)), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async4(void* dst, const void* src, bool valid) {
  const int sz = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(smem_addr(dst)), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int D, int DS>
__device__ __forceinline__ void load_tile_async(bf16* dst, const bf16* src, int ld, int valid, int rows) {
  constexpr int VPR = D / 8;
  for (int idx = threadIdx.x; idx < rows * VPR; idx += blockDim.x) {
    const int r = idx / VPR, c = (idx % VPR) * 8;
    const bool ok = r < valid;
    cp_async16(dst + (size_t)r * DS + c, ok ? src + (size_t)r * ld + c : src, ok);
  }
}

// acc[t] (16 x 8 per key tile t) += A(16 x D rows i0.. of As) * B^T, B = rows of Bs (keys), both [.][DS]
template <int D, int DS, int NKT>
__device__ __forceinline__ void qk_mma(float (*acc)[4], const bf16* As, int i0, const bf16* Bs, int lane) {
#pragma unroll
  for (int kk = 0; kk < D / 16; ++kk) {
    uint32_t af[4];
    ldsm_x4(af, As + (size_t)(i0 + (lane & 15)) * DS + kk * 16 + (lane >> 4) * 8);
#pragma unroll
    for (int t = 0; t < NKT; t += 2) {
      uint32_t bfr[4];
      ldsm_x4(bfr, Bs + (size_t)(t * 8 + (lane & 7) + (lane >> 4) * 8) * DS + kk * 16 + ((lane >> 3) & 1) * 8);
      mma16816(acc[t], af, bfr[0], bfr[1]);
      mma16816(acc[t + 1], af, bfr[2], bfr[3]);
    }
  }
}

// ---- dropout masks --------------------------------------------------------------------------------------
// One Philox call yields the 8 x 16 random bits of one 8-key tile of one dropout row (element (row, j):
// counter (row * lkp + j) >> 3, see common.cuh).  In the accumulator layout those 8 elements sit in the 4
// lanes of a quad, so instead of every lane running the generator for its 2 columns the warp computes each
// (row, tile) ONCE, keeps the 8 keep-bits as a byte in shared memory and the lanes pick their bits.
__device__ __forceinline__ uint32_t drop_tile_bits(const DropSpec& dr, unsigned long long prow, int lkp, int t) {
  uint32_t w[4];
  rng_words(dr.rng, dr.site, (prow * (unsigned long long)lkp >> 3) + (unsigned long long)t, w);
  uint32_t m = 0;
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const uint32_t u = (w[e >> 1] >> (16 * (e & 1))) & 0xFFFFu;
    m |= (u >= dr.thresh ? 1u : 0u) << e;
  }
  return m;
}

// keep-bytes of a warp's 16 dropout rows x NT key tiles -> mw[16][NTP]; rows / tiles beyond the problem get 0xFF
template <int NT, int NTP>
__device__ __forceinline__ void warp_drop_bits(uint8_t* mw, const DropSpec& dr, unsigned long long prow_first,
                                               int rows_valid, int lkp, int tile0, int tiles_valid, int lane) {
  for (int c = lane; c < 16 * NT; c += 32) {
    const int r = c / NT, t = c % NT;
    uint32_t bits = 0xFFu;
    if (r < rows_valid && t < tiles_valid) bits = drop_tile_bits(dr, prow_first + r, lkp, tile0 + t);
    mw[r * NTP + t] = (uint8_t)bits;
  }
  __syncwarp();
}

// the keep-bytes of one row (NTP bytes, 8-byte aligned) as 32-bit words
template <int NTP>
__device__ __forceinline__ void load_drop_row(uint32_t (&bw)[NTP / 4], const uint8_t* row) {
  const uint2* p = reinterpret_cast<const uint2*>(row);
#pragma unroll
  for (int i = 0; i < NTP / 8; ++i) {
    const uint2 v = p[i];
    bw[2 * i] = v.x; bw[2 * i + 1] = v.y;
  }
}
// multipliers of this thread's columns (2q, 2q + 1) of key tile t
template <int NTP>
__device__ __forceinline__ void drop_pair(const uint32_t (&bw)[NTP / 4], int t, int q, float scale, float& m0,
                                          float& m1) {
  const uint32_t bits = bw[t >> 2] >> (8 * (t & 3) + 2 * q);
  m0 = (bits & 1u) ? scale : 0.f;
  m1 = (bits & 2u) ? scale : 0.f;
}

template <int D, int NKT>
__global__ void __launch_bounds__(256) attn_fwd_mma(AttnArgs a, bf16* __restrict__ ctx, int ldo, float* __restrict__ lse_out) {
  constexpr int DS = D + 8, LKP = NKT * 8;
  extern __shared__ __align__(16) uint8_t smem[];
  const int h = blockIdx.x, b = blockIdx.y;
  const int Lq = a.Lq, Lk = a.Lk;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int LQP = nw * 16;
  bf16* Qs = reinterpret_cast<bf16*>(smem);
  bf16* Ks = Qs + (size_t)LQP * DS;
  bf16* Vs = Ks + (size_t)LKP * DS;
  constexpr int NTP = NKT <= 8 ? 8 : 16;
  uint8_t* mw = reinterpret_cast<uint8_t*>(Vs + (size_t)LKP * DS) + warp * 16 * NTP;
  // all three tiles in flight at once (cp.async): a register-staged copy loop serialises one global-load
  // latency per 16 bytes and thread, which was most of this CTA's life
  load_tile_async<D, DS>(Qs, a.q + (size_t)b * Lq * a.ldq + a.q_off + h * D, a.ldq, Lq, LQP);
  load_tile_async<D, DS>(Ks, a.kv + (size_t)b * Lk * a.ldkv + a.k_off + h * D, a.ldkv, Lk, LKP);
  load_tile_async<D, DS>(Vs, a.kv + (size_t)b * Lk * a.ldkv + a.v_off + h * D, a.ldkv, Lk, LKP);
  cp_async_commit();
  const int g = lane >> 2, q = lane & 3, i0 = warp * 16;
  const int lkp = (Lk + 7) & ~7;
  if (a.drop.thresh)     // while the tile loads are in flight
    warp_drop_bits<NKT, NTP>(mw, a.drop, (unsigned long long)(b * a.H + h) * Lq + i0, Lq - i0, lkp, 0, lkp >> 3, lane);
  cp_async_wait<0>();
  __syncthreads();
  float s[NKT][4];
#pragma unroll
  for (int t = 0; t < NKT; ++t) s[t][0] = s[t][1] = s[t][2] = s[t][3] = 0.f;
  qk_mma<D, DS, NKT>(s, Qs, i0, Ks, lane);
  const float* mrow = a.mask + (size_t)b * Lk;
  float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
  for (int t = 0; t < NKT; ++t) {
    const int j = t * 8 + 2 * q;
    const float ma = j < Lk ? mrow[j] : -INFINITY, mb = j + 1 < Lk ? mrow[j + 1] : -INFINITY;
    s[t][0] = fmaf(s[t][0], a.scale, ma); s[t][1] = fmaf(s[t][1], a.scale, mb);
    s[t][2] = fmaf(s[t][2], a.scale, ma); s[t][3] = fmaf(s[t][3], a.scale, mb);
    mx0 = fmaxf(mx0, fmaxf(s[t][0], s[t][1]));
    mx1 = fmaxf(mx1, fmaxf(s[t][2], s[t][3]));
  }
  mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
  mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
  float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
  for (int t = 0; t < NKT; ++t) {
    s[t][0] = __expf(s[t][0] - mx0); s[t][1] = __expf(s[t][1] - mx0);
    s[t][2] = __expf(s[t][2] - mx1); s[t][3] = __expf(s[t][3] - mx1);
    sum0 += s[t][0] + s[t][1];
    sum1 += s[t][2] + s[t][3];
  }
  sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1); sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
  sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1); sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
  const int r0 = i0 + g, r1 = i0 + g + 8;
  const unsigned long long prow0 = (unsigned long long)(b * a.H + h) * Lq + r0, prow1 = prow0 + 8;
  if (q == 0) {
    if (r0 < Lq) lse_out[prow0] = mx0 + __logf(sum0);
    if (r1 < Lq) lse_out[prow1] = mx1 + __logf(sum1);
  }
  const float inv0 = 1.0f / sum0, inv1 = 1.0f / sum1;
  uint32_t bw0[NTP / 4], bw1[NTP / 4];
  if (a.drop.thresh) {
    load_drop_row<NTP>(bw0, mw + g * NTP);
    load_drop_row<NTP>(bw1, mw + (g + 8) * NTP);
  }
#pragma unroll
  for (int t = 0; t < NKT; ++t) {
    float m00 = 1.f, m01 = 1.f, m10 = 1.f, m11 = 1.f;
    if (a.drop.thresh) {
      drop_pair<NTP>(bw0, t, q, a.drop.scale, m00, m01);
      drop_pair<NTP>(bw1, t, q, a.drop.scale, m10, m11);
    }
    s[t][0] *= inv0 * m00; s[t][1] *= inv0 * m01;
    s[t][2] *= inv1 * m10; s[t][3] *= inv1 * m11;
  }
  // O = P V : P from the accumulator registers (bf16), V^T fragments by transposing ldmatrix
  float o[D / 8][4];
#pragma unroll
  for (int c = 0; c < D / 8; ++c) o[c][0] = o[c][1] = o[c][2] = o[c][3] = 0.f;
#pragma unroll
  for (int kt = 0; kt < NKT / 2; ++kt) {
    uint32_t pf[4];
    pf[0] = pack_bf16(s[2 * kt][0], s[2 * kt][1]);
    pf[1] = pack_bf16(s[2 * kt][2], s[2 * kt][3]);
    pf[2] = pack_bf16(s[2 * kt + 1][0], s[2 * kt + 1][1]);
    pf[3] = pack_bf16(s[2 * kt + 1][2], s[2 * kt + 1][3]);
#pragma unroll
    for (int c = 0; c < D / 8; c += 2) {
      uint32_t vf[4];
      ldsm_x4_t(vf, Vs + (size_t)(kt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * DS + c * 8 + (lane >> 4) * 8);
      mma16816(o[c], pf, vf[0], vf[1]);
      mma16816(o[c + 1], pf, vf[2], vf[3]);
    }
  }
  bf16* o0 = ctx + (size_t)(b * Lq + r0) * ldo + h * D + 2 * q;
  bf16* o1 = ctx + (size_t)(b * Lq + r1) * ldo + h * D + 2 * q;
#pragma unroll
  for (int c = 0; c < D / 8; ++c) {
    if (r0 < Lq) *reinterpret_cast<uint32_t*>(o0 + c * 8) = pack_bf16(o[c][0], o[c][1]);
    if (r1 < Lq) *reinterpret_cast<uint32_t*>(o1 + c * 8) = pack_bf16(o[c][2], o[c][3]);
  }
}

// Backward.  P is recomputed from Q, K and the saved log-sum-exp.
//   dp_ij = m_ij (dO_i . V_j) ; D_i = sum_j p_ij dp_ij ; dS_ij = p_ij (dp_ij - D_i)
//   dQ = scale * dS K ;  dK = scale * dS^T Q ;  dV = (P m)^T dO
// Phase 1: one warp per 16 query rows (S, dP, dS, dQ; P m and dS go to shared memory as bf16).
// Phase 2: one warp per 16 key rows (dV, dK) with the transposed operands read by ldmatrix.trans.
template <int D, int NKT>
__global__ void __launch_bounds__(256) attn_bwd_mma(AttnArgs a, const bf16* __restrict__ dctx, int ldo,
                                                    const float* __restrict__ lse, bf16* __restrict__ dq, int lddq,
                                                    bf16* __restrict__ dkv, int lddkv) {
  constexpr int DS = D + 8, LKP = NKT * 8, PS = LKP + 8;
  extern __shared__ __align__(16) uint8_t smem[];
  const int h = blockIdx.x, b = blockIdx.y;
  const int Lq = a.Lq, Lk = a.Lk;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int LQP = nw * 16;
  bf16* Qs = reinterpret_cast<bf16*>(smem);
  bf16* dOs = Qs + (size_t)LQP * DS;
  bf16* Ks = dOs + (size_t)LQP * DS;
  bf16* Vs = Ks + (size_t)LKP * DS;
  bf16* Ps = Vs + (size_t)LKP * DS;      // [LQP][PS]  dropped probabilities
  bf16* Ss = Ps + (size_t)LQP * PS;      // [LQP][PS]  dS
  constexpr int NTP = NKT <= 8 ? 8 : 16;
  uint8_t* mw = reinterpret_cast<uint8_t*>(Ss + (size_t)LQP * PS) + warp * 16 * NTP;
  load_tile_async<D, DS>(Qs, a.q + (size_t)b * Lq * a.ldq + a.q_off + h * D, a.ldq, Lq, LQP);
  load_tile_async<D, DS>(dOs, dctx + (size_t)b * Lq * ldo + h * D, ldo, Lq, LQP);
  load_tile_async<D, DS>(Ks, a.kv + (size_t)b * Lk * a.ldkv + a.k_off + h * D, a.ldkv, Lk, LKP);
  load_tile_async<D, DS>(Vs, a.kv + (size_t)b * Lk * a.ldkv + a.v_off + h * D, a.ldkv, Lk, LKP);
  cp_async_commit();
  const int g = lane >> 2, q = lane & 3, i0 = warp * 16;
  const int lkp = (Lk + 7) & ~7;
  if (a.drop.thresh)
    warp_drop_bits<NKT, NTP>(mw, a.drop, (unsigned long long)(b * a.H + h) * Lq + i0, Lq - i0, lkp, 0, lkp >> 3, lane);
  cp_async_wait<0>();
  __syncthreads();
  {
    float s[NKT][4], dp[NKT][4];
#pragma unroll
    for (int t = 0; t < NKT; ++t) {
      s[t][0] = s[t][1] = s[t][2] = s[t][3] = 0.f;
      dp[t][0] = dp[t][1] = dp[t][2] = dp[t][3] = 0.f;
    }
    qk_mma<D, DS, NKT>(s, Qs, i0, Ks, lane);
    qk_mma<D, DS, NKT>(dp, dOs, i0, Vs, lane);
    const int r0 = i0 + g, r1 = i0 + g + 8;
    const unsigned long long prow0 = (unsigned long long)(b * a.H + h) * Lq + r0, prow1 = prow0 + 8;
    const float l0 = r0 < Lq ? lse[prow0] : 0.f, l1 = r1 < Lq ? lse[prow1] : 0.f;
    const float* mrow = a.mask + (size_t)b * Lk;
    float D0 = 0.f, D1 = 0.f;
    uint32_t pm[NKT][2];     // dropped P, packed, kept for the smem store
    uint32_t bw0[NTP / 4], bw1[NTP / 4];
    if (a.drop.thresh) {
      load_drop_row<NTP>(bw0, mw + g * NTP);
      load_drop_row<NTP>(bw1, mw + (g + 8) * NTP);
    }
#pragma unroll
    for (int t = 0; t < NKT; ++t) {
      const int j = t * 8 + 2 * q;
      const bool v0 = j < Lk, v1 = j + 1 < Lk;
      const float ma = v0 ? mrow[j] : 0.f, mb = v1 ? mrow[j + 1] : 0.f;
      float p00 = (v0 && r0 < Lq) ? __expf(fmaf(s[t][0], a.scale, ma) - l0) : 0.f;
      float p01 = (v1 && r0 < Lq) ? __expf(fmaf(s[t][1], a.scale, mb) - l0) : 0.f;
      float p10 = (v0 && r1 < Lq) ? __expf(fmaf(s[t][2], a.scale, ma) - l1) : 0.f;
      float p11 = (v1 && r1 < Lq) ? __expf(fmaf(s[t][3], a.scale, mb) - l1) : 0.f;
      float m00 = 1.f, m01 = 1.f, m10 = 1.f, m11 = 1.f;
      if (a.drop.thresh) {
        drop_pair<NTP>(bw0, t, q, a.drop.scale, m00, m01);
        drop_pair<NTP>(bw1, t, q, a.drop.scale, m10, m11);
      }
      dp[t][0] *= m00; dp[t][1] *= m01; dp[t][2] *= m10; dp[t][3] *= m11;
      D0 += p00 * dp[t][0] + p01 * dp[t][1];
      D1 += p10 * dp[t][2] + p11 * dp[t][3];
      pm[t][0] = pack_bf16(p00 * m00, p01 * m01);
      pm[t][1] = pack_bf16(p10 * m10, p11 * m11);
      s[t][0] = p00; s[t][1] = p01; s[t][2] = p10; s[t][3] = p11;
    }
    D0 += __shfl_xor_sync(0xffffffffu, D0, 1); D0 += __shfl_xor_sync(0xffffffffu, D0, 2);
    D1 += __shfl_xor_sync(0xffffffffu, D1, 1); D1 += __shfl_xor_sync(0xffffffffu, D1, 2);
    uint32_t dsf[NKT][2];
#pragma unroll
    for (int t = 0; t < NKT; ++t) {
      dsf[t][0] = pack_bf16(s[t][0] * (dp[t][0] - D0), s[t][1] * (dp[t][1] - D0));
      dsf[t][1] = pack_bf16(s[t][2] * (dp[t][2] - D1), s[t][3] * (dp[t][3] - D1));
      const int j = t * 8 + 2 * q;
      *reinterpret_cast<uint32_t*>(Ps + (size_t)r0 * PS + j) = pm[t][0];
      *reinterpret_cast<uint32_t*>(Ps + (size_t)r1 * PS + j) = pm[t][1];
      *reinterpret_cast<uint32_t*>(Ss + (size_t)r0 * PS + j) = dsf[t][0];
      *reinterpret_cast<uint32_t*>(Ss + (size_t)r1 * PS + j) = dsf[t][1];
    }
    // dQ = scale * dS K   (A = dS from registers, B[k = key][n = channel] by transposing ldmatrix on K)
    float o[D / 8][4];
#pragma unroll
    for (int c = 0; c < D / 8; ++c) o[c][0] = o[c][1] = o[c][2] = o[c][3] = 0.f;
#pragma unroll
    for (int kt = 0; kt < NKT / 2; ++kt) {
      uint32_t af[4] = {dsf[2 * kt][0], dsf[2 * kt][1], dsf[2 * kt + 1][0], dsf[2 * kt + 1][1]};
#pragma unroll
      for (int c = 0; c < D / 8; c += 2) {
        uint32_t kf[4];
        ldsm_x4_t(kf, Ks + (size_t)(kt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * DS + c * 8 + (lane >> 4) * 8);
        mma16816(o[c], af, kf[0], kf[1]);
        mma16816(o[c + 1], af, kf[2], kf[3]);
      }
    }
    bf16* q0 = dq + (size_t)(b * Lq + r0) * lddq + a.q_off + h * D + 2 * q;
    bf16* q1 = dq + (size_t)(b * Lq + r1) * lddq + a.q_off + h * D + 2 * q;
#pragma unroll
    for (int c = 0; c < D / 8; ++c) {
      if (r0 < Lq) *reinterpret_cast<uint32_t*>(q0 + c * 8) = pack_bf16(o[c][0] * a.scale, o[c][1] * a.scale);
      if (r1 < Lq) *reinterpret_cast<uint32_t*>(q1 + c * 8) = pack_bf16(o[c][2] * a.scale, o[c][3] * a.scale);
    }
  }
  __syncthreads();
  // phase 2: key tiles
  for (int jt = warp; jt < NKT / 2; jt += nw) {
    const int j0 = jt * 16;
    float dv[D / 8][4], dk[D / 8][4];
#pragma unroll
    for (int c = 0; c < D / 8; ++c) {
      dv[c][0] = dv[c][1] = dv[c][2] = dv[c][3] = 0.f;
      dk[c][0] = dk[c][1] = dk[c][2] = dk[c][3] = 0.f;
    }
    for (int it = 0; it < LQP / 16; ++it) {
      uint32_t pa[4], sa[4];
      const size_t off = (size_t)(it * 16 + (lane & 7) + (lane >> 4) * 8) * PS + j0 + ((lane >> 3) & 1) * 8;
      ldsm_x4_t(pa, Ps + off);
      ldsm_x4_t(sa, Ss + off);
#pragma unroll
      for (int c = 0; c < D / 8; c += 2) {
        uint32_t of[4], qf[4];
        const size_t boff = (size_t)(it * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * DS + c * 8 + (lane >> 4) * 8;
        ldsm_x4_t(of, dOs + boff);
        ldsm_x4_t(qf, Qs + boff);
        mma16816(dv[c], pa, of[0], of[1]);
        mma16816(dv[c + 1], pa, of[2], of[3]);
        mma16816(dk[c], sa, qf[0], qf[1]);
        mma16816(dk[c + 1], sa, qf[2], qf[3]);
      }
    }
    const int r0 = j0 + g, r1 = j0 + g + 8;
    bf16* k0 = dkv + (size_t)(b * Lk + r0) * lddkv + h * D + 2 * q;
    bf16* k1 = dkv + (size_t)(b * Lk + r1) * lddkv + h * D + 2 * q;
#pragma unroll
    for (int c = 0; c < D / 8; ++c) {
      if (r0 < Lk) {
        *reinterpret_cast<uint32_t*>(k0 + a.k_off + c * 8) = pack_bf16(dk[c][0] * a.scale, dk[c][1] * a.scale);
        *reinterpret_cast<uint32_t*>(k0 + a.v_off + c * 8) = pack_bf16(dv[c][0], dv[c][1]);
      }
      if (r1 < Lk) {
        *reinterpret_cast<uint32_t*>(k1 + a.k_off + c * 8) = pack_bf16(dk[c][2] * a.scale, dk[c][3] * a.scale);
        *reinterpret_cast<uint32_t*>(k1 + a.v_off + c * 8) = pack_bf16(dv[c][2], dv[c][3]);
      }
    }
  }
}

// ---- long sequences (more than 128 x 128 tokens per head) ------------------------------------------------
// The item-alignment finetune runs 256-token PV sequences (run_finetune_item_alignment.sh:40-43).  Same
// warp-level MMA building blocks, flash-style: a CTA owns 64 queries (forward, dQ) or 64 keys (dK / dV) and
// walks the other side in blocks through shared memory.
//   forward : online softmax over key blocks of 64 (running max / sum per row, O rescaled per block)
//   backward: D_i = sum_j p_ij dp_ij needs every key of row i, so the dQ kernel makes a first sweep for D
//             (stored in the second half of the lse buffer), a second for dQ; the dK / dV kernel computes the
//             TRANSPOSED tiles S^T = K Q^T and dP^T = V dO^T directly, so no probabilities are staged.
constexpr int LONG_BQ = 64, LONG_BK = 64;

template <int D>
__global__ void __launch_bounds__(128) attn_fwd_long(AttnArgs a, bf16* __restrict__ ctx, int ldo,
                                                     float* __restrict__ lse_out) {
  constexpr int DS = D + 8, NKT = LONG_BK / 8;
  extern __shared__ __align__(16) uint8_t smem[];
  const int h = blockIdx.x, b = blockIdx.y, q_base = blockIdx.z * LONG_BQ;
  const int Lq = a.Lq, Lk = a.Lk;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  bf16* Qs = reinterpret_cast<bf16*>(smem);
  bf16* KV = Qs + (size_t)LONG_BQ * DS;                             // [2 buffers][K | V][LONG_BK][DS]
  uint8_t* mw = reinterpret_cast<uint8_t*>(KV + (size_t)4 * LONG_BK * DS) + warp * 16 * 8;   // keep-bytes [16][8]
  const int nq = Lq - q_base < LONG_BQ ? Lq - q_base : LONG_BQ;
  auto load_block = [&](int k_base, int buf) {
    const int nk = Lk - k_base < LONG_BK ? Lk - k_base : LONG_BK;
    bf16* Kb = KV + (size_t)buf * 2 * LONG_BK * DS;
    load_tile_async<D, DS>(Kb, a.kv + (size_t)(b * Lk + k_base) * a.ldkv + a.k_off + h * D, a.ldkv, nk, LONG_BK);
    load_tile_async<D, DS>(Kb + (size_t)LONG_BK * DS, a.kv + (size_t)(b * Lk + k_base) * a.ldkv + a.v_off + h * D,
                           a.ldkv, nk, LONG_BK);
  };
  load_tile_async<D, DS>(Qs, a.q + (size_t)(b * Lq + q_base) * a.ldq + a.q_off + h * D, a.ldq, nq, LONG_BQ);
  load_block(0, 0);
  cp_async_commit();
  const int g = lane >> 2, q = lane & 3, i0 = warp * 16;
  const int r0 = q_base + i0 + g, r1 = r0 + 8;                      // query rows of this thread
  const unsigned long long prow0 = (unsigned long long)(b * a.H + h) * Lq + r0, prow1 = prow0 + 8;
  const float* mrow = a.mask + (size_t)b * Lk;
  const int lkp = (Lk + 7) & ~7;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  float o[D / 8][4];
#pragma unroll
  for (int c = 0; c < D / 8; ++c) o[c][0] = o[c][1] = o[c][2] = o[c][3] = 0.f;
  for (int k_base = 0, it = 0; k_base < Lk; k_base += LONG_BK, ++it) {
    if (k_base + LONG_BK < Lk) {
      load_block(k_base + LONG_BK, (it + 1) & 1);                   // buffer released by the barrier below
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    if (a.drop.thresh)     // this block's keep-bytes (the barrier at the end of the last block released mw)
      warp_drop_bits<NKT, 8>(mw, a.drop, prow0 - g, Lq - (q_base + i0), lkp, k_base >> 3, (lkp - k_base) >> 3, lane);
    __syncthreads();
    const bf16* Ks = KV + (size_t)(it & 1) * 2 * LONG_BK * DS;
    const bf16* Vs = Ks + (size_t)LONG_BK * DS;
    float s[NKT][4];
#pragma unroll
    for (int t = 0; t < NKT; ++t) s[t][0] = s[t][1] = s[t][2] = s[t][3] = 0.f;
    qk_mma<D, DS, NKT>(s, Qs, i0, Ks, lane);
    float bm0 = -INFINITY, bm1 = -INFINITY;
#pragma unroll
    for (int t = 0; t < NKT; ++t) {
      const int j = k_base + t * 8 + 2 * q;
      const float ma = j < Lk ? mrow[j] : -INFINITY, mb = j + 1 < Lk ? mrow[j + 1] : -INFINITY;
      s[t][0] = fmaf(s[t][0], a.scale, ma); s[t][1] = fmaf(s[t][1], a.scale, mb);
      s[t][2] = fmaf(s[t][2], a.scale, ma); s[t][3] = fmaf(s[t][3], a.scale, mb);
      bm0 = fmaxf(bm0, fmaxf(s[t][0], s[t][1]));
      bm1 = fmaxf(bm1, fmaxf(s[t][2], s[t][3]));
    }
    bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 1)); bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 2));
    bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 1)); bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 2));
    const float n0 = fmaxf(m0, bm0), n1 = fmaxf(m1, bm1);           // finite: every block holds a valid key
    const float c0 = __expf(m0 - n0), c1 = __expf(m1 - n1);         // exp(-inf) = 0 on the first block
    m0 = n0; m1 = n1;
    float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
    for (int t = 0; t < NKT; ++t) {
      s[t][0] = __expf(s[t][0] - n0); s[t][1] = __expf(s[t][1] - n0);
      s[t][2] = __expf(s[t][2] - n1); s[t][3] = __expf(s[t][3] - n1);
      sum0 += s[t][0] + s[t][1];
      sum1 += s[t][2] + s[t][3];
    }
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1); sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1); sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
    l0 = l0 * c0 + sum0;
    l1 = l1 * c1 + sum1;
#pragma unroll
    for (int c = 0; c < D / 8; ++c) { o[c][0] *= c0; o[c][1] *= c0; o[c][2] *= c1; o[c][3] *= c1; }
    if (a.drop.thresh) {
      uint32_t bw0[2], bw1[2];
      load_drop_row<8>(bw0, mw + g * 8);
      load_drop_row<8>(bw1, mw + (g + 8) * 8);
#pragma unroll
      for (int t = 0; t < NKT; ++t) {
        float m00, m01, m10, m11;
        drop_pair<8>(bw0, t, q, a.drop.scale, m00, m01);
        drop_pair<8>(bw1, t, q, a.drop.scale, m10, m11);
        s[t][0] *= m00; s[t][1] *= m01; s[t][2] *= m10; s[t][3] *= m11;
      }
    }
#pragma unroll
    for (int kt = 0; kt < NKT / 2; ++kt) {
      uint32_t pf[4];
      pf[0] = pack_bf16(s[2 * kt][0], s[2 * kt][1]);
      pf[1] = pack_bf16(s[2 * kt][2], s[2 * kt][3]);
      pf[2] = pack_bf16(s[2 * kt + 1][0], s[2 * kt + 1][1]);
      pf[3] = pack_bf16(s[2 * kt + 1][2], s[2 * kt + 1][3]);
#pragma unroll
      for (int c = 0; c < D / 8; c += 2) {
        uint32_t vf[4];
        ldsm_x4_t(vf, Vs + (size_t)(kt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * DS + c * 8 + (lane >> 4) * 8);
        mma16816(o[c], pf, vf[0], vf[1]);
        mma16816(o[c + 1], pf, vf[2], vf[3]);
      }
    }
    __syncthreads();                                                // every warp is done with this buffer
  }
  if (q == 0) {
    if (r0 < Lq) lse_out[prow0] = m0 + __logf(l0);
    if (r1 < Lq) lse_out[prow1] = m1 + __logf(l1);
  }
  const float inv0 = 1.0f / l0, inv1 = 1.0f / l1;
  bf16* o0 = ctx + (size_t)(b * Lq + r0) * ldo + h * D + 2 * q;
  bf16* o1 = ctx + (size_t)(b * Lq + r1) * ldo + h * D + 2 * q;
#pragma unroll
  for (int c = 0; c < D / 8; ++c) {
    if (r0 < Lq) *reinterpret_cast<uint32_t*>(o0 + c * 8) = pack_bf16(o[c][0] * inv0, o[c][1] * inv0);
    if (r1 < Lq) *reinterpret_cast<uint32_t*>(o1 + c * 8) = pack_bf16(o[c][2] * inv1, o[c][3] * inv1);
  }
}

// dQ for 64 queries of one (b, h); sweep 0 reduces D_i over all keys and stores it at dsum[prow] for
// the dK / dV kernel, sweep 1 forms dS = P (dP - D) block by block and accumulates dQ = scale dS K.
template <int D>
__global__ void __launch_bounds__(128) attn_bwd_long_dq(AttnArgs a, const bf16* __restrict__ dctx, int ldo,
                                                        const float* __restrict__ lse, float* __restrict__ dsum,
                                                        bf16* __restrict__ dq, int lddq) {
  constexpr int DS = D + 8, NKT = LONG_BK / 8;
  extern __shared__ __align__(16) uint8_t smem[];
  const int h = blockIdx.x, b = blockIdx.y, q_base = blockIdx.z * LONG_BQ;
  const int Lq = a.Lq, Lk = a.Lk;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  bf16* Qs = reinterpret_cast<bf16*>(smem);
  bf16* dOs = Qs + (size_t)LONG_BQ * DS;
  bf16* KV = dOs + (size_t)LONG_BQ * DS;                            // [2 buffers][K | V][LONG_BK][DS]
  uint8_t* mw = reinterpret_cast<uint8_t*>(KV + (size_t)4 * LONG_BK * DS) + warp * 16 * 8;   // keep-bytes [16][8]
  const int nq = Lq - q_base < LONG_BQ ? Lq - q_base : LONG_BQ;
  const int nkb = (Lk + LONG_BK - 1) / LONG_BK;
  auto load_block = [&](int kb, int buf) {
    const int k_base = kb * LONG_BK;
    const int nk = Lk - k_base < LONG_BK ? Lk - k_base : LONG_BK;
    bf16* Kb = KV + (size_t)buf * 2 * LONG_BK * DS;
    load_tile_async<D, DS>(Kb, a.kv + (size_t)(b * Lk + k_base) * a.ldkv + a.k_off + h * D, a.ldkv, nk, LONG_BK);
    load_tile_async<D, DS>(Kb + (size_t)LONG_BK * DS, a.kv + (size_t)(b * Lk + k_base) * a.ldkv + a.v_off + h * D,
                           a.ldkv, nk, LONG_BK);
  };
  load_tile_async<D, DS>(Qs, a.q + (size_t)(b * Lq + q_base) * a.ldq + a.q_off + h * D, a.ldq, nq, LONG_BQ);
  load_tile_async<D, DS>(dOs, dctx + (size_t)(b * Lq + q_base) * ldo + h * D, ldo, nq, LONG_BQ);
  load_block(0, 0);
  cp_async_commit();
  const int g = lane >> 2, q = lane & 3, i0 = warp * 16;
  const int r0 = q_base + i0 + g, r1 = r0 + 8;
  const bool ok0 = r0 < Lq, ok1 = r1 < Lq;
  const unsigned long long prow0 = (unsigned long long)(b * a.H + h) * Lq + r0, prow1 = prow0 + 8;
  const float ls0 = ok0 ? lse[prow0] : 0.f, ls1 = ok1 ? lse[prow1] : 0.f;
  const float* mrow = a.mask + (size_t)b * Lk;
  const int lkp = (Lk + 7) & ~7;
  float D0 = 0.f, D1 = 0.f;
  float o[D / 8][4];
#pragma unroll
  for (int c = 0; c < D / 8; ++c) o[c][0] = o[c][1] = o[c][2] = o[c][3] = 0.f;
  int it = 0;                                                       // block counter over both sweeps
  for (int sweep = 0; sweep < 2; ++sweep) {
    for (int kb = 0; kb < nkb; ++kb, ++it) {
      const int k_base = kb * LONG_BK;
      if (it + 1 < 2 * nkb) {
        load_block(kb + 1 < nkb ? kb + 1 : 0, (it + 1) & 1);        // sweep 1 starts again at block 0
        cp_async_commit();
        cp_async_wait<1>();
      } else {
        cp_async_wait<0>();
      }
      if (a.drop.thresh)
        warp_drop_bits<NKT, 8>(mw, a.drop, prow0 - g, Lq - (q_base + i0), lkp, k_base >> 3, (lkp - k_base) >> 3, lane);
      __syncthreads();
      const bf16* Ks = KV + (size_t)(it & 1) * 2 * LONG_BK * DS;
      const bf16* Vs = Ks + (size_t)LONG_BK * DS;
      uint32_t bw0[2], bw1[2];
      if (a.drop.thresh) {
        load_drop_row<8>(bw0, mw + g * 8);
        load_drop_row<8>(bw1, mw + (g + 8) * 8);
      }
      float s[NKT][4], dp[NKT][4];
#pragma unroll
      for (int t = 0; t < NKT; ++t) {
        s[t][0] = s[t][1] = s[t][2] = s[t][3] = 0.f;
        dp[t][0] = dp[t][1] = dp[t][2] = dp[t][3] = 0.f;
      }
      qk_mma<D, DS, NKT>(s, Qs, i0, Ks, lane);
      qk_mma<D, DS, NKT>(dp, dOs, i0, Vs, lane);
#pragma unroll
      for (int t = 0; t < NKT; ++t) {
        const int j = k_base + t * 8 + 2 * q;
        const bool v0 = j < Lk, v1 = j + 1 < Lk;
        const float ma = v0 ? mrow[j] : 0.f, mb = v1 ? mrow[j + 1] : 0.f;
        const float p00 = (v0 && ok0) ? __expf(fmaf(s[t][0], a.scale, ma) - ls0) : 0.f;
        const float p01 = (v1 && ok0) ? __expf(fmaf(s[t][1], a.scale, mb) - ls0) : 0.f;
        const float p10 = (v0 && ok1) ? __expf(fmaf(s[t][2], a.scale, ma) - ls1) : 0.f;
        const float p11 = (v1 && ok1) ? __expf(fmaf(s[t][3], a.scale, mb) - ls1) : 0.f;
        float m00 = 1.f, m01 = 1.f, m10 = 1.f, m11 = 1.f;
        if (a.drop.thresh) {
          drop_pair<8>(bw0, t, q, a.drop.scale, m00, m01);
          drop_pair<8>(bw1, t, q, a.drop.scale, m10, m11);
        }
        dp[t][0] *= m00; dp[t][1] *= m01; dp[t][2] *= m10; dp[t][3] *= m11;
        if (sweep == 0) {
          D0 += p00 * dp[t][0] + p01 * dp[t][1];
          D1 += p10 * dp[t][2] + p11 * dp[t][3];
        } else {
          s[t][0] = p00 * (dp[t][0] - D0); s[t][1] = p01 * (dp[t][1] - D0);
          s[t][2] = p10 * (dp[t][2] - D1); s[t][3] = p11 * (dp[t][3] - D1);
        }
      }
      if (sweep == 1) {
#pragma unroll
        for (int kt = 0; kt < NKT / 2; ++kt) {
          uint32_t af[4];
          af[0] = pack_bf16(s[2 * kt][0], s[2 * kt][1]);
          af[1] = pack_bf16(s[2 * kt][2], s[2 * kt][3]);
          af[2] = pack_bf16(s[2 * kt + 1][0], s[2 * kt + 1][1]);
          af[3] = pack_bf16(s[2 * kt + 1][2], s[2 * kt + 1][3]);
#pragma unroll
          for (int c = 0; c < D / 8; c += 2) {
            uint32_t kf[4];
            ldsm_x4_t(kf, Ks + (size_t)(kt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * DS + c * 8 + (lane >> 4) * 8);
            mma16816(o[c], af, kf[0], kf[1]);
            mma16816(o[c + 1], af, kf[2], kf[3]);
          }
        }
      }
      __syncthreads();                                              // every warp is done with this buffer
    }
    if (sweep == 0) {
      D0 += __shfl_xor_sync(0xffffffffu, D0, 1); D0 += __shfl_xor_sync(0xffffffffu, D0, 2);
      D1 += __shfl_xor_sync(0xffffffffu, D1, 1); D1 += __shfl_xor_sync(0xffffffffu, D1, 2);
      if (q == 0) {
        if (ok0) dsum[prow0] = D0;
        if (ok1) dsum[prow1] = D1;
      }
    }
  }
  bf16* q0 = dq + (size_t)(b * Lq + r0) * lddq + a.q_off + h * D + 2 * q;
  bf16* q1 = dq + (size_t)(b * Lq + r1) * lddq + a.q_off + h * D + 2 * q;
#pragma unroll
  for (int c = 0; c < D / 8; ++c) {
    if (ok0) *reinterpret_cast<uint32_t*>(q0 + c * 8) = pack_bf16(o[c][0] * a.scale, o[c][1] * a.scale);
    if (ok1) *reinterpret_cast<uint32_t*>(q1 + c * 8) = pack_bf16(o[c][2] * a.scale, o[c][3] * a.scale);
  }
}

// dK / dV for 64 keys of one (b, h): a warp owns 16 keys and walks the queries in blocks of 32, computing
// S^T (keys x queries) and dP^T on the tensor cores so the accumulator fragments are already the A
// operands of dV += P^T dO and dK += dS^T Q.
constexpr int LONG_BQ2 = 32;
template <int D>
__global__ void __launch_bounds__(128) attn_bwd_long_dkv(AttnArgs a, const bf16* __restrict__ dctx, int ldo,
                                                         const float* __restrict__ lse,
                                                         const float* __restrict__ dsum, bf16* __restrict__ dkv,
                                                         int lddkv) {
  constexpr int DS = D + 8, NQT = LONG_BQ2 / 8;
  extern __shared__ __align__(16) uint8_t smem[];
  const int h = blockIdx.x, b = blockIdx.y, k_base = blockIdx.z * LONG_BK;
  const int Lq = a.Lq, Lk = a.Lk;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  bf16* Ks = reinterpret_cast<bf16*>(smem);
  bf16* Vs = Ks + (size_t)LONG_BK * DS;
  bf16* QO = Vs + (size_t)LONG_BK * DS;                             // [2 buffers][Q | dO][LONG_BQ2][DS]
  float* stats = reinterpret_cast<float*>(QO + (size_t)4 * LONG_BQ2 * DS);   // [2 buffers][lse | D][LONG_BQ2]
  uint8_t* msk = reinterpret_cast<uint8_t*>(stats + 4 * LONG_BQ2);           // [2 buffers][LONG_BQ2][8] keep-bytes
  const int nk = Lk - k_base < LONG_BK ? Lk - k_base : LONG_BK;
  const int lkp = (Lk + 7) & ~7;
  const unsigned long long pbase = (unsigned long long)(b * a.H + h) * Lq;
  // keep-bytes of a query block x this CTA's 8 key tiles, one generator call per (query, tile); written one
  // block ahead, so the barriers of the pipeline publish them
  auto gen_mask = [&](int q_base, int buf) {
    for (int c = threadIdx.x; c < LONG_BQ2 * 8; c += 128) {
      const int i = q_base + (c >> 3), T = (k_base >> 3) + (c & 7);
      uint32_t bits = 0xFFu;
      if (i < Lq && T * 8 < Lk) bits = drop_tile_bits(a.drop, pbase + i, lkp, T);
      msk[buf * LONG_BQ2 * 8 + c] = (uint8_t)bits;
    }
  };
  auto load_block = [&](int q_base, int buf) {
    const int nq = Lq - q_base < LONG_BQ2 ? Lq - q_base : LONG_BQ2;
    bf16* Qb = QO + (size_t)buf * 2 * LONG_BQ2 * DS;
    load_tile_async<D, DS>(Qb, a.q + (size_t)(b * Lq + q_base) * a.ldq + a.q_off + h * D, a.ldq, nq, LONG_BQ2);
    load_tile_async<D, DS>(Qb + (size_t)LONG_BQ2 * DS, dctx + (size_t)(b * Lq + q_base) * ldo + h * D, ldo, nq,
                           LONG_BQ2);
    if (threadIdx.x < 2 * LONG_BQ2) {
      const int i = threadIdx.x & (LONG_BQ2 - 1);
      const float* src = threadIdx.x < LONG_BQ2 ? lse : dsum;
      const bool ok = i < nq;
      cp_async4(stats + buf * 2 * LONG_BQ2 + threadIdx.x, ok ? src + pbase + q_base + i : src, ok);
    }
  };
  load_tile_async<D, DS>(Ks, a.kv + (size_t)(b * Lk + k_base) * a.ldkv + a.k_off + h * D, a.ldkv, nk, LONG_BK);
  load_tile_async<D, DS>(Vs, a.kv + (size_t)(b * Lk + k_base) * a.ldkv + a.v_off + h * D, a.ldkv, nk, LONG_BK);
  load_block(0, 0);
  cp_async_commit();
  if (a.drop.thresh) gen_mask(0, 0);
  const int g = lane >> 2, q = lane & 3, j0 = warp * 16;
  const int jr0 = k_base + j0 + g, jr1 = jr0 + 8;                   // key rows of this thread
  const float* mrow = a.mask + (size_t)b * Lk;
  const float mk0 = jr0 < Lk ? mrow[jr0] : 0.f, mk1 = jr1 < Lk ? mrow[jr1] : 0.f;
  float dv[D / 8][4], dk[D / 8][4];
#pragma unroll
  for (int c = 0; c < D / 8; ++c) {
    dv[c][0] = dv[c][1] = dv[c][2] = dv[c][3] = 0.f;
    dk[c][0] = dk[c][1] = dk[c][2] = dk[c][3] = 0.f;
  }
  for (int q_base = 0, it = 0; q_base < Lq; q_base += LONG_BQ2, ++it) {
    if (q_base + LONG_BQ2 < Lq) {
      load_block(q_base + LONG_BQ2, (it + 1) & 1);                  // buffer released by the barrier below
      cp_async_commit();
      if (a.drop.thresh) gen_mask(q_base + LONG_BQ2, (it + 1) & 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const bf16* Qs = QO + (size_t)(it & 1) * 2 * LONG_BQ2 * DS;
    const bf16* dOs = Qs + (size_t)LONG_BQ2 * DS;
    const float* lse_s = stats + (it & 1) * 2 * LONG_BQ2;
    const float* d_s = lse_s + LONG_BQ2;
    const uint8_t* mk = msk + (it & 1) * LONG_BQ2 * 8 + warp * 2;   // this warp's two key tiles
    float s[NQT][4], dp[NQT][4];
#pragma unroll
    for (int t = 0; t < NQT; ++t) {
      s[t][0] = s[t][1] = s[t][2] = s[t][3] = 0.f;
      dp[t][0] = dp[t][1] = dp[t][2] = dp[t][3] = 0.f;
    }
    qk_mma<D, DS, NQT>(s, Ks, j0, Qs, lane);       // rows = keys, columns = queries
    qk_mma<D, DS, NQT>(dp, Vs, j0, dOs, lane);
#pragma unroll
    for (int t = 0; t < NQT; ++t) {
      uint32_t kb[2] = {0xFFFFu, 0xFFFFu};                          // keep-bytes of queries 2q, 2q + 1 of tile t
      if (a.drop.thresh) {
        kb[0] = *reinterpret_cast<const uint16_t*>(mk + (t * 8 + 2 * q) * 8);
        kb[1] = *reinterpret_cast<const uint16_t*>(mk + (t * 8 + 2 * q + 1) * 8);
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int il = t * 8 + 2 * q + (e & 1), i = q_base + il;
        const int j = (e & 2) ? jr1 : jr0;
        const bool ok = i < Lq && j < Lk;
        const float p = ok ? __expf(fmaf(s[t][e], a.scale, (e & 2) ? mk1 : mk0) - lse_s[il]) : 0.f;
        float m = 1.f;                                              // key jr0: bit g of byte 0, jr1: of byte 1
        if (a.drop.thresh) m = ((kb[e & 1] >> (g + ((e & 2) ? 8 : 0))) & 1u) ? a.drop.scale : 0.f;
        const float dpm = dp[t][e] * m;
        s[t][e] = p * m;                              // dropped probability (transposed)
        dp[t][e] = p * (dpm - d_s[il]);               // dS (transposed)
      }
    }
#pragma unroll
    for (int kt = 0; kt < NQT / 2; ++kt) {
      uint32_t pa[4], sa[4];
      pa[0] = pack_bf16(s[2 * kt][0], s[2 * kt][1]);
      pa[1] = pack_bf16(s[2 * kt][2], s[2 * kt][3]);
      pa[2] = pack_bf16(s[2 * kt + 1][0], s[2 * kt + 1][1]);
      pa[3] = pack_bf16(s[2 * kt + 1][2], s[2 * kt + 1][3]);
      sa[0] = pack_bf16(dp[2 * kt][0], dp[2 * kt][1]);
      sa[1] = pack_bf16(dp[2 * kt][2], dp[2 * kt][3]);
      sa[2] = pack_bf16(dp[2 * kt + 1][0], dp[2 * kt + 1][1]);
      sa[3] = pack_bf16(dp[2 * kt + 1][2], dp[2 * kt + 1][3]);
#pragma unroll
      for (int c = 0; c < D / 8; c += 2) {
        uint32_t of[4], qf[4];
        const size_t boff = (size_t)(kt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * DS + c * 8 + (lane >> 4) * 8;
        ldsm_x4_t(of, dOs + boff);
        ldsm_x4_t(qf, Qs + boff);
        mma16816(dv[c], pa, of[0], of[1]);
        mma16816(dv[c + 1], pa, of[2], of[3]);
        mma16816(dk[c], sa, qf[0], qf[1]);
        mma16816(dk[c + 1], sa, qf[2], qf[3]);
      }
    }
    __syncthreads();                                                // every warp is done with this buffer
  }
  bf16* k0 = dkv + (size_t)(b * Lk + jr0) * lddkv + h * D + 2 * q;
  bf16* k1 = dkv + (size_t)(b * Lk + jr1) * lddkv + h * D + 2 * q;
#pragma unroll
  for (int c = 0; c < D / 8; ++c) {
    if (jr0 < Lk) {
      *reinterpret_cast<uint32_t*>(k0 + a.k_off + c * 8) = pack_bf16(dk[c][0] * a.scale, dk[c][1] * a.scale);
      *reinterpret_cast<uint32_t*>(k0 + a.v_off + c * 8) = pack_bf16(dv[c][0], dv[c][1]);
    }
    if (jr1 < Lk) {
      *reinterpret_cast<uint32_t*>(k1 + a.k_off + c * 8) = pack_bf16(dk[c][2] * a.scale, dk[c][3] * a.scale);
      *reinterpret_cast<uint32_t*>(k1 + a.v_off + c * 8) = pack_bf16(dv[c][2], dv[c][3]);
    }
  }
}
