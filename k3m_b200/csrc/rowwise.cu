// HBM-bound row-wise kernels: LayerNorm forward / backward (+ dropout, + bias-gradient column
// sums), column sums, casts and small elementwise helpers.  One warp owns one row; every global
// access is a 16-byte vector; statistics and reductions are fp32.
#include "common.cuh"

namespace {

// ---------------------------------------------------------------------------------------------
// LayerNorm forward: y = (x - mean) * rstd * gamma + beta   [TF style: eps inside the sqrt,
// biased variance — vilbert_k3m.py:328-332], optional dropout on y (embeddings).
// NV = 8-element vectors per lane (H <= 256 * NV).
// Grouped form: up to 8 independent LayerNorms (the six streams of a connection stage, a text and an image layer)
// in ONE launch; NV is sized for the widest of them.  A CTA (four rows) belongs to one problem.
struct LnFwdProb {
  const act_t *x, *gamma, *beta;
  act_t* y;
  float *mean, *rstd;
  int M, H;
  DropSpec drop;
  int cta_begin;
};
struct LnFwdGroup {
  int n;
  float eps;
  LnFwdProb p[8];
};
// Warps walk the concatenated row list of all problems (row = warp, warp + #warps, ...) and load the NEXT row
// (kept packed, NV x 16 bytes per lane) before they reduce the current one: CTAs that live for a single 1.5 KB row
// spent their time being scheduled (1.9 TB/s over the 113 launches of a step), two rows in flight per warp hide
// the HBM latency instead.
template <int NV>
__global__ void __launch_bounds__(256) ln_fwd_group_kernel(const __grid_constant__ LnFwdGroup g) {
  const int lane = threadIdx.x & 31;
  const int nwarps = gridDim.x * 8;
  int total = 0;
  for (int i = 0; i < g.n; ++i) total += g.p[i].M;
  int r = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= total) return;
  auto locate = [&](int row, int& pi, int& local) {
    pi = 0;
    local = row;
    while (pi + 1 < g.n && local >= g.p[pi].M) {
      local -= g.p[pi].M;
      ++pi;
    }
  };
  int pi, lr;
  locate(r, pi, lr);
  Pack8 cur[NV], nxt[NV];
  {
    const LnFwdProb& P = g.p[pi];
    const act_t* xr = P.x + (size_t)lr * P.H;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (lane + 32 * i) * 8;
      if (c < P.H) cur[i] = load_pack8(xr + c);
    }
  }
  while (true) {
    const int rn = r + nwarps;
    int pn = 0, ln_ = 0;
    const bool more = rn < total;
    if (more) {
      locate(rn, pn, ln_);
      const LnFwdProb& Q = g.p[pn];
      const act_t* xr = Q.x + (size_t)ln_ * Q.H;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = (lane + 32 * i) * 8;
        if (c < Q.H) nxt[i] = load_pack8(xr + c);
      }
    }
    {
      const LnFwdProb& P = g.p[pi];
      const int H = P.H;
      float v[NV][8];
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = (lane + 32 * i) * 8;
        if (c < H) {
          unpack8(cur[i], v[i]);
#pragma unroll
          for (int j = 0; j < 8; ++j) s += v[i][j];
        }
      }
      const float mean = warp_sum(s) / (float)H;
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = (lane + 32 * i) * 8;
        if (c < H) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float d = v[i][j] - mean;
            q += d * d;
          }
        }
      }
      const float rstd = rsqrtf(warp_sum(q) / (float)H + g.eps);
      if (lane == 0) {
        P.mean[lr] = mean;
        P.rstd[lr] = rstd;
      }
      act_t* yr = P.y + (size_t)lr * H;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = (lane + 32 * i) * 8;
        if (c < H) {
          float gm[8], b[8], o[8];
          ld8(P.gamma + c, gm);
          ld8(P.beta + c, b);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * gm[j] + b[j];
          if (P.drop.thresh) {
            uint32_t w[4];
            rng_words(P.drop.rng, P.drop.site, ((unsigned long long)lr * H + c) >> 3, w);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] *= drop_mult(P.drop, w, j);
          }
          st8(yr + c, o);
        }
      }
    }
    if (!more) break;
    r = rn; pi = pn; lr = ln_;
#pragma unroll
    for (int i = 0; i < NV; ++i) cur[i] = nxt[i];
  }
}

// ---------------------------------------------------------------------------------------------
// LayerNorm backward.  dy is optionally masked first by the post-LN dropout (drop_out).
// dx -> act_t; if drop_in is set a second output dxm = dx * mask_in * scale is written (gradient of
// the GEMM output that was dropped out before the residual add).  Column sums: dgamma += dy*xhat,
// dbeta += dy, dbias += dxm  (fp32 atomics, one set per CTA after a grid-stride loop over rows).
// cluster helpers: the column sums of the CTAs of a cluster are combined through distributed shared memory
__device__ __forceinline__ uint32_t cluster_rank_x() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float ld_dsmem(const float* local, uint32_t rank) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(local), ra;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(a), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
  return v;
}

// CL = CTAs per cluster (1: no cluster).  With CL > 1 the per-CTA column sums meet in distributed shared
// memory and each CTA of the cluster issues the atomics of 1/CL of the columns: CL x fewer same-address
// atomics.  An experiment that did not pay (see k3m_ln_bwd below); selectable with K3M_LN_CLUSTER=1.
// SACC: the three column-sum sets live in this warp's slice of shared memory (read-modify-written per row, 16-byte
// accesses, no bank conflicts: lane owns its columns) instead of 3 x NV x 8 registers per thread — 128-210 registers
// allowed one or two CTAs per SM; ~70 allow three, i.e. 24 rows in flight per SM for this latency-bound kernel.
__device__ __forceinline__ void sacc8(float* p, const float v[8]) {
  float4 a = *reinterpret_cast<float4*>(p), b = *reinterpret_cast<float4*>(p + 4);
  a.x += v[0]; a.y += v[1]; a.z += v[2]; a.w += v[3]; b.x += v[4]; b.y += v[5]; b.z += v[6]; b.w += v[7];
  *reinterpret_cast<float4*>(p) = a;
  *reinterpret_cast<float4*>(p + 4) = b;
}
template <int NV, int CL, bool SACC = false>
__device__ __forceinline__ void ln_bwd_body(const act_t* __restrict__ dy, const act_t* __restrict__ x,
                                            const act_t* __restrict__ gamma, const float* __restrict__ mean_in,
                                            const float* __restrict__ rstd_in, act_t* __restrict__ dx,
                                            act_t* __restrict__ dxm, float* __restrict__ dgamma,
                                            float* __restrict__ dbeta, float* __restrict__ dbias, int M, int H,
                                            const DropSpec& drop_out, const DropSpec& drop_in, int cta, int nctas) {
  // Register budget: two CTAs (16 warps) per SM keep enough rows in flight to cover HBM latency; the
  // row's dy / x stay packed as loaded (act_t) between the two passes and gamma is re-read (L1) per row,
  // so the live state is the 3 x NV x 8 fp32 column-sum accumulators.
  extern __shared__ float sred[];  // [8 warps][3][H]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NA = SACC ? 1 : NV;
  float ag[NA][8], ab[NA][8], ac[NA][8];
#pragma unroll
  for (int i = 0; i < NA; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) ag[i][j] = ab[i][j] = ac[i][j] = 0.f;
  float* s0 = sred + (size_t)warp * 3 * H;
  if (SACC) {
    for (int c = lane * 4; c < 3 * H; c += 128) *reinterpret_cast<float4*>(s0 + c) = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
  }
  for (int row = cta * 8 + warp; row < M; row += nctas * 8) {
    const float mean = mean_in[row], rstd = rstd_in[row];
    Pack8 pd[NV], px[NV];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (lane + 32 * i) * 8;
      if (c < H) {
        pd[i] = load_pack8(dy + (size_t)row * H + c);
        px[i] = load_pack8(x + (size_t)row * H + c);
      }
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (lane + 32 * i) * 8;
      if (c < H) {
        float d[8], xv[8], gm[8];
        unpack8(pd[i], d);
        unpack8(px[i], xv);
        ld8(gamma + c, gm);
        if (drop_out.thresh) {
          uint32_t w[4];
          rng_words(drop_out.rng, drop_out.site, ((unsigned long long)row * H + c) >> 3, w);
#pragma unroll
          for (int j = 0; j < 8; ++j) d[j] *= drop_mult(drop_out, w, j);
        }
        float dxh[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = (xv[j] - mean) * rstd;
          const float g = d[j] * gm[j];
          dxh[j] = d[j] * xh;
          s1 += g;
          s2 += g * xh;
        }
        if (SACC) {
          sacc8(s0 + c, dxh);
          sacc8(s0 + H + c, d);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            ag[i][j] += dxh[j];
            ab[i][j] += d[j];
          }
        }
      }
    }
    s1 = warp_sum(s1) / (float)H;
    s2 = warp_sum(s2) / (float)H;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (lane + 32 * i) * 8;
      if (c < H) {
        float d[8], xv[8], gm[8], o[8];
        unpack8(pd[i], d);
        unpack8(px[i], xv);
        ld8(gamma + c, gm);
        if (drop_out.thresh) {
          uint32_t w[4];
          rng_words(drop_out.rng, drop_out.site, ((unsigned long long)row * H + c) >> 3, w);
#pragma unroll
          for (int j = 0; j < 8; ++j) d[j] *= drop_mult(drop_out, w, j);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (d[j] * gm[j] - s1 - (xv[j] - mean) * rstd * s2) * rstd;
        st8(dx + (size_t)row * H + c, o);
        if (drop_in.thresh) {
          uint32_t w[4];
          rng_words(drop_in.rng, drop_in.site, ((unsigned long long)row * H + c) >> 3, w);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] *= drop_mult(drop_in, w, j);
          st8(dxm + (size_t)row * H + c, o);
        }
        if (dbias) {
          if (SACC) {
            sacc8(s0 + 2 * H + c, o);
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) ac[i][j] += o[j];
          }
        }
      }
    }
  }
  // cross-warp reduction of the three column-sum sets, then one atomic per column per CTA
  if (!SACC) {
#pragma unroll
    for (int i = 0; i < NA; ++i) {
      const int c = (lane + 32 * i) * 8;
      if (c < H) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          s0[c + j] = ag[i][j];
          s0[H + c + j] = ab[i][j];
          s0[2 * H + c + j] = ac[i][j];
        }
      }
    }
  }
  __syncthreads();
  const int nsets = dbias ? 3 : 2;
  for (int idx = threadIdx.x; idx < nsets * H; idx += blockDim.x) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sred[(size_t)w * 3 * H + idx];
    if (CL > 1) {
      sred[idx] = t;                 // slot of warp 0: this thread is the only reader and writer of element idx
    } else {
      float* dst = idx < H ? dgamma + idx : (idx < 2 * H ? dbeta + (idx - H) : dbias + (idx - 2 * H));
      atomicAdd(dst, t);
    }
  }
  if (CL > 1) {
    cluster_barrier();               // every CTA's totals are in its sred[0 .. nsets*H)
    const uint32_t rank = cluster_rank_x();
    const int n = nsets * H, chunk = (n + CL - 1) / CL;
    const int hi = (int)(rank + 1) * chunk < n ? (int)(rank + 1) * chunk : n;
    for (int idx = (int)rank * chunk + threadIdx.x; idx < hi; idx += blockDim.x) {
      float t = 0.f;
#pragma unroll
      for (int r = 0; r < CL; ++r) t += ld_dsmem(sred + idx, (uint32_t)r);
      float* dst = idx < H ? dgamma + idx : (idx < 2 * H ? dbeta + (idx - H) : dbias + (idx - 2 * H));
      atomicAdd(dst, t);
    }
    cluster_barrier();               // nobody leaves while its shared memory is still being read
  }
}

template <int NV, int CL>
__global__ void __launch_bounds__(256, NV >= 4 ? 1 : 2) ln_bwd_kernel(const act_t* __restrict__ dy, const act_t* __restrict__ x,
                                                        const act_t* __restrict__ gamma, const float* __restrict__ mean_in,
                                                        const float* __restrict__ rstd_in, act_t* __restrict__ dx,
                                                        act_t* __restrict__ dxm, float* __restrict__ dgamma,
                                                        float* __restrict__ dbeta, float* __restrict__ dbias, int M, int H,
                                                        DropSpec drop_out, DropSpec drop_in) {
  ln_bwd_body<NV, CL>(dy, x, gamma, mean_in, rstd_in, dx, dxm, dgamma, dbeta, dbias, M, H, drop_out, drop_in,
                      (int)blockIdx.x, (int)gridDim.x);
}

// Grouped form (see LnFwdGroup): every problem owns a contiguous range of CTAs, sized by its share of the rows.
struct LnBwdProb {
  const act_t *dy, *x, *gamma;
  const float *mean, *rstd;
  act_t *dx, *dxm;
  float *dgamma, *dbeta, *dbias;
  int M, H;
  DropSpec drop_out, drop_in;
  int cta_begin, nctas;
};
struct LnBwdGroup {
  int n;
  LnBwdProb p[8];
};
template <int NV>
__global__ void __launch_bounds__(256, NV >= 4 ? 1 : 2) ln_bwd_group_kernel(const __grid_constant__ LnBwdGroup g) {
  int pi = 0;
  while (pi + 1 < g.n && (int)blockIdx.x >= g.p[pi + 1].cta_begin) ++pi;
  const LnBwdProb& P = g.p[pi];
  ln_bwd_body<NV, 1>(P.dy, P.x, P.gamma, P.mean, P.rstd, P.dx, P.dxm, P.dgamma, P.dbeta, P.dbias, P.M, P.H, P.drop_out,
                     P.drop_in, (int)blockIdx.x - P.cta_begin, P.nctas);
}
template <int NV>
__global__ void __launch_bounds__(256, NV >= 4 ? 2 : 3) ln_bwd_group_sacc_kernel(const __grid_constant__ LnBwdGroup g) {
  int pi = 0;
  while (pi + 1 < g.n && (int)blockIdx.x >= g.p[pi + 1].cta_begin) ++pi;
  const LnBwdProb& P = g.p[pi];
  ln_bwd_body<NV, 1, true>(P.dy, P.x, P.gamma, P.mean, P.rstd, P.dx, P.dxm, P.dgamma, P.dbeta, P.dbias, P.M, P.H,
                           P.drop_out, P.drop_in, (int)blockIdx.x - P.cta_begin, P.nctas);
}

// Tile form of the backward (H > 512): a CTA walks tiles of eight rows that arrive in shared memory through cp.async,
// double-buffered (the next tile is in flight while the current one is computed: the row-per-warp form above stalls a
// quarter of its time on the global loads at the top of every row, profiles/r03k_rowwise_stalls.txt).
//   phase 1: warp w reduces row w of the tile (s1 = sum g, s2 = sum g * xhat, g = dy * gamma) — same lane / column
//            mapping and operation order as above, so s1 / s2 and dx are bit-equal to that form
//   phase 2: thread (v, half) owns the 8 columns of vector v for rows half, half + 2, ..: dx / dxm of its pieces, and the
//            three column sums of its columns in REGISTERS for the whole life of the CTA (gamma too) — no shared-memory
//            read-modify-write per row (another quarter of the stalls), no gamma reload
// Column sums leave as one atomic per column and CTA, as before.
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
template <int NV>
__device__ __forceinline__ void ln_bwd_tile_body(const act_t* __restrict__ dy, const act_t* __restrict__ x,
                                                 const act_t* __restrict__ gamma, const float* __restrict__ mean_in,
                                                 const float* __restrict__ rstd_in, act_t* __restrict__ dx,
                                                 act_t* __restrict__ dxm, float* __restrict__ dgamma,
                                                 float* __restrict__ dbeta, float* __restrict__ dbias, int M, int H,
                                                 const DropSpec& drop_out, const DropSpec& drop_in, int cta, int nctas) {
  constexpr int R = 8;
  extern __shared__ __align__(16) float sred[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(sred);
  const size_t row_bytes = (size_t)H * sizeof(act_t), tile_bytes = 2 * R * row_bytes;      // dy rows, then x rows
  act_t* sgamma = reinterpret_cast<act_t*>(smem + 2 * tile_bytes);
  float* stats = reinterpret_cast<float*>(smem + 2 * tile_bytes + row_bytes);                // [R][4]: mean, rstd, s1, s2
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int nvec = H >> 3, v = t & 127, half = t >> 7, c0 = v * 8;
  const bool owner = v < nvec;
  const int ntiles = (M + R - 1) / R;
  const int cpr = (int)(row_bytes / 16);             // 16-byte chunks per row
  auto issue = [&](int tile, int buf) {
    if (tile < ntiles) {
      uint8_t* dst = smem + (size_t)buf * tile_bytes;
      for (int idx = t; idx < 2 * R * cpr; idx += 256) {
        const int which = idx / (R * cpr), rem = idx % (R * cpr), r = rem / cpr, ch = rem % cpr;
        const int row = tile * R + r;
        if (row < M) {
          const uint8_t* src = reinterpret_cast<const uint8_t*>(which ? x : dy) + (size_t)row * row_bytes + (size_t)ch * 16;
          cp_async16(dst + ((size_t)which * R + r) * row_bytes + (size_t)ch * 16, src);
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  issue(cta, 0);
  for (int i = t; i < cpr; i += 256)
    reinterpret_cast<uint4*>(sgamma)[i] = reinterpret_cast<const uint4*>(gamma)[i];
  float gm[8], ag[8], ab[8], ac[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) gm[j] = ag[j] = ab[j] = ac[j] = 0.f;
  if (owner) ld8(gamma + c0, gm);
  int buf = 0;
  for (int tile = cta; tile < ntiles; tile += nctas, buf ^= 1) {
    issue(tile + nctas, buf ^ 1);                 // that buffer was released by the barrier that ended the last iteration
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncthreads();
    const act_t* sdy = reinterpret_cast<const act_t*>(smem + (size_t)buf * tile_bytes);
    const act_t* sx = reinterpret_cast<const act_t*>(smem + (size_t)buf * tile_bytes + R * row_bytes);
    {                                             // ---- phase 1: row statistics, one row per warp
      const int row = tile * R + warp;
      if (row < M) {
        const float mean = mean_in[row], rstd = rstd_in[row];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const int c = (lane + 32 * i) * 8;
          if (c < H) {
            float d[8], xv[8], g8[8];
            ld8(sdy + (size_t)warp * H + c, d);
            ld8(sx + (size_t)warp * H + c, xv);
            ld8(sgamma + c, g8);
            if (drop_out.thresh) {
              uint32_t w[4];
              rng_words(drop_out.rng, drop_out.site, ((unsigned long long)row * H + c) >> 3, w);
#pragma unroll
              for (int j = 0; j < 8; ++j) d[j] *= drop_mult(drop_out, w, j);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float xh = (xv[j] - mean) * rstd;
              const float g = d[j] * g8[j];
              s1 += g;
              s2 += g * xh;
            }
          }
        }
        s1 = warp_sum(s1) / (float)H;
        s2 = warp_sum(s2) / (float)H;
        if (lane == 0) {
          stats[warp * 4 + 0] = mean; stats[warp * 4 + 1] = rstd; stats[warp * 4 + 2] = s1; stats[warp * 4 + 3] = s2;
        }
      }
    }
    __syncthreads();
    if (owner) {                                  // ---- phase 2: my 8 columns of rows half, half + 2, ...
#pragma unroll 1
      for (int r = half; r < R; r += 2) {
        const int row = tile * R + r;
        if (row >= M) break;
        const float mean = stats[r * 4 + 0], rstd = stats[r * 4 + 1], s1 = stats[r * 4 + 2], s2 = stats[r * 4 + 3];
        float d[8], xv[8], o[8];
        ld8(sdy + (size_t)r * H + c0, d);
        ld8(sx + (size_t)r * H + c0, xv);
        if (drop_out.thresh) {
          uint32_t w[4];
          rng_words(drop_out.rng, drop_out.site, ((unsigned long long)row * H + c0) >> 3, w);
#pragma unroll
          for (int j = 0; j < 8; ++j) d[j] *= drop_mult(drop_out, w, j);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = (xv[j] - mean) * rstd;
          ag[j] += d[j] * xh;
          ab[j] += d[j];
          o[j] = (d[j] * gm[j] - s1 - xh * s2) * rstd;
        }
        st8(dx + (size_t)row * H + c0, o);
        if (drop_in.thresh) {
          uint32_t w[4];
          rng_words(drop_in.rng, drop_in.site, ((unsigned long long)row * H + c0) >> 3, w);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] *= drop_mult(drop_in, w, j);
          st8(dxm + (size_t)row * H + c0, o);
        }
        if (dbias) {
#pragma unroll
          for (int j = 0; j < 8; ++j) ac[j] += o[j];
        }
      }
    }
    __syncthreads();                              // stats and this tile's buffer are free again
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  // the two halves' column sums meet in shared memory (tile buffers are free now): [2][3][H] floats
  if (owner) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sred[(size_t)(half * 3 + 0) * H + c0 + j] = ag[j];
      sred[(size_t)(half * 3 + 1) * H + c0 + j] = ab[j];
      sred[(size_t)(half * 3 + 2) * H + c0 + j] = ac[j];
    }
  }
  __syncthreads();
  const int nsets = dbias ? 3 : 2;
  for (int idx = t; idx < nsets * H; idx += 256) {
    const float tot = sred[idx] + sred[(size_t)3 * H + idx];
    float* dst = idx < H ? dgamma + idx : (idx < 2 * H ? dbeta + (idx - H) : dbias + (idx - 2 * H));
    atomicAdd(dst, tot);
  }
}
template <int NV>
__global__ void __launch_bounds__(256, 2) ln_bwd_group_tile_kernel(const __grid_constant__ LnBwdGroup g) {
  int pi = 0;
  while (pi + 1 < g.n && (int)blockIdx.x >= g.p[pi + 1].cta_begin) ++pi;
  const LnBwdProb& P = g.p[pi];
  ln_bwd_tile_body<NV>(P.dy, P.x, P.gamma, P.mean, P.rstd, P.dx, P.dxm, P.dgamma, P.dbeta, P.dbias, P.M, P.H, P.drop_out,
                       P.drop_in, (int)blockIdx.x - P.cta_begin, P.nctas);
}

// ---------------------------------------------------------------------------------------------
// out[n] += sum_m x[m, n]   (x act_t or fp32, row stride ld)
// generic: one column per lane (any ld / alignment)
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ x, int ld, float* __restrict__ out, int M, int N,
                                                     int rows_per_cta) {
  __shared__ float sred[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + lane;
  const int r0 = blockIdx.y * rows_per_cta;
  int r1 = r0 + rows_per_cta;
  if (r1 > M) r1 = M;
  float acc = 0.f;
  if (col < N)
    for (int r = r0 + warp; r < r1; r += 8) acc += ld_f(x + (size_t)r * ld + col);
  sred[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && col < N) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sred[w][lane];
    atomicAdd(out + col, t);
  }
}

// act_t, 16-byte aligned rows: a warp reads 512 contiguous bytes of a row per instruction (8 columns per
// lane), four rows in flight per warp; 256 columns per CTA, rows split over blockIdx.y
__device__ __forceinline__ void colsum8_body(const act_t* __restrict__ x, int ld, float* __restrict__ out, int M, int N,
                                             int rows_per_cta, int bx, int by) {
  __shared__ float sred[8][256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = bx * 256 + lane * 8;
  const int r0 = by * rows_per_cta;
  int r1 = r0 + rows_per_cta;
  if (r1 > M) r1 = M;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (col < N) {
    int r = r0 + warp;
    for (; r + 24 < r1; r += 32) {
      float v[4][8];
#pragma unroll
      for (int u = 0; u < 4; ++u) ld8(x + (size_t)(r + 8 * u) * ld + col, v[u]);
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += v[u][j];
    }
    for (; r < r1; r += 8) {
      float v[8];
      ld8(x + (size_t)r * ld + col, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) sred[warp][lane * 8 + j] = acc[j];
  __syncthreads();
  const int c = bx * 256 + threadIdx.x;
  if (c < N) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sred[w][threadIdx.x];
    atomicAdd(out + c, t);
  }
}
__global__ void __launch_bounds__(256) colsum8_kernel(const act_t* __restrict__ x, int ld, float* __restrict__ out, int M,
                                                      int N, int rows_per_cta) {
  colsum8_body(x, ld, out, M, N, rows_per_cta, (int)blockIdx.x, (int)blockIdx.y);
}
// grouped form: the bias gradients of up to 8 independent problems (the streams of a lock-step stage) in one launch
struct ColsumProb {
  const act_t* x;
  float* out;
  int ld, M, N, rows_per, gx, cta_begin;
};
struct ColsumGroup {
  int n;
  ColsumProb p[8];
};
__global__ void __launch_bounds__(256) colsum8_group_kernel(const __grid_constant__ ColsumGroup g) {
  int pi = 0;
  while (pi + 1 < g.n && (int)blockIdx.x >= g.p[pi + 1].cta_begin) ++pi;
  const ColsumProb& P = g.p[pi];
  const int l = (int)blockIdx.x - P.cta_begin;
  colsum8_body(P.x, P.ld, P.out, P.M, P.N, P.rows_per, l % P.gx, l / P.gx);
}

// ---------------------------------------------------------------------------------------------
template <typename TI, typename TO>
__global__ void cast_kernel(const TI* __restrict__ x, TO* __restrict__ y, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    st_f(y + i, ld_f(x + i));
}

// out = wa*a + wb*b + wc*c   (c optional)
template <typename TI, typename TO>
__global__ void lincomb3_kernel(const TI* a, const TI* b, const TI* c, float wa, float wb, float wc, TO* out, int rows,
                                int cols, int lda, int ldb, int ldc, int ldo) {
  const size_t n = (size_t)rows * cols;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols), k = (int)(i % cols);
    float v = wa * ld_f(a + (size_t)r * lda + k) + wb * ld_f(b + (size_t)r * ldb + k);
    if (c) v += wc * ld_f(c + (size_t)r * ldc + k);
    st_f(out + (size_t)r * ldo + k, v);
  }
}

// y += alpha * x
template <typename TX, typename TY>
__global__ void axpy_kernel(const TX* x, TY* y, float alpha, int rows, int cols, int ldx, int ldy) {
  const size_t n = (size_t)rows * cols;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols), k = (int)(i % cols);
    TY* p = y + (size_t)r * ldy + k;
    st_f(p, ld_f(p) + alpha * ld_f(x + (size_t)r * ldx + k));
  }
}

// 8 elements per thread through 16 / 32-byte accesses (cols, row strides multiples of 8, aligned bases)
template <typename TX, typename TY>
__global__ void axpy8_kernel(const TX* __restrict__ x, TY* __restrict__ y, float alpha, int rows, int cols8, int ldx,
                             int ldy) {
  const int n = rows * cols8;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int r = i / cols8, k = (i - r * cols8) * 8;
    float a[8], b[8];
    ld8(x + (size_t)r * ldx + k, a);
    ld8(y + (size_t)r * ldy + k, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) b[j] = fmaf(alpha, a[j], b[j]);
    st8(y + (size_t)r * ldy + k, b);
  }
}

__global__ void scale_kernel(float* x, float alpha, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) x[i] *= alpha;
}

// x[r, :] *= *alpha (device scalar): upstream loss gradients reach the Python class API as device tensors
template <typename T>
__global__ void scale_dev_kernel(T* x, const float* __restrict__ alpha, int rows, int cols, int ld) {
  const float a = *alpha;
  const size_t n = (size_t)rows * cols;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    T* p = x + (i / cols) * (size_t)ld + (i % cols);
    st_f(p, ld_f(p) * a);
  }
}

// dx = dy * gelu'(pre)
__global__ void dgelu_kernel(const act_t* __restrict__ dy, const act_t* __restrict__ pre, act_t* __restrict__ dx, size_t n8) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    float a[8], b[8];
    ld8(dy + i * 8, a);
    ld8(pre + i * 8, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] *= dgelu_f(b[j]);
    st8(dx + i * 8, a);
  }
}

// (1 - mask) * -10000   (vilbert_k3m.py:2561)
__global__ void mask_additive_kernel(const long long* __restrict__ mask, float* __restrict__ out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (1.0f - (float)mask[i]) * -10000.0f;
}

inline int grid_for(size_t n, int block) {
  size_t g = (n + block - 1) / block;
  if (g > 148 * 16) g = 148 * 16;
  if (g < 1) g = 1;
  return (int)g;
}

inline DropSpec make_drop(const void* rng, unsigned int site, float p) {
  DropSpec d;
  d.rng = reinterpret_cast<const unsigned long long*>(rng);
  d.site = site;
  d.thresh = (rng && p > 0.f) ? (unsigned int)(p * 65536.0f + 0.5f) : 0u;
  d.scale = 1.0f / (1.0f - p);
  return d;
}

}  // namespace

extern "C" int k3m_ln_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd, int M,
                          int H, float eps, const void* rng, unsigned int site, float drop_p, void* stream) {
  K3M_REQUIRE(M > 0 && H > 0 && H % 8 == 0 && H <= 4096, "ln_fwd: need H %% 8 == 0 and H <= 4096 (H=%d)", H);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  DropSpec d = make_drop(rng, site, drop_p);
  // one problem of the grouped (persistent, prefetching) kernel
  LnFwdGroup g;
  g.n = 1;
  g.eps = eps;
  LnFwdProb& P = g.p[0];
  P.x = (const act_t*)x; P.gamma = (const act_t*)gamma; P.beta = (const act_t*)beta; P.y = (act_t*)y;
  P.mean = mean; P.rstd = rstd; P.M = M; P.H = H; P.drop = d; P.cta_begin = 0;
  for (int i = 1; i < 8; ++i) g.p[i] = g.p[0];
  int ctas = (M + 7) / 8;
  if (ctas > 148 * 4) ctas = 148 * 4;
  const int nv = (H + 255) / 256;
  if (nv <= 1) ln_fwd_group_kernel<1><<<ctas, 256, 0, st>>>(g);
  else if (nv <= 2) ln_fwd_group_kernel<2><<<ctas, 256, 0, st>>>(g);
  else if (nv <= 3) ln_fwd_group_kernel<3><<<ctas, 256, 0, st>>>(g);
  else if (nv <= 4) ln_fwd_group_kernel<4><<<ctas, 256, 0, st>>>(g);
  else if (nv <= 8) ln_fwd_group_kernel<8><<<ctas, 256, 0, st>>>(g);
  else ln_fwd_group_kernel<16><<<ctas, 256, 0, st>>>(g);
  return k3m_check_launch("k3m_ln_fwd");
}

extern "C" int k3m_ln_bwd_grouped(const void* descs, int n, void* stream);
struct LnBwdDesc1 {      // == K3MLnBwdDesc (declared again further down for the grouped entry)
  const void *dy, *x, *gamma;
  const float *mean, *rstd;
  void *dx, *dxm;
  float *dgamma, *dbeta, *dbias;
  int M, H;
  const void* rng;
  unsigned int site_out;
  float p_out;
  unsigned int site_in;
  float p_in;
};

extern "C" int k3m_ln_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                          void* dx, void* dxm, float* dgamma, float* dbeta, float* dbias, int M, int H, const void* rng,
                          unsigned int site_out, float p_out, unsigned int site_in, float p_in, void* stream) {
  K3M_REQUIRE(M > 0 && H > 0 && H % 8 == 0 && H <= 1024, "ln_bwd: need H %% 8 == 0 and H <= 1024 (H=%d)", H);
  static const int use_cluster0 = getenv("K3M_LN_CLUSTER") ? atoi(getenv("K3M_LN_CLUSTER")) : 0;
  if (!use_cluster0) {       // one problem of the grouped kernel (shared-memory column accumulators)
    LnBwdDesc1 d = {dy, x, gamma, mean, rstd, dx, dxm, dgamma, dbeta, dbias, M, H, rng, site_out, p_out, site_in, p_in};
    return k3m_ln_bwd_grouped(&d, 1, stream);
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  DropSpec dout = make_drop(rng, site_out, p_out), din = make_drop(rng, site_in, p_in);
  K3M_REQUIRE(!din.thresh || dxm, "ln_bwd: drop_in needs the dxm output");
  int grid = (M + 7) / 8;
  if (grid > 148 * 2) grid = 148 * 2;            // two resident CTAs per SM
  // measured: with clusters of 8 the 113 launches of a step take 4.58 ms instead of 3.20 ms (cluster
  // co-scheduling and two cluster barriers cost more than the 8x fewer atomics save) -> off by default
  static const int use_cluster = getenv("K3M_LN_CLUSTER") ? atoi(getenv("K3M_LN_CLUSTER")) : 0;
  constexpr int CL = 8;
  const bool cl = use_cluster && grid >= 2 * CL;
  if (cl) grid = (grid + CL - 1) / CL * CL;      // whole clusters (CTAs without rows add zeros)
  const size_t smem = (size_t)8 * 3 * H * sizeof(float);
  const int nv = (H + 255) / 256;
#define LN_BWD(NV)                                                                                                   \
  do {                                                                                                               \
    static bool cfg = false;                                                                                         \
    if (!cfg) {                                                                                                      \
      cudaFuncSetAttribute(ln_bwd_kernel<NV, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 3 * 1024 * 4);     \
      cudaFuncSetAttribute(ln_bwd_kernel<NV, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 3 * 1024 * 4);    \
      cfg = true;                                                                                                    \
    }                                                                                                                \
    if (cl) {                                                                                                        \
      cudaLaunchConfig_t lc = {};                                                                                    \
      lc.gridDim = dim3(grid);                                                                                       \
      lc.blockDim = dim3(256);                                                                                       \
      lc.dynamicSmemBytes = smem;                                                                                    \
      lc.stream = st;                                                                                                \
      cudaLaunchAttribute at[1];                                                                                     \
      at[0].id = cudaLaunchAttributeClusterDimension;                                                                \
      at[0].val.clusterDim.x = CL;                                                                                   \
      at[0].val.clusterDim.y = 1;                                                                                    \
      at[0].val.clusterDim.z = 1;                                                                                    \
      lc.attrs = at;                                                                                                 \
      lc.numAttrs = 1;                                                                                               \
      cudaLaunchKernelEx(&lc, ln_bwd_kernel<NV, CL>, (const act_t*)dy, (const act_t*)x, (const act_t*)gamma, mean, rstd, \
                         (act_t*)dx, (act_t*)dxm, dgamma, dbeta, dbias, M, H, dout, din);                              \
    } else {                                                                                                         \
      ln_bwd_kernel<NV, 1><<<grid, 256, smem, st>>>((const act_t*)dy, (const act_t*)x, (const act_t*)gamma, mean, rstd,  \
                                                    (act_t*)dx, (act_t*)dxm, dgamma, dbeta, dbias, M, H, dout, din);   \
    }                                                                                                                \
  } while (0)
  if (nv <= 1) LN_BWD(1);
  else if (nv <= 2) LN_BWD(2);
  else if (nv <= 3) LN_BWD(3);
  else LN_BWD(4);
#undef LN_BWD
  return k3m_check_launch("k3m_ln_bwd");
}

struct LnFwdDesc {      // == K3MLnFwdDesc of include/k3m.h
  const void *x, *gamma, *beta;
  void* y;
  float *mean, *rstd;
  int M, H;
  const void* rng;
  unsigned int site;
  float drop_p;
};
struct LnBwdDesc {      // == K3MLnBwdDesc
  const void *dy, *x, *gamma;
  const float *mean, *rstd;
  void *dx, *dxm;
  float *dgamma, *dbeta, *dbias;
  int M, H;
  const void* rng;
  unsigned int site_out;
  float p_out;
  unsigned int site_in;
  float p_in;
};

extern "C" int k3m_ln_fwd_grouped(const void* descs, int n, float eps, void* stream) {
  const LnFwdDesc* d = reinterpret_cast<const LnFwdDesc*>(descs);
  K3M_REQUIRE(n >= 1 && n <= 8, "ln_fwd_grouped: 1..8 problems per launch (got %d)", n);
  LnFwdGroup g;
  g.n = n;
  g.eps = eps;
  int ctas = 0, hmax = 0;
  for (int i = 0; i < n; ++i) {
    K3M_REQUIRE(d[i].M > 0 && d[i].H > 0 && d[i].H % 8 == 0 && d[i].H <= 4096, "ln_fwd_grouped: need H %% 8 == 0 and H <= 4096 (H=%d)", d[i].H);
    LnFwdProb& P = g.p[i];
    P.x = (const act_t*)d[i].x; P.gamma = (const act_t*)d[i].gamma; P.beta = (const act_t*)d[i].beta;
    P.y = (act_t*)d[i].y; P.mean = d[i].mean; P.rstd = d[i].rstd; P.M = d[i].M; P.H = d[i].H;
    P.drop = make_drop(d[i].rng, d[i].site, d[i].drop_p);
    P.cta_begin = 0;
    ctas += d[i].M;
    if (d[i].H > hmax) hmax = d[i].H;
  }
  for (int i = n; i < 8; ++i) g.p[i] = g.p[0];
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int nv = (hmax + 255) / 256;
  // rows per warp >= 2 once there are enough rows: at most 148 SMs x 4 CTAs of 8 warps
  ctas = (ctas + 7) / 8;
  if (ctas > 148 * 4) ctas = 148 * 4;
  if (nv <= 1) ln_fwd_group_kernel<1><<<ctas, 256, 0, st>>>(g);
  else if (nv <= 2) ln_fwd_group_kernel<2><<<ctas, 256, 0, st>>>(g);
  else if (nv <= 3) ln_fwd_group_kernel<3><<<ctas, 256, 0, st>>>(g);
  else if (nv <= 4) ln_fwd_group_kernel<4><<<ctas, 256, 0, st>>>(g);
  else if (nv <= 8) ln_fwd_group_kernel<8><<<ctas, 256, 0, st>>>(g);
  else ln_fwd_group_kernel<16><<<ctas, 256, 0, st>>>(g);
  return k3m_check_launch("k3m_ln_fwd_grouped");
}

extern "C" int k3m_ln_bwd_grouped(const void* descs, int n, void* stream) {
  const LnBwdDesc* d = reinterpret_cast<const LnBwdDesc*>(descs);
  K3M_REQUIRE(n >= 1 && n <= 8, "ln_bwd_grouped: 1..8 problems per launch (got %d)", n);
  LnBwdGroup g;
  g.n = n;
  long rows = 0;
  int hmax = 0;
  for (int i = 0; i < n; ++i) {
    K3M_REQUIRE(d[i].M > 0 && d[i].H > 0 && d[i].H % 8 == 0 && d[i].H <= 1024, "ln_bwd_grouped: need H %% 8 == 0 and H <= 1024 (H=%d)", d[i].H);
    rows += d[i].M;
    if (d[i].H > hmax) hmax = d[i].H;
  }
  const int nv = (hmax + 255) / 256;
  static const int sacc = getenv("K3M_LN_BWD_SACC") ? atoi(getenv("K3M_LN_BWD_SACC")) : 1;
  // tile form (cp.async double buffering, column sums in registers) for the wide problems of the step (H = 768 / 1024):
  // a third fewer instructions and no exposed row loads, but 109 registers (two CTAs per SM instead of three), three
  // block barriers per tile and a quarter of the phase-2 threads idle at H = 768 — measured EQUAL to the row-per-warp
  // form on the step (17.54 / 17.69 vs 17.59 / 17.64 ms, profiles/r03l_ln_bwd_tile_ab.txt; the GPU suite passes with
  // either; forced to 80 registers / three CTAs per SM it is 0.3 ms SLOWER, profiles/r03m_ln_bwd_tile_ab.txt).  Kept
  // selectable (K3M_LN_BWD_TILE=1) as the record of that experiment; off by default.
  static const int tile_mode = getenv("K3M_LN_BWD_TILE") ? atoi(getenv("K3M_LN_BWD_TILE")) : 0;
  static const int tile_ctas = getenv("K3M_LN_BWD_TILE_CTAS") ? atoi(getenv("K3M_LN_BWD_TILE_CTAS")) : 2;
  int hmin = hmax;
  for (int i = 0; i < n; ++i) hmin = d[i].H < hmin ? d[i].H : hmin;
  const bool tiled = tile_mode && hmin > 512;
  const int per_sm = tiled ? tile_ctas : (sacc ? (nv >= 4 ? 2 : 3) : (nv >= 4 ? 1 : 2));
  const int target = 148 * per_sm * ((sacc || tiled) ? 1 : 2);      // one (register accumulators: two) wave(s) of resident CTAs
  int ctas = 0;
  for (int i = 0; i < n; ++i) {
    LnBwdProb& P = g.p[i];
    P.dy = (const act_t*)d[i].dy; P.x = (const act_t*)d[i].x; P.gamma = (const act_t*)d[i].gamma;
    P.mean = d[i].mean; P.rstd = d[i].rstd; P.dx = (act_t*)d[i].dx; P.dxm = (act_t*)d[i].dxm;
    P.dgamma = d[i].dgamma; P.dbeta = d[i].dbeta; P.dbias = d[i].dbias; P.M = d[i].M; P.H = d[i].H;
    P.drop_out = make_drop(d[i].rng, d[i].site_out, d[i].p_out);
    P.drop_in = make_drop(d[i].rng, d[i].site_in, d[i].p_in);
    K3M_REQUIRE(!P.drop_in.thresh || P.dxm, "ln_bwd_grouped: drop_in needs the dxm output");
    int c = (int)(((long)target * d[i].M + rows - 1) / rows);
    const int cmax = (d[i].M + 7) / 8;
    if (c > cmax) c = cmax;
    if (c < 1) c = 1;
    P.cta_begin = ctas;
    P.nctas = c;
    ctas += c;
  }
  for (int i = n; i < 8; ++i) g.p[i] = g.p[0];
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t smem = (size_t)8 * 3 * hmax * sizeof(float);
#define LN_BWD_G(NV)                                                                                              \
  do {                                                                                                            \
    static bool cfg = false;                                                                                      \
    if (!cfg) {                                                                                                   \
      cudaFuncSetAttribute(ln_bwd_group_kernel<NV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 3 * 1024 * 4); \
      cfg = true;                                                                                                 \
    }                                                                                                             \
    if (sacc) {                                                                                                   \
      static bool cfg2 = false;                                                                                   \
      if (!cfg2) {                                                                                                \
        cudaFuncSetAttribute(ln_bwd_group_sacc_kernel<NV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 3 * 1024 * 4); \
        cfg2 = true;                                                                                              \
      }                                                                                                           \
      ln_bwd_group_sacc_kernel<NV><<<ctas, 256, smem, st>>>(g);                                                   \
    } else {                                                                                                      \
      ln_bwd_group_kernel<NV><<<ctas, 256, smem, st>>>(g);                                                        \
    }                                                                                                             \
  } while (0)
  if (tiled) {
    const size_t tsmem = (size_t)33 * hmax * sizeof(act_t) + 128;      // two tiles of 2 x 8 rows, gamma, row statistics
    static bool cfg3 = false;
    if (!cfg3) {
      cudaFuncSetAttribute(ln_bwd_group_tile_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 33 * 1024 * (int)sizeof(act_t) + 128);
      cudaFuncSetAttribute(ln_bwd_group_tile_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 33 * 1024 * (int)sizeof(act_t) + 128);
      cfg3 = true;
    }
    if (nv <= 3) ln_bwd_group_tile_kernel<3><<<ctas, 256, tsmem, st>>>(g);
    else ln_bwd_group_tile_kernel<4><<<ctas, 256, tsmem, st>>>(g);
    return k3m_check_launch("k3m_ln_bwd_grouped");
  }
  if (nv <= 1) LN_BWD_G(1);
  else if (nv <= 2) LN_BWD_G(2);
  else if (nv <= 3) LN_BWD_G(3);
  else LN_BWD_G(4);
#undef LN_BWD_G
  return k3m_check_launch("k3m_ln_bwd_grouped");
}

extern "C" int k3m_colsum(const void* x, int ld, int x_f32, float* out, int M, int N, void* stream) {
  K3M_REQUIRE(M > 0 && N > 0, "colsum: empty");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!x_f32 && N % 8 == 0 && ld % 8 == 0 && (((uintptr_t)x) & 15) == 0) {
    const int gx = (N + 255) / 256;
    int gy = (148 * 4 + gx - 1) / gx;             // ~4 CTAs per SM
    if (gy > (M + 63) / 64) gy = (M + 63) / 64;   // at least 64 rows per CTA
    if (gy < 1) gy = 1;
    const int rows_per = (M + gy - 1) / gy;
    colsum8_kernel<<<dim3(gx, gy), 256, 0, st>>>((const act_t*)x, ld, out, M, N, rows_per);
    return k3m_check_launch("k3m_colsum");
  }
  int gy = (M + 255) / 256;
  if (gy > 64) gy = 64;
  const int rows_per = (M + gy - 1) / gy;
  dim3 grid((N + 31) / 32, gy);
  if (x_f32) colsum_kernel<float><<<grid, 256, 0, st>>>((const float*)x, ld, out, M, N, rows_per);
  else colsum_kernel<act_t><<<grid, 256, 0, st>>>((const act_t*)x, ld, out, M, N, rows_per);
  return k3m_check_launch("k3m_colsum");
}

struct ColsumDesc {      // == K3MColsumDesc of include/k3m.h
  const void* x;
  int ld;
  float* out;
  int M, N;
};
extern "C" int k3m_colsum_grouped(const void* descs, int n, void* stream) {
  const ColsumDesc* d = reinterpret_cast<const ColsumDesc*>(descs);
  K3M_REQUIRE(n >= 1 && n <= 8, "colsum_grouped: 1..8 problems per launch (got %d)", n);
  ColsumGroup g;
  g.n = n;
  int total_cols = 0;
  for (int i = 0; i < n; ++i) total_cols += (d[i].N + 255) / 256;
  int ctas = 0;
  for (int i = 0; i < n; ++i) {
    K3M_REQUIRE(d[i].M > 0 && d[i].N > 0 && d[i].N % 8 == 0 && d[i].ld % 8 == 0 && (((uintptr_t)d[i].x) & 15) == 0,
                "colsum_grouped: activation-type input with N %% 8 == 0, ld %% 8 == 0, 16-byte aligned");
    ColsumProb& P = g.p[i];
    P.x = (const act_t*)d[i].x; P.out = d[i].out; P.ld = d[i].ld; P.M = d[i].M; P.N = d[i].N;
    P.gx = (d[i].N + 255) / 256;
    int gy = (148 * 4 + total_cols - 1) / total_cols;      // ~4 CTAs per SM over the whole group
    if (gy > (d[i].M + 63) / 64) gy = (d[i].M + 63) / 64;
    if (gy < 1) gy = 1;
    P.rows_per = (d[i].M + gy - 1) / gy;
    P.cta_begin = ctas;
    ctas += P.gx * gy;
  }
  for (int i = n; i < 8; ++i) g.p[i] = g.p[0];
  colsum8_group_kernel<<<ctas, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(g);
  return k3m_check_launch("k3m_colsum_grouped");
}

extern "C" int k3m_cast(const void* x, int x_f32, void* y, int y_f32, long long n, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int g = grid_for((size_t)n, 256);
  if (x_f32 && !y_f32) cast_kernel<float, act_t><<<g, 256, 0, st>>>((const float*)x, (act_t*)y, (size_t)n);
  else if (!x_f32 && y_f32) cast_kernel<act_t, float><<<g, 256, 0, st>>>((const act_t*)x, (float*)y, (size_t)n);
  else if (x_f32) cast_kernel<float, float><<<g, 256, 0, st>>>((const float*)x, (float*)y, (size_t)n);
  else cast_kernel<act_t, act_t><<<g, 256, 0, st>>>((const act_t*)x, (act_t*)y, (size_t)n);
  return k3m_check_launch("k3m_cast");
}

extern "C" int k3m_lincomb3(const void* a, int lda, const void* b, int ldb, const void* c, int ldc, float wa, float wb,
                            float wc, int in_f32, void* out, int ldo, int out_f32, int rows, int cols, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int g = grid_for((size_t)rows * cols, 256);
  if (in_f32 && out_f32)
    lincomb3_kernel<float, float><<<g, 256, 0, st>>>((const float*)a, (const float*)b, (const float*)c, wa, wb, wc,
                                                     (float*)out, rows, cols, lda, ldb, ldc, ldo);
  else if (!in_f32 && !out_f32)
    lincomb3_kernel<act_t, act_t><<<g, 256, 0, st>>>((const act_t*)a, (const act_t*)b, (const act_t*)c, wa, wb, wc,
                                                   (act_t*)out, rows, cols, lda, ldb, ldc, ldo);
  else if (in_f32)
    lincomb3_kernel<float, act_t><<<g, 256, 0, st>>>((const float*)a, (const float*)b, (const float*)c, wa, wb, wc,
                                                    (act_t*)out, rows, cols, lda, ldb, ldc, ldo);
  else
    lincomb3_kernel<act_t, float><<<g, 256, 0, st>>>((const act_t*)a, (const act_t*)b, (const act_t*)c, wa, wb, wc,
                                                    (float*)out, rows, cols, lda, ldb, ldc, ldo);
  return k3m_check_launch("k3m_lincomb3");
}

extern "C" int k3m_axpy(const void* x, int ldx, int x_f32, void* y, int ldy, int y_f32, float alpha, int rows, int cols,
                        void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  auto ok = [](const void* q, int ld, int f32) { return (((uintptr_t)q) & (f32 ? 31 : 15)) == 0 && ld % 8 == 0; };
  if (cols % 8 == 0 && ok(x, ldx, x_f32) && ok(y, ldy, y_f32) && (size_t)rows * cols < (1u << 31)) {
    const int g8 = grid_for((size_t)rows * cols / 8, 256);
    if (x_f32 && y_f32) axpy8_kernel<float, float><<<g8, 256, 0, st>>>((const float*)x, (float*)y, alpha, rows, cols / 8, ldx, ldy);
    else if (!x_f32 && !y_f32) axpy8_kernel<act_t, act_t><<<g8, 256, 0, st>>>((const act_t*)x, (act_t*)y, alpha, rows, cols / 8, ldx, ldy);
    else if (x_f32) axpy8_kernel<float, act_t><<<g8, 256, 0, st>>>((const float*)x, (act_t*)y, alpha, rows, cols / 8, ldx, ldy);
    else axpy8_kernel<act_t, float><<<g8, 256, 0, st>>>((const act_t*)x, (float*)y, alpha, rows, cols / 8, ldx, ldy);
    return k3m_check_launch("k3m_axpy");
  }
  const int g = grid_for((size_t)rows * cols, 256);
  if (x_f32 && y_f32) axpy_kernel<float, float><<<g, 256, 0, st>>>((const float*)x, (float*)y, alpha, rows, cols, ldx, ldy);
  else if (!x_f32 && !y_f32) axpy_kernel<act_t, act_t><<<g, 256, 0, st>>>((const act_t*)x, (act_t*)y, alpha, rows, cols, ldx, ldy);
  else if (x_f32) axpy_kernel<float, act_t><<<g, 256, 0, st>>>((const float*)x, (act_t*)y, alpha, rows, cols, ldx, ldy);
  else axpy_kernel<act_t, float><<<g, 256, 0, st>>>((const act_t*)x, (float*)y, alpha, rows, cols, ldx, ldy);
  return k3m_check_launch("k3m_axpy");
}

extern "C" int k3m_scale(float* x, float alpha, long long n, void* stream) {
  scale_kernel<<<grid_for((size_t)n, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(x, alpha, (size_t)n);
  return k3m_check_launch("k3m_scale");
}

extern "C" int k3m_scale_dev(void* x, int ld, int x_f32, const float* alpha, int rows, int cols, void* stream) {
  K3M_REQUIRE(alpha != nullptr && rows >= 0 && cols >= 0 && ld >= cols, "scale_dev: bad arguments");
  if (rows == 0 || cols == 0) return K3M_OK;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int g = grid_for((size_t)rows * cols, 256);
  if (x_f32) scale_dev_kernel<float><<<g, 256, 0, st>>>((float*)x, alpha, rows, cols, ld);
  else scale_dev_kernel<act_t><<<g, 256, 0, st>>>((act_t*)x, alpha, rows, cols, ld);
  return k3m_check_launch("k3m_scale_dev");
}

extern "C" int k3m_dgelu_mul(const void* dy, const void* pre, void* dx, long long n, void* stream) {
  K3M_REQUIRE(n % 8 == 0, "dgelu_mul: n must be a multiple of 8");
  dgelu_kernel<<<grid_for((size_t)n / 8, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      (const act_t*)dy, (const act_t*)pre, (act_t*)dx, (size_t)n / 8);
  return k3m_check_launch("k3m_dgelu_mul");
}

extern "C" int k3m_mask_additive(const long long* mask, float* out, int n, void* stream) {
  mask_additive_kernel<<<(n + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(mask, out, n);
  return k3m_check_launch("k3m_mask_additive");
}
