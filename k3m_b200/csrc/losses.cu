// Loss kernels (masked-LM cross entropy, masked-region KL, ITM value), the batched structure
// aggregator + link-prediction loss, and multi-tensor AdamW.
#include <stdlib.h>

#include "common.cuh"

namespace {

#define ST(s) reinterpret_cast<cudaStream_t>(s)

// ---- label counts ---------------------------------------------------------------------------
__global__ void count_kernel(const long long* __restrict__ lab, int n, long long value, int equal, float* __restrict__ out) {
  __shared__ float red[32];
  float c = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) c += ((lab[i] == value) == (equal != 0)) ? 1.f : 0.f;
  c = block_sum(c, red);
  if (threadIdx.x == 0) out[0] = c;
}

// ---- masked-LM cross entropy (CrossEntropyLoss(ignore_index=-1), vilbert_k3m.py:2817-2826) -----------
// One CTA per row.  loss += (lse - logit[label]) / count ; dlogits = (softmax - onehot) * gscale / count
// written in place over the logits (rows without a label become zero).
__global__ void __launch_bounds__(256) xent_kernel(act_t* __restrict__ logits, int ld, const long long* __restrict__ labels,
                                                   const float* __restrict__ count, int V, float gscale,
                                                   float* __restrict__ loss, int write_grad) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  act_t* lr = logits + (size_t)row * ld;
  const long long lab = labels[row];
  if (lab < 0 || lab >= V) {
    // ignore_index = -1 (`vilbert_k3m.py:2207`); any other label outside [0, V) makes torch's CrossEntropyLoss
    // raise — no exception can leave a captured graph, so the loss turns NaN instead of reading out of bounds
    if ((lab < -1 || lab >= V) && threadIdx.x == 0) atomicAdd(loss, __int_as_float(0x7fc00000));
    if (write_grad)
      for (int c = threadIdx.x; c < ld; c += blockDim.x) lr[c] = act_from(0.f);
    return;
  }
  float mx = -3.0e38f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) mx = fmaxf(mx, to_f(lr[c]));
  mx = block_max(mx, red);
  float s = 0.f;
  for (int c = threadIdx.x; c < V; c += blockDim.x) s += __expf(to_f(lr[c]) - mx);
  s = block_sum(s, red);
  const float lse = mx + __logf(s);
  const float inv_cnt = 1.0f / count[0];
  if (threadIdx.x == 0) atomicAdd(loss, (lse - to_f(lr[lab])) * inv_cnt);
  if (write_grad) {
    const float g = gscale * inv_cnt;
    __syncthreads();
    for (int c = threadIdx.x; c < ld; c += blockDim.x) {
      float v = 0.f;
      if (c < V) v = (__expf(to_f(lr[c]) - lse) - (c == lab ? 1.f : 0.f)) * g;
      lr[c] = act_from(v);
    }
  }
}

// Same contract, one pass: the whole row (V <= 256 * 8 * NVT) is held packed in registers — one 16-byte-vector read and
// one write of the row instead of three scalar sweeps (2-byte loads, 64 B per warp instruction: 0.2 of the HBM roof).
// Needs V % 8 == 0 and 16-byte aligned rows; ld - V (< 8 by the engine's padding rule, any value here) tail columns are
// zeroed by the scalar loop at the end.
template <int NVT>
__global__ void __launch_bounds__(256) xent_row_kernel(act_t* __restrict__ logits, int ld, const long long* __restrict__ labels,
                                                       const float* __restrict__ count, int V, float gscale,
                                                       float* __restrict__ loss, int write_grad) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  act_t* lr = logits + (size_t)row * ld;
  const long long lab = labels[row];
  const int nvec = V >> 3;
  if (lab < 0 || lab >= V) {
    if ((lab < -1 || lab >= V) && threadIdx.x == 0) atomicAdd(loss, __int_as_float(0x7fc00000));
    if (write_grad) {
      float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int i = threadIdx.x; i < nvec; i += 256) st8(lr + i * 8, z);
      for (int c = V + threadIdx.x; c < ld; c += 256) lr[c] = act_from(0.f);
    }
    return;
  }
  Pack8 pk[NVT];
  float mx = -3.0e38f;
#pragma unroll
  for (int t = 0; t < NVT; ++t) {
    const int i = threadIdx.x + t * 256;
    if (i < nvec) pk[t] = load_pack8(lr + i * 8);
  }
#pragma unroll
  for (int t = 0; t < NVT; ++t) {
    const int i = threadIdx.x + t * 256;
    if (i < nvec) {
      float v[8];
      unpack8(pk[t], v);
#pragma unroll
      for (int j = 0; j < 8; ++j) mx = fmaxf(mx, v[j]);
    }
  }
  mx = block_max(mx, red);
  float s = 0.f;
#pragma unroll
  for (int t = 0; t < NVT; ++t) {
    const int i = threadIdx.x + t * 256;
    if (i < nvec) {
      float v[8];
      unpack8(pk[t], v);
#pragma unroll
      for (int j = 0; j < 8; ++j) s += __expf(v[j] - mx);
    }
  }
  s = block_sum(s, red);
  const float lse = mx + __logf(s);
  const float inv_cnt = 1.0f / count[0];
  if (threadIdx.x == 0) atomicAdd(loss, (lse - to_f(lr[lab])) * inv_cnt);
  if (write_grad) {
    const float g = gscale * inv_cnt;
    __syncthreads();                      // lr[lab] has been read by thread 0
    const int lv = (int)(lab >> 3), lj = (int)(lab & 7);
#pragma unroll
    for (int t = 0; t < NVT; ++t) {
      const int i = threadIdx.x + t * 256;
      if (i < nvec) {
        float v[8];
        unpack8(pk[t], v);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = (__expf(v[j] - lse) - ((i == lv && j == lj) ? 1.f : 0.f)) * g;
        st8(lr + i * 8, v);
      }
    }
    for (int c = V + threadIdx.x; c < ld; c += 256) lr[c] = act_from(0.f);
  }
}

// ---- masked-region KL (KLDivLoss(none)(log_softmax, target) * (label==1) / count, :2754-2760) ------
// One CTA per (item, region) row of the fp32 logits [B*Lv1, ld]; row 0 of every item is the global
// region (no target).  dlogits (act_t, same ld) = (softmax * sum(t) - t) * gscale / count.
__global__ void __launch_bounds__(256) kl_kernel(const float* __restrict__ logits, int ld, const float* __restrict__ target,
                                                 const long long* __restrict__ label, const float* __restrict__ count, int T,
                                                 int Lv1, float gscale, float* __restrict__ loss, act_t* __restrict__ dlogits) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  const int b = row / Lv1, i = row % Lv1;
  act_t* dr = dlogits ? dlogits + (size_t)row * ld : nullptr;
  const bool active = i > 0 && label[(size_t)b * (Lv1 - 1) + (i - 1)] == 1;
  if (!active) {
    if (dr)
      for (int c = threadIdx.x; c < ld; c += blockDim.x) dr[c] = act_from(0.f);
    return;
  }
  const float* lr = logits + (size_t)row * ld;
  const float* tr = target + ((size_t)b * (Lv1 - 1) + (i - 1)) * T;
  float mx = -3.0e38f;
  for (int c = threadIdx.x; c < T; c += blockDim.x) mx = fmaxf(mx, lr[c]);
  mx = block_max(mx, red);
  float s = 0.f, ts = 0.f;
  for (int c = threadIdx.x; c < T; c += blockDim.x) {
    s += __expf(lr[c] - mx);
    ts += tr[c];
  }
  s = block_sum(s, red);
  ts = block_sum(ts, red);
  const float lse = mx + __logf(s);
  float kl = 0.f;
  for (int c = threadIdx.x; c < T; c += blockDim.x) {
    const float t = tr[c];
    if (t > 0.f) kl += t * (__logf(t) - (lr[c] - lse));
  }
  kl = block_sum(kl, red);
  const float inv_cnt = 1.0f / count[0];
  if (threadIdx.x == 0) atomicAdd(loss, kl * inv_cnt);
  if (dr) {
    const float g = gscale * inv_cnt;
    for (int c = threadIdx.x; c < ld; c += blockDim.x) {
      float v = 0.f;
      if (c < T) v = (__expf(lr[c] - lse) * ts - tr[c]) * g;
      dr[c] = act_from(v);
    }
  }
}

// ---- ITM value: CE(Linear_{H->2}(dropout(pool_t + pool_pv + pool_v)), 1 - ((l1+l2+l3)==0)) -----------
__global__ void __launch_bounds__(128) itm_kernel(const float* pt, const float* ppv, const float* pv, const act_t* w,
                                                  const act_t* bias, const long long* l1, const long long* l2,
                                                  const long long* l3, int B, int H, DropSpec drop, float* loss) {
  __shared__ float red[32];
  const int b = blockIdx.x;
  float a0 = 0.f, a1 = 0.f;
  for (int c = threadIdx.x; c < H; c += blockDim.x) {
    float x = pt[(size_t)b * H + c] + ppv[(size_t)b * H + c] + pv[(size_t)b * H + c];
    if (drop.thresh) x *= drop_mult_at(drop, (unsigned long long)b * ((H + 7) / 8 * 8) + c);
    a0 = fmaf(x, ld_f(w + c), a0);
    a1 = fmaf(x, ld_f(w + H + c), a1);
  }
  a0 = block_sum(a0, red) + ld_f(bias);
  a1 = block_sum(a1, red) + ld_f(bias + 1);
  if (threadIdx.x == 0) {
    const int lab = ((l1[b] + l2[b] + l3[b]) == 0) ? 0 : 1;
    const float mx = fmaxf(a0, a1);
    const float lse = mx + __logf(__expf(a0 - mx) + __expf(a1 - mx));
    atomicAdd(loss, (lse - (lab ? a1 : a0)) / (float)B);
  }
}

// ---- item-alignment head (K3MForItemAlignment.forward "ce" branch, vilbert_k3m.py:3443-3450) ----------
// logits fp32 [B, 2] (row stride ld) -> probs = softmax(logits) [B, 2]; loss = mean CE(logits, labels);
// dlogits (act_t, row stride ldd, columns 0..1) = grad_scale * (probs - onehot) / B
__global__ void __launch_bounds__(128) pair_cls_kernel(const float* __restrict__ logits, int ld,
                                                       const long long* __restrict__ labels, int B, float grad_scale,
                                                       float* __restrict__ loss, float* __restrict__ probs,
                                                       act_t* __restrict__ dlogits, int ldd) {
  __shared__ float red[32];
  float acc = 0.f;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    const float a0 = logits[(size_t)b * ld], a1 = logits[(size_t)b * ld + 1];
    const float mx = fmaxf(a0, a1);
    const float e0 = __expf(a0 - mx), e1 = __expf(a1 - mx);
    const float inv = 1.0f / (e0 + e1);
    const float p0 = e0 * inv, p1 = e1 * inv;
    const int lab = labels[b] != 0 ? 1 : 0;
    acc += (mx + __logf(e0 + e1)) - (lab ? a1 : a0);
    if (probs) {
      probs[(size_t)b * 2] = p0;
      probs[(size_t)b * 2 + 1] = p1;
    }
    if (dlogits) {
      const float s = grad_scale / (float)B;
      dlogits[(size_t)b * ldd] = act_from(s * (p0 - (lab ? 0.f : 1.f)));
      dlogits[(size_t)b * ldd + 1] = act_from(s * (p1 - (lab ? 1.f : 0.f)));
    }
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) *loss = acc / (float)B;
}

// y = x * keep-mask * 1/(1-p): nn.Dropout applied outside a GEMM epilogue (ClassificationHead, :2176-2180);
// the same (site, element index) regenerates the mask for the backward pass
template <typename TX, typename TY>
__global__ void dropout_kernel(const TX* __restrict__ x, int ldx, TY* __restrict__ y, int ldy, int rows, int cols,
                               int ld_drop, DropSpec drop) {
  const size_t n = (size_t)rows * cols;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols), c = (int)(i % cols);
    float v = ld_f(x + (size_t)r * ldx + c);
    if (drop.thresh) v *= drop_mult_at(drop, (unsigned long long)r * ld_drop + c);
    st_f(y + (size_t)r * ldy + c, v);
  }
}

// ---- structure aggregator (structure_aggregator, vilbert_k3m.py:2437-2467) ---------------------------
// X[p] = [c_init[item] | mean(seq[r0], seq[r1]) | mean(seq[r2], seq[r3])]  (act_t for the W1 GEMM);
// PV[p] = the two means in fp32 for the LPM norms.
__global__ void pair_gather_fwd_kernel(const float* __restrict__ c_init, const act_t* __restrict__ seq, int lds,
                                       const int* __restrict__ item, const int* __restrict__ rows, act_t* __restrict__ X,
                                       float* __restrict__ PV, int P, int H) {
  const size_t n = (size_t)P * H;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int p = (int)(i / H), h = (int)(i % H);
    const int* r = rows + (size_t)p * 4;
    const float pm = 0.5f * (to_f(seq[(size_t)r[0] * lds + h]) + to_f(seq[(size_t)r[1] * lds + h]));
    const float vm = 0.5f * (to_f(seq[(size_t)r[2] * lds + h]) + to_f(seq[(size_t)r[3] * lds + h]));
    act_t* x = X + (size_t)p * 3 * H;
    x[h] = act_from(c_init[(size_t)item[p] * H + h]);
    x[H + h] = act_from(pm);
    x[2 * H + h] = act_from(vm);
    PV[(size_t)p * 2 * H + h] = pm;
    PV[(size_t)p * 2 * H + H + h] = vm;
  }
}

// one CTA per pair (pairs of one item may share sequence rows -> read-modify-write must not race
// across CTAs: sequence rows get fp32 atomics on a scratch [rows, H] buffer that the caller adds).
__global__ void pair_gather_bwd_kernel(const float* __restrict__ dX, const float* __restrict__ dPV, const int* __restrict__ item,
                                       const int* __restrict__ rows, float* __restrict__ dc_init, float* __restrict__ dseq32,
                                       int P, int H) {
  const size_t n = (size_t)P * H;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int p = (int)(i / H), h = (int)(i % H);
    const int* r = rows + (size_t)p * 4;
    const float* dx = dX + (size_t)p * 3 * H;
    atomicAdd(dc_init + (size_t)item[p] * H + h, dx[h]);
    const float gp = 0.5f * (dx[H + h] + dPV[(size_t)p * 2 * H + h]);
    const float gv = 0.5f * (dx[2 * H + h] + dPV[(size_t)p * 2 * H + H + h]);
    atomicAdd(dseq32 + (size_t)r[0] * H + h, gp);
    atomicAdd(dseq32 + (size_t)r[1] * H + h, gp);
    atomicAdd(dseq32 + (size_t)r[2] * H + h, gv);
    atomicAdd(dseq32 + (size_t)r[3] * H + h, gv);
  }
}

// Leading items without any property/value pair: the reference's `t` is unbound there, its bare `except`
// sets t = c_initial[i] (vilbert_k3m.py:2452-2456) and later items without pairs keep reusing that row.
// The plan holds one pseudo pair row per `structure_aggregator` call for it: fb = (row of t, item) or (-1, -1).
// Forward: t[row] = c_init[item] (bypassing struc_w1).  Backward: the row's gradient goes straight to
// dc_init[item] and is zeroed, so nothing flows into struc_w1 / the gathered sequence rows from it.
__global__ void struct_fallback_fwd_kernel(float* __restrict__ t, const float* __restrict__ c_init,
                                           const int* __restrict__ fb, int H) {
  const int row = fb[2 * blockIdx.x], item = fb[2 * blockIdx.x + 1];
  if (row < 0) return;
  for (int h = threadIdx.x; h < H; h += blockDim.x) t[(size_t)row * H + h] = c_init[(size_t)item * H + h];
}
__global__ void struct_fallback_bwd_kernel(float* __restrict__ dt, float* __restrict__ dc_init,
                                           const int* __restrict__ fb, int H) {
  const int row = fb[2 * blockIdx.x], item = fb[2 * blockIdx.x + 1];
  if (row < 0) return;
  for (int h = threadIdx.x; h < H; h += blockDim.x) {
    atomicAdd(dc_init + (size_t)item * H + h, dt[(size_t)row * H + h]);
    dt[(size_t)row * H + h] = 0.f;
  }
}

// one CTA per item: scores b_j = w2 . leaky_relu(t_j) + b2 over the item's pair range, softmax,
// s = sum_j att_j t_j.  att is stored [B, maxn].
__global__ void __launch_bounds__(256) struct_attn_fwd_kernel(const float* __restrict__ t, const act_t* __restrict__ w2,
                                                              const act_t* __restrict__ b2, const int* __restrict__ seg,
                                                              float* __restrict__ s_out, float* __restrict__ att, int maxn, int H) {
  extern __shared__ float sc[];  // [maxn]
  __shared__ float red[32];
  const int i = blockIdx.x, s0 = seg[2 * i], n = seg[2 * i + 1] - s0;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int j = warp; j < n; j += nw) {
    const float* tr = t + (size_t)(s0 + j) * H;
    float a = 0.f;
    for (int c = lane; c < H; c += 32) {
      const float v = tr[c];
      a = fmaf(v > 0.f ? v : 0.01f * v, ld_f(w2 + c), a);
    }
    a = warp_sum(a);
    if (lane == 0) sc[j] = a + ld_f(b2);
  }
  __syncthreads();
  float mx = -3.0e38f;
  for (int j = threadIdx.x; j < n; j += blockDim.x) mx = fmaxf(mx, sc[j]);
  mx = block_max(mx, red);
  float sum = 0.f;
  for (int j = threadIdx.x; j < n; j += blockDim.x) sum += __expf(sc[j] - mx);
  sum = block_sum(sum, red);
  __syncthreads();
  for (int j = threadIdx.x; j < n; j += blockDim.x) {
    const float a = __expf(sc[j] - mx) / sum;
    sc[j] = a;
    att[(size_t)i * maxn + j] = a;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < H; c += blockDim.x) {
    float acc = 0.f;
    for (int j = 0; j < n; ++j) acc = fmaf(sc[j], t[(size_t)(s0 + j) * H + c], acc);
    s_out[(size_t)i * H + c] = acc;
  }
}

__global__ void __launch_bounds__(256) struct_attn_bwd_kernel(const float* __restrict__ ds, const float* __restrict__ t,
                                                              const float* __restrict__ att, const act_t* __restrict__ w2,
                                                              const int* __restrict__ seg, float* __restrict__ dt,
                                                              float* __restrict__ dw2, float* __restrict__ db2, int maxn, int H) {
  extern __shared__ float sm[];  // da[maxn], dsc[maxn]
  __shared__ float red[32];
  float* da = sm;
  float* dsc = sm + maxn;
  const int i = blockIdx.x, s0 = seg[2 * i], n = seg[2 * i + 1] - s0;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const float* dsr = ds + (size_t)i * H;
  for (int j = warp; j < n; j += nw) {
    const float* tr = t + (size_t)(s0 + j) * H;
    float a = 0.f;
    for (int c = lane; c < H; c += 32) a = fmaf(tr[c], dsr[c], a);
    a = warp_sum(a);
    if (lane == 0) da[j] = a;
  }
  __syncthreads();
  float dot = 0.f;
  for (int j = threadIdx.x; j < n; j += blockDim.x) dot += att[(size_t)i * maxn + j] * da[j];
  dot = block_sum(dot, red);
  float bsum = 0.f;
  for (int j = threadIdx.x; j < n; j += blockDim.x) {
    const float a = att[(size_t)i * maxn + j];
    dsc[j] = a * (da[j] - dot);
    bsum += dsc[j];
  }
  bsum = block_sum(bsum, red);
  if (threadIdx.x == 0) atomicAdd(db2, bsum);
  __syncthreads();
  for (int c = threadIdx.x; c < H; c += blockDim.x) {
    const float wc = ld_f(w2 + c), g = dsr[c];
    float dwc = 0.f;
    for (int j = 0; j < n; ++j) {
      const float tv = t[(size_t)(s0 + j) * H + c];
      const float a = att[(size_t)i * maxn + j];
      atomicAdd(dt + (size_t)(s0 + j) * H + c, a * g + dsc[j] * wc * (tv > 0.f ? 1.f : 0.01f));
      dwc = fmaf(dsc[j], tv > 0.f ? tv : 0.01f * tv, dwc);
    }
    atomicAdd(dw2 + c, dwc);
  }
}

// ---- link-prediction loss (vilbert_k3m.py:2469-2502) --------------------------------------------------
// one CTA per term (i, j, ni, nv): pos = |c[i] + p[j] - v[j]|, neg = |c[ni] + p[j] - v[nv]|,
// loss += max(0, neg - pos + margin) / T, and the gradient goes straight to dc_final / dPV.
__global__ void __launch_bounds__(128) lpm_kernel(const float* __restrict__ c_final, const float* __restrict__ PV,
                                                  const int* __restrict__ terms, int T_cap,
                                                  const int* __restrict__ n_valid, int H, float margin,
                                                  float* __restrict__ loss, float* __restrict__ dc, float* __restrict__ dPV) {
  __shared__ float red[32];
  const int tix = blockIdx.x;
  // the term list may be padded to a fixed capacity (CUDA-graph replay); the real count lives on the device
  const int T = n_valid ? *n_valid : T_cap;
  if (tix >= T) return;
  const int i = terms[4 * tix], j = terms[4 * tix + 1], ni = terms[4 * tix + 2], nv = terms[4 * tix + 3];
  const float* ci = c_final + (size_t)i * H;
  const float* cn = c_final + (size_t)ni * H;
  const float* pj = PV + (size_t)j * 2 * H;
  const float* vj = pj + H;
  const float* vn = PV + (size_t)nv * 2 * H + H;
  float sp = 0.f, sn = 0.f;
  for (int c = threadIdx.x; c < H; c += blockDim.x) {
    const float up = ci[c] + pj[c] - vj[c], un = cn[c] + pj[c] - vn[c];
    sp = fmaf(up, up, sp);
    sn = fmaf(un, un, sn);
  }
  sp = block_sum(sp, red);
  sn = block_sum(sn, red);
  const float npos = sqrtf(sp), nneg = sqrtf(sn);
  const float hinge = nneg - npos + margin;
  if (hinge <= 0.f) return;
  const float invT = 1.0f / (float)T;
  if (threadIdx.x == 0) atomicAdd(loss, hinge * invT);
  const float gp = -invT / fmaxf(npos, 1e-30f), gn = invT / fmaxf(nneg, 1e-30f);
  for (int c = threadIdx.x; c < H; c += blockDim.x) {
    const float up = (ci[c] + pj[c] - vj[c]) * gp, un = (cn[c] + pj[c] - vn[c]) * gn;
    atomicAdd(dc + (size_t)i * H + c, up);
    atomicAdd(dc + (size_t)ni * H + c, un);
    atomicAdd(dPV + (size_t)j * 2 * H + c, up + un);
    atomicAdd(dPV + (size_t)j * 2 * H + H + c, -up);
    atomicAdd(dPV + (size_t)nv * 2 * H + H + c, -un);
  }
}

// ---- multi-tensor AdamW over a flat arena range (torch.optim.AdamW semantics, pretrain.py:919-924) ----
// hyper (device fp32[8]) = lr, beta1, beta2, eps, weight_decay, step, grad_scale, unused.  Reading the
// hyper-parameters from device memory keeps the kernel node valid across CUDA-graph replays.
// G16: the (all-reduced) gradient is read from a bf16 buffer `g16` instead of the fp32 arena `g`, which is still zeroed.
template <bool G16>
__global__ void __launch_bounds__(256) adamw_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                                                    float* __restrict__ v, bf16* __restrict__ p16,
                                                    const float* __restrict__ hyper, size_t n4, int zero_grad,
                                                    const bf16* __restrict__ g16) {
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4], step = hyper[5], gs = hyper[6];
  const float bc1 = 1.0f - powf(b1, step), bc2 = 1.0f - powf(b2, step);
  const float step_size = lr / bc1, inv_sqrt_bc2 = rsqrtf(bc2), decay = 1.0f - lr * wd;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i], gg;
    if (G16) {
      const uint2 gu = reinterpret_cast<const uint2*>(g16)[i];
      const float2 ga2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gu.x));
      const float2 gb2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gu.y));
      gg = make_float4(ga2.x, ga2.y, gb2.x, gb2.y);
    } else {
      gg = reinterpret_cast<float4*>(g)[i];
    }
    float4 mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    float* pa = &pp.x; float* ga = &gg.x; float* ma = &mm.x; float* va = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gr = ga[k] * gs;
      pa[k] *= decay;
      ma[k] = b1 * ma[k] + (1.0f - b1) * gr;
      va[k] = b2 * va[k] + (1.0f - b2) * gr * gr;
      pa[k] -= step_size * ma[k] / (sqrtf(va[k]) * inv_sqrt_bc2 + eps);
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __nv_bfloat162 lo = __floats2bfloat162_rn(pp.x, pp.y), hi = __floats2bfloat162_rn(pp.z, pp.w);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&lo);
    u.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(p16)[i] = u;
  }
}

// "Background" AdamW: the same update, shaped to run UNDER the GEMMs of backward instead of between them.  A GEMM
// CTA holds 55 K of an SM's 64 K registers and all but 2 KB of its shared memory, so a kernel only co-resides
// if it fits into what is left: 10 240 registers and ONE more CTA (the 1 KB of shared memory the system reserves
// per CTA is all that is free) — one 256-thread CTA of <= 40 registers per SM, one float4 per array in flight
// per thread.  At that occupancy it streams ~3 TB/s instead of ~6 — which costs
// nothing while the tensor-bound GEMMs it hides under leave HBM idle (measured before: the 2.4 ms of AdamW per
// step were fully exposed, because neither kernel could start on an SM the other one occupied).
__global__ void __launch_bounds__(256, 6) adamw_bg_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                                                           float* __restrict__ v, bf16* __restrict__ p16,
                                                           const float* __restrict__ hyper, size_t n4, int zero_grad) {
  const float b1 = hyper[1], b2 = hyper[2], gs = hyper[6];
  const float step = hyper[5];
  const float step_size = hyper[0] / (1.0f - powf(b1, step)), inv_sqrt_bc2 = rsqrtf(1.0f - powf(b2, step));
  const float decay = 1.0f - hyper[0] * hyper[4], eps = hyper[3];
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i], gg = reinterpret_cast<float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    float* pa = &pp.x; float* ga = &gg.x; float* ma = &mm.x; float* va = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gr = ga[k] * gs;
      pa[k] *= decay;
      ma[k] = b1 * ma[k] + (1.0f - b1) * gr;
      va[k] = b2 * va[k] + (1.0f - b2) * gr * gr;
      pa[k] -= step_size * ma[k] / (sqrtf(va[k]) * inv_sqrt_bc2 + eps);
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __nv_bfloat162 lo = __floats2bfloat162_rn(pp.x, pp.y), hi = __floats2bfloat162_rn(pp.z, pp.w);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&lo);
    u.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(p16)[i] = u;
  }
}

__device__ __forceinline__ uint4 pack8_bf16(const float v[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  return u;
}

// ---- data parallel: gradient reduce-scatter + AdamW + weight all-gather as ONE kernel over NVLink peer memory ----
// Every rank owns 1/world of each gradient bucket (ZeRO-1 style).  For its shard the kernel
//   * loads the bf16 gradient of every rank straight from that rank's HBM (peer-to-peer loads over NVLink /
//     NVSwitch; `g[r]` = base of rank r's symmetric bf16 gradient arena) and sums in fp32,
//   * applies torch.optim.AdamW's update to its fp32 master / moments (local HBM, 1/world of the optimizer traffic
//     of a replicated update),
//   * stores the new bf16 weight into EVERY rank's weight arena (`w[r]`, peer-to-peer stores),
// which replaces all-reduce + a replicated AdamW: the same NVLink bytes as a ring all-reduce, no staging, no
// NCCL kernel waiting for an SM, and 8x less optimizer traffic at 8 GPUs.  The caller brackets it with barriers
// (all ranks' gradient casts are visible before / all weight stores before the next forward).
struct DpPeers {
  const bf16* g[16];
  bf16* w[16];
  int world;
};
// MC: the symmetric arenas also have a MULTICAST mapping (NVSwitch, NVLS): one multimem.ld_reduce returns the sum of
// all ranks' gradients (added inside the switch, fp32 accumulation, bf16 result) and one multimem.st writes the new
// weights into every rank's arena — 1/world of the peer-load bytes and one store instead of `world`.
template <bool MC>
__global__ void __launch_bounds__(256) adamw_dp_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                                                       const float* __restrict__ hyper, const __grid_constant__ DpPeers peers,
                                                       size_t offset, size_t n8, const bf16* __restrict__ g_mc,
                                                       bf16* __restrict__ w_mc) {
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4], step = hyper[5], gs = hyper[6];
  const float bc1 = 1.0f - powf(b1, step), bc2 = 1.0f - powf(b2, step);
  const float step_size = lr / bc1, inv_sqrt_bc2 = rsqrtf(bc2), decay = 1.0f - lr * wd;
  const int world = peers.world;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    const size_t e = offset + i * 8;                  // element index inside the arena
    float g[8];
    if (MC) {
      uint4 u;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                   : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "l"(g_mc + e) : "memory");
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = __bfloat1622float2(h[k]);
        g[2 * k] = f.x; g[2 * k + 1] = f.y;
      }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = 0.f;
      for (int r = 0; r < world; ++r) {
        float t[8];
        ld8(peers.g[r] + e, t);
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] += t[k];
      }
    }
    float pa[8], ma[8], va[8];
    ld8(p + i * 8, pa);
    ld8(m + i * 8, ma);
    ld8(v + i * 8, va);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float gr = g[k] * gs;
      pa[k] *= decay;
      ma[k] = b1 * ma[k] + (1.0f - b1) * gr;
      va[k] = b2 * va[k] + (1.0f - b2) * gr * gr;
      pa[k] -= step_size * ma[k] / (sqrtf(va[k]) * inv_sqrt_bc2 + eps);
    }
    st8(p + i * 8, pa);
    st8(m + i * 8, ma);
    st8(v + i * 8, va);
    const uint4 wv = pack8_bf16(pa);
    if (MC) {
      asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};"
                   :: "l"(w_mc + e), "r"(wv.x), "r"(wv.y), "r"(wv.z), "r"(wv.w) : "memory");
    } else {
      for (int r = 0; r < world; ++r) *reinterpret_cast<uint4*>(peers.w[r] + e) = wv;
    }
  }
}

// gradient bucket fp32 -> bf16 send buffer, zeroing the fp32 arena for the next step in the same pass
__global__ void __launch_bounds__(256) cast_zero_kernel(float* __restrict__ g, bf16* __restrict__ out, size_t n8) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    float t[8];
    ld8(g + i * 8, t);
    st8(out + i * 8, t);
    reinterpret_cast<float4*>(g + i * 8)[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    reinterpret_cast<float4*>(g + i * 8)[1] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

}  // namespace

extern "C" int k3m_cast_zero(float* g32, void* out16, long long n, void* stream) {
  K3M_REQUIRE(n > 0 && n % 8 == 0 && (((uintptr_t)g32) & 31) == 0 && (((uintptr_t)out16) & 15) == 0,
              "cast_zero: range must be a multiple of 8 elements, 32-byte aligned");
  const size_t n8 = (size_t)n / 8;
  size_t g = (n8 + 255) / 256;
  if (g > 148 * 16) g = 148 * 16;
  cast_zero_kernel<<<(int)g, 256, 0, ST(stream)>>>(g32, (bf16*)out16, n8);
  return k3m_check_launch("k3m_cast_zero");
}

extern "C" int k3m_adamw_dp(float* p32, float* m, float* v, const float* hyper, const void* const* g16_peers,
                            void* const* w16_peers, int world, long long offset, long long n, const void* g16_mc,
                            void* w16_mc, void* stream) {
  K3M_REQUIRE(world >= 1 && world <= 16, "adamw_dp: 1..16 ranks (got %d)", world);
  K3M_REQUIRE(n > 0 && n % 8 == 0 && offset % 8 == 0, "adamw_dp: shard offset / length must be multiples of 8 elements");
  K3M_REQUIRE((((uintptr_t)p32 | (uintptr_t)m | (uintptr_t)v) & 31) == 0, "adamw_dp: misaligned shard");
  DpPeers peers;
  peers.world = world;
  for (int r = 0; r < 16; ++r) {
    peers.g[r] = (const bf16*)g16_peers[r < world ? r : 0];
    peers.w[r] = (bf16*)w16_peers[r < world ? r : 0];
    K3M_REQUIRE((((uintptr_t)peers.g[r] | (uintptr_t)peers.w[r]) & 15) == 0, "adamw_dp: misaligned peer arena");
  }
  const size_t n8 = (size_t)n / 8;
  size_t g = (n8 + 255) / 256;
  if (g > 148 * 8) g = 148 * 8;
  if (g16_mc != nullptr && w16_mc != nullptr) {
    K3M_REQUIRE((((uintptr_t)g16_mc | (uintptr_t)w16_mc) & 15) == 0, "adamw_dp: misaligned multicast arena");
    adamw_dp_kernel<true><<<(int)g, 256, 0, ST(stream)>>>(p32, m, v, hyper, peers, (size_t)offset, n8, (const bf16*)g16_mc,
                                                          (bf16*)w16_mc);
  } else {
    adamw_dp_kernel<false><<<(int)g, 256, 0, ST(stream)>>>(p32, m, v, hyper, peers, (size_t)offset, n8, nullptr, nullptr);
  }
  return k3m_check_launch("k3m_adamw_dp");
}

extern "C" int k3m_count_labels(const long long* labels, int n, long long value, int equal, float* out, void* stream) {
  count_kernel<<<1, 1024, 0, ST(stream)>>>(labels, n, value, equal, out);
  return k3m_check_launch("k3m_count_labels");
}

extern "C" int k3m_xent_fwd_bwd(void* logits, int ld, const long long* labels, const float* count, int M, int V,
                                float grad_scale, float* loss, int write_grad, void* stream) {
  K3M_REQUIRE(M > 0 && V > 0 && ld >= V, "xent: bad shape");
  const bool rows_ok = V % 8 == 0 && (((uintptr_t)logits) & 15) == 0 && ((size_t)ld * sizeof(act_t)) % 16 == 0;
  if (rows_ok && V <= 256 * 8 * 4)
    xent_row_kernel<4><<<M, 256, 0, ST(stream)>>>((act_t*)logits, ld, labels, count, V, grad_scale, loss, write_grad);
  else if (rows_ok && V <= 256 * 8 * 12)
    xent_row_kernel<12><<<M, 256, 0, ST(stream)>>>((act_t*)logits, ld, labels, count, V, grad_scale, loss, write_grad);
  else
    xent_kernel<<<M, 256, 0, ST(stream)>>>((act_t*)logits, ld, labels, count, V, grad_scale, loss, write_grad);
  return k3m_check_launch("k3m_xent_fwd_bwd");
}

extern "C" int k3m_kl_fwd_bwd(const float* logits, int ld, const float* target, const long long* image_label,
                              const float* count, int B, int Lv1, int T, float grad_scale, float* loss, void* dlogits,
                              void* stream) {
  K3M_REQUIRE(B > 0 && Lv1 >= 2 && ld >= T, "kl: bad shape");
  kl_kernel<<<B * Lv1, 256, 0, ST(stream)>>>(logits, ld, target, image_label, count, T, Lv1, grad_scale, loss, (act_t*)dlogits);
  return k3m_check_launch("k3m_kl_fwd_bwd");
}

extern "C" int k3m_itm_loss(const float* pool_t, const float* pool_pv, const float* pool_v, const void* w, const void* b,
                            const long long* l1, const long long* l2, const long long* l3, int B, int H, const void* rng,
                            unsigned int site, float drop_p, float* loss, void* stream) {
  DropSpec d;
  d.rng = (const unsigned long long*)rng;
  d.site = site;
  d.thresh = (rng && drop_p > 0.f) ? (unsigned int)(drop_p * 65536.0f + 0.5f) : 0u;
  d.scale = 1.0f / (1.0f - drop_p);
  itm_kernel<<<B, 128, 0, ST(stream)>>>(pool_t, pool_pv, pool_v, (const act_t*)w, (const act_t*)b, l1, l2, l3, B, H, d, loss);
  return k3m_check_launch("k3m_itm_loss");
}

extern "C" int k3m_pair_cls_loss(const float* logits, int ld, const long long* labels, int B, float grad_scale,
                                 float* loss, float* probs, void* dlogits, int ldd, void* stream) {
  K3M_REQUIRE(B > 0 && ld >= 2, "pair_cls_loss: bad shape");
  pair_cls_kernel<<<1, 128, 0, ST(stream)>>>(logits, ld, labels, B, grad_scale, loss, probs, (act_t*)dlogits, ldd);
  return k3m_check_launch("k3m_pair_cls_loss");
}

extern "C" int k3m_dropout(const void* x, int ldx, int x_f32, void* y, int ldy, int y_f32, int rows, int cols,
                           const void* rng, unsigned int site, float drop_p, void* stream) {
  K3M_REQUIRE(rows > 0 && cols > 0, "dropout: empty");
  DropSpec d;
  d.rng = reinterpret_cast<const unsigned long long*>(rng);
  d.site = site;
  d.thresh = (rng && drop_p > 0.f) ? (unsigned int)(drop_p * 65536.0f + 0.5f) : 0u;
  d.scale = 1.0f / (1.0f - drop_p);
  const int ld_drop = (cols + 7) / 8 * 8;
  size_t n = (size_t)rows * cols;
  int g = (int)((n + 255) / 256);
  if (g > 148 * 8) g = 148 * 8;
  if (x_f32 && y_f32) dropout_kernel<float, float><<<g, 256, 0, ST(stream)>>>((const float*)x, ldx, (float*)y, ldy, rows, cols, ld_drop, d);
  else if (x_f32) dropout_kernel<float, act_t><<<g, 256, 0, ST(stream)>>>((const float*)x, ldx, (act_t*)y, ldy, rows, cols, ld_drop, d);
  else if (y_f32) dropout_kernel<act_t, float><<<g, 256, 0, ST(stream)>>>((const act_t*)x, ldx, (float*)y, ldy, rows, cols, ld_drop, d);
  else dropout_kernel<act_t, act_t><<<g, 256, 0, ST(stream)>>>((const act_t*)x, ldx, (act_t*)y, ldy, rows, cols, ld_drop, d);
  return k3m_check_launch("k3m_dropout");
}

extern "C" int k3m_pair_gather_fwd(const float* c_init, const void* seq, int lds, const int* pair_item, const int* pair_rows,
                                   void* X, float* PV, int P, int H, void* stream) {
  K3M_REQUIRE(P > 0, "pair_gather: no pairs");
  size_t n = (size_t)P * H;
  int g = (int)((n + 255) / 256);
  if (g > 148 * 8) g = 148 * 8;
  pair_gather_fwd_kernel<<<g, 256, 0, ST(stream)>>>(c_init, (const act_t*)seq, lds, pair_item, pair_rows, (act_t*)X, PV, P, H);
  return k3m_check_launch("k3m_pair_gather_fwd");
}

extern "C" int k3m_pair_gather_bwd(const float* dX, const float* dPV, const int* pair_item, const int* pair_rows,
                                   float* dc_init, float* dseq32, int P, int H, void* stream) {
  size_t n = (size_t)P * H;
  int g = (int)((n + 255) / 256);
  if (g > 148 * 8) g = 148 * 8;
  pair_gather_bwd_kernel<<<g, 256, 0, ST(stream)>>>(dX, dPV, pair_item, pair_rows, dc_init, dseq32, P, H);
  return k3m_check_launch("k3m_pair_gather_bwd");
}

extern "C" int k3m_struct_fallback_fwd(float* t, const float* c_init, const int* fb, int n, int H, void* stream) {
  K3M_REQUIRE(n > 0 && H > 0, "struct_fallback: empty");
  struct_fallback_fwd_kernel<<<n, 256, 0, ST(stream)>>>(t, c_init, fb, H);
  return k3m_check_launch("k3m_struct_fallback_fwd");
}

extern "C" int k3m_struct_fallback_bwd(float* dt, float* dc_init, const int* fb, int n, int H, void* stream) {
  K3M_REQUIRE(n > 0 && H > 0, "struct_fallback: empty");
  struct_fallback_bwd_kernel<<<n, 256, 0, ST(stream)>>>(dt, dc_init, fb, H);
  return k3m_check_launch("k3m_struct_fallback_bwd");
}

extern "C" int k3m_struct_attn_fwd(const float* t, const void* w2, const void* b2, const int* seg, float* s_out, float* att,
                                   int B, int maxn, int H, void* stream) {
  K3M_REQUIRE(maxn > 0 && maxn <= 8192, "struct_attn: bad maxn");
  struct_attn_fwd_kernel<<<B, 256, maxn * sizeof(float), ST(stream)>>>(t, (const act_t*)w2, (const act_t*)b2, seg, s_out, att, maxn, H);
  return k3m_check_launch("k3m_struct_attn_fwd");
}

extern "C" int k3m_struct_attn_bwd(const float* ds, const float* t, const float* att, const void* w2, const int* seg,
                                   float* dt, float* dw2, float* db2, int B, int maxn, int H, void* stream) {
  struct_attn_bwd_kernel<<<B, 256, 2 * maxn * sizeof(float), ST(stream)>>>(ds, t, att, (const act_t*)w2, seg, dt, dw2, db2, maxn, H);
  return k3m_check_launch("k3m_struct_attn_bwd");
}

extern "C" int k3m_lpm_fwd_bwd(const float* c_final, const float* PV, const int* terms, int T, const int* n_valid,
                               int H, float margin, float* loss, float* dc_final, float* dPV, void* stream) {
  if (T <= 0) return K3M_OK;
  lpm_kernel<<<T, 128, 0, ST(stream)>>>(c_final, PV, terms, T, n_valid, H, margin, loss, dc_final, dPV);
  return k3m_check_launch("k3m_lpm_fwd_bwd");
}

extern "C" int k3m_adamw_bg(float* p32, float* g32, float* m, float* v, void* p16, const float* hyper, long long n,
                            int zero_grad, void* stream) {
  K3M_REQUIRE(n % 4 == 0, "adamw: range length must be a multiple of 4");
  K3M_REQUIRE((((uintptr_t)p32 | (uintptr_t)g32 | (uintptr_t)m | (uintptr_t)v) & 15) == 0 && (((uintptr_t)p16) & 7) == 0,
              "adamw: misaligned range");
  const size_t n4 = (size_t)n / 4;
  size_t g = (n4 + 255) / 256;
  static const int per_sm = getenv("K3M_ADAMW_BG_CTAS") ? atoi(getenv("K3M_ADAMW_BG_CTAS")) : 1;
  if (g > (size_t)148 * per_sm) g = (size_t)148 * per_sm;
  adamw_bg_kernel<<<(int)g, 256, 0, ST(stream)>>>(p32, g32, m, v, (bf16*)p16, hyper, n4, zero_grad);
  return k3m_check_launch("k3m_adamw_bg");
}

extern "C" int k3m_adamw(float* p32, float* g32, float* m, float* v, void* p16, const float* hyper, long long n,
                         int zero_grad, void* stream) {
  K3M_REQUIRE(n % 4 == 0, "adamw: range length must be a multiple of 4");
  K3M_REQUIRE((((uintptr_t)p32 | (uintptr_t)g32 | (uintptr_t)m | (uintptr_t)v) & 15) == 0 && (((uintptr_t)p16) & 7) == 0,
              "adamw: misaligned range");
  const size_t n4 = (size_t)n / 4;
  size_t g = (n4 + 255) / 256;
  if (g > 148 * 16) g = 148 * 16;
  adamw_kernel<false><<<(int)g, 256, 0, ST(stream)>>>(p32, g32, m, v, (bf16*)p16, hyper, n4, zero_grad, nullptr);
  return k3m_check_launch("k3m_adamw");
}

// gradient in act_t (`g16`, e.g. the all-reduced send buffer of a data-parallel step); `g32` is the fp32 arena
// range to zero for the next step
extern "C" int k3m_adamw_g16(float* p32, const void* g16, float* g32, float* m, float* v, void* p16, const float* hyper,
                             long long n, int zero_grad, void* stream) {
  K3M_REQUIRE(n % 4 == 0, "adamw: range length must be a multiple of 4");
  K3M_REQUIRE((((uintptr_t)p32 | (uintptr_t)g32 | (uintptr_t)m | (uintptr_t)v) & 15) == 0 && (((uintptr_t)p16 | (uintptr_t)g16) & 7) == 0,
              "adamw: misaligned range");
  const size_t n4 = (size_t)n / 4;
  size_t g = (n4 + 255) / 256;
  if (g > 148 * 16) g = 148 * 16;
  adamw_kernel<true><<<(int)g, 256, 0, ST(stream)>>>(p32, g32, m, v, (bf16*)p16, hyper, n4, zero_grad, (const bf16*)g16);
  return k3m_check_launch("k3m_adamw_g16");
}
