// GPU-side batch preprocessing (SURVEY §8 f, row N2): what `BertPreprocessBatch` and the loader's
// `__iter__` do per sample in Python (`vilbert_k3m/datasets/concept_cap_dataset_struc.py:532-933`, `:372-413`)
// as four small kernels over a padded raw batch that is already in HBM.  Integer / byte work, HBM-bound:
// nothing here goes near the tensor cores.  The random decisions are INPUTS (one double per title token and
// per region, as `random.random()` hands them out), so results are bit-identical to the reference given
// the same draws; in production the draws come from a device generator.
//
// Floating point follows numpy's float32 evaluation order with explicit round-to-nearest intrinsics (no
// FMA contraction), and float64 where numpy promotes (`sum / int64 count`, the `prob / 0.15` chain).
#include "common.cuh"

namespace {

constexpr int PV_COLON = 131, PV_SEMI = 132;   // ":" and ";" of the PV string (`:785-790`)

// ---- title: mask_word (`:763-783`) + specials + padding (`:664-690`) ------------------------------------
__device__ void prep_title_body(int b, const long long* __restrict__ tokens, const int* __restrict__ n_tok,
                                  const double* __restrict__ u, const long long* __restrict__ rnd, int L,
                                  long long cls_id, long long sep_id, long long mask_id,
                                  long long* __restrict__ ids, long long* __restrict__ att,
                                  long long* __restrict__ seg, long long* __restrict__ labels) {
  const int n = n_tok[b] < L - 2 ? n_tok[b] : L - 2;
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    long long id = 0, lab = -1, m = 0;
    if (i == 0) { id = cls_id; m = 1; }
    else if (i <= n) {
      const long long t = tokens[(size_t)b * L + i - 1];
      double prob = u[(size_t)b * L + i - 1];
      id = t; m = 1;
      if (prob < 0.15) {
        prob /= 0.15;
        if (prob < 0.8) id = mask_id;
        else if (prob < 0.9) id = rnd[(size_t)b * L + i - 1];
        lab = t;
      }
    } else if (i == n + 1) { id = sep_id; m = 1; }
    ids[(size_t)b * L + i] = id;
    att[(size_t)b * L + i] = m;
    seg[(size_t)b * L + i] = 0;
    labels[(size_t)b * L + i] = lab;
  }
}

// ---- PV: mask_word_pv (`:815-839`) + specials + index_pv (`:785-813`) -----------------------------------
// The pairing rules are sequential list surgery (drop-first, zip to the shorter list, early stop), so one
// thread walks the row; the row is at most a few hundred tokens and the kernel runs B rows in parallel.
__device__ void prep_pv_body(int b, int* sh, const long long* __restrict__ tokens, const int* __restrict__ n_tok, int L,
                               int max_num_pv, long long cls_id, long long sep_id, long long mask_id,
                               long long* __restrict__ ids, long long* __restrict__ att, long long* __restrict__ seg,
                               long long* __restrict__ labels, long long* __restrict__ index_p,
                               long long* __restrict__ index_v) {
  int* p131 = sh;                             // [L] positions of 131, [L + 1] positions of 132
  int* p132 = sh + L;
  const int n = n_tok[b] < L - 2 ? n_tok[b] : L - 2;
  const long long* tk = tokens + (size_t)b * L;
  long long* o_ids = ids + (size_t)b * L;
  long long* o_lab = labels + (size_t)b * L;
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    long long id = 0, m = 0;
    if (i == 0) { id = cls_id; m = 1; }
    else if (i <= n) { id = tk[i - 1]; m = 1; }
    else if (i == n + 1) { id = sep_id; m = 1; }
    o_ids[i] = id;
    att[(size_t)b * L + i] = m;
    seg[(size_t)b * L + i] = 0;
    o_lab[i] = -1;
  }
  for (int i = threadIdx.x; i < 2 * max_num_pv; i += blockDim.x) {
    index_p[(size_t)b * 2 * max_num_pv + i] = 0;
    index_v[(size_t)b * 2 * max_num_pv + i] = 0;
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  int n131 = 0, n132 = 0;
  for (int i = 0; i < n; ++i) {               // positions in the list WITHOUT specials
    if (tk[i] == PV_COLON) p131[n131++] = i;
    if (tk[i] == PV_SEMI) p132[n132++] = i;
  }
  // mask_word_pv: every value span except the first pair's becomes [MASK]; the label is the token the
  // span held WHEN it was visited (a second visit of an overlapping span records the mask id, as the
  // reference's in-place edit does)
  if (n132 == n131 - 1) p132[n132++] = n;
  const int first = n132 > 1 ? 1 : 0;
  int pairs = n131 < n132 ? n131 : n132;
  for (int k = first; k < pairs; ++k)
    for (int i = p131[k] + 1; i < p132[k]; ++i) {
      o_lab[i + 1] = o_ids[i + 1];
      o_ids[i + 1] = mask_id;
    }
  // index_pv reads the row AFTER masking and WITH the specials (`:671-674`): a malformed string whose
  // span swallowed a separator changes what it finds
  n131 = n132 = 0;
  for (int i = 0; i < n + 2; ++i) {
    if (o_ids[i] == PV_COLON) p131[n131++] = i;
    if (o_ids[i] == PV_SEMI) p132[n132++] = i;
  }
  if (n132 == n131) {}
  else if (n132 == n131 - 1) n131 -= 1;
  else n131 = n132 = 0;
  pairs = n131 < n132 ? n131 : n132;
  int pv_begin = 1, cnt = 0;
  for (int k = 0; k < pairs && cnt < max_num_pv; ++k, ++cnt) {
    const int i131 = p131[k], i132 = p132[k];
    index_p[((size_t)b * max_num_pv + cnt) * 2 + 0] = pv_begin;
    index_p[((size_t)b * max_num_pv + cnt) * 2 + 1] = i131;
    index_v[((size_t)b * max_num_pv + cnt) * 2 + 0] = i131 + 1;
    index_v[((size_t)b * max_num_pv + cnt) * 2 + 1] = i132;
    pv_begin = i132 + 1;
  }
}

// ---- regions: image_loc (`:585-596`), iou (`:40-78`), mask_region (`:887-933`), loader prepend (`:388-397`)
__device__ __forceinline__ float box_area(const float* bx) {
  return __fmul_rn(__fadd_rn(__fsub_rn(bx[2], bx[0]), 1.0f), __fadd_rn(__fsub_rn(bx[3], bx[1]), 1.0f));
}

__device__ void prep_regions_body(int b, float* shf, const float* __restrict__ boxes, const int* __restrict__ num_boxes,
                                    const float* __restrict__ whp, const double* __restrict__ u, int R,
                                    float* __restrict__ image_loc, long long* __restrict__ image_mask,
                                    long long* __restrict__ image_label, unsigned char* __restrict__ zero_flag,
                                    int* __restrict__ count0) {
  float* bx = shf;                            // [R][4] boxes, then int sel[R], masked[R]
  int* sel = reinterpret_cast<int*>(shf + 4 * R);
  int* msk = sel + R;
  __shared__ int cnt;
  const int nb = num_boxes[b];
  const float w = whp[b * 3 + 0], h = whp[b * 3 + 1], wh = whp[b * 3 + 2];
  if (threadIdx.x == 0) cnt = 0;
  for (int r = threadIdx.x; r < R; r += blockDim.x) {
    int s = 0, z = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) bx[r * 4 + k] = r < nb ? boxes[((size_t)b * R + r) * 4 + k] : 0.f;
    if (r < nb) {
      double prob = u[(size_t)b * R + r];
      if (prob < 0.15) {
        prob /= 0.15;
        s = 1;
        z = prob < 0.9;
      }
    }
    sel[r] = s;
    zero_flag[(size_t)b * R + r] = (unsigned char)z;
    image_label[(size_t)b * R + r] = s ? 1 : -1;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < R; j += blockDim.x) {
    int m = 0;
    if (j < nb) {
      const float aj = box_area(bx + 4 * j);
      for (int i = 0; i < nb; ++i) {
        if (!sel[i]) continue;
        float iw = __fadd_rn(__fsub_rn(fminf(bx[4 * i + 2], bx[4 * j + 2]), fmaxf(bx[4 * i + 0], bx[4 * j + 0])), 1.0f);
        float ih = __fadd_rn(__fsub_rn(fminf(bx[4 * i + 3], bx[4 * j + 3]), fmaxf(bx[4 * i + 1], bx[4 * j + 1])), 1.0f);
        if (iw < 0.f) iw = 0.f;
        if (ih < 0.f) ih = 0.f;
        const float inter = __fmul_rn(iw, ih);
        const float ua = __fsub_rn(__fadd_rn(box_area(bx + 4 * i), aj), inter);
        if (__fdiv_rn(inter, ua) > 0.4f) m = 1;
      }
    }
    msk[j] = m;
    // location row j + 1: area fraction from the RAW corners first, then the corners scaled (`:586-595`)
    const float x1 = bx[4 * j], y1 = bx[4 * j + 1], x2 = bx[4 * j + 2], y2 = bx[4 * j + 3];
    float* o = image_loc + ((size_t)b * (R + 1) + j + 1) * 5;
    o[4] = __fdiv_rn(__fmul_rn(__fsub_rn(y2, y1), __fsub_rn(x2, x1)), wh);
    o[0] = __fdiv_rn(x1, w); o[1] = __fdiv_rn(y1, h); o[2] = __fdiv_rn(x2, w); o[3] = __fdiv_rn(y2, h);
    image_mask[(size_t)b * (R + 1) + j + 1] = j < nb ? 1 : 0;
  }
  __syncthreads();
  int local = 0;
  for (int j = threadIdx.x; j < R; j += blockDim.x) local += msk[j] == 0;
  atomicAdd(&cnt, local);
  __syncthreads();
  if (threadIdx.x == 0) {
    count0[b] = cnt > 0 ? cnt : 1;            // `sum_count[sum_count == 0] = 1` (`:384-385`)
    float* g = image_loc + (size_t)b * (R + 1) * 5;
    g[0] = 0.f; g[1] = 0.f; g[2] = 1.f; g[3] = 1.f; g[4] = 1.f;
    image_mask[(size_t)b * (R + 1)] = 1;
  }
}

// One launch for the three row-wise jobs (blockIdx.y + job0: 0 title, 1 PV string, 2 regions): they are
// microseconds of work each, so a launch apiece was most of their time.
struct PrepArgs {
  const long long *tokens, *rnd, *tokens_pv;
  const int *n_tok, *n_tok_pv, *num_boxes;
  const double *u_tok, *u_reg;
  const float *boxes, *whp;
  long long *ids, *att, *seg, *labels, *ids_pv, *att_pv, *seg_pv, *labels_pv, *index_p, *index_v;
  float* image_loc;
  long long *image_mask, *image_label;
  unsigned char* zero_flag;
  int* count0;
  long long cls_id, sep_id, mask_id;
  int Lt, Lpv, R, max_num_pv;
};

__global__ void __launch_bounds__(128) prep_rows_kernel(PrepArgs a, int job0) {
  extern __shared__ __align__(16) unsigned char dyn[];
  const int b = blockIdx.x, job = job0 + blockIdx.y;
  if (job == 0)
    prep_title_body(b, a.tokens, a.n_tok, a.u_tok, a.rnd, a.Lt, a.cls_id, a.sep_id, a.mask_id, a.ids, a.att, a.seg,
                    a.labels);
  else if (job == 1)
    prep_pv_body(b, reinterpret_cast<int*>(dyn), a.tokens_pv, a.n_tok_pv, a.Lpv, a.max_num_pv, a.cls_id, a.sep_id,
                 a.mask_id, a.ids_pv, a.att_pv, a.seg_pv, a.labels_pv, a.index_p, a.index_v);
  else
    prep_regions_body(b, reinterpret_cast<float*>(dyn), a.boxes, a.num_boxes, a.whp, a.u_reg, a.R, a.image_loc,
                      a.image_mask, a.image_label, a.zero_flag, a.count0);
}

// features: zero the selected regions (`:906-907`), sum over regions in order (numpy reduces a
// non-contiguous axis slice by slice), global feature = float32(float64 sum / int64 count) in front
__global__ void prep_feat_kernel(const float* __restrict__ feat, const unsigned char* __restrict__ zero_flag,
                                 const int* __restrict__ count0, int R, int F, float* __restrict__ out) {
  const int b = blockIdx.y, c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= F) return;
  float s = 0.f;
  for (int r = 0; r < R; ++r) {
    const float v = zero_flag[(size_t)b * R + r] ? 0.f : feat[((size_t)b * R + r) * F + c];
    out[((size_t)b * (R + 1) + r + 1) * F + c] = v;
    s = __fadd_rn(s, v);
  }
  out[(size_t)b * (R + 1) * F + c] = (float)((double)s / (double)count0[b]);
}

// same, 4 columns per thread through 16-byte accesses and 8 rows of loads in flight before the (ordered) adds
__global__ void __launch_bounds__(128) prep_feat4_kernel(const float4* __restrict__ feat,
                                                         const unsigned char* __restrict__ zero_flag,
                                                         const int* __restrict__ count0, int R, int F4,
                                                         float4* __restrict__ out) {
  const int b = blockIdx.y, c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= F4) return;
  const unsigned char* zf = zero_flag + (size_t)b * R;
  const float4* src = feat + (size_t)b * R * F4 + c;
  float4* dst = out + ((size_t)b * (R + 1) + 1) * F4 + c;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int r0 = 0; r0 < R; r0 += 8) {
    float4 v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r0 + k < R && !zf[r0 + k]) v[k] = __ldcs(src + (size_t)(r0 + k) * F4);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (r0 + k < R) {
        __stcs(dst + (size_t)(r0 + k) * F4, v[k]);
        s.x = __fadd_rn(s.x, v[k].x); s.y = __fadd_rn(s.y, v[k].y);
        s.z = __fadd_rn(s.z, v[k].z); s.w = __fadd_rn(s.w, v[k].w);
      }
    }
  }
  const double n = (double)count0[b];
  out[(size_t)b * (R + 1) * F4 + c] = make_float4((float)((double)s.x / n), (float)((double)s.y / n),
                                                  (float)((double)s.z / n), (float)((double)s.w / n));
}

int launch_feat(const float* feat, const unsigned char* zero_flag, const int* count0, int B, int R, int F,
                float* image_feat, cudaStream_t st) {
  if (F % 4 == 0 && (((uintptr_t)feat | (uintptr_t)image_feat) & 15) == 0)
    prep_feat4_kernel<<<dim3((F / 4 + 127) / 128, B), 128, 0, st>>>(reinterpret_cast<const float4*>(feat), zero_flag,
                                                                    count0, R, F / 4,
                                                                    reinterpret_cast<float4*>(image_feat));
  else
    prep_feat_kernel<<<dim3((F + 255) / 256, B), 256, 0, st>>>(feat, zero_flag, count0, R, F, image_feat);
  return k3m_check_launch("k3m_prep_regions(features)");
}

size_t rows_smem(int Lpv, int R) {
  const size_t a = (size_t)(2 * Lpv + 1) * sizeof(int), b = (size_t)R * (4 * sizeof(float) + 2 * sizeof(int));
  return a > b ? a : b;
}

}  // namespace

extern "C" int k3m_prep_title(const long long* tokens, const int* n_tok, const double* u, const long long* rnd, int B,
                              int L, long long cls_id, long long sep_id, long long mask_id, long long* ids,
                              long long* att, long long* seg, long long* labels, void* stream) {
  K3M_REQUIRE(B > 0 && L >= 2, "prep_title: empty batch");
  PrepArgs a = {};
  a.tokens = tokens; a.n_tok = n_tok; a.u_tok = u; a.rnd = rnd; a.Lt = L;
  a.cls_id = cls_id; a.sep_id = sep_id; a.mask_id = mask_id;
  a.ids = ids; a.att = att; a.seg = seg; a.labels = labels;
  prep_rows_kernel<<<dim3(B, 1), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a, 0);
  return k3m_check_launch("k3m_prep_title");
}

extern "C" int k3m_prep_pv(const long long* tokens, const int* n_tok, int B, int L, int max_num_pv, long long cls_id,
                           long long sep_id, long long mask_id, long long* ids, long long* att, long long* seg,
                           long long* labels, long long* index_p, long long* index_v, void* stream) {
  K3M_REQUIRE(B > 0 && L >= 2 && max_num_pv > 0, "prep_pv: empty batch");
  const size_t smem = rows_smem(L, 0);
  K3M_REQUIRE(smem <= 48 * 1024, "prep_pv: sequence too long (L=%d)", L);
  PrepArgs a = {};
  a.tokens_pv = tokens; a.n_tok_pv = n_tok; a.Lpv = L; a.max_num_pv = max_num_pv;
  a.cls_id = cls_id; a.sep_id = sep_id; a.mask_id = mask_id;
  a.ids_pv = ids; a.att_pv = att; a.seg_pv = seg; a.labels_pv = labels; a.index_p = index_p; a.index_v = index_v;
  prep_rows_kernel<<<dim3(B, 1), 128, smem, reinterpret_cast<cudaStream_t>(stream)>>>(a, 1);
  return k3m_check_launch("k3m_prep_pv");
}

extern "C" int k3m_prep_regions(const float* boxes, const int* num_boxes, const float* whp, const double* u,
                                const float* feat, int B, int R, int F, float* image_feat, float* image_loc,
                                long long* image_mask, long long* image_label, unsigned char* zero_flag, int* count0,
                                void* stream) {
  K3M_REQUIRE(B > 0 && R > 0 && F > 0, "prep_regions: empty batch");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t smem = rows_smem(0, R);
  K3M_REQUIRE(smem <= 48 * 1024, "prep_regions: too many regions (R=%d)", R);
  PrepArgs a = {};
  a.boxes = boxes; a.num_boxes = num_boxes; a.whp = whp; a.u_reg = u; a.R = R;
  a.image_loc = image_loc; a.image_mask = image_mask; a.image_label = image_label;
  a.zero_flag = zero_flag; a.count0 = count0;
  prep_rows_kernel<<<dim3(B, 1), 128, smem, st>>>(a, 2);
  int rc = k3m_check_launch("k3m_prep_regions");
  if (rc) return rc;
  return launch_feat(feat, zero_flag, count0, B, R, F, image_feat, st);
}

extern "C" int k3m_prep_batch(const long long* tokens, const int* n_tok, const double* u_tok, const long long* rnd_tok,
                              const long long* tokens_pv, const int* n_tok_pv, const float* boxes,
                              const int* num_boxes, const float* whp, const double* u_reg, const float* feat, int B,
                              int Lt, int Lpv, int R, int F, int max_num_pv, long long cls_id, long long sep_id,
                              long long mask_id, long long* ids, long long* att, long long* seg, long long* labels,
                              long long* ids_pv, long long* att_pv, long long* seg_pv, long long* labels_pv,
                              long long* index_p, long long* index_v, float* image_feat, float* image_loc,
                              long long* image_mask, long long* image_label, unsigned char* zero_flag, int* count0,
                              void* stream) {
  K3M_REQUIRE(B > 0 && Lt >= 2 && Lpv >= 2 && R > 0 && F > 0 && max_num_pv > 0, "prep_batch: empty batch");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t smem = rows_smem(Lpv, R);
  K3M_REQUIRE(smem <= 48 * 1024, "prep_batch: sequence / region count too large (Lpv=%d R=%d)", Lpv, R);
  PrepArgs a = {};
  a.tokens = tokens; a.n_tok = n_tok; a.u_tok = u_tok; a.rnd = rnd_tok; a.Lt = Lt;
  a.tokens_pv = tokens_pv; a.n_tok_pv = n_tok_pv; a.Lpv = Lpv; a.max_num_pv = max_num_pv;
  a.boxes = boxes; a.num_boxes = num_boxes; a.whp = whp; a.u_reg = u_reg; a.R = R;
  a.cls_id = cls_id; a.sep_id = sep_id; a.mask_id = mask_id;
  a.ids = ids; a.att = att; a.seg = seg; a.labels = labels;
  a.ids_pv = ids_pv; a.att_pv = att_pv; a.seg_pv = seg_pv; a.labels_pv = labels_pv;
  a.index_p = index_p; a.index_v = index_v;
  a.image_loc = image_loc; a.image_mask = image_mask; a.image_label = image_label;
  a.zero_flag = zero_flag; a.count0 = count0;
  prep_rows_kernel<<<dim3(B, 3), 128, smem, st>>>(a, 0);
  int rc = k3m_check_launch("k3m_prep_batch");
  if (rc) return rc;
  return launch_feat(feat, zero_flag, count0, B, R, F, image_feat, st);
}

// ---- k3m_prep_regions without any allocation on the caller's side ------------------------------------------------
// For callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): the
// two scratch ranges of k3m_prep_regions (zero_flag [B, R] bytes, count0 [B] ints) live in a file-scope device array
// instead of caller memory.  Every GPU that loads the library holds K3M_PREP_SCRATCH_BYTES (1 MiB + 16 KiB) until
// exit.  Layout: bytes [0, K3M_PREP_MAX_BR) zero_flag, then K3M_PREP_MAX_B ints count0.
// The rows pass writes zero_flag[b, r] for EVERY r < R and count0[b] for every b before the feature pass (next launch
// on the same stream) reads them, so nothing has to start at zero and nothing is cleared; what the previous call left
// behind is overwritten or never read.  All calls on a device share the array: two calls that overlap on the GPU
// (different streams, no dependency) race on it.
#include <mutex>

#define K3M_PREP_MAX_B 4096
#define K3M_PREP_MAX_BR (1 << 20)
#define K3M_PREP_SCRATCH_BYTES (K3M_PREP_MAX_BR + K3M_PREP_MAX_B * 4)

static __device__ __align__(256) unsigned char k3m_prep_scratch[K3M_PREP_SCRATCH_BYTES];

namespace {
constexpr int MAX_DEVICES = 64;
unsigned char* g_prep_scratch_addr[MAX_DEVICES];     // device address of k3m_prep_scratch, per device ordinal
std::mutex g_prep_scratch_mu;

// In host code `k3m_prep_scratch` is only a handle for the runtime, not a device address: ask the runtime, once per
// device, with that device current, and cache the answer.
int prep_scratch_addr(unsigned char** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) {
    k3m_set_error("k3m_prep_regions_noalloc: cudaGetDevice -> %d (%s)", dev, cudaGetErrorString(e));
    return K3M_ERR_CUDA;
  }
  std::lock_guard<std::mutex> lock(g_prep_scratch_mu);
  if (!g_prep_scratch_addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, k3m_prep_scratch);
    if (e != cudaSuccess || !p) {
      k3m_set_error("k3m_prep_regions_noalloc: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
      return K3M_ERR_CUDA;
    }
    g_prep_scratch_addr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_prep_scratch_addr[dev];
  return K3M_OK;
}
}  // namespace

extern "C" int k3m_prep_regions_noalloc(const float* boxes, const int* num_boxes, const float* whp, const double* u,
                                        const float* feat, int B, int R, int F, float* image_feat, float* image_loc,
                                        long long* image_mask, long long* image_label, void* stream) {
  K3M_REQUIRE(B > 0 && R > 0 && F > 0, "prep_regions_noalloc: empty batch");
  K3M_REQUIRE(B <= K3M_PREP_MAX_B && (long long)B * R <= K3M_PREP_MAX_BR,
              "prep_regions_noalloc: B=%d R=%d exceeds the static scratch (B <= %d, B*R <= %d); use k3m_prep_regions "
              "with caller scratch", B, R, K3M_PREP_MAX_B, K3M_PREP_MAX_BR);
  const size_t smem = rows_smem(0, R);
  K3M_REQUIRE(smem <= 48 * 1024, "prep_regions_noalloc: too many regions (R=%d)", R);
  unsigned char* scratch = nullptr;
  if (int rc = prep_scratch_addr(&scratch)) return rc;
  unsigned char* zero_flag = scratch;
  int* count0 = reinterpret_cast<int*>(scratch + K3M_PREP_MAX_BR);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  PrepArgs a = {};
  a.boxes = boxes; a.num_boxes = num_boxes; a.whp = whp; a.u_reg = u; a.R = R;
  a.image_loc = image_loc; a.image_mask = image_mask; a.image_label = image_label;
  a.zero_flag = zero_flag; a.count0 = count0;
  prep_rows_kernel<<<dim3(B, 1), 128, smem, st>>>(a, 2);
  if (int rc = k3m_check_launch("k3m_prep_regions_noalloc")) return rc;
  return launch_feat(feat, zero_flag, count0, B, R, F, image_feat, st);
}
