/* libk3m_sm100a.so — C ABI of the B200-native K3M pretraining-step kernels.
 *
 * The reference (sunzeyeah/K3M) has no native code on this path: its "interface" is the Python
 * class API of vilbert_k3m/vilbert_k3m.py, and every arithmetic op is an ATen library call.  Each
 * entry point below therefore names the reference Python lines it replaces.  Conventions:
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer owned by the caller;
 *   - "act" tensors are bf16 (row-major, explicit row stride `ld*` in elements); statistics,
 *     losses and gradients of parameters are fp32;
 *   - asynchronous on `stream` (a cudaStream_t), no allocation, no host synchronisation,
 *     CUDA-graph capturable; re-entrant across streams;
 *   - return 0 on success, <0 on error (K3M_ERR_*); k3m_last_error_string() describes the last
 *     error of the calling thread.  No exceptions cross the ABI.
 *   - dropout: `rng` -> two device uint64 {seed, step}; `site` numbers the dropout application
 *     inside the step; p == 0 or rng == NULL disables it.  Masks are regenerated in backward
 *     from the same (seed, step, site, element index) — Philox4x32-7, 16 bits per element.
 */
#ifndef K3M_H_
#define K3M_H_

#ifdef __cplusplus
extern "C" {
#endif

#define K3M_OK 0
#define K3M_ERR_ARG -1
#define K3M_ERR_CUDA -2
#define K3M_ERR_UNSUPPORTED -3

/* epilogue activations of k3m_gemm_bf16 */
#define K3M_ACT_NONE 0
#define K3M_ACT_GELU 1    /* aux_out (optional) receives the pre-activation; out = gelu(v)   */
#define K3M_ACT_DGELU 2   /* out = v * gelu'(aux_in)                                         */
#define K3M_ACT_SIGMOID 3
#define K3M_ACT_TANH 4
#define K3M_ACT_DTANH 5   /* out = v * (1 - aux_in^2)                                        */
/* BertIntermediate (vilbert_k3m.py:515-518, gelu :119-125) with the derivative computed where Phi(v) already is:
 * forward  out = gelu(v), aux_out = gelu'(v);  backward (autograd of :518 through the FFN-down dgrad)  out = v * aux_in */
#define K3M_ACT_GELU_GRAD 6
#define K3M_ACT_MULAUX 7

/* GEMM operand layouts */
#define K3M_GEMM_NT 0 /* A[M,K] , B[N,K]  : Y = X W^T    (nn.Linear forward)                   */
#define K3M_GEMM_NN 1 /* A[M,K] , B[K,N]  : dX = dY W    (input gradient)                      */
#define K3M_GEMM_TN 2 /* A[K,M] , B[K,N]  : dW = dY^T X  (weight gradient)                     */

int k3m_get_version(void);
/* debug aid: cudaStreamIsCapturing status of `stream` (0 none, 1 active, 2 invalidated, < 0 query error)  */
int k3m_debug_capture_status(void* stream);
const char* k3m_last_error_string(void);

/* tcgen05/TMEM/TMA GEMM with fused epilogue.  Replaces every nn.Linear on the path
 * (vilbert_k3m.py:440-446, 485-489, 515-518, 528-532, 764-783, 986-996, 1795-1799, 1836-1839,
 * 1921-1924, 2153-2161, 2347-2360) together with the elementwise ops the reference runs after
 * it (gelu :125, nn.Dropout, residual add) and their autograd backward.
 *   v = acc (+ bias[n]); activation; dropout(p); v += addend[m,n];
 *   out = accumulate ? out + v (fp32 atomics, split-K) : v.                                     */
int k3m_gemm_bf16(const void* a, int lda, const void* b, int ldb, int layout, int M, int N, int K,
                  void* out, int ldc, int out_f32, int accumulate, const void* bias, int act,
                  const void* aux_in, int ld_aux_in, void* aux_out, int ld_aux_out,
                  const void* addend, int ld_add, int addend_f32, const void* rng,
                  unsigned int drop_site, float drop_p, void* stream);

/* Grouped form: n (1..8) independent problems of ONE layout in one persistent launch — the tile scheduler
 * walks the concatenated tile list (longest-k problems first), so the small M ~ 2304 problems of the three
 * connection layers of a stage (vilbert_k3m.py:1030-1111, called at :1686-1750 once per pass), a text and an
 * image layer, or a stage's weight gradients fill whole waves of CTA pairs together.  Each K3MGemmDesc holds
 * the arguments of k3m_gemm_bf16 (same meaning, same epilogues).                                            */
typedef struct K3MGemmDesc {
  const void* a; int lda; const void* b; int ldb; int layout; int M, N, K;
  void* out; int ldc; int out_f32; int accumulate; const void* bias; int act;
  const void* aux_in; int ld_aux_in; void* aux_out; int ld_aux_out;
  const void* addend; int ld_add; int addend_f32; const void* rng; unsigned int drop_site; float drop_p;
} K3MGemmDesc;
int k3m_gemm_grouped_bf16(const K3MGemmDesc* problems, int n, void* stream);

/* Persistent GEMM grids use at most `n` SMs (0 = all): a data-parallel step leaves a few SMs to the collective's
 * CTAs, which otherwise wait for an SM that a persistent GEMM CTA holds.                                      */
void k3m_gemm_set_sm_limit(int n);

/* Debug aid: a description of the kernel variant the calling thread's last GEMM call dispatched
 * ("bn=256 ctas=2 ew=4 ... epi=F_ADD ...").  Parity tests assert that they ran the benchmark's variant. */
const char* k3m_gemm_last_variant(void);

/* BertLayerNorm (vilbert_k3m.py:319-332): TF style, biased variance, eps inside the sqrt.      */
int k3m_ln_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
               int M, int H, float eps, const void* rng, unsigned int site, float drop_p, void* stream);
/* backward of LN (+ the dropouts around it); dgamma/dbeta/dbias are accumulated (+=).           */
int k3m_ln_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
               void* dx, void* dxm, float* dgamma, float* dbeta, float* dbias, int M, int H,
               const void* rng, unsigned int site_out, float p_out, unsigned int site_in, float p_in,
               void* stream);
/* Grouped forms: n (1..8) independent LayerNorms in one launch (the six streams of a connection stage, a text and
 * an image layer); hidden sizes may differ.  Same arithmetic as k3m_ln_fwd / k3m_ln_bwd per problem.          */
typedef struct K3MLnFwdDesc {
  const void *x, *gamma, *beta; void* y; float *mean, *rstd; int M, H;
  const void* rng; unsigned int site; float drop_p;
} K3MLnFwdDesc;
typedef struct K3MLnBwdDesc {
  const void *dy, *x, *gamma; const float *mean, *rstd; void *dx, *dxm; float *dgamma, *dbeta, *dbias; int M, H;
  const void* rng; unsigned int site_out; float p_out; unsigned int site_in; float p_in;
} K3MLnBwdDesc;
int k3m_ln_fwd_grouped(const K3MLnFwdDesc* problems, int n, float eps, void* stream);
int k3m_ln_bwd_grouped(const K3MLnBwdDesc* problems, int n, void* stream);
int k3m_colsum(const void* x, int ld, int x_f32, float* out, int M, int N, void* stream);
/* grouped form: n (1..8) column sums in one launch; x in the activation type, N % 8 == 0, ld % 8 == 0       */
typedef struct K3MColsumDesc { const void* x; int ld; float* out; int M, N; } K3MColsumDesc;
int k3m_colsum_grouped(const K3MColsumDesc* problems, int n, void* stream);
int k3m_cast(const void* x, int x_f32, void* y, int y_f32, long long n, void* stream);
int k3m_lincomb3(const void* a, int lda, const void* b, int ldb, const void* c, int ldc, float wa,
                 float wb, float wc, int in_f32, void* out, int ldo, int out_f32, int rows, int cols,
                 void* stream);
int k3m_axpy(const void* x, int ldx, int x_f32, void* y, int ldy, int y_f32, float alpha, int rows,
             int cols, void* stream);
int k3m_scale(float* x, float alpha, long long n, void* stream);
/* x[rows, cols] (row stride ld) *= *alpha, alpha a DEVICE fp32 scalar: how the upstream gradient of a
 * loss (autograd's grad_output of `loss.backward()`, pretrain.py:1062) reaches the stored d(loss)/d(logits) */
int k3m_scale_dev(void* x, int ld, int x_f32, const float* alpha, int rows, int cols, void* stream);
int k3m_dgelu_mul(const void* dy, const void* pre, void* dx, long long n, void* stream);
/* (1 - mask) * -10000 (vilbert_k3m.py:2561, 2570, 2580)                                          */
int k3m_mask_additive(const long long* mask, float* out, int n, void* stream);

/* BertSelfAttention.forward (vilbert_k3m.py:449-464) and both directions of
 * BertBiAttention.forward (:786-824): softmax(q k^T / sqrt(d) + mask) -> dropout -> @ v.
 * `lse` is [2, B*H*Lq] fp32: row 0 receives the log-sum-exp of every query row (forward), row 1 is
 * scratch of the backward pass (the per-row sum p . dp when more than 128 x 128 tokens per head
 * make the backward a two-kernel, blocked pass).  Any Lq / Lk for d in {16, 64, 96, 128}.         */
int k3m_attn_fwd(const void* q, int ldq, int q_off, const void* kv, int ldkv, int k_off, int v_off,
                 const float* mask, int B, int H, int Lq, int Lk, int d, void* ctx, int ldo, float* lse,
                 const void* rng, unsigned int site, float drop_p, void* stream);
int k3m_attn_bwd(const void* dctx, int ldo, const void* q, int ldq, int q_off, const void* kv, int ldkv,
                 int k_off, int v_off, const float* mask, float* lse, void* dq, int lddq, void* dkv,
                 int lddkv, int B, int H, int Lq, int Lk, int d, const void* rng, unsigned int site,
                 float drop_p, void* stream);

/* BertEmbeddings.forward gathers (vilbert_k3m.py:361-376); backward skips padding_idx 0 (:343).   */
int k3m_embed_fwd(const long long* ids, const long long* type_ids, const void* word, const void* pos,
                  const void* typ, void* out, int M, int L, int H, void* stream);
int k3m_embed_bwd(const void* dx, const long long* ids, const long long* type_ids, float* dword,
                  float* dpos, float* dtyp, int M, int L, int H, void* stream);
/* image_location_embeddings, a K=5 projection (vilbert_k3m.py:2155)                              */
int k3m_loc_embed_fwd(const float* loc, const void* w, const void* b, void* out, int M, int N, void* stream);
int k3m_loc_embed_bwd(const void* dy, const float* loc, float* dw, float* db, int M, int N, void* stream);

/* pre_sampling_sequence (vilbert_k3m.py:2331-2374): relu + concat, hard Gumbel pick, backward.    */
int k3m_relu_cat3(const void* a, int lda, const void* b, int ldb, const void* c, int ldc, void* out,
                  int M, int H, void* stream);
int k3m_fuse_select_fwd(const void* cat, const float* gates, const float* noise, const void* rng,
                        unsigned int site, void* out, int ldo, int M, int H, void* stream);
int k3m_fuse_select_bwd(const void* dout, int ldd, const void* cat, const float* gates, const float* noise,
                        const void* rng, unsigned int site, void* dgates, void* dcat, int M, int H,
                        void* stream);
int k3m_relu_bwd_split3(const void* dcat, const void* cat, void* da, int lda, void* db, int ldb, void* dc,
                        int ldc, int M, int H, void* stream);
/* pre_sampling_sequence_soft (vilbert_k3m.py:2300-2329), if_pre_sampling = 2: alpha = sigmoid(score GEMM)
 * comes from k3m_gemm_bf16 (ACT_SIGMOID); gated = cat * alpha feeds the soft_* GEMM.  bf16, contiguous,
 * n elements (multiple of 8).  Backward: dcat = dgated * alpha, dscore = dgated * cat * alpha * (1 - alpha). */
int k3m_gate_mul(const void* cat, const void* alpha, void* out, long long n, void* stream);
int k3m_gate_bwd(const void* dgated, const void* cat, const void* alpha, void* dcat, void* dscore,
                 long long n, void* stream);
/* Masked-LM head on the labelled rows only (vilbert_k3m.py:1836-1839 decoder on every position + :2817-2826
 * CrossEntropyLoss(ignore_index=-1): unlabelled rows give neither loss nor gradient).  out[i] = src[idx[i]]
 * (idx < 0: zero row); backward dst[idx[i]] += dsrc[i] (unique indices).  Bit-exact row copies.            */
int k3m_gather_rows(const void* src, int ld, const int* idx, void* out, int n, int H, void* stream);
int k3m_scatter_add_rows(const void* dsrc, const int* idx, void* dst, int ld, int n, int H, void* stream);
/* token-mean pooling over positions 1.. (vilbert_k3m.py:2405-2409)                               */
int k3m_mean_pool_fwd(const void* seq, int ld, void* out, int out_f32, int B, int L, int H, void* stream);
int k3m_mean_pool_bwd(const float* dpool, void* dseq, int ld, int B, int L, int H, void* stream);

/* structure_aggregator (vilbert_k3m.py:2413-2467) and the LPM loss (:2469-2502) over the host plan */
int k3m_pair_gather_fwd(const float* c_init, const void* seq, int lds, const int* pair_item,
                        const int* pair_rows, void* X, float* PV, int P, int H, void* stream);
int k3m_pair_gather_bwd(const float* dX, const float* dPV, const int* pair_item, const int* pair_rows,
                        float* dc_init, float* dseq32, int P, int H, void* stream);
/* Leading items without a property/value pair (NameError -> bare except -> t = c_initial[i],
 * vilbert_k3m.py:2452-2456): fb int32 [n, 2] = (row of t, item) or (-1, -1) per aggregator call.
 * fwd: t[row] = c_init[item].  bwd: dc_init[item] += dt[row]; dt[row] = 0.                        */
int k3m_struct_fallback_fwd(float* t, const float* c_init, const int* fb, int n, int H, void* stream);
int k3m_struct_fallback_bwd(float* dt, float* dc_init, const int* fb, int n, int H, void* stream);
int k3m_struct_attn_fwd(const float* t, const void* w2, const void* b2, const int* seg, float* s_out,
                        float* att, int B, int maxn, int H, void* stream);
int k3m_struct_attn_bwd(const float* ds, const float* t, const float* att, const void* w2, const int* seg,
                        float* dt, float* dw2, float* db2, int B, int maxn, int H, void* stream);
/* `terms` holds T rows; if n_valid (device int, may be NULL) is given only the first *n_valid count
 * (fixed-capacity term lists for CUDA-graph replay).                                              */
int k3m_lpm_fwd_bwd(const float* c_final, const float* PV, const int* terms, int T, const int* n_valid,
                    int H, float margin, float* loss, float* dc_final, float* dPV, void* stream);

/* loss block (vilbert_k3m.py:2739-2832)                                                          */
int k3m_count_labels(const long long* labels, int n, long long value, int equal, float* out, void* stream);
int k3m_xent_fwd_bwd(void* logits, int ld, const long long* labels, const float* count, int M, int V,
                     float grad_scale, float* loss, int write_grad, void* stream);
int k3m_kl_fwd_bwd(const float* logits, int ld, const float* target, const long long* image_label,
                   const float* count, int B, int Lv1, int T, float grad_scale, float* loss, void* dlogits,
                   void* stream);
int k3m_itm_loss(const float* pool_t, const float* pool_pv, const float* pool_v, const void* w, const void* b,
                 const long long* l1, const long long* l2, const long long* l3, int B, int H, const void* rng,
                 unsigned int site, float drop_p, float* loss, void* stream);

/* item-alignment head, "ce" branch of K3MForItemAlignment.forward (vilbert_k3m.py:3443-3450): logits fp32
 * [B, 2] -> probs = softmax [B, 2], loss = mean cross-entropy, dlogits (bf16, row stride ldd) =
 * grad_scale * (probs - onehot) / B.  probs / dlogits may be NULL.                                  */
int k3m_pair_cls_loss(const float* logits, int ld, const long long* labels, int B, float grad_scale,
                      float* loss, float* probs, void* dlogits, int ldd, void* stream);
/* nn.Dropout outside a GEMM epilogue (ClassificationHead.forward, vilbert_k3m.py:2176-2180): y = x *
 * keep * 1/(1-p); element (r, c) uses counter r * round_up(cols, 8) + c of `site`, so a GEMM epilogue
 * with the same site and N = cols regenerates the mask in backward.                                 */
int k3m_dropout(const void* x, int ldx, int x_f32, void* y, int ldy, int y_f32, int rows, int cols,
                const void* rng, unsigned int site, float drop_p, void* stream);

/* torch.optim.AdamW step (pretrain.py:919-924) over a flat range; hyper = device fp32[8]
 * {lr, beta1, beta2, eps, weight_decay, step, grad_scale, -}.                                     */
int k3m_adamw(float* p32, float* g32, float* m, float* v, void* p16, const float* hyper, long long n,
              int zero_grad, void* stream);
/* The same update with the gradient read from a bf16 buffer (the all-reduced send buffer of a data-parallel step,
 * pretrain.py:166 DDP); g32 is the fp32 arena range that is zeroed for the next step.                        */
int k3m_adamw_g16(float* p32, const void* g16, float* g32, float* m, float* v, void* p16, const float* hyper,
                  long long n, int zero_grad, void* stream);
/* Data parallel (pretrain.py:166 DDP all-reduce + :919-924 AdamW), fused over NVLink peer memory: every rank owns
 * 1/world of a gradient bucket.  k3m_cast_zero: fp32 gradient range -> this rank's bf16 send arena, zeroing the
 * fp32 range.  k3m_adamw_dp: for the shard [offset, offset + n) of the arena: sum the bf16 gradients of all ranks
 * (peer-to-peer loads from g16_peers[r], host array of `world` device pointers to each rank's send arena), AdamW on
 * the local fp32 master / moments (p32, m, v point AT the shard), store the new bf16 weights into every rank's
 * weight arena (w16_peers[r]).  The caller synchronises the ranks before (casts visible) and after (weights
 * visible).                                                                                                   */
int k3m_cast_zero(float* g32, void* out16, long long n, void* stream);
int k3m_adamw_dp(float* p32, float* m, float* v, const float* hyper, const void* const* g16_peers,
                 void* const* w16_peers, int world, long long offset, long long n, const void* g16_mc,
                 void* w16_mc, void* stream);
/* g16_mc / w16_mc: multicast (NVLS) addresses of the two symmetric arenas, or NULL: with them the gradient sum is
 * one multimem.ld_reduce (added inside NVSwitch) and the weight broadcast one multimem.st.                     */
/* The same update sized to co-reside with a running GEMM (two 128-thread, <= 32-register CTAs per SM, no shared
 * memory): used for the gradient buckets that become final while backward is still running.               */
int k3m_adamw_bg(float* p32, float* g32, float* m, float* v, void* p16, const float* hyper, long long n,
              int zero_grad, void* stream);

/* ---- batch preprocessing on the GPU: BertPreprocessBatch + the loader's prepend ----------
 * (vilbert_k3m/datasets/concept_cap_dataset_struc.py).  Inputs are a padded raw batch in device memory and the
 * random draws the reference takes from random.random() (doubles: one per title token, one per region) and
 * np.random.randint (replacement ids); given the same draws the outputs equal the reference's bit for bit.
 * All id / mask / label / index outputs are int64 like the tensors the reference's loader yields.
 *
 * title: mask_word :763-783, specials + padding :664-690.  tokens [B, L] ids WITHOUT specials (n_tok[b] <= L-2
 * of them), u / rnd [B, L] indexed like tokens -> ids / att / seg / labels [B, L].                          */
int k3m_prep_title(const long long* tokens, const int* n_tok, const double* u, const long long* rnd, int B,
                   int L, long long cls_id, long long sep_id, long long mask_id, long long* ids, long long* att,
                   long long* seg, long long* labels, void* stream);
/* PV string: mask_word_pv :815-839 (every value span but the first pair's -> [MASK]), then index_pv :785-813
 * on the masked row with specials -> index_p / index_v [B, max_num_pv, 2] (zero padded).                     */
int k3m_prep_pv(const long long* tokens, const int* n_tok, int B, int L, int max_num_pv, long long cls_id,
                long long sep_id, long long mask_id, long long* ids, long long* att, long long* seg,
                long long* labels, long long* index_p, long long* index_v, void* stream);
/* regions: location normalisation :585-596, iou :40-78, mask_region :887-933, global feature / whole-image
 * box / mask prepend :372-413.  boxes [B, R, 4] pixel corners (num_boxes[b] valid), whp [B, 3] = {w, h,
 * float32(double w * double h)}, u [B, R], feat [B, R, F] (rows >= num_boxes zero) -> image_feat [B, R+1, F],
 * image_loc [B, R+1, 5], image_mask [B, R+1], image_label [B, R]; zero_flag [B, R] bytes and count0 [B] ints
 * are scratch (which regions were zeroed; the divisor of the global feature).                               */
int k3m_prep_regions(const float* boxes, const int* num_boxes, const float* whp, const double* u,
                     const float* feat, int B, int R, int F, float* image_feat, float* image_loc,
                     long long* image_mask, long long* image_label, unsigned char* zero_flag, int* count0,
                     void* stream);
/* k3m_prep_regions for callers that cannot allocate.  NO counterpart in the reference: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.
 * Same inputs, outputs and results (bit for bit) as k3m_prep_regions, but the zero_flag / count0 scratch lives in a
 * static device array of the library (1 MiB + 16 KiB per GPU that loads it, held until exit) and the call performs
 * no cudaMalloc / cudaFree / allocating library call.  Limits: B <= 4096 and B * R <= 1048576, beyond that the
 * call returns an argument error (use k3m_prep_regions with caller scratch).  All calls on a device share the
 * array: calls that overlap on the GPU (different streams without a dependency) race; calls on one stream, or
 * ordered by events, are safe.  Nothing needs clearing between calls (every scratch byte read is written first). */
int k3m_prep_regions_noalloc(const float* boxes, const int* num_boxes, const float* whp, const double* u,
                             const float* feat, int B, int R, int F, float* image_feat, float* image_loc,
                             long long* image_mask, long long* image_label, void* stream);

/* all of the above in two launches (title / PV / regions rows in one, the feature pass after it): what
 * k3m_b200/preprocess.py calls per batch.  Arguments as in the three entry points above.                      */
int k3m_prep_batch(const long long* tokens, const int* n_tok, const double* u_tok, const long long* rnd_tok,
                   const long long* tokens_pv, const int* n_tok_pv, const float* boxes, const int* num_boxes,
                   const float* whp, const double* u_reg, const float* feat, int B, int Lt, int Lpv, int R, int F,
                   int max_num_pv, long long cls_id, long long sep_id, long long mask_id, long long* ids,
                   long long* att, long long* seg, long long* labels, long long* ids_pv, long long* att_pv,
                   long long* seg_pv, long long* labels_pv, long long* index_p, long long* index_v,
                   float* image_feat, float* image_loc, long long* image_mask, long long* image_label,
                   unsigned char* zero_flag, int* count0, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* K3M_H_ */
