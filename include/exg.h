/* exg.h — C ABI of libexg.so, the sm_100a CUDA side of exoground_b200.
 *
 * The reference (arjunrs1/ExoGround) is 100% Python/PyTorch and has no FFI; its boundary for this
 * path is the Python signatures of model/tfm_model.py, model/tan_model.py, model/exo_ground_model.py,
 * train/loss.py and train/loss_egoexo4d.py.  The Python mirror in exoground_b200/{model,train}
 * keeps those signatures; every arithmetic step underneath goes through the entry points below
 * (bound with ctypes in exoground_b200/ops/lib.py).  Each entry cites the
 * reference op(s) it replaces.
 *
 * Conventions
 *   - plain pointers and sizes only; all pointers are DEVICE pointers owned by the caller;
 *   - every entry returns 0 on success, <0 for EXG_E* argument errors, >0 for a cudaError_t;
 *     exg_last_error() gives a thread-local message;
 *   - entries only enqueue work on `stream` (a cudaStream_t passed as void*): no allocation,
 *     no synchronisation, safe under CUDA-graph capture;
 *   - dtype tags: EXG_F32 / EXG_BF16.  "T" below means the compute dtype of the call.
 */
#ifndef EXG_H_
#define EXG_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EXG_F32 0
#define EXG_BF16 1

#define EXG_EINVAL (-1)
#define EXG_EUNSUPPORTED (-2)

int exg_version(void);
const char* exg_last_error(void);
/* number of kernels this library has launched since load (process-wide; for bench gpu_launches) */
long long exg_launch_count(void);
/* Reserve n SMs for kernels of other streams (NCCL while a gradient all-reduce overlaps backward): the persistent
 * kernels launched from now on use (SM count - n) CTAs, so none of them waits behind a communication CTA.  Returns
 * the previous reservation; 0 restores full-width grids.  Replaces nothing in the reference (DDP's bucket overlap,
 * train/main_egoexo4d_distributed.py:611-614, relies on non-persistent cuBLAS kernels). */
int exg_reserve_sms(int n);

/* Per-kernel-class timing for bench.py's roofline: between exg_profile_begin() and exg_profile_end() every
 * launch is bracketed by a CUDA-event pair on its stream.  exg_profile_end synchronises those events and
 * fills, per class c < exg_profile_classes(): ms[c] = summed device time, work[c] = summed algorithmic work
 * (FLOPs for gemm / attention classes, bytes for the row-wise / loss / optimizer classes), launches[c]. */
int exg_profile_classes(void);
const char* exg_profile_class_name(int cls);
int exg_profile_begin(void);
int exg_profile_end(double* ms, double* work, long long* launches);

/* Row addressing helper used by the row-wise kernels: logical row r of a [rows, C] operand lives at
 * physical row  (r / group) * group_stride + (r % group) + offset.  {rows,rows,0} == contiguous.
 * Lets kernels read/write the video / text slices of a joint [B, T+N, C] buffer in place
 * (replaces torch.cat / slicing at model/tan_model.py:195-203). */
typedef struct { long group; long group_stride; long offset; } exg_rowmap;

/* ---------------------------------------------------------------- GEMM ------------------------
 * C[b] = epilogue( alpha * op(A[b]) (M x K) * op(B[b]) (K x N) )            for b in [0,batch)
 *   transA=0: A stored [M,K] (K contiguous, row stride lda);  transA=1: A stored [K,M] (M contiguous)
 *   transB=0: B stored [N,K] (K contiguous, row stride ldb)  -- nn.Linear weight layout;
 *   transB=1: B stored [K,N] (N contiguous)
 * Replaces nn.Linear / F.linear / torch.einsum / torch.bmm at model/tfm_model.py:21-27,62-69,
 * model/tan_model.py:42-43,112,132,256,274 and their autograd backward.
 * bf16 inputs with 16-byte aligned operands run on tcgen05 tensor cores (TMA -> smem -> tcgen05.mma
 * -> TMEM -> epilogue); everything else runs on the fp32 SIMT kernel.
 */
#define EXG_EPI_NONE 0       /* C = v                      (v = alpha*acc + bias[n] if bias) */
#define EXG_EPI_QGELU 1      /* aux_out = v ; C = v*sigmoid(1.702 v)   (model/tfm_model.py:11-13) */
#define EXG_EPI_RESID 2      /* C(fp32) = aux(fp32)[m,n] + v           (model/tfm_model.py:36-37) */
#define EXG_EPI_QGELU_BWD 3  /* C = v * dQuickGELU(aux[m,n]), aux = saved pre-activation (T)  */
#define EXG_EPI_RELU 4       /* C = max(v,0)                           (model/vi_encoder.py:217) */
#define EXG_EPI_RELU_BWD 5   /* C = v * (aux[m,n] > 0), aux = forward output (T) */

#define EXG_GEMM_AUTO 0
#define EXG_GEMM_SIMT 1
#define EXG_GEMM_TC 2      /* one tile per CTA, register epilogue (every epilogue, aux operands) */
#define EXG_GEMM_TC2 3     /* persistent, TMEM double-buffered, TMA-store epilogue (NONE/QGELU/RELU) */

typedef struct {
  const void* A; const void* B; void* C;
  int M, N, K;
  long lda, ldb, ldc;
  int transA, transB;
  int batch; long strideA, strideB, strideC;     /* elements */
  int in_dtype;                                  /* dtype of A and B */
  int out_dtype;                                 /* dtype of C (and aux_out) */
  int epilogue;
  const float* bias;                             /* [N] fp32 or NULL */
  const void* aux; long ldaux; long strideAux;   /* see epilogues */
  void* aux_out;                                 /* QGELU pre-activation, same ld/stride as C */
  float alpha;
  int impl;                                      /* EXG_GEMM_AUTO | _SIMT | _TC */
} exg_gemm_params;

int exg_gemm(const exg_gemm_params* p, void* stream);
/* 1 if the tcgen05 path accepts these parameters (alignment / dtype), else 0 */
int exg_gemm_tc_supported(const exg_gemm_params* p);

/* ---------------------------------------------------------------- row-wise kernels ------------
 * LayerNorm forward (nn.LayerNorm eps=1e-5; model/tfm_model.py:22,28,63,70,71; tan_model.py:44-48):
 *   y[map_y(r)] = (x[r]-mean)*rstd*gamma + beta  (+ add[r % add_period] if add)
 * x: x_dtype [rows,C] contiguous;  y: y_dtype;  mean/rstd: fp32 [rows] (may be NULL);
 * y2 (optional second output, y2_dtype, contiguous) receives the same values. */
int exg_layernorm_fwd(const void* x, int x_dtype, const float* gamma, const float* beta,
                      void* y, int y_dtype, exg_rowmap map_y, void* y2, int y2_dtype,
                      const float* add, long add_period,
                      float* mean, float* rstd, long rows, int C, float eps, void* stream);
/* Residual-stream update fused with the following LayerNorm (model/tfm_model.py:36-37 `x = x + attn(...)`,
 * `x = x + mlp(...)` followed by ln_2 / the next block's ln_1):
 *   sum_out[r] = resid[r] + x[r]   (fp32; may alias resid)  ;  y[map_y(r)] = LN(sum_out[r]) * gamma + beta */
int exg_add_layernorm_fwd(const void* x, int x_dtype, const float* resid, float* sum_out,
                          const float* gamma, const float* beta, void* y, int y_dtype, exg_rowmap map_y,
                          long rows, int C, float eps, void* stream);
/* LayerNorm backward.  dy = dy1[map1(r)] (+ dy2[map2(r)] if dy2);
 *   dx[r] = (dres ? dres[r] : 0) + LN'(dy)        written fp32 to dx and (optional) dtype copy dxT
 *   dgamma[C] += sum_r dy*xhat ; dbeta[C] += sum_r dy    (fp32 atomics: caller zero-initialises)
 *   dxsum[C] (optional) += sum_r dx[r]: the bias gradient of the Linear that feeds this residual stream
 *   (out_proj / c_proj of model/tfm_model.py:36-37), produced without another pass over dx */
int exg_layernorm_bwd(const void* dy1, int dy1_dtype, exg_rowmap map1,
                      const void* dy2, int dy2_dtype, exg_rowmap map2,
                      const void* x, int x_dtype, const float* gamma,
                      const float* mean, const float* rstd, const float* dres,
                      float* dx, void* dxT, int dxT_dtype,
                      float* dgamma, float* dbeta, float* dxsum, long rows, int C, void* stream);
/* out[map_out(r)] = x[map_x(r)] + (add ? add[r % add_period] : 0) (+ x2[map_x2(r)] if x2); fp32 in,
 * out_dtype out.  Replaces `video_embed + ln_position_init(pos)` (tan_model.py:161,193) and torch.cat. */
int exg_add_rows(const float* x, exg_rowmap map_x, const float* x2, exg_rowmap map_x2,
                 const float* add, long add_period, void* out, int out_dtype, exg_rowmap map_out,
                 void* out2, int out2_dtype, long rows, int C, void* stream);
/* y[r] = x[map_x(r)] / ||x[map_x(r)]||_2  (NO eps: tan_model.py:110-111); inv_norm[r] saved (fp32). */
int exg_l2norm_fwd(const void* x, int x_dtype, exg_rowmap map_x, void* y, int y_dtype,
                   float* inv_norm, long rows, int C, void* stream);
/* dx[map_dx(r)] = (dy[r] - y[r]*<dy[r],y[r]>) * inv_norm[r]   (+= if accumulate) ; dx dtype dx_dtype */
int exg_l2norm_bwd(const void* dy, int dy_dtype, const void* y, int y_dtype, const float* inv_norm,
                   void* dx, int dx_dtype, exg_rowmap map_dx, int accumulate,
                   long rows, int C, void* stream);
/* out[n] (+)= sum_m x[m,n]  — bias gradients and broadcast-add gradients */
int exg_colsum(const void* x, int x_dtype, long M, long N, long ldx, float* out, int accumulate,
               void* stream);
/* out[n] (+)= sum_m row_w[m] * x[m,n]   (weight gradient of a one-output Linear) ; N, ldx multiples of 4 */
int exg_wcolsum(const void* x, int x_dtype, const float* row_w, long M, long N, long ldx, float* out,
                int accumulate, void* stream);
/* y[r] = <x[r,:], w> (+ bias[0])   (nn.Linear(C, 1): the alignability head, model/tan_model.py:64-68) ; w fp32 [C] */
int exg_rowdot(const void* x, int x_dtype, const float* w, const float* bias, float* y, long rows, int C,
               void* stream);
/* out[r,:] = g[r] * w   (input gradient of exg_rowdot) */
int exg_outer_rows(const float* g, const float* w, void* out, int out_dtype, long rows, int C, void* stream);
/* elementwise dtype conversion of n elements (fp32 master weights -> bf16 shadow copies) */
int exg_cast(const void* src, int src_dtype, void* dst, int dst_dtype, long n, void* stream);
/* mean over `S` stage slabs: out[r] = (1/S) sum_s x[s*stage_stride + r]   (exo_ground_model.py:153-154) */
int exg_stage_mean(const void* x, int x_dtype, long stage_stride, int S, float* out, void* outT,
                   int outT_dtype, long n, void* stream);

/* backward of exg_stage_mean: out[s*stage_stride + i] = alpha * g[i] for s < S (alpha = 1/S) */
int exg_bcast_rows(const float* g, float alpha, void* out, int out_dtype, long stage_stride, int S, long n,
                   void* stream);

/* ---------------------------------------------------------------- attention -------------------
 * softmax(q k^T * scale + key_padding_mask) v per (batch, head); replaces nn.MultiheadAttention's
 * scaled-dot-product core at model/tfm_model.py:32,75,79 (need_weights=False).
 * q: rows b*Sq+i, row stride ldq, head h at column h*d;  k,v likewise with Sk / ldk / ldv.
 * key_pad: uint8 [B,Sk] (1 = ignore key) or NULL.  lse: fp32 [B,H,Sq] (log-sum-exp of scaled scores).
 */
int exg_attn_fwd(const void* q, const void* k, const void* v, int dtype,
                 long ldq, long ldk, long ldv, const uint8_t* key_pad,
                 void* o, long ldo, float* lse, int B, int H, int Sq, int Sk, int d, float scale,
                 void* stream);
/* delta: fp32 workspace of exg_attn_bwd_workspace(B,H,Sq,Sk,d) elements ([B,H,Sq] row statistics, plus the fp32 dQ
 * accumulator of the multi-tile tensor-core backward when Sq or Sk exceeds 128).  dq/dk/dv use the same strides as
 * q/k/v. */
long exg_attn_bwd_workspace(int B, int H, int Sq, int Sk, int d);
int exg_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout,
                 int dtype, long ldq, long ldk, long ldv, long ldo, const uint8_t* key_pad,
                 const float* lse, float* delta, void* dq, void* dk, void* dv,
                 int B, int H, int Sq, int Sk, int d, float scale, void* stream);

/* ---------------------------------------------------------------- MIL-NCE ---------------------
 * Masked bidirectional MIL-NCE over stage-major raw cosine logits L[S, R=B*T, Cc=B*N]
 * (train/loss.py:234-278).  Positives: column (b,k) is positive for row (a,t) iff a==b and
 * tgt[b,k,t]; columns with col_valid==0 (padded sentences, loss.py:244) are skipped.
 * fwd: exp-sums with a fixed shift (|logit*inv_temp| <= inv_temp since |cos|<=1):
 *   row_all[s,r]=sum_c e, row_pos[s,r]=sum_{c pos} e, col_all[s,c]=sum_r e, col_pos[s,c]=sum_{r pos} e,
 *   e = exp(L*inv_temp - inv_temp).  All four outputs are zeroed by the call and accumulated with atomics.
 *   Rows are read with 16-byte loads when B*N is a multiple of 8 (element-wise otherwise).
 * bwd: G[s,r,c] = inv_temp * ( wr[r]*(e/row_all - pos*e/row_pos) + wc[c]*(e/col_all - pos*e/col_pos) )
 *   written in g_dtype; invalid columns get 0.  wr/wc: fp32 per-row / per-column weights (0 = unused).
 * own_row_mask (optional, uint8 [B*T]): rows r = (a,t) with own_row_mask[r] != 0 drop their OWN-clip block
 *   (columns (a,k)) from every sum, i.e. those logits count as -inf.  This reproduces the in-place -6e4 fill of the
 *   padded video seconds that the reference's self-labelling branch leaves in the logits (train/loss.py:100-106).
 * tgt_bits (optional): the same targets packed by exg_nce_pack_targets as [B*T, ceil(N/32)] uint32 rows
 *   (bit n of row (b,t) = tgt[b,n,t]); with it, bf16 logits and B*N % 8 == 0 the calls take the TMA streaming
 *   kernels (nce_tma.cu), otherwise the generic ones.
 * Bc, clip0: column clips and the column clip owned by row clip 0.  Bc == B, clip0 == 0 is the single-GPU square
 *   case above.  With the text features all-gathered over ranks (cross-GPU negatives) a rank holds the row block
 *   L[S, B*T, Bc*N] of the global logits: its row clip a owns column clip a + clip0 (clip0 = rank * B), tgt /
 *   tgt_bits / own_row_mask / wr / row sums stay LOCAL ([B,...]), col_valid / wc / column sums are GLOBAL ([Bc*N]);
 *   the caller sums col_all / col_pos over ranks between fwd and bwd.
 */
int exg_nce_pack_targets(const uint8_t* tgt, uint32_t* tgt_bits, int B, int T, int N, void* stream);
int exg_nce_fwd(const void* logits, int dtype, const uint8_t* tgt, const uint32_t* tgt_bits, const uint8_t* col_valid,
                const uint8_t* own_row_mask, float* row_all, float* row_pos, float* col_all, float* col_pos,
                int S, int B, int T, int N, int Bc, int clip0, float inv_temp, void* stream);
int exg_nce_bwd(const void* logits, int dtype, const uint8_t* tgt, const uint32_t* tgt_bits, const uint8_t* col_valid,
                const uint8_t* own_row_mask, const float* row_all, const float* row_pos, const float* col_all, const float* col_pos,
                const float* wr, const float* wc, void* G, int g_dtype,
                int S, int B, int T, int N, int Bc, int clip0, float inv_temp, void* stream);
/* own_max[b,k] = max_t L[stage, (b,t), (b + clip0,k)]   (train/loss.py:283-286) */
int exg_own_max(const void* logits, int dtype, float* own_max, int stage, int B, int T, int N, int Bc, int clip0,
                void* stream);

/* Scalar tail of the TAN loss (reference train/loss.py:251-360) after exg_nce_fwd, two launches instead of ~300
 * elementwise / reduction / sort launches.
 *   exg_nce_logdiff: rows_x = [2,S,R] (row_all | row_pos), cols_x = [2,S,Cc] as written by exg_nce_fwd;
 *     vr_x[r] = sum_s log(row_all) - log(max(row_pos, 1e-38)),  tc_x[c] = sum_s log(max(col_all, 1e-38)) - log(max(col_pos, 1e-38))
 *   exg_tan_loss_tail (one CTA): tgt_bits = exg_nce_pack_targets of the [B,N,T] targets (already AND-ed with the
 *     valid sentences), valid [B*N],
 *     own_max_x [B*N] = exg_own_max of the last stage, la = alignability logits [B*N] (null without the head),
 *     abs_c = mean absolute text position [B*N] or null.  loss_threshold <= 0 disables the quantile re-weighting
 *     (loss.py:280-307); use_head enables the median-split BCE (loss.py:309-360).
 *     out[0..10] = loss, loss-dual, loss-joint, loss-dual-all, loss-joint-all, loss-total, loss-joint-bce,
 *                  alignability_top1, #rows, #columns behind the weights, #valid sentences.
 *     wr_x [B*T], wc_x [B*N] = d loss / d (row / column log-ratio) for exg_nce_bwd; dla [B*N] = d loss / d la. */
int exg_nce_logdiff(const float* rows_d, const float* cols_d, int E, const float* rows_j, const float* cols_j, int D,
                    int R, int Cc, float* vr_d, float* tc_d, float* vr_j, float* tc_j, void* stream);
size_t exg_tan_loss_tail_workspace(int B, int T, int N);
int exg_tan_loss_tail(const uint32_t* tgt_bits, const uint8_t* valid, const float* vr_d, const float* tc_d,
                      const float* vr_j, const float* tc_j, const float* own_max_d, const float* own_max_j,
                      const float* la, const float* abs_c, int B, int T, int N, int E, int D, float inv_temp,
                      float loss_threshold, int use_head, float nce_weight, float* wr_d, float* wc_d, float* wr_j,
                      float* wc_j, float* dla, float* out, void* ws, size_t ws_bytes, void* stream);
/* Allocation-free variant of exg_tan_loss_tail.  No counterpart in the reference: added
 * for callers that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  Same arguments and results as exg_tan_loss_tail minus (ws, ws_bytes): the scratch lives in a
 * file-scope device array of exg_tan_loss_tail_static_max_bytes() = 1 MiB that every GPU loading the library holds
 * until exit; the call performs no cudaMalloc / cudaFree / allocating library call.  Sizes whose
 * exg_tan_loss_tail_workspace(B,T,N) exceeds that maximum return EXG_EINVAL with a message naming both numbers.  All
 * calls on one device share the array (the kernel writes every element before reading it, so no clearing is needed
 * between calls): calls that overlap on the GPU - different streams without ordering - race. */
size_t exg_tan_loss_tail_static_max_bytes(void);
int exg_tan_loss_tail_static(const uint32_t* tgt_bits, const uint8_t* valid, const float* vr_d, const float* tc_d,
                             const float* vr_j, const float* tc_j, const float* own_max_d, const float* own_max_j,
                             const float* la, const float* abs_c, int B, int T, int N, int E, int D, float inv_temp,
                             float loss_threshold, int use_head, float nce_weight, float* wr_d, float* wc_d,
                             float* wr_j, float* wc_j, float* dla, float* out, void* stream);

/* ---------------------------------------------------------------- ExoGround losses ------------
 * Row-wise softmax cross-entropy with the target on the diagonal: the per-clip T x T InfoNCE of
 * compute_info_nce_loss (model/exo_ground_model.py:206-234, keystep_ground_model.py:221-249,
 * train/loss_egoexo4d.py:325-347) after the normalised-feature GEMM.  Row r has target column r % period.
 *   fwd: lse_rows[r] = logsumexp(logits[r,:]); loss_rows[r] = lse - logits[r, r % period]
 *   bwd: G[r,c] = scale * (softmax(logits[r,:])[c] - [c == r % period]) */
int exg_xent_diag_fwd(const void* logits, int dtype, float* loss_rows, float* lse_rows, long rows, int cols,
                      int period, void* stream);
int exg_xent_diag_bwd(const void* logits, int dtype, const float* lse_rows, float scale, void* G, int g_dtype,
                      long rows, int cols, int period, void* stream);
/* Fused cross-view InfoNCE (train/loss_egoexo4d.py:269-323) and the L1 / cosine monitors (:201-233):
 * per (b,t): o = out_feat[b,t], p = ego[b,pos_idx[b,t],t], n = ego[b,neg_idx[b,t],t],
 * s = same_feat[b,same_idx[b,t]] (optional); logits = cos(o,.) * inv_temp over {p, n, s} ({p, s} if only_same);
 * loss = -log_softmax[0]; l1 = mean |o_hat - p_hat|; pos_cos / neg_cos = cos(o,p) / cos(o,n).
 * stats: fp32 [B,T,8] workspace carried to the backward, which writes d loss / d out_feat scaled by grad_loss[b,t]. */
int exg_xview_nce_fwd(const void* out_feat, int dtype, const float* ego, const long* pos_idx, const long* neg_idx,
                      const float* same_feat, const long* same_idx, int only_same, int B, int V, int T, int D,
                      float inv_temp, float* loss, float* l1, float* pos_cos, float* neg_cos, float* stats,
                      void* stream);
int exg_xview_nce_bwd(const void* out_feat, int dtype, const float* ego, const long* pos_idx, const long* neg_idx,
                      const float* same_feat, const long* same_idx, int only_same, int B, int V, int T, int D,
                      float inv_temp, const float* stats, const float* grad_loss, void* grad_out, int g_dtype,
                      void* stream);
/* Grounding regression loss (train/loss_egoexo4d.py:73-145) in one launch, no host round trip:
 * preds fp32 [M,2]; gt_a / gt_b fp32 [M] (center, duration | start, end); text_pad uint8 [M] (1 = padded).
 * stats (device) = {mean|p0-g0|, mean|p1-g1|, mean IoU, n_valid, count(IoU > thresholds[k])...}; iou [M] (0 for
 * padded rows); dpreds [M,2] = d[(1 - mean IoU) + L1 terms] / d preds.  thresholds is a HOST array (<= 8). */
int exg_ground_loss(const float* preds, const float* gt_a, const float* gt_b, const uint8_t* text_pad, long M,
                    int use_center_duration, float eps, const float* thresholds, int n_thresh, float* stats,
                    float* iou, float* dpreds, void* stream);

/* Out-projection GEMM with the residual add and the LayerNorm that follow it fused into the epilogue
 * (model/tfm_model.py:22,28,34-38: `x = x + attn(..)` / `x = x + mlp(..)` and the LayerNorm the next sub-block opens with):
 *   s = resid + A @ W^T + bias  (fp32 [M, 512]),   y = LayerNorm(s) * gamma + beta  (bf16 [M, 512]).
 * A bf16 [M, K] (lda), W bf16 [512, K] (ldb: nn.Linear layout), bias fp32 [512] or NULL, resid / s_out fp32 [M, 512]
 * contiguous, y_out bf16 [M, 512] contiguous.  One 128 x 512 tcgen05 tile per CTA (the width C = 512 is one TMEM
 * accumulator), thread = row.  exg_gemm_add_ln_supported: 1 if the operands qualify (16-byte aligned rows, K >= 64). */
int exg_gemm_add_ln_supported(const void* A, const void* W, int M, int K, long lda, long ldb);
int exg_gemm_add_ln(const void* A, const void* W, const float* bias, const float* resid, const float* gamma,
                    const float* beta, float eps, float* s_out, void* y_out, int M, int K, long lda, long ldb,
                    void* stream);

/* Sliding-window alignment evaluation, device side (eval/eval_zeroshot_align.py:200-237): stitches the per-window
 * similarity blocks sj / sd [W, T, N] fp32 (joint / dual logits, already divided by 0.07) of one video into
 * sim [K, vlen] = overlap-averaged (joint + dual) / 2 with the reference's `sim == 0 -> -6e4` rule, prob = softmax over
 * time, pred = argmax over time of prob (first index among equals), row_max = max over time of sim.  win int32 [W, 4] =
 * (first second, seconds, first sentence, last sentence) per window; windows are summed in index order (the order of
 * the reference's loop), W <= 1024. */
int exg_align_finish(const float* sj, const float* sd, const int* win, int W, int T, int N, int K, int vlen,
                     float* sim, float* prob, long long* pred, float* row_max, void* stream);

/* Development aid: per-role pipeline event log of the persistent tcgen05 GEMM (csrc/gemm_tc2.cu; read by
 * scripts/gemm_trace.py).  buf: device uint64 [11][cap] or NULL (off); cta: which CTA of each launch records. */
int exg_gemm_trace(void* buf, int cap, int cta);

/* ---------------------------------------------------------------- optimizer -------------------
 * AdamW (torch.optim.AdamW semantics, train/main.py:418) on a flat fp32 segment; optional bf16
 * shadow copy of the updated weights.  grad_scale multiplies the gradient first (1/world_size). */
int exg_adamw(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, void* shadow_bf16,
              long n, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
              float grad_scale, void* stream);
/* Same update with {lr, 1 - beta1^step, sqrt(1 - beta2^step)} read from device memory (hyper[3]) at run time: a launch
 * captured into a CUDA graph then follows the schedule of the step it is replayed for (train/step.py launches the
 * AdamW of a finished transformer stack from inside the captured backward). */
int exg_adamw_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, void* shadow_bf16, long n,
                  const float* hyper, float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                  void* stream);
/* Per-parameter gradient clipping (utils/train_utils.py:3-13 clip_gradients, called at train/main.py:108-110 and
 * train/main_egoexo4d_distributed.py:176-181) on the flat gradient buffer, no host round trip.  Parameters start on
 * 64-element boundaries; blk2seg[b] (int32, device) = parameter index of 64-element block b; n % 64 == 0.
 *   exg_grad_sumsq      : seg_sumsq[s] = sum of squares of parameter s's gradient (zeroes seg_sumsq first);
 *   exg_grad_clip_scale : grad *= clip / (||grad|| * grad_scale + 1e-6) where that factor is < 1 (in place: the
 *                         reference's semantics, for callers that read .grad afterwards);
 *   exg_adamw_clip      : AdamW that applies the same factor to the gradient as it reads it (no extra write pass).
 *                         hyper (device, optional) as in exg_adamw_dev; otherwise lr / step are host values. */
int exg_grad_sumsq(const float* grad, const int* blk2seg, long n, float* seg_sumsq, int n_seg, void* stream);
int exg_grad_clip_scale(float* grad, const int* blk2seg, long n, const float* seg_sumsq, float clip, float grad_scale,
                        void* stream);
int exg_adamw_clip(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, void* shadow_bf16, long n,
                   const float* hyper, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                   float grad_scale, const int* blk2seg, const float* seg_sumsq, float clip, void* stream);
/* target = m*target + (1-m)*online   (TwinTemporalAligner._momentum_update, tan_model.py:334-338) */
int exg_ema(float* target, const float* online, long n, float m, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* EXG_H_ */
