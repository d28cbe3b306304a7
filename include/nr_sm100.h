/* nr_sm100.h — C ABI of libnr_sm100.so: the B200 (sm_100a) kernels behind the
 * NRMS/NAML train-and-eval hot path of patngnw/NewsRecommendation.
 *
 * Boundary contract:
 *  - extern "C", plain pointers and sizes, no torch types.  Every pointer is a
 *    DEVICE pointer unless its name ends in _host.  The caller (the PyTorch
 *    host code in newsrecommendation_b200/) owns every buffer, including the
 *    workspace; the library allocates nothing that outlives a call.
 *  - All work is enqueued on `stream` (a cudaStream_t passed as void*); no
 *    internal synchronisation.  Thread safe: no mutable globals except the
 *    thread-local error string.
 *  - Return 0 = ok; non-zero = error, text via nr_last_error().  Shapes
 *    outside the compiled envelope are an ERROR, never a fallback.
 *  - "reference" citations are file:line under /root/reference/ of the code
 *    each entry point replaces.
 *
 * Row sources.  Several ops read an activation X of logical shape
 * (n, S, D) through an `nr_rows_t`:
 *   mode NR_ROWS_DENSE : X = dense (n*S, D) row-major fp32
 *   mode NR_ROWS_TOKEN : X[i,s,:] = table[ids[i*ids_stride + s], :]      (A1:
 *                        nn.Embedding word lookup, src/model/NRMS.py:28)
 *   mode NR_ROWS_NEWS  : X[i,s,:] = table[ids[i*ids_stride], s*D:(s+1)*D] (A2:
 *                        per-news title row, src/model/NAML.py:47-50)
 * followed, when training with drop_p > 0, by inverted dropout whose keep mask
 * is the counter-based stream (seed, stream id, element (row i*S + s, column d))
 * exported by nr_dropout_keep_mask (F.dropout at NRMS.py:28-30, NAML.py:51-53).
 */
#ifndef NR_SM100_H
#define NR_SM100_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NR_ABI_VERSION 1

enum { NR_ROWS_DENSE = 0, NR_ROWS_TOKEN = 1, NR_ROWS_NEWS = 2 };
/* arithmetic mode of the GEMM-shaped parts: fp32 SIMT (parity mode, rel 1e-5)
 * or bf16 operands on tcgen05 tensor cores with fp32 TMEM accumulation. */
enum { NR_PREC_FP32 = 0, NR_PREC_BF16 = 1 };
/* dropout stream ids */
enum { NR_STREAM_IN = 1, NR_STREAM_OUT = 2 };

typedef struct nr_rows {
    int32_t mode;          /* NR_ROWS_* */
    int32_t ids_stride;    /* int32 elements between consecutive sequences in ids */
    const float* dense;    /* NR_ROWS_DENSE */
    const float* table;    /* NR_ROWS_TOKEN / NR_ROWS_NEWS */
    const int32_t* ids;
    int64_t table_rows;    /* row count of `table` (>= 1).  Enforced on the device: an id outside [0, table_rows)
                            * reads row 0, the padding row, and gets no gradient (the reference's nn.Embedding
                            * raises IndexError; the host drivers validate their id files up front) */
    float drop_p;          /* 0 => no dropout */
    uint64_t seed;
} nr_rows_t;

/* ---- library ------------------------------------------------------------ */
int nr_abi_version(void);
const char* nr_last_error(void);
/* sm count / compute capability of the current device */
int nr_device_info(int32_t* sm_count, int32_t* cc_major, int32_t* cc_minor);

/* number of kernels this library has launched in this process (bench.py: gpu_launches) */
long long nr_launch_count(void);
/* per-kernel CUDA-event timing on the launch stream: enable, run, then report.  The report
 * synchronises the device and writes "name launches total_ms" lines; returns bytes needed. */
int nr_prof_enable(int on);
size_t nr_prof_report(char* buf, size_t cap);

/* Device-resident seed salt: salt_dev points to ONE uint64 in device memory (or NULL to clear).  Every
 * dropout stream is then keyed by seed + *salt_dev, read when the kernel runs.  Lets a train step that
 * was captured in a CUDA graph (host seeds frozen into the launch arguments) draw fresh masks on every
 * replay: bump the value from inside the graph.  nr_dropout_keep_mask applies the salt too. */
int nr_set_seed_salt(const void* salt_dev);

/* *salt_dev += 1 on `stream` (the first node of a captured train step: every replay draws fresh masks) */
int nr_seed_salt_bump(void* salt_dev, void* stream);
/* stream-ordered memset / two-piece concatenation (memset / memcpy nodes when captured): the host code uses them so
 * that a captured step holds no ATen kernels (zero_grad of the flat gradient bucket; the candidate|history id
 * concatenation of NRMS.py:87,90 / NAML.py:122,125 fused into one encoder call) */
int nr_zero_bytes(void* dst, size_t bytes, void* stream);
int nr_concat2(const void* a, size_t bytes_a, const void* b, size_t bytes_b, void* out, void* stream);

/* The keep mask of a dropout stream in the plain (rows x row_width) layout of the activation it is applied to:
 * keep[i] = 1 if element i = (row i / row_width, column i % row_width) is kept at rate p.  Streams are counter based
 * (Philox4x32-7, one call = eight 16-bit uniforms; csrc/nr_common.cuh): the input stream NR_STREAM_IN indexes
 * (row, column) of the gathered (n*S x D) rows (seg_width = row_width = D); the output stream NR_STREAM_OUT indexes
 * (row, head, d) of the (n*S x heads*dk) context (row_width = H, seg_width = dk).  Used by the tests to replay the
 * kernels' own masks through the oracle (F.dropout, src/model/NRMS.py:28-34, NAML.py:51-53). */
int nr_dropout_keep_mask(uint8_t* keep, int64_t count, float p, uint64_t seed, int32_t stream_id,
                         int32_t row_width, int32_t seg_width, void* stream);

/* ---- A1/A2: embedding row gather (bit exact) ---------------------------------
 * out[i,:] = table[ids[i],:], width fp32 elements per row (width % 4 == 0).
 * reference: nn.Embedding lookups src/model/NRMS.py:28, src/model/NAML.py:48;
 * eval gathers src/dataset.py:68,72. */
int nr_gather_rows(const float* table, int64_t table_rows, const int32_t* ids,
                   int64_t count, int32_t width, float* out, void* stream);

/* ---- A6: AttentionPooling (src/model/model_utils.py:7-31) ----------------------
 * x (n,S,H), mask (n,S) or NULL, w1 (Q,H), b1 (Q), w2 (Q), b2 (1) -> out (n,H).
 * e_save (n,S,Q) and alpha_save (n,S) are written for backward. */
size_t nr_attention_pooling_workspace(int64_t n, int32_t S, int32_t H, int32_t Q, int32_t prec);
int nr_attention_pooling_fwd(const float* x, const float* mask, const float* w1, const float* b1,
                             const float* w2, const float* b2, float* out, float* e_save,
                             float* alpha_save, int64_t n, int32_t S, int32_t H, int32_t Q,
                             int32_t prec, void* ws, size_t ws_bytes, void* stream);
/* g (n,H) -> dx (n,S,H) written; dw1,db1,dw2,db2 ACCUMULATED (+=). dx may be NULL. */
int nr_attention_pooling_bwd(const float* g, const float* x, const float* w1, const float* w2,
                             const float* e_save, const float* alpha_save, float* dx, float* dw1,
                             float* db1, float* dw2, float* db2, int64_t n, int32_t S, int32_t H,
                             int32_t Q, int32_t prec, void* ws, size_t ws_bytes, void* stream);

/* ---- A4/A5: MultiHeadSelfAttention with Q=K=V=X (model_utils.py:34-95) ----------
 * X through `rows` (n,S,Dm); key_mask (n,S) or NULL; w* (H,Dm), b* (H), H=heads*dk.
 * ctx (n,S,H) = heads concatenated, after output dropout (stream NR_STREAM_OUT,
 * rate out_drop_p; src/model/NRMS.py:32-34).  qkv_save (n,S,3H) kept for backward. */
size_t nr_mhsa_workspace(int64_t n, int32_t S, int32_t Dm, int32_t H, int32_t prec);
int nr_mhsa_fwd(const nr_rows_t* rows, const float* key_mask, const float* wq, const float* bq,
                const float* wk, const float* bk, const float* wv, const float* bv, float* ctx,
                float* qkv_save, int64_t n, int32_t S, int32_t Dm, int32_t heads, int32_t dk,
                float out_drop_p, int32_t prec, void* ws, size_t ws_bytes, void* stream);
/* dctx (n,S,H) -> weight and bias grads ACCUMULATED; dx_dense (n,S,Dm) written if non-NULL;
 * dtable (same shape as rows.table) ACCUMULATED by scatter-add if non-NULL (token / news
 * rows; row 0 = padding_idx gets no gradient, src/model/NRMS.py:73). */
int nr_mhsa_bwd(const nr_rows_t* rows, const float* key_mask, const float* wq, const float* wk,
                const float* wv, const float* qkv_save, const float* dctx, float* dwq, float* dbq,
                float* dwk, float* dbk, float* dwv, float* dbv, float* dx_dense, float* dtable,
                int64_t n, int32_t S, int32_t Dm, int32_t heads, int32_t dk, float out_drop_p,
                int32_t prec, void* ws, size_t ws_bytes, void* stream);

/* ---- A7: Conv1d(k=3, pad=1) along tokens (src/model/NAML.py:27-32,54) ------------
 * X through `rows` (n,L,D); w (H,D,3) as nn.Conv1d.weight, b (H) -> y (n,L,H). */
size_t nr_conv1d_k3_workspace(int64_t n, int32_t L, int32_t D, int32_t H, int32_t prec);
int nr_conv1d_k3_fwd(const nr_rows_t* rows, const float* w, const float* b, float* y, int64_t n,
                     int32_t L, int32_t D, int32_t H, int32_t prec, void* ws, size_t ws_bytes,
                     void* stream);
/* dy -> dw, db ACCUMULATED; dx_dense (n,L,D) written if non-NULL (dense rows);
 * dtable ACCUMULATED by scatter-add if non-NULL (token / news rows; row 0 skipped). */
int nr_conv1d_k3_bwd(const nr_rows_t* rows, const float* w, const float* dy, float* dw, float* db,
                     float* dx_dense, float* dtable, int64_t n, int32_t L, int32_t D, int32_t H,
                     int32_t prec, void* ws, size_t ws_bytes, void* stream);

/* ---- A8: NAML side view: dense(emb[ids]) (src/model/NAML.py:60-68) ---------------
 * out[i*out_stride : +H] = w (H,E) . emb[ids[i*ids_stride]] + b. */
size_t nr_embed_dense_workspace(int64_t emb_rows, int32_t H);
/* ws (>= nr_embed_dense_workspace bytes, or NULL): lets a small vocabulary (emb_rows <= n/2) go through the
 * transformed table T = emb . w^T + b, the batch then only gathers rows of T */
int nr_embed_dense_fwd(const float* emb, int64_t emb_rows, const int32_t* ids, int32_t ids_stride,
                       const float* w, const float* b, float* out, int64_t out_stride, int64_t n,
                       int32_t E, int32_t H, void* ws, size_t ws_bytes, void* stream);
/* dout (row stride dout_stride) -> dw, db, demb ACCUMULATED (emb row 0 = padding_idx skipped) */
int nr_embed_dense_bwd(const float* emb, int64_t emb_rows, const int32_t* ids, int32_t ids_stride,
                       const float* w, const float* dout, int64_t dout_stride, float* dw, float* db,
                       float* demb, int64_t n, int32_t E, int32_t H, void* ws, size_t ws_bytes,
                       void* stream);

/* ---- A9/A10: pad_doc blend (src/model/NRMS.py:59-60, src/model/NAML.py:94-95) ----
 * out = v*m + pad*(1-m);  v (n,T,H), m (n,T), pad (H). */
int nr_pad_blend_fwd(const float* v, const float* mask, const float* pad, float* out, int64_t n,
                     int32_t T, int32_t H, void* stream);
/* dv written (= dout*m); dpad ACCUMULATED (= sum dout*(1-m)) */
int nr_pad_blend_bwd(const float* dout, const float* mask, float* dv, float* dpad, int64_t n,
                     int32_t T, int32_t H, void* stream);

/* ---- A11: candidate scoring + CrossEntropyLoss (NRMS.py:93-94, NAML.py:128-129) --
 * cand (B,C,H), user (B,H), label (B) int64 -> score (B,C), loss (1) mean CE,
 * hits (1) = #argmax==label (utils.acc, src/utils.py:36-40). */
int nr_score_ce_fwd(const float* cand, const float* user, const int64_t* label, float* score,
                    float* loss, float* hits, float* per_sample_ws, int64_t B, int32_t C,
                    int32_t H, void* stream);
/* Allocation-free variant of nr_score_ce_fwd.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same inputs, outputs and arithmetic as nr_score_ce_fwd, but the (B,2) fp32
 * per-sample scratch lives in a file-scope `static __device__` array inside the library instead of a caller buffer:
 * NR_SCORE_CE_NOALLOC_MAX_B * 2 floats = 128 KB on every GPU that loads the library, held until the process exits.
 * B > NR_SCORE_CE_NOALLOC_MAX_B is an error (use nr_score_ce_fwd).  The call performs no cudaMalloc / cudaFree and
 * calls no allocating library routine.  All calls on a device SHARE the one array: calls that can overlap on the GPU
 * (different streams, no ordering between them) race on it; calls on one stream, or otherwise ordered, are safe.
 * The array keeps the previous call's contents; every element the call reads is written by the call first. */
#define NR_SCORE_CE_NOALLOC_MAX_B 16384
int nr_score_ce_fwd_noalloc(const float* cand, const float* user, const int64_t* label, float* score,
                            float* loss, float* hits, int64_t B, int32_t C, int32_t H, void* stream);
/* gloss (1) upstream grad of loss -> dcand (B,C,H), duser (B,H) written */
int nr_score_ce_bwd(const float* cand, const float* user, const int64_t* label, const float* score,
                    const float* gloss, float* dcand, float* duser, int64_t B, int32_t C, int32_t H,
                    void* stream);

/* ---- A12: fused Adam over a flat fp32 bucket (torch.optim.Adam defaults, main.py:76)
 * g is multiplied by grad_scale first (1/world_size folded in after the all-reduce). */
int nr_adam_step(float* p, const float* g, float* m, float* v, int64_t count, float lr, float beta1,
                 float beta2, float eps, int32_t step, float grad_scale, void* stream);

/* A12 at N > 1 — gradient exchange fused with the optimizer (src/main.py:82,109-110: DDP gradient mean + Adam.step):
 * Adam over the slice [lo, hi) of the flat buckets with the step count read from DEVICE memory (t = *step_dev + 1,
 * int64 count of completed steps, bumped by the caller afterwards with nr_seed_salt_bump: the launch can be replayed
 * from a CUDA graph and several slices of one step can be updated by separate launches).  With the multicast pointers given (both buckets in symmetric
 * memory mapped through an NVLink multicast object), g is read with multimem.ld_reduce (summed over the ranks in the
 * NVSwitch) and the updated parameters are written with multimem.st into every rank's copy: reduce-scatter + Adam +
 * all-gather in ONE kernel.  p / m / v are this rank's local buffers (m, v are only touched inside the slice).  The
 * caller brackets the launch with cross-rank barriers.  g_mc == p_mc == NULL: plain local Adam on g. */
int nr_adam_step_dev(const float* g, const float* g_mc, float* p, float* p_mc, float* m, float* v, int64_t lo,
                     int64_t hi, float lr, float beta1, float beta2, float eps, const void* step_dev,
                     float grad_scale, void* stream);

/* ---- A14: eval impression scoring (src/dataset.py:64-74, src/main.py:253) ---------
 * scores[j] = table[cand_idx[j],:] . user_vecs[imp(j),:], impressions as CSR offsets (I+1). */
int nr_eval_score(const float* table, int64_t table_rows, const float* user_vecs,
                  const int64_t* offsets, const int32_t* cand_idx, float* scores, int64_t I,
                  int64_t total_cands, int32_t H, void* stream);
/* ---- A15: ranking metrics (src/metrics.py:1-23, src/main.py:250-258) ---------------
 * per impression i: metrics[i*4..] = AUC, MRR, nDCG@5, nDCG@10 (fp64); valid[i]=0 for the
 * skipped impressions (labels all 0 or all 1).  Ties ranked higher-index-first. */
int nr_eval_rank_metrics(const float* scores, const int8_t* labels, const int64_t* offsets,
                         double* metrics, uint8_t* valid, int64_t I, void* stream);
/* sums[0..3] = sum over valid impressions, sums[4] = #valid (deterministic order) */
int nr_eval_reduce_metrics(const double* metrics, const uint8_t* valid, double* sums, int64_t I,
                           void* stream);

/* ---- bf16 tensor-core building block (tcgen05.mma, TMA-fed, fp32 accumulation in TMEM) ------
 * C[m,n] = bias[n] + sum_{t<taps} sum_{k<K} A[m + shifts[t], k] * B[n, t*Kpad + k],  Kpad = ceil64(K)
 * A (a_rows x K, leading dim lda) and B (N x taps*Kpad, leading dim ldb) are bf16, K contiguous;
 * rows of A outside [0,a_rows) read as zero.  taps=1: plain x.W^T GEMM (nn.Linear, model_utils.py:21,
 * 89-91); taps=3, shifts {-1,0,1}: Conv1d(k=3,pad=1) (NAML.py:27-32).  C is bf16 (c_bf16=1) or fp32.
 * shifts_host is a HOST pointer. */
int nr_tc_gemm_nt(const void* A, int64_t a_rows, int32_t K, int64_t lda, const void* B, int32_t N,
                  int64_t ldb, const float* bias, void* C, int64_t ldc, int32_t c_bf16, int64_t M,
                  int32_t taps, const int32_t* shifts_host, void* stream);

/* C[tap][i,j] (fp32, Mo x ldc, taps stacked tap_stride apart) += sum_r A[r,i] * B[r + shifts[tap], j]
 * A (rows x Mo, lda), B (rows x No, ldb) bf16 row-major: the split-K weight-gradient GEMM
 * (dW = dY^T X of nn.Linear / nn.Conv1d backward; autograd of model_utils.py:21,89-91, NAML.py:54). */
int nr_tc_gemm_tn(const void* A, int32_t Mo, int64_t lda, const void* B, int32_t No, int64_t ldb,
                  int64_t rows, int32_t taps, const int32_t* shifts_host, float* C, int64_t ldc,
                  int64_t tap_stride, void* stream);

/* ---- bf16 title encoder: the whole NewsEncoder title path / NRMS UserEncoder in one call ------
 *   NR_ENC_MHSA: rows -> [dropout] -> MultiHeadSelfAttention -> [dropout] -> AttentionPooling
 *                (src/model/NRMS.py:27-35 news encoder; NRMS.py:55-62 user encoder with rows dense)
 *   NR_ENC_CONV: rows -> [dropout] -> Conv1d(k=3,pad=1) -> AttentionPooling   (src/model/NAML.py:47-56)
 * GEMMs run on tcgen05 tensor cores (bf16 operands, fp32 accumulation in TMEM); softmax-like
 * normalisations, tanh and all reductions are fp32.  `mask` (n,S) or NULL is used both as the MHSA
 * key mask and as the pooling mask (NRMS.py:56-57).  `save` keeps the bf16 activations for backward
 * (opaque, caller owned, nr_title_encoder_save_bytes); `ws` is scratch.
 *   NR_ENC_POOL: dense rows (n,S,H) -> AttentionPooling only (model_utils.py:7-31): NAML's view
 *                attention (NAML.py:74) and user pooling (NAML.py:91-96); Dm must equal H. */
enum { NR_ENC_MHSA = 0, NR_ENC_CONV = 1, NR_ENC_POOL = 2 };
typedef struct nr_title_encoder {
    int32_t kind, S, Dm, H, Q, heads;
    const float* w[3];      /* MHSA: W_Q, W_K, W_V (H,Dm);  CONV: w[0] = Conv1d.weight (H,Dm,3) */
    const float* b[3];      /* matching biases (H) */
    const float* p_w1;      /* AttentionPooling att_fc1.weight (Q,H) */
    const float* p_b1;      /* (Q) */
    const float* p_w2;      /* att_fc2.weight (Q) */
    const float* p_b2;      /* (1) */
    const float* mask;      /* (n,S) fp32 or NULL */
    float out_drop_p;       /* MHSA output dropout (NRMS.py:32-34); 0 otherwise */
} nr_title_encoder_t;
typedef struct nr_title_encoder_grads {   /* all ACCUMULATED (+=); any pointer may be NULL = not wanted */
    float* dw[3];
    float* db[3];
    float* dp_w1; float* dp_b1; float* dp_w2; float* dp_b2;
    float* dtable;          /* token / news rows: dense embedding-table gradient (row 0 skipped) */
    float* dx_dense;        /* dense rows: (n,S,Dm) gradient, WRITTEN */
    void* dtable_ready;     /* optional cudaEvent_t recorded on `stream` as soon as dtable / dx_dense is final (MHSA kind:
                             * before the weight-gradient GEMM), so a caller can overlap the exchange of that part of
                             * its gradient bucket with the rest of the backward; NULL = not wanted */
} nr_title_encoder_grads_t;
size_t nr_title_encoder_save_bytes(const nr_title_encoder_t* enc, int64_t n);
size_t nr_title_encoder_workspace(const nr_title_encoder_t* enc, int64_t n);
/* Inference (eval mode, the news-vector cache of src/main.py:190-196): nr_title_encoder_fwd with save == NULL keeps
 * nothing for a backward and runs the whole MHSA title encoder after the row staging as ONE fused tcgen05 kernel
 * (QKV projection -> per-head attention straight from TMEM -> fc1 -> pooling; csrc/nr_fused_enc.cuh).  `ws` must then
 * hold nr_title_encoder_infer_bytes(enc, n) bytes; 0 = this encoder is outside the fused envelope (pass a save buffer). */
size_t nr_title_encoder_infer_bytes(const nr_title_encoder_t* enc, int64_t n);
/* out (n,H) fp32 */
int nr_title_encoder_fwd(const nr_title_encoder_t* enc, const nr_rows_t* rows, int64_t n, float* out,
                         void* save, size_t save_bytes, void* ws, size_t ws_bytes, void* stream);
/* g (n,H) fp32 = dL/dout.  MHSA kind over NR_ROWS_TOKEN rows (the NRMS title encoder, NRMS.py:28,73): dqkv is kept only
 * for the rows of real tokens (id != 0 = padding_idx); dtable, dw[] contract over those rows, db[] is summed over all
 * rows inside the attention backward in a fixed order.  NR_DX_GATHER=0 in the environment (read per call) keeps every
 * padded row; results agree to the order of fp32 reductions (db: to bf16 rounding of dqkv). */
int nr_title_encoder_bwd(const nr_title_encoder_t* enc, const nr_rows_t* rows, int64_t n, const float* g,
                         const void* save, size_t save_bytes, const nr_title_encoder_grads_t* grads,
                         void* ws, size_t ws_bytes, void* stream);

/* ---- device-side batch assembly --------------------------------------------------
 * The behaviors files are tokenised ONCE on the host into a device-resident index:
 *   hist_off int64 (I+1), hist_idx int32 (sum of history lengths; news indices, unknown id -> 0),
 *   pos int32 (I), neg int32 (I,K).
 * nr_batch_train does, for B lines at once, what DatasetTrain.line_mapper does per line
 * (src/dataset.py:26-49): pad_to_fix_len (:17-24, keep the LAST T clicks, pad the FRONT with 0, mask
 * 0/1 fp32), candidate order neg[:label] + pos + neg[label:] (:45-46) with the caller's label
 * (random.randint(0, npratio), :45) and the news_combined[...] feature gathers (:40,:47).
 *   sample_ids int64 (B): rows of the index; label int64 (B); news_combined int32 (news_rows, F)
 *   -> history int32 (B,T,F), mask fp32 (B,T), candidate int32 (B,1+K,F).   Bit exact. */
int nr_batch_train(const int64_t* hist_off, const int32_t* hist_idx, const int32_t* pos, const int32_t* neg,
                   const int64_t* sample_ids, const int64_t* label, const int32_t* news_combined,
                   int64_t news_rows, int64_t B, int32_t T, int32_t K, int32_t F, int32_t* history,
                   float* mask, int32_t* candidate, void* stream);
/* DatasetTest.line_mapper's history half (src/dataset.py:64-68): padded click indices + mask for B
 * impressions (imp_ids int64 (B) or NULL = rows 0..B-1) -> out_idx int32 (B,T), out_mask fp32 (B,T);
 * the candidate half is the CSR gather nr_eval_score already consumes. */
int nr_batch_eval_history(const int64_t* hist_off, const int32_t* hist_idx, const int64_t* imp_ids,
                          int64_t news_rows, int64_t B, int32_t T, int32_t* out_idx, float* out_mask,
                          void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NR_SM100_H */
