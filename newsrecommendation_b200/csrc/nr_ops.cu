// nr_ops.cu — op-level C ABI: AttentionPooling, MultiHeadSelfAttention, Conv1d(k=3),
// NAML side-view dense.  fp32 parity mode: SIMT GEMMs (nr_gemm_simt.cuh) + tail kernels.
#include "nr_gemm_simt.cuh"
#include "nr_tails.cuh"

using namespace nr;

namespace nr {
int set_seed_salt_ops(const void* p) { return set_seed_salt_this_tu(p); }
}  // namespace nr

namespace {

template <int DK>
int launch_attn_fwd(const float* qkv, const float* mask, float* ctx, int64_t n, int S, int heads,
                    float p, uint64_t seed, cudaStream_t st) {
    const size_t smem = attn_fwd_smem(S, DK);
    NR_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<DK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int groups = (heads + 3) / 4;
    ProfScope ps("attn_fwd", st);
    attn_fwd_kernel<DK><<<(unsigned)(n * groups), 128, smem, st>>>(
        qkv, mask, ctx, S, heads, 1.0f / sqrtf((float)DK), p > 0.f ? drop_threshold(p) : 0u,
        p > 0.f ? 1.f / (1.f - p) : 1.f, seed);
    return check_last("attn_fwd");
}
template <int DK>
int launch_attn_bwd(const float* qkv, const float* mask, const float* dctx, float* dqkv, int64_t n, int S,
                    int heads, float p, uint64_t seed, cudaStream_t st) {
    const size_t smem = attn_bwd_smem(S, DK);
    NR_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<DK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int groups = (heads + 3) / 4;
    ProfScope ps("attn_bwd", st);
    attn_bwd_kernel<DK><<<(unsigned)(n * groups), 128, smem, st>>>(
        qkv, mask, dctx, dqkv, S, heads, 1.0f / sqrtf((float)DK), p > 0.f ? drop_threshold(p) : 0u,
        p > 0.f ? 1.f / (1.f - p) : 1.f, seed);
    return check_last("attn_bwd");
}

#define NR_DISPATCH_DK(dk, CALL)                                                       \
    switch (dk) {                                                                      \
        case 8: { constexpr int DK = 8; return CALL; }                                 \
        case 16: { constexpr int DK = 16; return CALL; }                               \
        case 20: { constexpr int DK = 20; return CALL; }                               \
        case 32: { constexpr int DK = 32; return CALL; }                               \
        case 40: { constexpr int DK = 40; return CALL; }                               \
        default: set_error("attention: head dim %d not in compiled set {8,16,20,32,40}", dk); return 1; \
    }

int attn_fwd(const float* qkv, const float* mask, float* ctx, int64_t n, int S, int heads, int dk, float p,
             uint64_t seed, cudaStream_t st) {
    NR_DISPATCH_DK(dk, launch_attn_fwd<DK>(qkv, mask, ctx, n, S, heads, p, seed, st));
}
int attn_bwd(const float* qkv, const float* mask, const float* dctx, float* dqkv, int64_t n, int S, int heads,
             int dk, float p, uint64_t seed, cudaStream_t st) {
    NR_DISPATCH_DK(dk, launch_attn_bwd<DK>(qkv, mask, dctx, dqkv, n, S, heads, p, seed, st));
}

int require_fp32(int prec, const char* who) {
    NR_REQUIRE(prec == NR_PREC_FP32, "%s: precision mode %d is not built into this library (fp32 only)", who, prec);
    return 0;
}

}  // namespace

extern "C" {

// ============================ AttentionPooling ==============================================
size_t nr_attention_pooling_workspace(int64_t n, int32_t S, int32_t H, int32_t Q, int32_t prec) {
    (void)H; (void)prec;
    return Workspace::pad((size_t)n * S * Q * sizeof(float));   // dz in backward
}

int nr_attention_pooling_fwd(const float* x, const float* mask, const float* w1, const float* b1,
                             const float* w2, const float* b2, float* out, float* e_save, float* alpha_save,
                             int64_t n, int32_t S, int32_t H, int32_t Q, int32_t prec, void* ws,
                             size_t ws_bytes, void* stream) {
    (void)ws; (void)ws_bytes;
    if (int e = require_fp32(prec, "nr_attention_pooling_fwd")) return e;
    NR_REQUIRE(x && w1 && b1 && w2 && b2 && out && e_save && alpha_save, "nr_attention_pooling_fwd: NULL pointer");
    NR_REQUIRE(n >= 0 && S >= 1 && S <= 4096 && (H & 3) == 0 && (Q & 3) == 0 && H > 0 && Q > 0,
               "nr_attention_pooling_fwd: unsupported shape S=%d H=%d Q=%d", S, H, Q);
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t M = n * S;
    if (int e = gemm_nt(DenseView{x, H, M, H}, DenseView{w1, H, Q, H}, EpiBiasTanh{e_save, Q, b1}, M, Q, H, st,
                        "pool.fc1"))
        return e;
    ProfScope ps("pool_tail_fwd", st);
    pool_tail_fwd_kernel<<<(unsigned)n, 128, S * sizeof(float), st>>>(x, e_save, mask, w2, b2, out, alpha_save,
                                                                      S, H, Q);
    return check_last("pool_tail_fwd");
}

int nr_attention_pooling_bwd(const float* g, const float* x, const float* w1, const float* w2,
                             const float* e_save, const float* alpha_save, float* dx, float* dw1, float* db1,
                             float* dw2, float* db2, int64_t n, int32_t S, int32_t H, int32_t Q, int32_t prec,
                             void* ws, size_t ws_bytes, void* stream) {
    if (int e = require_fp32(prec, "nr_attention_pooling_bwd")) return e;
    NR_REQUIRE(g && x && w1 && w2 && e_save && alpha_save && dw1 && db1 && dw2 && db2,
               "nr_attention_pooling_bwd: NULL pointer");
    NR_REQUIRE(n >= 0 && S >= 1 && S <= 4096 && (H & 3) == 0 && (Q & 3) == 0, "nr_attention_pooling_bwd: bad shape");
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    Workspace W(ws, ws_bytes);
    const int64_t M = n * S;
    float* dz = W.take<float>((size_t)M * Q);
    NR_REQUIRE(dz, "nr_attention_pooling_bwd: workspace too small (%zu bytes)", ws_bytes);
    int64_t grid = n < 4 * (int64_t)sm_count() ? n : 4 * (int64_t)sm_count();
    ProfScope ps("pool_tail_bwd", st);
    pool_tail_bwd_kernel<<<(unsigned)grid, 256, (S + Q) * sizeof(float), st>>>(g, x, e_save, alpha_save, w2, dz,
                                                                               dx, dw2, db2, n, S, H, Q);
    if (int e = check_last("pool_tail_bwd")) return e;
    if (dx)   // dx += dz . W1
        if (int e = gemm_nt(DenseView{dz, Q, M, Q}, TransView{w1, Q, H}, EpiAccum{dx, H}, M, H, Q, st, "pool.dx"))
            return e;
    // dW1 (Q,H) += dz^T . x
    if (int e = gemm_tn(DenseView{dz, Q, M, Q}, DenseView{x, H, M, H}, EpiAtomic{dw1, H}, Q, H, M, st, "pool.dw1"))
        return e;
    return colsum(dz, M, Q, Q, db1, db1, db1, Q, st);
}

// ============================ MultiHeadSelfAttention ==========================================
size_t nr_mhsa_workspace(int64_t n, int32_t S, int32_t Dm, int32_t H, int32_t prec) {
    (void)Dm; (void)prec;
    return Workspace::pad((size_t)n * S * 3 * H * sizeof(float));   // dqkv in backward
}

int nr_mhsa_fwd(const nr_rows_t* rows, const float* key_mask, const float* wq, const float* bq,
                const float* wk, const float* bk, const float* wv, const float* bv, float* ctx,
                float* qkv_save, int64_t n, int32_t S, int32_t Dm, int32_t heads, int32_t dk,
                float out_drop_p, int32_t prec, void* ws, size_t ws_bytes, void* stream) {
    (void)ws; (void)ws_bytes;
    if (int e = require_fp32(prec, "nr_mhsa_fwd")) return e;
    if (int e = validate_rows(rows, "nr_mhsa_fwd")) return e;
    NR_REQUIRE(wq && bq && wk && bk && wv && bv && ctx && qkv_save, "nr_mhsa_fwd: NULL pointer");
    NR_REQUIRE(n >= 0 && S >= 1 && S <= 64 && (Dm & 3) == 0 && Dm > 0 && heads > 0,
               "nr_mhsa_fwd: unsupported shape S=%d (max 64) Dm=%d heads=%d", S, Dm, heads);
    NR_REQUIRE(out_drop_p >= 0.f && out_drop_p < 1.f, "nr_mhsa_fwd: out_drop_p outside [0,1)");
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int H = heads * dk;
    const int64_t M = n * S;
    const RowsView xv = make_rows_view(rows, n, S, Dm);
    if (int e = gemm_nt(xv, Stack3View{wq, wk, wv, H, 3 * H, Dm}, EpiBias3{qkv_save, 3 * H, bq, bk, bv, H}, M,
                        3 * H, Dm, st, "mhsa.qkv"))
        return e;
    return attn_fwd(qkv_save, key_mask, ctx, n, S, heads, dk, out_drop_p, rows->seed, st);
}

int nr_mhsa_bwd(const nr_rows_t* rows, const float* key_mask, const float* wq, const float* wk,
                const float* wv, const float* qkv_save, const float* dctx, float* dwq, float* dbq, float* dwk,
                float* dbk, float* dwv, float* dbv, float* dx_dense, float* dtable, int64_t n, int32_t S,
                int32_t Dm, int32_t heads, int32_t dk, float out_drop_p, int32_t prec, void* ws,
                size_t ws_bytes, void* stream) {
    if (int e = require_fp32(prec, "nr_mhsa_bwd")) return e;
    if (int e = validate_rows(rows, "nr_mhsa_bwd")) return e;
    NR_REQUIRE(wq && wk && wv && qkv_save && dctx && dwq && dbq && dwk && dbk && dwv && dbv,
               "nr_mhsa_bwd: NULL pointer");
    NR_REQUIRE(n >= 0 && S >= 1 && S <= 64 && (Dm & 3) == 0, "nr_mhsa_bwd: unsupported shape");
    NR_REQUIRE(!(dtable && rows->mode == NR_ROWS_DENSE), "nr_mhsa_bwd: dtable given for dense rows");
    NR_REQUIRE(!(dx_dense && dtable), "nr_mhsa_bwd: give dx_dense or dtable, not both");
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int H = heads * dk;
    const int64_t M = n * S;
    Workspace W(ws, ws_bytes);
    float* dqkv = W.take<float>((size_t)M * 3 * H);
    NR_REQUIRE(dqkv, "nr_mhsa_bwd: workspace too small (%zu bytes)", ws_bytes);
    if (int e = attn_bwd(qkv_save, key_mask, dctx, dqkv, n, S, heads, dk, out_drop_p, rows->seed, st)) return e;
    const RowsView xv = make_rows_view(rows, n, S, Dm);
    const DenseView dq{dqkv, 3 * H, M, 3 * H};
    // [dWq; dWk; dWv] (3H, Dm) += dqkv^T . X
    if (int e = gemm_tn(dq, xv, EpiAtomic3{dwq, dwk, dwv, H, Dm}, 3 * H, Dm, M, st, "mhsa.dw")) return e;
    if (int e = colsum(dqkv, M, 3 * H, 3 * H, dbq, dbk, dbv, H, st)) return e;
    const TransStack3View wt{wq, wk, wv, H, Dm};
    if (dx_dense) return gemm_nt(dq, wt, EpiDropStore{dx_dense, xv}, M, Dm, 3 * H, st, "mhsa.dx");
    if (dtable) return gemm_nt(dq, wt, EpiScatterTok{dtable, xv}, M, Dm, 3 * H, st, "mhsa.dtable");
    return 0;
}

}  // extern "C"
