// nr_ops2.cu — Conv1d(k=3,pad=1) as an implicit GEMM over the im2col view, and the NAML
// category/subcategory dense views.  fp32 parity mode.
#include "nr_gemm_simt.cuh"
#include "nr_tails.cuh"

using namespace nr;

namespace nr {
int set_seed_salt_ops2(const void* p) { return set_seed_salt_this_tu(p); }
}  // namespace nr

extern "C" {

size_t nr_conv1d_k3_workspace(int64_t n, int32_t L, int32_t D, int32_t H, int32_t prec) {
    (void)n; (void)L; (void)D; (void)H; (void)prec;
    return 256;
}

int nr_conv1d_k3_fwd(const nr_rows_t* rows, const float* w, const float* b, float* y, int64_t n, int32_t L,
                     int32_t D, int32_t H, int32_t prec, void* ws, size_t ws_bytes, void* stream) {
    (void)ws; (void)ws_bytes;
    NR_REQUIRE(prec == NR_PREC_FP32, "nr_conv1d_k3_fwd: precision mode %d not built (fp32 only)", prec);
    if (int e = validate_rows(rows, "nr_conv1d_k3_fwd")) return e;
    NR_REQUIRE(w && b && y, "nr_conv1d_k3_fwd: NULL pointer");
    NR_REQUIRE(n >= 0 && L >= 1 && D > 0 && (D & 3) == 0 && H > 0 && (H & 3) == 0,
               "nr_conv1d_k3_fwd: unsupported shape L=%d D=%d H=%d", L, D, H);
    if (n == 0) return 0;
    const int64_t M = n * L;
    const ConvView a{make_rows_view(rows, n, L, D), M, 3 * D};
    return gemm_nt(a, ConvWView{w, H, D}, EpiBias{y, H, b}, M, H, 3 * D, (cudaStream_t)stream, "conv.fwd");
}

int nr_conv1d_k3_bwd(const nr_rows_t* rows, const float* w, const float* dy, float* dw, float* db,
                     float* dx_dense, float* dtable, int64_t n, int32_t L, int32_t D, int32_t H, int32_t prec,
                     void* ws, size_t ws_bytes, void* stream) {
    (void)ws; (void)ws_bytes;
    NR_REQUIRE(prec == NR_PREC_FP32, "nr_conv1d_k3_bwd: precision mode %d not built (fp32 only)", prec);
    if (int e = validate_rows(rows, "nr_conv1d_k3_bwd")) return e;
    NR_REQUIRE(w && dy && dw && db, "nr_conv1d_k3_bwd: NULL pointer");
    NR_REQUIRE(n >= 0 && L >= 1 && D > 0 && (D & 3) == 0 && H > 0 && (H & 3) == 0, "nr_conv1d_k3_bwd: bad shape");
    NR_REQUIRE(!(dtable && rows->mode == NR_ROWS_DENSE), "nr_conv1d_k3_bwd: dtable given for dense rows");
    NR_REQUIRE(!(dx_dense && dtable), "nr_conv1d_k3_bwd: give dx_dense or dtable, not both");
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t M = n * L;
    const RowsView xv = make_rows_view(rows, n, L, D);
    const ConvView xa{xv, M, 3 * D};
    const DenseView dyv{dy, H, M, H};
    // dW (H, 3D in (t,c) order) += dy^T . im2col(X)
    if (int e = gemm_tn(dyv, xa, EpiAtomicConvW{dw, D}, H, 3 * D, M, st, "conv.dw")) return e;
    if (int e = colsum(dy, M, H, H, db, db, db, H, st)) return e;
    if (dx_dense || dtable) {
        // dx[l,c] = sum_u sum_o dy[l+u-1, o] * w[o][c][2-u]: the same im2col over dy
        RowsView dv;
        dv.mode = NR_ROWS_DENSE; dv.dense = dy; dv.table = nullptr; dv.ids = nullptr; dv.ids_stride = 0;
        dv.S = L; dv.D = H; dv.rows = M; dv.thr = 0u; dv.scale = 1.f; dv.seed = 0;
        const ConvView da{dv, M, 3 * H};
        if (dx_dense) return gemm_nt(da, ConvWTView{w, H, D}, EpiDropStore{dx_dense, xv}, M, D, 3 * H, st, "conv.dx");
        return gemm_nt(da, ConvWTView{w, H, D}, EpiScatterTok{dtable, xv}, M, D, 3 * H, st, "conv.dtable");
    }
    return 0;
}

int nr_embed_dense_fwd(const float* emb, int64_t emb_rows, const int32_t* ids, int32_t ids_stride,
                       const float* w, const float* b, float* out, int64_t out_stride, int64_t n, int32_t E,
                       int32_t H, void* stream) {
    NR_REQUIRE(emb && ids && w && b && out, "nr_embed_dense_fwd: NULL pointer");
    NR_REQUIRE(emb_rows > 0 && ids_stride >= 1 && E > 0 && (E & 3) == 0 && H > 0 && (H & 3) == 0 &&
                   out_stride >= H && (out_stride & 3) == 0,
               "nr_embed_dense_fwd: unsupported shape E=%d H=%d", E, H);
    if (n == 0) return 0;
    return gemm_nt(EmbedView{emb, ids, ids_stride, n, E}, DenseView{w, E, H, E}, EpiBias{out, out_stride, b}, n, H,
                   E, (cudaStream_t)stream, "embed_dense.fwd");
}

int nr_embed_dense_bwd(const float* emb, int64_t emb_rows, const int32_t* ids, int32_t ids_stride,
                       const float* w, const float* dout, int64_t dout_stride, float* dw, float* db, float* demb,
                       int64_t n, int32_t E, int32_t H, void* ws, size_t ws_bytes, void* stream) {
    (void)ws; (void)ws_bytes; (void)emb_rows;
    NR_REQUIRE(emb && ids && w && dout && dw && db && demb, "nr_embed_dense_bwd: NULL pointer");
    NR_REQUIRE(E > 0 && (E & 3) == 0 && H > 0 && (H & 3) == 0 && (dout_stride & 3) == 0, "nr_embed_dense_bwd: bad shape");
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const DenseView dv{dout, dout_stride, n, H};
    if (int e = gemm_tn(dv, EmbedView{emb, ids, ids_stride, n, E}, EpiAtomic{dw, E}, H, E, n, st, "embed_dense.dw"))
        return e;
    if (int e = colsum(dout, n, dout_stride, H, db, db, db, H, st)) return e;
    return gemm_nt(dv, TransView{w, H, E}, EpiScatterEmb{demb, ids, ids_stride, E}, n, E, H, st, "embed_dense.demb");
}

}  // extern "C"
