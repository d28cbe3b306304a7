// nr_ops2.cu — Conv1d(k=3,pad=1) as an implicit GEMM over the im2col view, and the NAML
// category/subcategory dense views.  fp32 parity mode.
#include "nr_gemm_simt.cuh"
#include "nr_tails.cuh"

using namespace nr;

namespace nr {
int set_seed_salt_ops2(const void* p) { return set_seed_salt_this_tu(p); }
}  // namespace nr

namespace nr {
__global__ void table_gather_kernel(const float* __restrict__ T, int64_t rows, const int32_t* __restrict__ ids,
                                    int ids_stride, float* __restrict__ out, int64_t out_stride, int64_t n, int h4) {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n * h4; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = t / h4;
        const int q = (int)(t - i * h4);
        int64_t id = __ldg(ids + i * ids_stride);
        if (id < 0 || id >= rows) id = 0;
        *(float4*)(out + i * out_stride + 4 * q) = __ldg((const float4*)(T + id * (int64_t)h4 * 4) + q);
    }
}
__global__ void table_scatter_kernel(const float* __restrict__ dout, int64_t dout_stride, const int32_t* __restrict__ ids,
                                     int ids_stride, int64_t rows, float* __restrict__ dT, int64_t n, int h4) {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n * h4; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = t / h4;
        const int q = (int)(t - i * h4);
        int64_t id = __ldg(ids + i * ids_stride);
        if (id < 0 || id >= rows) id = 0;
        atomicAdd((float4*)(dT + id * (int64_t)h4 * 4) + q, ldg4(dout + i * dout_stride + 4 * q));
    }
}
// Small tables (17 categories): every one of the n rows lands on one of a handful of table rows, i.e. ~n / rows
// same-address reductions per element in L2.  Each CTA first reduces its slice of the rows into a private copy of
// a column tile of dT in shared memory (shared-memory atomics), then adds that copy to dT once.
// grid = (column tiles, row slices); block = cw4 float4 columns x (256 / cw4) row lanes.
__global__ void __launch_bounds__(256)
table_scatter_small_kernel(const float* __restrict__ dout, int64_t dout_stride, const int32_t* __restrict__ ids,
                           int ids_stride, int rows, float* __restrict__ dT, int64_t n, int h4, int cw4, int slice) {
    extern __shared__ __align__(16) float acc[];          // [rows][cw4 * 4]
    const int lanes = 256 / cw4;
    const int cq = threadIdx.x % cw4, rl = threadIdx.x / cw4;
    const int q = blockIdx.x * cw4 + cq;                   // float4 column
    for (int i = threadIdx.x; i < rows * cw4 * 4; i += 256) acc[i] = 0.f;
    __syncthreads();
    const int64_t i0 = (int64_t)blockIdx.y * slice;
    const int64_t i1 = i0 + slice < n ? i0 + slice : n;
    if (rl < lanes && q < h4) {
        for (int64_t i = i0 + rl; i < i1; i += lanes) {
            int id = __ldg(ids + i * ids_stride);
            if (id < 0 || id >= rows) id = 0;
            const float4 v = ldg4(dout + i * dout_stride + 4 * q);
            float* a = acc + (id * cw4 + cq) * 4;
            atomicAdd(a, v.x); atomicAdd(a + 1, v.y); atomicAdd(a + 2, v.z); atomicAdd(a + 3, v.w);
        }
    }
    __syncthreads();
    for (int t = threadIdx.x; t < rows * cw4; t += 256) {
        const int r = t / cw4, c = t - r * cw4;
        const int qq = blockIdx.x * cw4 + c;
        if (qq >= h4) continue;
        const float4 v = *(const float4*)(acc + t * 4);
        if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) atomicAdd((float4*)(dT + (int64_t)r * h4 * 4) + qq, v);
    }
}
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

// dense(emb[ids]) is linear, so for a small vocabulary (17 categories, 264 sub-categories against 7,040 news per
// step) the layer is applied to the TABLE once — T = emb . w^T + b, (emb_rows, H) — and the batch only gathers rows
// of T; backward scatter-adds dout into dT (emb_rows, H) and runs the three gradient products on emb_rows rows
// instead of n.  Same fp32 arithmetic, different summation order.
size_t nr_embed_dense_workspace(int64_t emb_rows, int32_t H) {
    return emb_rows > 0 && H > 0 ? (size_t)emb_rows * H * sizeof(float) + 256 : 256;
}

static bool use_table(int64_t emb_rows, int64_t n, const void* ws, size_t ws_bytes, int H) {
    return ws && emb_rows * 2 <= n && ws_bytes >= (size_t)emb_rows * H * sizeof(float) + 256;
}

int nr_embed_dense_fwd(const float* emb, int64_t emb_rows, const int32_t* ids, int32_t ids_stride,
                       const float* w, const float* b, float* out, int64_t out_stride, int64_t n, int32_t E,
                       int32_t H, void* ws, size_t ws_bytes, void* stream) {
    NR_REQUIRE(emb && ids && w && b && out, "nr_embed_dense_fwd: NULL pointer");
    NR_REQUIRE(emb_rows > 0 && ids_stride >= 1 && E > 0 && (E & 3) == 0 && H > 0 && (H & 3) == 0 &&
                   out_stride >= H && (out_stride & 3) == 0,
               "nr_embed_dense_fwd: unsupported shape E=%d H=%d", E, H);
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (use_table(emb_rows, n, ws, ws_bytes, H)) {
        float* T = (float*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
        if (int e = gemm_nt(DenseView{emb, E, emb_rows, E}, DenseView{w, E, H, E}, EpiBias{T, H, b}, emb_rows, H, E, st,
                            "embed_dense.table"))
            return e;
        const int h4 = H >> 2;
        int64_t blocks = (n * h4 + 255) / 256;
        if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
        ProfScope ps("embed_dense.gather", st);
        table_gather_kernel<<<(unsigned)blocks, 256, 0, st>>>(T, emb_rows, ids, ids_stride, out, out_stride, n, h4);
        return check_last("embed_dense.gather");
    }
    return gemm_nt(EmbedView{emb, ids, ids_stride, n, E, emb_rows}, DenseView{w, E, H, E}, EpiBias{out, out_stride, b}, n, H,
                   E, st, "embed_dense.fwd");
}

int nr_embed_dense_bwd(const float* emb, int64_t emb_rows, const int32_t* ids, int32_t ids_stride,
                       const float* w, const float* dout, int64_t dout_stride, float* dw, float* db, float* demb,
                       int64_t n, int32_t E, int32_t H, void* ws, size_t ws_bytes, void* stream) {
    NR_REQUIRE(emb && ids && w && dout && dw && db && demb, "nr_embed_dense_bwd: NULL pointer");
    NR_REQUIRE(E > 0 && (E & 3) == 0 && H > 0 && (H & 3) == 0 && (dout_stride & 3) == 0, "nr_embed_dense_bwd: bad shape");
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (use_table(emb_rows, n, ws, ws_bytes, H)) {
        float* dT = (float*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
        NR_CUDA(cudaMemsetAsync(dT, 0, (size_t)emb_rows * H * sizeof(float), st));
        const int h4 = H >> 2;
        int64_t blocks = (n * h4 + 255) / 256;
        if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
        {
            ProfScope ps("embed_dense.scatter", st);
            const int cw4 = (h4 % 20 == 0) ? 20 : 16;
            if (emb_rows <= 64 && n >= 64 * emb_rows) {
                const int tiles = (h4 + cw4 - 1) / cw4;
                int slices = 2 * sm_count() / tiles;
                if (slices < 1) slices = 1;
                const int slice = (int)((n + slices - 1) / slices);
                slices = (int)((n + slice - 1) / slice);
                table_scatter_small_kernel<<<dim3(tiles, slices), 256, (size_t)emb_rows * cw4 * 16, st>>>(
                    dout, dout_stride, ids, ids_stride, (int)emb_rows, dT, n, h4, cw4, slice);
            } else {
                table_scatter_kernel<<<(unsigned)blocks, 256, 0, st>>>(dout, dout_stride, ids, ids_stride, emb_rows, dT, n, h4);
            }
            if (int e = check_last("embed_dense.scatter")) return e;
        }
        const DenseView dv{dT, H, emb_rows, H};
        if (int e = gemm_tn(dv, DenseView{emb, E, emb_rows, E}, EpiAtomic{dw, E}, H, E, emb_rows, st, "embed_dense.dw"))
            return e;
        if (int e = colsum(dT, emb_rows, H, H, db, db, db, H, st)) return e;
        if (emb_rows < 2) return 0;
        // row 0 = padding_idx: no gradient (src/model/NAML.py:19,22)
        return gemm_nt(DenseView{dT + H, H, emb_rows - 1, H}, TransView{w, H, E}, EpiAtomic{demb + E, E}, emb_rows - 1, E, H,
                       st, "embed_dense.demb", 8);
    }
    const DenseView dv{dout, dout_stride, n, H};
    if (int e = gemm_tn(dv, EmbedView{emb, ids, ids_stride, n, E, emb_rows}, EpiAtomic{dw, E}, H, E, n, st, "embed_dense.dw"))
        return e;
    if (int e = colsum(dout, n, dout_stride, H, db, db, db, H, st)) return e;
    return gemm_nt(dv, TransView{w, H, E}, EpiScatterEmb{demb, ids, ids_stride, E, emb_rows}, n, E, H, st, "embed_dense.demb");
}

}  // extern "C"
