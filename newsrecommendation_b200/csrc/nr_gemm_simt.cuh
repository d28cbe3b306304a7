// nr_gemm_simt.cuh — fp32 CUDA-core GEMM (the parity-mode arithmetic, rel 1e-5).
//
//   C[i, j] = sum_k  A(i, k) * B(j, k)        i < M, j < N, k < K
//
// Operands are read through "views" (nr_common.cuh) so that the embedding
// gather, dropout and the conv im2col never materialise in HBM.  An operand is
// either K-contiguous ("NT" use: view rows = tile rows, view cols = k) or
// R-contiguous ("TN" use: view rows = k, view cols = tile rows; used for the
// weight-gradient GEMMs whose contraction runs over the n*S activation rows).
// Split-K over gridDim.z (atomic epilogues only).
//
// Tile 128 x 64 x 16, 256 threads, 8x4 accumulators per thread, register
// prefetch + double-buffered shared memory (one __syncthreads per k-tile).
#pragma once
#include "nr_common.cuh"

namespace nr {

constexpr int GBM = 128, GBN = 64, GBK = 16, GTM = 8, GTN = 4;
constexpr int GTHREADS = (GBM / GTM) * (GBN / GTN);   // 256

// ---- epilogues: called once per (row, 4 consecutive columns) ----------------------------
struct EpiBias {          // C = acc + bias
    float* C; int64_t ldc; const float* bias;
    __device__ __forceinline__ void operator()(int64_t i, int j, float4 a) const {
        if (bias) { const float4 b = ldg4(bias + j); a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
        *(float4*)(C + i * ldc + j) = a;
    }
};
struct EpiBias3 {         // columns [0,each) [each,2each) [2each,3each) take bias b0 b1 b2
    float* C; int64_t ldc; const float* b0; const float* b1; const float* b2; int each;
    __device__ __forceinline__ void operator()(int64_t i, int j, float4 a) const {
        const int w = j / each; const int jj = j - w * each;
        const float4 b = ldg4((w == 0 ? b0 : (w == 1 ? b1 : b2)) + jj);
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        *(float4*)(C + i * ldc + j) = a;
    }
};
struct EpiBiasTanh {      // C = tanh(acc + bias)   (AttentionPooling fc1, model_utils.py:21-22)
    float* C; int64_t ldc; const float* bias;
    __device__ __forceinline__ void operator()(int64_t i, int j, float4 a) const {
        const float4 b = ldg4(bias + j);
        a.x = tanhf(a.x + b.x); a.y = tanhf(a.y + b.y); a.z = tanhf(a.z + b.z); a.w = tanhf(a.w + b.w);
        *(float4*)(C + i * ldc + j) = a;
    }
};
struct EpiAccum {         // C += acc (each element owned by exactly one thread)
    float* C; int64_t ldc;
    __device__ __forceinline__ void operator()(int64_t i, int j, float4 a) const {
        float4* p = (float4*)(C + i * ldc + j);
        float4 c = *p; c.x += a.x; c.y += a.y; c.z += a.z; c.w += a.w; *p = c;
    }
};
struct EpiAtomic {        // C += acc, split-K safe
    float* C; int64_t ldc;
    __device__ __forceinline__ void operator()(int64_t i, int j, float4 a) const {
        float* p = C + i * ldc + j;
        atomicAdd(p, a.x); atomicAdd(p + 1, a.y); atomicAdd(p + 2, a.z); atomicAdd(p + 3, a.w);
    }
};
struct EpiAtomic3 {       // rows split over three gradient buffers (dW_Q, dW_K, dW_V)
    float* p0; float* p1; float* p2; int each; int64_t ldc;
    __device__ __forceinline__ void operator()(int64_t i, int j, float4 a) const {
        const int w = (int)i / each; const int ii = (int)i - w * each;
        float* p = (w == 0 ? p0 : (w == 1 ? p1 : p2)) + (int64_t)ii * ldc + j;
        atomicAdd(p, a.x); atomicAdd(p + 1, a.y); atomicAdd(p + 2, a.z); atomicAdd(p + 3, a.w);
    }
};
struct EpiAtomicConvW {   // (o, t*D + c) -> nn.Conv1d.weight grad layout dw[o][c][t]
    float* dw; int D;
    __device__ __forceinline__ void operator()(int64_t o, int j, float4 a) const {
        const int t = j / D; const int c = j - t * D;
        float* p = dw + (o * D + c) * 3 + t;
        atomicAdd(p, a.x); atomicAdd(p + 3, a.y); atomicAdd(p + 6, a.z); atomicAdd(p + 9, a.w);
    }
};
// dX of a token-gathered activation: undo input dropout, scatter-add into the dense
// embedding gradient; id 0 = padding_idx receives nothing (src/model/NRMS.py:73, B7).
struct EpiScatterTok {
    float* dtable; RowsView x;
    __device__ __forceinline__ void operator()(int64_t r, int c, float4 a) const {
        const int64_t seq = r / x.S; const int s = (int)(r - seq * x.S);
        const int64_t id = x.id_of(seq, s);
        if (id == 0) return;
        if (x.thr) a = dropout4(a, in_stream_index(r, c, x.D), x.seed, NR_STREAM_IN, x.thr, x.scale);
        float* p = x.mode == NR_ROWS_TOKEN ? dtable + id * x.D + c
                                           : dtable + id * ((int64_t)x.S * x.D) + (int64_t)s * x.D + c;
        atomicAdd(p, a.x); atomicAdd(p + 1, a.y); atomicAdd(p + 2, a.z); atomicAdd(p + 3, a.w);
    }
};
// dE of a small embedding table looked up by one id column (category / subcategory views)
struct EpiScatterEmb {
    float* demb; const int32_t* ids; int ids_stride; int E; int64_t emb_rows;
    __device__ __forceinline__ void operator()(int64_t r, int c, float4 a) const {
        const int64_t id = __ldg(ids + r * ids_stride);
        if (id == 0 || (uint64_t)id >= (uint64_t)emb_rows) return;     // padding_idx=0, src/model/NAML.py:19,22; out of range = padding
        float* p = demb + id * E + c;
        atomicAdd(p, a.x); atomicAdd(p + 1, a.y); atomicAdd(p + 2, a.z); atomicAdd(p + 3, a.w);
    }
};
struct EpiDropStore {     // dense dX with the input-dropout mask re-applied
    float* C; RowsView x;
    __device__ __forceinline__ void operator()(int64_t r, int c, float4 a) const {
        if (x.thr) a = dropout4(a, in_stream_index(r, c, x.D), x.seed, NR_STREAM_IN, x.thr, x.scale);
        *(float4*)(C + r * x.D + c) = a;
    }
};

// nn.Conv1d.weight (H, D, 3) read as B(o, k = t*D + c) = w[o][c][t]
struct ConvWView {
    const float* w; int H, D;
    __device__ __forceinline__ float4 load4(int64_t o, int k) const {
        if (o >= H || k >= 3 * D) return make_float4(0.f, 0.f, 0.f, 0.f);
        const int t = k / D; const int c = k - t * D;
        const float* p = w + (o * D + c) * 3 + t;
        return make_float4(__ldg(p), __ldg(p + 3), __ldg(p + 6), __ldg(p + 9));
    }
};
// same weight read as B(c, k = u*H + o) = w[o][c][2-u]  (dX of the conv)
struct ConvWTView {
    const float* w; int H, D;
    __device__ __forceinline__ float4 load4(int64_t c, int k) const {
        if (c >= D || k >= 3 * H) return make_float4(0.f, 0.f, 0.f, 0.f);
        const int u = k / H; const int o = k - u * H;
        const float* p = w + ((int64_t)o * D + c) * 3 + (2 - u);
        const int st = D * 3;
        return make_float4(__ldg(p), __ldg(p + st), __ldg(p + 2 * st), __ldg(p + 3 * st));
    }
};
// transposed read of a row-major (R x Ccols) matrix: B(j, k) = w[k][j]  (dX = dY . W)
struct TransView {
    const float* w; int R; int Ccols;   // logical view: rows = Ccols, cols = R
    __device__ __forceinline__ float4 load4(int64_t j, int k) const {
        if (j >= Ccols || k >= R) return make_float4(0.f, 0.f, 0.f, 0.f);
        const float* p = w + (int64_t)k * Ccols + j;
        return make_float4(__ldg(p), __ldg(p + Ccols), __ldg(p + 2 * Ccols), __ldg(p + 3 * Ccols));
    }
};
struct TransStack3View {  // B(j, k) over the stacked [W_Q; W_K; W_V] (3*each x Ccols), transposed
    const float* p0; const float* p1; const float* p2; int each; int Ccols;
    __device__ __forceinline__ float4 load4(int64_t j, int k) const {
        if (j >= Ccols || k >= 3 * each) return make_float4(0.f, 0.f, 0.f, 0.f);
        const int w = k / each; const int kk = k - w * each;
        const float* p = (w == 0 ? p0 : (w == 1 ? p1 : p2)) + (int64_t)kk * Ccols + j;
        return make_float4(__ldg(p), __ldg(p + Ccols), __ldg(p + 2 * Ccols), __ldg(p + 3 * Ccols));
    }
};

template <bool A_KCONTIG, bool B_KCONTIG, class AV, class BV, class Epi>
__global__ void __launch_bounds__(GTHREADS, 2)
gemm_simt_kernel(const AV a, const BV b, const Epi epi, const int64_t M, const int N, const int64_t K,
                 const int64_t k_chunk) {
    __shared__ __align__(16) float As[2][GBK][GBM + 4];
    __shared__ __align__(16) float Bs[2][GBK][GBN + 4];
    const int tid = threadIdx.x;
    const int tx = tid % (GBN / GTN), ty = tid / (GBN / GTN);
    const int64_t m0 = (int64_t)blockIdx.x * GBM;
    const int n0 = blockIdx.y * GBN;
    const int64_t k_begin = (int64_t)blockIdx.z * k_chunk;
    const int64_t k_end = (k_begin + k_chunk < K) ? k_begin + k_chunk : K;
    if (k_begin >= k_end) return;

    constexpr int A_F4 = GBM * GBK / 4 / GTHREADS;   // 2
    constexpr int B_F4 = GBN * GBK / 4 / GTHREADS;   // 1
    float4 ra[A_F4], rb[B_F4];

    auto fetch = [&](int64_t k0) {
#pragma unroll
        for (int i = 0; i < A_F4; ++i) {
            const int f = tid + i * GTHREADS;
            if (A_KCONTIG) {
                const int r = f / (GBK / 4), kq = f % (GBK / 4);
                const int64_t k = k0 + kq * 4;
                ra[i] = (k < k_end) ? a.load4(m0 + r, (int)k) : make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
                const int kk = f / (GBM / 4), rq = f % (GBM / 4);
                const int64_t k = k0 + kk;
                ra[i] = (k < k_end && m0 + rq * 4 < M) ? a.load4(k, (int)(m0 + rq * 4))
                                                       : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
#pragma unroll
        for (int i = 0; i < B_F4; ++i) {
            const int f = tid + i * GTHREADS;
            if (B_KCONTIG) {
                const int r = f / (GBK / 4), kq = f % (GBK / 4);
                const int64_t k = k0 + kq * 4;
                rb[i] = (k < k_end) ? b.load4(n0 + r, (int)k) : make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
                const int kk = f / (GBN / 4), rq = f % (GBN / 4);
                const int64_t k = k0 + kk;
                rb[i] = (k < k_end && n0 + rq * 4 < N) ? b.load4(k, n0 + rq * 4)
                                                       : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    };
    auto stash = [&](int buf) {
#pragma unroll
        for (int i = 0; i < A_F4; ++i) {
            const int f = tid + i * GTHREADS;
            if (A_KCONTIG) {
                const int r = f / (GBK / 4), kq = f % (GBK / 4);
                As[buf][kq * 4 + 0][r] = ra[i].x; As[buf][kq * 4 + 1][r] = ra[i].y;
                As[buf][kq * 4 + 2][r] = ra[i].z; As[buf][kq * 4 + 3][r] = ra[i].w;
            } else {
                const int kk = f / (GBM / 4), rq = f % (GBM / 4);
                *(float4*)&As[buf][kk][rq * 4] = ra[i];
            }
        }
#pragma unroll
        for (int i = 0; i < B_F4; ++i) {
            const int f = tid + i * GTHREADS;
            if (B_KCONTIG) {
                const int r = f / (GBK / 4), kq = f % (GBK / 4);
                Bs[buf][kq * 4 + 0][r] = rb[i].x; Bs[buf][kq * 4 + 1][r] = rb[i].y;
                Bs[buf][kq * 4 + 2][r] = rb[i].z; Bs[buf][kq * 4 + 3][r] = rb[i].w;
            } else {
                const int kk = f / (GBN / 4), rq = f % (GBN / 4);
                *(float4*)&Bs[buf][kk][rq * 4] = rb[i];
            }
        }
    };

    float acc[GTM][GTN];
#pragma unroll
    for (int i = 0; i < GTM; ++i)
#pragma unroll
        for (int j = 0; j < GTN; ++j) acc[i][j] = 0.f;

    fetch(k_begin);
    stash(0);
    __syncthreads();
    int buf = 0;
    for (int64_t k0 = k_begin; k0 < k_end; k0 += GBK) {
        const bool more = k0 + GBK < k_end;
        if (more) fetch(k0 + GBK);
#pragma unroll
        for (int kk = 0; kk < GBK; ++kk) {
            const float4 a0 = *(const float4*)&As[buf][kk][ty * GTM];
            const float4 a1 = *(const float4*)&As[buf][kk][ty * GTM + 4];
            const float4 b0 = *(const float4*)&Bs[buf][kk][tx * GTN];
            const float av[GTM] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[GTN] = {b0.x, b0.y, b0.z, b0.w};
#pragma unroll
            for (int i = 0; i < GTM; ++i)
#pragma unroll
                for (int j = 0; j < GTN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        if (more) {
            stash(buf ^ 1);
            __syncthreads();
            buf ^= 1;
        }
    }
    const int j = n0 + tx * GTN;
    if (j < N) {
#pragma unroll
        for (int i = 0; i < GTM; ++i) {
            const int64_t r = m0 + ty * GTM + i;
            if (r < M) epi(r, j, make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));
        }
    }
}

// NT: C(M x N) = A(M x K) . B(N x K)^T, both operands K-contiguous views.
template <class AV, class BV, class Epi>
// splits > 1 (ACCUMULATING epilogues only, e.g. EpiAtomic): the contraction is cut into `splits` k-chunks on
// blockIdx.z — for the few-tile GEMMs over the category tables, whose latency is one CTA walking all of K.
int gemm_nt(const AV& a, const BV& b, const Epi& e, int64_t M, int N, int K, cudaStream_t st,
            const char* what, int splits = 1) {
    if (M == 0 || N == 0) return 0;
    if ((N & 3) || (K & 3)) { set_error("%s: N=%d K=%d must be multiples of 4", what, N, K); return 1; }
    int k_chunk = K;
    if (splits > 1) {
        k_chunk = ((K + splits - 1) / splits + GBK - 1) / GBK * GBK;
        splits = (K + k_chunk - 1) / k_chunk;
    } else {
        splits = 1;
    }
    dim3 grid((unsigned)((M + GBM - 1) / GBM), (unsigned)((N + GBN - 1) / GBN), (unsigned)splits);
    ProfScope ps(what, st);
    gemm_simt_kernel<true, true, AV, BV, Epi><<<grid, GTHREADS, 0, st>>>(a, b, e, M, N, K, k_chunk);
    return check_last(what);
}

// TN: C(M x N) += sum over k < K of A(k, i) * B(k, j); split-K over the long contraction.
template <class AV, class BV, class Epi>
int gemm_tn(const AV& a, const BV& b, const Epi& e, int M, int N, int64_t K, cudaStream_t st,
            const char* what) {
    if (M == 0 || N == 0 || K == 0) return 0;
    if ((N & 3) || (M & 3)) { set_error("%s: M=%d N=%d must be multiples of 4", what, M, N); return 1; }
    const int tiles = ((M + GBM - 1) / GBM) * ((N + GBN - 1) / GBN);
    int64_t splits = (4 * (int64_t)sm_count() + tiles - 1) / tiles;
    const int64_t max_splits = (K + 4 * GBK - 1) / (4 * GBK);
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > 65535) splits = 65535;
    int64_t k_chunk = (K + splits - 1) / splits;
    k_chunk = (k_chunk + GBK - 1) / GBK * GBK;
    splits = (K + k_chunk - 1) / k_chunk;
    dim3 grid((unsigned)((M + GBM - 1) / GBM), (unsigned)((N + GBN - 1) / GBN), (unsigned)splits);
    ProfScope ps(what, st);
    gemm_simt_kernel<false, false, AV, BV, Epi><<<grid, GTHREADS, 0, st>>>(a, b, e, M, N, K, k_chunk);
    return check_last(what);
}

}  // namespace nr
