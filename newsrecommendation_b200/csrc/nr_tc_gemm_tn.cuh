// nr_tc_gemm_tn.cuh — split-K tcgen05 "TN" GEMM for the weight gradients (bf16 x bf16 -> fp32 in TMEM).
//
//   C[tap][i, j] += sum_{r < rows}  A[r, i] * B[r + shift[tap], j]        i < Mo, j < No
//
// The contraction runs over the activation ROWS r (n*S token rows), so both operands are
// "MN-major" for the tensor core: A (rows x Mo) and B (rows x No) are row-major bf16 and a TMA box of
// [64 rows x 64 columns] lands in shared memory as 64 k-rows of 128 bytes (SWIZZLE_128B) — exactly the
// canonical MN-major SW128 layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units, with
// SBO = 1024 B (8 k-rows) and LBO = 8192 B (next 64-column block = next TMA box).
// taps = 3 with shifts {-1,0,+1} gives the Conv1d(k=3) weight gradient dW[o][c][t] (NAML.py:27-32).
//
// One CTA = one (128 x Nt) output tile (Nt <= 512: up to two MMAs of N <= 256 per k-step) over one
// slice of the rows (split-K over blockIdx.y); the fp32 partial tile is added to the gradient with
// atomics by the epilogue functor.  192 threads: warp 0 TMA producer, warp 1 MMA issuer (owns TMEM),
// warps 2..5 epilogue.
#pragma once
#include "nr_common.cuh"
#include "nr_tc.cuh"

namespace nr {

constexpr int TN_BM = 128, TN_BK = 64, TN_THREADS = 192, TN_BOX_BYTES = 64 * 128;

struct TnParams {
    int Mo, No;            // output extent (A columns, B columns)
    int Nt;                // n tile (<= 512, multiple of 16)
    int n0w, n1w;          // widths of the two MMAs: n0w = min(Nt,256), n1w = Nt - n0w
    int nb;                // 64-column boxes per B tile
    int m_tiles, n_tiles, taps;
    int shift[3];
    int total_chunks, chunks_per_split;
    int stages;
    // GATHER kernels (taps == 1, one n tile): A holds *k_count compact rows; compact row r pairs with B row list[r].
    // Four extra producer warps copy the 64 gathered B rows of a stage with 16-byte cp.async into the MN-major
    // SWIZZLE_128B boxes a TMA load would have produced (rows past the count are zero filled); A still arrives by TMA.
    const int32_t* list;
    const int32_t* k_count;
    const __nv_bfloat16* b_ptr;
    int64_t ldb;
};
constexpr int TN_GATHER_WARPS = 4;

template <class Epi, bool GATHER = false>
__global__ void __launch_bounds__(TN_THREADS + (GATHER ? 32 * TN_GATHER_WARPS : 0), 1)
gemm_tn_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const TnParams p, const Epi epi_in) {
    int total_chunks = p.total_chunks, per_split = p.chunks_per_split;
    int64_t k_rows = 0;
    if (GATHER) {           // the contraction length is a device-side count: split it evenly over blockIdx.y
        k_rows = (int64_t)__ldg(p.k_count);
        total_chunks = (int)((k_rows + TN_BK - 1) / TN_BK);
        per_split = (total_chunks + (int)gridDim.y - 1) / (int)gridDim.y;
    }
    const int c_begin = blockIdx.y * per_split;
    int c_end = c_begin + per_split;
    if (c_end > total_chunks) c_end = total_chunks;
    if (c_begin >= c_end) return;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const int stages = p.stages;
    const uint32_t a_bytes = 2 * TN_BOX_BYTES, b_bytes = (uint32_t)p.nb * TN_BOX_BYTES;
    const uint32_t stage_bytes = a_bytes + b_bytes;
    uint64_t* bars = (uint64_t*)(smem + (size_t)stages * stage_bytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + stages;
    uint64_t* tfull = bars + 2 * stages;
    uint32_t* tmem_slot = (uint32_t*)(tfull + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int tile = blockIdx.x;
    const int mt = tile % p.m_tiles; tile /= p.m_tiles;
    const int nt = tile % p.n_tiles;
    const int tap = tile / p.n_tiles;
    const int i0 = mt * TN_BM, j0 = nt * p.Nt;
    const int shift = p.shift[tap];

    if (warp == 0 && lane == 0) {
        tc::prefetch_tmap(&tmA);
        tc::prefetch_tmap(&tmB);
        // full: one arrive (+ tx bytes) from the TMA lane; GATHER adds one cp.async-completion arrive per producer thread
        for (int s = 0; s < stages; ++s) { tc::mbar_init(full + s, GATHER ? 1 + 32 * TN_GATHER_WARPS : 1); tc::mbar_init(empty + s, 1); }
        tc::mbar_init(tfull, 1);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            for (int c = c_begin; c < c_end; ++c) {
                const int r = c * TN_BK;
                tc::mbar_wait(empty + s, ph ^ 1);
                uint8_t* sa = smem + (size_t)s * stage_bytes;
                tc::mbar_arrive_expect_tx(full + s, GATHER ? a_bytes : stage_bytes);
                tc::tma_load_2d(sa, &tmA, full + s, i0, r);
                tc::tma_load_2d(sa + TN_BOX_BYTES, &tmA, full + s, i0 + 64, r);
                if (!GATHER)
                    for (int b = 0; b < p.nb; ++b)
                        tc::tma_load_2d(sa + a_bytes + (size_t)b * TN_BOX_BYTES, &tmB, full + s, j0 + 64 * b, r + shift);
                if (++s == stages) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        const uint32_t idesc0 = tc::umma_idesc_bf16(TN_BM, p.n0w, 1, 1);
        const uint32_t idesc1 = tc::umma_idesc_bf16(TN_BM, p.n1w > 0 ? p.n1w : 16, 1, 1);
        int s = 0; uint32_t ph = 0;
        for (int c = c_begin; c < c_end; ++c) {
            tc::mbar_wait(full + s, ph);
            if (GATHER) tc::fence_proxy_async_smem();        // cp.async wrote B through the generic proxy
            tc::tc_fence_after();
            if (lane == 0) {
                const uint32_t sa = tc::smem_u32(smem + (size_t)s * stage_bytes);
                const uint64_t da = tc::umma_desc_sw128(sa, TN_BOX_BYTES, 1024);
                const uint64_t db0 = tc::umma_desc_sw128(sa + a_bytes, TN_BOX_BYTES, 1024);
                const uint64_t db1 = tc::umma_desc_sw128(sa + a_bytes + 4 * TN_BOX_BYTES, TN_BOX_BYTES, 1024);
#pragma unroll
                for (int k = 0; k < TN_BK / 16; ++k) {   // 16 k-rows = 2048 B  (>>4 => +128)
                    const uint32_t accum = (c > c_begin || k > 0) ? 1u : 0u;
                    tc::umma_bf16(tmem_base, da + 128 * k, db0 + 128 * k, idesc0, accum);
                    if (p.n1w > 0) tc::umma_bf16(tmem_base + 256, da + 128 * k, db1 + 128 * k, idesc1, accum);
                }
                tc::umma_commit(empty + s);
                if (c == c_end - 1) tc::umma_commit(tfull);
            }
            __syncwarp();
            if (++s == stages) { s = 0; ph ^= 1; }
        }
    } else if (GATHER && warp >= 6) {
        // ===== B-row gather producers: warp g copies k-rows 16g..16g+15 of every stage; lane l moves the 16-byte chunk
        // (l & 7) of rows 16g + 4j + (l >> 3), j = 0..3, in each of the nb 64-column boxes (a warp instruction reads
        // 4 rows x 128 contiguous bytes).  The row indices of the next stage are fetched while this one is issued.
        const int g = warp - 6;
        const int c = lane & 7, r4 = lane >> 3;
        const __nv_bfloat16* bcol = p.b_ptr + j0 + c * 8;
        auto rows_of = [&](int chunk, int32_t (&rr)[4]) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t i = (int64_t)chunk * TN_BK + 16 * g + 4 * j + r4;
                rr[j] = i < k_rows ? __ldg(p.list + i) : -1;
            }
        };
        int32_t cur[4], nxt[4];
        rows_of(c_begin, cur);
        int s = 0; uint32_t ph = 0;
        for (int ch = c_begin; ch < c_end; ++ch) {
            if (ch + 1 < c_end) rows_of(ch + 1, nxt);
            const uint32_t sb = tc::smem_u32(smem + (size_t)s * stage_bytes) + a_bytes;
            tc::mbar_wait(empty + s, ph ^ 1);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int row = 16 * g + 4 * j + r4;
                const uint32_t d = sb + row * 128 + ((c ^ (row & 7)) << 4);
                const __nv_bfloat16* src = bcol + (int64_t)(cur[j] < 0 ? 0 : cur[j]) * p.ldb;
                const uint32_t nbytes = cur[j] < 0 ? 0u : 16u;
                for (int b = 0; b < p.nb; ++b) tc::cp_async16_zfill(d + b * TN_BOX_BYTES, src + 64 * b, nbytes);
            }
            tc::cp_async_mbar_arrive_noinc(full + s);
#pragma unroll
            for (int j = 0; j < 4; ++j) cur[j] = nxt[j];
            if (++s == stages) { s = 0; ph ^= 1; }
        }
    } else {
        Epi epi = epi_in;
        const int q = warp & 3;
        const int i = i0 + q * 32 + lane;
        tc::mbar_wait(tfull, 0);
        tc::tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16);
        const int ncols = (p.No - j0 < p.Nt) ? p.No - j0 : p.Nt;
        for (int c = 0; c < ncols; c += 32) {
            float v[32];
            tc::tmem_ld32(t_row + (uint32_t)c, v);
            tc::tmem_ld_wait();
            if (i < p.Mo) epi(i, tap, j0 + c, v, (ncols - c < 32) ? ncols - c : 32);
        }
        tc::tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, 512);
    }
}

// A: (rows x Mo) bf16, leading dimension lda; B: (rows x No) bf16, leading dimension ldb.
template <class Epi>
int gemm_tc_tn(const __nv_bfloat16* A, int Mo, int64_t lda, const __nv_bfloat16* B, int No, int64_t ldb,
               int64_t rows, int taps, const int* shifts, const Epi& epi, cudaStream_t st, const char* what,
               const int32_t* list = nullptr, const int32_t* k_count = nullptr) {
    if (rows == 0 || Mo == 0 || No == 0) return 0;
    NR_REQUIRE(taps >= 1 && taps <= 3, "%s: taps %d outside [1,3]", what, taps);
    NR_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0, "%s: leading dimensions must be multiples of 8 elements", what);
    NR_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "%s: operands must be 16-byte aligned", what);
    TnParams p;
    p.Mo = Mo; p.No = No; p.taps = taps;
    for (int t = 0; t < 3; ++t) p.shift[t] = t < taps ? shifts[t] : 0;
    const int n16 = (No + 15) / 16 * 16;
    p.n_tiles = (n16 + 511) / 512;
    p.Nt = ((n16 + p.n_tiles - 1) / p.n_tiles + 15) / 16 * 16;
    p.n0w = p.Nt < 256 ? p.Nt : 256;
    p.n1w = p.Nt - p.n0w;
    p.nb = (p.Nt + 63) / 64;
    p.m_tiles = (Mo + TN_BM - 1) / TN_BM;
    p.total_chunks = (int)((rows + TN_BK - 1) / TN_BK);
    const int tiles = p.m_tiles * p.n_tiles * taps;
    int splits = sm_count() / tiles;
    if (splits < 1) splits = 1;
    // every split ends in Mo x No atomics: keep at least 8 k-chunks (512 rows) of MMA work behind each of them
    if (splits > (p.total_chunks + 7) / 8) splits = (p.total_chunks + 7) / 8;
    if (splits < 1) splits = 1;
    p.chunks_per_split = (p.total_chunks + splits - 1) / splits;
    splits = (p.total_chunks + p.chunks_per_split - 1) / p.chunks_per_split;
    const size_t stage_bytes = (size_t)(2 + p.nb) * TN_BOX_BYTES;
    int stages = (int)((227 * 1024 - 1024 - 256) / stage_bytes);
    if (stages > 6) stages = 6;
    NR_REQUIRE(stages >= 2, "%s: tile does not fit shared memory", what);
    p.stages = stages;
    p.list = list; p.k_count = k_count; p.b_ptr = B; p.ldb = ldb;
    NR_REQUIRE(!list || (k_count && taps == 1 && p.n_tiles == 1 && ldb >= (int64_t)p.nb * 64),
               "%s: row-gather mode needs a count pointer, taps == 1, one n tile and ldb >= %d", what, p.nb * 64);
    CUtensorMap tmA, tmB;
    if (int e = make_tmap_bf16_2d(&tmA, A, (uint64_t)rows, (uint64_t)Mo, (uint64_t)lda, TN_BK, 64)) return e;
    if (int e = make_tmap_bf16_2d(&tmB, B, (uint64_t)rows, (uint64_t)No, (uint64_t)ldb, TN_BK, 64)) return e;
    const size_t smem = 1024 + stages * stage_bytes + 256;
    if (list) {
        NR_CUDA(cudaFuncSetAttribute(gemm_tn_tc_kernel<Epi, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ProfScope ps(what, st);
        gemm_tn_tc_kernel<Epi, true><<<dim3((unsigned)tiles, (unsigned)splits), TN_THREADS + 32 * TN_GATHER_WARPS, smem, st>>>(
            tmA, tmB, p, epi);
        return check_last(what);
    }
    NR_CUDA(cudaFuncSetAttribute(gemm_tn_tc_kernel<Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ProfScope ps(what, st);
    gemm_tn_tc_kernel<Epi><<<dim3((unsigned)tiles, (unsigned)splits), TN_THREADS, smem, st>>>(tmA, tmB, p, epi);
    return check_last(what);
}

// ---- epilogues: (output row i, tap, first column j, 32 values, valid count) -----------------------
struct TnEpiPlain {       // C[tap] (Mo x ldc) fp32 += acc
    float* C; int64_t ldc; int64_t tap_stride;
    __device__ __forceinline__ void operator()(int i, int tap, int j, const float* v, int nc) const {
        float* p = C + tap * tap_stride + (int64_t)i * ldc + j;
        if ((((uintptr_t)p) & 15) == 0) {       // 16-byte vector reductions: a quarter of the L2 atomic transactions
#pragma unroll
            for (int t = 0; t < 32; t += 4) {
                if (t + 4 <= nc) atomicAdd((float4*)(p + t), make_float4(v[t], v[t + 1], v[t + 2], v[t + 3]));
                else {
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (t + u < nc) atomicAdd(p + t + u, v[t + u]);
                }
            }
            return;
        }
#pragma unroll
        for (int t = 0; t < 32; ++t)
            if (t < nc) atomicAdd(p + t, v[t]);
    }
};

}  // namespace nr
