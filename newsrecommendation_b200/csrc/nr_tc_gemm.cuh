// nr_tc_gemm.cuh — persistent, warp-specialised tcgen05 GEMM (bf16 x bf16 -> fp32 in TMEM).
//
//   C[m, n] = sum_{tap < taps} sum_{k < Kt}  A[m + shift[tap], k] * B[n, tap*Kt + k]
//
// A (rows x Kt) and B (N x taps*Kt) are bf16, K-contiguous, fetched by TMA (SWIZZLE_128B boxes of
// 64 k-elements).  taps = 1 is a plain "NT" GEMM; taps = 3 with shifts {-1,0,+1} is the
// Conv1d(k=3,pad=1) implicit GEMM over a token-major activation whose titles are separated by a
// zero row (TMA zero-fills rows outside the tensor).
//
// CTA = 320 threads: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane; owns TMEM),
// warps 2..5 = epilogue group 0, warps 6..9 = epilogue group 1 (TMEM -> registers -> functor).
// Tile 128 x BN (BN <= 256, BN % 16 == 0), two TMEM accumulators (2 x 256 columns): accumulator a is
// drained by epilogue group a, so two tile epilogues and the MMAs of a third tile overlap (the
// epilogue of these K <= 1200 GEMMs costs more cycles than their MMAs).
// Persistent: grid = #SMs, tiles assigned round-robin with n fastest (A tile reuse through L2).
// DYNM instantiations read the number of rows that exist from device memory (the compact backward).
// Two variants that cut L2 reads of the weight operand (a B-resident slice in shared memory; TMA-multicast CTA pairs)
// were built, parity-tested and measured no faster on B200; they are not in the tree.
#pragma once
#include <stdlib.h>
#include <type_traits>
#include "nr_common.cuh"
#include "nr_tc.cuh"

namespace nr {

constexpr int TC_BM = 128, TC_BK = 64, TC_THREADS = 320, TC_MAX_BN = 256;
constexpr int TC_STAGE_PITCH = 40;                                  // epilogue staging: 80-byte rows (conflict-free 16-byte accesses)

struct TcGemmParams {
    int64_t M;        // output rows
    int N;            // output cols
    int BN;           // n tile
    int Kt;           // k extent per tap (multiple of 64 after padding; OOB zero-filled)
    int taps;
    int shift[3];     // A row shift per tap
    int stages;
    int m_tiles, n_tiles;
    // DYNM kernels: the number of rows that exist is read from device memory (*m_count <= M), e.g. the count of
    // real-token rows of a step; tiles past it are never scheduled
    const int32_t* m_count;
};

// per-warp epilogue staging: 32 x TC_STAGE_PITCH bf16 unless the functor asks for more (Epi::kStagePerWarp bytes)
template <class E, class = void> struct EpiStage { static constexpr int bytes = 32 * TC_STAGE_PITCH * 2; };
template <class E> struct EpiStage<E, std::void_t<decltype(E::kStagePerWarp)>> { static constexpr int bytes = E::kStagePerWarp; };
inline size_t tc_stage_bytes(int BN) { return (size_t)TC_BM * TC_BK * 2 + (size_t)BN * TC_BK * 2; }
inline size_t tc_smem_bytes(int BN, int stages, int epi_bytes) { return 1024 + stages * tc_stage_bytes(BN) + 256 + epi_bytes; }
inline bool tc_debug() {
    static const bool on = [] { const char* e = getenv("NR_TC_DEBUG"); return e && e[0] == '1'; }();
    return on;
}
// Epilogue functor protocol (one thread owns one output row of the tile; ALL 32 lanes of the warp call
// every hook, `cx.live` says whether this lane's row exists):
//   begin(cx, row) ; chunk(cx, row, col0, v[32], ncols) for col0 = n0, n0+32, ... ; end(cx, row)
struct EpiCtx {
    __nv_bfloat16* stage;   // this warp's [32][TC_STAGE_PITCH] bf16 staging tile in shared memory
    int lane;
    int64_t row0;           // first row of the warp (lane l owns row0 + l)
    int64_t M;              // rows that exist
    bool live;
    int n0, n1;             // column range [n0, n1) of the tile being drained (set before begin())
};

// A thread-per-row epilogue writes 16 bytes per lane into 32 DIFFERENT rows per store instruction (32
// L1 wavefronts, half-filled sectors).  This helper transposes the warp's 32 x 32 bf16 chunk through
// shared memory so that every store instruction covers 8 rows x 64 contiguous bytes instead.
// second half of store_rows_bf16: the lanes have written their rows into cx.stage (row = lane, 32 bf16)
__device__ __forceinline__ void flush_rows_bf16(const EpiCtx& cx, __nv_bfloat16* C, int64_t ldc, int col0, int nc);

__device__ __forceinline__ void store_rows_bf16(const EpiCtx& cx, __nv_bfloat16* C, int64_t ldc, int col0, const float* o,
                                                int nc) {
    uint4* srow = (uint4*)(cx.stage + cx.lane * TC_STAGE_PITCH);
#pragma unroll
    for (int j = 0; j < 4; ++j)
        srow[j] = make_uint4(tc::pack_bf16(o[8 * j], o[8 * j + 1]), tc::pack_bf16(o[8 * j + 2], o[8 * j + 3]),
                             tc::pack_bf16(o[8 * j + 4], o[8 * j + 5]), tc::pack_bf16(o[8 * j + 6], o[8 * j + 7]));
    flush_rows_bf16(cx, C, ldc, col0, nc);
}

__device__ __forceinline__ void flush_rows_bf16(const EpiCtx& cx, __nv_bfloat16* C, int64_t ldc, int col0, int nc) {
    __syncwarp();
    const int c = (cx.lane & 3) * 8;
    if (nc == 32 && cx.row0 + 32 <= cx.M) {          // full chunk: straight-line code
        const __nv_bfloat16* src = cx.stage + (cx.lane >> 2) * TC_STAGE_PITCH + c;
        __nv_bfloat16* dst = C + (cx.row0 + (cx.lane >> 2)) * ldc + col0 + c;
#pragma unroll
        for (int j = 0; j < 4; ++j) *(uint4*)(dst + (int64_t)8 * j * ldc) = *(const uint4*)(src + 8 * j * TC_STAGE_PITCH);
        __syncwarp();
        return;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int r = 8 * j + (cx.lane >> 2);
        if (cx.row0 + r < cx.M && c < nc) {
            const __nv_bfloat16* src = cx.stage + r * TC_STAGE_PITCH + c;
            __nv_bfloat16* dst = C + (cx.row0 + r) * ldc + col0 + c;
            if (c + 8 <= nc) *(uint4*)dst = *(const uint4*)src;
            else
                for (int t = 0; t < nc - c; ++t) dst[t] = src[t];
        }
    }
    __syncwarp();
}

// EW = epilogue warps: 8 (two groups of 4, one per accumulator) or 16 (two groups per accumulator, each draining
// half of the tile's 32-column chunks; 18 warps -> at most 112 registers per thread).  Functors that carry per-row
// state across chunks (Epi::kRowState) need the whole row in one thread and stay at EW = 8.
template <class Epi, int EW, bool DYNM = false>
__global__ void __launch_bounds__(64 + 32 * EW, 1)   // EW = 8: 10 warps, 3 on one SM sub-partition -> 168 registers is the ceiling
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const TcGemmParams p, const Epi epi_in) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // 1024-B alignment for SWIZZLE_128B atoms
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const int stages = p.stages;
    const uint32_t a_bytes = TC_BM * TC_BK * 2, b_bytes = (uint32_t)p.BN * TC_BK * 2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t M = DYNM ? (int64_t)__ldg(p.m_count) : p.M;
    const int total_tiles = (DYNM ? (int)((M + TC_BM - 1) / TC_BM) : p.m_tiles) * p.n_tiles;
    const int kc = p.Kt / TC_BK;
    const int k_iters = p.taps * kc;
    const uint32_t stage_bytes = a_bytes + b_bytes;
    uint8_t* ring = smem;                                           // operand ring
    uint64_t* bars = (uint64_t*)(ring + (size_t)stages * stage_bytes);
    uint64_t* full = bars;                  // [stages]   TMA -> MMA
    uint64_t* empty = bars + stages;        // [stages]   MMA -> TMA
    uint64_t* tfull = bars + 2 * stages;    // [2]        MMA -> epilogue
    uint64_t* tempty = tfull + 2;           // [2]        epilogue -> MMA
    uint32_t* tmem_slot = (uint32_t*)(tempty + 2);
    __nv_bfloat16* epi_stage = (__nv_bfloat16*)((uint8_t*)bars + 256);      // 8 x [32][TC_STAGE_PITCH]
    // tile sequence of this CTA: it-th tile -> (m tile, n tile)
    auto tile_of = [&](int it, int& mt, int& nt) -> bool {
        const int64_t tile = (int64_t)blockIdx.x + (int64_t)it * gridDim.x;
        if (tile >= total_tiles) return false;
        mt = (int)(tile / p.n_tiles);
        nt = (int)(tile % p.n_tiles);
        return true;
    };

    if (warp == 0 && lane == 0) {
        tc::prefetch_tmap(&tmA);
        tc::prefetch_tmap(&tmB);
        for (int s = 0; s < stages; ++s) { tc::mbar_init(full + s, 1); tc::mbar_init(empty + s, 1); }
        for (int a = 0; a < 2; ++a) { tc::mbar_init(tfull + a, 1); tc::mbar_init(tempty + a, EW / 2); }
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            int mt, nt;
            for (int it = 0; tile_of(it, mt, nt); ++it) {
                const int m0 = mt * TC_BM, n0 = nt * p.BN;
                for (int i = 0; i < k_iters; ++i) {
                    const int tap = i / kc, kk = (i - tap * kc) * TC_BK;
                    tc::mbar_wait(empty + s, ph ^ 1);
                    uint8_t* sa = ring + (size_t)s * stage_bytes;
                    tc::mbar_arrive_expect_tx(full + s, stage_bytes);
                    tc::tma_load_2d(sa, &tmA, full + s, kk, m0 + p.shift[tap]);
                    tc::tma_load_2d(sa + a_bytes, &tmB, full + s, i * TC_BK, n0);
                    if (++s == stages) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        const uint32_t idesc = tc::umma_idesc_bf16(TC_BM, p.BN, 0, 0);
        int s = 0; uint32_t ph = 0;
        int acc = 0; uint32_t aph = 0;
        int mt, nt;
        for (int it = 0; tile_of(it, mt, nt); ++it) {
            tc::mbar_wait(tempty + acc, aph ^ 1);
            tc::tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)acc * TC_MAX_BN;
            for (int i = 0; i < k_iters; ++i) {
                tc::mbar_wait(full + s, ph);
                tc::tc_fence_after();
                if (lane == 0) {
                    const uint32_t sa = tc::smem_u32(ring + (size_t)s * stage_bytes);
                    const uint64_t da = tc::umma_desc_sw128(sa, 16, 1024);
                    const uint64_t db = tc::umma_desc_sw128(sa + a_bytes, 16, 1024);
#pragma unroll
                    for (int k = 0; k < TC_BK / 16; ++k)   // +32 B per 16-element k step (>>4 => +2)
                        tc::umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (i | k) ? 1u : 0u);
                    tc::umma_commit(empty + s);                      // smem slot free when MMAs retire
                    if (i == k_iters - 1) tc::umma_commit(tfull + acc);   // accumulator ready
                }
                __syncwarp();
                if (++s == stages) { s = 0; ph ^= 1; }
            }
            if (++acc == 2) { acc = 0; aph ^= 1; }
        }
    } else {
        // ===== epilogue: group g = warps 2+4g .. 5+4g drains accumulator g; warp w may touch TMEM
        // lanes [32*(w%4), +32) =====
        Epi epi = epi_in;
        const int q = warp & 3;
        const int grp4 = (warp - 2) >> 2;            // EW = 8: 0..1; EW = 16: 0..3
        const int grp = grp4 & 1;                    // accumulator
        const int half = grp4 >> 1;                  // which half of the chunks (EW = 16)
        int mt, nt;
        for (int it = grp; tile_of(it, mt, nt); it += 2) {
            const int64_t m0 = (int64_t)mt * TC_BM;
            const int n0 = nt * p.BN;
            const int64_t row = m0 + q * 32 + lane;
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)grp * TC_MAX_BN;
            const int ncols = (p.N - n0 < p.BN) ? p.N - n0 : p.BN;
            int c_lo = 0, c_hi = ncols;
            if (EW == 16) {
                const int split = ((ncols + 31) / 32 + 1) / 2 * 32;      // first half takes the extra chunk
                if (half) c_lo = split < ncols ? split : ncols; else c_hi = split < ncols ? split : ncols;
            }
            EpiCtx cx;
            cx.stage = (__nv_bfloat16*)((uint8_t*)epi_stage + (warp - 2) * EpiStage<Epi>::bytes);
            cx.lane = lane;
            cx.row0 = m0 + q * 32;
            cx.M = M;
            cx.live = row < M;
            cx.n0 = n0 + c_lo; cx.n1 = n0 + c_hi;
            // begin() before the accumulator wait: its global loads (per-row scalars, the first chunk's
            // operands) are in flight while the MMA warp finishes the tile
            epi.begin(cx, row);
            tc::mbar_wait(tfull + grp, (uint32_t)(it >> 1) & 1u);
            tc::tc_fence_after();
            // two register buffers: the TMEM load of chunk i+1 is in flight while chunk i is processed
            if (EW == 16) {
                // 18 warps leave 96 registers per thread: one TMEM buffer; the other three epilogue groups of the SM
                // sub-partition cover the TMEM round trip instead of a second register buffer
                float va[32];
                for (int c = c_lo; c < c_hi; c += 32) {
                    tc::tmem_ld32(t_row + (uint32_t)c, va);
                    tc::tmem_ld_wait();
                    epi.chunk(cx, row, n0 + c, va, (c_hi - c < 32) ? c_hi - c : 32);
                }
            } else {
                float va[32], vb[32];
                if (c_lo < c_hi) tc::tmem_ld32(t_row + (uint32_t)c_lo, va);
                for (int c = c_lo; c < c_hi; c += 64) {
                    tc::tmem_ld_wait();
                    const bool has_b = c + 32 < c_hi;
                    if (has_b) tc::tmem_ld32(t_row + (uint32_t)(c + 32), vb);
                    epi.chunk(cx, row, n0 + c, va, (c_hi - c < 32) ? c_hi - c : 32);
                    if (has_b) {
                        tc::tmem_ld_wait();
                        if (c + 64 < c_hi) tc::tmem_ld32(t_row + (uint32_t)(c + 64), va);
                        epi.chunk(cx, row, n0 + c + 32, vb, (c_hi - c - 32 < 32) ? c_hi - c - 32 : 32);
                    }
                }
            }
            epi.end(cx, row);
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(tempty + grp);
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, 512);
    }
}

// ---- host launcher -----------------------------------------------------------------------------
// A: (a_rows x Kt_real) bf16 with leading dimension lda; B: (N x b_cols) bf16 with leading dimension ldb
// (b_cols = taps * Kt_pad).  Kt_pad = Kt_real rounded up to 64: A columns past Kt_real are zero filled
// by TMA, so B's padding columns may hold anything finite (they multiply zeros).
template <class Epi, int EW = 8>
int gemm_tc(const __nv_bfloat16* A, int64_t a_rows, int Kt_real, int64_t lda, const __nv_bfloat16* B, int N,
            int64_t ldb, int taps, const int* shifts, int64_t M, const Epi& epi, cudaStream_t st,
            const char* what, const int32_t* m_count = nullptr) {
    if (M == 0 || N == 0) return 0;
    NR_REQUIRE(taps >= 1 && taps <= 3, "%s: taps %d outside [1,3]", what, taps);
    NR_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0, "%s: leading dimensions must be multiples of 8 elements", what);
    NR_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "%s: operands must be 16-byte aligned", what);
    TcGemmParams p;
    p.M = M; p.N = N; p.taps = taps;
    p.m_count = m_count;
    p.Kt = (Kt_real + TC_BK - 1) / TC_BK * TC_BK;
    for (int t = 0; t < 3; ++t) p.shift[t] = t < taps ? shifts[t] : 0;
    // n tile: the widest multiple of 16 <= 256 that splits N evenly enough
    const int n16 = (N + 15) / 16 * 16;
    int nt = (n16 + TC_MAX_BN - 1) / TC_MAX_BN;
    p.BN = ((n16 + nt - 1) / nt + 15) / 16 * 16;
    p.n_tiles = (N + p.BN - 1) / p.BN;
    p.m_tiles = (int)((M + TC_BM - 1) / TC_BM);
    const size_t budget = 227 * 1024 - 1024 - 256 - (size_t)EW * EpiStage<Epi>::bytes;
    int stages = (int)(budget / tc_stage_bytes(p.BN));
    if (stages > 8) stages = 8;
    NR_REQUIRE(stages >= 2, "%s: tile does not fit shared memory", what);
    int grid = sm_count();
    const int tiles = p.m_tiles * p.n_tiles;
    if (grid > tiles) grid = tiles;
    p.stages = stages;
    CUtensorMap tmA, tmB;
    if (int e = make_tmap_bf16_2d(&tmA, A, (uint64_t)a_rows, (uint64_t)Kt_real, (uint64_t)lda, TC_BM, TC_BK)) return e;
    if (int e = make_tmap_bf16_2d(&tmB, B, (uint64_t)N, (uint64_t)taps * p.Kt, (uint64_t)ldb, (uint32_t)p.BN, TC_BK))
        return e;
    const size_t smem = tc_smem_bytes(p.BN, stages, EW * EpiStage<Epi>::bytes);
    if (m_count) {
        NR_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<Epi, EW, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ProfScope ps(what, st);
        gemm_tc_kernel<Epi, EW, true><<<grid, 64 + 32 * EW, smem, st>>>(tmA, tmB, p, epi);
        return check_last(what);
    }
    NR_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<Epi, EW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (tc_debug())
        fprintf(stderr, "[nr] %s M=%lld N=%d K=%d taps=%d BN=%d tiles=%dx%d stages=%d grid=%d smem=%zu\n", what,
                (long long)M, N, Kt_real, taps, p.BN, p.m_tiles, p.n_tiles, stages, grid, smem);
    ProfScope ps(what, st);
    gemm_tc_kernel<Epi, EW><<<grid, 64 + 32 * EW, smem, st>>>(tmA, tmB, p, epi);
    return check_last(what);
}

// ---- epilogues -----------------------------------------------------------------------------------
// All register-array indexing is compile-time (fully unrolled, predicated tails) so v[] never
// spills to local memory.  bias pointers must be 16-byte aligned (col0 is a multiple of 16).
__device__ __forceinline__ void load_bias32(const float* bias, int col0, int nc, float* b) {
    if (bias && nc == 32) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 t = ldg4(bias + col0 + 4 * j);
            b[4 * j] = t.x; b[4 * j + 1] = t.y; b[4 * j + 2] = t.z; b[4 * j + 3] = t.w;
        }
        return;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        if (bias && 4 * j < nc) {
            if (4 * j + 4 <= nc) t = ldg4(bias + col0 + 4 * j);
            else {
                t.x = __ldg(bias + col0 + 4 * j);
                if (4 * j + 1 < nc) t.y = __ldg(bias + col0 + 4 * j + 1);
                if (4 * j + 2 < nc) t.z = __ldg(bias + col0 + 4 * j + 2);
            }
        }
        b[4 * j] = t.x; b[4 * j + 1] = t.y; b[4 * j + 2] = t.z; b[4 * j + 3] = t.w;
    }
}
// 32 fp32 values -> bf16 row segment (16-byte stores where the tail allows)
__device__ __forceinline__ void store_bf16_32(__nv_bfloat16* out, const float* v, int nc) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (8 * j + 8 <= nc) {
            *(uint4*)(out + 8 * j) = make_uint4(tc::pack_bf16(v[8 * j], v[8 * j + 1]), tc::pack_bf16(v[8 * j + 2], v[8 * j + 3]),
                                                tc::pack_bf16(v[8 * j + 4], v[8 * j + 5]), tc::pack_bf16(v[8 * j + 6], v[8 * j + 7]));
        } else {
#pragma unroll
            for (int t = 0; t < 8; ++t)
                if (8 * j + t < nc) out[8 * j + t] = __float2bfloat16(v[8 * j + t]);
        }
    }
}
// C (bf16, leading dimension ldc, ldc % 8 == 0) = acc + bias[col]
struct TcEpiBiasBf16 {
    __nv_bfloat16* C; int64_t ldc; const float* bias; int N;
    __device__ __forceinline__ void begin(const EpiCtx&, int64_t) {}
    __device__ __forceinline__ void end(const EpiCtx&, int64_t) {}
    __device__ __forceinline__ void chunk(const EpiCtx& cx, int64_t, int col0, const float* v, int nc) {
        float b[32], o[32];
        load_bias32(bias, col0, nc, b);
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] = v[j] + b[j];
        store_rows_bf16(cx, C, ldc, col0, o, nc);
    }
};
// C (bf16) = acc   (bias already inside the product: ones column of A times a bias column of B)
struct TcEpiStoreBf16 {
    __nv_bfloat16* C; int64_t ldc;
    __device__ __forceinline__ void begin(const EpiCtx&, int64_t) {}
    __device__ __forceinline__ void end(const EpiCtx&, int64_t) {}
    __device__ __forceinline__ void chunk(const EpiCtx& cx, int64_t, int col0, const float* v, int nc) {
        store_rows_bf16(cx, C, ldc, col0, v, nc);
    }
};
// C (fp32, ldc % 4 == 0) = acc + bias[col]
struct TcEpiBiasF32 {
    float* C; int64_t ldc; const float* bias; int N;
    __device__ __forceinline__ void begin(const EpiCtx&, int64_t) {}
    __device__ __forceinline__ void end(const EpiCtx&, int64_t) {}
    __device__ __forceinline__ void chunk(const EpiCtx& cx, int64_t row, int col0, const float* v, int nc) {
        if (!cx.live) return;
        float b[32];
        load_bias32(bias, col0, nc, b);
        float* out = C + row * ldc + col0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (4 * j + 4 <= nc) {
                *(float4*)(out + 4 * j) = make_float4(v[4 * j] + b[4 * j], v[4 * j + 1] + b[4 * j + 1],
                                                      v[4 * j + 2] + b[4 * j + 2], v[4 * j + 3] + b[4 * j + 3]);
            } else {
#pragma unroll
                for (int t = 0; t < 4; ++t)
                    if (4 * j + t < nc) out[4 * j + t] = v[4 * j + t] + b[4 * j + t];
            }
        }
    }
};

}  // namespace nr
