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
#pragma once
#include "nr_common.cuh"
#include "nr_tc.cuh"

namespace nr {

constexpr int TC_BM = 128, TC_BK = 64, TC_THREADS = 320, TC_MAX_BN = 256;

struct TcGemmParams {
    int64_t M;        // output rows
    int N;            // output cols
    int BN;           // n tile
    int Kt;           // k extent per tap (multiple of 64 after padding; OOB zero-filled)
    int taps;
    int shift[3];     // A row shift per tap
    int stages;
    int m_tiles, n_tiles;
};

inline size_t tc_stage_bytes(int BN) { return (size_t)TC_BM * TC_BK * 2 + (size_t)BN * TC_BK * 2; }
inline size_t tc_smem_bytes(int BN, int stages) { return 1024 + stages * tc_stage_bytes(BN) + 256; }

// Epilogue functor protocol (one thread owns one output row of the tile):
//   begin(row) ; chunk(row, col0, v[32], ncols) for col0 = n0, n0+32, ... ; end(row)
template <class Epi>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const TcGemmParams p, const Epi epi_in) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // 1024-B alignment for SWIZZLE_128B atoms
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const int stages = p.stages;
    const uint32_t a_bytes = TC_BM * TC_BK * 2, b_bytes = (uint32_t)p.BN * TC_BK * 2;
    const uint32_t stage_bytes = a_bytes + b_bytes;
    uint64_t* bars = (uint64_t*)(smem + (size_t)stages * stage_bytes);
    uint64_t* full = bars;                  // [stages]   TMA -> MMA
    uint64_t* empty = bars + stages;        // [stages]   MMA -> TMA
    uint64_t* tfull = bars + 2 * stages;    // [2]        MMA -> epilogue
    uint64_t* tempty = tfull + 2;           // [2]        epilogue -> MMA
    uint32_t* tmem_slot = (uint32_t*)(tempty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total_tiles = p.m_tiles * p.n_tiles;
    const int kc = p.Kt / TC_BK;
    const int k_iters = p.taps * kc;

    if (warp == 0 && lane == 0) {
        tc::prefetch_tmap(&tmA);
        tc::prefetch_tmap(&tmB);
        for (int s = 0; s < stages; ++s) { tc::mbar_init(full + s, 1); tc::mbar_init(empty + s, 1); }
        for (int a = 0; a < 2; ++a) { tc::mbar_init(tfull + a, 1); tc::mbar_init(tempty + a, 4); }
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
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const int m0 = (tile / p.n_tiles) * TC_BM, n0 = (tile % p.n_tiles) * p.BN;
                for (int i = 0; i < k_iters; ++i) {
                    const int tap = i / kc, kk = (i - tap * kc) * TC_BK;
                    tc::mbar_wait(empty + s, ph ^ 1);
                    uint8_t* sa = smem + (size_t)s * stage_bytes;
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
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            tc::mbar_wait(tempty + acc, aph ^ 1);
            tc::tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)acc * TC_MAX_BN;
            for (int i = 0; i < k_iters; ++i) {
                tc::mbar_wait(full + s, ph);
                tc::tc_fence_after();
                if (lane == 0) {
                    const uint32_t sa = tc::smem_u32(smem + (size_t)s * stage_bytes);
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
        const int grp = (warp - 2) >> 2;
        for (int it = grp;; it += 2) {
            const int64_t tile = (int64_t)blockIdx.x + (int64_t)it * gridDim.x;
            if (tile >= total_tiles) break;
            const int64_t m0 = (tile / p.n_tiles) * TC_BM;
            const int n0 = (int)(tile % p.n_tiles) * p.BN;
            const int64_t row = m0 + q * 32 + lane;
            tc::mbar_wait(tfull + grp, (uint32_t)(it >> 1) & 1u);
            tc::tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)grp * TC_MAX_BN;
            const int ncols = (p.N - n0 < p.BN) ? p.N - n0 : p.BN;
            const bool live = row < p.M;
            if (live) epi.begin(row);
            for (int c = 0; c < ncols; c += 32) {
                float v[32];
                tc::tmem_ld32(t_row + (uint32_t)c, v);
                tc::tmem_ld_wait();
                if (live) epi.chunk(row, n0 + c, v, (ncols - c < 32) ? ncols - c : 32);
            }
            if (live) epi.end(row);
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
template <class Epi>
int gemm_tc(const __nv_bfloat16* A, int64_t a_rows, int Kt_real, int64_t lda, const __nv_bfloat16* B, int N,
            int64_t ldb, int taps, const int* shifts, int64_t M, const Epi& epi, cudaStream_t st,
            const char* what) {
    if (M == 0 || N == 0) return 0;
    NR_REQUIRE(taps >= 1 && taps <= 3, "%s: taps %d outside [1,3]", what, taps);
    NR_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0, "%s: leading dimensions must be multiples of 8 elements", what);
    NR_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "%s: operands must be 16-byte aligned", what);
    TcGemmParams p;
    p.M = M; p.N = N; p.taps = taps;
    p.Kt = (Kt_real + TC_BK - 1) / TC_BK * TC_BK;
    for (int t = 0; t < 3; ++t) p.shift[t] = t < taps ? shifts[t] : 0;
    // n tile: the widest multiple of 16 <= 256 that splits N evenly enough
    const int n16 = (N + 15) / 16 * 16;
    int nt = (n16 + TC_MAX_BN - 1) / TC_MAX_BN;
    p.BN = ((n16 + nt - 1) / nt + 15) / 16 * 16;
    p.n_tiles = (N + p.BN - 1) / p.BN;
    p.m_tiles = (int)((M + TC_BM - 1) / TC_BM);
    int stages = (int)((227 * 1024 - 1024 - 256) / tc_stage_bytes(p.BN));
    if (stages > 8) stages = 8;
    p.stages = stages;
    NR_REQUIRE(stages >= 2, "%s: tile does not fit shared memory", what);
    CUtensorMap tmA, tmB;
    if (int e = make_tmap_bf16_2d(&tmA, A, (uint64_t)a_rows, (uint64_t)Kt_real, (uint64_t)lda, TC_BM, TC_BK)) return e;
    if (int e = make_tmap_bf16_2d(&tmB, B, (uint64_t)N, (uint64_t)taps * p.Kt, (uint64_t)ldb, (uint32_t)p.BN, TC_BK))
        return e;
    const size_t smem = tc_smem_bytes(p.BN, stages);
    NR_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = sm_count();
    const int tiles = p.m_tiles * p.n_tiles;
    if (grid > tiles) grid = tiles;
    ProfScope ps(what, st);
    gemm_tc_kernel<Epi><<<grid, TC_THREADS, smem, st>>>(tmA, tmB, p, epi);
    return check_last(what);
}

// ---- epilogues -----------------------------------------------------------------------------------
// All register-array indexing is compile-time (fully unrolled, predicated tails) so v[] never
// spills to local memory.  bias pointers must be 16-byte aligned (col0 is a multiple of 16).
__device__ __forceinline__ void load_bias32(const float* bias, int col0, int nc, float* b) {
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
    __device__ __forceinline__ void begin(int64_t) {}
    __device__ __forceinline__ void end(int64_t) {}
    __device__ __forceinline__ void chunk(int64_t row, int col0, const float* v, int nc) {
        float b[32], o[32];
        load_bias32(bias, col0, nc, b);
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] = v[j] + b[j];
        store_bf16_32(C + row * ldc + col0, o, nc);
    }
};
// C (fp32, ldc % 4 == 0) = acc + bias[col]
struct TcEpiBiasF32 {
    float* C; int64_t ldc; const float* bias; int N;
    __device__ __forceinline__ void begin(int64_t) {}
    __device__ __forceinline__ void end(int64_t) {}
    __device__ __forceinline__ void chunk(int64_t row, int col0, const float* v, int nc) {
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
