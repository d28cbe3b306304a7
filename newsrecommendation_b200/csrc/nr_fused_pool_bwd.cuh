// nr_fused_pool_bwd.cuh — backward of the title-level AttentionPooling (model_utils.py:13-31) as ONE
// persistent tcgen05 kernel per 128-row tile (4 titles of pitch 32).  It replaces two launches of round 1:
//   pool_tail_bwd_*   g, Y, E, alpha -> dz (+ dw2, db1, db2)        (read Y, E; wrote dz)
//   enc.pool_dx       dz . W1 + alpha g, output-dropout keep bits -> dY   (read dz back; epilogue-bound GEMM)
//
//   d alpha_s = g . y_s ;  da_s = alpha_s (d alpha_s - sum_t alpha_t d alpha_t) ;  dz[s,q] = da_s w2[q] (1 - e[s,q]^2)
//   dw2 += sum_s da_s e_s ;  db1 += sum_s dz[s,:] ;  db2 += sum_s da_s ;  dY[s,:] = (dz[s,:] . W1 + alpha_s g) [* keep / (1-p)]
//
// Per tile: the E tile (128 x Q bf16) arrives by TMA in the K-major SWIZZLE_128B layout; the compute warps read their
// Y rows straight from global memory for d alpha (the only use of Y here), turn E into dz IN PLACE, and the tile is at
// once the A operand of the tcgen05 GEMM against W1^T blocks (streamed through a 3-stage ring) and the source of the
// TMA store of dz that the dW1 GEMM reads afterwards.  The epilogue (thread = row, TMEM -> registers) adds alpha g,
// applies the saved output-dropout keep bits and writes dY with 256-bit row stores.  E tiles are double buffered, the
// two n passes of a tile use the two TMEM accumulators alternately.
// CTA = 576 threads: warps 0..15 compute (title = warp % 4, the title's four warps split rows / columns / chunks),
// warp 16 = TMA producer, warp 17 = MMA issuer.
#pragma once
#include "nr_fused_enc.cuh"

namespace nr {
namespace fp {

constexpr int FP_THREADS = fe::FE_THREADS, FP_CW = fe::FE_CW, FP_WSTAGES = 3;
constexpr int FP_ETILE = 4 * fe::FE_YBLK;                    // 64 KB: Q <= 256 columns as 4 blocks of 64
constexpr int FP_WSTAGE = 256 * TC_BK * 2;                   // 32 KB: room for the widest W1^T box (BN <= 256 rows)

struct PoolBwdParams {
    int64_t R, n;
    int S, H, Q;
    int BN, G;               // n passes over H: G chunks of BN columns (BN % 16 == 0, <= 256)
    int QB;                  // 64-column blocks of E / dz in use = ceil(Q / 64)
    int tiles;
    const bf16* Y; const float* g; const float* alpha; const float* w2;
    const uint8_t* keep; float dscale;
    bf16* dY; float* dw2; float* db1; float* db2;
};

inline size_t pool_bwd_smem_bytes() { return 1024 + 2 * (size_t)FP_ETILE + FP_WSTAGES * (size_t)FP_WSTAGE + 256 + 1024; }

template <int DUMMY = 0>
__global__ void __launch_bounds__(FP_THREADS, 1)
fused_pool_bwd_kernel(const __grid_constant__ CUtensorMap tmE, const __grid_constant__ CUtensorMap tmW,
                      const __grid_constant__ CUtensorMap tmZ, const PoolBwdParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* etile = smem;                                   // [2][FP_ETILE]
    uint8_t* wring = smem + 2 * FP_ETILE;                    // [FP_WSTAGES][FP_WSTAGE]
    uint64_t* bars = (uint64_t*)(wring + FP_WSTAGES * FP_WSTAGE);
    uint64_t* efull = bars;             // [2] TMA -> compute (E tile landed)
    uint64_t* tfree = bars + 2;         // [2] MMA warp -> TMA (tile buffer reusable: MMAs retired + dz store has read it)
    uint64_t* zfull = bars + 4;         // [2] compute -> MMA (dz complete; FP_CW warps)
    uint64_t* wfull = bars + 6;         // [3] TMA -> MMA
    uint64_t* wempty = bars + 9;        // [3] MMA -> TMA
    uint64_t* tfull = bars + 12;        // [2] MMA -> compute (accumulator ready)
    uint64_t* tempty = bars + 14;       // [2] compute -> MMA (accumulator drained; FP_CW warps)
    uint32_t* tmem_slot = (uint32_t*)(bars + 16);
    float* sda = (float*)((uint8_t*)bars + 256);             // [128] d alpha per row of the tile
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t wbytes = (uint32_t)p.BN * TC_BK * 2;

    if (warp == fe::FE_TMA_WARP && lane == 0) {
        tc::prefetch_tmap(&tmE); tc::prefetch_tmap(&tmW); tc::prefetch_tmap(&tmZ);
        for (int b = 0; b < 2; ++b) {
            tc::mbar_init(efull + b, 1); tc::mbar_init(tfree + b, 2); tc::mbar_init(zfull + b, FP_CW);
            tc::mbar_init(tfull + b, 1); tc::mbar_init(tempty + b, FP_CW);
        }
        for (int s = 0; s < FP_WSTAGES; ++s) { tc::mbar_init(wfull + s, 1); tc::mbar_init(wempty + s, 1); }
        tc::fence_barrier_init();
    }
    if (warp == fe::FE_MMA_WARP) tc::tmem_alloc(tmem_slot, 512);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == fe::FE_TMA_WARP) {
        // ===== TMA producer: E tile of the NEXT tile first, then this tile's W1^T blocks (G passes x QB k blocks) =====
        if (lane == 0) {
            auto load_e = [&](int tile, int it) {
                const int b = it & 1;
                tc::mbar_wait(tfree + b, (((uint32_t)it >> 1) & 1u) ^ 1u);
                tc::mbar_arrive_expect_tx(efull + b, (uint32_t)p.QB * fe::FE_YBLK);
                for (int kb = 0; kb < p.QB; ++kb)
                    tc::tma_load_2d(etile + (size_t)b * FP_ETILE + (size_t)kb * fe::FE_YBLK, &tmE, efull + b, kb * TC_BK, tile * TC_BM);
            };
            int s = 0; uint32_t ph = 0, it = 0;
            if ((int)blockIdx.x < p.tiles) load_e(blockIdx.x, 0);
            for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
                if (tile + (int)gridDim.x < p.tiles) load_e(tile + gridDim.x, it + 1);
                for (int g = 0; g < p.G; ++g)
                    for (int kb = 0; kb < p.QB; ++kb) {
                        tc::mbar_wait(wempty + s, ph ^ 1);
                        tc::mbar_arrive_expect_tx(wfull + s, wbytes);
                        tc::tma_load_2d(wring + (size_t)s * FP_WSTAGE, &tmW, wfull + s, kb * TC_BK, g * p.BN);
                        if (++s == FP_WSTAGES) { s = 0; ph ^= 1; }
                    }
            }
        }
    } else if (warp == fe::FE_MMA_WARP) {
        // ===== MMA issuer =====
        const uint32_t idesc = tc::umma_idesc_bf16(TC_BM, p.BN, 0, 0);
        int s = 0; uint32_t ph = 0, pc = 0, it = 0;
        for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
            const int b = it & 1;
            uint8_t* et = etile + (size_t)b * FP_ETILE;
            tc::mbar_wait(zfull + b, (it >> 1) & 1u);
            tc::tc_fence_after();
            if (lane == 0) {                // dz (bf16) -> HBM for the dW1 GEMM: one bulk tensor store per 64-column block
                for (int kb = 0; kb < p.QB; ++kb) fe::tma_store_2d(&tmZ, et + (size_t)kb * fe::FE_YBLK, kb * TC_BK, tile * TC_BM);
                fe::bulk_commit();
            }
            for (int g = 0; g < p.G; ++g, ++pc) {
                const uint32_t acc = pc & 1u;
                tc::mbar_wait(tempty + acc, ((pc >> 1) & 1u) ^ 1u);
                tc::tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * 256u;
                for (int kb = 0; kb < p.QB; ++kb) {
                    tc::mbar_wait(wfull + s, ph);
                    tc::tc_fence_after();
                    if (lane == 0) {
                        const uint64_t da = tc::umma_desc_sw128(tc::smem_u32(et + (size_t)kb * fe::FE_YBLK), 16, 1024);
                        const uint64_t db = tc::umma_desc_sw128(tc::smem_u32(wring + (size_t)s * FP_WSTAGE), 16, 1024);
                        // columns >= Q of the last block are zero on both sides (TMA zero fill / zero-padded pack)
                        const int ksteps = (p.Q - kb * TC_BK >= TC_BK) ? TC_BK / 16 : (p.Q - kb * TC_BK + 15) / 16;
                        for (int k = 0; k < ksteps; ++k) tc::umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) ? 1u : 0u);
                        tc::umma_commit(wempty + s);
                        if (kb == p.QB - 1) {
                            tc::umma_commit(tfull + acc);
                            if (g == p.G - 1) tc::umma_commit(tfree + b);       // every MMA that reads this tile has retired
                        }
                    }
                    __syncwarp();
                    if (++s == FP_WSTAGES) { s = 0; ph ^= 1; }
                }
            }
            if (lane == 0) { fe::bulk_wait_read0(); tc::mbar_arrive(tfree + b); }  // ... and the dz store has read it
            __syncwarp();
        }
        if (lane == 0) fe::bulk_wait0();
    } else {
        // ===== compute warps =====
        const int q = warp & 3;                         // title of the tile = TMEM lane quadrant
        const int sub = warp >> 2;                      // which of the title's four warps
        const int S = p.S, H = p.H, Q = p.Q;
        const int HC = H >> 3;                          // 16-byte chunks per Y row
        const int wq = ((Q >> 1) + 3) / 4;              // bf16 pairs (words) of an E row per warp (<= 32: Q <= 256)
        const int wcol = sub * wq + lane;               // this lane's word in phase 2
        const bool wlive = lane < wq && 2 * wcol < Q;
        float w2a = 0.f, w2b = 0.f;
        if (wlive) { w2a = __ldg(p.w2 + 2 * wcol); w2b = __ldg(p.w2 + 2 * wcol + 1); }
        float acc_w2a = 0.f, acc_w2b = 0.f, acc_b1a = 0.f, acc_b1b = 0.f, acc_b2 = 0.f;
        // Software pipeline over the CTA's tiles: phases 1 + 2 (d alpha, dz) of tile t+1 run BEFORE the epilogue of tile t,
        // so both MMA passes of tile t (and the TMA traffic of tile t+1) proceed underneath instead of being waited for.
        struct Prev { int64_t row, title; float al; bool live; };
        auto phase12 = [&](int tile, uint32_t it, Prev& out) {
            const int b = it & 1;
            const uint32_t ebase = tc::smem_u32(etile + (size_t)b * FP_ETILE);
            const int64_t m0 = (int64_t)tile * TC_BM;
            const int64_t title = (int64_t)tile * 4 + q;
            const bool live = title < p.n;
            const int64_t row = m0 + q * 32 + lane;     // thread = row phases
            // ---- phase 1: d alpha_s = g . y_s for the 8 rows of this warp (lane = 16-byte chunk of the row, <= 2 per lane)
            float gv[2][8];
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int c = lane + 32 * k;
#pragma unroll
                for (int j = 0; j < 8; ++j) gv[k][j] = 0.f;
                if (live && c < HC) {
                    const float4 x0 = ldg4(p.g + title * H + c * 8), x1 = ldg4(p.g + title * H + c * 8 + 4);
                    gv[k][0] = x0.x; gv[k][1] = x0.y; gv[k][2] = x0.z; gv[k][3] = x0.w;
                    gv[k][4] = x1.x; gv[k][5] = x1.y; gv[k][6] = x1.z; gv[k][7] = x1.w;
                }
            }
#pragma unroll
            for (int hb = 0; hb < 2; ++hb) {            // two batches of 4 rows: 8 independent 16-byte loads per lane in flight
                uint4 yv[4][2];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int s = sub * 8 + hb * 4 + i;
                    const bf16* yr = p.Y + (m0 + q * 32 + s) * (int64_t)H;
#pragma unroll
                    for (int k = 0; k < 2; ++k) {
                        const int c = lane + 32 * k;
                        yv[i][k] = (live && s < S && c < HC) ? __ldg((const uint4*)yr + c) : make_uint4(0u, 0u, 0u, 0u);
                    }
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float part = 0.f;
#pragma unroll
                    for (int k = 0; k < 2; ++k) {
                        float f[8];
                        bf8_to_f(yv[i][k], f);
#pragma unroll
                        for (int j = 0; j < 8; ++j) part = fmaf(gv[k][j], f[j], part);
                    }
                    part = warp_sum(part);
                    if (lane == 0) sda[q * 32 + sub * 8 + hb * 4 + i] = part;
                }
            }
            fe::title_sync(1 + q);
            // ---- da_s = alpha_s (d alpha_s - sum_t alpha_t d alpha_t), lane = row of the title (every warp of the title)
            const float al = (live && lane < S && row < p.R) ? __ldg(p.alpha + row) : 0.f;
            out.row = row; out.title = title; out.al = al; out.live = live;
            const float dal = sda[q * 32 + lane];
            const float cs = warp_sum(al * dal);
            const float da = al * (dal - cs);
            if (sub == 0) acc_b2 += da;                 // summed over the lanes at the end
            // ---- phase 2: dz in place over the E tile (lane = bf16 pair of the row, rows walked)
            tc::mbar_wait(efull + b, (it >> 1) & 1u);
#pragma unroll 4
            for (int s = 0; s < 32; ++s) {
                const float d = __shfl_sync(0xffffffffu, da, s);
                if (wlive) {
                    const uint32_t a = ebase + fe::y_off(q * 32 + s, 2 * wcol);
                    uint32_t u;
                    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(u) : "r"(a));
                    const float2 e = bf2_to_f2(u);
                    const float za = d * w2a * (1.f - e.x * e.x), zb = d * w2b * (1.f - e.y * e.y);
                    acc_w2a = fmaf(d, e.x, acc_w2a); acc_w2b = fmaf(d, e.y, acc_w2b);
                    acc_b1a += za; acc_b1b += zb;
                    asm volatile("st.shared.b32 [%0], %1;" ::"r"(a), "r"(tc::pack_bf16(za, zb)) : "memory");
                }
            }
            tc::fence_proxy_async_smem();               // dz writes -> visible to the MMA / the TMA store
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(zfull + b);
            fe::title_sync(1 + q);                      // sda of this title is free for the next phase 1
        };
        uint32_t pc = 0;
        auto epilogue = [&](const Prev& pv) {
            const int64_t row = pv.row, title = pv.title;
            const bool live = pv.live;
            const float al = pv.al;
            // ---- epilogue of the G passes: dY = (acc + alpha g) * keep * dscale, pad rows zero; thread = row
            const int srow = lane;
            const bool real = live && srow < S && row < p.R;
            const float a_row = real ? al : 0.f;
            for (int g = 0; g < p.G; ++g, ++pc) {
                const uint32_t acc = pc & 1u;
                tc::mbar_wait(tfull + acc, (pc >> 1) & 1u);
                tc::tc_fence_after();
                const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * 256u;
                const int ncols = (H - g * p.BN < p.BN) ? H - g * p.BN : p.BN;
                for (int c = 32 * sub; c < ncols; c += 128) {
                    const int col0 = g * p.BN + c;
                    const int nc = (ncols - c < 32) ? ncols - c : 32;      // multiple of 16 (H % 16 == 0, BN % 16 == 0)
                    float v[32];
                    tc::tmem_ld32(t_row + (uint32_t)c, v);
                    uint32_t kb0 = 0xFFFFFFFFu, kb1 = 0xFFFFFFFFu;
                    if (p.keep && real) {       // one byte per 4 elements, indexed by the padded row
                        const uint32_t* kp = (const uint32_t*)(p.keep + ((row * H + col0) >> 2));
                        kb0 = __ldg(kp);
                        if (nc > 16) kb1 = __ldg(kp + 1);
                    }
                    if (!real) kb0 = kb1 = 0u;
                    const float* gr = p.g + title * H + col0;
                    tc::tmem_ld_wait();
                    if (row < p.R) {
#pragma unroll
                        for (int h2 = 0; h2 < 2; ++h2) {
                            if (16 * h2 < nc) {
                                uint32_t w[8];
#pragma unroll
                                for (int j = 0; j < 4; ++j) {
                                    const float4 gc = real ? ldg4(gr + 16 * h2 + 4 * j) : make_float4(0.f, 0.f, 0.f, 0.f);
                                    const uint32_t nib = ((h2 ? kb1 : kb0) >> (8 * j)) & 0xFu;
                                    const float* vv = v + 16 * h2 + 4 * j;
                                    const float t0 = (nib & 1u) ? fmaf(a_row, gc.x, vv[0]) * p.dscale : 0.f;
                                    const float t1 = (nib & 2u) ? fmaf(a_row, gc.y, vv[1]) * p.dscale : 0.f;
                                    const float t2 = (nib & 4u) ? fmaf(a_row, gc.z, vv[2]) * p.dscale : 0.f;
                                    const float t3 = (nib & 8u) ? fmaf(a_row, gc.w, vv[3]) * p.dscale : 0.f;
                                    w[2 * j] = tc::pack_bf16(t0, t1); w[2 * j + 1] = tc::pack_bf16(t2, t3);
                                }
                                asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p.dY + row * (int64_t)H + col0 + 16 * h2),
                                             "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]) : "memory");
                            }
                        }
                    }
                }
                tc::tc_fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(tempty + acc);
            }
        };
        {
            Prev cur{}, nxt{};
            uint32_t it = 0;
            int tile = blockIdx.x;
            if (tile < p.tiles) phase12(tile, 0, cur);
            for (; tile < p.tiles; tile += gridDim.x, ++it) {
                const int tn = tile + (int)gridDim.x;
                if (tn < p.tiles) phase12(tn, it + 1, nxt);
                epilogue(cur);
                cur = nxt;
            }
        }
        // ---- flush the column sums of this CTA (dw2, db1) and db2
        if (wlive) {
            atomicAdd(p.dw2 + 2 * wcol, acc_w2a); atomicAdd(p.dw2 + 2 * wcol + 1, acc_w2b);
            atomicAdd(p.db1 + 2 * wcol, acc_b1a); atomicAdd(p.db1 + 2 * wcol + 1, acc_b1b);
        }
        acc_b2 = warp_sum(acc_b2);
        if (sub == 0 && lane == 0) atomicAdd(p.db2, acc_b2);
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == fe::FE_MMA_WARP) {
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, 512);
    }
}

// host side ------------------------------------------------------------------------------------------------------
inline bool pool_bwd_supported(int SP, int H, int Q) {
    return SP == 32 && H % 16 == 0 && H >= 16 && H <= 512 && Q % 8 == 0 && Q >= 8 && Q <= 256;
}

// E (R x Q), w1t (H x QK, K-major transposed W1, zero beyond Q), dz (R x Q) bf16; see PoolBwdParams for the rest
inline int launch_fused_pool_bwd(const bf16* E, const bf16* w1t, int QK, bf16* dz, PoolBwdParams p, cudaStream_t st) {
    const int n16 = (p.H + 15) / 16 * 16;
    p.G = (n16 + 255) / 256;
    p.BN = ((n16 + p.G - 1) / p.G + 15) / 16 * 16;
    p.QB = (p.Q + 63) / 64;
    p.tiles = (int)((p.R + TC_BM - 1) / TC_BM);
    NR_REQUIRE(((uintptr_t)p.dY & 31) == 0, "fused pooling backward: dY must be 32-byte aligned");
    CUtensorMap tmE, tmW, tmZ;
    if (int e = make_tmap_bf16_2d(&tmE, E, (uint64_t)p.R, (uint64_t)p.Q, (uint64_t)p.Q, TC_BM, TC_BK)) return e;
    if (int e = make_tmap_bf16_2d(&tmW, w1t, (uint64_t)p.H, (uint64_t)QK, (uint64_t)QK, (uint32_t)p.BN, TC_BK)) return e;
    if (int e = make_tmap_bf16_2d(&tmZ, dz, (uint64_t)p.R, (uint64_t)p.Q, (uint64_t)p.Q, TC_BM, TC_BK)) return e;
    const size_t smem = pool_bwd_smem_bytes();
    auto kern = fused_pool_bwd_kernel<0>;
    NR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = sm_count();
    if (grid > p.tiles) grid = p.tiles;
    ProfScope ps("enc.fused_pool_bwd", st);
    kern<<<grid, FP_THREADS, smem, st>>>(tmE, tmW, tmZ, p);
    return check_last("enc.fused_pool_bwd");
}

}  // namespace fp
}  // namespace nr
