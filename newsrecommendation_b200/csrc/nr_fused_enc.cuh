// nr_fused_enc.cuh — the NRMS title encoder forward as ONE persistent tcgen05 kernel per 128-row tile
// (4 titles of pitch 32):
//
//     Xb tile --TMA--> QKV projection (tcgen05.mma, fp32 accumulators in TMEM, 4 heads = 240 columns per pass)
//        --> per-head exp-attention by the warps that own the title's TMEM lanes: Q, K, V fragments are read
//            STRAIGHT FROM TMEM in mma.sync fragment layout (tcgen05.ld.16x256b == the m16n8 C-fragment layout,
//            profiles/r2/tmem_frag_probe.txt), S = QK^T / P V on warp-level MMAs, softmax-like normaliser with quad
//            shuffles (model_utils.py:47-54: exp, no max subtraction, eps inside the denominator), output dropout
//        --> context tile Y (128 x H bf16) written into SHARED memory in the K-major SWIZZLE_128B layout
//        --> AttentionPooling fc1 (tcgen05.mma straight from that tile) + tanh + fc2 dot (model_utils.py:21-23)
//        --> exp / normalise over the title's 32 lanes (warp shuffles) and the weighted sum over the smem tile
//            (model_utils.py:24-30) --> out (n, H) fp32.
//
// Reference path replaced: src/model/NRMS.py:28-35 (NewsEncoder.forward), model_utils.py:78-95, :39-55, :13-31.
// In training mode the kernel also writes what the backward needs — qkv (bf16, head-major), Y (one TMA store of the
// smem tile), E = tanh(fc1), alpha and the output-dropout keep bits — each exactly once; in inference mode
// (TRAIN = false: the news-vector cache of src/main.py:190-196) nothing but `out` leaves the SM.  The un-fused
// pipeline (nr_encoder.cu) wrote qkv, read it back, wrote Y, read it twice and wrote E / logits from three launches.
//
// CTA = 576 threads: warps 0..15 = compute, warp 16 = TMA producer, warp 17 = MMA issuer (owns the 512 TMEM columns).
// Compute warp w may touch TMEM lanes [32 (w % 4), +32) = title (w % 4) of the tile; the FOUR warps of a title take
// one head of a pass each (and a quarter of the columns of every store / the fc1 epilogue / the pooling sum).  The
// first version had two warps per title: 2 warps per scheduler could not hide the MUFU / HMMA / shuffle latencies of
// the attention chain (issue slots 36 % busy, the kernel no faster than the separate launches).
// Accumulators: two 256-column TMEM buffers used strictly alternately by the G head-group passes and the fc1 pass
// of every tile, so pass i+1's MMAs run while pass i is drained.
#pragma once
#include "nr_attn_mma.cuh"
#include "nr_tc_gemm.cuh"

namespace nr {
namespace fe {

#ifdef NR_FE_PROBES
#define FE_DBG(p, bit) ((p).dbg & (bit))
#else
#define FE_DBG(p, bit) 0
#endif

constexpr int FE_CW = 16;                                    // compute warps
constexpr int FE_THREADS = 64 + 32 * FE_CW, FE_STAGES = 2;
// The warp scheduler prefers the HIGHEST warp id among eligible warps (B300_MICROARCH.md, "hi-wid-first").  With the
// producer and the MMA issuer as warps 0 / 1 they only got issue slots when every compute warp of their scheduler was
// stalled: the TMA / MMA pipeline ran in the gaps of the compute warps instead of beside them (measured: every
// microsecond of compute added to the kernel time, nothing overlapped).  They are the two top warps instead.
constexpr int FE_TMA_WARP = FE_CW, FE_MMA_WARP = FE_CW + 1;
constexpr int FE_A_BYTES = TC_BM * TC_BK * 2;                // 16 KB
constexpr int FE_B_BYTES = 240 * TC_BK * 2;                  // 30 KB: the widest B box (240 rows)
constexpr int FE_STAGE_BYTES = FE_A_BYTES + FE_B_BYTES;      // 46 KB (multiple of 1024)
constexpr int FE_YBLK = TC_BM * TC_BK * 2;                   // one 64-column block of the Y tile

template <int DK>
struct FCfg {
    static_assert(DK == 16 || DK == 20, "head width outside the compiled set of the fused encoder");
    static constexpr int KF = DK / 16, KH = (DK % 16) ? 1 : 0;
    static constexpr int NB = (DK + 7) / 8;                  // 8-column blocks per q / k / v slice
    static constexpr int HG = 4;                             // heads per pass = warps per title
    static constexpr int BN = HG * 3 * DK;                   // accumulator columns per pass (192 or 240; % 16 == 0)
};

struct FusedParams {
    int64_t R, n;            // padded rows (n * 32), titles
    int S, H, Q, heads;
    int G;                   // passes per tile: heads / HG (MHSA) or the n chunks of the Conv1d output (CONV)
    int BN;                  // accumulator columns per pass (MHSA: 4 heads x 3 dk; CONV: H split into G chunks, % 16 == 0)
    int taps;                // CONV: 3 (A row shifts -1, 0, +1 = Conv1d(k=3, pad=1), NAML.py:27-32); MHSA: 1
    int Kp;                  // CONV: k pitch of one tap in the packed weights
    int KB;                  // 64-column k blocks of Xb / the packed QKV weights
    int HKB;                 // 64-column blocks of the Y tile = ceil(H / 64)
    int NQ;                  // fc1 accumulator columns = round_up(Q, 16) (weight rows >= Q are zero filled by TMA)
    int tiles;
    float* out;              // (n, H)
    const float* mask;       // (n, S) or NULL: key mask and pooling mask (NRMS.py:56-57)
    const float* b1; const float* w2; const float* b2;
    float scale;             // 1 / sqrt(dk)
    // training-mode saves
    bf16* qkv; bf16* E; float* alpha; uint8_t* keep;
    uint32_t thr; float dscale; uint64_t seed;
    int dbg;                 // timing probes, only in builds with -DNR_FE_PROBES (make EXTRA=-DNR_FE_PROBES) and
                             // NR_FE_DBG set: 1 skip attention, 2 skip qkv / E stores, 4 skip dropout draws, 8 skip the
                             // pooling sum, 16 pure delay of (dbg >> 8) x 256 cycles per pass, 32 no loads / no MMAs.
                             // They produce WRONG results by design and are compiled out of the product.
};

inline size_t fused_smem_bytes(int HKB) {
    return 1024 + (size_t)FE_STAGES * FE_STAGE_BYTES + (size_t)HKB * FE_YBLK + 256 + 2048 + 2048;
}

__device__ __forceinline__ float tanh_fast(float x) {      // MUFU.TANH, rel. error ~2^-11 (< bf16 rounding)
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ---- tcgen05.ld.16x256b: 16 TMEM lanes x 8 columns per repeat; register 4b+j of lane l = (row l/4 + 8 (j>>1),
// column 8b + 2 (l%4) + (j&1)): the mma.m16n8 accumulator layout -----------------------------------------------
__device__ __forceinline__ void tmem_ld_frag1(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_frag2(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr) : "memory");
}
template <int NB>
__device__ __forceinline__ void tmem_ld_blocks(uint32_t taddr, uint32_t* r) {
    static_assert(NB == 2 || NB == 3, "8-column blocks per head slice");
    tmem_ld_frag2(taddr, r);
    if (NB == 3) tmem_ld_frag1(taddr + 16, r + 8);
}
__device__ __forceinline__ uint32_t packu(uint32_t a, uint32_t b) { return tc::pack_bf16(__uint_as_float(a), __uint_as_float(b)); }

// TMA store of one [128 rows x 64 columns] swizzled block of the Y tile (rows / columns outside the tensor are clipped)
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"((uint64_t)m), "r"(tc::smem_u32(src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void title_sync(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }

// byte offset of element (row, col) of the Y tile: 64-column blocks of [128 rows][128 B], 16-byte chunks XOR-swizzled
// by the row (the layout TMA SWIZZLE_128B writes and the K-major UMMA descriptor reads)
__device__ __forceinline__ uint32_t y_off(int row, int col) {
    return (uint32_t)((col >> 6) * FE_YBLK + row * 128 + ((((col & 63) >> 3) ^ (row & 7)) << 4) + (col & 7) * 2);
}
// 32 fp32 accumulator values of this lane's row -> 64 contiguous bytes of bf16 as two 256-bit stores (sm_100:
// STG.E.ENL2.256; one L1 wavefront per 32-byte sector and row instead of two half-sector 128-bit stores).  dst must be
// 32-byte aligned (qkv rows: pitch 3H * 2 bytes, H % 16 == 0; column chunks of 32).
__device__ __forceinline__ void store_row32_bf16_v8(bf16* dst, const float* v, int nc) {
#pragma unroll
    for (int j = 0; j < 2; ++j)
        if (16 * j < nc) {    // nc % 16 == 0
            uint32_t w[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) w[t] = tc::pack_bf16(v[16 * j + 2 * t], v[16 * j + 2 * t + 1]);
            asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst + 16 * j), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                         "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]) : "memory");
        }
}
// 32 fp32 accumulator values of this lane's row -> 64 contiguous bytes of bf16 (row-major global matrix)
__device__ __forceinline__ void store_row32_bf16(bf16* dst, const float* v, int nc) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
        if (8 * j < nc)     // nc % 8 == 0
            *(uint4*)(dst + 8 * j) = make_uint4(tc::pack_bf16(v[8 * j], v[8 * j + 1]), tc::pack_bf16(v[8 * j + 2], v[8 * j + 3]),
                                                tc::pack_bf16(v[8 * j + 4], v[8 * j + 5]), tc::pack_bf16(v[8 * j + 6], v[8 * j + 7]));
}

template <int DK, bool TRAIN, bool CONV>
__global__ void __launch_bounds__(FE_THREADS, 1)
fused_enc_fwd_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmY,
                     const FusedParams p) {
    using C = FCfg<DK>;
    constexpr int KF = C::KF, KH = C::KH, NB = C::NB, HG = C::HG, BN = C::BN;
    const int kpp = p.taps * p.KB;                  // ring stages per pass (CONV: 3 taps x KB k blocks)
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* ring = smem;
    uint8_t* ysm = smem + FE_STAGES * FE_STAGE_BYTES;
    uint64_t* bars = (uint64_t*)(ysm + (size_t)p.HKB * FE_YBLK);
    uint64_t* full = bars;                  // [2]  TMA -> MMA
    uint64_t* empty = bars + 2;             // [2]  MMA -> TMA
    uint64_t* tfull = bars + 4;             // [2]  MMA -> compute (accumulator ready)
    uint64_t* tempty = bars + 6;            // [2]  compute -> MMA (accumulator drained; FE_CW warps)
    uint64_t* yfull = bars + 8;             // [1]  compute -> MMA (Y tile complete; FE_CW warps)
    uint32_t* tmem_slot = (uint32_t*)(bars + 9);
    float* plog = (float*)((uint8_t*)bars + 256);                       // [4][128] partial pooling logits
    float* sb1 = plog + 512;                                            // [256] att_fc1.bias  (zero beyond Q)
    float* sw2 = sb1 + 256;                                             // [256] att_fc2.weight (zero beyond Q)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // All CTAs walk the SAME weight blocks; in lockstep they all ask the L2 for the same lines at the same moment
    // (measured on the first version: 0.57 ms in lockstep vs 0.25 ms staggered for the bare GEMM pipeline).  CTA c
    // therefore starts at head group c % G and walks the groups cyclically.  Only the ORDER OF THE PASSES is rotated:
    // the k blocks of one accumulation keep their order, because a title's result must not depend on which CTA
    // computes it (fp32 accumulation order; tests/test_gpu_fullsize.py checks scores bit for bit under batch
    // permutation).
    const int g_rot = (int)blockIdx.x % p.G;
    constexpr int y_rot = 0;
    if (warp == FE_TMA_WARP && lane == 0) {
        tc::prefetch_tmap(&tmA); tc::prefetch_tmap(&tmB); tc::prefetch_tmap(&tmW1);
        if (TRAIN) tc::prefetch_tmap(&tmY);
        for (int s = 0; s < FE_STAGES; ++s) { tc::mbar_init(full + s, 1); tc::mbar_init(empty + s, 1); }
        for (int a = 0; a < 2; ++a) { tc::mbar_init(tfull + a, 1); tc::mbar_init(tempty + a, FE_CW); }
        tc::mbar_init(yfull, FE_CW);
        tc::fence_barrier_init();
    }
    for (int i = threadIdx.x; i < 256; i += FE_THREADS) {
        sb1[i] = i < p.Q ? __ldg(p.b1 + i) : 0.f;
        sw2[i] = i < p.Q ? __ldg(p.w2 + i) : 0.f;
    }
    if (warp == FE_MMA_WARP) tc::tmem_alloc(tmem_slot, 512);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == FE_TMA_WARP) {
        // ===== TMA producer: per tile G x KB stages of (Xb block, W_qkv block), then HKB stages of a W1 block =====
        if (lane == 0 && !FE_DBG(p, 32)) {
            int s = 0; uint32_t ph = 0;
            for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
                const int m0 = tile * TC_BM;
                for (int gi = 0; gi < p.G; ++gi) {
                    const int g = (gi + g_rot) % p.G;
                    for (int ki = 0; ki < kpp; ++ki) {
                        const int tap = ki / p.KB, kb = ki - tap * p.KB;
                        tc::mbar_wait(empty + s, ph ^ 1);
                        uint8_t* sa = ring + (size_t)s * FE_STAGE_BYTES;
                        tc::mbar_arrive_expect_tx(full + s, FE_A_BYTES + (uint32_t)p.BN * TC_BK * 2);
                        // CONV: tap t reads the rows shifted by t - 1 (zero rows between titles, TMA zero fill off the ends)
                        tc::tma_load_2d(sa, &tmA, full + s, kb * TC_BK, CONV ? m0 + tap - 1 : m0);
                        tc::tma_load_2d(sa + FE_A_BYTES, &tmB, full + s, tap * p.Kp + kb * TC_BK, g * p.BN);
                        if (++s == FE_STAGES) { s = 0; ph ^= 1; }
                    }
                }
                for (int ki = 0; ki < p.HKB; ++ki) {
                    const int kb = (ki + y_rot) % p.HKB;
                    tc::mbar_wait(empty + s, ph ^ 1);
                    uint8_t* sa = ring + (size_t)s * FE_STAGE_BYTES;
                    tc::mbar_arrive_expect_tx(full + s, (uint32_t)p.NQ * TC_BK * 2);
                    tc::tma_load_2d(sa + FE_A_BYTES, &tmW1, full + s, kb * TC_BK, 0);
                    if (++s == FE_STAGES) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == FE_MMA_WARP) {
        // ===== MMA issuer =====
        const uint32_t idesc_qkv = tc::umma_idesc_bf16(TC_BM, p.BN, 0, 0);
        const uint32_t idesc_fc1 = tc::umma_idesc_bf16(TC_BM, p.NQ, 0, 0);
        int s = 0; uint32_t ph = 0;
        uint32_t pc = 0;                                // pass counter: accumulator pc & 1, use number pc >> 1
        int it = 0;
        for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
            const int m0 = tile * TC_BM;
            for (int g = 0; g < p.G; ++g, ++pc) {
                const uint32_t acc = pc & 1u;
                tc::mbar_wait(tempty + acc, ((pc >> 1) & 1u) ^ 1u);
                tc::tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * 256u;
                if (FE_DBG(p, 32)) {       // probe: no loads, no MMAs — the accumulator is handed over as it is
                    if (lane == 0) tc::umma_commit(tfull + acc);
                    __syncwarp();
                    continue;
                }
                for (int kb = 0; kb < kpp; ++kb) {
                    tc::mbar_wait(full + s, ph);
                    tc::tc_fence_after();
                    if (lane == 0) {
                        const uint32_t sa = tc::smem_u32(ring + (size_t)s * FE_STAGE_BYTES);
                        const uint64_t da = tc::umma_desc_sw128(sa, 16, 1024);
                        const uint64_t db = tc::umma_desc_sw128(sa + FE_A_BYTES, 16, 1024);
#pragma unroll
                        for (int k = 0; k < TC_BK / 16; ++k)
                            tc::umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc_qkv, (kb | k) ? 1u : 0u);
                        tc::umma_commit(empty + s);
                        if (kb == kpp - 1) {
                            // the previous tile's Y tile is still being read by its TMA store: the compute warps start
                            // overwriting it when this pass is signalled
                            if (TRAIN && g == 0 && it > 0) bulk_wait_read0();
                            tc::umma_commit(tfull + acc);
                        }
                    }
                    __syncwarp();
                    if (++s == FE_STAGES) { s = 0; ph ^= 1; }
                }
            }
            // fc1 pass: A = the Y tile in shared memory (written by the compute warps), B = W1 blocks from the ring
            {
                const uint32_t acc = pc & 1u;
                tc::mbar_wait(tempty + acc, ((pc >> 1) & 1u) ^ 1u);
                tc::mbar_wait(yfull, (uint32_t)it & 1u);
                tc::tc_fence_after();
                if (TRAIN && lane == 0) {               // Y (bf16) -> HBM for the backward: one bulk tensor store per block
                    for (int kb = 0; kb < p.HKB; ++kb) tma_store_2d(&tmY, ysm + (size_t)kb * FE_YBLK, kb * TC_BK, m0);
                    bulk_commit();
                }
                const uint32_t d_tmem = tmem_base + acc * 256u;
                if (FE_DBG(p, 32)) {
                    if (lane == 0) tc::umma_commit(tfull + acc);
                    __syncwarp();
                    ++pc;
                    continue;
                }
                for (int ki = 0; ki < p.HKB; ++ki) {
                    const int kb = (ki + y_rot) % p.HKB;
                    tc::mbar_wait(full + s, ph);
                    tc::tc_fence_after();
                    if (lane == 0) {
                        const uint64_t da = tc::umma_desc_sw128(tc::smem_u32(ysm + (size_t)kb * FE_YBLK), 16, 1024);
                        const uint64_t db = tc::umma_desc_sw128(tc::smem_u32(ring + (size_t)s * FE_STAGE_BYTES + FE_A_BYTES), 16, 1024);
                        const int ksteps = (p.H - kb * TC_BK >= TC_BK) ? TC_BK / 16 : (p.H - kb * TC_BK) / 16;
                        for (int k = 0; k < ksteps; ++k)
                            tc::umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc_fc1, (ki | k) ? 1u : 0u);
                        tc::umma_commit(empty + s);
                        if (ki == p.HKB - 1) tc::umma_commit(tfull + acc);
                    }
                    __syncwarp();
                    if (++s == FE_STAGES) { s = 0; ph ^= 1; }
                }
                ++pc;
            }
        }
        if (TRAIN && lane == 0) bulk_wait0();
    } else {
        // ===== compute warps =====
        const int q = warp & 3;                         // TMEM lane quadrant = title of the tile
        const int sub = warp >> 2;                      // which of the title's four warps
        const int g8 = lane >> 2, t4 = lane & 3;
        const float sl2 = p.scale * 1.4426950408889634f;
        const uint64_t key = TRAIN ? salted(p.seed) : 0;
        const int S = p.S, H = p.H, Q = p.Q;
        const uint32_t ybase = tc::smem_u32(ysm);
        uint32_t pc = 0;
        for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
            const int64_t m0 = (int64_t)tile * TC_BM;
            const int64_t title = (int64_t)tile * 4 + q;
            const bool live = title < p.n;
            const int64_t row = m0 + q * 32 + lane;     // this lane's row in the thread-per-row phases
            // key mask of this title for the columns this lane holds in a C fragment (cols nt*8 + 2t + e)
            float mk[4][2];
#pragma unroll
            for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int col = nt * 8 + 2 * t4 + e;
                    mk[nt][e] = col < S ? ((p.mask && live) ? __ldg(p.mask + title * S + col) : 1.f) : 0.f;
                }
            for (int gi = 0; gi < p.G; ++gi, ++pc) {
                const int g = (gi + g_rot) % p.G;
                const uint32_t acc = pc & 1u;
                tc::mbar_wait(tfull + acc, (pc >> 1) & 1u);
                tc::tc_fence_after();
                const uint32_t t_quad = tmem_base + ((uint32_t)(q * 32) << 16) + acc * 256u;
                if (CONV) {
                    // Conv1d output chunk (bias already inside the product) -> bf16 -> the swizzled Y tile; thread = row,
                    // the title's four warps take every fourth 32-column chunk
                    const int ncols = (H - g * p.BN < p.BN) ? H - g * p.BN : p.BN;
                    for (int c = 32 * sub; c < ncols; c += 128) {
                        float v[32];
                        tc::tmem_ld32(t_quad + (uint32_t)c, v);
                        tc::tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (c + 8 * j < ncols) {      // H % 8 == 0
                                const uint32_t a = ybase + y_off(q * 32 + lane, g * p.BN + c + 8 * j);
                                asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(tc::pack_bf16(v[8 * j], v[8 * j + 1])),
                                             "r"(tc::pack_bf16(v[8 * j + 2], v[8 * j + 3])), "r"(tc::pack_bf16(v[8 * j + 4], v[8 * j + 5])),
                                             "r"(tc::pack_bf16(v[8 * j + 6], v[8 * j + 7])) : "memory");
                            }
                    }
                }
                if (!CONV && TRAIN && !FE_DBG(p, 2)) {
                    // qkv (bf16, head-major columns g*BN + c) -> HBM: thread = row, 64 contiguous bytes per chunk
                    for (int c = 32 * sub; c < BN; c += 128) {
                        float v[32];
                        tc::tmem_ld32(t_quad + (uint32_t)c, v);
                        tc::tmem_ld_wait();
                        if (row < p.R) store_row32_bf16_v8(p.qkv + row * 3 * (int64_t)H + g * BN + c, v, BN - c < 32 ? BN - c : 32);
                    }
                }
                if (FE_DBG(p, 16)) {       // probe: a pure delay of (dbg >> 8) x 256 cycles per pass, nothing touched
                    const long long t0 = clock64();
                    while (clock64() - t0 < (long long)(p.dbg >> 8) * 256) {}
                }
                if (!CONV && sub < HG && !FE_DBG(p, 1)) {
                    const int head = g * HG + sub;
                    const uint32_t cb = (uint32_t)(sub * 3 * DK);
                    uint32_t rq[2][4 * NB], rk[2][4 * NB], rv[2][4 * NB];
#pragma unroll
                    for (int L = 0; L < 2; ++L) {
                        const uint32_t ta = t_quad + ((uint32_t)(16 * L) << 16) + cb;
                        tmem_ld_blocks<NB>(ta, rq[L]);
                        tmem_ld_blocks<NB>(ta + DK, rk[L]);
                        tmem_ld_blocks<NB>(ta + 2 * DK, rv[L]);
                    }
                    tc::tmem_ld_wait();
                    // ---- S = Q K^T
                    float sc[2][4][4];
                    amma::zero_c(sc);
                    {
                        uint32_t aq[2][KF + 1][4], bk[4][KF + 1][2];
#pragma unroll
                        for (int L = 0; L < 2; ++L) {
#pragma unroll
                            for (int ks = 0; ks < KF; ++ks) {
                                aq[L][ks][0] = packu(rq[L][8 * ks], rq[L][8 * ks + 1]);
                                aq[L][ks][1] = packu(rq[L][8 * ks + 2], rq[L][8 * ks + 3]);
                                aq[L][ks][2] = packu(rq[L][8 * ks + 4], rq[L][8 * ks + 5]);
                                aq[L][ks][3] = packu(rq[L][8 * ks + 6], rq[L][8 * ks + 7]);
                                bk[2 * L][ks][0] = packu(rk[L][8 * ks], rk[L][8 * ks + 1]);
                                bk[2 * L][ks][1] = packu(rk[L][8 * ks + 4], rk[L][8 * ks + 5]);
                                bk[2 * L + 1][ks][0] = packu(rk[L][8 * ks + 2], rk[L][8 * ks + 3]);
                                bk[2 * L + 1][ks][1] = packu(rk[L][8 * ks + 6], rk[L][8 * ks + 7]);
                            }
                            if (KH) {
                                // 8-wide half step: columns 16KF + 2t, +1 of the slice; those >= DK belong to the next
                                // slice and are zeroed in Q (one zeroed operand is enough: the other side is finite)
                                const bool ok = 16 * KF + 2 * t4 < DK;
                                aq[L][KF][0] = ok ? packu(rq[L][8 * KF], rq[L][8 * KF + 1]) : 0u;
                                aq[L][KF][1] = ok ? packu(rq[L][8 * KF + 2], rq[L][8 * KF + 3]) : 0u;
                                aq[L][KF][2] = 0u; aq[L][KF][3] = 0u;
                                bk[2 * L][KF][0] = packu(rk[L][8 * KF], rk[L][8 * KF + 1]);
                                bk[2 * L + 1][KF][0] = packu(rk[L][8 * KF + 2], rk[L][8 * KF + 3]);
                                bk[2 * L][KF][1] = 0u; bk[2 * L + 1][KF][1] = 0u;
                            }
                        }
#pragma unroll
                        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                            for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
                                for (int ks = 0; ks < KF; ++ks) amma::mma16816(sc[mt][nt], aq[mt][ks], bk[nt][ks]);
                                if (KH) amma::mma1688(sc[mt][nt], aq[mt][KF], bk[nt][KF][0]);
                            }
                    }
                    // ---- P = exp(S / sqrt(dk)) * key mask ; row sums (model_utils.py:47-53)
                    float z[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
#pragma unroll
                    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const float m = mk[nt][e & 1];
                                float pr = amma::exp2_fast(sc[mt][nt][e] * sl2) * m;
                                pr = m != 0.f ? pr : 0.f;           // select: inf * 0 guard for pad / masked columns
                                sc[mt][nt][e] = pr;
                                z[mt][e >> 1] += pr;
                            }
                    float inv[2][2];
#pragma unroll
                    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                        for (int h2 = 0; h2 < 2; ++h2)
                            inv[mt][h2] = __fdividef(p.dscale, amma::quad_sum(z[mt][h2]) + kEps);      // model_utils.py:53
                    // ---- O = P V : V fragments = 8x8 blocks of the TMEM rows transposed in registers
                    float o[2][NB][4];
                    {
                        uint32_t ap[2][2][4], bv[NB][2][2];
                        amma::c_to_a(sc, ap);
#pragma unroll
                        for (int L = 0; L < 2; ++L)
#pragma unroll
                            for (int b = 0; b < NB; ++b) {
                                bv[b][L][0] = amma::movm_t(packu(rv[L][4 * b], rv[L][4 * b + 1]));
                                bv[b][L][1] = amma::movm_t(packu(rv[L][4 * b + 2], rv[L][4 * b + 3]));
                            }
                        amma::mm_32xn<NB>(o, ap, bv);
                    }
                    // ---- normalise, output dropout (NRMS.py:32-34), bf16, into the swizzled Y tile.
                    // Keep bits: the head's (32 rows x 8-column block) groups of the output stream are drawn ONE Philox
                    // call per lane for a PAIR of (16-row half mt, block b) tiles — lane l draws row (l & 15) of tile
                    // 2j + (l >> 4) — and handed to the lanes that hold those elements by shuffles.
                    const int64_t kbase = (title * 32) * H + head * DK;     // keep-byte index (padded rows)
                    uint32_t kb[2][NB][2];                                  // [mt][b][row half]: 2 keep bits (cols 2t, 2t+1)
#pragma unroll
                    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                        for (int b = 0; b < NB; ++b) { kb[mt][b][0] = 3u; kb[mt][b][1] = 3u; }
                    if (TRAIN && p.thr && !FE_DBG(p, 4)) {
                        constexpr int NT2 = 2 * NB;                         // (mt, b) tiles of the head, tile index = mt * NB + b
#pragma unroll
                        for (int j = 0; j < (NT2 + 1) / 2; ++j) {
                            const int tile = 2 * j + (lane >> 4);
                            const int tmt = tile / NB, tb = tile - tmt * NB;
                            const int prow = 16 * tmt + (lane & 15);
                            uint32_t bits = 0u;
                            if (tile < NT2 && prow < S && live) {
                                bits = dropout8_bits_key(out_stream_index(title * S + prow, head, 8 * tb, p.heads, DK) >> 3, key,
                                                         NR_STREAM_OUT, p.thr);
                                if (8 * tb + 4 >= DK) bits &= 0xFu;
                                uint8_t* kp = p.keep + ((kbase + (int64_t)prow * H + 8 * tb) >> 2);
                                kp[0] = (uint8_t)(bits & 0xFu);
                                if (8 * tb + 4 < DK) kp[1] = (uint8_t)(bits >> 4);
                            }
#pragma unroll
                            for (int u = 0; u < 2; ++u) {                   // the two tiles of the pair
                                const int t2 = 2 * j + u;
                                if (t2 < NT2) {
                                    const int mt = t2 / NB, b = t2 - mt * NB;       // compile-time after unrolling
                                    const uint32_t lo = __shfl_sync(0xffffffffu, bits, 16 * u + g8);
                                    const uint32_t hi = __shfl_sync(0xffffffffu, bits, 16 * u + g8 + 8);
                                    kb[mt][b][0] = (lo >> (2 * t4)) & 3u;
                                    kb[mt][b][1] = (hi >> (2 * t4)) & 3u;
                                }
                            }
                        }
                    }
#pragma unroll
                    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                        for (int b = 0; b < NB; ++b) {
                            const uint32_t klo = kb[mt][b][0], khi = kb[mt][b][1];
                            const int d0 = 8 * b + 2 * t4;
                            if (d0 < DK) {
                                const int r0 = 16 * mt + g8, r1 = r0 + 8;
                                float x0 = o[mt][b][0] * inv[mt][0], x1 = o[mt][b][1] * inv[mt][0];
                                float x2 = o[mt][b][2] * inv[mt][1], x3 = o[mt][b][3] * inv[mt][1];
                                if (!(klo & 1u) || r0 >= S) x0 = 0.f;
                                if (!(klo & 2u) || r0 >= S) x1 = 0.f;
                                if (!(khi & 1u) || r1 >= S) x2 = 0.f;
                                if (!(khi & 2u) || r1 >= S) x3 = 0.f;
                                const int col = head * DK + d0;
                                asm volatile("st.shared.b32 [%0], %1;" ::"r"(ybase + y_off(q * 32 + r0, col)), "r"(tc::pack_bf16(x0, x1)) : "memory");
                                asm volatile("st.shared.b32 [%0], %1;" ::"r"(ybase + y_off(q * 32 + r1, col)), "r"(tc::pack_bf16(x2, x3)) : "memory");
                            }
                        }
                }
                tc::tc_fence_before();
                if (gi == p.G - 1) tc::fence_proxy_async_smem();    // this warp's Y writes -> visible to the fc1 MMA / TMA store
                __syncwarp();
                if (lane == 0) {
                    tc::mbar_arrive(tempty + acc);
                    if (gi == p.G - 1) tc::mbar_arrive(yfull);
                }
            }
            // ---- fc1 pass: e = tanh(acc + b1), logit = e . w2 + b2 (model_utils.py:21-23); thread = row
            {
                const uint32_t acc = pc & 1u;
                tc::mbar_wait(tfull + acc, (pc >> 1) & 1u);
                tc::tc_fence_after();
                const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * 256u;
                float part = 0.f;
                for (int c = 32 * sub; c < Q; c += 128) {
                    const int nc = (Q - c < 32) ? Q - c : 32;
                    float v[32];
                    tc::tmem_ld32(t_row + (uint32_t)c, v);
                    tc::tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {      // b1 / w2 from shared memory (broadcast reads; zero beyond Q)
                        const float4 b = *(const float4*)(sb1 + c + j), w = *(const float4*)(sw2 + c + j);
                        v[j] = tanh_fast(v[j] + b.x); v[j + 1] = tanh_fast(v[j + 1] + b.y);
                        v[j + 2] = tanh_fast(v[j + 2] + b.z); v[j + 3] = tanh_fast(v[j + 3] + b.w);
                        part = fmaf(v[j], w.x, part); part = fmaf(v[j + 1], w.y, part);
                        part = fmaf(v[j + 2], w.z, part); part = fmaf(v[j + 3], w.w, part);
                    }
                    if (TRAIN && row < p.R && !FE_DBG(p, 2)) store_row32_bf16(p.E + row * (int64_t)Q + c, v, nc);
                }
                plog[sub * 128 + q * 32 + lane] = part;
                tc::tc_fence_before();
                title_sync(1 + q);
                const float logit = plog[q * 32 + lane] + plog[128 + q * 32 + lane] + plog[256 + q * 32 + lane] +
                                    plog[384 + q * 32 + lane] + __ldg(p.b2);
                float pe = 0.f;
                if (lane < S) {
                    pe = expf(logit);
                    if (p.mask && live) pe *= __ldg(p.mask + title * S + lane);
                }
                const float al = pe / (warp_sum(pe) + kEps);                  // model_utils.py:29
                if (TRAIN && sub == 0 && row < p.R) p.alpha[row] = al;
                // out = sum_s alpha_s Y_s over this warp's quarter of the columns: bf16 pairs (words) lane, lane + 32
                const int wpr = H / 2, wq = (wpr + 3) / 4;
                const int w_lo = sub * wq, w_hi = (w_lo + wq < wpr) ? w_lo + wq : wpr;
                constexpr int WMAX = 2;                                       // wq <= 64  (H <= 512)
                float a0[WMAX], a1[WMAX];
#pragma unroll
                for (int k = 0; k < WMAX; ++k) { a0[k] = 0.f; a1[k] = 0.f; }
#pragma unroll 2
                for (int s = 0; s < (FE_DBG(p, 8) ? 0 : S); ++s) {
                    const float as = __shfl_sync(0xffffffffu, al, s);
#pragma unroll
                    for (int k = 0; k < WMAX; ++k) {
                        const int w = w_lo + lane + 32 * k;
                        if (w < w_hi) {
                            uint32_t u;
                            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(u) : "r"(ybase + y_off(q * 32 + s, 2 * w)));
                            const float2 f = bf2_to_f2(u);
                            a0[k] = fmaf(as, f.x, a0[k]);
                            a1[k] = fmaf(as, f.y, a1[k]);
                        }
                    }
                }
                if (live) {
#pragma unroll
                    for (int k = 0; k < WMAX; ++k) {
                        const int w = w_lo + lane + 32 * k;
                        if (w < w_hi) *(float2*)(p.out + title * H + 2 * w) = make_float2(a0[k], a1[k]);
                    }
                }
                // the other warps of the title still read its rows of the Y tile (and plog): nobody starts the next
                // tile's writes early
                title_sync(1 + q);
                tc::tc_fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(tempty + acc);
                ++pc;
            }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == FE_MMA_WARP) {
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, 512);
    }
}

// host side ------------------------------------------------------------------------------------------------------
struct FusedArgs {
    const bf16* xb; int64_t R; int ldx; int Kx;      // staged rows (R x ldx), Kx = Dm + 1 real columns
    const bf16* wb; int Kp;                         // MHSA: packed head-major QKV weights (3H x Kp); CONV: (H x 3 Kp), tap-major
    const bf16* w1b; int HK;                        // packed W1 (Q x HK)
    bf16* y;                                        // (R x H) training save, NULL in inference
};

inline bool fused_supported(int SP, int dk, int heads, int H, int Q, int Kp) {
    if (SP != 32 || (dk != 20 && dk != 16)) return false;
    if (heads % 4) return false;
    if (H % 16 || H > 448 || Q % 8 || Q > 240 || Q < 8) return false;
    return Kp % 64 == 0;
}
inline bool fused_conv_supported(int SP, int H, int Q, int Kp) {
    return SP == 32 && H % 16 == 0 && H >= 16 && H <= 448 && Q % 8 == 0 && Q >= 8 && Q <= 240 && Kp % 64 == 0;
}

template <int DK, bool TRAIN, bool CONV>
int launch_fused(const FusedArgs& a, FusedParams p, cudaStream_t st) {
    using C = FCfg<DK>;
    if (CONV) {
        const int n16 = (p.H + 15) / 16 * 16;
        p.G = (n16 + 239) / 240;
        p.BN = ((n16 + p.G - 1) / p.G + 15) / 16 * 16;
        p.taps = 3;
    } else {
        p.G = p.heads / C::HG;
        p.BN = C::BN;
        p.taps = 1;
    }
    p.Kp = a.Kp;
    p.KB = a.Kp / 64;
    p.HKB = (p.H + 63) / 64;
    p.NQ = (p.Q + 15) / 16 * 16;
    p.tiles = (int)((a.R + TC_BM - 1) / TC_BM);
    CUtensorMap tmA, tmB, tmW1, tmY;
    if (int e = make_tmap_bf16_2d(&tmA, a.xb, (uint64_t)a.R, (uint64_t)a.Kx, (uint64_t)a.ldx, TC_BM, TC_BK)) return e;
    if (CONV) {
        if (int e = make_tmap_bf16_2d(&tmB, a.wb, (uint64_t)p.H, (uint64_t)3 * a.Kp, (uint64_t)3 * a.Kp, (uint32_t)p.BN, TC_BK)) return e;
    } else {
        if (int e = make_tmap_bf16_2d(&tmB, a.wb, (uint64_t)3 * p.H, (uint64_t)a.Kp, (uint64_t)a.Kp, (uint32_t)p.BN, TC_BK)) return e;
    }
    if (int e = make_tmap_bf16_2d(&tmW1, a.w1b, (uint64_t)p.Q, (uint64_t)a.HK, (uint64_t)a.HK, (uint32_t)p.NQ, TC_BK)) return e;
    if (TRAIN) {
        if (int e = make_tmap_bf16_2d(&tmY, a.y, (uint64_t)a.R, (uint64_t)p.H, (uint64_t)p.H, TC_BM, TC_BK)) return e;
    } else {
        tmY = tmA;
    }
    const size_t smem = fused_smem_bytes(p.HKB);
    NR_REQUIRE(smem <= 227 * 1024, "fused title encoder: %zu bytes of shared memory do not fit", smem);
    auto kern = fused_enc_fwd_kernel<DK, TRAIN, CONV>;
    NR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = sm_count();
    if (grid > p.tiles) grid = p.tiles;
    ProfScope ps(TRAIN ? "enc.fused_fwd" : "enc.fused_infer", st);
    kern<<<grid, FE_THREADS, smem, st>>>(tmA, tmB, tmW1, tmY, p);
    return check_last("enc.fused_fwd");
}

}  // namespace fe
}  // namespace nr
