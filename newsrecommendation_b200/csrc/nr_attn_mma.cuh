// nr_attn_mma.cuh — per-head exp-attention of the title encoder on warp-level tensor-core MMAs
// (model_utils.py:39-55; SURVEY Appendix C for the gradients).
//
// One title is S <= 31 tokens and a head is DK <= 32 wide: the whole (32 x 32) score tile of one
// (title, head) lives in ONE warp's registers, so each product is a handful of
// mma.sync.m16n8k16 / m16n8k8 (bf16 in, fp32 accumulate) instead of ~10^3 scalar FMAs per lane.
// tcgen05 is the wrong tool at this size: its smallest tile is 128 x N x 16 out of shared memory into
// TMEM, four titles per instruction with 3/4 of the tile wasted on cross-title blocks, and a TMEM round
// trip per product; the warp-level path keeps P, dS and all intermediates in registers.
//
//   forward : S = Q K^T,  P = exp(S/sqrt(dk)) * keymask,  O = (P / (rowsum P + eps)) V   [+ dropout]
//   backward: both orientations are computed (S and S^T, dA and dA^T) so that every operand of the
//             five gradient products is already in MMA fragment layout: dQ = dS K, dK = dS^T Q, dV = A^T dO.
//
// Layout: qkv (n*32, 3H) bf16, pitch SP = 32.  Two column orders, chosen by the caller through (hs, ms) = element
// strides between heads and between the q/k/v slices of a head: canonical q | k | v blocks (hs = DK, ms = H) or
// HEAD-MAJOR [head][q|k|v] (hs = 3DK, ms = DK), where the three slices a warp needs are 6*DK contiguous bytes per row
// instead of three 2*DK-byte pieces 2H bytes apart (fewer partially used 32-byte sectors).
// Shared tiles are [32 rows][PITCH] bf16 with PITCH chosen per head width so that ldmatrix is bank
// conflict free (48-byte rows for DK = 16/20, 80-byte rows for DK = 32, 16-byte rows for DK = 8) and the
// k extent is covered by DK/16 full k-steps plus one 8-wide half step (m16n8k8) when DK % 16 != 0.
#pragma once
#include "nr_enc_kernels.cuh"

namespace nr {
namespace amma {

template <int DK>
struct Cfg {
    static_assert(DK == 8 || DK == 16 || DK == 20 || DK == 32, "head width outside the compiled set");
    static constexpr int KF = DK / 16;                      // full 16-wide k-steps of Q K^T
    static constexpr int KH = (DK % 16) ? 1 : 0;            // one 8-wide half step (DK % 16 <= 8)
    static constexpr int NT = (DK + 7) / 8;                 // n-tiles of the (32 x DK) outputs
    static constexpr int PITCH = DK == 32 ? 40 : (DK == 8 ? 8 : 24);
    static constexpr int TILE = 32 * PITCH;
    static constexpr int SLOTS = DK / 4;                    // 8-byte chunks per lane per tile
};

__device__ __forceinline__ void ldsm_x4(const bf16* p, uint32_t* r) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(tc::smem_u32(p)));
}
__device__ __forceinline__ void ldsm_x2(const bf16* p, uint32_t* r) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];"
                 : "=r"(r[0]), "=r"(r[1]) : "r"(tc::smem_u32(p)));
}
__device__ __forceinline__ void ldsm_x2_trans(const bf16* p, uint32_t* r) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];"
                 : "=r"(r[0]), "=r"(r[1]) : "r"(tc::smem_u32(p)));
}
__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, const uint32_t* b) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void mma1688(float* c, const uint32_t* a, uint32_t b) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(b));
}

// Fragments of a row-major (32 x DK) shared tile X used as the A operand (rows = m, cols = k) and as the
// "row" B operand (B = X^T: rows = n, cols = k).  Full k-steps + optional half step.
template <int DK>
struct Frag {
    uint32_t a[2][Cfg<DK>::KF + 1][4];      // [mt][ks]: ks == KF is the half step (2 registers used)
    uint32_t b[4][Cfg<DK>::KF + 1][2];      // [nt][ks]: half step uses b[..][KF][0]
};
template <int DK>
__device__ __forceinline__ void load_a(const bf16* tile, int lane, Frag<DK>& f) {
    constexpr int P = Cfg<DK>::PITCH, KF = Cfg<DK>::KF, KH = Cfg<DK>::KH;
    const int row = (lane & 7) + ((lane >> 3) & 1) * 8;
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
        for (int ks = 0; ks < KF; ++ks) ldsm_x4(tile + (mt * 16 + row) * P + ks * 16 + (lane >> 4) * 8, f.a[mt][ks]);
        if (KH) ldsm_x2(tile + (mt * 16 + row) * P + KF * 16, f.a[mt][KF]);
    }
}
template <int DK>
__device__ __forceinline__ void load_b_rows(const bf16* tile, int lane, Frag<DK>& f) {
    constexpr int P = Cfg<DK>::PITCH, KF = Cfg<DK>::KF, KH = Cfg<DK>::KH;
#pragma unroll
    for (int np = 0; np < 2; ++np) {
#pragma unroll
        for (int ks = 0; ks < KF; ++ks) {
            uint32_t r[4];
            ldsm_x4(tile + (np * 16 + (lane & 7) + (lane >> 4) * 8) * P + ks * 16 + ((lane >> 3) & 1) * 8, r);
            f.b[2 * np][ks][0] = r[0]; f.b[2 * np][ks][1] = r[1];
            f.b[2 * np + 1][ks][0] = r[2]; f.b[2 * np + 1][ks][1] = r[3];
        }
        if (KH) {
            uint32_t r[2];
            ldsm_x2(tile + (np * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * P + KF * 16, r);
            f.b[2 * np][KF][0] = r[0];
            f.b[2 * np + 1][KF][0] = r[1];
        }
    }
}
// C (32 x 32) += A-fragments . B-row-fragments^T over the DK-wide contraction
template <int DK>
__device__ __forceinline__ void mm_32x32(float (&c)[2][4][4], const Frag<DK>& fa, const Frag<DK>& fb) {
    constexpr int KF = Cfg<DK>::KF, KH = Cfg<DK>::KH;
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
            for (int ks = 0; ks < KF; ++ks) mma16816(c[mt][nt], fa.a[mt][ks], fb.b[nt][ks]);
            if (KH) mma1688(c[mt][nt], fa.a[mt][KF], fb.b[nt][KF][0]);
        }
}
// B fragments of X for a row-major X (rows = k (32), cols = n): b[nt][ks][2], nt < NT, ks < 2
template <int DK>
__device__ __forceinline__ void load_b_cols(const bf16* tile, int lane, uint32_t (&b)[Cfg<DK>::NT][2][2]) {
    constexpr int P = Cfg<DK>::PITCH;
#pragma unroll
    for (int nt = 0; nt < Cfg<DK>::NT; ++nt)
#pragma unroll
        for (int ks = 0; ks < 2; ++ks)
            ldsm_x2_trans(tile + (ks * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * P + nt * 8, b[nt][ks]);
}
// (32 x 32) accumulator tile (C layout) -> A fragments of the same matrix, bf16
__device__ __forceinline__ void c_to_a(const float (&c)[2][4][4], uint32_t (&a)[2][2][4]) {
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) {
            a[mt][ks][0] = tc::pack_bf16(c[mt][2 * ks][0], c[mt][2 * ks][1]);
            a[mt][ks][1] = tc::pack_bf16(c[mt][2 * ks][2], c[mt][2 * ks][3]);
            a[mt][ks][2] = tc::pack_bf16(c[mt][2 * ks + 1][0], c[mt][2 * ks + 1][1]);
            a[mt][ks][3] = tc::pack_bf16(c[mt][2 * ks + 1][2], c[mt][2 * ks + 1][3]);
        }
}
// 8 x 8 bf16 block transpose inside the warp (fragment layout in, fragment layout out)
__device__ __forceinline__ uint32_t movm_t(uint32_t x) {
    uint32_t y;
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(y) : "r"(x));
    return y;
}
// A fragments of a (32 x 32) matrix X -> A fragments of X^T.  Register r of a[mt][ks] is the 8 x 8 block
// (rows mt*16 + 8*(r&1), cols ks*16 + 8*(r>>1)); the transpose swaps the tile indices and the two middle registers.
__device__ __forceinline__ void transpose_a(const uint32_t (&a)[2][2][4], uint32_t (&t)[2][2][4]) {
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) {
            t[mt][ks][0] = movm_t(a[ks][mt][0]);
            t[mt][ks][1] = movm_t(a[ks][mt][2]);
            t[mt][ks][2] = movm_t(a[ks][mt][1]);
            t[mt][ks][3] = movm_t(a[ks][mt][3]);
        }
}
__device__ __forceinline__ void zero_c(float (&c)[2][4][4]) {
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) c[mt][nt][e] = 0.f;
}
// C (32 x 8NT) = A (32 x 32) . B
template <int NT>
__device__ __forceinline__ void mm_32xn(float (&c)[2][NT][4], const uint32_t (&a)[2][2][4], const uint32_t (&b)[NT][2][2]) {
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
            for (int e = 0; e < 4; ++e) c[mt][nt][e] = 0.f;
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) mma16816(c[mt][nt], a[mt][ks], b[nt][ks]);
        }
}
// exp(x * scale) as one FMUL + MUFU.EX2: the caller passes c = scale * log2(e)
__device__ __forceinline__ float exp2_fast(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// quad reduction: lanes 4g..4g+3 hold partial sums of the same rows
__device__ __forceinline__ float quad_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    return v;
}
// accumulator (32 x 8NT, C layout) * mul -> fp32 staging tile os[32][DK]
template <int NT, int DK>
__device__ __forceinline__ void stage_out(const float (&c)[2][NT][4], float* os, int lane, float m00, float m01, float m10,
                                          float m11) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int col = nt * 8 + 2 * t;
            if (col < DK) {
                const float lo = mt == 0 ? m00 : m10, hi = mt == 0 ? m01 : m11;
                *(float2*)(os + (mt * 16 + g) * DK + col) = make_float2(c[mt][nt][0] * lo, c[mt][nt][1] * lo);
                *(float2*)(os + (mt * 16 + g + 8) * DK + col) = make_float2(c[mt][nt][2] * hi, c[mt][nt][3] * hi);
            }
        }
}

// Per-lane copy slots: a (32 x DK) bf16 head slice is 32*DK/4 8-byte chunks = DK/4 chunks per lane.
// Slot i of lane l is chunk idx = 32 i + l: row s = idx / (DK/4), 4-column group q4 = idx % (DK/4).
// The row and the shared/global element offsets are computed ONCE per kernel (no divisions in the loop).
template <int DK>
struct Slots {
    int s[Cfg<DK>::SLOTS];          // row
    int so[Cfg<DK>::SLOTS];         // element offset in a shared tile
    int c[Cfg<DK>::SLOTS];          // column (q4*4)
    int g1[Cfg<DK>::SLOTS];         // element offset in a global matrix of pitch H   (ctx, dctx)
    int g3[Cfg<DK>::SLOTS];         // element offset in a global matrix of pitch 3H  (qkv, dqkv)
    __device__ __forceinline__ void init(int lane, int H) {
#pragma unroll
        for (int i = 0; i < Cfg<DK>::SLOTS; ++i) {
            const int idx = i * 32 + lane;
            s[i] = idx / Cfg<DK>::SLOTS;
            c[i] = (idx - s[i] * Cfg<DK>::SLOTS) * 4;
            so[i] = s[i] * Cfg<DK>::PITCH + c[i];
            g1[i] = s[i] * H + c[i];
            g3[i] = s[i] * 3 * H + c[i];
        }
    }
};
// asynchronous copy of a head slice into a shared tile (columns >= DK stay zero): the copies of item
// i+1 are in flight while item i is computed (double buffer)
template <int DK>
__device__ __forceinline__ void prefetch_tile(bf16* tile, const bf16* src, const int (&off)[Cfg<DK>::SLOTS],
                                              const Slots<DK>& sl) {
    const uint32_t t0 = tc::smem_u32(tile);
#pragma unroll
    for (int i = 0; i < Cfg<DK>::SLOTS; ++i)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(t0 + 2u * (uint32_t)sl.so[i]), "l"(src + off[i])
                     : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
// fp32 staging tile -> bf16 rows of dst (rows >= S written as zero)
template <int DK>
__device__ __forceinline__ void store_tile(bf16* dst, const int (&off)[Cfg<DK>::SLOTS], const float* os, int S,
                                           const Slots<DK>& sl) {
#pragma unroll
    for (int i = 0; i < Cfg<DK>::SLOTS; ++i) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (sl.s[i] < S) v = *(const float4*)(os + sl.s[i] * DK + sl.c[i]);
        st_bf4(dst + off[i], v);
    }
}

// bf16 staging variants (backward: nothing but a scale is applied after the MMA, so the tile is rounded once,
// here, and copied out as 8-byte chunks): half the staging bytes of the fp32 tile -> one more CTA per SM
template <int NT, int DK>
__device__ __forceinline__ void stage_out_bf(const float (&c)[2][NT][4], bf16* os, int lane, float mul) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int col = nt * 8 + 2 * t;
            if (col < DK) {
                *(uint32_t*)(os + (mt * 16 + g) * DK + col) = tc::pack_bf16(c[mt][nt][0] * mul, c[mt][nt][1] * mul);
                *(uint32_t*)(os + (mt * 16 + g + 8) * DK + col) = tc::pack_bf16(c[mt][nt][2] * mul, c[mt][nt][3] * mul);
            }
        }
}
template <int NT, int DK>
__device__ __forceinline__ void stage_out_bf2(const float (&c)[2][NT][4], bf16* os, int lane, float m00, float m01, float m10,
                                              float m11) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int col = nt * 8 + 2 * t;
            if (col < DK) {
                const float lo = mt == 0 ? m00 : m10, hi = mt == 0 ? m01 : m11;
                *(uint32_t*)(os + (mt * 16 + g) * DK + col) = tc::pack_bf16(c[mt][nt][0] * lo, c[mt][nt][1] * lo);
                *(uint32_t*)(os + (mt * 16 + g + 8) * DK + col) = tc::pack_bf16(c[mt][nt][2] * hi, c[mt][nt][3] * hi);
            }
        }
}
template <int DK>
__device__ __forceinline__ void store_tile_bf(bf16* dst, const int (&off)[Cfg<DK>::SLOTS], const bf16* os, int S,
                                              const Slots<DK>& sl) {
#pragma unroll
    for (int i = 0; i < Cfg<DK>::SLOTS; ++i) {
        uint2 v = make_uint2(0u, 0u);
        if (sl.s[i] < S) v = *(const uint2*)(os + sl.s[i] * DK + sl.c[i]);
        *(uint2*)(dst + off[i]) = v;
    }
}

// compact variant: row s of the tile goes to row cm[slot] of a matrix of pitch ld3 that holds real-token rows only
// (cm < 0: pad token or pad slot, nothing is written)
template <int DK>
__device__ __forceinline__ void store_tile_bf_rows(bf16* base, const int (&cm)[Cfg<DK>::SLOTS], int64_t ld3, const bf16* os,
                                                   const Slots<DK>& sl) {
#pragma unroll
    for (int i = 0; i < Cfg<DK>::SLOTS; ++i)
        if (cm[i] >= 0) *(uint2*)(base + (int64_t)cm[i] * ld3 + sl.c[i]) = *(const uint2*)(os + sl.s[i] * DK + sl.c[i]);
}
// per-thread partial column sums of a (32 x 8NT) accumulator: b[nt][e] covers column nt*8 + 2t + e, this lane's 4 rows
template <int NT>
__device__ __forceinline__ void add_colsum(float (&b)[NT][2], const float (&c)[2][NT][4]) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) b[nt][e] += (c[0][nt][e] + c[0][nt][e + 2]) + (c[1][nt][e] + c[1][nt][e + 2]);
}

__host__ __device__ constexpr int pitch_of(int DK) { return DK == 32 ? 40 : (DK == 8 ? 8 : 24); }
__host__ __device__ constexpr size_t fwd_smem_per_warp(int DK) {
    return (size_t)2 * 3 * 32 * pitch_of(DK) * 2 + (size_t)32 * DK * 2;
}
__host__ __device__ constexpr size_t bwd_smem_per_warp(int DK) {
    return (size_t)2 * 4 * 32 * pitch_of(DK) * 2 + (size_t)32 * DK * 2;
}

template <int DK, bool MASK>
static __global__ void __launch_bounds__(128)
attn_fwd_mma_kernel(const bf16* __restrict__ qkv, const float* __restrict__ key_mask, bf16* __restrict__ ctx,
                    int64_t items, int S, int heads, float scale, uint32_t thr, float dscale, uint64_t seed,
                    uint8_t* __restrict__ keep, int hs, int ms) {
    constexpr int NT = Cfg<DK>::NT, TILE = Cfg<DK>::TILE;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t = lane & 3;
    const int H = heads * DK;
    const float sl2 = scale * 1.4426950408889634f;
    const uint64_t key = salted(seed);              // dropout key, once per kernel
    uint8_t* base = smem_raw + warp * fwd_smem_per_warp(DK);
    bf16* buf = (bf16*)base;                       // 2 stages x (Q | K | V)
    bf16* Os = buf + 6 * TILE;                   // bf16 staging tile (rounded once, after the final scale)
    for (int i = lane; i < 6 * TILE / 8; i += 32) ((uint4*)buf)[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncwarp();
    Slots<DK> sl;
    sl.init(lane, H);
    const int stride = (int)gridDim.x * 4;          // items = n * heads < 2^31 (host checked)
    int item = (int)blockIdx.x * 4 + warp;
    auto issue = [&](int it, int stage) {
        const int sq = it / heads;
        const bf16* src = qkv + (int64_t)sq * 32 * 3 * H + (it - sq * heads) * hs;
        bf16* b = buf + stage * 3 * TILE;
        prefetch_tile<DK>(b, src, sl.g3, sl);
        prefetch_tile<DK>(b + TILE, src + ms, sl.g3, sl);
        prefetch_tile<DK>(b + 2 * TILE, src + 2 * ms, sl.g3, sl);
        cp_async_commit();
    };
    if (item < items) issue(item, 0);
    for (int stage = 0; item < items; item += stride, stage ^= 1) {
        const int64_t seq = item / heads;
        const int head = item - (int)seq * heads;
        if ((int64_t)item + stride < items) issue(item + stride, stage ^ 1); else cp_async_commit();
        const bf16* Qs = buf + stage * 3 * TILE;
        const bf16* Ks = Qs + TILE;
        const bf16* Vs = Ks + TILE;
        float mk[4][2];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int col = nt * 8 + 2 * t + e;
                mk[nt][e] = col < S ? (MASK ? __ldg(key_mask + seq * S + col) : 1.f) : 0.f;
            }
        cp_async_wait<1>();
        __syncwarp();
        float acc[2][4][4];
        zero_c(acc);
        {
            Frag<DK> fq, fk;
            load_a<DK>(Qs, lane, fq);
            load_b_rows<DK>(Ks, lane, fk);
            mm_32x32<DK>(acc, fq, fk);
        }
        float z[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float m = mk[nt][e & 1];
                    float p = exp2_fast(acc[mt][nt][e] * sl2);
                    if (MASK) p *= m;
                    p = m != 0.f ? p : 0.f;                 // select, not a branch (inf * 0 guard for pad columns)
                    acc[mt][nt][e] = p;
                    z[mt][e >> 1] += p;
                }
        float inv[2][2];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int h = 0; h < 2; ++h) inv[mt][h] = 1.f / (quad_sum(z[mt][h]) + kEps);    // model_utils.py:53
        uint32_t ap[2][2][4], bv[NT][2][2];
        c_to_a(acc, ap);
        load_b_cols<DK>(Vs, lane, bv);
        float o[2][NT][4];
        mm_32xn<NT>(o, ap, bv);
        // 1/(1-p) of the output dropout is folded into the row normaliser: the epilogue only zeroes dropped elements
        stage_out_bf2<NT, DK>(o, Os, lane, inv[0][0] * dscale, inv[0][1] * dscale, inv[1][0] * dscale, inv[1][1] * dscale);
        __syncwarp();
        // per-item bases once; per-slot offsets (s*H + c) come precomputed from sl.g1
        const int64_t obase = seq * 32 * H + head * DK;              // ctx / keep-bit element index (padded rows)
        bf16* dst = ctx + obase;
#pragma unroll
        for (int i = 0; i < Cfg<DK>::SLOTS; ++i) {
            const int s = sl.s[i];
            uint2 v = make_uint2(0u, 0u);
            if (s < S) {
                v = *(const uint2*)(Os + s * DK + sl.c[i]);
                if (thr) {
                    // the 4 keep bits are saved (one byte per 4 elements, padded-row indexing) so that the
                    // backward epilogue reads a byte instead of re-running Philox
                    const uint32_t bits = dropout4_bits_key(out_stream_index(seq * S + s, head, sl.c[i], heads, DK), key, NR_STREAM_OUT, thr);
                    v.x &= ((bits & 1u) ? 0x0000ffffu : 0u) | ((bits & 2u) ? 0xffff0000u : 0u);
                    v.y &= ((bits & 4u) ? 0x0000ffffu : 0u) | ((bits & 8u) ? 0xffff0000u : 0u);
                    if (keep) keep[(obase + sl.g1[i]) >> 2] = (uint8_t)bits;
                }
            }
            *(uint2*)(dst + sl.g1[i]) = v;
        }
        __syncwarp();
    }
}

// COMPACT: dqkv holds the rows of real tokens only (rowmap: padded row -> compact row or -1; dx_rowlist_kernel), the
// layout enc.dx and enc.dwqkv then contract over; the bias gradients (column sums over ALL rows, which the ones column
// of Xb used to deliver through enc.dwqkv) are accumulated here from the fp32 fragments.  The host sizes the grid so
// that 4 * gridDim.x is a multiple of heads: a warp then sees ONE head, keeps its 6 NT partial sums in registers and
// leaves them in the dbpart scratch (4 * gridDim.x x 3 x DK floats) for db_reduce_kernel.
template <int DK, bool MASK, bool COMPACT = false>
static __global__ void __launch_bounds__(128)
attn_bwd_mma_kernel(const bf16* __restrict__ qkv, const float* __restrict__ key_mask, const bf16* __restrict__ dctx,
                    bf16* __restrict__ dqkv, int64_t items, int S, int heads, float scale, int hs, int ms,
                    const int32_t* __restrict__ rowmap = nullptr, float* __restrict__ dbq = nullptr,    // dbq: dbpart scratch
                    const uint8_t* __restrict__ allpad = nullptr) {
    constexpr int NT = Cfg<DK>::NT, TILE = Cfg<DK>::TILE;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int H = heads * DK;
    const float sl2 = scale * 1.4426950408889634f;
    uint8_t* base = smem_raw + warp * bwd_smem_per_warp(DK);
    bf16* buf = (bf16*)base;                       // 2 stages x (Q | K | V | dO)
    bf16* Os = buf + 8 * TILE;
    for (int i = lane; i < 8 * TILE / 8; i += 32) ((uint4*)buf)[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncwarp();
    Slots<DK> sl;
    sl.init(lane, H);
    const int stride = (int)gridDim.x * 4;          // items = n * heads < 2^31 (host checked)
    int item = (int)blockIdx.x * 4 + warp;
    // COMPACT without a key mask: a title whose 32 rows are all pad tokens (allpad[title], dx_rowlist_kernel) has
    // q = k = v = the biases in every row, so A is uniform, dA is constant along the keys, dS = A (dA - rd) = 0,
    // dQ = dK = 0 and sum_j dV_j = sum_i dO_i: such an item only needs its dO tile, for the V-bias column sums
    // (half of the titles of a MIND step are the padding news of short histories).
    auto skip_item = [&](int sq) -> bool { return COMPACT && !MASK && __ldg(allpad + sq) != 0; };
    auto issue = [&](int it, int stage) {
        const int sq = it / heads;
        const int hd = it - sq * heads;
        const bf16* src = qkv + (int64_t)sq * 32 * 3 * H + hd * hs;
        bf16* b = buf + stage * 4 * TILE;
        if (!skip_item(sq)) {
            prefetch_tile<DK>(b, src, sl.g3, sl);
            prefetch_tile<DK>(b + TILE, src + ms, sl.g3, sl);
            prefetch_tile<DK>(b + 2 * TILE, src + 2 * ms, sl.g3, sl);
        }
        prefetch_tile<DK>(b + 3 * TILE, dctx + (int64_t)sq * 32 * H + hd * DK, sl.g1, sl);
        cp_async_commit();
    };
    float bq[NT][2], bk2[NT][2], bv[NT][2];        // COMPACT: column sums of dQ, dK, dV over this warp's items
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) bq[nt][0] = bq[nt][1] = bk2[nt][0] = bk2[nt][1] = bv[nt][0] = bv[nt][1] = 0.f;
    if (item < items) issue(item, 0);
    for (int stage = 0; item < items; item += stride, stage ^= 1) {
        const int64_t seq = item / heads;
        const int head = item - (int)seq * heads;
        int cm[Cfg<DK>::SLOTS];
        if (COMPACT) {
#pragma unroll
            for (int i = 0; i < Cfg<DK>::SLOTS; ++i) cm[i] = __ldg(rowmap + seq * 32 + sl.s[i]);
        }
        if ((int64_t)item + stride < items) issue(item + stride, stage ^ 1); else cp_async_commit();
        const bf16* Qs = buf + stage * 4 * TILE;
        const bf16* Ks = Qs + TILE;
        const bf16* Vs = Ks + TILE;
        const bf16* Gs = Vs + TILE;
        float mcol[4][2];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int col = nt * 8 + 2 * t + e;
                mcol[nt][e] = col < S ? (MASK ? __ldg(key_mask + seq * S + col) : 1.f) : 0.f;
            }
        cp_async_wait<1>();
        __syncwarp();
        if (skip_item((int)seq)) {
            // V-bias column sums straight from the dO tile: lane (g, t) adds rows g, g+8, g+16, g+24 of its 2 NT columns
#pragma unroll
            for (int h = 0; h < 4; ++h) {
                const int r = g + 8 * h;
                if (r < S) {
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) {
                        const int col = nt * 8 + 2 * t;
                        if (col < DK) {
                            const float2 v2 = __bfloat1622float2(*(const __nv_bfloat162*)(Gs + r * Cfg<DK>::PITCH + col));
                            bv[nt][0] += v2.x; bv[nt][1] += v2.y;
                        }
                    }
                }
            }
            __syncwarp();
            continue;
        }
        bf16* dst = COMPACT ? dqkv + head * hs : dqkv + seq * 32 * 3 * H + head * hs;
        // ---- A = P/Z, dA = dO V^T, dS = A (dA - rd); dQ = dS K / sqrt(dk).  The key-major operands
        //      A^T and dS^T of dV = A^T dO and dK = dS^T Q are NOT recomputed (that was a second S^T = K Q^T,
        //      dA^T = V dO^T and exp pass): every 8 x 8 bf16 block of the packed fragments is transposed in
        //      registers with movmatrix, which turns the C-layout tile into A fragments of the transpose.
        uint32_t at[2][2][4], dst_t[2][2][4];
        {
            float s[2][4][4], da[2][4][4];
            zero_c(s); zero_c(da);
            {
                Frag<DK> fa, fb;
                load_a<DK>(Qs, lane, fa); load_b_rows<DK>(Ks, lane, fb);
                mm_32x32<DK>(s, fa, fb);
                load_a<DK>(Gs, lane, fa); load_b_rows<DK>(Vs, lane, fb);
                mm_32x32<DK>(da, fa, fb);
            }
            float z[2][2] = {{0.f, 0.f}, {0.f, 0.f}}, rd[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float m = mcol[nt][e & 1];
                        float p = exp2_fast(s[mt][nt][e] * sl2);
                        if (MASK) p *= m;
                        p = m != 0.f ? p : 0.f;
                        s[mt][nt][e] = p;
                        z[mt][e >> 1] += p;
                        rd[mt][e >> 1] = fmaf(p, da[mt][nt][e], rd[mt][e >> 1]);
                    }
            float inv[2][2];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int row = mt * 16 + g + 8 * h;
                    const float zs = quad_sum(z[mt][h]), rs = quad_sum(rd[mt][h]);   // shuffles: all lanes, no branch
                    inv[mt][h] = row < S ? 1.f / (zs + kEps) : 0.f;                  // pad query rows: A = dS = 0
                    rd[mt][h] = rs * inv[mt][h];
                }
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float a = s[mt][nt][e] * inv[mt][e >> 1];
                        s[mt][nt][e] = a;
                        da[mt][nt][e] = a * (da[mt][nt][e] - rd[mt][e >> 1]);
                    }
            uint32_t ads[2][2][4], bk[NT][2][2];
            c_to_a(da, ads);
            transpose_a(ads, dst_t);
            {
                uint32_t aa[2][2][4];
                c_to_a(s, aa);
                transpose_a(aa, at);
            }
            load_b_cols<DK>(Ks, lane, bk);
            float dq[2][NT][4];
            mm_32xn<NT>(dq, ads, bk);
            if (COMPACT) add_colsum<NT>(bq, dq);
            stage_out_bf<NT, DK>(dq, Os, lane, scale);
        }
        __syncwarp();
        if (COMPACT) store_tile_bf_rows<DK>(dst, cm, 3 * H, Os, sl); else store_tile_bf<DK>(dst, sl.g3, Os, S, sl);
        __syncwarp();
        {
            uint32_t bf[NT][2][2];
            float out[2][NT][4];
            load_b_cols<DK>(Qs, lane, bf);
            mm_32xn<NT>(out, dst_t, bf);                          // dK = dS^T Q / sqrt(dk)
            if (COMPACT) add_colsum<NT>(bk2, out);
            stage_out_bf<NT, DK>(out, Os, lane, scale);
            __syncwarp();
            if (COMPACT) store_tile_bf_rows<DK>(dst + ms, cm, 3 * H, Os, sl); else store_tile_bf<DK>(dst + ms, sl.g3, Os, S, sl);
            __syncwarp();
            load_b_cols<DK>(Gs, lane, bf);
            mm_32xn<NT>(out, at, bf);                             // dV = A^T dO
            if (COMPACT) add_colsum<NT>(bv, out);
            stage_out_bf<NT, DK>(out, Os, lane, 1.f);
        }
        __syncwarp();
        if (COMPACT) store_tile_bf_rows<DK>(dst + 2 * ms, cm, 3 * H, Os, sl); else store_tile_bf<DK>(dst + 2 * ms, sl.g3, Os, S, sl);
        __syncwarp();
    }
    if (COMPACT) {
        // rows 4 lanes apart hold the same columns: fold the 8 row groups, then lanes 0..3 write their 2 NT columns of
        // this warp's partial sums, dbpart[warp][q | k | v][DK] (zeros for a warp without items); db_reduce_kernel adds
        // the partials of each head in warp order, so the bias gradients do not depend on the order atomics land in
        float* mine = dbq + (int64_t)((int)blockIdx.x * 4 + warp) * 3 * DK;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                float a = bq[nt][e], b = bk2[nt][e], c = bv[nt][e];
#pragma unroll
                for (int o = 4; o < 32; o <<= 1) {
                    a += __shfl_xor_sync(0xFFFFFFFFu, a, o);
                    b += __shfl_xor_sync(0xFFFFFFFFu, b, o);
                    c += __shfl_xor_sync(0xFFFFFFFFu, c, o);
                }
                const int col = nt * 8 + 2 * t + e;
                if (g == 0 && col < DK) {
                    mine[col] = a * scale;
                    mine[DK + col] = b * scale;
                    mine[2 * DK + col] = c;
                }
            }
    }
}

// bias gradients of the COMPACT backward: db[m][head*DK + col] += sum over the warps w = head (mod heads) of
// part[w][m][col], in increasing w (a fixed order: run-to-run identical).  One thread per output element.
static __global__ void db_reduce_kernel(const float* __restrict__ part, int nwarps, int heads, int DK,
                                        float* __restrict__ dbq, float* __restrict__ dbk, float* __restrict__ dbv) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int H = heads * DK;
    if (i >= 3 * H) return;
    const int m = i / H, hc = i - m * H, head = hc / DK, col = hc - head * DK;
    float s = 0.f;
    const float* src = part + (int64_t)m * DK + col;
    for (int w = head; w < nwarps; w += 16 * heads) {        // 16 loads in flight, added in warp order
        float v[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) v[u] = w + u * heads < nwarps ? __ldg(src + (int64_t)(w + u * heads) * 3 * DK) : 0.f;
#pragma unroll
        for (int u = 0; u < 16; ++u) s += v[u];
    }
    float* db = m == 0 ? dbq : (m == 1 ? dbk : dbv);
    if (db) db[hc] += s;
}

// ---------------------------------------------------------------------------------------------------------------
// Blocked variant for sequences of up to 32*NB positions (the user encoder: 50 clicked titles, pitch 64).
// One warp still owns one (sequence, head) item, but the score matrix is walked as NB x NB blocks of 32 x 32
// with the same fragment helpers; tiles are [32*NB rows][PITCH].  The item count here is tiny (batch x heads),
// so the loads are not double buffered and the normalisers are computed in a first sweep over the key blocks
// and the products recomputed in a second one (registers stay at the level of the 32 x 32 kernels).
// ---------------------------------------------------------------------------------------------------------------
template <int NT>
__device__ __forceinline__ void mm_32xn_acc(float (&c)[2][NT][4], const uint32_t (&a)[2][2][4], const uint32_t (&b)[NT][2][2]) {
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) mma16816(c[mt][nt], a[mt][ks], b[nt][ks]);
}
template <int NT>
__device__ __forceinline__ void zero_o(float (&c)[2][NT][4]) {
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) c[mt][nt][e] = 0.f;
}
__host__ __device__ constexpr size_t fwd_blk_smem_per_warp(int DK, int NB) {
    return (size_t)3 * NB * 32 * pitch_of(DK) * 2 + (size_t)32 * DK * 4;
}
__host__ __device__ constexpr size_t bwd_blk_smem_per_warp(int DK, int NB) {
    return (size_t)4 * NB * 32 * pitch_of(DK) * 2 + (size_t)32 * DK * 4 + (size_t)2 * 32 * NB * 4;
}

template <int DK, bool MASK, int NB>
static __global__ void __launch_bounds__(128)
attn_fwd_blk_kernel(const bf16* __restrict__ qkv, const float* __restrict__ key_mask, bf16* __restrict__ ctx,
                    int64_t items, int S, int heads, float scale, uint32_t thr, float dscale, uint64_t seed,
                    uint8_t* __restrict__ keep, int hs, int ms) {
    constexpr int NT = Cfg<DK>::NT, TILE = Cfg<DK>::TILE, SPB = 32 * NB;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t = lane & 3;
    const int H = heads * DK;
    const float sl2 = scale * 1.4426950408889634f;
    const uint64_t key = salted(seed);              // dropout key, once per kernel
    bf16* buf = (bf16*)(smem_raw + warp * fwd_blk_smem_per_warp(DK, NB));      // Q | K | V, NB blocks each
    float* Os = (float*)(buf + 3 * NB * TILE);
    for (int i = lane; i < 3 * NB * TILE / 8; i += 32) ((uint4*)buf)[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncwarp();
    Slots<DK> sl;
    sl.init(lane, H);
    const int stride = (int)gridDim.x * 4;
    for (int item = (int)blockIdx.x * 4 + warp; item < items; item += stride) {
        const int64_t seq = item / heads;
        const int head = item - (int)seq * heads;
        const bf16* src = qkv + seq * SPB * 3 * H + head * hs;
#pragma unroll
        for (int m = 0; m < 3; ++m)
#pragma unroll
            for (int b = 0; b < NB; ++b)
                prefetch_tile<DK>(buf + (m * NB + b) * TILE, src + m * ms + (int64_t)b * 32 * 3 * H, sl.g3, sl);
        cp_async_commit();
        cp_async_wait<0>();
        __syncwarp();
        const bf16* Qs = buf;
        const bf16* Ks = buf + NB * TILE;
        const bf16* Vs = buf + 2 * NB * TILE;
#pragma unroll 1
        for (int qb = 0; qb < NB; ++qb) {
            Frag<DK> fq;
            load_a<DK>(Qs + qb * TILE, lane, fq);
            float acc[NB][2][4][4];
            float z[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
#pragma unroll
            for (int kb = 0; kb < NB; ++kb) {
                Frag<DK> fk;
                load_b_rows<DK>(Ks + kb * TILE, lane, fk);
                zero_c(acc[kb]);
                mm_32x32<DK>(acc[kb], fq, fk);
#pragma unroll
                for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int col = kb * 32 + nt * 8 + 2 * t + e;
                        const float m = col < S ? (MASK ? __ldg(key_mask + seq * S + col) : 1.f) : 0.f;
#pragma unroll
                        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                            for (int h = 0; h < 2; ++h) {
                                float p = exp2_fast(acc[kb][mt][nt][2 * h + e] * sl2);
                                if (MASK) p *= m;
                                p = m != 0.f ? p : 0.f;
                                acc[kb][mt][nt][2 * h + e] = p;
                                z[mt][h] += p;
                            }
                    }
            }
            float inv[2][2];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int h = 0; h < 2; ++h) inv[mt][h] = 1.f / (quad_sum(z[mt][h]) + kEps);    // model_utils.py:53
            float o[2][NT][4];
            zero_o<NT>(o);
#pragma unroll
            for (int kb = 0; kb < NB; ++kb) {
                uint32_t ap[2][2][4], bv[NT][2][2];
                c_to_a(acc[kb], ap);
                load_b_cols<DK>(Vs + kb * TILE, lane, bv);
                mm_32xn_acc<NT>(o, ap, bv);
            }
            stage_out<NT, DK>(o, Os, lane, inv[0][0], inv[0][1], inv[1][0], inv[1][1]);
            __syncwarp();
            const int64_t obase = (seq * SPB + qb * 32) * H + head * DK;
            bf16* dst = ctx + obase;
#pragma unroll
            for (int i = 0; i < Cfg<DK>::SLOTS; ++i) {
                const int s = sl.s[i], sa = qb * 32 + s;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (sa < S) {
                    v = *(const float4*)(Os + s * DK + sl.c[i]);
                    if (thr) {
                        const uint32_t bits = dropout4_bits_key(out_stream_index(seq * S + sa, head, sl.c[i], heads, DK), key, NR_STREAM_OUT, thr);
                        v = apply_keep4(v, bits, dscale);
                        if (keep) keep[(obase + sl.g1[i]) >> 2] = (uint8_t)bits;
                    }
                }
                st_bf4(dst + sl.g1[i], v);
            }
            __syncwarp();
        }
    }
}

template <int DK, bool MASK, int NB>
static __global__ void __launch_bounds__(128)
attn_bwd_blk_kernel(const bf16* __restrict__ qkv, const float* __restrict__ key_mask, const bf16* __restrict__ dctx,
                    bf16* __restrict__ dqkv, int64_t items, int S, int heads, float scale, int hs, int ms) {
    constexpr int NT = Cfg<DK>::NT, TILE = Cfg<DK>::TILE, SPB = 32 * NB;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int H = heads * DK;
    const float sl2 = scale * 1.4426950408889634f;
    bf16* buf = (bf16*)(smem_raw + warp * bwd_blk_smem_per_warp(DK, NB));      // Q | K | V | dO, NB blocks each
    float* Os = (float*)(buf + 4 * NB * TILE);
    float* zinv = Os + 32 * DK;
    float* rds = zinv + SPB;
    for (int i = lane; i < 4 * NB * TILE / 8; i += 32) ((uint4*)buf)[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncwarp();
    Slots<DK> sl;
    sl.init(lane, H);
    const int stride = (int)gridDim.x * 4;
    for (int item = (int)blockIdx.x * 4 + warp; item < items; item += stride) {
        const int64_t seq = item / heads;
        const int head = item - (int)seq * heads;
        const bf16* src = qkv + seq * SPB * 3 * H + head * hs;
        const bf16* gsrc = dctx + seq * SPB * H + head * DK;
#pragma unroll
        for (int b = 0; b < NB; ++b) {
#pragma unroll
            for (int m = 0; m < 3; ++m)
                prefetch_tile<DK>(buf + (m * NB + b) * TILE, src + m * ms + (int64_t)b * 32 * 3 * H, sl.g3, sl);
            prefetch_tile<DK>(buf + (3 * NB + b) * TILE, gsrc + (int64_t)b * 32 * H, sl.g1, sl);
        }
        cp_async_commit();
        cp_async_wait<0>();
        __syncwarp();
        const bf16* Qs = buf;
        const bf16* Ks = buf + NB * TILE;
        const bf16* Vs = buf + 2 * NB * TILE;
        const bf16* Gs = buf + 3 * NB * TILE;
        bf16* dst = dqkv + seq * SPB * 3 * H + head * hs;
        // ---- query-major: sweep 1 -> Z and rd per query row; sweep 2 -> dS blocks and dQ = sum_kb dS K / sqrt(dk)
#pragma unroll 1
        for (int qb = 0; qb < NB; ++qb) {
            Frag<DK> fq, fg;
            load_a<DK>(Qs + qb * TILE, lane, fq);
            load_a<DK>(Gs + qb * TILE, lane, fg);
            float z[2][2] = {{0.f, 0.f}, {0.f, 0.f}}, rd[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
            float inv[2][2];
            float dq[2][NT][4];
            zero_o<NT>(dq);
#pragma unroll 1
            for (int sweep = 0; sweep < 2; ++sweep) {
#pragma unroll 1
                for (int kb = 0; kb < NB; ++kb) {
                    float s[2][4][4], da[2][4][4];
                    zero_c(s); zero_c(da);
                    {
                        Frag<DK> fb;
                        load_b_rows<DK>(Ks + kb * TILE, lane, fb);
                        mm_32x32<DK>(s, fq, fb);
                        load_b_rows<DK>(Vs + kb * TILE, lane, fb);
                        mm_32x32<DK>(da, fg, fb);
                    }
#pragma unroll
                    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int col = kb * 32 + nt * 8 + 2 * t + e;
                            const float m = col < S ? (MASK ? __ldg(key_mask + seq * S + col) : 1.f) : 0.f;
#pragma unroll
                            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                                for (int h = 0; h < 2; ++h) {
                                    float p = exp2_fast(s[mt][nt][2 * h + e] * sl2);
                                    if (MASK) p *= m;
                                    p = m != 0.f ? p : 0.f;
                                    if (sweep == 0) {
                                        z[mt][h] += p;
                                        rd[mt][h] = fmaf(p, da[mt][nt][2 * h + e], rd[mt][h]);
                                    } else {
                                        s[mt][nt][2 * h + e] = p * inv[mt][h] * (da[mt][nt][2 * h + e] - rd[mt][h]);
                                    }
                                }
                        }
                    if (sweep == 1) {
                        uint32_t ads[2][2][4], bk[NT][2][2];
                        c_to_a(s, ads);
                        load_b_cols<DK>(Ks + kb * TILE, lane, bk);
                        mm_32xn_acc<NT>(dq, ads, bk);
                    }
                }
                if (sweep == 0) {
#pragma unroll
                    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const int row = qb * 32 + mt * 16 + g + 8 * h;
                            const float zs = quad_sum(z[mt][h]), rs = quad_sum(rd[mt][h]);   // shuffles: all lanes, no branch
                            inv[mt][h] = row < S ? 1.f / (zs + kEps) : 0.f;
                            rd[mt][h] = rs * inv[mt][h];
                            if (t == 0) { zinv[row] = inv[mt][h]; rds[row] = rd[mt][h]; }
                        }
                }
            }
            stage_out<NT, DK>(dq, Os, lane, scale, scale, scale, scale);
            __syncwarp();
            store_tile<DK>(dst + (int64_t)qb * 32 * 3 * H, sl.g3, Os, S - qb * 32, sl);
            __syncwarp();
        }
        // ---- key-major: dK = sum_qb dS^T Q / sqrt(dk), dV = sum_qb A^T dO
#pragma unroll 1
        for (int kb = 0; kb < NB; ++kb) {
            Frag<DK> fk, fv;
            load_a<DK>(Ks + kb * TILE, lane, fk);
            load_a<DK>(Vs + kb * TILE, lane, fv);
            float mrow[2][2];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int row = kb * 32 + mt * 16 + g + 8 * h;
                    mrow[mt][h] = row < S ? (MASK ? __ldg(key_mask + seq * S + row) : 1.f) : 0.f;
                }
            float ok[2][NT][4], ov[2][NT][4];
            zero_o<NT>(ok); zero_o<NT>(ov);
#pragma unroll 1
            for (int qb = 0; qb < NB; ++qb) {
                float s[2][4][4], da[2][4][4];
                zero_c(s); zero_c(da);
                {
                    Frag<DK> fb;
                    load_b_rows<DK>(Qs + qb * TILE, lane, fb);
                    mm_32x32<DK>(s, fk, fb);
                    load_b_rows<DK>(Gs + qb * TILE, lane, fb);
                    mm_32x32<DK>(da, fv, fb);
                }
#pragma unroll
                for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const float zv = zinv[qb * 32 + nt * 8 + 2 * t + e], rv = rds[qb * 32 + nt * 8 + 2 * t + e];
#pragma unroll
                        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                            for (int h = 0; h < 2; ++h) {
                                const float m = mrow[mt][h];
                                const float mz = MASK ? m * zv : (m != 0.f ? zv : 0.f);
                                float a = exp2_fast(s[mt][nt][2 * h + e] * sl2) * mz;
                                a = mz != 0.f ? a : 0.f;
                                s[mt][nt][2 * h + e] = a;                                     // A^T
                                da[mt][nt][2 * h + e] = a * (da[mt][nt][2 * h + e] - rv);     // dS^T
                            }
                    }
                uint32_t af[2][2][4], bf[NT][2][2];
                c_to_a(da, af);
                load_b_cols<DK>(Qs + qb * TILE, lane, bf);
                mm_32xn_acc<NT>(ok, af, bf);
                c_to_a(s, af);
                load_b_cols<DK>(Gs + qb * TILE, lane, bf);
                mm_32xn_acc<NT>(ov, af, bf);
            }
            stage_out<NT, DK>(ok, Os, lane, scale, scale, scale, scale);
            __syncwarp();
            store_tile<DK>(dst + ms + (int64_t)kb * 32 * 3 * H, sl.g3, Os, S - kb * 32, sl);
            __syncwarp();
            stage_out<NT, DK>(ov, Os, lane, 1.f, 1.f, 1.f, 1.f);
            __syncwarp();
            store_tile<DK>(dst + 2 * ms + (int64_t)kb * 32 * 3 * H, sl.g3, Os, S - kb * 32, sl);
            __syncwarp();
        }
    }
}

}  // namespace amma
}  // namespace nr
