// nr_encoder.cu — bf16 tensor-core title encoder (nr_title_encoder_fwd / _bwd): the NewsEncoder title
// path of NRMS (gather -> dropout -> MHSA -> dropout -> additive pooling, src/model/NRMS.py:27-35), of
// NAML (gather -> dropout -> Conv1d(k=3) -> additive pooling, src/model/NAML.py:47-56) and the NRMS
// UserEncoder (NRMS.py:55-62) as one pipeline of tcgen05 GEMMs (nr_tc_gemm.cuh, nr_tc_gemm_tn.cuh)
// with fused epilogues and fp32 SIMT tails (nr_enc_kernels.cuh).  Gradient math: SURVEY Appendix C.
#include <string.h>
#include "nr_enc_kernels.cuh"
#include "nr_attn_mma.cuh"
#include "nr_tc_gemm.cuh"
#include "nr_tc_gemm_tn.cuh"
#include "nr_fused_enc.cuh"

using namespace nr;

namespace nr {
int set_seed_salt_encoder(const void* p) { return set_seed_salt_this_tu(p); }
}  // namespace nr

namespace {

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

// ---- NT epilogues (protocol of nr_tc_gemm.cuh: one thread = one output row; register arrays are
// indexed at compile time only) -----------------------------------------------------------------------
__device__ __forceinline__ float tanh_fast(float x) {      // MUFU.TANH, rel. error ~2^-11 (< bf16 rounding)
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// 16 epilogue warps (two groups per accumulator) pay only for enc.pool_dx, whose epilogue is the heaviest (measured:
// pool_dx 0.163 -> 0.139 ms; dx (L2 reductions) and qkv (operand delivery) do not gain, fc1 loses: DESIGN.md section 4)
static bool tc_ew16(const char* what) { return !strcmp(what, "enc.pool_dx"); }
template <class Epi>
int gemm_tc_pick(const bf16* A, int64_t a_rows, int Kt_real, int64_t lda, const bf16* B, int N, int64_t ldb, int taps,
                 const int* shifts, int64_t M, const Epi& epi, cudaStream_t st, const char* what) {
    if (tc_ew16(what)) return gemm_tc<Epi, 16>(A, a_rows, Kt_real, lda, B, N, ldb, taps, shifts, M, epi, st, what);
    return gemm_tc<Epi, 8>(A, a_rows, Kt_real, lda, B, N, ldb, taps, shifts, M, epi, st, what);
}

// Titles (pitch 32) go through the fused title-encoder kernel (nr_fused_enc.cuh) in inference (save == NULL) and in
// training (0.48 ms vs 0.55 ms for the separate qkv / attention / fc1 / pooling launches).
// NR_ENC_FUSED=0 (read per call) selects the separate launches: the A/B switch of bench.py and of the parity matrix
// (tests/test_gpu_fused.py runs both paths against each other and the oracle).
static bool fused_enabled() {
    const char* e = getenv("NR_ENC_FUSED");
    return !(e && e[0] == '0');
}

// Head-major qkv columns ([head][q|k|v], see nr_attn_mma.cuh) whenever one of the warp-MMA kernels runs the
// attention; the SIMT fall-back kernels keep the canonical q | k | v blocks.  The weight packers, the dW epilogue
// and the attention launchers all ask this one predicate.
static bool qkv_head_major(int SP, int dk) {
    return (SP == 32 || SP == 64) && (dk == 8 || dk == 16 || dk == 20 || dk == 32);
}

// AttentionPooling fc1 + tanh + fc2 (model_utils.py:21-23): e (bf16) = tanh(acc + b1);
// logits[row] = e . w2 + b2 (fp32, from the unrounded tanh).  Needs the whole row in one n tile.
struct EpiTanhLogit {
    bf16* E; int64_t lde; const float* b1; const float* w2; const float* b2; float* logits;
    float acc;
    __device__ __forceinline__ void begin(const EpiCtx&, int64_t) { acc = 0.f; }
    __device__ __forceinline__ void end(const EpiCtx& cx, int64_t row) { if (cx.live) logits[row] = acc + __ldg(b2); }
    __device__ __forceinline__ void chunk(const EpiCtx& cx, int64_t, int col0, const float* v, int nc) {
        float b[32], w[32], o[32];
        load_bias32(b1, col0, nc, b);
        load_bias32(w2, col0, nc, w);      // zero beyond nc: padded columns add nothing to the logit
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            o[j] = tanh_fast(v[j] + b[j]);
            acc = fmaf(o[j], w[j], acc);
        }
        store_rows_bf16(cx, E, lde, col0, o, nc);
    }
};

// dY (bf16) = (dz . W1 + alpha_s g) [* output-dropout mask]; pad rows zero.  nc % 4 == 0 (H % 8 == 0).
struct EpiPoolDx {
    bf16* dY; float* dx32; int64_t ld; const float* alpha; const float* g; int S, SP, H;
    const uint8_t* keep; float dscale;      // saved output-dropout keep bits (NULL: no dropout)
    int64_t seq; int s; float a; bool real;
    // The epilogue has 8 warps per SM and nothing to hide a global-load round trip behind, so the lines the NEXT
    // chunk reads are pulled into L1 (prefetch.global.L1: no destination registers) while this one is processed.
    __device__ __forceinline__ void warm(int col0, int n1) const {
        if (!real || col0 >= n1) return;
        asm volatile("prefetch.global.L1 [%0];" ::"l"(g + seq * H + col0));
        if (keep) asm volatile("prefetch.global.L1 [%0];" ::"l"(keep + (((seq * SP + s) * H + col0) >> 2)));
    }
    __device__ __forceinline__ void begin(const EpiCtx& cx, int64_t row) {
        seq = row / SP; s = (int)(row - seq * SP);
        real = cx.live && s < S;
        a = real ? __ldg(alpha + row) : 0.f;
        warm(cx.n0, cx.n1);
    }
    __device__ __forceinline__ void end(const EpiCtx&, int64_t) {}
    __device__ __forceinline__ void chunk(const EpiCtx& cx, int64_t, int col0, const float* v, int nc) {
        warm(col0 + 32, cx.n1);
        const float* gr = g + seq * H + col0;
        // Branch-free body (selects and predicated loads only): a 4-bit keep nibble per column quad, forced to 0 for
        // pad rows / columns beyond the tile so that those come out as exact zeros.
        uint32_t kb0 = 0xFFFFFFFFu, kb1 = 0xFFFFFFFFu;
        if (keep && real) {     // one byte per 4 elements, indexed by the padded row (written by the attention forward)
            const uint32_t* kp = (const uint32_t*)(keep + (((seq * SP + s) * H + col0) >> 2));
            kb0 = __ldg(kp);
            if (nc > 16) kb1 = __ldg(kp + 1);
        }
        if (!real) kb0 = kb1 = 0u;
        if (nc < 32) {          // nc % 4 == 0: drop the nibbles of quads >= nc / 4
            const int q = nc >> 2;
            kb0 = q >= 4 ? kb0 : (kb0 & ((1u << (8 * q)) - 1u));
            kb1 = q <= 4 ? 0u : (kb1 & ((1u << (8 * (q - 4))) - 1u));
        }
        // all eight 16-byte loads of g are issued before the first use (one L1 round trip per chunk, not eight)
        float4 gc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) gc[j] = (real && 4 * j < nc) ? ldg4(gr + 4 * j) : make_float4(0.f, 0.f, 0.f, 0.f);
        float* out32 = dx32 ? dx32 + (seq * S + s) * ld + col0 : nullptr;       // POOL kind: fp32 (n,S,H), no pad rows
        uint2* srow = (uint2*)(cx.stage + cx.lane * TC_STAGE_PITCH);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t nib = (j < 4 ? kb0 : kb1) >> (8 * (j & 3));
            float4 t;
            t.x = (nib & 1u) ? fmaf(a, gc[j].x, v[4 * j]) * dscale : 0.f;
            t.y = (nib & 2u) ? fmaf(a, gc[j].y, v[4 * j + 1]) * dscale : 0.f;
            t.z = (nib & 4u) ? fmaf(a, gc[j].z, v[4 * j + 2]) * dscale : 0.f;
            t.w = (nib & 8u) ? fmaf(a, gc[j].w, v[4 * j + 3]) * dscale : 0.f;
            if (dx32) {
                if (real && 4 * j < nc) *(float4*)(out32 + 4 * j) = t;
            } else {
                srow[j] = make_uint2(tc::pack_bf16(t.x, t.y), tc::pack_bf16(t.z, t.w));
            }
        }
        if (!dx32) flush_rows_bf16(cx, dY, ld, col0, nc);      // pad rows are written as zero
    }
};

// dX of the encoder input: input-dropout mask re-applied, then either a dense fp32 store (n,S,Dm) or a
// scatter-add into the embedding-table gradient (id 0 = padding_idx skipped; NRMS.py:73, B7).
// A thread owns one row of the accumulator, so a direct scatter would issue 32 separate 16-byte reductions per warp
// instruction (32 LSU wavefronts, half-filled sectors).  Each 32-column chunk is transposed through the warp's
// staging tile instead: one instruction then covers 4 rows x 128 contiguous bytes.
struct EpiDx {
    static constexpr int kRowPitch = 36;                                 // floats: conflict-free 16-byte accesses both ways
    static constexpr int kStagePerWarp = 32 * kRowPitch * 4 + 32 * 8;    // [32][36] fp32 + the 32 row pointers
    float* dx_dense; float* dtable; RowsView x; int SP;
    const uint8_t* keep_in; int kpitch;     // input-dropout keep bits saved by enc.stage: 1 byte per 8 columns
    int64_t seq; int s; int64_t id;
    const int32_t* rowlist;                 // compact backward: output row i is padded row rowlist[i] (dx_rowlist_kernel)
    __device__ __forceinline__ void begin(const EpiCtx& cx, int64_t row) {
        if (rowlist) row = cx.live ? (int64_t)__ldg(rowlist + row) : 0;
        seq = row / SP; s = (int)(row - seq * SP);
        id = -1;
        if (!cx.live) s = x.S;          // rows past the end behave like pad rows
        if (s < x.S && x.mode != NR_ROWS_DENSE) id = x.id_of(seq, s);
        float* dst = nullptr;           // where this lane's row goes (NULL: pad slot, pad token or past the end)
        if (s < x.S && !(x.mode != NR_ROWS_DENSE && id == 0)) {
            const int64_t r = seq * x.S + s;
            dst = x.mode == NR_ROWS_DENSE ? dx_dense + r * x.D
                : (x.mode == NR_ROWS_TOKEN ? dtable + id * x.D : dtable + id * ((int64_t)x.S * x.D) + (int64_t)s * x.D);
        }
        __syncwarp();
        ((float**)((float*)cx.stage + 32 * kRowPitch))[cx.lane] = dst;
        __syncwarp();
    }
    __device__ __forceinline__ void end(const EpiCtx&, int64_t) {}
    __device__ __forceinline__ void chunk(const EpiCtx& cx, int64_t, int col0, const float* v, int nc) {
        float* tile = (float*)cx.stage;
        float* const* rowp = (float* const*)(tile + 32 * kRowPitch);
        const int64_t r = seq * x.S + s;
        uint32_t kb = 0xFFFFFFFFu;
        const bool mine = s < x.S && !(x.mode != NR_ROWS_DENSE && id == 0);
        if (mine && x.thr && keep_in) {        // col0 is a multiple of 16 (BN % 16 == 0): 2-byte aligned, not always 4
            const uint16_t* kp = (const uint16_t*)(keep_in + (seq * SP + s) * kpitch + (col0 >> 3));
            kb = (uint32_t)__ldg(kp);
            if (nc > 16) kb |= (uint32_t)__ldg(kp + 1) << 16;
        }
        float4* srow = (float4*)(tile + cx.lane * kRowPitch);
#pragma unroll
        for (int j = 0; j < 8; ++j) {      // nc % 4 == 0 (Dm % 4 == 0)
            float4 t = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            if (mine && x.thr && 4 * j < nc) {
                if (keep_in) t = apply_keep4(t, (kb >> (4 * j)) & 0xFu, x.scale);
                else t = dropout4(t, in_stream_index(r, col0, x.D) + 4 * j, x.seed, NR_STREAM_IN, x.thr, x.scale);
            }
            srow[j] = t;
        }
        __syncwarp();
        const int c = (cx.lane & 7) * 4;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int rr = (cx.lane >> 3) + 4 * j;
            float* dst = rowp[rr];
            if (dst && c < nc) {
                const float4 t = *(const float4*)(tile + rr * kRowPitch + c);
                if (x.mode == NR_ROWS_DENSE) *(float4*)(dst + col0 + c) = t;
                else atomicAdd((float4*)(dst + col0 + c), t);
            }
        }
        __syncwarp();
    }
};

// Rows of the padded (n, SP) row space that carry a real token (s < S and id != padding_idx 0): the only rows whose
// dX is kept (NRMS.py:73) and the only ones with a non-zero input row in dW = dqkv^T Xb.  62 % of the rows of a MIND
// step are pad tokens, so the backward keeps dqkv COMPACT: list[c] = padded row of compact row c, rowmap[row] = c or
// -1.  A block covers 1024 consecutive rows in order (a title's real rows stay adjacent); blocks land in the order
// of one atomic each, which does not matter: every output row is independent and the consumers are reductions.
// *count is zeroed by the caller.
__global__ void __launch_bounds__(1024) dx_rowlist_kernel(const RowsView x, const int SP, const int64_t R,
                                                          int32_t* __restrict__ list, int32_t* __restrict__ rowmap,
                                                          uint8_t* __restrict__ allpad, int32_t* __restrict__ count) {
    __shared__ int wsum[32];
    __shared__ int base;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // nn.Embedding.from_pretrained(..., padding_idx=0) keeps whatever row 0 of the pretrained matrix holds and only
    // drops its gradient (NRMS.py:71-73).  A pad token's input row is zero only if that row is: when it is not, every
    // position s < S stays in the list (dW needs it; enc.dx still skips id 0), and nothing is lost but the saving.
    bool nz = false;
    for (int c = threadIdx.x * 4; c < x.D; c += blockDim.x * 4) {
        const float4 v = ldg4(x.table + c);
        nz = nz || v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f;
    }
    const bool row0_live = __syncthreads_or(nz) != 0;
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool real = false;
    if (row < R) {
        const int64_t seq = row / SP;
        const int s = (int)(row - seq * SP);
        real = s < x.S && (row0_live || x.id_of(seq, s) != 0);
    }
    const unsigned b = __ballot_sync(0xFFFFFFFFu, real);
    if (lane == 0) wsum[warp] = __popc(b);
    // SP == 32 (compact_supported): a warp is exactly one title; a title without any listed row is ALL PAD
    if (lane == 0 && row < R) allpad[row / SP] = b == 0u;
    __syncthreads();
    if (warp == 0) {
        const int c = wsum[lane];
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += t;
        }
        wsum[lane] = incl - c;                       // exclusive offset of each warp
        if (lane == 31) base = incl ? atomicAdd(count, incl) : 0;
    }
    __syncthreads();
    const int c = base + wsum[warp] + __popc(b & ((1u << lane) - 1u));
    if (real) list[c] = (int32_t)row;
    if (row < R) rowmap[row] = real ? c : -1;
}
// rows [count, count + 64) of the compact dqkv are read by the last k chunk of enc.dwqkv: they must be finite zeros
__global__ void zero_tail_rows_kernel(bf16* __restrict__ a, const int32_t* __restrict__ count, int64_t rows, int ld) {
    const int64_t r = (int64_t)__ldg(count) + blockIdx.x;
    if (r >= rows) return;
    uint4* p = (uint4*)(a + r * ld);
    for (int i = threadIdx.x; i < ld / 8; i += blockDim.x) p[i] = make_uint4(0u, 0u, 0u, 0u);
}
// NR_DX_GATHER=0 (read per call) keeps the backward on all R rows: the A/B switch of bench.py and tests/test_gpu_fused.py
static bool dx_gather_enabled() {
    const char* e = getenv("NR_DX_GATHER");
    return !(e && e[0] == '0');
}

// ---- TN epilogues ----------------------------------------------------------------------------------------
// rows i of [dW_Q; dW_K; dW_V] (3H x Dm); column Dm of the product (ones column of Xb) = bias gradient
struct TnEpiQkv {
    float* dw0; float* dw1; float* dw2; float* db0; float* db1; float* db2; int H, Dm;
    int hm_dk;      // head width when the rows of dqkv^T . Xb arrive head-major (see qkv_head_major), else 0
    __device__ __forceinline__ void operator()(int i, int, int j, const float* v, int nc) const {
        if (hm_dk) i = qkv_canonical(i, hm_dk, H);
        const int w = i / H, ii = i - w * H;
        float* dw = w == 0 ? dw0 : (w == 1 ? dw1 : dw2);
        float* db = w == 0 ? db0 : (w == 1 ? db1 : db2);
        float* row = dw ? dw + (int64_t)ii * Dm : nullptr;
        const bool vec = row && (Dm & 3) == 0 && (((uintptr_t)dw) & 15) == 0;
#pragma unroll
        for (int t = 0; t < 32; t += 4) {
            const int c = j + t;
            if (vec && t + 4 <= nc && c + 4 <= Dm) {      // 16-byte vector reduction
                atomicAdd((float4*)(row + c), make_float4(v[t], v[t + 1], v[t + 2], v[t + 3]));
                continue;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (t + u < nc) {
                    if (c + u < Dm) { if (row) atomicAdd(row + c + u, v[t + u]); }
                    else if (c + u == Dm) { if (db) atomicAdd(db + ii, v[t + u]); }
                }
            }
        }
    }
};
// nn.Conv1d.weight gradient layout dw[o][c][t]; bias gradient = ones column of the unshifted tap
struct TnEpiConv {
    float* dw; float* db; int Dm;
    __device__ __forceinline__ void operator()(int o, int tap, int j, const float* v, int nc) const {
#pragma unroll
        for (int t = 0; t < 32; ++t) {
            const int c = j + t;
            if (t < nc) {
                if (c < Dm) { if (dw) atomicAdd(dw + ((int64_t)o * Dm + c) * 3 + tap, v[t]); }
                else if (c == Dm && tap == 1) { if (db) atomicAdd(db + o, v[t]); }
            }
        }
    }
};

// ---- buffer plan ---------------------------------------------------------------------------------------------
struct Plan {
    int SP, ldx, Kp, HK, QK, K3, Nw;
    int64_t R;
    // save
    size_t o_wb, o_w1b, o_w1t, o_wt, o_bias, o_keep, o_keep_in, o_xb, o_qkv, o_y, o_e, o_alpha, o_logits, save_bytes;
    // backward workspace
    size_t o_dz, o_dy, o_dqkv, o_rowlist, o_dbpart, ws_bytes;
};

// scratch of the compact attention backward: one 3 x dk slot of partial bias sums per warp of the largest grid
static size_t dbpart_bytes(int heads, int dk) { return (size_t)4 * (2 * sm_count() + heads) * 3 * dk * sizeof(float); }
Plan make_plan(const nr_title_encoder_t* e, int64_t n) {
    Plan p;
    const bool pool = e->kind == NR_ENC_POOL;
    p.SP = pool ? e->S : round_up(e->S + 1, 32);
    p.ldx = pool ? e->H : round_up(e->Dm + 1, 64);
    p.Kp = pool ? round_up(e->Dm, 64) : round_up(e->Dm + 1, 64);   // room for the bias column at Dm
    p.HK = round_up(e->H, 64);
    p.QK = round_up(e->Q, 64);
    p.K3 = round_up(3 * e->H, 64);
    p.Nw = e->kind == NR_ENC_MHSA ? 3 * e->H : e->H;
    p.R = n * p.SP;
    const int taps = e->kind == NR_ENC_MHSA ? 1 : 3;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o += (bytes + 1023) & ~size_t(1023); return r; };
    p.o_wb = take(pool ? 0 : (size_t)p.Nw * taps * p.Kp * 2);
    p.o_w1b = take((size_t)e->Q * p.HK * 2);
    // transposed operands of the backward dX GEMMs, packed by the forward in the same launch
    p.o_w1t = take((size_t)e->H * p.QK * 2);
    p.o_wt = take(pool ? 0 : (e->kind == NR_ENC_MHSA ? (size_t)e->Dm * p.K3 * 2 : (size_t)e->Dm * 3 * p.HK * 2));
    p.o_bias = take((size_t)3 * e->H * 4);
    p.o_keep = take(e->out_drop_p > 0.f ? (size_t)p.R * e->H / 4 + 64 : 0);     // output-dropout keep bits, 1 byte / 4 elements
    p.o_keep_in = take(pool ? 0 : (size_t)p.R * (p.ldx / 8) + 64);              // input-dropout keep bits, 1 byte / 8 columns
    p.o_xb = take(pool ? 0 : (size_t)p.R * p.ldx * 2);
    p.o_qkv = take(e->kind == NR_ENC_MHSA ? (size_t)p.R * 3 * e->H * 2 : 0);
    p.o_y = take((size_t)p.R * e->H * 2);
    p.o_e = take((size_t)p.R * e->Q * 2);
    p.o_alpha = take((size_t)p.R * 4);
    p.o_logits = take((size_t)p.R * 4);
    p.save_bytes = o;
    o = 0;
    p.o_dz = take((size_t)p.R * e->Q * 2);
    p.o_dy = take(pool ? 0 : (size_t)p.R * e->H * 2);
    p.o_dqkv = take(e->kind == NR_ENC_MHSA ? (size_t)p.R * 3 * e->H * 2 : 0);
    p.o_rowlist = take(e->kind == NR_ENC_MHSA ? 64 + (size_t)p.R * 8 + (size_t)n : 0);     // [count, pad to 64 bytes, list (R), rowmap (R), allpad (n bytes)]
    p.o_dbpart = take(e->kind == NR_ENC_MHSA ? dbpart_bytes(e->heads, e->H / e->heads) : 0);
    p.ws_bytes = o;
    return p;
}

int validate(const nr_title_encoder_t* e, const char* who) {
    NR_REQUIRE(e != nullptr, "%s: enc is NULL", who);
    NR_REQUIRE(e->kind >= NR_ENC_MHSA && e->kind <= NR_ENC_POOL, "%s: bad kind %d", who, e->kind);
    NR_REQUIRE(e->S >= 1 && (e->S <= 64 || e->kind == NR_ENC_POOL) && e->S <= 4096, "%s: S=%d outside the envelope", who, e->S);
    NR_REQUIRE(e->Dm >= 4 && (e->Dm & 3) == 0, "%s: Dm=%d must be a positive multiple of 4", who, e->Dm);
    NR_REQUIRE(e->H >= 8 && (e->H & 7) == 0, "%s: H=%d must be a positive multiple of 8", who, e->H);
    NR_REQUIRE(e->Q >= 8 && (e->Q & 7) == 0 && e->Q <= 256, "%s: Q=%d must be a multiple of 8 in [8,256]", who, e->Q);
    NR_REQUIRE(e->p_w1 && e->p_b1 && e->p_w2 && e->p_b2, "%s: pooling parameter is NULL", who);
    NR_REQUIRE(e->out_drop_p >= 0.f && e->out_drop_p < 1.f, "%s: out_drop_p outside [0,1)", who);
    if (e->kind == NR_ENC_MHSA) {
        NR_REQUIRE(e->heads >= 1 && e->H % e->heads == 0, "%s: H=%d not divisible by heads=%d", who, e->H, e->heads);
        for (int i = 0; i < 3; ++i) NR_REQUIRE(e->w[i] && e->b[i], "%s: W_Q/W_K/W_V parameter is NULL", who);
    } else if (e->kind == NR_ENC_CONV) {
        NR_REQUIRE(e->w[0] && e->b[0], "%s: Conv1d parameter is NULL", who);
        NR_REQUIRE(e->out_drop_p == 0.f, "%s: CONV encoder has no output dropout", who);
    } else {
        NR_REQUIRE(e->Dm == e->H, "%s: POOL kind needs Dm == H (got %d, %d)", who, e->Dm, e->H);
        NR_REQUIRE(e->out_drop_p == 0.f, "%s: POOL kind has no output dropout", who);
    }
    return 0;
}

int pack_multi(const PackArgs* a, bf16* const* out, int n, cudaStream_t st) {
    PackJobs j;
    int64_t most = 0;
    for (int i = 0; i < n; ++i) {
        j.a[i] = a[i]; j.out[i] = out[i];
        const int64_t total = (int64_t)a[i].rows * a[i].ldo;
        if (total > most) most = total;
    }
    const int blocks = (int)((most + 255) / 256 < 592 ? (most + 255) / 256 : 592);
    ProfScope ps("enc.pack_w", st);
    pack_weights_multi_kernel<<<dim3(blocks, n), 256, 0, st>>>(j);
    return check_last("enc.pack_w");
}

template <int DK>
int launch_attn_fwd(const bf16* qkv, const float* mask, bf16* ctx, int64_t n, int S, int SP, int heads, float p,
                    uint64_t seed, uint8_t* keep, cudaStream_t st) {
    const size_t smem = attn_bf16_fwd_smem(S, DK);
    NR_CUDA(cudaFuncSetAttribute(attn_fwd_bf16_kernel<DK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ProfScope ps("enc.attn_fwd", st);
    attn_fwd_bf16_kernel<DK><<<(unsigned)(n * ((heads + 3) / 4)), 128, smem, st>>>(
        qkv, mask, ctx, S, SP, heads, 1.0f / sqrtf((float)DK), p > 0.f ? drop_threshold(p) : 0u,
        p > 0.f ? 1.f / (1.f - p) : 1.f, seed, keep);
    return check_last("enc.attn_fwd");
}
template <int DK>
int launch_attn_bwd(const bf16* qkv, const float* mask, const bf16* dctx, bf16* dqkv, int64_t n, int S, int SP,
                    int heads, cudaStream_t st) {
    const size_t smem = attn_bf16_bwd_smem(S, DK);
    NR_CUDA(cudaFuncSetAttribute(attn_bwd_bf16_kernel<DK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ProfScope ps("enc.attn_bwd", st);
    attn_bwd_bf16_kernel<DK><<<(unsigned)(n * ((heads + 3) / 4)), 128, smem, st>>>(qkv, mask, dctx, dqkv, S, SP, heads,
                                                                                  1.0f / sqrtf((float)DK));
    return check_last("enc.attn_bwd");
}
// warp-level tensor-core attention (S <= 31 => pitch 32, head dim <= 32)
template <int DK, bool MASK>
int launch_attn_fwd_mma(const bf16* qkv, const float* mask, bf16* ctx, int64_t n, int S, int heads, float p,
                        uint64_t seed, uint8_t* keep, bool hm, cudaStream_t st) {
    const size_t smem = 4 * amma::fwd_smem_per_warp(DK);
    auto kern = amma::attn_fwd_mma_kernel<DK, MASK>;
    NR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // adjacent heads of a title share 32-byte sectors (40-byte head slices of a 2400-byte row), and the four warps of a
    // CTA fetch them at slightly different times: the L1 has to keep them, so the carve-out leaves it room (measured:
    // max-shared carve-out = 1.5x slower)
    const int cap = 3, carve = 70;      // resident CTAs, shared-memory carve-out (%): measured best
    NR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
    const int64_t items = n * heads;
    NR_REQUIRE(items < (int64_t)1 << 30, "attention: n*heads = %lld too large for one launch", (long long)items);
    int64_t grid = (items + 3) / 4;
    int per_sm = 1;
    NR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, smem));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > cap) per_sm = cap;
    if (grid > (int64_t)sm_count() * per_sm) grid = (int64_t)sm_count() * per_sm;    // persistent: one wave
    ProfScope ps("enc.attn_fwd", st);
    kern<<<(unsigned)grid, 128, smem, st>>>(qkv, mask, ctx, items, S, heads, 1.0f / sqrtf((float)DK),
                                            p > 0.f ? drop_threshold(p) : 0u, p > 0.f ? 1.f / (1.f - p) : 1.f, seed, keep,
                                            hm ? 3 * DK : DK, hm ? DK : heads * DK);
    return check_last("enc.attn_fwd");
}
// rowmap != NULL: compact dqkv rows + bias gradients db[0..2] from the kernel (amma::attn_bwd_mma_kernel COMPACT)
template <int DK, bool MASK, bool COMPACT = false>
int launch_attn_bwd_mma(const bf16* qkv, const float* mask, const bf16* dctx, bf16* dqkv, int64_t n, int S, int heads,
                        bool hm, cudaStream_t st, const int32_t* rowmap = nullptr, float* const* db = nullptr,
                        const uint8_t* allpad = nullptr) {
    const size_t smem = 4 * amma::bwd_smem_per_warp(DK);
    auto kern = amma::attn_bwd_mma_kernel<DK, MASK, COMPACT>;
    NR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // adjacent heads of a title share 32-byte sectors (40-byte head slices of a 2400-byte row), and the four warps of a
    // CTA fetch them at slightly different times: the L1 has to keep them, so the carve-out leaves it room (measured:
    // max-shared carve-out = 1.5x slower)
    const int cap = 2, carve = 50;      // resident CTAs, shared-memory carve-out (%): measured best
    NR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
    const int64_t items = n * heads;
    NR_REQUIRE(items < (int64_t)1 << 30, "attention: n*heads = %lld too large for one launch", (long long)items);
    int64_t grid = (items + 3) / 4;
    int per_sm = 1;
    NR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, smem));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > cap) per_sm = cap;
    if (grid > (int64_t)sm_count() * per_sm) grid = (int64_t)sm_count() * per_sm;
    if (COMPACT) {
        // a warp's items are 4 * grid apart: a multiple of heads keeps every warp on one head (its bias sums stay in
        // registers); surplus CTAs of a tiny launch simply find no item
        int g4 = heads, a = 4;
        while (a) { const int t = g4 % a; g4 = a; a = t; }            // gcd(heads, 4)
        const int m = heads / g4;
        grid = grid / m * m;
        if (grid < m) grid = m;
    }
    ProfScope ps("enc.attn_bwd", st);
    kern<<<(unsigned)grid, 128, smem, st>>>(qkv, mask, dctx, dqkv, items, S, heads, 1.0f / sqrtf((float)DK),
                                            hm ? 3 * DK : DK, hm ? DK : heads * DK, rowmap, COMPACT ? db[3] : nullptr, allpad);
    if (COMPACT)        // db[3] = per-warp partial sums (dbpart_bytes), folded in a fixed order
        amma::db_reduce_kernel<<<(3 * heads * DK + 127) / 128, 128, 0, st>>>(db[3], (int)grid * 4, heads, DK, db[0], db[1], db[2]);
    return check_last("enc.attn_bwd");
}
// blocked variant: 32 < pitch <= 64 (the user encoder's 50 clicked titles)
template <int DK, bool MASK>
int launch_attn_fwd_blk(const bf16* qkv, const float* mask, bf16* ctx, int64_t n, int S, int heads, float p,
                        uint64_t seed, uint8_t* keep, bool hm, cudaStream_t st) {
    const size_t smem = 4 * amma::fwd_blk_smem_per_warp(DK, 2);
    auto kern = amma::attn_fwd_blk_kernel<DK, MASK, 2>;
    NR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t items = n * heads;
    NR_REQUIRE(items < (int64_t)1 << 30, "attention: n*heads = %lld too large for one launch", (long long)items);
    int64_t grid = (items + 3) / 4;
    int per_sm = 1;
    NR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, smem));
    if (per_sm < 1) per_sm = 1;
    if (grid > (int64_t)sm_count() * per_sm) grid = (int64_t)sm_count() * per_sm;
    ProfScope ps("enc.attn_fwd", st);
    kern<<<(unsigned)grid, 128, smem, st>>>(qkv, mask, ctx, items, S, heads, 1.0f / sqrtf((float)DK),
                                            p > 0.f ? drop_threshold(p) : 0u, p > 0.f ? 1.f / (1.f - p) : 1.f, seed, keep,
                                            hm ? 3 * DK : DK, hm ? DK : heads * DK);
    return check_last("enc.attn_fwd");
}
template <int DK, bool MASK>
int launch_attn_bwd_blk(const bf16* qkv, const float* mask, const bf16* dctx, bf16* dqkv, int64_t n, int S, int heads,
                        bool hm, cudaStream_t st) {
    const size_t smem = 4 * amma::bwd_blk_smem_per_warp(DK, 2);
    auto kern = amma::attn_bwd_blk_kernel<DK, MASK, 2>;
    NR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t items = n * heads;
    NR_REQUIRE(items < (int64_t)1 << 30, "attention: n*heads = %lld too large for one launch", (long long)items);
    int64_t grid = (items + 3) / 4;
    int per_sm = 1;
    NR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, smem));
    if (per_sm < 1) per_sm = 1;
    if (grid > (int64_t)sm_count() * per_sm) grid = (int64_t)sm_count() * per_sm;
    ProfScope ps("enc.attn_bwd", st);
    kern<<<(unsigned)grid, 128, smem, st>>>(qkv, mask, dctx, dqkv, items, S, heads, 1.0f / sqrtf((float)DK),
                                            hm ? 3 * DK : DK, hm ? DK : heads * DK);
    return check_last("enc.attn_bwd");
}
#define NR_BLK_FWD(DKV) (mask ? launch_attn_fwd_blk<DKV, true>(qkv, mask, ctx, n, S, heads, p, seed, keep, hm, st) \
                              : launch_attn_fwd_blk<DKV, false>(qkv, mask, ctx, n, S, heads, p, seed, keep, hm, st))
#define NR_BLK_BWD(DKV) (mask ? launch_attn_bwd_blk<DKV, true>(qkv, mask, dctx, dqkv, n, S, heads, hm, st) \
                              : launch_attn_bwd_blk<DKV, false>(qkv, mask, dctx, dqkv, n, S, heads, hm, st))
#define NR_MMA_FWD(DKV) (mask ? launch_attn_fwd_mma<DKV, true>(qkv, mask, ctx, n, S, heads, p, seed, keep, hm, st) \
                              : launch_attn_fwd_mma<DKV, false>(qkv, mask, ctx, n, S, heads, p, seed, keep, hm, st))
#define NR_MMA_BWD(DKV) (rowmap ? (mask ? launch_attn_bwd_mma<DKV, true, true>(qkv, mask, dctx, dqkv, n, S, heads, hm, st, rowmap, db, allpad) \
                                        : launch_attn_bwd_mma<DKV, false, true>(qkv, mask, dctx, dqkv, n, S, heads, hm, st, rowmap, db, allpad)) \
                                : (mask ? launch_attn_bwd_mma<DKV, true>(qkv, mask, dctx, dqkv, n, S, heads, hm, st) \
                                        : launch_attn_bwd_mma<DKV, false>(qkv, mask, dctx, dqkv, n, S, heads, hm, st)))

#define NR_ENC_DISPATCH_DK(dk, CALL)                                                   \
    switch (dk) {                                                                      \
        case 8: { constexpr int DK = 8; return CALL; }                                 \
        case 16: { constexpr int DK = 16; return CALL; }                               \
        case 20: { constexpr int DK = 20; return CALL; }                               \
        case 32: { constexpr int DK = 32; return CALL; }                               \
        case 40: { constexpr int DK = 40; return CALL; }                               \
        default: set_error("title encoder: head dim %d not in compiled set {8,16,20,32,40}", dk); return 1; \
    }
int attn_fwd_b(const bf16* qkv, const float* mask, bf16* ctx, int64_t n, int S, int SP, int heads, int dk, float p,
               uint64_t seed, uint8_t* keep, cudaStream_t st) {
    const bool hm = qkv_head_major(SP, dk);
    if (SP == 32) {
        switch (dk) {
            case 8: return NR_MMA_FWD(8);
            case 16: return NR_MMA_FWD(16);
            case 20: return NR_MMA_FWD(20);
            case 32: return NR_MMA_FWD(32);
            default: break;
        }
    }
    if (SP == 64) {
        switch (dk) {
            case 8: return NR_BLK_FWD(8);
            case 16: return NR_BLK_FWD(16);
            case 20: return NR_BLK_FWD(20);
            case 32: return NR_BLK_FWD(32);
            default: break;
        }
    }
    NR_ENC_DISPATCH_DK(dk, launch_attn_fwd<DK>(qkv, mask, ctx, n, S, SP, heads, p, seed, keep, st));
}
// rowmap (pitch 32 with a warp-MMA head width only, see compact_supported): dqkv rows of real tokens only, db[0..2] filled
// here through the scratch db[3] (dbpart_bytes)
static bool compact_supported(int SP, int dk) { return SP == 32 && (dk == 8 || dk == 16 || dk == 20 || dk == 32); }
int attn_bwd_b(const bf16* qkv, const float* mask, const bf16* dctx, bf16* dqkv, int64_t n, int S, int SP, int heads,
               int dk, cudaStream_t st, const int32_t* rowmap = nullptr, float* const* db = nullptr,
               const uint8_t* allpad = nullptr) {
    const bool hm = qkv_head_major(SP, dk);
    NR_REQUIRE(!rowmap || compact_supported(SP, dk), "attention backward: compact rows need pitch 32 and a warp-MMA head width");
    if (SP == 32) {
        switch (dk) {
            case 8: return NR_MMA_BWD(8);
            case 16: return NR_MMA_BWD(16);
            case 20: return NR_MMA_BWD(20);
            case 32: return NR_MMA_BWD(32);
            default: break;
        }
    }
    if (SP == 64) {
        switch (dk) {
            case 8: return NR_BLK_BWD(8);
            case 16: return NR_BLK_BWD(16);
            case 20: return NR_BLK_BWD(20);
            case 32: return NR_BLK_BWD(32);
            default: break;
        }
    }
    NR_ENC_DISPATCH_DK(dk, launch_attn_bwd<DK>(qkv, mask, dctx, dqkv, n, S, SP, heads, st));
}

}  // namespace

extern "C" {

size_t nr_title_encoder_save_bytes(const nr_title_encoder_t* enc, int64_t n) {
    if (!enc || n <= 0) return 256;
    return make_plan(enc, n).save_bytes + 1024;
}
size_t nr_title_encoder_workspace(const nr_title_encoder_t* enc, int64_t n) {
    if (!enc || n <= 0) return 256;
    return make_plan(enc, n).ws_bytes + 1024;
}

static bool infer_supported(const nr_title_encoder_t* e) {
    const int SP = round_up(e->S + 1, 32), Kp = round_up(e->Dm + 1, 64);
    if (e->S > 31 || e->out_drop_p != 0.f) return false;
    if (e->kind == NR_ENC_CONV) return fe::fused_conv_supported(SP, e->H, e->Q, Kp);
    if (e->kind != NR_ENC_MHSA || e->heads < 1 || e->H % e->heads) return false;
    return fe::fused_supported(SP, e->H / e->heads, e->heads, e->H, e->Q, Kp);
}
struct InferPlan { int SP, ldx, Kp, HK; int64_t R; size_t o_wb, o_w1b, o_xb, bytes; };
static InferPlan make_infer_plan(const nr_title_encoder_t* e, int64_t n) {
    InferPlan p;
    p.SP = round_up(e->S + 1, 32); p.ldx = round_up(e->Dm + 1, 64); p.Kp = p.ldx; p.HK = round_up(e->H, 64);
    p.R = n * p.SP;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o += (bytes + 1023) & ~size_t(1023); return r; };
    p.o_wb = take((size_t)3 * e->H * p.Kp * 2);            // MHSA: (3H x Kp); CONV: (H x 3 Kp)
    p.o_w1b = take((size_t)e->Q * p.HK * 2);
    p.o_xb = take((size_t)p.R * p.ldx * 2);
    p.bytes = o;
    return p;
}
size_t nr_title_encoder_infer_bytes(const nr_title_encoder_t* enc, int64_t n) {
    if (!enc || n <= 0 || !infer_supported(enc)) return 0;
    return make_infer_plan(enc, n).bytes + 1024;
}

// save == NULL: inference.  Nothing is kept for a backward; the whole encoder after the row staging is ONE kernel.
static int title_encoder_infer(const nr_title_encoder_t* enc, const nr_rows_t* rows, int64_t n, float* out, void* ws,
                               size_t ws_bytes, cudaStream_t st) {
    const char* who = "nr_title_encoder_fwd(inference)";
    NR_REQUIRE(infer_supported(enc), "%s: save == NULL needs an MHSA or CONV title encoder inside the fused envelope (S <= 31; "
               "MHSA: head width 16 or 20, heads %% 4 == 0, no output dropout); pass a save buffer otherwise", who);
    NR_REQUIRE(rows->drop_p == 0.f, "%s: input dropout in inference mode", who);
    NR_REQUIRE(ws, "%s: workspace is NULL", who);
    const InferPlan p = make_infer_plan(enc, n);
    char* w = (char*)(((uintptr_t)ws + 1023) & ~(uintptr_t)1023);
    NR_REQUIRE((size_t)(w - (char*)ws) + p.bytes <= ws_bytes, "%s: workspace too small (%zu < %zu)", who, ws_bytes, p.bytes + 1024);
    const bool conv = enc->kind == NR_ENC_CONV;
    const int S = enc->S, Dm = enc->Dm, H = enc->H, Q = enc->Q, dk = conv ? 20 : H / enc->heads;
    bf16* wb = (bf16*)(w + p.o_wb);
    bf16* w1b = (bf16*)(w + p.o_w1b);
    bf16* xb = (bf16*)(w + p.o_xb);
    {
        PackArgs ja[2];
        bf16* jo[2];
        ja[0] = PackArgs{enc->p_w1, nullptr, nullptr, 0, Q, H, 0, Q, p.HK, 1}; jo[0] = w1b;
        if (conv) ja[1] = PackArgs{enc->w[0], nullptr, nullptr, 2, H, Dm, p.Kp, H, 3 * p.Kp, 1, enc->b[0], nullptr, nullptr};
        else ja[1] = PackArgs{enc->w[0], enc->w[1], enc->w[2], 0, H, Dm, 0, 3 * H, p.Kp, 3, enc->b[0], enc->b[1], enc->b[2], dk};
        jo[1] = wb;
        if (int e = pack_multi(ja, jo, 2, st)) return e;
    }
    {
        const RowsView xv = make_rows_view(rows, n, S, Dm);
        NR_REQUIRE(p.R < ((int64_t)1 << 31) && p.ldx <= 2048, "%s: n*pitch = %lld rows / ldx %d outside the staging envelope",
                   who, (long long)p.R, p.ldx);
        const int cx = p.ldx >> 3;
        const int ry = 256 / cx > 0 ? 256 / cx : 1;
        int64_t blocks = (p.R + ry - 1) / ry;
        if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
        ProfScope ps("enc.stage", st);
        stage_rows_kernel<<<(unsigned)blocks, dim3(cx, ry), 0, st>>>(xv, xb, (int)p.R, p.SP, p.ldx, nullptr);
        if (int e = check_last("enc.stage")) return e;
    }
    fe::FusedArgs fa{xb, p.R, p.ldx, Dm + 1, wb, p.Kp, w1b, p.HK, nullptr};
    fe::FusedParams fp{};
    fp.R = p.R; fp.n = n; fp.S = S; fp.H = H; fp.Q = Q; fp.heads = conv ? 4 : enc->heads;
    fp.out = out; fp.mask = enc->mask; fp.b1 = enc->p_b1; fp.w2 = enc->p_w2; fp.b2 = enc->p_b2;
    fp.scale = 1.0f / sqrtf((float)dk);
    fp.dscale = 1.f;
    if (conv) return fe::launch_fused<20, false, true>(fa, fp, st);
    if (dk == 20) return fe::launch_fused<20, false, false>(fa, fp, st);
    return fe::launch_fused<16, false, false>(fa, fp, st);
}

int nr_title_encoder_fwd(const nr_title_encoder_t* enc, const nr_rows_t* rows, int64_t n, float* out, void* save,
                         size_t save_bytes, void* ws, size_t ws_bytes, void* stream) {
    const char* who = "nr_title_encoder_fwd";
    if (int e = validate(enc, who)) return e;
    if (int e = validate_rows(rows, who)) return e;
    NR_REQUIRE(out, "%s: NULL pointer", who);
    NR_REQUIRE(n >= 0, "%s: n < 0", who);
    if (n == 0) return 0;
    if (!save) return title_encoder_infer(enc, rows, n, out, ws, ws_bytes, (cudaStream_t)stream);
    const Plan p = make_plan(enc, n);
    char* sv = (char*)(((uintptr_t)save + 1023) & ~(uintptr_t)1023);
    NR_REQUIRE((size_t)(sv - (char*)save) + p.save_bytes <= save_bytes, "%s: save buffer too small (%zu < %zu)", who,
               save_bytes, p.save_bytes + 1024);
    cudaStream_t st = (cudaStream_t)stream;
    const int S = enc->S, Dm = enc->Dm, H = enc->H, Q = enc->Q;
    bf16* wb = (bf16*)(sv + p.o_wb);
    bf16* w1b = (bf16*)(sv + p.o_w1b);
    bf16* xb = (bf16*)(sv + p.o_xb);
    bf16* qkv = (bf16*)(sv + p.o_qkv);
    bf16* y = (bf16*)(sv + p.o_y);
    bf16* eb = (bf16*)(sv + p.o_e);
    float* alpha = (float*)(sv + p.o_alpha);
    float* logits = (float*)(sv + p.o_logits);

    const bool pool = enc->kind == NR_ENC_POOL;
    if (pool) {
        NR_REQUIRE(rows->mode != NR_ROWS_NEWS && rows->drop_p == 0.f, "%s: POOL kind takes dense or token-gathered rows without dropout", who);
        xb = y;                     // the staged input IS the pooling input
    }
    // 1. weights -> bf16 GEMM operands (forward operands + the transposes backward needs), one launch
    {
        PackArgs ja[4];
        bf16* jo[4];
        int nj = 0;
        ja[nj] = PackArgs{enc->p_w1, nullptr, nullptr, 0, Q, H, 0, Q, p.HK, 1}; jo[nj++] = w1b;
        ja[nj] = PackArgs{enc->p_w1, nullptr, nullptr, 1, Q, H, 0, H, p.QK, 1}; jo[nj++] = (bf16*)(sv + p.o_w1t);
        if (enc->kind == NR_ENC_MHSA) {
            const int hm_dk = qkv_head_major(p.SP, H / enc->heads) ? H / enc->heads : 0;
            ja[nj] = PackArgs{enc->w[0], enc->w[1], enc->w[2], 0, H, Dm, 0, 3 * H, p.Kp, 3, enc->b[0], enc->b[1], enc->b[2], hm_dk};
            jo[nj++] = wb;
            ja[nj] = PackArgs{enc->w[0], enc->w[1], enc->w[2], 1, H, Dm, 0, Dm, p.K3, 3, nullptr, nullptr, nullptr, hm_dk};
            jo[nj++] = (bf16*)(sv + p.o_wt);
        } else if (!pool) {
            ja[nj] = PackArgs{enc->w[0], nullptr, nullptr, 2, H, Dm, p.Kp, H, 3 * p.Kp, 1, enc->b[0], nullptr, nullptr};
            jo[nj++] = wb;
            ja[nj] = PackArgs{enc->w[0], nullptr, nullptr, 3, H, Dm, p.HK, Dm, 3 * p.HK, 1}; jo[nj++] = (bf16*)(sv + p.o_wt);
        }
        if (int e = pack_multi(ja, jo, nj, st)) return e;
    }
    // 2. stage rows: gather + input dropout + bf16, padded pitch
    {
        const RowsView xv = make_rows_view(rows, n, S, Dm);
        NR_REQUIRE(p.R < ((int64_t)1 << 31) && p.ldx <= 2048, "%s: n*pitch = %lld rows / ldx %d outside the staging envelope",
                   who, (long long)p.R, p.ldx);
        const int cx = p.ldx >> 3;
        const int ry = 256 / cx > 0 ? 256 / cx : 1;
        int64_t blocks = (p.R + ry - 1) / ry;
        if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
        ProfScope ps("enc.stage", st);
        stage_rows_kernel<<<(unsigned)blocks, dim3(cx, ry), 0, st>>>(xv, xb, (int)p.R, p.SP, p.ldx,
                                                                     pool ? nullptr : (uint8_t*)(sv + p.o_keep_in));
        if (int e = check_last("enc.stage")) return e;
    }
    // 3. token mixer
    const int one = 0;
    if (enc->kind == NR_ENC_MHSA && fused_enabled() && fe::fused_supported(p.SP, H / enc->heads, enc->heads, H, Q, p.Kp)) {
        // titles (pitch 32): QKV projection, attention, fc1 and pooling in ONE persistent kernel (nr_fused_enc.cuh)
        fe::FusedArgs fa{xb, p.R, p.ldx, Dm + 1, wb, p.Kp, w1b, p.HK, y};
        fe::FusedParams fp{};
        fp.R = p.R; fp.n = n; fp.S = S; fp.H = H; fp.Q = Q; fp.heads = enc->heads;
        fp.out = out; fp.mask = enc->mask; fp.b1 = enc->p_b1; fp.w2 = enc->p_w2; fp.b2 = enc->p_b2;
        fp.scale = 1.0f / sqrtf((float)(H / enc->heads));
        fp.qkv = qkv; fp.E = eb; fp.alpha = alpha;
        fp.keep = enc->out_drop_p > 0.f ? (uint8_t*)(sv + p.o_keep) : nullptr;
        fp.thr = enc->out_drop_p > 0.f ? drop_threshold(enc->out_drop_p) : 0u;
        fp.dscale = enc->out_drop_p > 0.f ? 1.f / (1.f - enc->out_drop_p) : 1.f;
        fp.seed = rows->seed;
#ifdef NR_FE_PROBES
        if (const char* e = getenv("NR_FE_DBG")) fp.dbg = atoi(e);      // timing probes: development builds only
#endif
        if (H / enc->heads == 20) return fe::launch_fused<20, true, false>(fa, fp, st);
        return fe::launch_fused<16, true, false>(fa, fp, st);
    }
    if (enc->kind == NR_ENC_CONV && fused_enabled() && fe::fused_conv_supported(p.SP, H, Q, p.Kp)) {
        // NAML titles: Conv1d(k=3) passes write the Y tile in shared memory, then fc1 + pooling — one kernel
        fe::FusedArgs fa{xb, p.R, p.ldx, Dm + 1, wb, p.Kp, w1b, p.HK, y};
        fe::FusedParams fp{};
        fp.R = p.R; fp.n = n; fp.S = S; fp.H = H; fp.Q = Q; fp.heads = 4;
        fp.out = out; fp.mask = enc->mask; fp.b1 = enc->p_b1; fp.w2 = enc->p_w2; fp.b2 = enc->p_b2;
        fp.E = eb; fp.alpha = alpha; fp.dscale = 1.f;
        return fe::launch_fused<20, true, true>(fa, fp, st);
    }
    if (pool) {
    } else if (enc->kind == NR_ENC_MHSA) {
        // K = Dm + 1: the ones column of Xb times the bias column of the packed weights
        if (int e = gemm_tc_pick(xb, p.R, Dm + 1, p.ldx, wb, 3 * H, p.Kp, 1, &one, p.R, TcEpiStoreBf16{qkv, 3 * H}, st, "enc.qkv"))
            return e;
        if (int e = attn_fwd_b(qkv, enc->mask, y, n, S, p.SP, enc->heads, H / enc->heads, enc->out_drop_p, rows->seed,
                               enc->out_drop_p > 0.f ? (uint8_t*)(sv + p.o_keep) : nullptr, st))
            return e;
    } else {
        const int sh[3] = {-1, 0, 1};
        if (int e = gemm_tc_pick(xb, p.R, Dm + 1, p.ldx, wb, H, 3 * p.Kp, 3, sh, p.R, TcEpiStoreBf16{y, H}, st, "enc.conv"))
            return e;
    }
    // 4. additive attention: fc1 + tanh + fc2 in the GEMM epilogue, then the softmax-like tail
    if (int e = gemm_tc(y, p.R, H, H, w1b, Q, p.HK, 1, &one, p.R,
                        EpiTanhLogit{eb, Q, enc->p_b1, enc->p_w2, enc->p_b2, logits, 0.f}, st, "enc.fc1"))
        return e;
    ProfScope ps("enc.pool_fwd", st);
    // one warp per sequence when there are enough sequences to fill the machine with warps; a few hundred long
    // sequences (the user encoder: 128 x 50) are latency bound that way and keep the block-per-sequence kernel
    if (p.SP <= 64 && (H & 7) == 0 && H <= 1024 && n >= 16 * (int64_t)sm_count()) {
        auto launch = [&](auto kern) {
            int per_sm = 1;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, 0);
            int64_t grid = (n + 7) / 8;                       // persistent warps, exactly one resident wave
            if (grid > (int64_t)sm_count() * (per_sm > 0 ? per_sm : 1)) grid = (int64_t)sm_count() * (per_sm > 0 ? per_sm : 1);
            kern<<<(unsigned)grid, 256, 0, st>>>(y, logits, enc->mask, out, alpha, n, S, p.SP, H);
        };
        if (H <= 256) launch(pool_tail_fwd_warp_kernel<1>);
        else if (H <= 512) launch(pool_tail_fwd_warp_kernel<2>);
        else launch(pool_tail_fwd_warp_kernel<4>);
    } else {
        // few sequences (fewer than two per SM): 8 warps per block so that all of a sequence's rows are in flight at once
        const int threads = (S > 16 && n <= 2 * (int64_t)sm_count()) ? 256 : 128;
        pool_tail_fwd_bf16_kernel<<<(unsigned)n, threads, (p.SP + (threads / 32) * (size_t)H) * sizeof(float), st>>>(y, logits, enc->mask, out, alpha, S, p.SP, H);
    }
    return check_last("enc.pool_fwd");
}

int nr_title_encoder_bwd(const nr_title_encoder_t* enc, const nr_rows_t* rows, int64_t n, const float* g,
                         const void* save, size_t save_bytes, const nr_title_encoder_grads_t* gr, void* ws,
                         size_t ws_bytes, void* stream) {
    const char* who = "nr_title_encoder_bwd";
    if (int e = validate(enc, who)) return e;
    if (int e = validate_rows(rows, who)) return e;
    NR_REQUIRE(g && save && gr && ws, "%s: NULL pointer", who);
    NR_REQUIRE(gr->dp_w1 && gr->dp_b1 && gr->dp_w2 && gr->dp_b2, "%s: pooling gradient pointer is NULL", who);
    NR_REQUIRE(!(gr->dtable && rows->mode == NR_ROWS_DENSE), "%s: dtable given for dense rows", who);
    NR_REQUIRE(!(gr->dx_dense && rows->mode != NR_ROWS_DENSE), "%s: dx_dense given for gathered rows", who);
    if (n == 0) return 0;
    const Plan p = make_plan(enc, n);
    const char* sv = (const char*)(((uintptr_t)save + 1023) & ~(uintptr_t)1023);
    NR_REQUIRE((size_t)(sv - (const char*)save) + p.save_bytes <= save_bytes, "%s: save buffer too small", who);
    char* wsb = (char*)(((uintptr_t)ws + 1023) & ~(uintptr_t)1023);
    NR_REQUIRE((size_t)(wsb - (char*)ws) + p.ws_bytes <= ws_bytes, "%s: workspace too small (%zu < %zu)", who, ws_bytes,
               p.ws_bytes + 1024);
    cudaStream_t st = (cudaStream_t)stream;
    const int S = enc->S, Dm = enc->Dm, H = enc->H, Q = enc->Q;
    const bf16* xb = (const bf16*)(sv + p.o_xb);
    const bf16* qkv = (const bf16*)(sv + p.o_qkv);
    const bf16* y = (const bf16*)(sv + p.o_y);
    const bf16* eb = (const bf16*)(sv + p.o_e);
    const float* alpha = (const float*)(sv + p.o_alpha);
    bf16* dz = (bf16*)(wsb + p.o_dz);
    bf16* dy = (bf16*)(wsb + p.o_dy);
    bf16* dqkv = (bf16*)(wsb + p.o_dqkv);
    const bf16* w1t = (const bf16*)(sv + p.o_w1t);      // packed by the forward
    const bf16* wt = (const bf16*)(sv + p.o_wt);
    const bool pool = enc->kind == NR_ENC_POOL;
    const bool want_dx = gr->dtable || gr->dx_dense;
    const int one = 0;

    // 1. pooling tail backward: dz, dw2, db2, db1
    {
        ProfScope ps("enc.pool_bwd", st);
        if (p.SP <= 64 && (H & 7) == 0 && (Q & 7) == 0 && H <= 1024 && Q <= 512 && n >= 16 * (int64_t)sm_count()) {
            const size_t sm = (2 * Q + 1) * sizeof(float);
            auto launch = [&](auto kern) {
                int per_sm = 1;
                cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, sm);
                int64_t grid = (n + 7) / 8;                   // persistent warps, exactly one resident wave
                if (grid > (int64_t)sm_count() * (per_sm > 0 ? per_sm : 1)) grid = (int64_t)sm_count() * (per_sm > 0 ? per_sm : 1);
                kern<<<(unsigned)grid, 256, sm, st>>>(g, y, eb, alpha, enc->p_w2, dz, gr->dp_w2, gr->dp_b2, gr->dp_b1, n, S,
                                                     p.SP, H, Q);
            };
#define NR_POOL_BWD(KH, KQ) launch(pool_tail_bwd_warp_kernel<KH, KQ>)
            if (H <= 256) { if (Q <= 256) NR_POOL_BWD(1, 1); else NR_POOL_BWD(1, 2); }
            else if (H <= 512) { if (Q <= 256) NR_POOL_BWD(2, 1); else NR_POOL_BWD(2, 2); }
            else { if (Q <= 256) NR_POOL_BWD(4, 1); else NR_POOL_BWD(4, 2); }
#undef NR_POOL_BWD
        } else {
            int64_t grid = n < 8 * (int64_t)sm_count() ? n : 8 * (int64_t)sm_count();
            pool_tail_bwd_bf16_kernel<<<(unsigned)grid, 256, (H + p.SP + 2 * Q) * sizeof(float), st>>>(
                g, y, eb, alpha, enc->p_w2, dz, gr->dp_w2, gr->dp_b2, gr->dp_b1, n, S, p.SP, H, Q);
        }
        if (int e = check_last("enc.pool_bwd")) return e;
    }
    // 2. dW1 (Q,H) += dz^T . y
    if (int e = gemm_tc_tn(dz, Q, Q, y, H, H, p.R, 1, &one, TnEpiPlain{gr->dp_w1, H, 0}, st, "enc.dw1")) return e;
    // 3. dY = dz . W1 + alpha g   (output-dropout mask re-applied for MHSA)
    if (!pool || want_dx) {
        EpiPoolDx ep{dy, pool ? gr->dx_dense : nullptr, H, alpha, g, S, p.SP, H,
                     enc->out_drop_p > 0.f ? (const uint8_t*)(sv + p.o_keep) : nullptr,
                     enc->out_drop_p > 0.f ? 1.f / (1.f - enc->out_drop_p) : 1.f, 0, 0, 0.f, false};
        if (int e = gemm_tc_pick(dz, p.R, Q, Q, w1t, H, p.QK, 1, &one, p.R, ep, st, "enc.pool_dx")) return e;
    }
    if (pool) return 0;
    const RowsView xv = make_rows_view(rows, n, S, Dm);
    if (enc->kind == NR_ENC_MHSA) {
        // token rows: dqkv is kept COMPACT (rows of real tokens only; dx_rowlist_kernel).  attn_bwd writes it and sums the
        // bias gradients itself, enc.dx and enc.dwqkv contract over the compact rows.  Dense / per-title rows keep every row.
        const bool gather = rows->mode == NR_ROWS_TOKEN && p.R <= INT32_MAX && compact_supported(p.SP, H / enc->heads) &&
                            dx_gather_enabled();
        int32_t* cnt = (int32_t*)(wsb + p.o_rowlist);
        int32_t* list = cnt + 16;
        int32_t* rowmap = list + p.R;
        uint8_t* allpad = (uint8_t*)(rowmap + p.R);
        if (gather) {
            ProfScope ps("enc.dx_rows", st);
            NR_CUDA(cudaMemsetAsync(cnt, 0, 4, st));
            dx_rowlist_kernel<<<(unsigned)((p.R + 1023) / 1024), 1024, 0, st>>>(xv, p.SP, p.R, list, rowmap, allpad, cnt);
            zero_tail_rows_kernel<<<TN_BK, 128, 0, st>>>(dqkv, cnt, p.R, 3 * H);
            if (int e = check_last("enc.dx_rows")) return e;
        }
        float* const db[4] = {gr->db[0], gr->db[1], gr->db[2], (float*)(wsb + p.o_dbpart)};
        if (int e = attn_bwd_b(qkv, enc->mask, dy, dqkv, n, S, p.SP, enc->heads, H / enc->heads, st, gather ? rowmap : nullptr,
                               gather ? db : nullptr, gather ? allpad : nullptr))
            return e;
        // dX (the embedding-table gradient: 15 of the 19 MB of the NRMS gradient bucket) BEFORE the weight gradient:
        // the caller can start exchanging that part of the bucket while enc.dwqkv still runs (gr->dtable_ready)
        if (want_dx)
            if (int e = gemm_tc<EpiDx, 8>(dqkv, p.R, 3 * H, 3 * H, wt, Dm, p.K3, 1, &one, p.R,
                                EpiDx{gr->dx_dense, gr->dtable, xv, p.SP, (const uint8_t*)(sv + p.o_keep_in), p.ldx / 8, 0, 0, 0,
                                      gather ? list : nullptr}, st, "enc.dx", gather ? cnt : nullptr))
                return e;
        if (gr->dtable_ready) NR_CUDA(cudaEventRecord((cudaEvent_t)gr->dtable_ready, st));
        if (int e = gemm_tc_tn(dqkv, 3 * H, 3 * H, xb, Dm + 1, p.ldx, p.R, 1, &one,
                               TnEpiQkv{gr->dw[0], gr->dw[1], gr->dw[2], gather ? nullptr : gr->db[0], gather ? nullptr : gr->db[1],
                                        gather ? nullptr : gr->db[2], H, Dm,
                                        qkv_head_major(p.SP, H / enc->heads) ? H / enc->heads : 0}, st,
                               "enc.dwqkv", gather ? list : nullptr, gather ? cnt : nullptr))
            return e;
    } else {
        const int sh[3] = {-1, 0, 1};
        if (int e = gemm_tc_tn(dy, H, H, xb, Dm + 1, p.ldx, p.R, 3, sh, TnEpiConv{gr->dw[0], gr->db[0], Dm}, st,
                               "enc.dwconv"))
            return e;
        if (want_dx) {
            const int shx[3] = {1, 0, -1};
            if (int e = gemm_tc_pick(dy, p.R, H, H, wt, Dm, 3 * p.HK, 3, shx, p.R,
                                EpiDx{gr->dx_dense, gr->dtable, xv, p.SP, (const uint8_t*)(sv + p.o_keep_in), p.ldx / 8, 0, 0, 0, nullptr}, st, "enc.dx"))
                return e;
        }
    }
    return 0;
}

int nr_tc_gemm_tn(const void* A, int32_t Mo, int64_t lda, const void* B, int32_t No, int64_t ldb, int64_t rows,
                  int32_t taps, const int32_t* shifts_host, float* C, int64_t ldc, int64_t tap_stride, void* stream) {
    NR_REQUIRE(A && B && C, "nr_tc_gemm_tn: NULL pointer");
    NR_REQUIRE(Mo > 0 && No > 0 && rows >= 0, "nr_tc_gemm_tn: bad sizes");
    int sh[3] = {0, 0, 0};
    for (int t = 0; t < taps && t < 3; ++t) sh[t] = shifts_host ? shifts_host[t] : 0;
    return gemm_tc_tn((const bf16*)A, Mo, lda, (const bf16*)B, No, ldb, rows, taps, sh, TnEpiPlain{C, ldc, tap_stride},
                      (cudaStream_t)stream, "tc.gemm_tn");
}

}  // extern "C"
