// nr_enc_kernels.cuh — SIMT kernels of the bf16 title-encoder pipeline (nr_encoder.cu):
// row staging (gather + dropout + bf16 pack), weight packing, per-head exp-attention and the
// additive-attention tails, all over the PADDED token layout:
//
//   activation row of (sequence i, position s) = i*SP + s,  SP = pitch >= S+1 (multiple of 32);
//   rows s >= S are zero ("pad rows").  Zero pad rows make the Conv1d taps of NAML.py:27-32 read
//   zeros across title edges and contribute nothing to any weight gradient, and the 32-row pitch
//   aligns every title with one epilogue warp of the tcgen05 GEMMs.
//
// Arithmetic is fp32 in registers; only the HBM representation is bf16.
#pragma once
#include "nr_common.cuh"
#include "nr_tc.cuh"

namespace nr {

typedef __nv_bfloat16 bf16;

__device__ __forceinline__ float2 bf2_to_f2(uint32_t w) {
    return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xFFFF0000u));
}
__device__ __forceinline__ float4 ld_bf4(const bf16* p) {   // 8-byte aligned
    const uint2 w = __ldg((const uint2*)p);
    const float2 a = bf2_to_f2(w.x), b = bf2_to_f2(w.y);
    return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void st_bf4(bf16* p, float4 v) {
    *(uint2*)p = make_uint2(tc::pack_bf16(v.x, v.y), tc::pack_bf16(v.z, v.w));
}

// ---- staging: X (fp32, through a RowsView incl. input dropout) -> Xb (n*SP, ldx) bf16 ------------
// column Dm holds 1.0 on real rows (the "ones column": dW GEMMs then yield the bias gradient as
// output column Dm); columns (Dm, ldx) and pad rows are zero.  One thread = 8 output columns.
// blockDim = (ldx/8, rows per block): threadIdx.x = 8-column chunk, threadIdx.y = row slot, so no
// per-thread division by the row width and one id lookup (broadcast) per row.
static __global__ void __launch_bounds__(256)
stage_rows_kernel(RowsView x, bf16* __restrict__ xb, int R, int SP, int ldx, uint8_t* __restrict__ keep_in) {
    const int c = threadIdx.x * 8;
    const uint64_t key = x.thr ? salted(x.seed) : 0;
    for (int row = blockIdx.x * blockDim.y + threadIdx.y; row < R; row += gridDim.x * blockDim.y) {
        const int seq = row / SP;
        const int s = row - seq * SP;
        uint4 o = make_uint4(0u, 0u, 0u, 0u);
        if (s < x.S) {
            const int64_t r = (int64_t)seq * x.S + s;
            const float* src;
            bool pad_token = false;
            if (x.mode == NR_ROWS_DENSE) src = x.dense + r * x.D;
            else if (x.mode == NR_ROWS_TOKEN) { const int64_t id = x.id_of(seq, s); pad_token = id == 0; src = x.table + id * x.D; }
            else src = x.table + x.id_of(seq, s) * ((int64_t)x.S * x.D) + (int64_t)s * x.D;
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
            if (c < x.D) a = ldg4(src + c);
            if (c + 4 < x.D) b = ldg4(src + c + 4);
            // Pad tokens (id 0 = padding_idx: 62 % of MIND title positions): their embedding row is zero, dropout(0) = 0
            // whatever the mask, and their gradient is dropped (NRMS.py:73), so the Philox draw is skipped when the
            // loaded values really are all zero (a non-zero row 0 takes the ordinary path).
            const bool skip_draw = pad_token && a.x == 0.f && a.y == 0.f && a.z == 0.f && a.w == 0.f &&
                                   b.x == 0.f && b.y == 0.f && b.z == 0.f && b.w == 0.f;
            if (x.thr && skip_draw) {
                if (keep_in) keep_in[(int64_t)row * (ldx >> 3) + threadIdx.x] = 0;
            } else if (x.thr) {
                // the 8 keep bits of this thread's columns are saved (1 byte, padded-row indexing) so that the
                // enc.dx epilogue re-applies the mask from a 4-byte load instead of eight Philox calls per chunk
                // this thread's 8 columns are exactly one 8-element group of the input stream: ONE Philox call
                uint32_t bits = 0;
                if (c < x.D) bits = dropout8_bits_key(in_stream_index(r, c, x.D) >> 3, key, NR_STREAM_IN, x.thr);
                if (c + 4 >= x.D) bits &= 0xFu;
                a = apply_keep4(a, bits & 0xFu, x.scale);
                b = apply_keep4(b, bits >> 4, x.scale);
                if (keep_in) keep_in[(int64_t)row * (ldx >> 3) + threadIdx.x] = (uint8_t)bits;
            }
            if (c <= x.D && x.D < c + 8) {           // the ones column
                if (x.D < c + 4) a.x = 1.0f; else b.x = 1.0f;    // D % 4 == 0
            }
            o = make_uint4(tc::pack_bf16(a.x, a.y), tc::pack_bf16(a.z, a.w), tc::pack_bf16(b.x, b.y),
                           tc::pack_bf16(b.z, b.w));
        }
        *(uint4*)(xb + (int64_t)row * ldx + c) = o;
    }
}

// ---- weight packing (fp32 Parameters -> bf16 GEMM operands, zero padded) ---------------------------
// mode 0: out[j][k]          = W_{j / H}[j % H][k]          (3 stacked (H,Dm) matrices -> (3H, ldo));
//         with biases given, out[j][Dm] = b_{j / H}[j % H]: the product with the ones column of Xb
//         adds the bias inside the GEMM
// (mode 2 likewise puts the Conv1d bias at column Dm of the unshifted tap)
// mode 1: out[c][j]          = W_{j / H}[j % H][c]          (their transpose (Dm, ldo), ldo >= 3H)
// mode 2: out[o][t*ldt + c]  = w[o][c][t]                   (Conv1d weight (H,Dm,3) -> (H, 3*ldt))
// mode 3: out[c][u*ldt + o]  = w[o][c][u]                   (conv dX operand (Dm, 3*ldt), ldt >= H)
struct PackArgs {
    const float* w0; const float* w1; const float* w2;
    int mode, H, Dm, ldt, rows, ldo, nmat;
    const float* b0; const float* b1; const float* b2;     // optional biases (modes 0 and 2)
    int hm_dk;                                             // modes 0/1 with nmat = 3: head width when the 3H axis is HEAD-MAJOR
};
// index on a head-major 3H axis ([head][q|k|v][d]) -> canonical index ([q|k|v][head][d]) = row of [W_Q; W_K; W_V]
__host__ __device__ __forceinline__ int qkv_canonical(int r, int dk, int H) {
    const int hd = r / (3 * dk), rem = r - hd * 3 * dk;
    const int m = rem / dk, d = rem - m * dk;
    return m * H + hd * dk + d;
}
// up to four packing jobs in ONE launch (blockIdx.y = job): the forward packs its two operands and the two
// transposed operands the backward GEMMs need, so a step pays one ~5 us launch per encoder instead of four
struct PackJobs { PackArgs a[4]; bf16* out[4]; };
static __device__ __forceinline__ void pack_body(const PackArgs& a, bf16* __restrict__ out);
static __global__ void pack_weights_multi_kernel(const PackJobs j) { pack_body(j.a[blockIdx.y], j.out[blockIdx.y]); }
static __device__ __forceinline__ void pack_body(const PackArgs& a, bf16* __restrict__ out) {
    const int64_t total = (int64_t)a.rows * a.ldo;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x) {
        int r = (int)(t / a.ldo), c = (int)(t - (int64_t)r * a.ldo);
        float v = 0.f;
        if (a.hm_dk) {       // the packed 3H axis is head-major: read the canonical source row / column
            if (a.mode == 0) r = qkv_canonical(r, a.hm_dk, a.H);
            else if (a.mode == 1 && c < a.nmat * a.H) c = qkv_canonical(c, a.hm_dk, a.H);
        }
        if (a.mode == 0) {
            const int w = r / a.H;
            if (c < a.Dm) {
                const float* p = w == 0 ? a.w0 : (w == 1 ? a.w1 : a.w2);
                v = __ldg(p + (int64_t)(r - w * a.H) * a.Dm + c);
            } else if (c == a.Dm && a.b0) {
                const float* p = w == 0 ? a.b0 : (w == 1 ? a.b1 : a.b2);
                v = __ldg(p + (r - w * a.H));
            }
        } else if (a.mode == 1) {
            if (c < a.nmat * a.H) {
                const int w = c / a.H;
                const float* p = w == 0 ? a.w0 : (w == 1 ? a.w1 : a.w2);
                v = __ldg(p + (int64_t)(c - w * a.H) * a.Dm + r);
            }
        } else if (a.mode == 2) {
            const int tp = c / a.ldt, cc = c - tp * a.ldt;
            if (cc < a.Dm) v = __ldg(a.w0 + ((int64_t)r * a.Dm + cc) * 3 + tp);
            else if (cc == a.Dm && tp == 1 && a.b0) v = __ldg(a.b0 + r);
        } else {
            const int u = c / a.ldt, o = c - u * a.ldt;
            if (o < a.H) v = __ldg(a.w0 + ((int64_t)o * a.Dm + r) * 3 + u);
        }
        out[t] = __float2bfloat16(v);
    }
}

// ---- per-head exp-attention over bf16 qkv (model_utils.py:39-55; F8: no max subtraction) -------------
// qkv (n*SP, 3H): q at col h*DK, k at H + h*DK, v at 2H + h*DK.  One warp per (sequence, head), four
// heads per block; K and V of the head in shared memory (fp32); lane i owns query rows i, i+32, ...
// ctx (n*SP, H) = heads concatenated, output dropout (stream NR_STREAM_OUT, element index
// (seq*S+i)*H + col) applied; pad rows written as zero.
template <int DK>
static __global__ void __launch_bounds__(128)
attn_fwd_bf16_kernel(const bf16* __restrict__ qkv, const float* __restrict__ key_mask, bf16* __restrict__ ctx,
                     int S, int SP, int heads, float scale, uint32_t thr, float dscale, uint64_t seed,
                     uint8_t* __restrict__ keep) {
    extern __shared__ __align__(16) float smem[];
    const int H = heads * DK;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int groups = (heads + 3) / 4;
    const int64_t seq = blockIdx.x / groups;
    const int head = (blockIdx.x % groups) * 4 + warp;
    if (head >= heads) return;
    float* Ks = smem + warp * (2 * S * DK);
    float* Vs = Ks + S * DK;
    const bf16* base = qkv + seq * SP * 3 * H + head * DK;
    for (int idx = lane; idx < S * (DK / 4); idx += 32) {
        const int s = idx / (DK / 4), q4 = idx % (DK / 4);
        const bf16* row = base + (int64_t)s * 3 * H + q4 * 4;
        *(float4*)(Ks + s * DK + q4 * 4) = ld_bf4(row + H);
        *(float4*)(Vs + s * DK + q4 * 4) = ld_bf4(row + 2 * H);
    }
    __syncwarp();
    for (int i = lane; i < SP; i += 32) {
        bf16* out = ctx + (seq * SP + i) * H + head * DK;
        if (i >= S) {
#pragma unroll
            for (int d = 0; d < DK; d += 4) *(uint2*)(out + d) = make_uint2(0u, 0u);
            continue;
        }
        float q[DK], o[DK];
        const bf16* qr = base + (int64_t)i * 3 * H;
#pragma unroll
        for (int d = 0; d < DK; d += 4) {
            const float4 t = ld_bf4(qr + d);
            q[d] = t.x * scale; q[d + 1] = t.y * scale; q[d + 2] = t.z * scale; q[d + 3] = t.w * scale;
        }
#pragma unroll
        for (int d = 0; d < DK; ++d) o[d] = 0.f;
        float z = 0.f;
        for (int j = 0; j < S; ++j) {
            float dot = 0.f;
#pragma unroll
            for (int d = 0; d < DK; ++d) dot = fmaf(q[d], Ks[j * DK + d], dot);
            float p = __expf(dot);
            if (key_mask) p *= __ldg(key_mask + seq * S + j);
            z += p;
#pragma unroll
            for (int d = 0; d < DK; ++d) o[d] = fmaf(p, Vs[j * DK + d], o[d]);
        }
        const float inv = 1.f / (z + kEps);
#pragma unroll
        for (int d = 0; d < DK; d += 4) {
            float4 t = make_float4(o[d] * inv, o[d + 1] * inv, o[d + 2] * inv, o[d + 3] * inv);
            if (thr) {
                const uint32_t bits = dropout4_bits(out_stream_index(seq * S + i, head, d, heads, DK), seed, NR_STREAM_OUT, thr);
                t = apply_keep4(t, bits, dscale);
                if (keep) keep[((seq * SP + i) * H + head * DK + d) >> 2] = (uint8_t)bits;
            }
            st_bf4(out + d, t);
        }
    }
}

// backward: dctx here is the gradient w.r.t. the PRE-dropout context (the output
// dropout mask is applied by the producer of dctx).  Pass 1 (lane = query row): dQ, Z_i, rd_i;
// pass 2 (lane = key row): dK, dV.  dqkv pad rows are written as zero.
template <int DK>
static __global__ void __launch_bounds__(128)
attn_bwd_bf16_kernel(const bf16* __restrict__ qkv, const float* __restrict__ key_mask,
                     const bf16* __restrict__ dctx, bf16* __restrict__ dqkv, int S, int SP, int heads,
                     float scale) {
    extern __shared__ __align__(16) float smem[];
    const int H = heads * DK;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int groups = (heads + 3) / 4;
    const int64_t seq = blockIdx.x / groups;
    const int head = (blockIdx.x % groups) * 4 + warp;
    if (head >= heads) return;
    const int per_warp = 4 * S * DK + 2 * S + ((2 * S) & 3 ? 4 - ((2 * S) & 3) : 0);
    float* Qs = smem + warp * per_warp;
    float* Ks = Qs + S * DK;
    float* Vs = Ks + S * DK;
    float* Gs = Vs + S * DK;
    float* Zs = Gs + S * DK;
    float* Rs = Zs + S;
    const bf16* base = qkv + seq * SP * 3 * H + head * DK;
    const bf16* gbase = dctx + seq * SP * H + head * DK;
    for (int idx = lane; idx < S * (DK / 4); idx += 32) {
        const int s = idx / (DK / 4), q4 = idx % (DK / 4);
        const bf16* row = base + (int64_t)s * 3 * H + q4 * 4;
        *(float4*)(Qs + s * DK + q4 * 4) = ld_bf4(row);
        *(float4*)(Ks + s * DK + q4 * 4) = ld_bf4(row + H);
        *(float4*)(Vs + s * DK + q4 * 4) = ld_bf4(row + 2 * H);
        *(float4*)(Gs + s * DK + q4 * 4) = ld_bf4(gbase + (int64_t)s * H + q4 * 4);
    }
    __syncwarp();
    bf16* dbase = dqkv + seq * SP * 3 * H + head * DK;
    for (int i = lane; i < SP; i += 32) {
        bf16* dq = dbase + (int64_t)i * 3 * H;
        if (i >= S) {
#pragma unroll
            for (int d = 0; d < DK; d += 4) {
                *(uint2*)(dq + d) = make_uint2(0u, 0u);
                *(uint2*)(dq + H + d) = make_uint2(0u, 0u);
                *(uint2*)(dq + 2 * H + d) = make_uint2(0u, 0u);
            }
            continue;
        }
        float q[DK], go[DK], t1[DK], t2[DK];
#pragma unroll
        for (int d = 0; d < DK; ++d) { q[d] = Qs[i * DK + d]; go[d] = Gs[i * DK + d]; t1[d] = 0.f; t2[d] = 0.f; }
        float z = 0.f, rd = 0.f;
        for (int j = 0; j < S; ++j) {
            float dot = 0.f, dA = 0.f;
#pragma unroll
            for (int d = 0; d < DK; ++d) { dot = fmaf(q[d], Ks[j * DK + d], dot); dA = fmaf(go[d], Vs[j * DK + d], dA); }
            float p = __expf(dot * scale);
            if (key_mask) p *= __ldg(key_mask + seq * S + j);
            z += p;
            const float pd = p * dA;
            rd += pd;
#pragma unroll
            for (int d = 0; d < DK; ++d) { const float kv = Ks[j * DK + d]; t1[d] = fmaf(pd, kv, t1[d]); t2[d] = fmaf(p, kv, t2[d]); }
        }
        z += kEps;
        const float inv = 1.f / z;
        rd *= inv;
        Zs[i] = inv;
        Rs[i] = rd;
#pragma unroll
        for (int d = 0; d < DK; d += 4)
            st_bf4(dq + d, make_float4((t1[d] - rd * t2[d]) * inv * scale, (t1[d + 1] - rd * t2[d + 1]) * inv * scale,
                                       (t1[d + 2] - rd * t2[d + 2]) * inv * scale, (t1[d + 3] - rd * t2[d + 3]) * inv * scale));
    }
    __syncwarp();
    for (int j = lane; j < S; j += 32) {
        float k[DK], v[DK], dk[DK], dv[DK];
#pragma unroll
        for (int d = 0; d < DK; ++d) { k[d] = Ks[j * DK + d]; v[d] = Vs[j * DK + d]; dk[d] = 0.f; dv[d] = 0.f; }
        const float mj = key_mask ? __ldg(key_mask + seq * S + j) : 1.f;
        for (int i = 0; i < S; ++i) {
            float dot = 0.f, dA = 0.f;
#pragma unroll
            for (int d = 0; d < DK; ++d) { dot = fmaf(Qs[i * DK + d], k[d], dot); dA = fmaf(Gs[i * DK + d], v[d], dA); }
            const float a = __expf(dot * scale) * mj * Zs[i];
            const float dS = a * (dA - Rs[i]);
#pragma unroll
            for (int d = 0; d < DK; ++d) { dk[d] = fmaf(dS, Qs[i * DK + d], dk[d]); dv[d] = fmaf(a, Gs[i * DK + d], dv[d]); }
        }
        bf16* dkr = dbase + (int64_t)j * 3 * H + H;
        bf16* dvr = dkr + H;
#pragma unroll
        for (int d = 0; d < DK; d += 4) {
            st_bf4(dkr + d, make_float4(dk[d] * scale, dk[d + 1] * scale, dk[d + 2] * scale, dk[d + 3] * scale));
            st_bf4(dvr + d, make_float4(dv[d], dv[d + 1], dv[d + 2], dv[d + 3]));
        }
    }
}

inline size_t attn_bf16_fwd_smem(int S, int DK) { return (size_t)4 * 2 * S * DK * sizeof(float); }
inline size_t attn_bf16_bwd_smem(int S, int DK) {
    const int per_warp = 4 * S * DK + 2 * S + ((2 * S) & 3 ? 4 - ((2 * S) & 3) : 0);
    return (size_t)4 * per_warp * sizeof(float);
}

// ---- AttentionPooling tails over bf16 (model_utils.py:23-30) -----------------------------------------
// logits (n*SP) fp32 come from the fc1 GEMM epilogue (tanh, dot w2, + b2).  p = exp(a)*m;
// alpha = p/(sum p + eps); out = sum_s alpha_s y_s.  One block (128 or 256 threads) per sequence.
static __global__ void __launch_bounds__(256)
pool_tail_fwd_bf16_kernel(const bf16* __restrict__ y, const float* __restrict__ logits,
                          const float* __restrict__ mask, float* __restrict__ out,
                          float* __restrict__ alpha_save, int S, int SP, int H) {
    extern __shared__ float sp[];   // SP
    const int64_t seq = blockIdx.x;
    for (int s = threadIdx.x; s < SP; s += blockDim.x) {
        float p = 0.f;
        if (s < S) {
            p = expf(__ldg(logits + seq * SP + s));
            if (mask) p *= __ldg(mask + seq * S + s);
        }
        sp[s] = p;
    }
    __syncthreads();
    float z = 0.f;
    for (int s = 0; s < S; ++s) z += sp[s];
    const float inv = 1.f / (z + kEps);          // model_utils.py:29
    __syncthreads();
    for (int s = threadIdx.x; s < SP; s += blockDim.x) {
        const float a = sp[s] * inv;
        sp[s] = a;
        alpha_save[seq * SP + s] = a;
    }
    __syncthreads();
    // Few, long sequences (the user encoder: 128 x 50 rows): the block is alone on its SM, so what matters is how
    // many loads are in flight.  Warp w takes rows w, w+NW, ... (each lane 16-byte chunks c = lane, lane+32, ...),
    // all of a warp's row loads are independent, and the NW partial sums are combined through shared memory.
    if ((H & 7) == 0 && H <= 512) {
        float* part = sp + SP;                      // [NW][H] partial sums (dynamic smem sized by the launcher)
        const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, NW = blockDim.x >> 5;
        const int HC = H >> 3;
        float acc[2][8];
#pragma unroll
        for (int k = 0; k < 2; ++k)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[k][j] = 0.f;
        const bf16* yt = y + seq * SP * H;
#pragma unroll 4
        for (int s = w; s < S; s += NW) {
            const float a = sp[s];
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int c = lane + 32 * k;
                if (c < HC) {
                    float f[8];
                    const uint4 v = __ldg((const uint4*)(yt + (int64_t)s * H) + c);
                    const float2 x0 = bf2_to_f2(v.x), x1 = bf2_to_f2(v.y), x2 = bf2_to_f2(v.z), x3 = bf2_to_f2(v.w);
                    f[0] = x0.x; f[1] = x0.y; f[2] = x1.x; f[3] = x1.y; f[4] = x2.x; f[5] = x2.y; f[6] = x3.x; f[7] = x3.y;
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[k][j] = fmaf(a, f[j], acc[k][j]);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int c = lane + 32 * k;
            if (c < HC) {
#pragma unroll
                for (int j = 0; j < 8; ++j) part[w * H + c * 8 + j] = acc[k][j];
            }
        }
        __syncthreads();
        for (int h = threadIdx.x; h < H; h += blockDim.x) {
            float o = 0.f;
            for (int k = 0; k < NW; k += 2) o += part[k * H + h] + part[(k + 1) * H + h];     // NW is even; fixed order
            out[seq * H + h] = o;
        }
        return;
    }
    for (int h = threadIdx.x * 4; h < H; h += blockDim.x * 4) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        const bf16* yr = y + seq * SP * H + h;
        for (int s = 0; s < S; ++s) {
            const float4 v = ld_bf4(yr + (int64_t)s * H);
            const float a = sp[s];
            acc.x = fmaf(a, v.x, acc.x); acc.y = fmaf(a, v.y, acc.y); acc.z = fmaf(a, v.z, acc.z); acc.w = fmaf(a, v.w, acc.w);
        }
        *(float4*)(out + seq * H + h) = acc;
    }
}

// backward tail: dalpha_s = g.y_s ; da_s = alpha_s (dalpha_s - sum_t alpha_t dalpha_t) ;
// dz[s,q] = da_s w2[q] (1 - e[s,q]^2) (bf16, pad rows zero) ; dw2 += sum da_s e_s ; db2 += sum da_s ;
// db1 += sum_s dz[s,:].
// 256 threads per block, blocks stride over sequences.  Phase 1: warp w takes rows w, w+8, ... and reads
// them with 16-byte loads (coalesced) against g held in shared memory.  Phase 2: thread t owns the
// 4-column group q4 = t % (Q/4) and the rows rg, rg+RG, ... (RG = 256 / (Q/4) row groups), so the E read
// and the dz write are coalesced and spread over all threads; the column sums for dw2 / db1 stay in
// registers across sequences and are flushed once per block.
static __global__ void __launch_bounds__(256)
pool_tail_bwd_bf16_kernel(const float* __restrict__ g, const bf16* __restrict__ y, const bf16* __restrict__ e,
                          const float* __restrict__ alpha, const float* __restrict__ w2, bf16* __restrict__ dz,
                          float* __restrict__ dw2, float* __restrict__ db2, float* __restrict__ db1, int64_t n,
                          int S, int SP, int H, int Q) {
    extern __shared__ __align__(16) float sm[];   // gs[H] | da[SP] | red[2*Q]
    float* gs = sm;
    float* da = gs + H;
    float* red = da + SP;
    __shared__ float s_c;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int QC = Q >> 2;                       // 4-column groups
    const int RG = blockDim.x / QC;              // row groups (>= 1: Q <= 1024)
    const int q4 = (threadIdx.x % QC) * 4, rg = threadIdx.x / QC;
    const bool worker = rg < RG;
    float4 wv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (worker) wv = ldg4(w2 + q4);
    float4 a2 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a2;
    float db2_acc = 0.f;
    for (int q = threadIdx.x; q < 2 * Q; q += blockDim.x) red[q] = 0.f;
    for (int64_t seq = blockIdx.x; seq < n; seq += gridDim.x) {
        __syncthreads();
        for (int h = threadIdx.x * 4; h < H; h += blockDim.x * 4) *(float4*)(gs + h) = ldg4(g + seq * H + h);
        __syncthreads();
#pragma unroll 4      // the rows of a warp are independent: keep several rows of loads in flight (few, long sequences)
        for (int s = warp; s < S; s += nw) {
            const bf16* yr = y + (seq * SP + s) * H;
            float d = 0.f;
#pragma unroll 2
            for (int h = lane * 8; h < H; h += 256) {
                const uint4 w = __ldg((const uint4*)(yr + h));
                const float2 y0 = bf2_to_f2(w.x), y1 = bf2_to_f2(w.y), y2 = bf2_to_f2(w.z), y3 = bf2_to_f2(w.w);
                const float4 g0 = *(const float4*)(gs + h), g1 = *(const float4*)(gs + h + 4);
                d = fmaf(g0.x, y0.x, d); d = fmaf(g0.y, y0.y, d); d = fmaf(g0.z, y1.x, d); d = fmaf(g0.w, y1.y, d);
                d = fmaf(g1.x, y2.x, d); d = fmaf(g1.y, y2.y, d); d = fmaf(g1.z, y3.x, d); d = fmaf(g1.w, y3.y, d);
            }
            d = warp_sum(d);
            if (lane == 0) da[s] = d;
        }
        __syncthreads();
        if (warp == 0) {
            float c = 0.f;
            for (int s = lane; s < S; s += 32) c = fmaf(__ldg(alpha + seq * SP + s), da[s], c);
            c = warp_sum(c);
            if (lane == 0) s_c = c;
        }
        __syncthreads();
        const float c = s_c;
        __syncthreads();
        for (int s = threadIdx.x; s < S; s += blockDim.x) da[s] = __ldg(alpha + seq * SP + s) * (da[s] - c);
        __syncthreads();
        if (threadIdx.x == 0) {
            float t = 0.f;
            for (int s = 0; s < S; ++s) t += da[s];
            db2_acc += t;
        }
        if (worker) {
#pragma unroll 4
            for (int s = rg; s < SP; s += RG) {
                bf16* zr = dz + (seq * SP + s) * Q + q4;
                if (s >= S) { *(uint2*)zr = make_uint2(0u, 0u); continue; }
                const float4 ev = ld_bf4(e + (seq * SP + s) * Q + q4);
                const float d = da[s];
                const float4 z = make_float4(d * wv.x * (1.f - ev.x * ev.x), d * wv.y * (1.f - ev.y * ev.y),
                                             d * wv.z * (1.f - ev.z * ev.z), d * wv.w * (1.f - ev.w * ev.w));
                st_bf4(zr, z);
                a2.x = fmaf(d, ev.x, a2.x); a2.y = fmaf(d, ev.y, a2.y); a2.z = fmaf(d, ev.z, a2.z); a2.w = fmaf(d, ev.w, a2.w);
                a1.x += z.x; a1.y += z.y; a1.z += z.z; a1.w += z.w;
            }
        }
    }
    __syncthreads();
    if (worker) {
        atomicAdd(red + q4, a2.x); atomicAdd(red + q4 + 1, a2.y); atomicAdd(red + q4 + 2, a2.z); atomicAdd(red + q4 + 3, a2.w);
        atomicAdd(red + Q + q4, a1.x); atomicAdd(red + Q + q4 + 1, a1.y); atomicAdd(red + Q + q4 + 2, a1.z);
        atomicAdd(red + Q + q4 + 3, a1.w);
    }
    __syncthreads();
    for (int q = threadIdx.x; q < Q; q += blockDim.x) {
        atomicAdd(dw2 + q, red[q]);
        atomicAdd(db1 + q, red[Q + q]);
    }
    if (threadIdx.x == 0) atomicAdd(db2, db2_acc);
}


// ---- warp-per-sequence pooling tails (SP <= 64) --------------------------------------------------------------
// The block-per-sequence kernels above spend their time in __syncthreads (7 per sequence in backward) with
// half the warps idle; a title is only 30 x 400 bf16 = 24 KB.  Here ONE warp owns a sequence: lane l holds the
// per-row scalars of rows l and l+32 (logit, alpha, d alpha) in registers, walks the rows with 16-byte loads
// (chunk c = l + 32k of a row, KH = ceil(H/256) chunks per lane) and broadcasts the row scalar with a shuffle;
// no shared memory, no block barrier, several rows of loads in flight per lane.
__device__ __forceinline__ void bf8_to_f(const uint4 w, float* f) {
    const float2 a = bf2_to_f2(w.x), b = bf2_to_f2(w.y), c = bf2_to_f2(w.z), d = bf2_to_f2(w.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

template <int KH>
static __global__ void __launch_bounds__(256)
pool_tail_fwd_warp_kernel(const bf16* __restrict__ y, const float* __restrict__ logits, const float* __restrict__ mask,
                          float* __restrict__ out, float* __restrict__ alpha_save, int64_t n, int S, int SP, int H) {
    const int lane = threadIdx.x & 31;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int HC = H >> 3;
    for (int64_t seq = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; seq < n; seq += nw) {
        float p0 = 0.f, p1 = 0.f;
        if (lane < S) {
            p0 = expf(__ldg(logits + seq * SP + lane));
            if (mask) p0 *= __ldg(mask + seq * S + lane);
        }
        if (lane + 32 < S) {
            p1 = expf(__ldg(logits + seq * SP + lane + 32));
            if (mask) p1 *= __ldg(mask + seq * S + lane + 32);
        }
        const float inv = 1.f / (warp_sum(p0 + p1) + kEps);          // model_utils.py:29
        const float a0 = p0 * inv, a1 = p1 * inv;
        if (lane < SP) alpha_save[seq * SP + lane] = a0;
        if (lane + 32 < SP) alpha_save[seq * SP + lane + 32] = a1;
        float acc[KH][8];
#pragma unroll
        for (int k = 0; k < KH; ++k)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[k][j] = 0.f;
        const bf16* yt = y + seq * SP * H;
#pragma unroll 8
        for (int s = 0; s < S; ++s) {
            const float a = __shfl_sync(0xffffffffu, s < 32 ? a0 : a1, s & 31);
#pragma unroll
            for (int k = 0; k < KH; ++k) {
                const int c = lane + 32 * k;
                if (c < HC) {
                    float f[8];
                    bf8_to_f(__ldg((const uint4*)(yt + (int64_t)s * H) + c), f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[k][j] = fmaf(a, f[j], acc[k][j]);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < KH; ++k) {
            const int c = lane + 32 * k;
            if (c < HC) {
                float4* o = (float4*)(out + seq * H + c * 8);
                o[0] = make_float4(acc[k][0], acc[k][1], acc[k][2], acc[k][3]);
                o[1] = make_float4(acc[k][4], acc[k][5], acc[k][6], acc[k][7]);
            }
        }
    }
}

template <int KH, int KQ>
static __global__ void __launch_bounds__(256)
pool_tail_bwd_warp_kernel(const float* __restrict__ g, const bf16* __restrict__ y, const bf16* __restrict__ e,
                          const float* __restrict__ alpha, const float* __restrict__ w2, bf16* __restrict__ dz,
                          float* __restrict__ dw2, float* __restrict__ db2, float* __restrict__ db1, int64_t n,
                          int S, int SP, int H, int Q) {
    extern __shared__ __align__(16) float red[];     // [2Q + 1]: dw2 | db1 | db2 partials of this block
    const int lane = threadIdx.x & 31;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int HC = H >> 3, QC = Q >> 3;
    for (int q = threadIdx.x; q < 2 * Q + 1; q += blockDim.x) red[q] = 0.f;
    __syncthreads();
    float wv[KQ][8], a2[KQ][8], a1[KQ][8];
#pragma unroll
    for (int k = 0; k < KQ; ++k) {
        const int c = lane + 32 * k;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            wv[k][j] = c < QC ? __ldg(w2 + c * 8 + j) : 0.f;
            a2[k][j] = 0.f; a1[k][j] = 0.f;
        }
    }
    float db2_acc = 0.f;
    for (int64_t seq = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; seq < n; seq += nw) {
        float gv[KH][8];
#pragma unroll
        for (int k = 0; k < KH; ++k) {
            const int c = lane + 32 * k;
            if (c < HC) {
                const float4 x0 = ldg4(g + seq * H + c * 8), x1 = ldg4(g + seq * H + c * 8 + 4);
                gv[k][0] = x0.x; gv[k][1] = x0.y; gv[k][2] = x0.z; gv[k][3] = x0.w;
                gv[k][4] = x1.x; gv[k][5] = x1.y; gv[k][6] = x1.z; gv[k][7] = x1.w;
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) gv[k][j] = 0.f;
            }
        }
        // d alpha_s = g . y_s, kept by lane s % 32
        const bf16* yt = y + seq * SP * H;
        float d0 = 0.f, d1 = 0.f;
#pragma unroll 8
        for (int s = 0; s < S; ++s) {
            float part = 0.f;
#pragma unroll
            for (int k = 0; k < KH; ++k) {
                const int c = lane + 32 * k;
                if (c < HC) {
                    float f[8];
                    bf8_to_f(__ldg((const uint4*)(yt + (int64_t)s * H) + c), f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) part = fmaf(gv[k][j], f[j], part);
                }
            }
            part = warp_sum(part);
            if ((s & 31) == lane) { if (s < 32) d0 = part; else d1 = part; }
        }
        const float al0 = lane < S ? __ldg(alpha + seq * SP + lane) : 0.f;
        const float al1 = lane + 32 < S ? __ldg(alpha + seq * SP + lane + 32) : 0.f;
        const float c = warp_sum(fmaf(al0, d0, al1 * d1));
        const float da0 = al0 * (d0 - c), da1 = al1 * (d1 - c);
        db2_acc += da0 + da1;
        // dz rows (pad rows written as zero), column sums for dw2 / db1
        const bf16* et = e + seq * SP * Q;
        bf16* zt = dz + seq * SP * Q;
#pragma unroll 8
        for (int s = 0; s < SP; ++s) {
            const float d = __shfl_sync(0xffffffffu, s < 32 ? da0 : da1, s & 31);
#pragma unroll
            for (int k = 0; k < KQ; ++k) {
                const int cq = lane + 32 * k;
                if (cq < QC) {
                    uint4 o = make_uint4(0u, 0u, 0u, 0u);
                    if (s < S) {
                        float f[8], z[8];
                        bf8_to_f(__ldg((const uint4*)(et + (int64_t)s * Q) + cq), f);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            z[j] = d * wv[k][j] * (1.f - f[j] * f[j]);
                            a2[k][j] = fmaf(d, f[j], a2[k][j]);
                            a1[k][j] += z[j];
                        }
                        o = make_uint4(tc::pack_bf16(z[0], z[1]), tc::pack_bf16(z[2], z[3]), tc::pack_bf16(z[4], z[5]),
                                       tc::pack_bf16(z[6], z[7]));
                    }
                    *((uint4*)(zt + (int64_t)s * Q) + cq) = o;
                }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < KQ; ++k) {
        const int cq = lane + 32 * k;
        if (cq < QC) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                atomicAdd(red + cq * 8 + j, a2[k][j]);
                atomicAdd(red + Q + cq * 8 + j, a1[k][j]);
            }
        }
    }
    db2_acc = warp_sum(db2_acc);
    if (lane == 0) atomicAdd(red + 2 * Q, db2_acc);
    __syncthreads();
    for (int q = threadIdx.x; q < Q; q += blockDim.x) {
        atomicAdd(dw2 + q, red[q]);
        atomicAdd(db1 + q, red[Q + q]);
    }
    if (threadIdx.x == 0) atomicAdd(db2, red[2 * Q]);
}

}  // namespace nr
