// nr_tails.cuh — the per-sequence "tail" kernels around the GEMMs: additive-attention
// softmax + weighted sum, per-head exp-attention, column sums.  All fp32, exact reference
// formulas (no max subtraction, eps inside the denominator).
#pragma once
#include "nr_common.cuh"

namespace nr {

// ---- AttentionPooling tail (model_utils.py:23-30) ----------------------------------------
// e (n,S,Q) = tanh(fc1(x)) from the GEMM; logits a_s = e_s.w2 + b2; p = exp(a)*m;
// alpha = p/(sum p + eps); out = sum_s alpha_s x_s.   One block per sequence.
static __global__ void __launch_bounds__(128)
pool_tail_fwd_kernel(const float* __restrict__ x, const float* __restrict__ e,
                     const float* __restrict__ mask, const float* __restrict__ w2,
                     const float* __restrict__ b2, float* __restrict__ out,
                     float* __restrict__ alpha_save, int S, int H, int Q) {
    extern __shared__ float sp[];   // S
    const int64_t seq = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const float bias = __ldg(b2);
    for (int s = warp; s < S; s += nw) {
        const float* er = e + (seq * S + s) * Q;
        float d = 0.f;
        for (int q = lane; q < Q; q += 32) d = fmaf(__ldg(er + q), __ldg(w2 + q), d);
        d = warp_sum(d);
        if (lane == 0) {
            float p = expf(d + bias);
            if (mask) p *= __ldg(mask + seq * S + s);
            sp[s] = p;
        }
    }
    __syncthreads();
    float z = 0.f;
    for (int s = 0; s < S; ++s) z += sp[s];    // same order in every thread -> identical Z
    __syncthreads();
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
        const float a = sp[s] / (z + kEps);    // model_utils.py:29
        sp[s] = a;
        alpha_save[seq * S + s] = a;
    }
    __syncthreads();
    for (int h = threadIdx.x; h < H; h += blockDim.x) {
        float acc = 0.f;
        const float* xr = x + seq * S * H + h;
        for (int s = 0; s < S; ++s) acc = fmaf(sp[s], __ldg(xr + (int64_t)s * H), acc);
        out[seq * H + h] = acc;
    }
}

// backward tail: given g (n,H):  dalpha_s = g.x_s ; da_s = alpha_s (dalpha_s - sum_t alpha_t dalpha_t)
// dz[s,q] = da_s w2[q] (1 - e[s,q]^2) ; dx[s,:] = alpha_s g ; dw2 += sum da_s e_s ; db2 += sum da_s.
// Blocks stride over sequences so dw2/db2 need one atomic per block and column.
static __global__ void __launch_bounds__(256)
pool_tail_bwd_kernel(const float* __restrict__ g, const float* __restrict__ x,
                     const float* __restrict__ e, const float* __restrict__ alpha,
                     const float* __restrict__ w2, float* __restrict__ dz, float* __restrict__ dx,
                     float* __restrict__ dw2, float* __restrict__ db2, int64_t n, int S, int H, int Q) {
    extern __shared__ float sm[];   // da[S] | dw2_acc[Q]
    float* da = sm;
    float* dw2_acc = sm + S;
    __shared__ float s_c;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int q = threadIdx.x; q < Q; q += blockDim.x) dw2_acc[q] = 0.f;
    float db2_acc = 0.f;
    for (int64_t seq = blockIdx.x; seq < n; seq += gridDim.x) {
        __syncthreads();
        const float* gr = g + seq * H;
        for (int s = warp; s < S; s += nw) {
            const float* xr = x + (seq * S + s) * H;
            float d = 0.f;
            for (int h = lane; h < H; h += 32) d = fmaf(__ldg(gr + h), __ldg(xr + h), d);
            d = warp_sum(d);
            if (lane == 0) da[s] = d;
        }
        __syncthreads();
        if (warp == 0) {
            float c = 0.f;
            for (int s = lane; s < S; s += 32) c = fmaf(__ldg(alpha + seq * S + s), da[s], c);
            c = warp_sum(c);
            if (lane == 0) s_c = c;
        }
        __syncthreads();
        const float c = s_c;
        __syncthreads();
        for (int s = threadIdx.x; s < S; s += blockDim.x) {
            const float a = __ldg(alpha + seq * S + s);
            da[s] = a * (da[s] - c);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            float t = 0.f;
            for (int s = 0; s < S; ++s) t += da[s];
            db2_acc += t;
        }
        for (int q = threadIdx.x; q < Q; q += blockDim.x) {
            const float w = __ldg(w2 + q);
            float acc = 0.f;
            for (int s = 0; s < S; ++s) {
                const float ev = __ldg(e + (seq * S + s) * Q + q);
                dz[(seq * S + s) * Q + q] = da[s] * w * (1.f - ev * ev);
                acc = fmaf(da[s], ev, acc);
            }
            dw2_acc[q] += acc;
        }
        if (dx) {
            for (int h = threadIdx.x; h < H; h += blockDim.x) {
                const float gv = __ldg(gr + h);
                for (int s = 0; s < S; ++s)
                    dx[(seq * S + s) * H + h] = __ldg(alpha + seq * S + s) * gv;
            }
        }
    }
    __syncthreads();
    for (int q = threadIdx.x; q < Q; q += blockDim.x) atomicAdd(dw2 + q, dw2_acc[q]);
    if (threadIdx.x == 0) atomicAdd(db2, db2_acc);
}

// out[c] += sum_r Y[r*ld + c]; columns [0,each) [each,2each) [2each,3each) go to o0 o1 o2.
static __global__ void __launch_bounds__(128)
colsum_kernel(const float* __restrict__ Y, int64_t rows, int64_t ld, int cols, float* o0, float* o1,
              float* o2, int each) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cols) return;
    float acc = 0.f;
    for (int64_t r = blockIdx.y; r < rows; r += gridDim.y) acc += __ldg(Y + r * ld + c);
    const int w = c / each;
    float* o = (w == 0 ? o0 : (w == 1 ? o1 : o2)) + (c - w * each);
    atomicAdd(o, acc);
}

inline int colsum(const float* Y, int64_t rows, int64_t ld, int cols, float* o0, float* o1, float* o2,
                  int each, cudaStream_t st) {
    if (rows == 0 || cols == 0) return 0;
    const int gx = (cols + 127) / 128;
    int64_t gy = (4 * (int64_t)sm_count() + gx - 1) / gx;
    if (gy > rows) gy = rows;
    if (gy > 65535) gy = 65535;
    ProfScope ps("colsum", st);
    colsum_kernel<<<dim3(gx, (unsigned)gy), 128, 0, st>>>(Y, rows, ld, cols, o0, o1, o2, each);
    return check_last("colsum");
}

// ---- per-head exp-attention (model_utils.py:39-55) ------------------------------------------
// qkv (n,S,3H): q at col h*DK, k at H + h*DK, v at 2H + h*DK.  One warp per (sequence, head),
// four heads of one sequence per block.  K and V of the head live in shared memory; lane i
// owns query rows i, i+32.  S <= 64.
template <int DK>
static __global__ void __launch_bounds__(128)
attn_fwd_kernel(const float* __restrict__ qkv, const float* __restrict__ key_mask,
                float* __restrict__ ctx, int S, int heads, float scale, uint32_t thr, float dscale,
                uint64_t seed) {
    extern __shared__ __align__(16) float smem[];
    const int H = heads * DK;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int groups = (heads + 3) / 4;
    const int64_t seq = blockIdx.x / groups;
    const int head = (blockIdx.x % groups) * 4 + warp;
    if (head >= heads) return;
    float* Ks = smem + warp * (2 * S * DK);
    float* Vs = Ks + S * DK;
    const float* base = qkv + seq * S * 3 * H;
    for (int idx = lane; idx < S * (DK / 4); idx += 32) {
        const int s = idx / (DK / 4), q4 = idx % (DK / 4);
        const float* row = base + (int64_t)s * 3 * H + head * DK + q4 * 4;
        *(float4*)(Ks + s * DK + q4 * 4) = ldg4(row + H);
        *(float4*)(Vs + s * DK + q4 * 4) = ldg4(row + 2 * H);
    }
    __syncwarp();
    for (int i = lane; i < S; i += 32) {
        float q[DK], o[DK];
        const float* qr = base + (int64_t)i * 3 * H + head * DK;
#pragma unroll
        for (int d = 0; d < DK; d += 4) {
            const float4 t = ldg4(qr + d);
            q[d] = t.x; q[d + 1] = t.y; q[d + 2] = t.z; q[d + 3] = t.w;
        }
#pragma unroll
        for (int d = 0; d < DK; ++d) o[d] = 0.f;
        float z = 0.f;
        for (int j = 0; j < S; ++j) {
            float dot = 0.f;
#pragma unroll
            for (int d = 0; d < DK; ++d) dot = fmaf(q[d], Ks[j * DK + d], dot);
            float p = expf(dot * scale);
            if (key_mask) p *= __ldg(key_mask + seq * S + j);
            z += p;
#pragma unroll
            for (int d = 0; d < DK; ++d) o[d] = fmaf(p, Vs[j * DK + d], o[d]);
        }
        const float inv = 1.f / (z + kEps);
        float* out = ctx + (seq * S + i) * H + head * DK;
#pragma unroll
        for (int d = 0; d < DK; d += 4) {
            float4 t = make_float4(o[d] * inv, o[d + 1] * inv, o[d + 2] * inv, o[d + 3] * inv);
            if (thr) t = dropout4(t, out_stream_index(seq * S + i, head, d, heads, DK), seed, NR_STREAM_OUT, thr, dscale);
            *(float4*)(out + d) = t;
        }
    }
}

// backward: recompute P from q,k; see SURVEY Appendix C.  Pass 1 (lane = query row) yields dQ,
// Z_i and rd_i = sum_j A_ij dA_ij; pass 2 (lane = key row) yields dK, dV.
template <int DK>
static __global__ void __launch_bounds__(128)
attn_bwd_kernel(const float* __restrict__ qkv, const float* __restrict__ key_mask,
                const float* __restrict__ dctx, float* __restrict__ dqkv, int S, int heads, float scale,
                uint32_t thr, float dscale, uint64_t seed) {
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
    float* Gs = Vs + S * DK;     // dO (with output dropout undone)
    float* Zs = Gs + S * DK;
    float* Rs = Zs + S;
    const float* base = qkv + seq * S * 3 * H;
    for (int idx = lane; idx < S * (DK / 4); idx += 32) {
        const int s = idx / (DK / 4), q4 = idx % (DK / 4);
        const float* row = base + (int64_t)s * 3 * H + head * DK + q4 * 4;
        *(float4*)(Qs + s * DK + q4 * 4) = ldg4(row);
        *(float4*)(Ks + s * DK + q4 * 4) = ldg4(row + H);
        *(float4*)(Vs + s * DK + q4 * 4) = ldg4(row + 2 * H);
        const int64_t e0 = (seq * S + s) * H + head * DK + q4 * 4;
        float4 gv = ldg4(dctx + e0);
        if (thr) gv = dropout4(gv, out_stream_index(seq * S + s, head, q4 * 4, heads, DK), seed, NR_STREAM_OUT, thr, dscale);
        *(float4*)(Gs + s * DK + q4 * 4) = gv;
    }
    __syncwarp();
    float* dbase = dqkv + seq * S * 3 * H;
    for (int i = lane; i < S; i += 32) {
        float q[DK], go[DK], t1[DK], t2[DK];
#pragma unroll
        for (int d = 0; d < DK; ++d) { q[d] = Qs[i * DK + d]; go[d] = Gs[i * DK + d]; t1[d] = 0.f; t2[d] = 0.f; }
        float z = 0.f, rd = 0.f;
        for (int j = 0; j < S; ++j) {
            float dot = 0.f, dA = 0.f;
#pragma unroll
            for (int d = 0; d < DK; ++d) { dot = fmaf(q[d], Ks[j * DK + d], dot); dA = fmaf(go[d], Vs[j * DK + d], dA); }
            float p = expf(dot * scale);
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
        Zs[i] = z;
        Rs[i] = rd;
        float* dq = dbase + (int64_t)i * 3 * H + head * DK;
#pragma unroll
        for (int d = 0; d < DK; d += 4) {
            *(float4*)(dq + d) = make_float4((t1[d] - rd * t2[d]) * inv * scale, (t1[d + 1] - rd * t2[d + 1]) * inv * scale,
                                             (t1[d + 2] - rd * t2[d + 2]) * inv * scale, (t1[d + 3] - rd * t2[d + 3]) * inv * scale);
        }
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
            const float a = expf(dot * scale) * mj / Zs[i];
            const float dS = a * (dA - Rs[i]);
#pragma unroll
            for (int d = 0; d < DK; ++d) { dk[d] = fmaf(dS, Qs[i * DK + d], dk[d]); dv[d] = fmaf(a, Gs[i * DK + d], dv[d]); }
        }
        float* dkr = dbase + (int64_t)j * 3 * H + H + head * DK;
        float* dvr = dkr + H;
#pragma unroll
        for (int d = 0; d < DK; d += 4) {
            *(float4*)(dkr + d) = make_float4(dk[d] * scale, dk[d + 1] * scale, dk[d + 2] * scale, dk[d + 3] * scale);
            *(float4*)(dvr + d) = make_float4(dv[d], dv[d + 1], dv[d + 2], dv[d + 3]);
        }
    }
}

inline size_t attn_fwd_smem(int S, int DK) { return (size_t)4 * 2 * S * DK * sizeof(float); }
inline size_t attn_bwd_smem(int S, int DK) {
    const int per_warp = 4 * S * DK + 2 * S + ((2 * S) & 3 ? 4 - ((2 * S) & 3) : 0);
    return (size_t)4 * per_warp * sizeof(float);
}

}  // namespace nr
