// nr_eval.cu — eval path on device: CSR impression scoring from the cached news-vector
// table (src/dataset.py:64-74, src/main.py:253) and per-impression ranking metrics
// (src/metrics.py, sklearn roc_auc_score; src/main.py:250-258).
#include "nr_common.cuh"

namespace nr {

// one warp per candidate: scores[j] = table[cand_idx[j]] . user_vecs[imp(j)]
__global__ void __launch_bounds__(256)
eval_score_kernel(const float* __restrict__ table, const float* __restrict__ user,
                  const int64_t* __restrict__ offsets, const int32_t* __restrict__ cand,
                  float* __restrict__ scores, int64_t I, int64_t total, int H, int64_t table_rows) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t j = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; j < total; j += warps) {
        // impression i with offsets[i] <= j < offsets[i+1]
        int64_t lo = 0, hi = I;
        while (hi - lo > 1) {
            const int64_t mid = (lo + hi) >> 1;
            if (__ldg(offsets + mid) <= j) lo = mid; else hi = mid;
        }
        int64_t cid = __ldg(cand + j);
        if ((uint64_t)cid >= (uint64_t)table_rows) cid = 0;          // unknown ids score the padding news (dataset.py:14-15)
        const float4* c = (const float4*)(table + cid * H);
        const float4* u = (const float4*)(user + lo * H);
        float s = 0.f;
        for (int q = lane; q < H / 4; q += 32) {
            const float4 a = __ldg(c + q), b = __ldg(u + q);
            s = fmaf(a.x, b.x, s); s = fmaf(a.y, b.y, s); s = fmaf(a.z, b.z, s); s = fmaf(a.w, b.w, s);
        }
        s = warp_sum(s);
        if (lane == 0) scores[j] = s;
    }
}

// One WARP per impression: the user vector stays in registers (KV float4 per lane, H <= 128 KV), the impression's
// candidates are walked linearly from the CSR offsets (no per-candidate search), four candidate rows in flight per
// lane; the four dot products are reduced together and lanes 0..3 store one score each.
template <int KV>
__global__ void __launch_bounds__(256)
eval_score_warp_kernel(const float* __restrict__ table, const float* __restrict__ user,
                       const int64_t* __restrict__ offsets, const int32_t* __restrict__ cand,
                       float* __restrict__ scores, int64_t I, int H, int64_t table_rows) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int HV = H >> 2;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < I; i += warps) {
        float4 u[KV];
        const float4* up = (const float4*)(user + i * H);
#pragma unroll
        for (int k = 0; k < KV; ++k) u[k] = lane + 32 * k < HV ? __ldg(up + lane + 32 * k) : make_float4(0.f, 0.f, 0.f, 0.f);
        const int64_t lo = __ldg(offsets + i), hi = __ldg(offsets + i + 1);
        for (int64_t j0 = lo; j0 < hi; j0 += 4) {
            float s[4];
            int32_t id[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                id[c] = j0 + c < hi ? __ldg(cand + j0 + c) : 0;
                if ((uint64_t)(int64_t)id[c] >= (uint64_t)table_rows) id[c] = 0;   // unknown ids score the padding news (dataset.py:14-15)
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float4* cp = (const float4*)(table + (int64_t)id[c] * H);
                float a = 0.f;
#pragma unroll
                for (int k = 0; k < KV; ++k) {
                    if (lane + 32 * k < HV) {
                        const float4 v = __ldg(cp + lane + 32 * k);
                        a = fmaf(v.x, u[k].x, a); a = fmaf(v.y, u[k].y, a); a = fmaf(v.z, u[k].z, a); a = fmaf(v.w, u[k].w, a);
                    }
                }
                s[c] = a;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                for (int c = 0; c < 4; ++c) s[c] += __shfl_xor_sync(0xffffffffu, s[c], o);
            }
            const float mine = lane == 0 ? s[0] : (lane == 1 ? s[1] : (lane == 2 ? s[2] : s[3]));
            if (lane < 4 && j0 + lane < hi) scores[j0 + lane] = mine;
        }
    }
}

// One block per impression.  rank_i = 1 + #{j : s_j > s_i or (s_j == s_i and j > i)}
// (descending order, ties higher-index-first = stable argsort reversed).
// AUC = sum over positives of (#neg with smaller score + 0.5 #neg with equal score) / (n+ n-).
__global__ void __launch_bounds__(128)
rank_metrics_kernel(const float* __restrict__ scores, const int8_t* __restrict__ labels,
                    const int64_t* __restrict__ offsets, double* __restrict__ metrics,
                    uint8_t* __restrict__ valid, int64_t I) {
    __shared__ double red[4][4];
    __shared__ int s_npos;
    for (int64_t imp = blockIdx.x; imp < I; imp += gridDim.x) {
        const int64_t lo = offsets[imp];
        const int C = (int)(offsets[imp + 1] - lo);
        const float* s = scores + lo;
        const int8_t* y = labels + lo;
        int npos = 0;
        for (int i = threadIdx.x; i < C; i += blockDim.x) npos += y[i] != 0;
        npos = __reduce_add_sync(0xffffffffu, npos);
        __syncthreads();
        if (threadIdx.x == 0) s_npos = 0;
        __syncthreads();
        if ((threadIdx.x & 31) == 0) atomicAdd(&s_npos, npos);
        __syncthreads();
        npos = s_npos;
        if (npos == 0 || npos == C) {          // src/main.py:250-251
            if (threadIdx.x == 0) {
                valid[imp] = 0;
                for (int k = 0; k < 4; ++k) metrics[imp * 4 + k] = 0.0;
            }
            continue;
        }
        double auc = 0.0, mrr = 0.0, d5 = 0.0, d10 = 0.0;
        for (int i = threadIdx.x; i < C; i += blockDim.x) {
            if (y[i] == 0) continue;
            const float si = s[i];
            int greater = 0, less_neg = 0, eq_neg = 0;
            for (int j = 0; j < C; ++j) {
                const float sj = s[j];
                greater += (sj > si) || (sj == si && j > i);
                if (y[j] == 0) { less_neg += sj < si; eq_neg += sj == si; }
            }
            const int rank = greater + 1;
            auc += (double)less_neg + 0.5 * (double)eq_neg;
            mrr += 1.0 / (double)rank;
            const double g = 1.0 / log2((double)rank + 1.0);   // (2^1 - 1) / log2(pos + 2), pos = rank-1
            if (rank <= 5) d5 += g;
            if (rank <= 10) d10 += g;
        }
        auc = warp_sum(auc); mrr = warp_sum(mrr); d5 = warp_sum(d5); d10 = warp_sum(d10);
        const int w = threadIdx.x >> 5;
        if ((threadIdx.x & 31) == 0) { red[w][0] = auc; red[w][1] = mrr; red[w][2] = d5; red[w][3] = d10; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double t[4] = {0, 0, 0, 0};
            for (int ww = 0; ww < 4; ++ww) for (int k = 0; k < 4; ++k) t[k] += red[ww][k];
            double i5 = 0.0, i10 = 0.0;
            for (int r = 0; r < npos && r < 10; ++r) {
                const double g = 1.0 / log2((double)r + 2.0);
                if (r < 5) i5 += g;
                i10 += g;
            }
            metrics[imp * 4 + 0] = t[0] / ((double)npos * (double)(C - npos));
            metrics[imp * 4 + 1] = t[1] / (double)npos;
            metrics[imp * 4 + 2] = t[2] / i5;
            metrics[imp * 4 + 3] = t[3] / i10;
            valid[imp] = 1;
        }
        __syncthreads();
    }
}

// deterministic single-block tree reduction over impressions
__global__ void __launch_bounds__(1024)
reduce_metrics_kernel(const double* __restrict__ metrics, const uint8_t* __restrict__ valid,
                      double* __restrict__ sums, int64_t I) {
    __shared__ double sh[5][1024];
    double t[5] = {0, 0, 0, 0, 0};
    for (int64_t i = threadIdx.x; i < I; i += blockDim.x) {
        if (valid[i]) {
            for (int k = 0; k < 4; ++k) t[k] += metrics[i * 4 + k];
            t[4] += 1.0;
        }
    }
    for (int k = 0; k < 5; ++k) sh[k][threadIdx.x] = t[k];
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o)
            for (int k = 0; k < 5; ++k) sh[k][threadIdx.x] += sh[k][threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x < 5) sums[threadIdx.x] = sh[threadIdx.x][0];
}

}  // namespace nr

using namespace nr;

extern "C" {

int nr_eval_score(const float* table, int64_t table_rows, const float* user_vecs, const int64_t* offsets,
                  const int32_t* cand_idx, float* scores, int64_t I, int64_t total_cands, int32_t H,
                  void* stream) {
    NR_REQUIRE(table && user_vecs && offsets && cand_idx && scores, "nr_eval_score: NULL pointer");
    NR_REQUIRE(table_rows > 0 && I >= 0 && total_cands >= 0 && H > 0 && (H & 3) == 0, "nr_eval_score: bad sizes");
    if (I == 0 || total_cands == 0) return 0;
    ProfScope ps("eval_score", (cudaStream_t)stream);
    if (H <= 512) {
        int64_t wb = (I + 7) / 8;
        if (wb > (int64_t)sm_count() * 8) wb = (int64_t)sm_count() * 8;
        const int kv = (H / 4 + 31) / 32;
        cudaStream_t st = (cudaStream_t)stream;
        if (kv <= 1) eval_score_warp_kernel<1><<<(unsigned)wb, 256, 0, st>>>(table, user_vecs, offsets, cand_idx, scores, I, H, table_rows);
        else if (kv == 2) eval_score_warp_kernel<2><<<(unsigned)wb, 256, 0, st>>>(table, user_vecs, offsets, cand_idx, scores, I, H, table_rows);
        else if (kv == 3) eval_score_warp_kernel<3><<<(unsigned)wb, 256, 0, st>>>(table, user_vecs, offsets, cand_idx, scores, I, H, table_rows);
        else eval_score_warp_kernel<4><<<(unsigned)wb, 256, 0, st>>>(table, user_vecs, offsets, cand_idx, scores, I, H, table_rows);
        return check_last("nr_eval_score");
    }
    int64_t blocks = (total_cands * 32 + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    eval_score_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(table, user_vecs, offsets, cand_idx,
                                                                         scores, I, total_cands, H, table_rows);
    return check_last("nr_eval_score");
}

int nr_eval_rank_metrics(const float* scores, const int8_t* labels, const int64_t* offsets, double* metrics,
                         uint8_t* valid, int64_t I, void* stream) {
    NR_REQUIRE(scores && labels && offsets && metrics && valid, "nr_eval_rank_metrics: NULL pointer");
    if (I == 0) return 0;
    int64_t blocks = I < (int64_t)sm_count() * 16 ? I : (int64_t)sm_count() * 16;
    ProfScope ps("rank_metrics", (cudaStream_t)stream);
    rank_metrics_kernel<<<(unsigned)blocks, 128, 0, (cudaStream_t)stream>>>(scores, labels, offsets, metrics,
                                                                           valid, I);
    return check_last("nr_eval_rank_metrics");
}

int nr_eval_reduce_metrics(const double* metrics, const uint8_t* valid, double* sums, int64_t I, void* stream) {
    NR_REQUIRE(metrics && valid && sums, "nr_eval_reduce_metrics: NULL pointer");
    ProfScope ps("reduce_metrics", (cudaStream_t)stream);
    reduce_metrics_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(metrics, valid, sums, I);
    return check_last("nr_eval_reduce_metrics");
}

}  // extern "C"
