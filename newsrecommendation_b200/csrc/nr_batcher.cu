// nr_batcher.cu — device-side batch assembly: what DatasetTrain.line_mapper /
// DatasetTest.line_mapper do per TSV line in Python (src/dataset.py:17-49, 64-74), done for a whole batch
// from a pre-tokenised, device-resident index of the same behaviors files.  Pure integer work, bit exact.
//
//   history slot t of sample b (T = user_log_length):   pad = T - min(len, T)
//       t <  pad : news index 0, mask 0                 (front padding,            dataset.py:19-20)
//       t >= pad : hist[len - (T - t)], mask 1          (the LAST T clicks kept,   dataset.py:19 x[-T:])
//   candidate slot j (label = slot of the positive):    neg[:label] + pos + neg[label:]   (dataset.py:45-46)
//   features: news_combined[idx] rows of F int32        (dataset.py:40, 47)
//
// One warp per output row (a news slot): the index is resolved by lane 0 and broadcast, the F-wide feature
// row is copied with coalesced 4-byte loads (F = 30..32 for MIND, so one row is one 128-byte line).
#include "nr_common.cuh"

namespace nr {
namespace {

__global__ void batch_train_kernel(const int64_t* __restrict__ hist_off, const int32_t* __restrict__ hist_idx,
                                   const int32_t* __restrict__ pos, const int32_t* __restrict__ neg,
                                   const int64_t* __restrict__ sample_ids, const int64_t* __restrict__ label,
                                   const int32_t* __restrict__ news, int64_t news_rows, int64_t B, int T, int K,
                                   int F, int32_t* __restrict__ history, float* __restrict__ mask,
                                   int32_t* __restrict__ candidate) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int slots = T + 1 + K;
    for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < B * slots; w += warps) {
        const int64_t b = w / slots;
        const int t = (int)(w - b * slots);
        int32_t idx = 0;
        if (lane == 0) {
            const int64_t i = sample_ids[b];
            if (t < T) {
                const int64_t o0 = hist_off[i], len = hist_off[i + 1] - o0;
                const int64_t kept = len < T ? len : T;
                const int pad = T - (int)kept;
                const bool real = t >= pad;
                if (real) idx = hist_idx[o0 + len - (T - t)];
                mask[b * T + t] = real ? 1.f : 0.f;
            } else {
                const int j = t - T;
                const int64_t lb = label[b];
                idx = j == lb ? pos[i] : neg[i * K + (j < lb ? j : j - 1)];
            }
            if (idx < 0 || idx >= news_rows) idx = 0;      // unknown ids map to the pad row (dataset.py:14-15)
        }
        idx = __shfl_sync(0xffffffffu, idx, 0);
        const int32_t* src = news + (int64_t)idx * F;
        int32_t* dst = t < T ? history + (b * T + t) * F : candidate + (b * (1 + K) + (t - T)) * F;
        for (int c = lane; c < F; c += 32) dst[c] = __ldg(src + c);
    }
}

__global__ void batch_eval_history_kernel(const int64_t* __restrict__ hist_off, const int32_t* __restrict__ hist_idx,
                                          const int64_t* __restrict__ imp_ids, int64_t news_rows, int64_t B, int T,
                                          int32_t* __restrict__ out_idx, float* __restrict__ out_mask) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < B * T; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = e / T;
        const int t = (int)(e - b * T);
        const int64_t i = imp_ids ? imp_ids[b] : b;
        const int64_t o0 = hist_off[i], len = hist_off[i + 1] - o0;
        const int64_t kept = len < T ? len : T;
        const bool real = t >= T - (int)kept;
        int32_t idx = real ? hist_idx[o0 + len - (T - t)] : 0;
        if (idx < 0 || idx >= news_rows) idx = 0;
        out_idx[e] = idx;
        out_mask[e] = real ? 1.f : 0.f;
    }
}

}  // namespace
}  // namespace nr

using namespace nr;

extern "C" {

int nr_batch_train(const int64_t* hist_off, const int32_t* hist_idx, const int32_t* pos, const int32_t* neg,
                   const int64_t* sample_ids, const int64_t* label, const int32_t* news_combined, int64_t news_rows,
                   int64_t B, int32_t T, int32_t K, int32_t F, int32_t* history, float* mask, int32_t* candidate,
                   void* stream) {
    NR_REQUIRE(B >= 0 && T > 0 && K > 0 && F > 0 && news_rows > 0, "nr_batch_train: bad sizes");
    if (B == 0) return 0;                 // an empty batch has no buffers to point at
    NR_REQUIRE(hist_off && hist_idx && pos && neg && sample_ids && label && news_combined && history && mask && candidate,
               "nr_batch_train: NULL pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t rows = B * (T + 1 + K);
    int64_t blocks = (rows + 7) / 8;                         // 8 warps per block, one warp per news slot
    if (blocks > (int64_t)sm_count() * 8) blocks = (int64_t)sm_count() * 8;
    ProfScope ps("batch.train", st);
    batch_train_kernel<<<(unsigned)blocks, 256, 0, st>>>(hist_off, hist_idx, pos, neg, sample_ids, label, news_combined,
                                                         news_rows, B, T, K, F, history, mask, candidate);
    return check_last("batch.train");
}

int nr_batch_eval_history(const int64_t* hist_off, const int32_t* hist_idx, const int64_t* imp_ids, int64_t news_rows,
                          int64_t B, int32_t T, int32_t* out_idx, float* out_mask, void* stream) {
    NR_REQUIRE(B >= 0 && T > 0 && news_rows > 0, "nr_batch_eval_history: bad sizes");
    if (B == 0) return 0;
    NR_REQUIRE(hist_off && hist_idx && out_idx && out_mask, "nr_batch_eval_history: NULL pointer");
    cudaStream_t st = (cudaStream_t)stream;
    int64_t blocks = (B * T + 255) / 256;
    if (blocks > (int64_t)sm_count() * 8) blocks = (int64_t)sm_count() * 8;
    ProfScope ps("batch.eval_history", st);
    batch_eval_history_kernel<<<(unsigned)blocks, 256, 0, st>>>(hist_off, hist_idx, imp_ids, news_rows, B, T, out_idx,
                                                                out_mask);
    return check_last("batch.eval_history");
}

}  // extern "C"
