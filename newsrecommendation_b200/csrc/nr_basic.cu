// nr_basic.cu — library plumbing and the HBM-bound elementwise/gather ops.
#include <stdarg.h>
#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>
#include "nr_common.cuh"

namespace nr {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_last(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: kernel launch failed: %s", what, cudaGetErrorString(e));
        return 3;
    }
    return 0;
}

// ---- launch counter + optional per-kernel event timing ------------------------------------
static std::atomic<long long> g_launches{0};
static std::atomic<int> g_prof_on{0};
struct ProfRec { const char* name; cudaEvent_t a, b; };
static std::mutex g_prof_mu;
static std::vector<ProfRec*> g_prof_recs;

ProfScope::ProfScope(const char* n, cudaStream_t s) : name(n), st(s), slot(nullptr) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    ProfRec* r = new ProfRec{n, nullptr, nullptr};
    cudaEventCreate(&r->a);
    cudaEventCreate(&r->b);
    cudaEventRecord(r->a, s);
    slot = r;
}
ProfScope::~ProfScope() {
    if (!slot) return;
    ProfRec* r = (ProfRec*)slot;
    cudaEventRecord(r->b, st);
    std::lock_guard<std::mutex> lk(g_prof_mu);
    g_prof_recs.push_back(r);
}

RowsView make_rows_view(const nr_rows_t* r, int64_t n, int S, int D) {
    RowsView v;
    v.mode = r->mode;
    v.dense = r->dense;
    v.table = r->table;
    v.ids = r->ids;
    v.ids_stride = r->ids_stride;
    v.S = S;
    v.D = D;
    v.rows = n * S;
    v.table_rows = r->table_rows;
    v.thr = r->drop_p > 0.f ? drop_threshold(r->drop_p) : 0u;
    v.scale = r->drop_p > 0.f ? 1.0f / (1.0f - r->drop_p) : 1.0f;
    v.seed = r->seed;
    return v;
}

int validate_rows(const nr_rows_t* r, const char* who) {
    NR_REQUIRE(r != nullptr, "%s: rows is NULL", who);
    NR_REQUIRE(r->mode >= NR_ROWS_DENSE && r->mode <= NR_ROWS_NEWS, "%s: bad rows.mode %d", who, r->mode);
    if (r->mode == NR_ROWS_DENSE) {
        NR_REQUIRE(r->dense != nullptr, "%s: rows.dense is NULL", who);
    } else {
        NR_REQUIRE(r->table != nullptr && r->ids != nullptr, "%s: rows.table/ids is NULL", who);
        NR_REQUIRE(r->ids_stride >= 1, "%s: rows.ids_stride must be >= 1", who);
        NR_REQUIRE(r->table_rows >= 1, "%s: rows.table_rows must be the row count of rows.table (got %lld)", who,
                   (long long)r->table_rows);
    }
    NR_REQUIRE(r->drop_p >= 0.f && r->drop_p < 1.f, "%s: drop_p %f outside [0,1)", who, r->drop_p);
    return 0;
}

// ---- kernels -----------------------------------------------------------------------------
// A1/A2 gather: one float4 per thread; consecutive threads walk one row -> 128-bit coalesced.
__global__ void gather_rows_kernel(const float* __restrict__ table, const int32_t* __restrict__ ids,
                                   float* __restrict__ out, int64_t total4, int w4, int64_t table_rows) {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total4;
         t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = t / w4;
        const int q = (int)(t - i * w4);
        int64_t id = __ldg(ids + i);
        if ((uint64_t)id >= (uint64_t)table_rows) id = 0;        // out-of-range ids read the padding row
        ((float4*)out)[t] = __ldg((const float4*)table + id * w4 + q);
    }
}

// keep[i] for the plain (rows x row_width) layout of a dropout stream: element i = (row, col); the stream index of
// (row, col) follows nr_common.cuh (rows padded to 8 elements; with seg_width = head width for the output stream)
__global__ void keep_mask_kernel(uint8_t* keep, int64_t count, uint32_t thr, uint64_t seed, uint32_t sid, int row_width,
                                 int seg_width) {
    const uint64_t key = salted(seed);
    const int nseg = row_width / seg_width;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / row_width;
        const int col = (int)(i - row * row_width);
        const int seg = col / seg_width, d = col - seg * seg_width;
        const uint64_t e = (uint64_t)((row * nseg + seg) * up8(seg_width) + d);
        keep[i] = (dropout8_bits_key(e >> 3, key, sid, thr) >> (e & 7)) & 1u;
    }
}

__global__ void pad_blend_fwd_kernel(const float* __restrict__ v, const float* __restrict__ mask,
                                     const float* __restrict__ pad, float* __restrict__ out,
                                     int64_t rows, int h4) {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < rows * h4;
         t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = t / h4;
        const int q = (int)(t - r * h4);
        const float m = __ldg(mask + r);
        const float4 a = __ldg((const float4*)v + t);
        const float4 p = __ldg((const float4*)pad + q);
        const float im = 1.f - m;   // v*m + pad*(1-m), same association as NRMS.py:60
        ((float4*)out)[t] = make_float4(a.x * m + p.x * im, a.y * m + p.y * im, a.z * m + p.z * im,
                                        a.w * m + p.w * im);
    }
}

// dv = dout*m ; dpad += sum_r dout*(1-m).
// Fast path (H % 4 == 0, H/4 <= 256): a block is RS row slots x H/4 column threads; every thread streams its
// 4-column group of rows slot, slot + RS*grid, ... with 16-byte accesses (RS*H/4 threads in flight per block instead
// of H/4: the kernel is latency-bound, 10 MB in and out), the slots' partial column sums meet in shared memory and
// slot 0 issues the block's H atomics (the serial chain per dpad address is the block count, so one block per SM).
constexpr int PADB_THREADS = 512;
__global__ void __launch_bounds__(PADB_THREADS) pad_blend_bwd_slots_kernel(
        const float* __restrict__ dout, const float* __restrict__ mask, float* __restrict__ dv,
        float* __restrict__ dpad, int64_t rows, int H) {
    __shared__ float4 part[PADB_THREADS];
    const int tpr = H >> 2, RS = PADB_THREADS / tpr;
    const int slot = threadIdx.x / tpr, cg = threadIdx.x - slot * tpr;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (slot < RS) {
        const int c = cg * 4;
#pragma unroll 4
        for (int64_t r = (int64_t)blockIdx.x * RS + slot; r < rows; r += (int64_t)gridDim.x * RS) {
            const float m = __ldg(mask + r), im = 1.f - m;
            const float4 g = ldg4(dout + r * H + c);
            *(float4*)(dv + r * H + c) = make_float4(g.x * m, g.y * m, g.z * m, g.w * m);
            acc.x = fmaf(g.x, im, acc.x); acc.y = fmaf(g.y, im, acc.y); acc.z = fmaf(g.z, im, acc.z); acc.w = fmaf(g.w, im, acc.w);
        }
    }
    part[threadIdx.x] = acc;
    __syncthreads();
    if (slot == 0) {
        for (int s2 = 1; s2 < RS; ++s2) {           // fixed slot order
            const float4 o = part[s2 * tpr + cg];
            acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w;
        }
        const int c = cg * 4;
        atomicAdd(dpad + c, acc.x); atomicAdd(dpad + c + 1, acc.y); atomicAdd(dpad + c + 2, acc.z); atomicAdd(dpad + c + 3, acc.w);
    }
}

// general shapes: grid.x blocks each own a strided set of rows, a thread owns one column
__global__ void pad_blend_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ mask,
                                     float* __restrict__ dv, float* __restrict__ dpad, int64_t rows, int H) {
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
        float acc = 0.f;
        for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
            const float m = __ldg(mask + r);
            const float g = __ldg(dout + r * H + c);
            dv[r * H + c] = g * m;
            acc += g * (1.f - m);
        }
        atomicAdd(dpad + c, acc);
    }
}

// score + CE: one warp per sample.  Each score is computed once (16-byte loads when H % 4 == 0) and kept by lane
// k % 32, so the log-sum-exp needs no second pass over the vectors; C > 32 walks the candidates in groups of 32
// with a running maximum.  The dot products are taken 8 candidates at a time so that a lane has 8 independent
// loads in flight per step of h (the kernel is pure latency at B = 128: 1 MB of vectors, one warp per sample).
__device__ __forceinline__ void dot_rows8(const float* __restrict__ c, const float* __restrict__ u, int H, int cnt,
                                          int lane, float (&out)[8]) {
    float s[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] = 0.f;
    if ((H & 3) == 0) {
        for (int h = lane * 4; h < H; h += 128) {
            const float4 v = ldg4(u + h);
            float4 a[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = j < cnt ? ldg4(c + (int64_t)j * H + h) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                s[j] = fmaf(a[j].x, v.x, s[j]); s[j] = fmaf(a[j].y, v.y, s[j]);
                s[j] = fmaf(a[j].z, v.z, s[j]); s[j] = fmaf(a[j].w, v.w, s[j]);
            }
        }
    } else {
        for (int h = lane; h < H; h += 32) {
            const float v = __ldg(u + h);
#pragma unroll
            for (int j = 0; j < 8; ++j) s[j] = fmaf(j < cnt ? __ldg(c + (int64_t)j * H + h) : 0.f, v, s[j]);
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) out[j] = warp_sum(s[j]);
}
__global__ void score_ce_fwd_kernel(const float* __restrict__ cand, const float* __restrict__ user,
                                    const int64_t* __restrict__ label, float* __restrict__ score,
                                    float* __restrict__ per_sample, int64_t B, int C, int H) {
    const int lane = threadIdx.x & 31;
    const int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (b >= B) return;
    const float* u = user + b * H;
    float mx = -INFINITY, picked = 0.f, z = 0.f;
    int arg = 0;
    const int lab = (int)label[b];
    for (int k0 = 0; k0 < C; k0 += 32) {
        float mine = -INFINITY, gmx = mx;
        const int kn = C - k0 < 32 ? C - k0 : 32;
        for (int j0 = 0; j0 < kn; j0 += 8) {
            float s8[8];
            const int cnt = kn - j0 < 8 ? kn - j0 : 8;
            dot_rows8(cand + (b * C + k0 + j0) * H, u, H, cnt, lane, s8);
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) {
                if (jj < cnt) {
                    const int j = j0 + jj, k = k0 + j;
                    const float s = s8[jj];
                    if (lane == j) { mine = s; score[b * C + k] = s; }
                    if (s > gmx) { gmx = s; arg = k; }      // first maximum, as torch.argmax
                    if (k == lab) picked = s;
                }
            }
        }
        z = z * expf(mx - gmx) + warp_sum(lane < kn ? expf(mine - gmx) : 0.f);     // exp(-inf) = 0 on the first group
        mx = gmx;
    }
    if (lane == 0) {
        per_sample[b * 2 + 0] = (mx + logf(z)) - picked;
        per_sample[b * 2 + 1] = (arg == lab) ? 1.f : 0.f;
    }
}

// fixed-order reduction of the per-sample losses -> mean loss, hit count (deterministic)
__global__ void score_ce_reduce_kernel(const float* __restrict__ per_sample, float* loss, float* hits,
                                       int64_t B) {
    __shared__ double sl[256], sh[256];
    double l = 0.0, h = 0.0;
    for (int64_t b = threadIdx.x; b < B; b += blockDim.x) {
        l += per_sample[b * 2];
        h += per_sample[b * 2 + 1];
    }
    sl[threadIdx.x] = l;
    sh[threadIdx.x] = h;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) { sl[threadIdx.x] += sl[threadIdx.x + o]; sh[threadIdx.x] += sh[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { *loss = (float)(sl[0] / (double)B); *hits = (float)sh[0]; }
}

// ds = (softmax(s) - onehot) * gloss / B ; dcand = ds*u ; duser = sum_k ds_k c_k.  block per sample.
__global__ void score_ce_bwd_kernel(const float* __restrict__ cand, const float* __restrict__ user,
                                    const int64_t* __restrict__ label, const float* __restrict__ score,
                                    const float* __restrict__ gloss, float* __restrict__ dcand,
                                    float* __restrict__ duser, int64_t B, int C, int H) {
    extern __shared__ float ds[];
    const int64_t b = blockIdx.x;
    if (threadIdx.x == 0) {
        float mx = -INFINITY;
        for (int k = 0; k < C; ++k) mx = fmaxf(mx, score[b * C + k]);
        float z = 0.f;
        for (int k = 0; k < C; ++k) z += expf(score[b * C + k] - mx);
        const float g = gloss[0] / (float)B;
        const int lab = (int)label[b];
        for (int k = 0; k < C; ++k)
            ds[k] = (expf(score[b * C + k] - mx) / z - (k == lab ? 1.f : 0.f)) * g;
    }
    __syncthreads();
    for (int h = threadIdx.x; h < H; h += blockDim.x) {
        const float u = __ldg(user + b * H + h);
        float du = 0.f;
        for (int k = 0; k < C; ++k) {
            const float d = ds[k];
            dcand[(b * C + k) * H + h] = d * u;
            du = fmaf(d, __ldg(cand + (b * C + k) * H + h), du);
        }
        duser[b * H + h] = du;
    }
}

// torch.optim.Adam (amsgrad off, wd 0): same operation order as torch's single-tensor path.
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, int64_t count, float lr_over_bc1, float beta1,
                            float beta2, float eps, float inv_sqrt_bc2, float gscale) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count;
         i += (int64_t)gridDim.x * blockDim.x) {
        const float gi = g[i] * gscale;
        const float mi = m[i] * beta1 + gi * (1.f - beta1);
        const float vi = v[i] * beta2 + (gi * gi) * (1.f - beta2);
        m[i] = mi;
        v[i] = vi;
        const float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
        p[i] = p[i] - lr_over_bc1 * (mi / denom);
    }
}

static inline unsigned grid_for(int64_t work, int threads, int per_sm = 8) {
    int64_t g = (work + threads - 1) / threads;
    const int64_t cap = (int64_t)sm_count() * per_sm;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (unsigned)g;
}

}  // namespace nr

using namespace nr;

extern "C" {

int nr_abi_version(void) { return NR_ABI_VERSION; }

long long nr_launch_count(void) { return nr::g_launches.load(); }

int nr_prof_enable(int on) {
    nr::g_prof_on.store(on ? 1 : 0);
    return 0;
}

// Synchronises the device, sums the recorded event pairs by kernel name and clears them.
// Writes lines "name launches total_ms\n" into buf (NUL terminated); returns bytes needed.
size_t nr_prof_report(char* buf, size_t cap) {
    cudaDeviceSynchronize();
    std::vector<nr::ProfRec*> recs;
    {
        std::lock_guard<std::mutex> lk(nr::g_prof_mu);
        recs.swap(nr::g_prof_recs);
    }
    std::map<std::string, std::pair<long long, double>> agg;
    for (nr::ProfRec* r : recs) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r->a, r->b) == cudaSuccess) {
            auto& e = agg[r->name];
            e.first += 1;
            e.second += ms;
        }
        cudaEventDestroy(r->a);
        cudaEventDestroy(r->b);
        delete r;
    }
    std::string out;
    char line[256];
    for (auto& kv : agg) {
        snprintf(line, sizeof(line), "%s %lld %.6f\n", kv.first.c_str(), kv.second.first, kv.second.second);
        out += line;
    }
    if (buf && cap) {
        size_t n = out.size() < cap - 1 ? out.size() : cap - 1;
        memcpy(buf, out.data(), n);
        buf[n] = 0;
    }
    return out.size() + 1;
}
const char* nr_last_error(void) { return nr::g_err; }

int nr_device_info(int32_t* sms, int32_t* major, int32_t* minor) {
    int dev = 0;
    NR_CUDA(cudaGetDevice(&dev));
    int a = 0, b = 0, c = 0;
    NR_CUDA(cudaDeviceGetAttribute(&a, cudaDevAttrMultiProcessorCount, dev));
    NR_CUDA(cudaDeviceGetAttribute(&b, cudaDevAttrComputeCapabilityMajor, dev));
    NR_CUDA(cudaDeviceGetAttribute(&c, cudaDevAttrComputeCapabilityMinor, dev));
    if (sms) *sms = a;
    if (major) *major = b;
    if (minor) *minor = c;
    return 0;
}

int nr_set_seed_salt(const void* salt_dev) {
    if (int e = nr::set_seed_salt_this_tu(salt_dev)) { nr::set_error("nr_set_seed_salt: cudaMemcpyToSymbol failed"); return e; }
    if (nr::set_seed_salt_ops(salt_dev) || nr::set_seed_salt_ops2(salt_dev) || nr::set_seed_salt_encoder(salt_dev)) {
        nr::set_error("nr_set_seed_salt: cudaMemcpyToSymbol failed");
        return 2;
    }
    return 0;
}

int nr_dropout_keep_mask(uint8_t* keep, int64_t count, float p, uint64_t seed, int32_t sid, int32_t row_width,
                         int32_t seg_width, void* stream) {
    NR_REQUIRE(keep && count >= 0 && p >= 0.f && p < 1.f, "nr_dropout_keep_mask: bad arguments");
    NR_REQUIRE(row_width >= 1 && seg_width >= 1 && row_width % seg_width == 0,
               "nr_dropout_keep_mask: row_width %d must be a positive multiple of seg_width %d", row_width, seg_width);
    if (count == 0) return 0;
    ProfScope ps_keep_mask("keep_mask", (cudaStream_t)stream);
    keep_mask_kernel<<<grid_for(count, 256), 256, 0, (cudaStream_t)stream>>>(
        keep, count, drop_threshold(p), seed, (uint32_t)sid, row_width, seg_width);
    return check_last("nr_dropout_keep_mask");
}

int nr_gather_rows(const float* table, int64_t table_rows, const int32_t* ids, int64_t count,
                   int32_t width, float* out, void* stream) {
    if (count == 0) return 0;
    NR_REQUIRE(table && ids && out, "nr_gather_rows: NULL pointer");
    NR_REQUIRE(width > 0 && (width & 3) == 0, "nr_gather_rows: width %d must be a positive multiple of 4", width);
    NR_REQUIRE(table_rows > 0 && count >= 0, "nr_gather_rows: bad sizes");
    if (count == 0) return 0;
    const int64_t total4 = count * (width / 4);
    ProfScope ps_gather_rows("gather_rows", (cudaStream_t)stream);
    gather_rows_kernel<<<grid_for(total4, 256, 16), 256, 0, (cudaStream_t)stream>>>(table, ids, out, total4,
                                                                                    width / 4, table_rows);
    return check_last("nr_gather_rows");
}

int nr_pad_blend_fwd(const float* v, const float* mask, const float* pad, float* out, int64_t n,
                     int32_t T, int32_t H, void* stream) {
    NR_REQUIRE(v && mask && pad && out, "nr_pad_blend_fwd: NULL pointer");
    NR_REQUIRE(H > 0 && (H & 3) == 0, "nr_pad_blend_fwd: H %d must be a multiple of 4", H);
    if (n * T == 0) return 0;
    ProfScope ps_pad_blend_fwd("pad_blend_fwd", (cudaStream_t)stream);
    pad_blend_fwd_kernel<<<grid_for(n * T * (H / 4), 256), 256, 0, (cudaStream_t)stream>>>(v, mask, pad, out,
                                                                                         n * T, H / 4);
    return check_last("nr_pad_blend_fwd");
}

int nr_pad_blend_bwd(const float* dout, const float* mask, float* dv, float* dpad, int64_t n, int32_t T,
                     int32_t H, void* stream) {
    NR_REQUIRE(dout && mask && dv && dpad, "nr_pad_blend_bwd: NULL pointer");
    if (n * T == 0) return 0;
    // every block ends with H atomics onto the same H addresses of dpad: the serial chain per address is the
    // block count, so one block per SM (not more) and deep unrolling inside
    int64_t g = n * T < sm_count() ? n * T : sm_count();
    ProfScope ps_pad_blend_bwd("pad_blend_bwd", (cudaStream_t)stream);
    if ((H & 3) == 0 && H / 4 <= PADB_THREADS / 2)
        pad_blend_bwd_slots_kernel<<<(unsigned)g, PADB_THREADS, 0, (cudaStream_t)stream>>>(dout, mask, dv, dpad, n * T, H);
    else
        pad_blend_bwd_kernel<<<(unsigned)g, 128, 0, (cudaStream_t)stream>>>(dout, mask, dv, dpad, n * T, H);
    return check_last("nr_pad_blend_bwd");
}

int nr_score_ce_fwd(const float* cand, const float* user, const int64_t* label, float* score, float* loss,
                    float* hits, float* per_sample_ws, int64_t B, int32_t C, int32_t H, void* stream) {
    NR_REQUIRE(cand && user && label && score && loss && hits && per_sample_ws, "nr_score_ce_fwd: NULL pointer");
    NR_REQUIRE(B > 0 && C > 0 && H > 0, "nr_score_ce_fwd: bad sizes");
    const int threads = B <= 2 * (int64_t)sm_count() ? 32 : 128;      // small batches: one warp per block, spread over the SMs
    ProfScope ps_score_ce_fwd("score_ce_fwd", (cudaStream_t)stream);
    score_ce_fwd_kernel<<<(unsigned)((B * 32 + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(
        cand, user, label, score, per_sample_ws, B, C, H);
    if (int e = check_last("nr_score_ce_fwd")) return e;
    ProfScope ps_score_ce_reduce("score_ce_reduce", (cudaStream_t)stream);
    score_ce_reduce_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(per_sample_ws, loss, hits, B);
    return check_last("nr_score_ce_fwd(reduce)");
}

int nr_score_ce_bwd(const float* cand, const float* user, const int64_t* label, const float* score,
                    const float* gloss, float* dcand, float* duser, int64_t B, int32_t C, int32_t H,
                    void* stream) {
    NR_REQUIRE(cand && user && label && score && gloss && dcand && duser, "nr_score_ce_bwd: NULL pointer");
    NR_REQUIRE(B > 0 && C > 0 && C <= 1024 && H > 0, "nr_score_ce_bwd: bad sizes");
    ProfScope ps_score_ce_bwd("score_ce_bwd", (cudaStream_t)stream);
    score_ce_bwd_kernel<<<(unsigned)B, 128, C * sizeof(float), (cudaStream_t)stream>>>(
        cand, user, label, score, gloss, dcand, duser, B, C, H);
    return check_last("nr_score_ce_bwd");
}

int nr_adam_step(float* p, const float* g, float* m, float* v, int64_t count, float lr, float beta1,
                 float beta2, float eps, int32_t step, float grad_scale, void* stream) {
    NR_REQUIRE(p && g && m && v, "nr_adam_step: NULL pointer");
    NR_REQUIRE(step >= 1, "nr_adam_step: step is 1-based");
    if (count == 0) return 0;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    ProfScope ps_adam("adam", (cudaStream_t)stream);
    adam_kernel<<<grid_for(count, 256), 256, 0, (cudaStream_t)stream>>>(
        p, g, m, v, count, (float)((double)lr / bc1), beta1, beta2, eps, (float)(1.0 / sqrt(bc2)), grad_scale);
    return check_last("nr_adam_step");
}

}  // extern "C"

// ---- allocation-free variant of nr_score_ce_fwd (no counterpart in the reference; added
// for callers that cannot allocate: CUDA-graph capture, tight serving loops) -----------------------
// nr_score_ce_fwd needs a caller-provided (B,2) fp32 scratch (per-sample loss / hit).  Here that scratch is a
// file-scope device array of NR_SCORE_CE_NOALLOC_MAX_B * 2 floats = 128 KB, held by every GPU that loads the library
// until the process exits.  In HOST code the array's name is only a stand-in, not a device address: the address is
// taken once per device with cudaGetSymbolAddress (that device current) and passed to the kernels as a pointer.
namespace nr {
static __device__ __align__(256) float g_score_ce_scratch[2 * NR_SCORE_CE_NOALLOC_MAX_B];
static std::mutex g_scratch_mu;
static float* g_scratch_ptr[64] = {nullptr};      // per device ordinal
static int score_ce_scratch(float** out) {
    int dev = -1;
    NR_CUDA(cudaGetDevice(&dev));
    NR_REQUIRE(dev >= 0 && dev < 64, "nr_score_ce_fwd_noalloc: device ordinal %d outside [0,64)", dev);
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        float* p = nullptr;
        NR_CUDA(cudaGetSymbolAddress((void**)&p, g_score_ce_scratch));      // address on the CURRENT device
        NR_REQUIRE(p != nullptr, "nr_score_ce_fwd_noalloc: cudaGetSymbolAddress returned NULL");
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return 0;
}
}  // namespace nr

extern "C" int nr_score_ce_fwd_noalloc(const float* cand, const float* user, const int64_t* label, float* score,
                                       float* loss, float* hits, int64_t B, int32_t C, int32_t H, void* stream) {
    NR_REQUIRE(cand && user && label && score && loss && hits, "nr_score_ce_fwd_noalloc: NULL pointer");
    NR_REQUIRE(B > 0 && C > 0 && H > 0, "nr_score_ce_fwd_noalloc: bad sizes");
    NR_REQUIRE(B <= (int64_t)NR_SCORE_CE_NOALLOC_MAX_B,
               "nr_score_ce_fwd_noalloc: B=%lld exceeds the built-in scratch (NR_SCORE_CE_NOALLOC_MAX_B = %d samples); "
               "use nr_score_ce_fwd with a caller-owned workspace", (long long)B, (int)NR_SCORE_CE_NOALLOC_MAX_B);
    float* per_sample = nullptr;
    if (int e = nr::score_ce_scratch(&per_sample)) return e;
    // every element [0, 2B) of the scratch is overwritten by the first kernel before the second reads it: nothing has
    // to start at zero, the previous call's contents are irrelevant.  No ProfScope here (it may create events).
    nr::g_launches.fetch_add(2, std::memory_order_relaxed);
    const int threads = B <= 2 * (int64_t)nr::sm_count() ? 32 : 128;     // same launch shape as nr_score_ce_fwd
    nr::score_ce_fwd_kernel<<<(unsigned)((B * 32 + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(
        cand, user, label, score, per_sample, B, C, H);
    if (int e = nr::check_last("nr_score_ce_fwd_noalloc")) return e;
    nr::score_ce_reduce_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(per_sample, loss, hits, B);
    return nr::check_last("nr_score_ce_fwd_noalloc(reduce)");
}

// ---- small stream-ordered utilities so that a captured train step contains no ATen kernels ---------------------
namespace nr {
__global__ void salt_bump_kernel(unsigned long long* salt) { *salt += 1ull; }
}  // namespace nr
extern "C" int nr_seed_salt_bump(void* salt_dev, void* stream) {
    NR_REQUIRE(salt_dev, "nr_seed_salt_bump: NULL pointer");
    nr::ProfScope ps("salt_bump", (cudaStream_t)stream);
    nr::salt_bump_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((unsigned long long*)salt_dev);
    return nr::check_last("nr_seed_salt_bump");
}
extern "C" int nr_zero_bytes(void* dst, size_t bytes, void* stream) {
    if (bytes == 0) return 0;
    NR_REQUIRE(dst, "nr_zero_bytes: NULL pointer");
    NR_CUDA(cudaMemsetAsync(dst, 0, bytes, (cudaStream_t)stream));
    return 0;
}
extern "C" int nr_concat2(const void* a, size_t bytes_a, const void* b, size_t bytes_b, void* out, void* stream) {
    NR_REQUIRE(out && (a || !bytes_a) && (b || !bytes_b), "nr_concat2: NULL pointer");
    if (bytes_a) NR_CUDA(cudaMemcpyAsync(out, a, bytes_a, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    if (bytes_b) NR_CUDA(cudaMemcpyAsync((char*)out + bytes_a, b, bytes_b, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
}
