// nr_common.cuh — shared device/host helpers for libnr_sm100.so
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/nr_sm100.h"

namespace nr {

// ---- error plumbing -----------------------------------------------------------
void set_error(const char* fmt, ...);
int check_last(const char* what);   // cudaGetLastError -> error code + message

// Per-kernel CUDA-event timing (off by default; bench.py turns it on for one profiled pass).
// A ProfScope placed around a launch records start/stop events on the launch stream and
// counts the launch; nr_prof_report sums the event pairs by name.
struct ProfScope {
    const char* name;
    cudaStream_t st;
    void* slot;
    ProfScope(const char* name, cudaStream_t st);
    ~ProfScope();
};

#define NR_REQUIRE(cond, ...)                      \
    do {                                           \
        if (!(cond)) {                             \
            nr::set_error(__VA_ARGS__);            \
            return 1;                              \
        }                                          \
    } while (0)

#define NR_CUDA(call)                                                               \
    do {                                                                            \
        cudaError_t e_ = (call);                                                    \
        if (e_ != cudaSuccess) {                                                    \
            nr::set_error("%s failed: %s", #call, cudaGetErrorString(e_));          \
            return 2;                                                               \
        }                                                                           \
    } while (0)

constexpr float kEps = 1e-8f;   // src/model/model_utils.py:29,53

inline int sm_count() {
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess)
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n;
}

// bump allocator over the caller-provided workspace
struct Workspace {
    char* base;
    size_t bytes, used;
    Workspace(void* p, size_t b) : base((char*)p), bytes(b), used(0) {}
    template <typename T>
    T* take(size_t count) {
        size_t need = (count * sizeof(T) + 255) & ~size_t(255);
        if (used + need > bytes) return nullptr;
        T* r = (T*)(base + used);
        used += need;
        return r;
    }
    static size_t pad(size_t b) { return (b + 255) & ~size_t(255); }
};

// ---- Philox4x32-7 counter RNG: dropout keep masks ------------------------------
// 7 rounds is the fewest that passes BigCrush (Salmon et al., SC'11, table 2); the customary 10 is safety margin a
// dropout mask does not need, and the rounds are most of the instructions of the staging kernel.
// Element e of stream (seed, sid): word (e & 3) of philox(counter = e >> 2).
#ifndef NR_PHILOX_ROUNDS
#define NR_PHILOX_ROUNDS 7
#endif
__host__ __device__ __forceinline__ void philox_round(uint32_t& c0, uint32_t& c1, uint32_t& c2,
                                                      uint32_t& c3, uint32_t k0, uint32_t k1) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1;
    c3 = (uint32_t)p0;
    c0 = n0;
    c2 = n2;
}

__host__ __device__ __forceinline__ uint4 philox4(uint64_t seed, uint32_t sid, uint64_t ctr) {
    uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32), c2 = sid, c3 = 0x6e72u;
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int i = 0; i < NR_PHILOX_ROUNDS; ++i) {
        philox_round(c0, c1, c2, c3, k0, k1);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

__host__ __device__ __forceinline__ uint32_t drop_threshold(float p) {
    double t = (double)p * 4294967296.0;
    if (t <= 0.0) return 0u;
    if (t >= 4294967295.0) return 4294967295u;
    return (uint32_t)t;
}

// Device-resident seed salt (nr_set_seed_salt): every dropout stream is keyed by seed + *salt.  A train
// step captured in a CUDA graph bakes its host seeds into the kernel arguments; bumping the salt from
// inside the graph gives every replay fresh masks while forward and backward of one replay still agree.
// One copy of the pointer per translation unit (no relocatable device code); nr_set_seed_salt sets all.
static __device__ const unsigned long long* g_seed_salt = nullptr;
inline int set_seed_salt_this_tu(const void* p) {
    return cudaMemcpyToSymbol(g_seed_salt, &p, sizeof(p)) == cudaSuccess ? 0 : 2;
}
__device__ __forceinline__ uint64_t salted(uint64_t seed) {
    const unsigned long long* p = g_seed_salt;
    return p ? seed + __ldg(p) : seed;
}

// ---- dropout streams ---------------------------------------------------------------------------------------------
// One Philox4x32-7 call yields 128 bits = EIGHT 16-bit uniforms: stream element e uses 16-bit field (e & 7) of
// philox(counter = e >> 3) and is kept when field >= thr16 = floor(p * 65536) (p = 0.2: keep probability 0.800003).
// The 32-bit-per-element version of round 1 spent a third of the staging / attention-epilogue instructions on Philox.
// Stream element indices are laid out so that the 8-element groups line up with what one thread handles:
//   input dropout  (NR_STREAM_IN,  element (row r, column c) of an (R x D) activation):   e = r * up8(D) + c
//   output dropout (NR_STREAM_OUT, element (row r, head h, d) of an (R x heads*dk) one):  e = (r * heads + h) * up8(dk) + d
// (nr_dropout_keep_mask exports either stream in the plain (R x width) layout for the tests.)
__host__ __device__ __forceinline__ int64_t up8(int64_t v) { return (v + 7) & ~(int64_t)7; }
__host__ __device__ __forceinline__ uint64_t in_stream_index(int64_t r, int c, int D) { return (uint64_t)(r * up8(D) + c); }
__host__ __device__ __forceinline__ uint64_t out_stream_index(int64_t r, int h, int d, int heads, int dk) {
    return (uint64_t)((r * heads + h) * up8(dk) + d);
}
__host__ __device__ __forceinline__ uint32_t thr16_of(uint32_t thr32) { return thr32 >> 16; }

// 8 keep bits of group G (elements 8G .. 8G+7): bit j = element 8G + j kept
__host__ __device__ __forceinline__ uint32_t dropout8_bits_key(uint64_t G, uint64_t key, uint32_t sid, uint32_t thr32) {
    const uint4 r = philox4(key, sid, G);
    const uint32_t t = thr16_of(thr32);
    return ((r.x & 0xFFFFu) >= t ? 1u : 0u) | ((r.x >> 16) >= t ? 2u : 0u) | ((r.y & 0xFFFFu) >= t ? 4u : 0u) |
           ((r.y >> 16) >= t ? 8u : 0u) | ((r.z & 0xFFFFu) >= t ? 16u : 0u) | ((r.z >> 16) >= t ? 32u : 0u) |
           ((r.w & 0xFFFFu) >= t ? 64u : 0u) | ((r.w >> 16) >= t ? 128u : 0u);
}
// 4 keep bits of the quad starting at element e0 (e0 % 4 == 0): bit j = element e0 + j kept
__host__ __device__ __forceinline__ uint32_t dropout4_bits_key(uint64_t e0, uint64_t key, uint32_t sid, uint32_t thr32) {
    return (dropout8_bits_key(e0 >> 3, key, sid, thr32) >> (e0 & 4)) & 0xFu;
}
__device__ __forceinline__ uint32_t dropout4_bits(uint64_t e0, uint64_t seed, uint32_t sid, uint32_t thr) {
    return dropout4_bits_key(e0, salted(seed), sid, thr);
}
__device__ __forceinline__ float4 apply_keep4(float4 v, uint32_t bits, float scale) {
    v.x = (bits & 1u) ? v.x * scale : 0.f;
    v.y = (bits & 2u) ? v.y * scale : 0.f;
    v.z = (bits & 4u) ? v.z * scale : 0.f;
    v.w = (bits & 8u) ? v.w * scale : 0.f;
    return v;
}
// Applies inverted dropout to 4 consecutive elements starting at stream element e0 (e0 % 4 == 0).
__device__ __forceinline__ float4 dropout4(float4 v, uint64_t e0, uint64_t seed, uint32_t sid, uint32_t thr, float scale) {
    return apply_keep4(v, dropout4_bits(e0, seed, sid, thr), scale);
}

// ---- warp / block reductions -------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg((const float4*)p); }

// ---- matrix views: row-major logical matrices read 4 columns at a time ------------------
// Every view exposes   float4 load4(int64_t row, int col)   with zero fill outside
// [0,rows) x [0,cols); col % 4 == 0 and cols % 4 == 0 always hold.

struct DenseView {
    const float* p;
    int64_t ld;
    int64_t rows;
    int cols;
    __device__ __forceinline__ float4 load4(int64_t r, int c) const {
        if (r >= rows || c >= cols) return make_float4(0.f, 0.f, 0.f, 0.f);
        return ldg4(p + r * ld + c);
    }
};

// up to three stacked (rows_each x cols) matrices addressed as one (3*rows_each x cols):
// W_Q / W_K / W_V are separate Parameters in the reference (model_utils.py:67-69).
struct Stack3View {
    const float* p0;
    const float* p1;
    const float* p2;
    int rows_each;
    int rows;
    int cols;
    __device__ __forceinline__ float4 load4(int64_t r, int c) const {
        if (r >= rows || c >= cols) return make_float4(0.f, 0.f, 0.f, 0.f);
        const int which = (int)r / rows_each;
        const int rr = (int)r - which * rows_each;
        const float* p = which == 0 ? p0 : (which == 1 ? p1 : p2);
        return ldg4(p + (int64_t)rr * cols + c);
    }
};

// Activation rows (n*S, D) behind an nr_rows_t, with optional input dropout.
struct RowsView {
    int mode;
    const float* dense;
    const float* table;
    const int32_t* ids;
    int ids_stride;
    int S, D;
    int64_t rows;   // n*S
    int64_t table_rows;   // ids outside [0, table_rows) read (and skip the gradient of) row 0, the padding row
    uint32_t thr;   // 0 => no dropout
    float scale;
    uint64_t seed;
    __device__ __forceinline__ int64_t id_of(int64_t seq, int s) const {
        const int64_t id = mode == NR_ROWS_TOKEN ? (int64_t)__ldg(ids + seq * ids_stride + s)
                                                 : (int64_t)__ldg(ids + seq * ids_stride);
        return (uint64_t)id < (uint64_t)table_rows ? id : 0;
    }
    // element (row r = seq*S+s, col c)
    __device__ __forceinline__ float4 load4(int64_t r, int c) const {
        if (r >= rows || c >= D) return make_float4(0.f, 0.f, 0.f, 0.f);
        float4 v;
        if (mode == NR_ROWS_DENSE) {
            v = ldg4(dense + r * D + c);
        } else {
            const int64_t seq = r / S;
            const int s = (int)(r - seq * S);
            const int64_t id = id_of(seq, s);
            v = mode == NR_ROWS_TOKEN ? ldg4(table + id * D + c)
                                      : ldg4(table + id * ((int64_t)S * D) + (int64_t)s * D + c);
        }
        if (thr) v = dropout4(v, in_stream_index(r, c, D), seed, NR_STREAM_IN, thr, scale);
        return v;
    }
};

// im2col view of a RowsView for Conv1d(k=3,pad=1): (n*L, 3*D), col = t*D + c -> X[l+t-1, c]
struct ConvView {
    RowsView x;
    int64_t rows;
    int cols;   // 3*D
    __device__ __forceinline__ float4 load4(int64_t r, int c) const {
        if (r >= rows || c >= cols) return make_float4(0.f, 0.f, 0.f, 0.f);
        const int t = c / x.D;
        const int cc = c - t * x.D;
        const int64_t seq = r / x.S;
        const int l = (int)(r - seq * x.S) + t - 1;
        if (l < 0 || l >= x.S) return make_float4(0.f, 0.f, 0.f, 0.f);
        return x.load4(seq * x.S + l, cc);
    }
};

// rows of a small embedding table gathered by one id column: (n, E)
struct EmbedView {
    const float* emb;
    const int32_t* ids;
    int ids_stride;
    int64_t rows;
    int cols;
    int64_t emb_rows;     // ids outside [0, emb_rows) read row 0
    __device__ __forceinline__ float4 load4(int64_t r, int c) const {
        if (r >= rows || c >= cols) return make_float4(0.f, 0.f, 0.f, 0.f);
        int64_t id = __ldg(ids + r * ids_stride);
        if ((uint64_t)id >= (uint64_t)emb_rows) id = 0;
        return ldg4(emb + id * cols + c);
    }
};

RowsView make_rows_view(const nr_rows_t* r, int64_t n, int S, int D);
// per-translation-unit salt setters (defined next to the kernels that draw dropout masks)
int set_seed_salt_ops(const void* p);
int set_seed_salt_ops2(const void* p);
int set_seed_salt_encoder(const void* p);
int validate_rows(const nr_rows_t* r, const char* who);

}  // namespace nr
