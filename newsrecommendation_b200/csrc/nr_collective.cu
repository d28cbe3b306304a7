// nr_collective.cu — the gradient exchange of the data-parallel train step fused with the optimizer
// (src/main.py:82,109-110: DistributedDataParallel's gradient mean followed by optim.Adam.step).
//
// The flat fp32 gradient and parameter buckets of every rank live in SYMMETRIC memory mapped through an NVLink
// multicast object (NVLS).  One kernel per rank then does, for the rank's 1/W slice of the bucket:
//
//     g   = multimem.ld_reduce.add(grad bucket)      reduce-scatter: the NVSwitch sums the W ranks' values in flight
//     p   = Adam(p, g / W, m, v)                     torch.optim.Adam defaults; moments exist only on the owner rank
//     multimem.st(parameter bucket) = p              all-gather: one store lands in every rank's copy
//
// i.e. reduce-scatter + Adam + all-gather in ONE launch with 2/W of an all-reduce's NVLink traffic per GPU and no
// second pass over the bucket; every rank receives bit-identical parameters by construction (only the owner computes
// them).  The launch is bracketed by two cross-rank barriers issued by the host code (torch symmetric-memory signal
// pads): gradients complete everywhere before, updated parameters visible everywhere after.  All of it is plain
// stream-ordered launches, so the whole optimisation step replays from one CUDA graph; the Adam step count is read
// from device memory for that reason (*step_dev = number of COMPLETED steps; bumped by the caller after the update, so
// that a part of the bucket can be exchanged early, on a side stream, with the same count).
#include "nr_common.cuh"

namespace nr {

__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
    return r;
}
__device__ __forceinline__ void multimem_st(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
// same operation order as adam_kernel (nr_basic.cu) = torch's single-tensor Adam: bit-identical for identical g
__device__ __forceinline__ float adam1(float p, float g, float& m, float& v, float lr_over_bc1, float b1, float b2, float eps,
                                       float inv_sqrt_bc2) {
    m = m * b1 + g * (1.f - b1);
    v = v * b2 + (g * g) * (1.f - b2);
    const float denom = sqrtf(v) * inv_sqrt_bc2 + eps;
    return p - lr_over_bc1 * (m / denom);
}

// MC = true: g read through the multicast mapping (reduced over ranks), p written through it (broadcast).
// MC = false: local buckets (single GPU, or the step count on the device for a fully captured step).
template <bool MC>
__global__ void __launch_bounds__(256)
adam_dev_kernel(const float* __restrict__ g_src, float* p_dst, const float* p_local,       // p_dst == p_local when !MC
                float* __restrict__ m, float* __restrict__ v, int64_t lo, int64_t hi, float lr, float b1, float b2,
                float eps, const long long* __restrict__ step_dev, float grad_scale) {
    // bias corrections from the device-resident step count, in double like the host path (nr_adam_step), once per block
    __shared__ float s_bc[2];
    if (threadIdx.x == 0) {
        const double t = (double)(__ldg(step_dev) + 1);      // *step_dev = completed steps; the caller bumps it afterwards
        s_bc[0] = (float)((double)lr / (1.0 - pow((double)b1, t)));
        s_bc[1] = (float)(1.0 / sqrt(1.0 - pow((double)b2, t)));
    }
    __syncthreads();
    const float lr1 = s_bc[0], bc2 = s_bc[1];
    for (int64_t i = lo + 4 * ((int64_t)blockIdx.x * blockDim.x + threadIdx.x); i < hi; i += 4 * (int64_t)gridDim.x * blockDim.x) {
        float4 g = MC ? multimem_ld_reduce_add(g_src + i) : *(const float4*)(g_src + i);
        g.x *= grad_scale; g.y *= grad_scale; g.z *= grad_scale; g.w *= grad_scale;
        const float4 p = *(const float4*)(p_local + i);
        float4 mm = *(float4*)(m + i), vv = *(float4*)(v + i), o;
        o.x = adam1(p.x, g.x, mm.x, vv.x, lr1, b1, b2, eps, bc2);
        o.y = adam1(p.y, g.y, mm.y, vv.y, lr1, b1, b2, eps, bc2);
        o.z = adam1(p.z, g.z, mm.z, vv.z, lr1, b1, b2, eps, bc2);
        o.w = adam1(p.w, g.w, mm.w, vv.w, lr1, b1, b2, eps, bc2);
        *(float4*)(m + i) = mm;
        *(float4*)(v + i) = vv;
        if (MC) multimem_st(p_dst + i, o); else *(float4*)(p_dst + i) = o;
    }
}

}  // namespace nr

using namespace nr;

extern "C" int nr_adam_step_dev(const float* g, const float* g_mc, float* p, float* p_mc, float* m, float* v, int64_t lo,
                                int64_t hi, float lr, float beta1, float beta2, float eps, const void* step_dev,
                                float grad_scale, void* stream) {
    NR_REQUIRE(p && m && v && step_dev && (g || g_mc), "nr_adam_step_dev: NULL pointer");
    NR_REQUIRE((g_mc == nullptr) == (p_mc == nullptr), "nr_adam_step_dev: give both multicast pointers or neither");
    NR_REQUIRE(lo >= 0 && hi >= lo && (lo & 3) == 0 && (hi & 3) == 0, "nr_adam_step_dev: slice [%lld, %lld) must be 4-float aligned",
               (long long)lo, (long long)hi);
    if (hi == lo) return 0;
    const int64_t n4 = (hi - lo) / 4;
    int64_t blocks = (n4 + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope ps(g_mc ? "adam_reduce_scatter_gather" : "adam", st);
    if (g_mc)
        adam_dev_kernel<true><<<(unsigned)blocks, 256, 0, st>>>(g_mc, p_mc, p, m, v, lo, hi, lr, beta1, beta2, eps,
                                                                 (const long long*)step_dev, grad_scale);
    else
        adam_dev_kernel<false><<<(unsigned)blocks, 256, 0, st>>>(g, p, p, m, v, lo, hi, lr, beta1, beta2, eps,
                                                                  (const long long*)step_dev, grad_scale);
    return check_last("nr_adam_step_dev");
}
