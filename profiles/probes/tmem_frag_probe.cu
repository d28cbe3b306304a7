// tmem_frag_probe.cu — development probe (not part of the library): which TMEM (lane, column) does each register
// of tcgen05.ld.16x256b.{x1,x2} hold, at lane bases +0 / +16 of a warp's quadrant and at column offsets that are
// not multiples of 8?  The fused title-encoder kernel (csrc/nr_fused_enc.cuh) reads Q/K/V straight from the QKV
// accumulator in mma.sync fragment layout with these loads; this probe pins the layout it assumes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o build/tmem_frag_probe tmem_frag_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128) probe(float* out) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(64u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot;
    // fill: lane (row) r = 32*warp + lane, column c -> r*100 + c, 64 columns, with 32x32b stores (thread = lane)
    const uint32_t trow = base + ((uint32_t)(warp * 32) << 16);
    for (int c0 = 0; c0 < 64; c0 += 8) {
        uint32_t v[8];
        for (int j = 0; j < 8; ++j) v[j] = __float_as_uint((float)((warp * 32 + lane) * 100 + c0 + j));
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(trow + c0), "r"(v[0]), "r"(v[1]),
                     "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // case k: (lane base offset, column offset); x2 = 16 columns -> 8 registers
    const int lb[4] = {0, 16, 0, 16};
    const int co[4] = {0, 0, 20, 44};
    for (int k = 0; k < 4; ++k) {
        uint32_t r[8];
        const uint32_t ta = base + ((uint32_t)(warp * 32 + lb[k]) << 16) + (uint32_t)co[k];
        asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(ta) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int j = 0; j < 8; ++j) out[((k * 4 + warp) * 32 + lane) * 8 + j] = __uint_as_float(r[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(64u) : "memory");
}

int main() {
    float* d; float h[4 * 4 * 32 * 8];
    cudaMalloc(&d, sizeof(h));
    cudaMemset(d, 0, sizeof(h));
    probe<<<1, 128>>>(d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("PROBE CUDA ERROR: %s\n", cudaGetErrorString(e)); return 1; }
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    const int lb[4] = {0, 16, 0, 16}, co[4] = {0, 0, 20, 44};
    int bad = 0;
    for (int k = 0; k < 4; ++k)
        for (int w = 0; w < 4; ++w)
            for (int l = 0; l < 32; ++l)
                for (int j = 0; j < 8; ++j) {
                    // assumed: register j of thread l: column block b = j/4, row g + 8*((j>>1)&1), column 8b + 2t + (j&1)
                    const int g = l >> 2, t = l & 3, b = j >> 2;
                    const int row = w * 32 + lb[k] + g + 8 * ((j >> 1) & 1);
                    const int col = co[k] + 8 * b + 2 * t + (j & 1);
                    const float want = (float)(row * 100 + col), got = h[((k * 4 + w) * 32 + l) * 8 + j];
                    if (want != got) { if (bad < 24) printf("MISMATCH case %d warp %d lane %d reg %d: got %.0f want %.0f\n", k, w, l, j, got, want); ++bad; }
                }
    printf("tmem 16x256b.x2 fragment layout: %s (%d mismatches)\n", bad ? "DIFFERENT FROM ASSUMED" : "AS ASSUMED (mma m16n8 C layout)", bad);
    for (int k = 0; k < 4; ++k) {       // raw dump of warp 1 so that a different layout can be read off
        printf("case %d (lane base +%d, col %d), warp 1:\n", k, lb[k], co[k]);
        for (int l = 0; l < 32; ++l) {
            printf("  lane %2d:", l);
            for (int j = 0; j < 8; ++j) printf(" %6.0f", h[((k * 4 + 1) * 32 + l) * 8 + j]);
            printf("\n");
        }
    }
    return bad ? 2 : 0;
}
