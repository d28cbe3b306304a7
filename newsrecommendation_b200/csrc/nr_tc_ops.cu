// nr_tc_ops.cu — bf16 tensor-core (tcgen05) entry points of the C ABI.
#include "nr_tc_gemm.cuh"

namespace nr {

int make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                      uint32_t box_rows, uint32_t box_cols) {
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstride[1] = {ld_elems * 2};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    // libcuda is reached through the runtime (cudaGetDriverEntryPoint), not linked: the library must
    // load on a box without a driver (build check, symbol tests) and fail only when a kernel is asked for.
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t ce = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (ce != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
            set_error("cuTensorMapEncodeTiled not available from the driver (%s)", cudaGetErrorString(ce));
            return 4;
        }
        encode = (EncodeFn)fn;
    }
    CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim,
                                        gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (CUresult %d) rows=%llu cols=%llu ld=%llu box=%ux%u", (int)r,
                  (unsigned long long)rows, (unsigned long long)cols,
                  (unsigned long long)ld_elems, box_rows, box_cols);
        return 4;
    }
    return 0;
}

}  // namespace nr

using namespace nr;

extern "C" {

int nr_tc_gemm_nt(const void* A, int64_t a_rows, int32_t K, int64_t lda, const void* B, int32_t N, int64_t ldb,
                  const float* bias, void* C, int64_t ldc, int32_t c_bf16, int64_t M, int32_t taps,
                  const int32_t* shifts_host, void* stream) {
    NR_REQUIRE(A && B && C, "nr_tc_gemm_nt: NULL pointer");
    NR_REQUIRE(M >= 0 && N > 0 && K > 0, "nr_tc_gemm_nt: bad sizes");
    NR_REQUIRE(ldc >= N && (ldc % (c_bf16 ? 8 : 4)) == 0 && ((uintptr_t)C & 15) == 0,
               "nr_tc_gemm_nt: C rows must be 16-byte aligned (ldc=%lld: multiple of %d elements)", (long long)ldc,
               c_bf16 ? 8 : 4);
    int sh[3] = {0, 0, 0};
    for (int t = 0; t < taps && t < 3; ++t) sh[t] = shifts_host ? shifts_host[t] : 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (c_bf16)
        return gemm_tc((const __nv_bfloat16*)A, a_rows, K, lda, (const __nv_bfloat16*)B, N, ldb, taps, sh, M,
                       TcEpiBiasBf16{(__nv_bfloat16*)C, ldc, bias, N}, st, "tc.gemm_nt");
    return gemm_tc((const __nv_bfloat16*)A, a_rows, K, lda, (const __nv_bfloat16*)B, N, ldb, taps, sh, M,
                   TcEpiBiasF32{(float*)C, ldc, bias, N}, st, "tc.gemm_nt");
}

}  // extern "C"
