import ctypes as C, sys, os
sys.path.insert(0, '/root/repo')
import torch
from newsrecommendation_b200 import _lib as L
M=4096
for (N,K,lda) in [(1200,301,320),(1200,320,320),(304,301,320)]:
    A = (torch.randn(M, lda, device='cuda') * 0.5).to(torch.bfloat16)
    B = (torch.randn(N, 320, device='cuda') * 0.1).to(torch.bfloat16)
    out = torch.empty(M, N, dtype=torch.bfloat16, device='cuda')
    sh = (C.c_int32 * 3)(0,0,0)
    try:
        L.call('nr_tc_gemm_nt', A.data_ptr(), M, K, A.stride(0), B.data_ptr(), N, B.stride(0), None, out.data_ptr(), out.stride(0), 1, M, 1, C.cast(sh, C.c_void_p), L.stream())
        torch.cuda.synchronize()
        ref = A[:, :K].float() @ B[:, :K].float().t()
        print(N,K,'ok', (out.float()-ref).abs().max().item())
    except Exception as e:
        print(N,K,'FAIL',str(e)[:100]); break
