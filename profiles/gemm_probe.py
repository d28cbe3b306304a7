"""Times the tcgen05 NT GEMM building block (plain bf16-store epilogue, no bias) on the shapes of the NRMS / NAML
step, to separate 'mainloop + plain store' from what each fused epilogue adds.   python profiles/gemm_probe.py"""
import ctypes as C
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from newsrecommendation_b200 import _lib as L

M = 7040 * 32
SHAPES = [('qkv', 1200, 301, 320, 1), ('dx', 300, 1200, 1200, 1), ('fc1', 200, 400, 400, 1), ('pool_dx', 400, 200, 200, 1),
          ('conv', 400, 301, 320, 3), ('conv_dx', 300, 400, 400, 3)]
for name, N, K, lda, taps in SHAPES:
    A = (torch.randn(M, lda, device='cuda') * 0.5).to(torch.bfloat16)
    Kp = (K + 63) // 64 * 64
    B = (torch.randn(N, taps * Kp, device='cuda') * 0.1).to(torch.bfloat16)
    out = torch.empty(M, (N + 7) // 8 * 8, dtype=torch.bfloat16, device='cuda')      # 16-byte aligned rows
    sh = (C.c_int32 * 3)(*([-1, 0, 1] if taps == 3 else [0, 0, 0]))
    def run():
        L.call('nr_tc_gemm_nt', A.data_ptr(), M, K, A.stride(0), B.data_ptr(), N, B.stride(0), None, out.data_ptr(),
               out.stride(0), 1, M, taps, C.cast(sh, C.c_void_p), L.stream())
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    fl = 2.0 * M * N * K * taps
    by = M * K * 2 + M * N * 2
    print(f'{name:8s} M={M} N={N} K={K} taps={taps}: {ms * 1e3:7.1f} us  {fl / ms / 1e9:7.1f} TF/s  {by / ms / 1e6:7.1f} GB/s (A in + C out)')
