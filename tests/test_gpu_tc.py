"""GPU parity of the tcgen05 (bf16 tensor-core) building blocks against torch fp64 matmul on the
same bf16-rounded operands.  Tolerance: fp32 accumulation of exact bf16 products -> rel 1e-5 for
fp32 output; bf16 output adds one rounding (2^-9 relative)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def tc_gemm(A, B, bias, M, taps=1, shifts=(0,), c_bf16=False, K=None):
    import ctypes as C
    from newsrecommendation_b200 import _lib as L
    N = B.shape[0]
    K = K if K is not None else A.shape[1]
    out = torch.empty(M, N, dtype=torch.bfloat16 if c_bf16 else torch.float32, device='cuda')
    sh = (C.c_int32 * 3)(*(list(shifts) + [0, 0, 0])[:3])
    L.call('nr_tc_gemm_nt', A.data_ptr(), A.shape[0], K, A.stride(0), B.data_ptr(), N, B.stride(0),
           None if bias is None else bias.data_ptr(), out.data_ptr(), out.stride(0), int(c_bf16), M, taps,
           C.cast(sh, C.c_void_p), L.stream())
    return out


@pytest.mark.parametrize('M,N,K', [(1000, 1200, 320), (128, 16, 64), (4096 + 77, 200, 400), (333, 400, 208),
                                   (20000, 320, 1200),
                                   (40000, 1200, 320), (38000 + 5, 200, 400), (40000, 400, 200)])   # B-resident mode
def test_tc_gemm_nt(M, N, K):
    g = torch.Generator(device='cpu').manual_seed(M + N + K)
    A = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).cuda()
    B = (torch.randn(N, K, generator=g) * 0.1).to(torch.bfloat16).cuda()
    bias = torch.randn(N, generator=g).cuda()
    ref = (A.double() @ B.double().t() + bias.double()).cpu()
    out = tc_gemm(A, B, bias, M).cpu()
    scale = ref.abs().max().item()
    np.testing.assert_allclose(out.numpy(), ref.numpy(), rtol=1e-5, atol=2e-6 * scale * np.sqrt(K))
    outb = tc_gemm(A, B, bias, M, c_bf16=True).float().cpu()
    np.testing.assert_allclose(outb.numpy(), ref.numpy(), rtol=1e-2, atol=1e-2 * scale)
    out0 = tc_gemm(A, B, None, M).cpu()
    np.testing.assert_allclose(out0.numpy(), (ref - bias.double().cpu()).numpy(), rtol=1e-5,
                               atol=2e-6 * scale * np.sqrt(K))


def test_tc_gemm_k_not_multiple_of_64_and_strided():
    # K = 300 inside rows padded to 320 (the word-embedding layout): TMA zero-fills past K
    M, N, K, ld = 700, 400, 300, 320
    g = torch.Generator(device='cpu').manual_seed(5)
    Af = torch.full((M, ld), 7.0)          # garbage in the padding columns must not leak in
    Af[:, :K] = torch.randn(M, K, generator=g) * 0.4
    A = Af.to(torch.bfloat16).cuda()
    B = (torch.randn(N, ld, generator=g) * 0.1).to(torch.bfloat16).cuda()
    ref = (A[:, :K].double() @ B[:, :K].double().t()).cpu()
    out = tc_gemm(A, B, None, M, K=K).cpu()
    np.testing.assert_allclose(out.numpy(), ref.numpy(), rtol=1e-5, atol=1e-4)


def test_tc_conv_taps():
    # Conv1d(k=3,pad=1) as 3 shifted GEMM taps over titles separated by one zero row
    n, Lw, D, H, Dp = 37, 30, 300, 400, 320
    g = torch.Generator(device='cpu').manual_seed(9)
    x = torch.randn(n, Lw, D, generator=g) * 0.4
    w = torch.randn(H, D, 3, generator=g) * 0.05
    b = torch.randn(H, generator=g) * 0.1
    P = Lw + 1
    X = torch.zeros(n * P, Dp)
    X.view(n, P, Dp)[:, :Lw, :D] = x
    Xb = X.to(torch.bfloat16).cuda()
    Wp = torch.zeros(H, 3 * Dp)
    for t in range(3):
        Wp[:, t * Dp:t * Dp + D] = w[:, :, t]
    Wb = Wp.to(torch.bfloat16).cuda()
    out = tc_gemm(Xb, Wb, b.cuda(), n * P, taps=3, shifts=(-1, 0, 1), K=Dp).cpu().view(n, P, H)[:, :Lw]
    xr = Xb.float().cpu().view(n, P, Dp)[:, :Lw, :D].double()
    wr = Wb.float().cpu().double()
    ref = torch.nn.functional.conv1d(xr.transpose(1, 2), torch.stack([wr[:, t * Dp:t * Dp + D] for t in range(3)], 2),
                                     b.double(), padding=1).transpose(1, 2)
    np.testing.assert_allclose(out.numpy(), ref.numpy(), rtol=1e-5, atol=2e-4)


def test_tc_gemm_nt_rejects_misaligned_output():
    """bf16 C rows must be 16-byte aligned (the epilogue stores 8 bf16 at a time): an error, not a fault."""
    A = torch.zeros(256, 64, dtype=torch.bfloat16, device='cuda')
    B = torch.zeros(300, 64, dtype=torch.bfloat16, device='cuda')
    with pytest.raises(RuntimeError, match='16-byte aligned'):
        tc_gemm(A, B, None, 256, c_bf16=True)          # out (256, 300): 600-byte rows


@pytest.mark.parametrize('M,N,K,taps', [(76000 + 55, 1200, 320, 1), (20000, 304, 1200, 1), (640 + 5, 200, 400, 1),
                                        (30000, 400, 320, 3)])
def test_tc_gemm_nt_multicast_pairs(M, N, K, taps, monkeypatch):
    """NR_TC_MC=1 (opt-in): clusters of two CTAs on adjacent m tiles, each TMA-multicasting half of the weight box
    (odd m-tile counts leave the last pair with one dead CTA).  Checked on GPU in fp32 against torch matmul."""
    monkeypatch.setenv('NR_TC_MC', '1')
    g = torch.Generator(device='cpu').manual_seed(M + N + K)
    A = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).cuda()
    B = (torch.randn(N, taps * K, generator=g) * 0.1).to(torch.bfloat16).cuda()
    bias = torch.randn(N, generator=g).cuda()
    shifts = (-1, 0, 1) if taps == 3 else (0,)
    out = tc_gemm(A, B, bias, M, taps=taps, shifts=shifts, K=K)
    Af = A.float()
    ref = bias.float()[None, :].repeat(M, 1)
    for t, sh in enumerate(shifts):
        As = torch.zeros_like(Af)
        if sh == 0:
            As = Af
        elif sh < 0:
            As[-sh:] = Af[:sh]
        else:
            As[:-sh] = Af[sh:]
        ref += As @ B[:, t * K:(t + 1) * K].float().t()
    scale = ref.abs().max().item()
    err = (out - ref).abs().max().item()
    assert err <= 3e-5 * scale * np.sqrt(K * taps), (err, scale)
