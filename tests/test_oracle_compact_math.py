"""The mathematics the compact title-encoder backward rests on, checked on the CPU oracle in
fp64 (reference math: src/model/model_utils.py:39-55,78-95; padding_idx: src/model/NRMS.py:71-73):

  1. with a zero embedding row 0, the W_Q/W_K/W_V WEIGHT gradients are sums over the rows of real tokens only, while
     the BIAS gradients are sums over all rows (so dqkv of pad-token rows is needed for the biases and nothing else);
  2. a title made of pad tokens only has dQ = dK = 0 and sum_j dV_j = sum_i dO_i per head (what skip_item uses);
  3. with a NON-zero row 0 (from_pretrained keeps it) pad-token rows do feed the weight gradients: the row list has to
     keep them, which is what dx_rowlist_kernel's row-0 check does.
No GPU: this pins the algebra; tests/test_gpu_fused.py pins the kernels against the all-rows backward."""
import numpy as np
import torch

from oracle import nr_oracle as O

H, HEADS, D, L = 40, 4, 12, 9


def _setup(row0_zero, seed=0):
    g = torch.Generator().manual_seed(seed)
    table = torch.randn(23, D, generator=g, dtype=torch.float64) * 0.3
    if row0_zero:
        table[0] = 0
    ws = [(torch.randn(H, D, generator=g, dtype=torch.float64) * 0.2).requires_grad_(True) for _ in range(3)]
    bs = [(torch.randn(H, generator=g, dtype=torch.float64) * 0.1).requires_grad_(True) for _ in range(3)]
    ids = torch.randint(1, 23, (6, L), generator=g)
    ids[0, 4:] = 0
    ids[1] = 0                      # a title of pad tokens only
    ids[2, ::2] = 0                 # holes inside a title
    ids[4, 1:] = 0
    return table, ws, bs, ids, g


def _qkv_grads(table, ws, bs, ids, g):
    """dL/dq, dL/dk, dL/dv (n,S,H) and the parameter gradients for a random linear loss on the attention output."""
    x = O.embedding_lookup(table, ids)
    n, S, _ = x.shape
    dk = H // HEADS
    qq = x @ ws[0].t() + bs[0]
    kk = x @ ws[1].t() + bs[1]
    vv = x @ ws[2].t() + bs[2]
    for t in (qq, kk, vv):
        t.retain_grad()
    sp = lambda t: t.view(n, S, HEADS, dk).transpose(1, 2)
    ctx = O.scaled_dot_product_attention(sp(qq), sp(kk), sp(vv), dk).transpose(1, 2).reshape(n, S, H)
    w = torch.randn(ctx.shape, generator=g, dtype=torch.float64)
    (ctx * w).sum().backward()
    return x.detach(), (qq.grad, kk.grad, vv.grad), [p.grad for p in ws], [p.grad for p in bs], w


def test_weight_gradients_need_real_token_rows_only_biases_need_all():
    table, ws, bs, ids, g = _setup(row0_zero=True)
    x, dqkv, dws, dbs, _ = _qkv_grads(table, ws, bs, ids, g)
    real = (ids != 0).reshape(-1)
    assert 0 < real.sum() < real.numel()
    X = x.reshape(-1, D)
    for m in range(3):
        G = dqkv[m].reshape(-1, H)
        np.testing.assert_allclose(dws[m].numpy(), (G[real].t() @ X[real]).numpy(), rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(dbs[m].numpy(), G.sum(0).numpy(), rtol=1e-12, atol=1e-14)
        if m != 1:      # (the key-bias gradient is analytically zero: a key shift cancels in the normalised weights)
            assert (G.sum(0) - G[real].sum(0)).abs().max() > 1e-6      # pad rows DO carry bias gradient
    # ... up to the 1e-8 in the normaliser (model_utils.py:53), which leaves O(1e-9)
    assert dbs[1].abs().max() < 1e-7 * dbs[0].abs().max()


def test_all_pad_title_has_zero_dq_dk_and_dv_sums_to_do():
    table, ws, bs, ids, g = _setup(row0_zero=True, seed=1)
    _, dqkv, _, _, w = _qkv_grads(table, ws, bs, ids, g)
    t = 1                                                  # ids[1] is all padding
    # eps = 1e-8 in the normaliser (model_utils.py:53) makes the rows of A sum to 1 - O(1e-9): compare at that level
    scale = dqkv[0][0].abs().max()
    assert dqkv[0][t].abs().max() < 1e-7 * scale and dqkv[1][t].abs().max() < 1e-7 * scale
    np.testing.assert_allclose(dqkv[2][t].sum(0).numpy(), w[t].sum(0).numpy(), rtol=1e-7, atol=1e-9)
    t = 0                                                  # a title with real tokens does not have that property
    assert dqkv[0][t].abs().max() > 1e-6


def test_nonzero_row0_makes_pad_rows_count_for_the_weights():
    table, ws, bs, ids, g = _setup(row0_zero=False, seed=2)
    x, dqkv, dws, _, _ = _qkv_grads(table, ws, bs, ids, g)
    real = (ids != 0).reshape(-1)
    X, G = x.reshape(-1, D), dqkv[2].reshape(-1, H)
    full = G.t() @ X
    np.testing.assert_allclose(dws[2].numpy(), full.numpy(), rtol=1e-12, atol=1e-14)
    assert (full - G[real].t() @ X[real]).abs().max() > 1e-4
