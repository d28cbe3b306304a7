"""The fused NRMS title-encoder kernel (csrc/nr_fused_enc.cuh: QKV projection -> attention from TMEM -> fc1 ->
pooling in one persistent tcgen05 kernel) against the separate-launch bf16 pipeline and the CPU oracle
(src/model/NRMS.py:28-35, model_utils.py:13-31,39-55,78-95).

  * inference variant (save == NULL; what the news-vector cache runs, main.py:190-196): vs oracle at the bf16
    tolerance (rel 1e-2 of the output scale) and vs the separate launches (same rounding points: 2e-3)
  * training variant (the default; NR_ENC_FUSED=0 selects the separate launches): loss / score / every gradient vs the separate launches, with and without
    dropout (both paths draw the same Philox stream), and vs the oracle at drop_rate 0."""
import os

import numpy as np
import pytest
import torch

import synth
from oracle import nr_oracle as O
from test_gpu_ops import T
from test_gpu_models import build_model

pytestmark = pytest.mark.gpu


@pytest.fixture
def fused_training():
    """fused training is the default; the test toggles NR_ENC_FUSED (read per call) to get the separate launches"""
    old = os.environ.get('NR_ENC_FUSED')
    os.environ['NR_ENC_FUSED'] = '1'
    yield
    if old is None:
        del os.environ['NR_ENC_FUSED']
    else:
        os.environ['NR_ENC_FUSED'] = old


def _titles(rs, n, L, V):
    ids = rs.randint(1, V, size=(n, L)).astype(np.int32)
    tl = rs.randint(1, L + 1, size=n)
    ids[np.arange(L)[None, :] >= tl[:, None]] = 0
    ids[0] = 0                                   # the all-pad news (B4)
    return ids


@pytest.mark.parametrize('n', [1, 3, 4, 5, 131, 1000])
def test_fused_inference_vs_oracle_and_separate_launches(n):
    kind, kw, Pn, bt, model = build_model('nrms_ulm0', precision='bf16')
    model.eval()
    rs = np.random.RandomState(n)
    ids = _titles(rs, n, kw['L'], kw['V'])
    x = T(ids).cuda()
    with torch.no_grad():
        v_fused = model.news_encoder(x)                       # no grad -> save == NULL -> fused kernel
    from newsrecommendation_b200 import _lib as L
    n0 = L.lib.nr_launch_count()
    with torch.no_grad():
        model.news_encoder(x)
    assert L.lib.nr_launch_count() - n0 == 3                  # pack + stage + ONE encoder kernel
    os.environ['NR_ENC_FUSED'] = '0'
    try:
        v_sep = model.news_encoder(x).detach()                # grad mode + switch -> saving path, separate launches
    finally:
        del os.environ['NR_ENC_FUSED']
    P = {k: T(v) for k, v in Pn.items()}
    v_or = O.nrms_news_encoder(T(ids), P, 20).numpy()
    scale = float(np.abs(v_or).max())
    np.testing.assert_allclose(v_fused.cpu().numpy(), v_or, rtol=1e-2, atol=1e-2 * scale)
    np.testing.assert_allclose(v_fused.cpu().numpy(), v_sep.cpu().numpy(), rtol=2e-3, atol=2e-3 * scale)
    assert torch.isfinite(v_fused).all()


def _run(model, bt, seed=None):
    hist, mask, cand, label = [T(a).cuda() for a in bt]
    if seed is not None:
        torch.manual_seed(seed)
        from newsrecommendation_b200 import ops
        import itertools
        ops._seed_counter = itertools.count(1)               # same dropout seeds for both runs
    model.zero_grad(set_to_none=True)
    loss, score = model(hist, mask, cand, label)
    loss.backward()
    return loss.item(), score.detach().cpu().numpy(), {k: p.grad.detach().cpu().numpy().copy()
                                                       for k, p in model.named_parameters() if p.grad is not None}


@pytest.mark.parametrize('tag,drop', [('nrms_ulm0', 0.0), ('nrms_ulm1', 0.0), ('nrms_ulm0', 0.2)])
def test_fused_training_vs_separate_launches(tag, drop, fused_training):
    kind, kw, Pn, bt, model = build_model(tag, precision='bf16', drop_rate=drop)
    model.train()
    l_f, s_f, g_f = _run(model, bt, seed=5)
    os.environ['NR_ENC_FUSED'] = '0'
    l_s, s_s, g_s = _run(model, bt, seed=5)
    os.environ['NR_ENC_FUSED'] = '1'
    assert abs(l_f - l_s) <= 2e-3 * abs(l_s)
    np.testing.assert_allclose(s_f, s_s, rtol=0, atol=3e-3 * float(np.abs(s_s).max()))
    assert g_f.keys() == g_s.keys()
    gmax = max(float(np.abs(g).max()) for g in g_s.values())
    for k in g_s:
        m = float(np.abs(g_s[k]).max())
        # tensors whose gradient is analytically zero (W_K.bias: the normaliser cancels a key shift) hold rounding
        # noise only: they are held to the noise floor of the step, not to their own scale
        np.testing.assert_allclose(g_f[k], g_s[k], rtol=0, atol=2e-2 * m + 1e-5 * gmax, err_msg=k)
        if m > 1e-3 * gmax:
            a, b = g_f[k].ravel().astype(np.float64), g_s[k].ravel().astype(np.float64)
            assert a @ b / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-30) > 0.999, k
    if drop == 0.0:
        P = {k: T(v) for k, v in Pn.items()}
        hist, mask, cand, label = [T(a) for a in bt]
        lo, so = O.nrms_forward(hist, mask, cand, label, P, 20, kw['user_log_mask'])
        assert abs(l_f - lo.item()) <= 1e-2 * abs(lo.item())
        np.testing.assert_allclose(s_f, so.numpy(), rtol=1e-2, atol=1e-2 * float(so.abs().max()))


@pytest.mark.parametrize('tag', ['naml_ulm0_c00', 'naml_ulm0_c11'])
def test_fused_conv_encoder_vs_oracle_and_separate_launches(tag):
    """NAML title path (gather -> Conv1d(k=3) -> AttentionPooling, NAML.py:47-56) through the fused kernel's CONV mode:
    inference and training variants vs the oracle and vs the separate conv / fc1 / pooling launches, incl. gradients."""
    kind, kw, Pn, bt, model = build_model(tag, precision='bf16')
    rs = np.random.RandomState(3)
    N = kw['N']
    F = 1 + int(kw['use_cat']) + int(kw['use_sub'])
    x = np.zeros((N + 1, F), np.int32)
    x[:, 0] = np.arange(N + 1)
    if kw['use_cat']:
        x[1:, 1] = rs.randint(1, synth.NCAT + 1, size=N)
    if kw['use_sub']:
        x[1:, F - 1] = rs.randint(1, synth.NSUB + 1, size=N)
    xd = T(x).cuda()
    P = {k: T(v) for k, v in Pn.items()}
    v_or = O.naml_news_encoder(T(x), P, kw['L'], synth.D, kw['use_cat'], kw['use_sub']).numpy()
    scale = float(np.abs(v_or).max())
    model.eval()
    with torch.no_grad():
        v_inf = model.news_encoder(xd)
    os.environ['NR_ENC_FUSED'] = '0'
    try:
        v_sep = model.news_encoder(xd).detach()
    finally:
        del os.environ['NR_ENC_FUSED']
    np.testing.assert_allclose(v_inf.cpu().numpy(), v_or, rtol=1e-2, atol=1e-2 * scale)
    np.testing.assert_allclose(v_inf.cpu().numpy(), v_sep.cpu().numpy(), rtol=2e-3, atol=2e-3 * scale)
    # training variant: whole model step, fused vs separate launches
    model.train()
    l_f, s_f, g_f = _run(model, bt, seed=7)
    os.environ['NR_ENC_FUSED'] = '0'
    try:
        l_s, s_s, g_s = _run(model, bt, seed=7)
    finally:
        del os.environ['NR_ENC_FUSED']
    assert abs(l_f - l_s) <= 2e-3 * abs(l_s)
    np.testing.assert_allclose(s_f, s_s, rtol=0, atol=3e-3 * float(np.abs(s_s).max()))
    gmax = max(float(np.abs(g).max()) for g in g_s.values())
    for k in g_s:
        m = float(np.abs(g_s[k]).max())
        np.testing.assert_allclose(g_f[k], g_s[k], rtol=0, atol=2e-2 * m + 1e-5 * gmax, err_msg=k)


@pytest.mark.parametrize('drop', [0.0, 0.2])
def test_dx_over_real_token_rows_vs_all_rows(drop):
    """The title-encoder backward keeps dqkv COMPACT (rows of real tokens only: attn_bwd writes them through a row map,
    enc.dx and enc.dwqkv contract over them, csrc/nr_encoder.cu); NR_DX_GATHER=0 (read per call) keeps every padded
    row.  Each dX row is the same dot product in both modes, so the embedding gradient (NRMS.py:73, padding_idx row
    stays zero) differs only by the order of the fp32 reductions into the table; the W_Q/W_K/W_V weight gradients lose
    only all-zero input rows.  Their BIAS gradients are summed inside attn_bwd from the fp32 fragments instead of from
    the bf16-rounded dqkv (the ones column of enc.dwqkv): equal to bf16 rounding (these sums cancel heavily: both routes sit
    3-4 % of max|g| from the fp32 oracle for W_V.bias, the compact one closer; profiles/r2/bias_check.py)."""
    kind, kw, Pn, bt, model = build_model('nrms_ulm0', precision='bf16', drop_rate=drop)
    model.train()
    ids = np.concatenate([bt[0].reshape(-1, kw['L']), bt[2].reshape(-1, kw['L'])])
    assert 0.05 < (ids == 0).mean() < 0.95                    # the case has pad tokens AND real ones
    old = os.environ.pop('NR_DX_GATHER', None)
    try:
        l_g, s_g, g_g = _run(model, bt, seed=7)
        os.environ['NR_DX_GATHER'] = '0'
        l_a, s_a, g_a = _run(model, bt, seed=7)
    finally:
        os.environ.pop('NR_DX_GATHER', None)
        if old is not None:
            os.environ['NR_DX_GATHER'] = old
    assert l_g == l_a and np.array_equal(s_g, s_a)            # the forward is untouched
    k = 'news_encoder.embedding_matrix.weight'
    gmax = float(np.abs(g_a[k]).max())
    assert gmax > 0 and not g_g[k][0].any() and not g_a[k][0].any()
    np.testing.assert_allclose(g_g[k], g_a[k], rtol=0, atol=2e-6 * gmax)
    touched = np.unique(ids[ids != 0])
    assert np.abs(g_g[k][touched]).sum(1).min() > 0           # every real token received its gradient
    untouched = np.setdiff1d(np.arange(g_g[k].shape[0]), touched)
    assert not g_g[k][untouched].any()
    gall = max(float(np.abs(v).max()) for v in g_a.values())
    for name in g_a:                                          # nothing else moves
        m = float(np.abs(g_a[name]).max())
        bias = name.endswith(('W_Q.bias', 'W_K.bias', 'W_V.bias')) and name.startswith('news_encoder')
        atol = 2.5e-2 * m + 1e-4 * gall if bias else 1e-5 * m + 1e-9       # W_K.bias is analytically zero: rounding noise
        np.testing.assert_allclose(g_g[name], g_a[name], rtol=0, atol=atol, err_msg=name)


@pytest.mark.parametrize('case', ['all_pad', 'no_pad', 'one_title', 'one_real_token', 'ragged_517', 'row0_nonzero'])
def test_compact_backward_edge_cases(case):
    """The compact title-encoder backward at the edges of its row list: no real token at all (count 0: enc.dx and
    enc.dwqkv schedule no tile, the bias gradients still come out of attn_bwd), no pad token, a single title (fewer
    items than the attention grid has warps), a single real token (one compact row, 63 zero-filled gather rows), and a
    ragged count that is a multiple of neither 64 nor 128 — each against the all-rows backward (NR_DX_GATHER=0).
    `row0_nonzero`: `from_pretrained(..., padding_idx=0)` (NRMS.py:71-73) keeps a non-zero pretrained row 0, so a pad
    token's input is NOT zero and the W_Q/W_K/W_V gradients need its rows: the row list then keeps every position.
    Titles that are all padding take the attention-backward shortcut (dQ = dK = 0, sum dV = sum dO) in the other cases."""
    kind, kw, Pn, bt, model = build_model('nrms_ulm0', precision='bf16', drop_rate=0.0)
    model.train()
    L, V = kw['L'], kw['V']
    rs = np.random.RandomState(11)
    n = {'all_pad': 37, 'no_pad': 37, 'one_title': 1, 'one_real_token': 9, 'ragged_517': 64, 'row0_nonzero': 21}[case]
    ids = rs.randint(1, V, size=(n, L)).astype(np.int32)
    if case == 'all_pad':
        ids[:] = 0
    elif case == 'one_real_token':
        ids[:] = 0
        ids[5, 3] = 7
    elif case == 'ragged_517':
        ids[:] = 0
        flat = ids.reshape(-1)
        flat[rs.permutation(flat.size)[:517]] = rs.randint(1, V, size=517)      # holes inside titles too
    elif case == 'one_title':
        ids[0, 20:] = 0
    elif case == 'row0_nonzero':
        ids[:, 9:] = 0
        ids[3] = 0                                               # an all-pad title that is NOT a shortcut case here
        with torch.no_grad():
            model.news_encoder.embedding_matrix.weight[0] = torch.from_numpy(rs.randn(300).astype(np.float32) * 0.1).cuda()
    x = T(ids).cuda()
    w = torch.from_numpy(rs.randn(n, 400).astype(np.float32)).cuda()

    def grads(gather):
        old = os.environ.pop('NR_DX_GATHER', None)
        if not gather:
            os.environ['NR_DX_GATHER'] = '0'
        try:
            model.zero_grad(set_to_none=True)
            out = model.news_encoder(x)
            (out * w).sum().backward()
            return out.detach().cpu().numpy(), {k: p.grad.detach().cpu().numpy().copy()
                                                for k, p in model.news_encoder.named_parameters() if p.grad is not None}
        finally:
            os.environ.pop('NR_DX_GATHER', None)
            if old is not None:
                os.environ['NR_DX_GATHER'] = old
    o_c, g_c = grads(True)
    o_a, g_a = grads(False)
    assert np.array_equal(o_c, o_a) and g_c.keys() == g_a.keys() and len(g_c) >= 10
    gall = max(float(np.abs(v).max()) for v in g_a.values())
    emb = 'embedding_matrix.weight'
    real = np.unique(ids[ids != 0])
    assert not g_c[emb][0].any()
    assert not np.delete(g_c[emb], real, axis=0).any()
    if real.size:
        assert np.abs(g_c[emb][real]).sum(1).min() > 0
    if case == 'row0_nonzero':      # pad rows feed dW: a backward that dropped them would be off by far more than 1e-5
        assert np.abs(g_c['multi_head_self_attn.W_V.weight']).max() > 0
    for k in g_a:
        m = float(np.abs(g_a[k]).max())
        bias = k.endswith(('W_Q.bias', 'W_K.bias', 'W_V.bias'))
        atol = 2.5e-2 * m + 1e-4 * gall if bias else 1e-5 * m + 1e-7 * gall
        np.testing.assert_allclose(g_c[k], g_a[k], rtol=0, atol=atol, err_msg=k)
    if case == 'all_pad':      # zero input rows: no weight gradient in either mode, the biases do get one
        assert not g_c['multi_head_self_attn.W_Q.weight'].any() and not g_c['multi_head_self_attn.W_V.weight'].any()
        assert np.abs(g_c['multi_head_self_attn.W_V.bias']).max() > 0
