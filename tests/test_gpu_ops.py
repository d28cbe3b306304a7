"""GPU parity, op level: libnr_sm100.so (through the C ABI / ops shims) vs the CPU oracle
and the reference-generated golden fixtures.  fp32 mode: rel 1e-5 (BASELINE north_star)."""
import os

import numpy as np
import pytest
import torch

import synth
from oracle import nr_oracle as O

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), 'golden')
RTOL, ATOL = 1e-5, 3e-6


def T(a):
    return torch.from_numpy(np.asarray(a))


def close(a, b, rtol=RTOL, atol=ATOL, msg=''):
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    b = b.detach().cpu().numpy() if torch.is_tensor(b) else np.asarray(b)
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol, err_msg=msg)


def gclose(a, b, msg=''):
    """gradients: tolerance relative to the tensor's own scale"""
    b = b.detach().cpu().numpy()
    sc = max(1e-12, float(np.abs(b).max()))
    close(a, b, rtol=1e-4, atol=2e-5 * sc + 1e-6, msg=msg)


@pytest.fixture(scope='module')
def ops():
    from newsrecommendation_b200 import ops as o
    return o


def pool_w(seed):
    P = {}
    synth._pool(np.random.RandomState(seed), 'p', P)
    return [T(P[k]) for k in ('p.att_fc1.weight', 'p.att_fc1.bias', 'p.att_fc2.weight', 'p.att_fc2.bias')]


def mhsa_w(seed, dm):
    P = {}
    synth._mhsa(np.random.RandomState(seed), 'm', dm, P)
    return [T(P[f'm.{n}.{s}']) for n in ('W_Q', 'W_K', 'W_V') for s in ('weight', 'bias')]


def test_gather_bit_exact(ops):
    rs = np.random.RandomState(0)
    for rows, width in ((97, 300), (41, 9000), (5, 4)):
        table = T(rs.standard_normal((rows, width)).astype(np.float32))
        ids = T(rs.randint(0, rows, size=(7, 13)).astype(np.int32))
        out = ops.gather_rows(table.cuda(), ids.cuda())
        assert torch.equal(out.cpu(), table[ids.long()])
    assert ops.gather_rows(table.cuda(), torch.zeros(0, dtype=torch.int32, device='cuda')).shape == (0, 4)


@pytest.mark.parametrize('masked', [False, True])
def test_attention_pooling_golden_and_grads(ops, masked):
    g = np.load(os.path.join(G, 'ops.npz'))
    w = pool_w(1)
    x = T((np.random.RandomState(2).standard_normal((6, 30, 400)) * 0.5).astype(np.float32))
    m = T((np.random.RandomState(3).rand(6, 30) > 0.3).astype(np.float32)); m[0] = 0
    mask = m if masked else None
    xd = x.cuda().requires_grad_(True)
    wd = [t.cuda().requires_grad_(True) for t in w]
    out = ops.attention_pooling(xd, None if mask is None else mask.cuda(), *wd)
    close(out, g['pool_out_mask' if masked else 'pool_out_nomask'])
    if masked:
        assert (out[0] == 0).all()                  # all-masked row -> exact zeros (B1)
    xo = x.clone().requires_grad_(True)
    wo = [t.clone().requires_grad_(True) for t in w]
    ref = O.attention_pooling(xo, *wo, mask=mask)
    close(out, ref)
    gup = T(np.random.RandomState(9).standard_normal((6, 400)).astype(np.float32))
    out.backward(gup.cuda())
    ref.backward(gup)
    gclose(xd.grad, xo.grad, 'dx')
    for a, b, n in zip(wd, wo, ('dw1', 'db1', 'dw2', 'db2')):
        gclose(a.grad, b.grad, n)


@pytest.mark.parametrize('S,H,Q,n', [(3, 400, 200, 11), (50, 400, 200, 4), (1, 8, 4, 2), (30, 64, 32, 130)])
def test_attention_pooling_shapes(ops, S, H, Q, n):
    rs = np.random.RandomState(S + n)
    x = T((rs.standard_normal((n, S, H)) * 0.5).astype(np.float32))
    w = [T(rs.uniform(-0.1, 0.1, sh).astype(np.float32)) for sh in ((Q, H), (Q,), (1, Q), (1,))]
    out = ops.attention_pooling(x.cuda(), None, *[t.cuda() for t in w])
    close(out, O.attention_pooling(x, *w))


@pytest.mark.parametrize('masked', [False, True])
def test_mhsa_dense_golden_and_grads(ops, masked):
    g = np.load(os.path.join(G, 'ops.npz'))
    w = mhsa_w(4, 300)
    xs = T((np.random.RandomState(5).standard_normal((5, 30, 300)) * 0.4).astype(np.float32))
    ms = T((np.random.RandomState(6).rand(5, 30) > 0.3).astype(np.float32)); ms[1] = 0
    mask = ms if masked else None
    xd = xs.cuda().requires_grad_(True)
    wd = [t.cuda().requires_grad_(True) for t in w]
    out = ops.mhsa(xd, None, None if mask is None else mask.cuda(), *wd, mode=0, S=30, Dm=300, heads=20)
    close(out, g['mhsa_out_mask' if masked else 'mhsa_out_nomask'])
    xo = xs.clone().requires_grad_(True)
    wo = [t.clone().requires_grad_(True) for t in w]
    ref = O.multi_head_self_attention(xo, *wo, n_heads=20, mask=mask)
    gup = T(np.random.RandomState(10).standard_normal((5, 30, 400)).astype(np.float32))
    out.backward(gup.cuda())
    ref.backward(gup)
    gclose(xd.grad, xo.grad, 'dx')
    for a, b, n in zip(wd, wo, ('wq', 'bq', 'wk', 'bk', 'wv', 'bv')):
        gclose(a.grad, b.grad, n)


def test_mhsa_user_shape_T50(ops):
    w = mhsa_w(11, 400)
    x = T((np.random.RandomState(12).standard_normal((3, 50, 400)) * 0.3).astype(np.float32))
    m = T((np.random.RandomState(13).rand(3, 50) > 0.5).astype(np.float32))
    out = ops.mhsa(x.cuda(), None, m.cuda(), *[t.cuda() for t in w], mode=0, S=50, Dm=400, heads=20)
    close(out, O.multi_head_self_attention(x, *w, n_heads=20, mask=m))


def test_mhsa_token_rows_with_dropout_replay(ops):
    """gather + input dropout + output dropout fused; the kernel's own keep masks are
    exported (nr_dropout_keep_mask) and replayed through the oracle -> exact train-mode parity."""
    from newsrecommendation_b200 import _lib as L
    rs = np.random.RandomState(21)
    V, n, S, D, H, p, seed = 53, 9, 30, 300, 400, 0.2, 1234567
    table = T((rs.standard_normal((V, D)) * 0.4).astype(np.float32)); table[0] = 0
    ids = T(rs.randint(0, V, size=(n, S)).astype(np.int32)); ids[:, 20:] = 0
    w = mhsa_w(22, D)
    td = table.cuda().requires_grad_(True)
    wd = [t.cuda().requires_grad_(True) for t in w]
    out = ops.mhsa(td, ids.cuda(), None, *wd, mode=L.ROWS_TOKEN, S=S, Dm=D, heads=20, in_drop=p, out_drop=p,
                   seed=seed)
    k1 = ops.dropout_keep_mask(n * S * D, p, seed, L.STREAM_IN, 'cuda', D).cpu().view(n, S, D)
    k2 = ops.dropout_keep_mask(n * S * H, p, seed, L.STREAM_OUT, 'cuda', H, H // 20).cpu().view(n, S, H)
    assert 0.75 < k1.float().mean() < 0.85 and 0.75 < k2.float().mean() < 0.85
    to = table.clone().requires_grad_(True)
    wo = [t.clone().requires_grad_(True) for t in w]
    x = O.dropout(O.embedding_lookup(to, ids), k1, p)
    ref = O.dropout(O.multi_head_self_attention(x, *wo, n_heads=20), k2, p)
    close(out, ref)
    gup = T(rs.standard_normal((n, S, H)).astype(np.float32))
    out.backward(gup.cuda())
    ref.backward(gup)
    gclose(td.grad, to.grad, 'dtable')
    assert (td.grad[0] == 0).all()                  # padding_idx row: no gradient (B7)
    for a, b, nm in zip(wd, wo, ('wq', 'bq', 'wk', 'bk', 'wv', 'bv')):
        gclose(a.grad, b.grad, nm)


def test_conv_golden_and_grads(ops):
    g = np.load(os.path.join(G, 'ops.npz'))
    r7 = np.random.RandomState(7)
    cw = T(synth._u(r7, (400, 300, 3), 0.033)); cb = T(synth._u(r7, (400,), 0.033))
    xc = T((r7.standard_normal((4, 30, 300)) * 0.4).astype(np.float32))
    xd = xc.cuda().requires_grad_(True)
    wd, bd = cw.cuda().requires_grad_(True), cb.cuda().requires_grad_(True)
    y = ops.conv1d_k3(xd, None, wd, bd, mode=0, L_=30, D=300)
    close(y, g['conv_out'])
    xo, wo, bo = xc.clone().requires_grad_(True), cw.clone().requires_grad_(True), cb.clone().requires_grad_(True)
    ref = O.conv1d_k3(xo, wo, bo)
    gup = T(r7.standard_normal((4, 30, 400)).astype(np.float32))
    y.backward(gup.cuda()); ref.backward(gup)
    gclose(xd.grad, xo.grad, 'dx'); gclose(wd.grad, wo.grad, 'dw'); gclose(bd.grad, bo.grad, 'db')


def test_conv_news_rows_dropout_and_table_grad(ops):
    from newsrecommendation_b200 import _lib as L
    rs = np.random.RandomState(31)
    N, n, Lw, D, H, p, seed = 17, 6, 20, 300, 400, 0.2, 99
    table = T((rs.standard_normal((N, Lw * D)) * 0.4).astype(np.float32)); table[0] = 0
    ids = T(rs.randint(0, N, size=(n, 3)).astype(np.int32)); ids[0, 0] = 0
    cw = T(synth._u(rs, (H, D, 3), 0.033)); cb = T(synth._u(rs, (H,), 0.033))
    td = table.cuda().requires_grad_(True)
    wd, bd = cw.cuda().requires_grad_(True), cb.cuda().requires_grad_(True)
    y = ops.conv1d_k3(td, ids.cuda(), wd, bd, mode=L.ROWS_NEWS, L_=Lw, D=D, in_drop=p, seed=seed)
    k1 = ops.dropout_keep_mask(n * Lw * D, p, seed, L.STREAM_IN, 'cuda', D).cpu().view(n, Lw, D)
    to, wo, bo = table.clone().requires_grad_(True), cw.clone().requires_grad_(True), cb.clone().requires_grad_(True)
    x = O.dropout(O.embedding_lookup(to, ids[:, 0]).reshape(n, Lw, D), k1, p)
    ref = O.conv1d_k3(x, wo, bo)
    close(y, ref)
    gup = T(rs.standard_normal((n, Lw, H)).astype(np.float32))
    y.backward(gup.cuda()); ref.backward(gup)
    gclose(td.grad, to.grad, 'dtable'); gclose(wd.grad, wo.grad, 'dw'); gclose(bd.grad, bo.grad, 'db')
    assert (td.grad[0] == 0).all()


def test_out_of_range_ids_read_the_padding_row():
    """nr_rows_t.table_rows is enforced on the device.  An id outside [0, table_rows) behaves like id 0 (the
    padding row: zeros in, no gradient out) in the gather, the encoders (forward and table gradient) and the eval
    scoring — no out-of-bounds read or atomic."""
    from newsrecommendation_b200 import _lib as L, ops
    rs = np.random.RandomState(5)
    V, D, n, S = 50, 300, 6, 30
    table = rs.standard_normal((V, D)).astype(np.float32); table[0] = 0
    ids = rs.randint(1, V, size=(n, S)).astype(np.int32)
    bad = ids.copy(); bad[1, 3] = V + 7; bad[2, 0] = -3; bad[4, 29] = 2 ** 30
    ref = ids.copy(); ref[1, 3] = 0; ref[2, 0] = 0; ref[4, 29] = 0
    tb = T(table).cuda()
    assert torch.equal(ops.gather_rows(tb, T(bad).cuda()), ops.gather_rows(tb, T(ref).cuda()))
    w = [T((rs.standard_normal(s) * 0.05).astype(np.float32)).cuda() for s in ((400, D), (400,), (400, D), (400,), (400, D), (400,))]
    pool = [T((rs.standard_normal(s) * 0.05).astype(np.float32)).cuda() for s in ((200, 400), (200,), (1, 200), (1,))]
    outs = []
    for x in (bad, ref):
        t = tb.clone().requires_grad_(True)
        y = ops.title_encoder(L.ENC_MHSA, t, T(x).cuda(), None, w, pool, mode=L.ROWS_TOKEN, S=S, Dm=D, heads=20)
        y.square().sum().backward()
        outs.append((y.detach(), t.grad.detach()))
    assert torch.equal(outs[0][0], outs[1][0])
    close(outs[0][1], outs[1][1].cpu().numpy(), rtol=1e-4, atol=1e-5 * float(outs[1][1].abs().max()))   # atomics: order differs
    assert (outs[0][1][0] == 0).all()
    # eval scoring
    cache = T(rs.standard_normal((V, 400)).astype(np.float32)).cuda()
    user = T(rs.standard_normal((2, 400)).astype(np.float32)).cuda()
    off = T(np.array([0, 3, 7], np.int64)).cuda()
    sc = []
    for c in ([1, V + 1, 4, 5, -1, 6, 7], [1, 0, 4, 5, 0, 6, 7]):
        cand = T(np.array(c, np.int32)).cuda()
        s = torch.empty(7, dtype=torch.float32, device='cuda')
        L.call('nr_eval_score', L.ptr(cache), V, L.ptr(user), L.ptr(off), L.ptr(cand), L.ptr(s), 2, 7, 400, L.stream())
        sc.append(s)
    assert torch.equal(sc[0], sc[1])


def test_dropout_stream_statistics():
    """16-bit-threshold Philox streams (one call = 8 elements): keep rate, serial correlations at the lags the layout
    could expose (inside a Philox call: 1..7; across calls: 8, 16; across rows / heads: 300, 304, 400, 24) and
    independence between the input and output streams and between seeds.  5-sigma bounds on 4 M elements."""
    from newsrecommendation_b200 import _lib as L, ops
    N, p = 1 << 22, 0.2
    a = ops.dropout_keep_mask(N, p, 1234, L.STREAM_IN, 'cuda', 304).float()
    b = ops.dropout_keep_mask(N, p, 1234, L.STREAM_OUT, 'cuda', 400, 20).float()
    c = ops.dropout_keep_mask(N, p, 1235, L.STREAM_IN, 'cuda', 304).float()
    d = ops.dropout_keep_mask(N, p, 1234, L.STREAM_IN, 'cuda', 300).float()     # the real MIND width: rows padded to 304 in stream space
    sig = (0.8 * 0.2 / N) ** 0.5
    for m in (a, b, c, d):
        assert abs(m.mean().item() - (1 - 13107 / 65536)) < 5 * sig          # thr16 = floor(0.2 * 65536)
    def corr(x, y):
        x, y = x - x.mean(), y - y.mean()
        return (x * y).mean().item() / (x.std().item() * y.std().item())
    lim = 5 / N ** 0.5
    for m in (a, b, d):
        for lag in (1, 2, 3, 4, 5, 6, 7, 8, 16, 24, 300, 304, 400):
            assert abs(corr(m[:-lag], m[lag:])) < lim, lag
    assert abs(corr(a, b)) < lim and abs(corr(a, c)) < lim
    # 8-element groups: the 256 keep patterns of a group follow the product law (chi-square over patterns by popcount)
    g = a.view(-1, 8).sum(1).long()
    cnt = torch.bincount(g, minlength=9).double().cpu().numpy()
    from math import comb
    q = 1 - 13107 / 65536
    exp = np.array([comb(8, k) * q ** k * (1 - q) ** (8 - k) for k in range(9)]) * (N // 8)
    chi2 = float(((cnt - exp) ** 2 / exp).sum())
    assert chi2 < 40.0, chi2                                                  # 8 dof: P(chi2 > 40) ~ 3e-6
