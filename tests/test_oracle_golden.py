"""Pin the oracle (oracle/nr_oracle.py) against outputs of the unmodified
reference (fixtures written by tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest
import torch

import synth
from oracle import nr_oracle as O

G = os.path.join(os.path.dirname(__file__), 'golden')


def T(a):
    return torch.from_numpy(np.asarray(a))


def test_ops_pool_mhsa_conv():
    g = np.load(os.path.join(G, 'ops.npz'))
    # fp32 BLAS sums in a thread-count dependent order: outputs are O(1), so 5e-6 absolute is the same 1e-5 bar
    ATOL = 5e-6
    P = {}
    synth._pool(np.random.RandomState(1), 'p', P)
    x = T((np.random.RandomState(2).standard_normal((6, 30, 400)) * 0.5).astype(np.float32))
    m = T((np.random.RandomState(3).rand(6, 30) > 0.3).astype(np.float32)); m[0] = 0
    w = [T(P[k]) for k in ('p.att_fc1.weight', 'p.att_fc1.bias', 'p.att_fc2.weight', 'p.att_fc2.bias')]
    np.testing.assert_allclose(O.attention_pooling(x, *w).numpy(), g['pool_out_nomask'], rtol=1e-5, atol=ATOL)
    out = O.attention_pooling(x, *w, mask=m).numpy()
    np.testing.assert_allclose(out, g['pool_out_mask'], rtol=1e-5, atol=ATOL)
    assert (out[0] == 0).all()          # all-masked row -> exact zeros (B1)
    P = {}
    synth._mhsa(np.random.RandomState(4), 'm', 300, P)
    xs = T((np.random.RandomState(5).standard_normal((5, 30, 300)) * 0.4).astype(np.float32))
    ms = T((np.random.RandomState(6).rand(5, 30) > 0.3).astype(np.float32)); ms[1] = 0
    w = [T(P[f'm.{n}.{s}']) for n in ('W_Q', 'W_K', 'W_V') for s in ('weight', 'bias')]
    np.testing.assert_allclose(O.multi_head_self_attention(xs, *w, n_heads=20).numpy(),
                               g['mhsa_out_nomask'], rtol=1e-5, atol=ATOL)
    np.testing.assert_allclose(O.multi_head_self_attention(xs, *w, n_heads=20, mask=ms).numpy(),
                               g['mhsa_out_mask'], rtol=1e-5, atol=ATOL)
    r7 = np.random.RandomState(7)
    cw = synth._u(r7, (400, 300, 3), 0.033); cb = synth._u(r7, (400,), 0.033)
    xc = T((r7.standard_normal((4, 30, 300)) * 0.4).astype(np.float32))
    np.testing.assert_allclose(O.conv1d_k3(xc, T(cw), T(cb)).numpy(), g['conv_out'], rtol=1e-5, atol=ATOL)


def oracle_model(tag, P, bt):
    kind, kw = synth.MODEL_CASES[tag]
    hist, mask, cand, label = [T(a) for a in bt]
    if kind == 'nrms':
        return O.nrms_forward(hist, mask, cand, label, P, 20, kw['user_log_mask'])
    return O.naml_forward(hist, mask, cand, label, P, kw['L'], synth.D, kw['use_cat'], kw['use_sub'],
                          kw['user_log_mask'])


@pytest.mark.parametrize('tag', sorted(synth.MODEL_CASES))
def test_model_forward_backward(tag):
    g = np.load(os.path.join(G, tag + '.npz'))
    kind, kw, Pn, bt = synth.case_inputs(tag)
    assert list(g['keys']) == list(Pn.keys())      # state_dict key contract
    P = {k: T(v).clone().requires_grad_(True) for k, v in Pn.items()}
    loss, score = oracle_model(tag, P, bt)
    np.testing.assert_allclose(loss.item(), g['loss'], rtol=1e-5)
    np.testing.assert_allclose(score.detach().numpy(), g['score'], rtol=1e-5, atol=2e-5)
    loss.backward()
    for k in Pn:
        if 'gs:' + k not in g.files:
            continue
        gr = P[k].grad
        gr = torch.zeros_like(P[k]) if gr is None else gr
        s, st = synth.grad_digest(gr.numpy())
        scale = max(1e-12, float(np.abs(g['gs:' + k]).max()))
        np.testing.assert_allclose(s, g['gs:' + k], rtol=1e-4, atol=2e-5 * scale + 1e-6, err_msg=k)
        np.testing.assert_allclose(st[1], g['gt:' + k][1], rtol=1e-4, atol=1e-6, err_msg=k)
    if kind == 'nrms':     # padding_idx row never receives gradient (B7)
        assert (P['news_encoder.embedding_matrix.weight'].grad[0] == 0).all()
    # eval-mode sub-module parity
    hist, mask = T(bt[0]), T(bt[1])
    F = hist.shape[-1]
    with torch.no_grad():
        Pd = {k: v.detach() for k, v in P.items()}
        if kind == 'nrms':
            nv = O.nrms_news_encoder(hist.reshape(-1, F), Pd, 20)
            uv = O.nrms_user_encoder(nv.reshape(hist.shape[0], hist.shape[1], -1), mask, Pd, 20, kw['user_log_mask'])
            pv = O.nrms_news_encoder(torch.zeros(1, F, dtype=torch.int32), Pd, 20)
        else:
            enc = lambda x: O.naml_news_encoder(x, Pd, kw['L'], synth.D, kw['use_cat'], kw['use_sub'])
            nv = enc(hist.reshape(-1, F))
            uv = O.naml_user_encoder(nv.reshape(hist.shape[0], hist.shape[1], -1), mask, Pd, kw['user_log_mask'])
            pv = enc(torch.zeros(1, F, dtype=torch.int32))
    np.testing.assert_allclose(nv.numpy(), g['hist_vecs'], rtol=1e-5, atol=5e-6)
    np.testing.assert_allclose(uv.numpy(), g['user_vec'], rtol=1e-5, atol=5e-6)
    np.testing.assert_allclose(pv.numpy(), g['pad_news_vec'], rtol=1e-5, atol=5e-6)
    assert np.abs(pv.numpy()).max() > 0            # pad news vector is non-zero (B4)
    if kw['user_log_mask']:
        assert (uv.numpy()[0] == 0).all()          # empty history -> exact zero (B1)
    else:
        assert np.abs(uv.numpy()[0]).max() > 0     # pad_doc-only user vector (B3)


def test_metrics_vs_reference():
    g = np.load(os.path.join(G, 'metrics.npz'))
    off = g['offsets']
    for i in range(len(off) - 1):
        l = g['labels'][off[i]:off[i + 1]]; s = g['scores'][off[i]:off[i + 1]]
        r = O.impression_metrics(l, s)
        np.testing.assert_allclose(r, g['expected'][i], rtol=0, atol=1e-12)
    off = g['tied_offsets']
    for i in range(len(off) - 1):
        l = g['tied_labels'][off[i]:off[i + 1]]; s = g['tied_scores'][off[i]:off[i + 1]]
        assert abs(O.auc_score(l, s) - g['tied_auc'][i]) < 1e-12
    assert O.impression_metrics(np.zeros(4, np.int8), np.ones(4, np.float32)) is None
    assert O.impression_metrics(np.ones(4, np.int8), np.ones(4, np.float32)) is None


def test_adam_vs_torch():
    g = np.load(os.path.join(G, 'adam.npz'))
    ra = np.random.RandomState(8)
    p = T(ra.standard_normal(1000).astype(np.float32)).clone()
    m = torch.zeros_like(p); v = torch.zeros_like(p)
    for s in range(3):
        gr = T(ra.standard_normal(1000).astype(np.float32))
        O.adam_step(p, gr, m, v, s + 1, 3e-4)
        np.testing.assert_allclose(p.numpy(), g[f'p{s + 1}'], rtol=1e-6, atol=1e-7)


def test_pad_history():
    idx, m = O.pad_history([5, 6, 7], 5)
    assert idx.tolist() == [0, 0, 5, 6, 7] and m.tolist() == [0, 0, 1, 1, 1]
    idx, m = O.pad_history(list(range(1, 9)), 5)
    assert idx.tolist() == [4, 5, 6, 7, 8] and m.tolist() == [1] * 5


def test_oracle_ref_recipe_is_the_unmodified_reference():
    """oracle/make_ref.py: the files under oracle/_ref (what bench.py times as the CPU baseline, `kind: "reference"`) are
    byte-identical copies of the reference's modules (checked against /root/reference when it is present, against the
    recipe's own manifest otherwise), and the loaded modules agree with the oracle port on a golden case."""
    import hashlib
    import json
    import os
    from oracle import make_ref
    if not make_ref.available():
        if not os.path.isdir(make_ref.REF_SRC):
            pytest.skip('oracle/_ref absent and /root/reference not present')
        assert make_ref.make(verbose=False)
    man = json.load(open(os.path.join(make_ref.OUT, 'MANIFEST.json')))['sha256']
    for rel in make_ref.FILES:
        h = hashlib.sha256(open(os.path.join(make_ref.OUT, rel), 'rb').read()).hexdigest()
        assert h == man[rel], rel
        src = os.path.join(make_ref.REF_SRC, rel)
        if os.path.exists(src):
            assert h == hashlib.sha256(open(src, 'rb').read()).hexdigest(), rel
    NRMS, NAML, M = make_ref.load()
    import argparse
    kind, kw, P, bt = synth.case_inputs('nrms_ulm1')
    a = argparse.Namespace(drop_rate=0.0, news_dim=400, num_attention_heads=20, word_embedding_dim=300,
                           news_query_vector_dim=200, user_query_vector_dim=200, user_log_mask=True, user_log_length=50,
                           num_words_title=kw['L'], npratio=4, freeze_embedding=False)
    m = NRMS.Model(a, P['news_encoder.embedding_matrix.weight'])
    m.load_state_dict({k: torch.from_numpy(v) for k, v in P.items()})
    m.eval()
    hist, mask, cand, label = [torch.from_numpy(x) for x in bt]
    with torch.no_grad():
        loss, score = m(hist, mask, cand, label)
    lo, so = O.nrms_forward(hist, mask, cand, label, {k: torch.from_numpy(v) for k, v in P.items()}, 20, True)
    np.testing.assert_allclose(score.numpy(), so.numpy(), rtol=1e-5, atol=5e-6)
    np.testing.assert_allclose(loss.item(), lo.item(), rtol=1e-5)
