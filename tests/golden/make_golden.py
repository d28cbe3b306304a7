"""Generate golden fixtures by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
Writes small .npz fixtures next to this file.  The reference modules are
imported from /root/reference/src (never copied).  Inputs and weights come
from tests/golden/synth.py (frozen numpy RandomState streams), so the
fixtures hold only the reference's OUTPUTS: loss, score, news/user vectors
and gradient digests.
"""
import argparse
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import synth  # noqa: E402

REF = '/root/reference/src'


def ref_args(**kw):
    a = argparse.Namespace(
        drop_rate=0.0, news_dim=400, num_attention_heads=20, word_embedding_dim=300,
        news_query_vector_dim=200, user_query_vector_dim=200, user_log_mask=False,
        user_log_length=50, num_words_title=30, npratio=4, freeze_embedding=False,
        use_category=False, use_subcategory=False, category_emb_dim=100)
    for k, v in kw.items():
        setattr(a, k, v)
    return a


def run_case(tag, NRMS, NAML):
    kind, kw, P, (hist, mask, cand, label) = synth.case_inputs(tag)
    if kind == 'nrms':
        a = ref_args(user_log_mask=kw['user_log_mask'], num_words_title=kw['L'])
        model = NRMS.Model(a, P['news_encoder.embedding_matrix.weight'])
    else:
        a = ref_args(user_log_mask=kw['user_log_mask'], num_words_title=kw['L'],
                     use_category=kw['use_cat'], use_subcategory=kw['use_sub'],
                     freeze_embedding=True)
        model = NAML.Model(a, P['news_encoder.title_embeddings.weight'], synth.NCAT, synth.NSUB)
    sd = model.state_dict()
    assert list(sd.keys()) == list(P.keys()), (list(sd.keys()), list(P.keys()))
    model.load_state_dict({k: torch.from_numpy(v) for k, v in P.items()})
    model.train()   # drop_rate = 0 -> deterministic
    loss, score = model(torch.from_numpy(hist), torch.from_numpy(mask),
                        torch.from_numpy(cand), torch.from_numpy(label))
    loss.backward()
    d = {'loss': loss.detach().numpy(), 'score': score.detach().numpy(),
         'keys': np.array(list(sd.keys()))}
    for k, p in model.named_parameters():
        if p.grad is None:
            continue
        s, st = synth.grad_digest(p.grad.numpy())
        d['gs:' + k] = s
        d['gt:' + k] = st
    model.eval()
    with torch.no_grad():
        F = hist.shape[-1]
        nv = model.news_encoder(torch.from_numpy(hist.reshape(-1, F)))
        d['hist_vecs'] = nv.numpy()
        d['user_vec'] = model.user_encoder(nv.reshape(hist.shape[0], hist.shape[1], -1),
                                           torch.from_numpy(mask)).numpy()
        d['pad_news_vec'] = model.news_encoder(torch.zeros(1, F, dtype=torch.int32)).numpy()
    np.savez_compressed(os.path.join(HERE, tag + '.npz'), **d)


def main():
    sys.path.insert(0, REF)
    mu = importlib.import_module('model.model_utils')
    NRMS = importlib.import_module('model.NRMS')
    NAML = importlib.import_module('model.NAML')
    metrics = importlib.import_module('metrics')
    rs = np.random.RandomState(0)

    # ---- op level (weights from synth helpers; outputs only) -------------------
    ops = {}
    P = {}
    synth._pool(np.random.RandomState(1), 'p', P)
    pool = mu.AttentionPooling(400, 200)
    pool.load_state_dict({k[2:]: torch.from_numpy(v) for k, v in P.items()})
    x = torch.from_numpy((np.random.RandomState(2).standard_normal((6, 30, 400)) * 0.5).astype(np.float32))
    m = torch.from_numpy((np.random.RandomState(3).rand(6, 30) > 0.3).astype(np.float32)); m[0] = 0
    ops['pool_out_nomask'] = pool(x).detach().numpy()
    ops['pool_out_mask'] = pool(x, m).detach().numpy()
    P = {}
    synth._mhsa(np.random.RandomState(4), 'm', 300, P)
    mh = mu.MultiHeadSelfAttention(300, 20, 20, 20)
    mh.load_state_dict({k[2:]: torch.from_numpy(v) for k, v in P.items()})
    xs = torch.from_numpy((np.random.RandomState(5).standard_normal((5, 30, 300)) * 0.4).astype(np.float32))
    ms = torch.from_numpy((np.random.RandomState(6).rand(5, 30) > 0.3).astype(np.float32)); ms[1] = 0
    ops['mhsa_out_nomask'] = mh(xs, xs, xs).detach().numpy()
    ops['mhsa_out_mask'] = mh(xs, xs, xs, ms).detach().numpy()
    r7 = np.random.RandomState(7)
    conv = torch.nn.Conv1d(300, 400, 3, padding=1)
    cw = synth._u(r7, (400, 300, 3), 0.033); cb = synth._u(r7, (400,), 0.033)
    conv.load_state_dict({'weight': torch.from_numpy(cw), 'bias': torch.from_numpy(cb)})
    xc = torch.from_numpy((r7.standard_normal((4, 30, 300)) * 0.4).astype(np.float32))
    ops['conv_out'] = conv(xc.transpose(1, 2)).transpose(1, 2).detach().numpy()
    np.savez_compressed(os.path.join(HERE, 'ops.npz'), **ops)

    # ---- model level -----------------------------------------------------------
    for tag in synth.MODEL_CASES:
        run_case(tag, NRMS, NAML)

    # ---- ranking metrics: reference metrics.py (+ sklearn AUC), incl. ties -----
    mets = {}
    cases = []
    for i in range(64):
        c = int(rs.randint(2, 60))
        lab = (rs.rand(c) < 0.2).astype(np.int64)
        if lab.sum() == 0:
            lab[int(rs.randint(0, c))] = 1
        if lab.sum() == c:
            lab[0] = 0
        cases.append((lab, rs.standard_normal(c).astype(np.float32)))
    off = np.zeros(len(cases) + 1, np.int64)
    off[1:] = np.cumsum([len(c[0]) for c in cases])
    mets['offsets'] = off
    mets['labels'] = np.concatenate([c[0] for c in cases]).astype(np.int8)
    mets['scores'] = np.concatenate([c[1] for c in cases])
    mets['expected'] = np.array([[metrics.roc_auc_score(l, s), metrics.mrr_score(l, s),
                                  metrics.ndcg_score(l, s, k=5), metrics.ndcg_score(l, s, k=10)]
                                 for l, s in cases], dtype=np.float64)
    tl, ts, ta = [], [], []
    for i in range(32):   # tied scores: only AUC is order-independent
        c = int(rs.randint(2, 40))
        lab = (rs.rand(c) < 0.3).astype(np.int64); lab[0] = 1; lab[-1] = 0
        sc = rs.randint(0, 4, size=c).astype(np.float32)
        tl.append(lab); ts.append(sc); ta.append(metrics.roc_auc_score(lab, sc))
    toff = np.zeros(len(tl) + 1, np.int64)
    toff[1:] = np.cumsum([len(l) for l in tl])
    mets['tied_offsets'] = toff
    mets['tied_labels'] = np.concatenate(tl).astype(np.int8)
    mets['tied_scores'] = np.concatenate(ts)
    mets['tied_auc'] = np.array(ta, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, 'metrics.npz'), **mets)

    # ---- Adam: torch.optim.Adam, 3 steps ---------------------------------------
    ra = np.random.RandomState(8)
    p = torch.nn.Parameter(torch.from_numpy(ra.standard_normal(1000).astype(np.float32)))
    opt = torch.optim.Adam([p], lr=3e-4)
    ad = {}
    for s in range(3):
        p.grad = torch.from_numpy(ra.standard_normal(1000).astype(np.float32))
        opt.step()
        ad[f'p{s + 1}'] = p.detach().numpy().copy()
    np.savez_compressed(os.path.join(HERE, 'adam.npz'), **ad)
    tot = 0
    for f in sorted(os.listdir(HERE)):
        if f.endswith('.npz'):
            tot += os.path.getsize(os.path.join(HERE, f))
            print(f, os.path.getsize(os.path.join(HERE, f)))
    print('total', tot)


if __name__ == '__main__':
    main()
