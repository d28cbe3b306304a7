"""Deterministic small-case inputs and weights shared by make_golden.py (which
feeds them to the imported reference) and by the tests (which feed the same
arrays to the oracle and to the CUDA path).  Uses numpy's legacy RandomState,
whose streams are frozen, so fixtures only need to hold OUTPUTS."""
import numpy as np

H, Q, HEADS, D, CATD = 400, 200, 20, 300, 100
NCAT, NSUB = 17, 264


def _u(rs, shape, bound):
    return rs.uniform(-bound, bound, size=shape).astype(np.float32)


def _pool(rs, prefix, P):
    P[prefix + '.att_fc1.weight'] = _u(rs, (Q, H), 0.08)
    P[prefix + '.att_fc1.bias'] = _u(rs, (Q,), 0.05)
    P[prefix + '.att_fc2.weight'] = _u(rs, (1, Q), 0.12)
    P[prefix + '.att_fc2.bias'] = _u(rs, (1,), 0.07)


def _mhsa(rs, prefix, d_model, P):
    b = float(np.sqrt(6.0 / (d_model + H)))
    for n in ('W_Q', 'W_K', 'W_V'):
        P[f'{prefix}.{n}.weight'] = _u(rs, (H, d_model), b)
        P[f'{prefix}.{n}.bias'] = _u(rs, (H,), 0.05)


def nrms_weights(seed, V):
    """state_dict (22 keys) for NRMS with a (V,300) word table."""
    rs = np.random.RandomState(seed)
    P = {}
    emb = (rs.standard_normal((V, D)) * 0.4).astype(np.float32)
    emb[0] = 0
    P['news_encoder.embedding_matrix.weight'] = emb
    _mhsa(rs, 'news_encoder.multi_head_self_attn', D, P)
    _pool(rs, 'news_encoder.attn', P)
    P['user_encoder.pad_doc'] = _u(rs, (1, H), 1.0)
    _mhsa(rs, 'user_encoder.multi_head_self_attn', H, P)
    _pool(rs, 'user_encoder.attn', P)
    return P


def naml_weights(seed, N, L, use_cat, use_sub):
    """state_dict for NAML in the reference's registration order (NAML.py:16-33)."""
    rs = np.random.RandomState(seed)
    P = {}
    t = (rs.standard_normal((N + 1, L * D)) * 0.4).astype(np.float32)
    t[0] = 0
    P['news_encoder.title_embeddings.weight'] = t
    if use_cat:
        e = (rs.standard_normal((NCAT + 1, CATD))).astype(np.float32); e[0] = 0
        P['news_encoder.category_emb.weight'] = e
        P['news_encoder.category_dense.weight'] = _u(rs, (H, CATD), 0.1)
        P['news_encoder.category_dense.bias'] = _u(rs, (H,), 0.1)
    if use_sub:
        e = (rs.standard_normal((NSUB + 1, CATD))).astype(np.float32); e[0] = 0
        P['news_encoder.subcategory_emb.weight'] = e
        P['news_encoder.subcategory_dense.weight'] = _u(rs, (H, CATD), 0.1)
        P['news_encoder.subcategory_dense.bias'] = _u(rs, (H,), 0.1)
    if use_cat or use_sub:
        _pool(rs, 'news_encoder.final_attn', P)
    P['news_encoder.cnn.weight'] = _u(rs, (H, D, 3), 0.033)
    P['news_encoder.cnn.bias'] = _u(rs, (H,), 0.033)
    _pool(rs, 'news_encoder.attn', P)
    P['user_encoder.pad_doc'] = _u(rs, (1, H), 1.0)
    _pool(rs, 'user_encoder.attn', P)
    return P


def batch(seed, B, T, C, L, n_rows, F=None):
    """history / mask / candidate / label with the reference's dtypes.
    F=None: NRMS token ids (.., L) with tail title padding (id 0).
    F in 1..3: NAML columns [news idx, category, subcategory]."""
    rs = np.random.RandomState(seed)
    if F is None:
        hist = rs.randint(1, n_rows, size=(B, T, L)).astype(np.int32)
        cand = rs.randint(1, n_rows, size=(B, C, L)).astype(np.int32)
        tl = rs.randint(3, L + 1, size=(B, T))
        cl = rs.randint(3, L + 1, size=(B, C))
        ar = np.arange(L)
        hist[ar[None, None, :] >= tl[..., None]] = 0
        cand[ar[None, None, :] >= cl[..., None]] = 0
    else:
        hist = np.zeros((B, T, F), np.int32)
        cand = np.zeros((B, C, F), np.int32)
        hist[..., 0] = rs.randint(1, n_rows, size=(B, T))
        cand[..., 0] = rs.randint(1, n_rows, size=(B, C))
        if F > 1:
            hist[..., 1] = rs.randint(1, NCAT + 1, size=(B, T))
            cand[..., 1] = rs.randint(1, NCAT + 1, size=(B, C))
        if F > 2:
            hist[..., 2] = rs.randint(1, NSUB + 1, size=(B, T))
            cand[..., 2] = rs.randint(1, NSUB + 1, size=(B, C))
    mask = np.ones((B, T), np.float32)
    for b in range(B):
        k = 0 if b == 0 else int(rs.randint(1, T + 1))   # user 0 has an empty history (B1/B3)
        mask[b, :T - k] = 0
        hist[b, :T - k] = 0                               # front padding (dataset.py:17-24)
    label = rs.randint(0, C, size=(B,)).astype(np.int64)
    return hist, mask, cand, label


def grad_digest(g):
    """Compact, order-sensitive summary of a gradient tensor: <=4096 strided
    samples + sum + L2 norm (full tensor when small)."""
    g = np.asarray(g, dtype=np.float32).reshape(-1)
    step = max(1, g.size // 4096)
    return g[::step][:4096].copy(), np.array([g.astype(np.float64).sum(),
                                              np.sqrt((g.astype(np.float64) ** 2).sum())])


MODEL_CASES = {
    # tag: (kind, kwargs)
    'nrms_ulm0': ('nrms', dict(user_log_mask=False, V=97, L=30)),
    'nrms_ulm1': ('nrms', dict(user_log_mask=True, V=97, L=30)),
    'naml_ulm0_c11': ('naml', dict(user_log_mask=False, use_cat=True, use_sub=True, N=41, L=20)),
    'naml_ulm1_c11': ('naml', dict(user_log_mask=True, use_cat=True, use_sub=True, N=41, L=20)),
    'naml_ulm0_c00': ('naml', dict(user_log_mask=False, use_cat=False, use_sub=False, N=41, L=30)),
    'naml_ulm1_c10': ('naml', dict(user_log_mask=True, use_cat=True, use_sub=False, N=41, L=20)),
}
B_, T_, C_ = 3, 50, 5


def case_inputs(tag):
    kind, kw = MODEL_CASES[tag]
    seed = sorted(MODEL_CASES).index(tag) + 10
    if kind == 'nrms':
        P = nrms_weights(seed, kw['V'])
        bt = batch(seed + 100, B_, T_, C_, kw['L'], kw['V'])
    else:
        P = naml_weights(seed, kw['N'], kw['L'], kw['use_cat'], kw['use_sub'])
        F = 1 + int(kw['use_cat']) + int(kw['use_sub'])
        bt = batch(seed + 100, B_, T_, C_, kw['L'], kw['N'] + 1, F=F)
    return kind, kw, P, bt
