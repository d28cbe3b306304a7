"""Synthetic MIND-shaped inputs: datasets / GloVe / BPEmb are unavailable
offline, so benchmarks and large-size tests use these generators.  Shapes, dtypes and
padding conventions are exactly what the reference's host pipeline emits
(src/preprocess.py:50-72, src/dataset.py:17-49,64-74)."""
import argparse

import numpy as np


def default_args(**kw):
    """Hot-path flags with the defaults of src/parameters.py:32-59."""
    a = argparse.Namespace(
        drop_rate=0.2, news_dim=400, num_attention_heads=20, word_embedding_dim=300,
        news_query_vector_dim=200, user_query_vector_dim=200, user_log_mask=False,
        user_log_length=50, num_words_title=20, npratio=4, freeze_embedding=False,
        use_category=False, use_subcategory=False, category_emb_dim=100, batch_size=32)
    for k, v in kw.items():
        setattr(a, k, v)
    return a


def _zipf_ids(rs, n_ids, size):
    """ids in [1, n_ids) with Zipf(1.0) popularity."""
    r = np.arange(1, n_ids, dtype=np.float64)
    p = 1.0 / r
    p /= p.sum()
    cdf = np.cumsum(p)
    u = rs.random_sample(size)
    return (np.searchsorted(cdf, u) + 1).clip(1, n_ids - 1).astype(np.int32)


def word_table(V, D=300, seed=0):
    """(V, D) fp32 ~ N(0, 0.4^2), row 0 (pad) = 0."""
    rs = np.random.RandomState(seed)
    t = (rs.standard_normal((V, D)) * 0.4).astype(np.float32)
    t[0] = 0
    return t


def news_titles(N, L, V, seed=1):
    """(N+1, L) int32 token ids, title length ~ clip(N(11.5,4),3,L), zero padded; row 0 = pad news."""
    rs = np.random.RandomState(seed)
    ids = _zipf_ids(rs, V, (N + 1, L))
    tl = np.clip(np.rint(rs.normal(11.5, 4.0, size=N + 1)), 3, L).astype(np.int64)
    ids[np.arange(L)[None, :] >= tl[:, None]] = 0
    ids[0] = 0
    return ids


def title_table(N, L, D=300, seed=2):
    """NAML fork layout: (N+1, L*D) fp32 per-news title embeddings (zero rows past the title
    length), row 0 = 0."""
    rs = np.random.RandomState(seed)
    t = (rs.standard_normal((N + 1, L, D)) * 0.4).astype(np.float32)
    tl = np.clip(np.rint(rs.normal(11.5, 4.0, size=N + 1)), 3, L).astype(np.int64)
    t[np.arange(L)[None, :] >= tl[:, None]] = 0
    t[0] = 0
    return t.reshape(N + 1, L * D)


def naml_features(N, n_cat=17, n_sub=264, seed=3):
    """news_combined for NAML with both side views: (N+1, 3) int32 [news idx, cat, subcat]."""
    rs = np.random.RandomState(seed)
    f = np.zeros((N + 1, 3), np.int32)
    f[:, 0] = np.arange(N + 1)
    f[1:, 1] = rs.randint(1, n_cat + 1, size=N)
    f[1:, 2] = rs.randint(1, n_sub + 1, size=N)
    return f


def _histories(rs, count, N, T):
    hl = np.minimum(T, np.rint(rs.lognormal(3.0, 1.0, size=count))).astype(np.int64)
    hl[rs.random_sample(count) < 0.02] = 0
    idx = np.zeros((count, T), np.int32)
    mask = np.zeros((count, T), np.float32)
    draw = _zipf_ids(rs, N + 1, (count, T))
    keep = np.arange(T)[None, :] >= (T - hl)[:, None]          # front padding (dataset.py:17-24)
    idx[keep] = draw[keep]
    mask[keep] = 1.0
    return idx, mask


def train_batches(news_combined, n_batches, B, T=50, K=4, seed=4):
    """List of (history (B,T,F) int32, mask (B,T) fp32, candidate (B,1+K,F) int32, label (B,) int64),
    i.e. the default-collated output of DatasetTrain.line_mapper (dataset.py:26-49)."""
    rs = np.random.RandomState(seed)
    N = news_combined.shape[0] - 1
    out = []
    for _ in range(n_batches):
        hidx, mask = _histories(rs, B, N, T)
        cidx = _zipf_ids(rs, N + 1, (B, 1 + K))
        label = rs.randint(0, 1 + K, size=B).astype(np.int64)
        out.append((news_combined[hidx], mask, news_combined[cidx], label))
    return out


def train_index(N, I, K=4, seed=6):
    """Tokenised behaviors_np{K}_{rank}.tsv (batcher.TrainIndex arrays): raw click histories of length
    ~ lognormal(3.0, 1.0) (un-truncated; 2 % empty), one positive and K negatives per line."""
    rs = np.random.RandomState(seed)
    lens = np.minimum(np.rint(rs.lognormal(3.0, 1.0, size=I)), 400).astype(np.int64)
    lens[rs.random_sample(I) < 0.02] = 0
    off = np.zeros(I + 1, np.int64)
    off[1:] = np.cumsum(lens)
    return (off, _zipf_ids(rs, N + 1, int(off[-1])), _zipf_ids(rs, N + 1, I), _zipf_ids(rs, N + 1, (I, K)))


def eval_impressions(N, I, T=50, seed=5):
    """CSR impressions: hist_idx (I,T) int32, hist_mask (I,T) fp32, offsets (I+1) int64,
    cand_idx int32, labels int8.  C ~ clip(lognormal(3.2,.8),2,300); 1-3 positives; a 0.5 %
    slice has single-class labels to exercise the skip rule (main.py:250-251)."""
    rs = np.random.RandomState(seed)
    hidx, mask = _histories(rs, I, N, T)
    C = np.clip(np.rint(rs.lognormal(3.2, 0.8, size=I)), 2, 300).astype(np.int64)
    off = np.zeros(I + 1, np.int64)
    off[1:] = np.cumsum(C)
    cand = _zipf_ids(rs, N + 1, int(off[-1]))
    labels = np.zeros(int(off[-1]), np.int8)
    npos = rs.randint(1, 4, size=I)
    skip = rs.random_sample(I) < 0.005
    for i in range(I):
        if skip[i]:
            continue
        k = min(int(npos[i]), int(C[i]) - 1)
        labels[off[i] + rs.choice(int(C[i]), size=k, replace=False)] = 1
    return hidx, mask, off, cand, labels
