"""CPU oracle for the NRMS/NAML train-and-eval hot path.

TEST INFRASTRUCTURE ONLY.  This module is a CPU restatement (torch-CPU tensor
algebra + numpy for the ranking metrics) of the algorithm implemented by the
reference `patngnw/NewsRecommendation`.  It is the *checker* for the CUDA
product path: only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s
`cpu_baseline` / `--impl reference` legs may import it.  Nothing under
`newsrecommendation_b200/` imports it and the product path never falls back
to it.

Parity pinning: the reference has no tests / golden vectors of its own
(SURVEY.md §4), so this oracle is pinned against outputs of the reference
itself, imported unmodified from /root/reference/src in the build container by
`tests/golden/make_golden.py`; the resulting fixtures are committed under
`tests/golden/*.npz` and checked by `tests/test_oracle_golden.py`.

Every function cites the reference file:line it restates (paths relative to
/root/reference/).  All functions are pure: weights come in as a flat
`dict[str, Tensor]` whose keys are the reference's `state_dict` keys.

Dropout: the reference uses torch's Philox stream which a CUDA kernel cannot
reproduce.  Here dropout takes an explicit keep-mask (0/1 tensor) so that the
CUDA path's own counter-based mask can be exported and replayed -> exact
parity in train mode.  mask=None == eval mode / drop_rate 0.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import torch

Tensor = torch.Tensor
EPS = 1e-8  # src/model/model_utils.py:29,53


# ----------------------------------------------------------------------------
# op layer  (src/model/model_utils.py)
# ----------------------------------------------------------------------------
def dropout(x: Tensor, keep: Optional[Tensor], p: float) -> Tensor:
    """Inverted dropout, F.dropout semantics (src/model/NRMS.py:28-34,
    src/model/NAML.py:51-53): y = x * keep / (1-p).  keep=None -> identity."""
    if keep is None or p == 0.0:
        return x
    return x * keep.to(x.dtype) * (1.0 / (1.0 - p))


def embedding_lookup(table: Tensor, ids: Tensor, padding_idx: int = 0) -> Tensor:
    """nn.Embedding(..., padding_idx=0) lookup (src/model/NRMS.py:70-73,28;
    src/model/NAML.py:19,22,104-107): a plain row gather in forward; in
    backward the padding row receives NO gradient."""
    ids = ids.long()
    rows = table[ids]
    if table.requires_grad:
        pad = (ids == padding_idx).reshape(ids.shape + (1,) * (rows.dim() - ids.dim()))
        rows = torch.where(pad, rows.detach(), rows)
    return rows


def attention_pooling(x: Tensor, w1: Tensor, b1: Tensor, w2: Tensor, b2: Tensor,
                      mask: Optional[Tensor] = None, return_alpha: bool = False):
    """AttentionPooling.forward, src/model/model_utils.py:13-31.

    x (n,S,H); w1 (Q,H); b1 (Q); w2 (1,Q); b2 (1); mask (n,S) or None -> (n,H)
    exp() WITHOUT max subtraction, multiplicative mask, eps inside the
    denominator (model_utils.py:24-29)."""
    e = torch.tanh(x @ w1.t() + b1)                       # :21-22
    a = torch.exp(e @ w2.reshape(-1, 1) + b2)             # :23-24  (n,S,1)
    if mask is not None:
        a = a * mask.to(a.dtype).unsqueeze(2)             # :26-27
    alpha = a / (a.sum(dim=1, keepdim=True) + EPS)        # :29
    out = (x * alpha).sum(dim=1)                          # :30  bmm(x^T, alpha)
    if return_alpha:
        return out, alpha.squeeze(2)
    return out


def scaled_dot_product_attention(q: Tensor, k: Tensor, v: Tensor, d_k: int,
                                 mask: Optional[Tensor] = None) -> Tensor:
    """ScaledDotProductAttention.forward, src/model/model_utils.py:39-55.
    q,k,v (n,h,S,d); mask (n,h,S) is a KEY mask broadcast over query rows."""
    s = torch.exp((q @ k.transpose(-1, -2)) / math.sqrt(d_k))   # :47-48
    if mask is not None:
        s = s * mask.to(s.dtype).unsqueeze(-2)                   # :50-51
    attn = s / (s.sum(dim=-1, keepdim=True) + EPS)               # :53
    return attn @ v                                              # :54


def multi_head_self_attention(x: Tensor, wq, bq, wk, bk, wv, bv, n_heads: int,
                              mask: Optional[Tensor] = None) -> Tensor:
    """MultiHeadSelfAttention.forward, src/model/model_utils.py:78-95
    (Q=K=V=x as at every call site).  x (n,S,d_model) -> (n,S,n_heads*d_k)."""
    n, S, _ = x.shape
    d_k = wq.shape[0] // n_heads
    hm = None
    if mask is not None:
        hm = mask.unsqueeze(1).expand(-1, n_heads, -1)           # :86-87
    q = (x @ wq.t() + bq).view(n, S, n_heads, d_k).transpose(1, 2)   # :89
    k = (x @ wk.t() + bk).view(n, S, n_heads, d_k).transpose(1, 2)   # :90
    v = (x @ wv.t() + bv).view(n, S, n_heads, d_k).transpose(1, 2)   # :91
    ctx = scaled_dot_product_attention(q, k, v, d_k, hm)             # :93
    return ctx.transpose(1, 2).reshape(n, S, n_heads * d_k)          # :94


def conv1d_k3(x: Tensor, w: Tensor, b: Tensor) -> Tensor:
    """nn.Conv1d(Cin->Cout, kernel_size=3, padding=1) applied along the token
    axis, src/model/NAML.py:27-32 and call :54 (with its two transposes folded).
    x (n,L,Cin); w (Cout,Cin,3); b (Cout) -> (n,L,Cout).
    y[l,o] = b[o] + sum_t sum_c w[o,c,t] * x[l+t-1,c], x = 0 outside [0,L)."""
    n, L, _ = x.shape
    xp = torch.zeros(n, L + 2, x.shape[2], dtype=x.dtype)
    xp[:, 1:L + 1] = x
    y = b.to(x.dtype).expand(n, L, -1).clone()
    for t in range(3):
        y = y + xp[:, t:t + L] @ w[:, :, t].t()
    return y


def pad_doc_blend(v: Tensor, mask: Tensor, pad_doc: Tensor) -> Tensor:
    """src/model/NRMS.py:59-60 / src/model/NAML.py:94-95."""
    m = mask.to(v.dtype).unsqueeze(-1)
    return v * m + pad_doc.unsqueeze(0) * (1 - m)


def score_and_loss(cand_vecs: Tensor, user_vec: Tensor, label: Tensor):
    """src/model/NRMS.py:93-94 / src/model/NAML.py:128-129: bmm score then
    nn.CrossEntropyLoss() (mean over batch)."""
    score = (cand_vecs * user_vec.unsqueeze(1)).sum(-1)
    lse = torch.logsumexp(score, dim=1)
    picked = score.gather(1, label.long().view(-1, 1)).squeeze(1)
    loss = (lse - picked).mean()
    return loss, score


def train_accuracy(label: Tensor, score: Tensor) -> float:
    """utils.acc, src/utils.py:36-40."""
    return float((score.argmax(dim=-1) == label).float().mean())


# ----------------------------------------------------------------------------
# NRMS  (src/model/NRMS.py)
# ----------------------------------------------------------------------------
def _pool_w(P: Dict[str, Tensor], prefix: str):
    return (P[prefix + '.att_fc1.weight'], P[prefix + '.att_fc1.bias'],
            P[prefix + '.att_fc2.weight'], P[prefix + '.att_fc2.bias'])


def _mhsa_w(P: Dict[str, Tensor], prefix: str):
    return (P[prefix + '.W_Q.weight'], P[prefix + '.W_Q.bias'],
            P[prefix + '.W_K.weight'], P[prefix + '.W_K.bias'],
            P[prefix + '.W_V.weight'], P[prefix + '.W_V.bias'])


def nrms_news_encoder(ids: Tensor, P: Dict[str, Tensor], n_heads: int, p: float = 0.0,
                      keep1: Optional[Tensor] = None, keep2: Optional[Tensor] = None) -> Tensor:
    """NRMS NewsEncoder.forward, src/model/NRMS.py:23-36, always mask=None
    (call sites NRMS.py:87,90; main.py:194).  ids (n,L) int -> (n,H).
    keep1 (n,L,D) / keep2 (n,L,H): dropout keep masks (None = eval)."""
    emb = P['news_encoder.embedding_matrix.weight']
    x = dropout(embedding_lookup(emb, ids), keep1, p)                                   # :28-30
    h = multi_head_self_attention(x, *_mhsa_w(P, 'news_encoder.multi_head_self_attn'),
                                  n_heads=n_heads)                           # :31
    h = dropout(h, keep2, p)                                                 # :32-34
    return attention_pooling(h, *_pool_w(P, 'news_encoder.attn'))            # :35


def nrms_user_encoder(news_vecs: Tensor, log_mask: Tensor, P: Dict[str, Tensor],
                      n_heads: int, user_log_mask: bool) -> Tensor:
    """NRMS UserEncoder.forward, src/model/NRMS.py:49-63."""
    mw = _mhsa_w(P, 'user_encoder.multi_head_self_attn')
    pw = _pool_w(P, 'user_encoder.attn')
    if user_log_mask:                                                        # :55-57
        h = multi_head_self_attention(news_vecs, *mw, n_heads=n_heads, mask=log_mask)
        return attention_pooling(h, *pw, mask=log_mask)
    v = pad_doc_blend(news_vecs, log_mask, P['user_encoder.pad_doc'])        # :59-60
    h = multi_head_self_attention(v, *mw, n_heads=n_heads)                   # :61
    return attention_pooling(h, *pw)                                         # :62


def nrms_forward(history: Tensor, history_mask: Tensor, candidate: Tensor, label: Tensor,
                 P: Dict[str, Tensor], n_heads: int, user_log_mask: bool, p: float = 0.0,
                 keep: Optional[Dict[str, Tensor]] = None):
    """NRMS Model.forward, src/model/NRMS.py:79-95.  Returns (loss, score).
    keep: optional dict with 'cand1','cand2','hist1','hist2' dropout keep masks."""
    B, T, L = history.shape
    C = candidate.shape[1]
    k = keep or {}
    cv = nrms_news_encoder(candidate.reshape(-1, L), P, n_heads, p,
                           k.get('cand1'), k.get('cand2')).reshape(B, C, -1)   # :86-87
    hv = nrms_news_encoder(history.reshape(-1, L), P, n_heads, p,
                           k.get('hist1'), k.get('hist2')).reshape(B, T, -1)   # :89-90
    u = nrms_user_encoder(hv, history_mask, P, n_heads, user_log_mask)         # :92
    return score_and_loss(cv, u, label)                                        # :93-94


# ----------------------------------------------------------------------------
# NAML  (src/model/NAML.py)
# ----------------------------------------------------------------------------
def naml_news_encoder(x: Tensor, P: Dict[str, Tensor], L: int, D: int, use_category: bool,
                      use_subcategory: bool, p: float = 0.0,
                      keep1: Optional[Tensor] = None, word_ids: bool = False, La: int = 0) -> Tensor:
    """NAML NewsEncoder.forward, src/model/NAML.py:35-75 (fork layout: column 0
    of x is the NEWS index into the per-news title table (N+1, L*D), NAML.py:47-50).
    word_ids=True is the upstream/commented layout NAML.py:40-45:
    columns [0,L) are token ids into a (V,D) word table, side columns follow.
    La > 0 adds the ABSTRACT view (BASELINE config 3 "+abstract").  The reference has NO code for it
    (parameters.py:43 is an unused flag): this is the title view's chain (NAML.py:47-55) applied to a
    second text field with its own weights `news_encoder.abstract_*`, stacked second — parity UNPINNED by the
    reference for this view."""
    if word_ids:
        tok = x[:, :L].long()                                                # :40
        emb = embedding_lookup(P['news_encoder.title_embeddings.weight'], tok)             # :41
        start = L
    else:
        idx = x[:, 0].long()                                                 # :47-48
        emb = embedding_lookup(P['news_encoder.title_embeddings.weight'], idx).reshape(-1, L, D)   # :48-50
        start = 1
    w = dropout(emb, keep1, p)                                               # :51-53
    c = conv1d_k3(w, P['news_encoder.cnn.weight'], P['news_encoder.cnn.bias'])   # :54
    title = attention_pooling(c, *_pool_w(P, 'news_encoder.attn'))           # :55
    views = [title]
    if La > 0:                                                               # by analogy with :47-55 (no reference code)
        if word_ids:
            aemb = embedding_lookup(P['news_encoder.abstract_embeddings.weight'], x[:, start:start + La].long())
            start += La
        else:
            aemb = embedding_lookup(P['news_encoder.abstract_embeddings.weight'], x[:, start].long()).reshape(-1, La, D)
            start += 1
        ca = conv1d_k3(aemb, P['news_encoder.abstract_cnn.weight'], P['news_encoder.abstract_cnn.bias'])
        views.append(attention_pooling(ca, *_pool_w(P, 'news_encoder.abstract_attn')))
    if use_category:                                                         # :60-64
        cat = x[:, start].long()
        views.append(embedding_lookup(P['news_encoder.category_emb.weight'], cat)
                     @ P['news_encoder.category_dense.weight'].t()
                     + P['news_encoder.category_dense.bias'])
        start += 1
    if use_subcategory:                                                      # :65-68
        sub = x[:, start].long()
        views.append(embedding_lookup(P['news_encoder.subcategory_emb.weight'], sub)
                     @ P['news_encoder.subcategory_dense.weight'].t()
                     + P['news_encoder.subcategory_dense.bias'])
    if len(views) == 1:                                                      # :70-71
        return views[0]
    return attention_pooling(torch.stack(views, dim=1),
                             *_pool_w(P, 'news_encoder.final_attn'))         # :73-74


def naml_user_encoder(news_vecs: Tensor, log_mask: Tensor, P: Dict[str, Tensor],
                      user_log_mask: bool) -> Tensor:
    """NAML UserEncoder.forward, src/model/NAML.py:85-97."""
    pw = _pool_w(P, 'user_encoder.attn')
    if user_log_mask:
        return attention_pooling(news_vecs, *pw, mask=log_mask)              # :91-92
    v = pad_doc_blend(news_vecs, log_mask, P['user_encoder.pad_doc'])        # :94-95
    return attention_pooling(v, *pw)                                         # :96


def naml_forward(history: Tensor, history_mask: Tensor, candidate: Tensor, label: Tensor,
                 P: Dict[str, Tensor], L: int, D: int, use_category: bool, use_subcategory: bool,
                 user_log_mask: bool, p: float = 0.0, keep: Optional[Dict[str, Tensor]] = None,
                 word_ids: bool = False, La: int = 0):
    """NAML Model.forward, src/model/NAML.py:113-130.  Returns (loss, score)."""
    B, T, F = history.shape
    C = candidate.shape[1]
    k = keep or {}
    cv = naml_news_encoder(candidate.reshape(-1, F), P, L, D, use_category, use_subcategory,
                           p, k.get('cand1'), word_ids, La).reshape(B, C, -1)     # :121-122
    hv = naml_news_encoder(history.reshape(-1, F), P, L, D, use_category, use_subcategory,
                           p, k.get('hist1'), word_ids, La).reshape(B, T, -1)     # :124-125
    u = naml_user_encoder(hv, history_mask, P, user_log_mask)                 # :127
    return score_and_loss(cv, u, label)                                       # :128-129


# ----------------------------------------------------------------------------
# optimizer  (src/main.py:76 -> torch.optim.Adam defaults)
# ----------------------------------------------------------------------------
def adam_step(param: Tensor, grad: Tensor, m: Tensor, v: Tensor, step: int, lr: float,
              beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8) -> None:
    """One torch.optim.Adam update (amsgrad off, weight_decay 0, bias-corrected),
    in place, for the optimizer built at src/main.py:76.  `step` is 1-based."""
    m.mul_(beta1).add_(grad, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    param.addcdiv_(m, denom, value=-lr / bc1)


# ----------------------------------------------------------------------------
# ranking metrics  (src/metrics.py; sklearn.metrics.roc_auc_score)
# ----------------------------------------------------------------------------
def _order_desc(score: np.ndarray) -> np.ndarray:
    """Descending order as `np.argsort(y_score)[::-1]` (src/metrics.py:6,20).
    numpy's default sort is not stable so the reference's order among TIED
    scores is unspecified; we fix it as a stable ascending sort reversed
    (= higher index first among equals), which is what the reference yields
    for fully tied impressions."""
    return np.argsort(score, kind='stable')[::-1]


def auc_score(y_true: np.ndarray, y_score: np.ndarray) -> float:
    """sklearn.metrics.roc_auc_score for binary labels (src/metrics.py:1,
    call src/main.py:255) in its Mann-Whitney rank-sum form with average
    ranks for ties."""
    y = np.asarray(y_true).astype(np.int64)
    s = np.asarray(y_score)
    pos = s[y == 1]
    neg = s[y == 0]
    if len(pos) == 0 or len(neg) == 0:
        raise ValueError('AUC undefined for single-class labels')
    less = (neg[None, :] < pos[:, None]).sum()
    eq = (neg[None, :] == pos[:, None]).sum()
    return float((less + 0.5 * eq) / (len(pos) * len(neg)))


def dcg_score(y_true: np.ndarray, y_score: np.ndarray, k: int = 10) -> float:
    """src/metrics.py:5-10."""
    order = _order_desc(np.asarray(y_score))
    yt = np.take(np.asarray(y_true), order[:k])
    gains = 2.0 ** yt - 1
    discounts = np.log2(np.arange(len(yt)) + 2)
    return float(np.sum(gains / discounts))


def ndcg_score(y_true: np.ndarray, y_score: np.ndarray, k: int = 10) -> float:
    """src/metrics.py:13-16 (ideal DCG = labels sorted by themselves)."""
    return dcg_score(y_true, y_score, k) / dcg_score(y_true, y_true, k)


def mrr_score(y_true: np.ndarray, y_score: np.ndarray) -> float:
    """src/metrics.py:19-23: mean reciprocal rank over ALL positives."""
    order = _order_desc(np.asarray(y_score))
    yt = np.take(np.asarray(y_true), order)
    rr = yt / (np.arange(len(yt)) + 1)
    return float(np.sum(rr) / np.sum(yt))


def impression_metrics(labels: np.ndarray, scores: np.ndarray):
    """The per-impression body of src/main.py:249-263.  Returns None for the
    skipped impressions (labels all 0 or all 1, main.py:250-251)."""
    m = labels.mean()
    if m == 0 or m == 1:
        return None
    return (auc_score(labels, scores), mrr_score(labels, scores),
            ndcg_score(labels, scores, 5), ndcg_score(labels, scores, 10))


# ----------------------------------------------------------------------------
# eval loop  (src/main.py:177-277, src/dataset.py:17-24,56-78)
# ----------------------------------------------------------------------------
def pad_history(hist: Sequence[int], T: int) -> Tuple[np.ndarray, np.ndarray]:
    """DatasetTrain.pad_to_fix_len(padding_front=True), src/dataset.py:17-24:
    keep the LAST T items, pad the FRONT with 0, mask 0/1 float32."""
    hist = list(hist)
    pad = [0] * (T - len(hist)) + hist[-T:]
    mask = [0] * (T - len(hist)) + [1] * min(T, len(hist))
    return np.asarray(pad, dtype=np.int32), np.asarray(mask, dtype=np.float32)


def build_news_cache(encode, news_combined: np.ndarray, batch: int = 128) -> np.ndarray:
    """src/main.py:185-198: news_encoder over EVERY row of news_combined
    (row 0 = pad news included) in batches -> (N+1, H) float32."""
    out = []
    with torch.no_grad():
        for i in range(0, news_combined.shape[0], batch):
            out.append(encode(torch.from_numpy(news_combined[i:i + batch])).float().numpy())
    return np.concatenate(out, axis=0)


def eval_impressions(user_encode, news_scoring: np.ndarray, hist_idx: np.ndarray,
                     hist_mask: np.ndarray, offsets: np.ndarray, cand_idx: np.ndarray,
                     labels: np.ndarray, batch: int = 128):
    """src/main.py:240-263 with the DatasetTest gathers of src/dataset.py:64-74.
    Impressions given as CSR (offsets, cand_idx, labels) + (I,T) history.
    Returns (per-impression metrics (I,4) float64 with NaN rows for skipped,
    scores flat float32)."""
    I = hist_idx.shape[0]
    mets = np.full((I, 4), np.nan, dtype=np.float64)
    scores = np.zeros(cand_idx.shape[0], dtype=np.float32)
    with torch.no_grad():
        for b0 in range(0, I, batch):
            b1 = min(I, b0 + batch)
            log_vecs = torch.from_numpy(news_scoring[hist_idx[b0:b1]])          # dataset.py:67-68
            log_mask = torch.from_numpy(hist_mask[b0:b1])
            user_vecs = user_encode(log_vecs, log_mask).float().numpy()          # main.py:247
            for i in range(b0, b1):
                lo, hi = int(offsets[i]), int(offsets[i + 1])
                lab = labels[lo:hi].astype(np.int64)
                sc = np.dot(news_scoring[cand_idx[lo:hi]], user_vecs[i - b0])    # main.py:253
                scores[lo:hi] = sc
                r = impression_metrics(lab, sc)                                  # main.py:250-258
                if r is not None:
                    mets[i] = r
    return mets, scores


def summarize_metrics(mets: np.ndarray):
    """src/main.py:277 (single-process semantics: mean over SCORED impressions)
    plus the distributed-mode denominator of main.py:275 (all impressions)."""
    valid = ~np.isnan(mets[:, 0])
    sums = mets[valid].sum(axis=0)
    return {'sums': sums, 'scored': int(valid.sum()), 'total': int(mets.shape[0]),
            'mean_scored': sums / max(1, int(valid.sum())),
            'mean_total': sums / max(1, mets.shape[0])}


# ----------------------------------------------------------------------------
# batch assembly  (src/dataset.py:14-49, 64-74) — SURVEY §8f rank 1
# ----------------------------------------------------------------------------
def trans_to_nindex(news_index: Dict[str, int], nids: Sequence[str]):
    """DatasetTrain.trans_to_nindex, src/dataset.py:14-15: unknown ids -> 0."""
    return [news_index[i] if i in news_index else 0 for i in nids]


def train_line_mapper(line: str, news_index: Dict[str, int], news_combined: np.ndarray, T: int, label: int):
    """DatasetTrain.line_mapper, src/dataset.py:26-49, with the `random.randint(0, npratio)` draw of
    :45 passed in as `label`.  Returns (history features (T,F), mask (T,), candidate features
    (1+K,F), label)."""
    f = line.strip().split('\t')
    hist, mask = pad_history(trans_to_nindex(news_index, f[3].split()), T)          # :39
    pos = trans_to_nindex(news_index, f[4].split())                                 # :42
    neg = trans_to_nindex(news_index, f[5].split())                                 # :43
    sample = neg[:label] + pos + neg[label:]                                        # :46
    return news_combined[hist], mask, news_combined[np.asarray(sample, dtype=np.int64)], label


def test_line_mapper(line: str, news_index: Dict[str, int], T: int):
    """DatasetTest.line_mapper, src/dataset.py:64-74, index level: padded click indices (T,), mask (T,),
    candidate indices (C,), labels (C,)."""
    f = line.strip().split('\t')
    hist, mask = pad_history(trans_to_nindex(news_index, f[3].split()), T)          # :66-67
    imp = f[4].split()
    cand = trans_to_nindex(news_index, [i.split('-')[0] for i in imp])              # :70
    labels = np.array([int(i.split('-')[1]) for i in imp])                          # :71
    return hist, mask, np.asarray(cand, dtype=np.int32), labels


def batch_train_np(hist_off, hist_idx, pos, neg, sample_ids, label, news_combined, T):
    """Vectorised numpy form of train_line_mapper over a tokenised index (for sizes where the per-line
    Python loop is too slow); same integer results."""
    sample_ids = np.asarray(sample_ids, dtype=np.int64)
    label = np.asarray(label, dtype=np.int64)
    B, K = sample_ids.shape[0], neg.shape[1]
    o0 = hist_off[sample_ids]
    ln = hist_off[sample_ids + 1] - o0
    t = np.arange(T)[None, :]
    padn = (T - np.minimum(ln, T))[:, None]
    real = t >= padn
    src = o0[:, None] + ln[:, None] - (T - t)
    idx = np.where(real, hist_idx[np.where(real, src, 0)], 0).astype(np.int32)
    j = np.arange(1 + K)[None, :]
    lb = label[:, None]
    negj = np.where(j < lb, j, j - 1).clip(0, K - 1)
    cand = np.where(j == lb, pos[sample_ids][:, None], neg[sample_ids][np.arange(B)[:, None], negj]).astype(np.int32)
    return news_combined[idx], real.astype(np.float32), news_combined[cand], label
