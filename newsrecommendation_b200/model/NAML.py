"""NAML plugin (drop-in for src/model/NAML.py), fork layout: column 0 of the feature
row is the NEWS index into a per-news title table (N+1, L*D) (NAML.py:47-50), followed by
the category / subcategory columns.  `args.naml_word_ids=True` selects the upstream
word-id layout (commented-out code NAML.py:40-45): columns [0,L) are token
ids into a (V,D) word table.

`args.use_abstract=True` adds the ABSTRACT view of BASELINE config 3 ("+abstract"; flag `--num_words_abstract`,
parameters.py:43).  The reference has no code for it, so this is a composition of the reference's
own modules by analogy with the title view and the NAML paper: a second per-news table (N+1, La*D) (word-id
layout: La more token columns), Conv1d(k=3) + AttentionPooling with their own weights, stacked as the second
view in front of category / subcategory and pooled by `final_attn`.  Feature columns: [title | abstract |
category | subcategory].  New state_dict keys (`news_encoder.abstract_*`) follow the news encoder's reference keys,
which keep their names and order.  Parity for this view is against the same composition in the oracle (unpinned by
the reference, as it must be)."""
import torch
from torch import nn

from newsrecommendation_b200 import ops
from newsrecommendation_b200 import _lib as L
from .model_utils import AttentionPooling, encode_both, precision_of


class NewsEncoder(nn.Module):
    def __init__(self, args, embedding_matrix, num_category, num_subcategory, abstract_embeddings=None):
        super().__init__()
        prec = precision_of(args)
        self.prec = prec
        self.drop_rate = args.drop_rate
        self.num_words_title = args.num_words_title
        self.word_embedding_dim = args.word_embedding_dim
        self.use_category = args.use_category
        self.use_subcategory = args.use_subcategory
        self.word_ids = bool(getattr(args, 'naml_word_ids', False))
        self.use_abstract = bool(getattr(args, 'use_abstract', False))
        self.title_embeddings = embedding_matrix
        # registration order as NAML.py:19-33 so state_dict key order matches
        if args.use_category:
            self.category_emb = nn.Embedding(num_category + 1, args.category_emb_dim, padding_idx=0)
            self.category_dense = nn.Linear(args.category_emb_dim, args.news_dim)
        if args.use_subcategory:
            self.subcategory_emb = nn.Embedding(num_subcategory + 1, args.category_emb_dim, padding_idx=0)
            self.subcategory_dense = nn.Linear(args.category_emb_dim, args.news_dim)
        if args.use_category or args.use_subcategory or self.use_abstract:
            self.final_attn = AttentionPooling(args.news_dim, args.news_query_vector_dim, prec)
        self.cnn = nn.Conv1d(in_channels=args.word_embedding_dim, out_channels=args.news_dim, kernel_size=3,
                             padding=1)
        self.attn = AttentionPooling(args.news_dim, args.news_query_vector_dim, prec)
        if self.use_abstract:
            if abstract_embeddings is None:
                raise ValueError('use_abstract=True needs abstract_embeddings_weight (N+1, num_words_abstract * D)')
            self.num_words_abstract = args.num_words_abstract
            self.abstract_embeddings = abstract_embeddings
            self.abstract_cnn = nn.Conv1d(in_channels=args.word_embedding_dim, out_channels=args.news_dim, kernel_size=3,
                                          padding=1)
            self.abstract_attn = AttentionPooling(args.news_dim, args.news_query_vector_dim, prec)

    def _conv_view(self, table, ids, cnn, attn, S, mode, p):
        """gather -> dropout -> Conv1d(k=3) -> AttentionPooling over one text field (NAML.py:47-55)."""
        D = self.word_embedding_dim
        if self.prec == L.PREC_BF16:
            return ops.title_encoder(L.ENC_CONV, table, ids, None, (cnn.weight, cnn.bias), attn._p(), mode=mode, S=S, Dm=D,
                                     in_drop=p, seed=ops.next_seed() if p else 0)
        c = ops.conv1d_k3(table, ids, cnn.weight, cnn.bias, mode=mode, L_=S, D=D, in_drop=p,
                          seed=ops.next_seed() if p else 0, prec=self.prec)
        return attn(c)

    def forward(self, x, mask=None):
        if mask is not None:
            raise RuntimeError('title masks are never used by the reference (NAML.py:122,125); not built')
        x = x.reshape(-1, x.shape[-1])
        Lw, D = self.num_words_title, self.word_embedding_dim
        p = self.drop_rate if self.training else 0.0
        mode = L.ROWS_TOKEN if self.word_ids else L.ROWS_NEWS
        title = self._conv_view(self.title_embeddings.weight, x, self.cnn, self.attn, Lw, mode, p)
        views = [title]
        col = Lw if self.word_ids else 1
        if self.use_abstract:
            La = self.num_words_abstract
            w = La if self.word_ids else 1
            views.append(self._conv_view(self.abstract_embeddings.weight, x[:, col:col + w], self.abstract_cnn,
                                         self.abstract_attn, La, mode, p))
            col += w
        if self.use_category:
            views.append(ops.embed_dense(self.category_emb.weight, x, col, self.category_dense.weight,
                                         self.category_dense.bias))
            col += 1
        if self.use_subcategory:
            views.append(ops.embed_dense(self.subcategory_emb.weight, x, col, self.subcategory_dense.weight,
                                         self.subcategory_dense.bias))
        if len(views) == 1:
            return views[0]
        return self.final_attn(torch.stack(views, dim=1))


class UserEncoder(nn.Module):
    """NAML.py:78-97."""

    def __init__(self, args):
        super().__init__()
        self.args = args
        self.attn = AttentionPooling(args.news_dim, args.user_query_vector_dim, precision_of(args))
        self.pad_doc = nn.Parameter(torch.empty(1, args.news_dim).uniform_(-1, 1))

    def forward(self, news_vecs, log_mask=None):
        if self.args.user_log_mask:
            return self.attn(news_vecs, log_mask)
        return self.attn(ops.pad_blend(news_vecs, log_mask, self.pad_doc))

    def forward_gathered(self, table, idx, log_mask):
        """Eval: masked pooling straight over rows of the cached news-vector table (see NRMS.UserEncoder)."""
        if self.attn.prec != L.PREC_BF16 or not self.args.user_log_mask:
            return None
        return ops.title_encoder(L.ENC_POOL, table, idx, log_mask, (), self.attn._p(), mode=L.ROWS_TOKEN,
                                 S=idx.shape[1], Dm=table.shape[1])


class Model(nn.Module):
    def __init__(self, args, news_embeddings_weight, num_category, num_subcategory, abstract_embeddings_weight=None,
                 **kwargs):
        super().__init__()
        self.args = args
        table = torch.as_tensor(news_embeddings_weight).float()
        emb = nn.Embedding.from_pretrained(table, freeze=args.freeze_embedding, padding_idx=0)
        aemb = None
        if abstract_embeddings_weight is not None:
            aemb = nn.Embedding.from_pretrained(torch.as_tensor(abstract_embeddings_weight).float(),
                                                freeze=args.freeze_embedding, padding_idx=0)
        self.news_encoder = NewsEncoder(args, emb, num_category, num_subcategory, aemb)
        self.user_encoder = UserEncoder(args)
        self.last_hits = None

    def forward(self, history, history_mask, candidate, label):
        a = self.args
        F = history.shape[-1]
        cv, hv = encode_both(self.news_encoder, candidate.reshape(-1, F), history.reshape(-1, F))
        cv = cv.reshape(-1, 1 + a.npratio, a.news_dim)
        hv = hv.reshape(-1, a.user_log_length, a.news_dim)
        u = self.user_encoder(hv, history_mask)
        loss, score, hits = ops.score_ce(cv, u, label)
        self.last_hits = hits
        return loss, score
