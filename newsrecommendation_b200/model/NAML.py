"""NAML plugin (drop-in for src/model/NAML.py), fork layout: column 0 of the feature
row is the NEWS index into a per-news title table (N+1, L*D) (NAML.py:47-50), followed by
the category / subcategory columns.  `args.naml_word_ids=True` selects the upstream
word-id layout (commented-out code NAML.py:40-45): columns [0,L) are token
ids into a (V,D) word table."""
import torch
from torch import nn

from newsrecommendation_b200 import ops
from newsrecommendation_b200 import _lib as L
from .model_utils import AttentionPooling, encode_both, precision_of


class NewsEncoder(nn.Module):
    def __init__(self, args, embedding_matrix, num_category, num_subcategory):
        super().__init__()
        prec = precision_of(args)
        self.prec = prec
        self.drop_rate = args.drop_rate
        self.num_words_title = args.num_words_title
        self.word_embedding_dim = args.word_embedding_dim
        self.use_category = args.use_category
        self.use_subcategory = args.use_subcategory
        self.word_ids = bool(getattr(args, 'naml_word_ids', False))
        self.title_embeddings = embedding_matrix
        # registration order as NAML.py:19-33 so state_dict key order matches
        if args.use_category:
            self.category_emb = nn.Embedding(num_category + 1, args.category_emb_dim, padding_idx=0)
            self.category_dense = nn.Linear(args.category_emb_dim, args.news_dim)
        if args.use_subcategory:
            self.subcategory_emb = nn.Embedding(num_subcategory + 1, args.category_emb_dim, padding_idx=0)
            self.subcategory_dense = nn.Linear(args.category_emb_dim, args.news_dim)
        if args.use_category or args.use_subcategory:
            self.final_attn = AttentionPooling(args.news_dim, args.news_query_vector_dim, prec)
        self.cnn = nn.Conv1d(in_channels=args.word_embedding_dim, out_channels=args.news_dim, kernel_size=3,
                             padding=1)
        self.attn = AttentionPooling(args.news_dim, args.news_query_vector_dim, prec)

    def forward(self, x, mask=None):
        if mask is not None:
            raise RuntimeError('title masks are never used by the reference (NAML.py:122,125); not built')
        x = x.reshape(-1, x.shape[-1])
        Lw, D = self.num_words_title, self.word_embedding_dim
        p = self.drop_rate if self.training else 0.0
        mode = L.ROWS_TOKEN if self.word_ids else L.ROWS_NEWS
        if self.prec == L.PREC_BF16:
            title = ops.title_encoder(L.ENC_CONV, self.title_embeddings.weight, x, None,
                                      (self.cnn.weight, self.cnn.bias), self.attn._p(), mode=mode, S=Lw, Dm=D,
                                      in_drop=p, seed=ops.next_seed() if p else 0)
        else:
            c = ops.conv1d_k3(self.title_embeddings.weight, x, self.cnn.weight, self.cnn.bias, mode=mode, L_=Lw,
                              D=D, in_drop=p, seed=ops.next_seed() if p else 0, prec=self.prec)
            title = self.attn(c)
        views = [title]
        col = Lw if self.word_ids else 1
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


class Model(nn.Module):
    def __init__(self, args, news_embeddings_weight, num_category, num_subcategory, **kwargs):
        super().__init__()
        self.args = args
        table = torch.as_tensor(news_embeddings_weight).float()
        emb = nn.Embedding.from_pretrained(table, freeze=args.freeze_embedding, padding_idx=0)
        self.news_encoder = NewsEncoder(args, emb, num_category, num_subcategory)
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
