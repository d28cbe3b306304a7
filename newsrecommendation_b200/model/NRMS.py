"""NRMS plugin (drop-in for src/model/NRMS.py): Model(args, embedding_matrix, ...),
model(history, history_mask, candidate, label) -> (loss, score), .news_encoder(ids),
.user_encoder(vecs, mask); 22 state_dict keys as the reference."""
import torch
from torch import nn

from newsrecommendation_b200 import ops
from newsrecommendation_b200 import _lib as L
from .model_utils import AttentionPooling, MultiHeadSelfAttention, encode_both, precision_of


def _heads(args):
    dk = args.news_dim // args.num_attention_heads
    if dk * args.num_attention_heads != args.news_dim:
        raise ValueError('news_dim must equal num_attention_heads * dim_per_head')
    return dk


class NewsEncoder(nn.Module):
    """word gather -> dropout -> MHSA -> dropout -> additive pooling (NRMS.py:8-36).
    The gather and both dropouts run inside the MHSA kernels."""

    def __init__(self, args, embedding_matrix):
        super().__init__()
        prec = precision_of(args)
        self.prec = prec
        self.embedding_matrix = embedding_matrix
        self.drop_rate = args.drop_rate
        self.dim_per_head = _heads(args)
        self.multi_head_self_attn = MultiHeadSelfAttention(args.word_embedding_dim, args.num_attention_heads,
                                                           self.dim_per_head, self.dim_per_head, prec)
        self.attn = AttentionPooling(args.news_dim, args.news_query_vector_dim, prec)

    def forward(self, x, mask=None):
        if mask is not None:
            raise RuntimeError('title masks are never used by the reference (NRMS.py:87,90); not built')
        p = self.drop_rate if self.training else 0.0
        if self.prec == L.PREC_BF16:
            m = self.multi_head_self_attn
            return ops.title_encoder(L.ENC_MHSA, self.embedding_matrix.weight, x.reshape(-1, x.shape[-1]), None,
                                     m._w(), self.attn._p(), mode=L.ROWS_TOKEN, S=x.shape[-1], Dm=m.d_model,
                                     heads=m.n_heads, in_drop=p, out_drop=p, seed=ops.next_seed() if p else 0)
        h = self.multi_head_self_attn.forward_rows(self.embedding_matrix.weight, x.reshape(-1, x.shape[-1]),
                                                   L.ROWS_TOKEN, x.shape[-1], p, p, ops.next_seed() if p else 0)
        return self.attn(h)


class UserEncoder(nn.Module):
    """NRMS.py:39-63: masked MHSA + masked pooling, or pad_doc blend + unmasked."""

    def __init__(self, args):
        super().__init__()
        prec = precision_of(args)
        self.prec = prec
        self.args = args
        self.dim_per_head = _heads(args)
        self.multi_head_self_attn = MultiHeadSelfAttention(args.news_dim, args.num_attention_heads,
                                                           self.dim_per_head, self.dim_per_head, prec)
        self.attn = AttentionPooling(args.news_dim, args.user_query_vector_dim, prec)
        self.pad_doc = nn.Parameter(torch.empty(1, args.news_dim).uniform_(-1, 1))

    def forward(self, news_vecs, log_mask=None):
        if self.prec == L.PREC_BF16:
            m = self.multi_head_self_attn
            if self.args.user_log_mask:
                v, mask = news_vecs, log_mask
            else:
                v, mask = ops.pad_blend(news_vecs, log_mask, self.pad_doc), None
            return ops.title_encoder(L.ENC_MHSA, v, None, mask, m._w(), self.attn._p(), mode=L.ROWS_DENSE,
                                     S=v.shape[1], Dm=m.d_model, heads=m.n_heads)
        if self.args.user_log_mask:
            h = self.multi_head_self_attn(news_vecs, mask=log_mask)
            return self.attn(h, log_mask)
        v = ops.pad_blend(news_vecs, log_mask, self.pad_doc)
        return self.attn(self.multi_head_self_attn(v))


    SHORT = 31          # longest history the fused title-encoder kernel takes (one 32-row TMEM quadrant)

    def forward_gathered(self, table, idx, log_mask):
        """Eval (src/dataset.py:67-68 + main.py:247): user vectors straight from the cached news-vector table and the
        padded click indices — the (b,50,400) fp32 gather of the reference is never materialised, the staging kernel
        of the encoder reads the table rows itself.  Returns None when this mode needs the materialised rows
        (pad_doc blend, fp32 parity mode): the caller then gathers and calls forward().

        With `user_log_mask=True` a masked slot contributes nothing as a key (model_utils.py:50-51) and is dropped as a
        query by the masked pooling (:26-27), and histories are FRONT padded (dataset.py:17-24): a history with at most
        31 real clicks is therefore exactly a 31-slot sequence (its last 31 slots).  Those — two thirds of MIND-shaped
        histories — go through the fused inference kernel (one launch after the staging), the longer ones through the
        separate launches over all 50 slots."""
        if self.prec != L.PREC_BF16 or not self.args.user_log_mask:
            return None
        m = self.multi_head_self_attn
        T = idx.shape[1]

        def run(ids, mask):
            return ops.title_encoder(L.ENC_MHSA, table, ids, mask, m._w(), self.attn._p(), mode=L.ROWS_TOKEN,
                                     S=ids.shape[1], Dm=m.d_model, heads=m.n_heads)
        if T <= self.SHORT or self.SHORT <= 0 or torch.is_grad_enabled():
            return run(idx, log_mask)
        short = log_mask[:, :T - self.SHORT].sum(1) == 0          # nothing real before the last 31 slots
        si = short.nonzero(as_tuple=True)[0]
        li = (~short).nonzero(as_tuple=True)[0]
        out = torch.empty(idx.shape[0], table.shape[1], dtype=torch.float32, device=table.device)
        if si.numel():
            out[si] = run(idx[si][:, T - self.SHORT:].contiguous(), log_mask[si][:, T - self.SHORT:].contiguous())
        if li.numel():
            out[li] = run(idx[li].contiguous(), log_mask[li].contiguous())
        return out


class Model(nn.Module):
    def __init__(self, args, embedding_matrix, num_category=None, num_subcategory=None, **kwargs):
        # two extra positionals tolerated: main.py:64 passes them to every model
        super().__init__()
        self.args = args
        table = torch.as_tensor(embedding_matrix).float()
        emb = nn.Embedding.from_pretrained(table, freeze=args.freeze_embedding, padding_idx=0)
        self.news_encoder = NewsEncoder(args, emb)
        self.user_encoder = UserEncoder(args)
        self.last_hits = None

    def forward(self, history, history_mask, candidate, label):
        a = self.args
        Lw = a.num_words_title
        cv, hv = encode_both(self.news_encoder, candidate.reshape(-1, Lw), history.reshape(-1, Lw))
        cv = cv.reshape(-1, 1 + a.npratio, a.news_dim)
        hv = hv.reshape(-1, a.user_log_length, a.news_dim)
        u = self.user_encoder(hv, history_mask)
        loss, score, hits = ops.score_ce(cv, u, label)
        self.last_hits = hits      # #argmax==label of this batch (utils.acc numerator), device scalar
        return loss, score
