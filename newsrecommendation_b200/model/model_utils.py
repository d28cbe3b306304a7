"""Operator modules of the hot path, mirroring src/model/model_utils.py.

Parameters live in the same nn.Linear containers as the reference (so the
state_dict keys and default initialisation match); forward() hands the raw
parameter tensors to the CUDA library through newsrecommendation_b200.ops.
"""
import os

import torch
from torch import nn

from newsrecommendation_b200 import ops
from newsrecommendation_b200 import _lib as L


def precision_of(args):
    """fp32 (parity mode, rel 1e-5) or bf16 tensor-core mode (rel 1e-2).  Taken from
    `args.precision` when the caller's parser defines it, else $NR_PRECISION, else fp32."""
    p = getattr(args, 'precision', None) or os.environ.get('NR_PRECISION', 'fp32')
    if p not in ('fp32', 'bf16'):
        raise ValueError(f'precision must be fp32 or bf16, got {p!r}')
    return L.PREC_BF16 if p == 'bf16' else L.PREC_FP32


class AttentionPooling(nn.Module):
    """src/model/model_utils.py:7-31."""

    def __init__(self, emb_size, hidden_size, prec=L.PREC_FP32):
        super().__init__()
        self.att_fc1 = nn.Linear(emb_size, hidden_size)
        self.att_fc2 = nn.Linear(hidden_size, 1)
        self.prec = prec

    def forward(self, x, attn_mask=None):
        if self.prec == L.PREC_BF16:      # tensor-core path (title-level pooling is fused into ops.title_encoder)
            return ops.title_encoder(L.ENC_POOL, x, None, attn_mask, (), self._p(), mode=L.ROWS_DENSE,
                                     S=x.shape[1], Dm=x.shape[2])
        return ops.attention_pooling(x, attn_mask, self.att_fc1.weight, self.att_fc1.bias,
                                     self.att_fc2.weight, self.att_fc2.bias, L.PREC_FP32)

    def _p(self):
        return (self.att_fc1.weight, self.att_fc1.bias, self.att_fc2.weight, self.att_fc2.bias)


class ScaledDotProductAttention(nn.Module):
    """src/model/model_utils.py:34-55.  The reference only ever reaches this through
    MultiHeadSelfAttention with Q=K=V; the library fuses it there (nr_mhsa_fwd), so this
    class exists for import compatibility and records d_k."""

    def __init__(self, d_k):
        super().__init__()
        self.d_k = d_k

    def forward(self, Q, K, V, attn_mask=None):
        raise RuntimeError('ScaledDotProductAttention is fused into MultiHeadSelfAttention '
                           '(nr_mhsa_fwd); call that module instead')


class MultiHeadSelfAttention(nn.Module):
    """src/model/model_utils.py:58-95 (xavier-uniform weights, :73-76)."""

    def __init__(self, d_model, n_heads, d_k, d_v, prec=L.PREC_FP32):
        super().__init__()
        if d_k != d_v:
            raise ValueError('d_k != d_v is outside the compiled envelope')
        self.d_model, self.n_heads, self.d_k, self.d_v = d_model, n_heads, d_k, d_v
        self.W_Q = nn.Linear(d_model, d_k * n_heads)
        self.W_K = nn.Linear(d_model, d_k * n_heads)
        self.W_V = nn.Linear(d_model, d_v * n_heads)
        self.scaled_dot_product_attn = ScaledDotProductAttention(d_k)
        self.prec = prec
        for m in (self.W_Q, self.W_K, self.W_V):
            nn.init.xavier_uniform_(m.weight, gain=1)

    def _w(self):
        return (self.W_Q.weight, self.W_Q.bias, self.W_K.weight, self.W_K.bias, self.W_V.weight, self.W_V.bias)

    def forward(self, Q, K=None, V=None, mask=None):
        if (K is not None and K is not Q) or (V is not None and V is not Q):
            raise RuntimeError('only self-attention (Q is K is V) is built, as at every reference call site')
        return ops.mhsa(Q, None, mask, *self._w(), mode=L.ROWS_DENSE, S=Q.shape[1], Dm=self.d_model,
                        heads=self.n_heads, prec=self.prec)

    def forward_rows(self, table, ids, mode, S, in_drop, out_drop, seed):
        """Self-attention over rows gathered from an embedding table (gather + both
        dropouts fused; NRMS.py:28-34)."""
        return ops.mhsa(table, ids, None, *self._w(), mode=mode, S=S, Dm=self.d_model, heads=self.n_heads,
                        in_drop=in_drop, out_drop=out_drop, seed=seed, prec=self.prec)


def encode_both(news_encoder, cand_rows, hist_rows):
    """The reference encodes candidates and history in two calls (NRMS.py:87,90 / NAML.py:122,125);
    one call over the concatenated rows gives the same vectors with half the launches."""
    nc = cand_rows.shape[0]
    return _SplitRows.apply(news_encoder(_concat_rows(cand_rows, hist_rows)), nc)


def _concat_rows(a, b):
    """torch.cat([a, b], 0) as two stream-ordered device copies (memcpy nodes in a captured step, no ATen kernel)."""
    a, b = a.contiguous(), b.contiguous()
    if not (a.is_cuda and a.dtype == b.dtype and a.shape[1:] == b.shape[1:]):
        return torch.cat([a, b], dim=0)
    out = torch.empty((a.shape[0] + b.shape[0],) + tuple(a.shape[1:]), dtype=a.dtype, device=a.device)
    L.call('nr_concat2', L.ptr(a), a.numel() * a.element_size(), L.ptr(b), b.numel() * b.element_size(), L.ptr(out),
           L.stream())
    return out


class _SplitRows(torch.autograd.Function):
    """v -> (v[:k], v[k:]).  Autograd's own slicing backward materialises two zero-filled full-size gradients and
    adds them (2 fills + 2 copies + 1 add per step); the two incoming gradients are simply the halves of dv."""

    @staticmethod
    def forward(ctx, v, k):
        ctx.k, ctx.n, ctx.tail = k, v.shape[0], v.shape[1:]
        ctx.like = (v.dtype, v.device)
        return v[:k], v[k:]

    @staticmethod
    def backward(ctx, ga, gb):
        dt, dev = ctx.like
        if ga is None:
            ga = torch.zeros((ctx.k,) + tuple(ctx.tail), dtype=dt, device=dev)
        if gb is None:
            gb = torch.zeros((ctx.n - ctx.k,) + tuple(ctx.tail), dtype=dt, device=dev)
        return _concat_rows(ga, gb), None
