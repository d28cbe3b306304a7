"""torch.autograd.Function shims over the C ABI (include/nr_sm100.h).

Each Function is the host-side mirror of one reference operator:
  AttentionPooling / MultiHeadSelfAttention (src/model/model_utils.py),
  nn.Conv1d(k=3,pad=1) and the category/subcategory dense views (src/model/NAML.py),
  the pad_doc blend (NRMS.py:59-60), bmm score + CrossEntropyLoss (NRMS.py:93-94).
Every byte of arithmetic happens inside libnr_sm100.so; there is no eager/PyTorch
fallback: non-CUDA tensors raise (see _lib.ptr).
"""
import ctypes as C
import itertools

import torch

from . import _lib as L

_seed_counter = itertools.count(1)

# Fused weight-gradient accumulation.  engine.FlatAdam tags the Parameters it owns (`p._nr_fused_grad = True`): for
# those, a backward kernel adds the gradient straight into the existing `.grad` buffer (a view of the optimizer's flat
# gradient bucket) and the autograd Function returns None for it — no per-parameter zero-fill + add launches, and the
# dense (V,300) embedding gradient is never materialised twice.  Every other Parameter (e.g. a model driven by the
# reference's own optim.Adam + DistributedDataParallel loop, whose hooks need the returned gradients) gets an ordinary
# gradient tensor back.  The switch is per Parameter, not per process: two models with different optimizers coexist.
def _grad_target(param, like):
    """(buffer to accumulate into, value to return to autograd) for one parameter."""
    if param is None:
        return None, None
    if (getattr(param, '_nr_fused_grad', False) and isinstance(param, torch.nn.Parameter) and param.grad is not None
            and param.grad.is_contiguous() and param.grad.dtype == torch.float32
            and param.grad.shape == param.shape and param.grad.is_cuda):
        return param.grad, None
    buf = torch.zeros_like(like)
    return buf, buf


def next_seed():
    """Fresh dropout seed per forward call, derived from torch's seed so that
    `torch.manual_seed(args.seed)` (src/main.py:285) makes runs repeatable."""
    return (torch.initial_seed() * 0x9E3779B97F4A7C15 + next(_seed_counter) * 0xD1B54A32D192ED03) & (2**64 - 1)


def _zeros_like(t):
    return torch.zeros_like(t, memory_format=torch.contiguous_format)


def gather_rows(table, ids):
    """out[i,:] = table[ids[i],:]  (bit exact; A1/A2 and the eval gathers)."""
    table = L.f32(table)
    ids = ids.contiguous().to(torch.int32)
    width = table.shape[1]
    out = torch.empty(ids.shape + (width,), dtype=torch.float32, device=table.device)
    L.call('nr_gather_rows', L.ptr(table), table.shape[0], L.ptr(ids), ids.numel(), width, L.ptr(out),
           L.stream())
    return out


def dropout_keep_mask(count, p, seed, stream_id, device, row_width, seg_width=None):
    """Keep mask of a dropout stream in the (rows, row_width) layout of the activation (seg_width = head width for the
    output stream of the MHSA; include/nr_sm100.h)."""
    keep = torch.empty(count, dtype=torch.uint8, device=device)
    L.call('nr_dropout_keep_mask', L.ptr(keep), count, float(p), C.c_uint64(seed), stream_id, int(row_width),
           int(seg_width or row_width), L.stream())
    return keep


class AttentionPoolingFn(torch.autograd.Function):
    """model_utils.py:13-31."""

    @staticmethod
    def forward(ctx, x, mask, w1, b1, w2, b2, prec):
        x = L.f32(x)
        n, S, H = x.shape
        Q = w1.shape[0]
        mask = None if mask is None else L.f32(mask)
        w1c, b1c, w2c, b2c = L.f32(w1), L.f32(b1), L.f32(w2), L.f32(b2)
        out = torch.empty(n, H, dtype=torch.float32, device=x.device)
        e = torch.empty(n, S, Q, dtype=torch.float32, device=x.device)
        alpha = torch.empty(n, S, dtype=torch.float32, device=x.device)
        ws = L.workspace(0, x.device)
        L.call('nr_attention_pooling_fwd', L.ptr(x), L.ptr(mask), L.ptr(w1c), L.ptr(b1c), L.ptr(w2c), L.ptr(b2c),
               L.ptr(out), L.ptr(e), L.ptr(alpha), n, S, H, Q, prec, L.ptr(ws), ws.numel(), L.stream())
        ctx.save_for_backward(x, w1c, w2c, e, alpha)
        ctx.prec = prec
        ctx.shapes = (w1.shape, b1.shape, w2.shape, b2.shape)
        return out

    @staticmethod
    def backward(ctx, g):
        x, w1, w2, e, alpha = ctx.saved_tensors
        n, S, H = x.shape
        Q = w1.shape[0]
        g = L.f32(g)
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        dw1 = torch.zeros(Q, H, dtype=torch.float32, device=x.device)
        db1 = torch.zeros(Q, dtype=torch.float32, device=x.device)
        dw2 = torch.zeros(Q, dtype=torch.float32, device=x.device)
        db2 = torch.zeros(1, dtype=torch.float32, device=x.device)
        ws = L.workspace(L.lib.nr_attention_pooling_workspace(n, S, H, Q, ctx.prec), x.device)
        L.call('nr_attention_pooling_bwd', L.ptr(g), L.ptr(x), L.ptr(w1), L.ptr(w2), L.ptr(e), L.ptr(alpha),
               L.ptr(dx), L.ptr(dw1), L.ptr(db1), L.ptr(dw2), L.ptr(db2), n, S, H, Q, ctx.prec, L.ptr(ws),
               ws.numel(), L.stream())
        s1, sb1, s2, sb2 = ctx.shapes
        return dx, None, dw1.view(s1), db1.view(sb1), dw2.view(s2), db2.view(sb2), None


def attention_pooling(x, mask, w1, b1, w2, b2, prec=L.PREC_FP32):
    return AttentionPoolingFn.apply(x, mask, w1, b1, w2, b2, prec)


class _Rows:
    """Python-side description of an nr_rows_t (see include/nr_sm100.h)."""

    def __init__(self, mode, n, S, D, dense=None, table=None, ids=None, ids_stride=0, drop_p=0.0, seed=0):
        self.mode, self.n, self.S, self.D = mode, n, S, D
        self.dense, self.table, self.ids, self.ids_stride = dense, table, ids, ids_stride
        self.drop_p, self.seed = drop_p, seed

    def c(self):
        return L.make_rows(self.mode, dense=self.dense, table=self.table, ids=self.ids,
                           ids_stride=self.ids_stride, drop_p=self.drop_p, seed=self.seed)


def _make_rows(src, ids, mode, S, D, drop_p, seed):
    """src: dense activation (n,S,D) for ROWS_DENSE, else the embedding table."""
    if mode == L.ROWS_DENSE:
        src = L.f32(src)
        return _Rows(mode, src.shape[0], S, D, dense=src, drop_p=drop_p, seed=seed)
    ids = ids.contiguous()
    if ids.dtype != torch.int32:
        ids = ids.to(torch.int32)
    table = L.f32(src)
    return _Rows(mode, ids.shape[0], S, D, table=table, ids=ids, ids_stride=ids.shape[1], drop_p=drop_p, seed=seed)


class MHSAFn(torch.autograd.Function):
    """MultiHeadSelfAttention with Q=K=V=X (model_utils.py:58-95), X read through
    a row source (dense / token gather / per-news gather) with fused input dropout,
    and the NRMS output dropout (NRMS.py:32-34) fused into the context store."""

    @staticmethod
    def forward(ctx, src, ids, key_mask, wq, bq, wk, bk, wv, bv, mode, S, Dm, heads, in_drop, out_drop, seed,
                prec):
        rows = _make_rows(src, ids, mode, S, Dm, in_drop, seed)
        n = rows.n
        H = wq.shape[0]
        dk = H // heads
        key_mask = None if key_mask is None else L.f32(key_mask)
        ws_ = [L.f32(t) for t in (wq, bq, wk, bk, wv, bv)]
        dev = ws_[0].device
        out = torch.empty(n, S, H, dtype=torch.float32, device=dev)
        qkv = torch.empty(n, S, 3 * H, dtype=torch.float32, device=dev)
        ws = L.workspace(0, dev)
        r = rows.c()
        L.call('nr_mhsa_fwd', C.byref(r), L.ptr(key_mask), *[L.ptr(t) for t in ws_], L.ptr(out), L.ptr(qkv),
               n, S, Dm, heads, dk, float(out_drop), prec, L.ptr(ws), ws.numel(), L.stream())
        ctx.rows, ctx.key_mask, ctx.w, ctx.qkv = rows, key_mask, ws_, qkv
        ctx.cfg = (n, S, Dm, heads, dk, float(out_drop), prec)
        ctx.src_shape = src.shape
        return out

    @staticmethod
    def backward(ctx, dctx):
        rows, key_mask, w, qkv = ctx.rows, ctx.key_mask, ctx.w, ctx.qkv
        n, S, Dm, heads, dk, out_drop, prec = ctx.cfg
        H = heads * dk
        dev = qkv.device
        dctx = L.f32(dctx)
        dwq, dwk, dwv = (torch.zeros(H, Dm, dtype=torch.float32, device=dev) for _ in range(3))
        dbq, dbk, dbv = (torch.zeros(H, dtype=torch.float32, device=dev) for _ in range(3))
        dx = dtable = None
        if ctx.needs_input_grad[0]:
            if rows.mode == L.ROWS_DENSE:
                dx = torch.empty(n, S, Dm, dtype=torch.float32, device=dev)
            else:
                dtable = torch.zeros(ctx.src_shape, dtype=torch.float32, device=dev)
        ws = L.workspace(L.lib.nr_mhsa_workspace(n, S, Dm, H, prec), dev)
        r = rows.c()
        L.call('nr_mhsa_bwd', C.byref(r), L.ptr(key_mask), L.ptr(w[0]), L.ptr(w[2]), L.ptr(w[4]), L.ptr(qkv),
               L.ptr(dctx), L.ptr(dwq), L.ptr(dbq), L.ptr(dwk), L.ptr(dbk), L.ptr(dwv), L.ptr(dbv), L.ptr(dx),
               L.ptr(dtable), n, S, Dm, heads, dk, out_drop, prec, L.ptr(ws), ws.numel(), L.stream())
        dsrc = dx.view(ctx.src_shape) if dx is not None else dtable
        return (dsrc, None, None, dwq, dbq, dwk, dbk, dwv, dbv) + (None,) * 8


def mhsa(src, ids, key_mask, wq, bq, wk, bk, wv, bv, *, mode, S, Dm, heads, in_drop=0.0, out_drop=0.0, seed=0,
         prec=L.PREC_FP32):
    return MHSAFn.apply(src, ids, key_mask, wq, bq, wk, bk, wv, bv, mode, S, Dm, heads, in_drop, out_drop, seed,
                        prec)


class Conv1dK3Fn(torch.autograd.Function):
    """nn.Conv1d(D->H, kernel_size=3, padding=1) along the token axis (NAML.py:27-32,54)
    over a row source, with the input dropout of NAML.py:51-53 fused in."""

    @staticmethod
    def forward(ctx, src, ids, w, b, mode, Lw, D, in_drop, seed, prec):
        rows = _make_rows(src, ids, mode, Lw, D, in_drop, seed)
        n = rows.n
        H = w.shape[0]
        wc, bc = L.f32(w), L.f32(b)
        y = torch.empty(n, Lw, H, dtype=torch.float32, device=wc.device)
        ws = L.workspace(L.lib.nr_conv1d_k3_workspace(n, Lw, D, H, prec), wc.device)
        r = rows.c()
        L.call('nr_conv1d_k3_fwd', C.byref(r), L.ptr(wc), L.ptr(bc), L.ptr(y), n, Lw, D, H, prec, L.ptr(ws),
               ws.numel(), L.stream())
        ctx.rows, ctx.w = rows, wc
        ctx.cfg = (n, Lw, D, H, prec)
        ctx.src_shape = src.shape
        return y

    @staticmethod
    def backward(ctx, dy):
        rows, w = ctx.rows, ctx.w
        n, Lw, D, H, prec = ctx.cfg
        dev = w.device
        dy = L.f32(dy)
        dw = torch.zeros_like(w)
        db = torch.zeros(H, dtype=torch.float32, device=dev)
        dx = dtable = None
        if ctx.needs_input_grad[0]:
            if rows.mode == L.ROWS_DENSE:
                dx = torch.empty(n, Lw, D, dtype=torch.float32, device=dev)
            else:
                dtable = torch.zeros(ctx.src_shape, dtype=torch.float32, device=dev)
        ws = L.workspace(L.lib.nr_conv1d_k3_workspace(n, Lw, D, H, prec), dev)
        r = rows.c()
        L.call('nr_conv1d_k3_bwd', C.byref(r), L.ptr(w), L.ptr(dy), L.ptr(dw), L.ptr(db), L.ptr(dx), L.ptr(dtable),
               n, Lw, D, H, prec, L.ptr(ws), ws.numel(), L.stream())
        dsrc = dx.view(ctx.src_shape) if dx is not None else dtable
        return (dsrc, None, dw, db) + (None,) * 6


def conv1d_k3(src, ids, w, b, *, mode, L_, D, in_drop=0.0, seed=0, prec=L.PREC_FP32):
    return Conv1dK3Fn.apply(src, ids, w, b, mode, L_, D, in_drop, seed, prec)


class EmbedDenseFn(torch.autograd.Function):
    """dense(emb[ids[:, col]]) — NAML category / subcategory views (NAML.py:60-68)."""

    @staticmethod
    def forward(ctx, emb, ids, col, w, b):
        embc, wc, bc = L.f32(emb), L.f32(w), L.f32(b)
        ids = ids.contiguous()
        if ids.dtype != torch.int32:
            ids = ids.to(torch.int32)
        n, F = ids.shape
        E, H = embc.shape[1], wc.shape[0]
        out = torch.empty(n, H, dtype=torch.float32, device=wc.device)
        idp = ids.data_ptr() + 4 * col
        ws = L.workspace(L.lib.nr_embed_dense_workspace(embc.shape[0], H), wc.device)
        L.call('nr_embed_dense_fwd', L.ptr(embc), embc.shape[0], idp, F, L.ptr(wc), L.ptr(bc), L.ptr(out), H, n,
               E, H, L.ptr(ws), ws.numel(), L.stream())
        ctx.saved = (embc, ids, col, wc)
        ctx.params = (emb, w, b)
        return out

    @staticmethod
    def backward(ctx, dout):
        emb, ids, col, w = ctx.saved
        n, F = ids.shape
        E, H = emb.shape[1], w.shape[0]
        # the gradient of torch.stack(views, 1) arrives as a row-strided slice: the kernels take the row stride, so
        # no contiguous copy (11 MB per view at B = 128)
        if not (dout.is_cuda and dout.dtype == torch.float32 and dout.dim() == 2 and dout.stride(1) == 1 and dout.stride(0) >= H
                and dout.stride(0) % 4 == 0 and dout.data_ptr() % 16 == 0):
            dout = L.f32(dout)
        ld = dout.stride(0)
        pe, pw, pb = ctx.params
        (bemb, demb), (bw, dw) = _grad_target(pe, emb), _grad_target(pw, w)
        bb, db = _grad_target(pb, torch.empty(H, dtype=torch.float32, device=w.device))
        ws = L.workspace(L.lib.nr_embed_dense_workspace(emb.shape[0], H), w.device)
        L.call('nr_embed_dense_bwd', L.ptr(emb), emb.shape[0], ids.data_ptr() + 4 * col, F, L.ptr(w), dout.data_ptr(),
               ld, L.ptr(bw), L.ptr(bb), L.ptr(bemb), n, E, H, L.ptr(ws), ws.numel(), L.stream())
        ctx.params = None
        return demb, None, None, dw, db


def embed_dense(emb, ids, col, w, b):
    return EmbedDenseFn.apply(emb, ids, col, w, b)


class PadBlendFn(torch.autograd.Function):
    """v*m + pad_doc*(1-m)  (NRMS.py:59-60 / NAML.py:94-95)."""

    @staticmethod
    def forward(ctx, v, mask, pad):
        v, mask, padc = L.f32(v), L.f32(mask), L.f32(pad)
        n, T, H = v.shape
        out = torch.empty_like(v)
        L.call('nr_pad_blend_fwd', L.ptr(v), L.ptr(mask), L.ptr(padc), L.ptr(out), n, T, H, L.stream())
        ctx.save_for_backward(mask)
        ctx.pad_shape = pad.shape
        ctx.pad_param = pad
        return out

    @staticmethod
    def backward(ctx, dout):
        (mask,) = ctx.saved_tensors
        dout = L.f32(dout)
        n, T, H = dout.shape
        dv = torch.empty_like(dout)
        buf, dpad = _grad_target(ctx.pad_param, torch.empty(ctx.pad_shape, dtype=torch.float32, device=dout.device))
        L.call('nr_pad_blend_bwd', L.ptr(dout), L.ptr(mask), L.ptr(dv), L.ptr(buf), n, T, H, L.stream())
        ctx.pad_param = None
        return dv, None, dpad


def pad_blend(v, mask, pad):
    return PadBlendFn.apply(v, mask, pad)


class ScoreCEFn(torch.autograd.Function):
    """bmm score + nn.CrossEntropyLoss (NRMS.py:93-94) + train accuracy hits (utils.py:36-40)."""

    @staticmethod
    def forward(ctx, cand, user, label):
        cand, user = L.f32(cand), L.f32(user)
        label = label.contiguous().to(torch.int64)
        B, Cn, H = cand.shape
        dev = cand.device
        score = torch.empty(B, Cn, dtype=torch.float32, device=dev)
        lh = torch.empty(2, dtype=torch.float32, device=dev)
        per = torch.empty(B, 2, dtype=torch.float32, device=dev)
        L.call('nr_score_ce_fwd', L.ptr(cand), L.ptr(user), L.ptr(label), L.ptr(score), lh.data_ptr(),
               lh.data_ptr() + 4, L.ptr(per), B, Cn, H, L.stream())
        ctx.save_for_backward(cand, user, label, score)
        ctx.mark_non_differentiable(score)
        loss, hits = lh[0], lh[1]
        ctx.mark_non_differentiable(hits)
        return loss, score, hits

    @staticmethod
    def backward(ctx, gloss, _gscore, _ghits):
        cand, user, label, score = ctx.saved_tensors
        B, Cn, H = cand.shape
        gl = L.f32(gloss).reshape(1)
        dcand = torch.empty_like(cand)
        duser = torch.empty_like(user)
        L.call('nr_score_ce_bwd', L.ptr(cand), L.ptr(user), L.ptr(label), L.ptr(score), L.ptr(gl), L.ptr(dcand),
               L.ptr(duser), B, Cn, H, L.stream())
        return dcand, duser, None


def score_ce(cand, user, label):
    return ScoreCEFn.apply(cand, user, label)


class TitleEncoderFn(torch.autograd.Function):
    """bf16 tensor-core title encoder (nr_title_encoder_fwd/_bwd): the whole
    gather -> dropout -> {MHSA -> dropout | Conv1d(k=3)} -> AttentionPooling chain of
    NRMS.py:27-35 / NAML.py:47-56, and the NRMS user encoder (NRMS.py:55-62) over dense rows."""

    @staticmethod
    def forward(ctx, src, ids, mask, w0, b0, w1, b1, w2, b2, p_w1, p_b1, p_w2, p_b2, cfg):
        kind, mode, S, Dm, heads, in_drop, out_drop, seed, infer = cfg
        rows = _make_rows(src, ids, mode, S, Dm, in_drop, seed)
        n = rows.n
        ws_ = [None if t is None else L.f32(t) for t in (w0, b0, w1, b1, w2, b2)]
        pool = [L.f32(t) for t in (p_w1, p_b1, p_w2, p_b2)]
        H, Q = (ws_[0].shape[0] if ws_[0] is not None else Dm), pool[0].shape[0]
        mask = None if mask is None else L.f32(mask)
        dev = pool[0].device
        enc = L.NrTitleEncoder()
        enc.kind, enc.S, enc.Dm, enc.H, enc.Q, enc.heads = kind, S, Dm, H, Q, heads
        for i in range(3):
            enc.w[i] = L.ptr(ws_[2 * i])
            enc.b[i] = L.ptr(ws_[2 * i + 1])
        enc.p_w1, enc.p_b1, enc.p_w2, enc.p_b2 = [L.ptr(t) for t in pool]
        enc.mask = L.ptr(mask)
        enc.out_drop_p = float(out_drop)
        out = torch.empty(n, H, dtype=torch.float32, device=dev)
        r = rows.c()
        if infer and in_drop == 0.0 and out_drop == 0.0:
            # inference (model.eval() under no_grad: the news-vector cache): nothing is saved, one fused kernel
            nb = L.lib.nr_title_encoder_infer_bytes(C.byref(enc), n)
            if nb:
                ws = L.workspace(nb, dev)
                L.call('nr_title_encoder_fwd', C.byref(enc), C.byref(r), n, L.ptr(out), None, 0, L.ptr(ws), ws.numel(),
                       L.stream())
                return out
        save = torch.empty(L.lib.nr_title_encoder_save_bytes(C.byref(enc), n), dtype=torch.uint8, device=dev)
        ws = L.workspace(0, dev)
        L.call('nr_title_encoder_fwd', C.byref(enc), C.byref(r), n, L.ptr(out), L.ptr(save), save.numel(),
               L.ptr(ws), ws.numel(), L.stream())
        ctx.enc, ctx.rows, ctx.keep = enc, rows, (ws_, pool, mask, save)
        ctx.params = ((w0, b0, w1, b1, w2, b2), (p_w1, p_b1, p_w2, p_b2), src)
        ctx.src_shape = src.shape
        ctx.n = n
        return out

    @staticmethod
    def backward(ctx, g):
        enc, rows, (ws_, pool, mask, save) = ctx.enc, ctx.rows, ctx.keep
        n = ctx.n
        dev = save.device
        g = L.f32(g)
        gr = L.NrTitleEncoderGrads()
        pw, pp, psrc = ctx.params
        tw = [_grad_target(p, t) for p, t in zip(pw, ws_)]
        tp = [_grad_target(p, t) for p, t in zip(pp, pool)]
        dws, dpool = [r for _, r in tw], [r for _, r in tp]
        for i in range(3):
            gr.dw[i] = L.ptr(tw[2 * i][0])
            gr.db[i] = L.ptr(tw[2 * i + 1][0])
        gr.dp_w1, gr.dp_b1, gr.dp_w2, gr.dp_b2 = [L.ptr(b) for b, _ in tp]
        dsrc = None
        if ctx.needs_input_grad[0]:
            if rows.mode == L.ROWS_DENSE:
                dsrc = torch.empty(ctx.src_shape, dtype=torch.float32, device=dev)
                gr.dx_dense = L.ptr(dsrc)
            else:
                buf, dsrc = _grad_target(psrc, rows.table)
                gr.dtable = L.ptr(buf)
                ev = getattr(psrc, '_nr_grad_ready_event', None)      # engine.FlatAdam: overlap the exchange of this slice
                if ev is not None:
                    gr.dtable_ready = ev.cuda_event
                    psrc._nr_grad_ready_recorded = True
        ws = L.workspace(L.lib.nr_title_encoder_workspace(C.byref(enc), n), dev)
        r = rows.c()
        L.call('nr_title_encoder_bwd', C.byref(enc), C.byref(r), n, L.ptr(g), L.ptr(save), save.numel(), C.byref(gr),
               L.ptr(ws), ws.numel(), L.stream())
        ctx.keep = ctx.params = None
        return (dsrc, None, None, *dws, *dpool, None)


def title_encoder(kind, src, ids, mask, weights, pool, *, mode, S, Dm, heads=1, in_drop=0.0, out_drop=0.0, seed=0):
    """weights: MHSA (W_Q, b_Q, W_K, b_K, W_V, b_V); CONV (conv.weight, conv.bias).
    pool: (att_fc1.weight, att_fc1.bias, att_fc2.weight, att_fc2.bias)."""
    w = list(weights) + [None] * (6 - len(weights))
    # grad mode is read HERE: inside autograd.Function.forward it is always off
    infer = not torch.is_grad_enabled()
    return TitleEncoderFn.apply(src, ids, mask, *w, *pool, (kind, mode, S, Dm, heads, in_drop, out_drop, seed, infer))
