"""Runtime around the model plugin: the train step of src/main.py:98-110 and the eval
path of src/main.py:185-277, both device-resident.

Train: parameters and gradients are re-homed into two flat fp32 buckets; one NCCL
all-reduce of the gradient bucket per step (DDP semantics, main.py:82,109) and one fused
Adam launch (torch.optim.Adam defaults, main.py:76) with the 1/world_size mean folded in.
`model.parameters()` remain ordinary leaf Parameters (views into the bucket), so the
reference's own `optim.Adam(model.parameters())` + DDP loop keeps working unchanged.

Eval: news-vector cache encoded in row shards + all-gather (reference: every rank encodes
everything, main.py:190-196); impressions as CSR, scored by gather+dot from the device
table; AUC/MRR/nDCG@5/10 per impression on device; 5-scalar all-reduce.
"""

import os

import torch
import torch.distributed as dist

from . import _lib as L
from . import ops


def _dist_on():
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def shard_rows(rows, world, rank):
    """Contiguous row range [lo, hi) of rank `rank` when `rows` rows are split over `world` ranks in
    equal slabs of `per` rows (the last slab may be short or empty)."""
    per = (rows + world - 1) // world
    return min(rows, rank * per), min(rows, (rank + 1) * per), per


def impression_shard(n_impressions, world, rank):
    """Impression i belongs to rank i % world (src/prepare_data.py:57-58)."""
    return range(rank, n_impressions, world)


def allreduce_gradients(flat_g):
    """DDP semantics (main.py:82,109): SUM all-reduce of the flat gradient bucket; returns the scale
    (1/world) the optimizer folds into its update so that the applied gradient is the rank mean."""
    if _dist_on():
        dist.all_reduce(flat_g, op=dist.ReduceOp.SUM)
        return 1.0 / dist.get_world_size()
    return 1.0


def _symmetric_buckets(count, dev):
    """(flat_p, flat_g, handle_p, handle_g) in NVLink symmetric memory with a multicast mapping, or None when the
    process group / fabric cannot provide one (then NCCL all-reduce + local Adam is the path)."""
    if not _dist_on() or os.environ.get('NR_FUSED_ALLREDUCE', '1') == '0':
        return None
    try:
        import torch.distributed._symmetric_memory as symm_mem
        p = symm_mem.empty(count, dtype=torch.float32, device=dev)
        g = symm_mem.empty(count, dtype=torch.float32, device=dev)
        hp = symm_mem.rendezvous(p, dist.group.WORLD)
        hg = symm_mem.rendezvous(g, dist.group.WORLD)
        ok = bool(hp.multicast_ptr) and bool(hg.multicast_ptr)
    except Exception as e:        # noqa: BLE001  (private torch API: any failure means "not available")
        import logging
        logging.getLogger(__name__).warning(f'symmetric-memory gradient exchange unavailable ({e!r}); using NCCL all-reduce')
        ok = False
    # every rank must take the same path
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if not int(flag.item()):
        return None
    p.zero_(); g.zero_()
    return p, g, hp, hg


class FlatAdam:
    """Flat-bucket Adam == torch.optim.Adam(params, lr) (betas .9/.999, eps 1e-8, wd 0).

    With a process group on an NVSwitch fabric the buckets live in symmetric memory and step() is barrier ->
    ONE kernel (multimem reduce-scatter + Adam on the rank's slice + multimem all-gather of the updated parameters,
    csrc/nr_collective.cu) -> barrier; Adam moments exist only for the rank's slice.  Everything in it is
    stream-ordered and capturable (`capturable` is True): GraphedTrainStep then replays the whole optimisation step.
    Otherwise: one NCCL all-reduce of the bucket + the local fused Adam."""

    def __init__(self, model, lr, betas=(0.9, 0.999), eps=1e-8):
        self.params = [p for p in model.parameters() if p.requires_grad]
        if not self.params:
            raise ValueError('no trainable parameters')
        dev = self.params[0].device
        if dev.type != 'cuda':
            raise RuntimeError('FlatAdam: move the model to a CUDA device first (no CPU path exists)')
        # every parameter starts on a 256-byte boundary (kernels use 128-bit loads on them)
        self.offsets, off = [], 0
        for p in self.params:
            self.offsets.append(off)
            off += (p.numel() + 63) // 64 * 64
        self.count = off
        self.symm = _symmetric_buckets(self.count, dev)
        if self.symm is not None:
            self.flat_p, self.flat_g, self.hdl_p, self.hdl_g = self.symm
            W, r = dist.get_world_size(), dist.get_rank()
            per = (self.count // 4 + W - 1) // W * 4                 # 4-float aligned slices
            self.lo, self.hi = min(self.count, r * per), min(self.count, (r + 1) * per)
        else:
            self.flat_p = torch.zeros(self.count, dtype=torch.float32, device=dev)
            self.flat_g = torch.zeros(self.count, dtype=torch.float32, device=dev)
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=dev)       # completed Adam steps, on the device (graph replay)
        # Early exchange: when one parameter dominates the bucket (NRMS: the dense (V,300) embedding gradient, 15 of
        # 19 MB) and sits at one end of it, its slice of the bucket is exchanged on a side stream as soon as the
        # backward has finished it (ops.TitleEncoderFn records `_nr_grad_ready_event` after enc.dx, before the
        # weight-gradient GEMM), overlapping the NVLink traffic with the rest of the backward.
        self.early = None
        if self.symm is not None and os.environ.get('NR_EXCHANGE_OVERLAP', '1') != '0':
            i = max(range(len(self.params)), key=lambda k: self.params[k].numel())
            big = self.params[i].numel() * 3 >= self.count or os.environ.get('NR_EXCHANGE_OVERLAP') == 'always'   # 'always': tests
            if big and i in (0, len(self.params) - 1):
                a = self.offsets[i]
                b = self.offsets[i + 1] if i + 1 < len(self.params) else self.count
                W, r = dist.get_world_size(), dist.get_rank()

                def sl(lo, hi):
                    per = ((hi - lo) // 4 + W - 1) // W * 4
                    return min(hi, lo + r * per), min(hi, lo + (r + 1) * per)
                self.early = {'param': self.params[i], 'slice': sl(a, b), 'rest': sl(b, self.count) if i == 0 else sl(0, a),
                              'side': torch.cuda.Stream(), 'ready': torch.cuda.Event(), 'join': torch.cuda.Event()}
                self.early['ready'].record()                      # materialise the CUDA event handle
                self.params[i]._nr_grad_ready_event = self.early['ready']
        self.m = torch.zeros_like(self.flat_p)
        self.v = torch.zeros_like(self.flat_p)
        for p, off in zip(self.params, self.offsets):
            n = p.numel()
            self.flat_p[off:off + n].copy_(p.data.reshape(-1))
            p.data = self.flat_p[off:off + n].view(p.shape)          # Parameter object unchanged
            p.grad = self.flat_g[off:off + n].view(p.shape)
            p._nr_fused_grad = True          # backward kernels add straight into the flat gradient bucket (ops._grad_target)
        self.lr, self.betas, self.eps = lr, betas, eps
        self.step_count = 0
        self.one = torch.ones((), dtype=torch.float32, device=dev)      # d(loss)/d(loss), see train_step

    def release(self):
        """Hand the Parameters back to ordinary autograd (another optimizer may take the model over)."""
        for p in self.params:
            p._nr_fused_grad = False

    def zero_grad(self):
        L.call('nr_zero_bytes', L.ptr(self.flat_g), self.flat_g.numel() * 4, L.stream())
        for p, off in zip(self.params, self.offsets):      # re-attach in case autograd replaced .grad
            g = self.flat_g[off:off + p.numel()].view(p.shape)
            if p.grad is None or p.grad.data_ptr() != g.data_ptr():
                p.grad = g

    def broadcast_params(self):
        """DDP constructor semantics (main.py:82): rank 0's weights everywhere."""
        if _dist_on():
            dist.broadcast(self.flat_p, src=0)

    @property
    def capturable(self):
        return self.symm is not None or not _dist_on()

    def _exchange(self, lo, hi, ch):
        b1, b2 = self.betas
        self.hdl_g.barrier(channel=ch)           # every rank's backward has written this part of its gradient bucket
        L.call('nr_adam_step_dev', None, self.hdl_g.multicast_ptr, L.ptr(self.flat_p), self.hdl_p.multicast_ptr,
               L.ptr(self.m), L.ptr(self.v), lo, hi, self.lr, b1, b2, self.eps, L.ptr(self.step_dev),
               1.0 / dist.get_world_size(), L.stream())
        self.hdl_p.barrier(channel=ch + 1)       # every rank's slice has landed in every copy of the parameters

    def step_on_stream(self):
        """The whole update as stream-ordered launches (no host-side state): capturable in a CUDA graph."""
        b1, b2 = self.betas
        if self.symm is None:
            L.call('nr_adam_step_dev', L.ptr(self.flat_g), None, L.ptr(self.flat_p), None, L.ptr(self.m), L.ptr(self.v),
                   0, self.count, self.lr, b1, b2, self.eps, L.ptr(self.step_dev), 1.0, L.stream())
        elif self.early is not None and getattr(self.early['param'], '_nr_grad_ready_recorded', False):
            e = self.early
            e['param']._nr_grad_ready_recorded = False
            main = torch.cuda.current_stream()
            e['side'].wait_event(e['ready'])     # recorded by the backward right after the dominant gradient was final
            with torch.cuda.stream(e['side']):
                self._exchange(*e['slice'], 2)
                e['join'].record()
            self._exchange(*e['rest'], 0)        # behind the whole backward, on the main stream
            main.wait_event(e['join'])
        else:
            self._exchange(self.lo, self.hi, 0)
        L.call('nr_seed_salt_bump', L.ptr(self.step_dev), L.stream())           # completed steps += 1

    def step(self):
        if self.symm is not None:
            self.step_count += 1
            self.step_on_stream()
            return
        scale = allreduce_gradients(self.flat_g)
        self.step_count += 1
        L.call('nr_adam_step', L.ptr(self.flat_p), L.ptr(self.flat_g), L.ptr(self.m), L.ptr(self.v), self.count,
               self.lr, self.betas[0], self.betas[1], self.eps, self.step_count, scale, L.stream())
        # keep the device-resident count (read by graph-replayed updates) in step with eager updates
        L.call('nr_seed_salt_bump', L.ptr(self.step_dev), L.stream())


def train_step(model, opt, history, history_mask, candidate, label):
    """One iteration of src/main.py:105-110.  Returns (loss, hits) device scalars."""
    opt.zero_grad()
    loss, _score = model(history, history_mask, candidate, label)
    loss.backward(opt.one)              # preallocated seed gradient: autograd's own ones_like would be an ATen fill
    opt.step()
    return loss.detach(), model.last_hits


class GraphedTrainStep:
    """The forward + backward of one train step (src/main.py:105-109) captured as ONE CUDA graph and
    replayed per batch; the gradient all-reduce and the fused Adam follow eagerly (two launches whose
    arguments — NCCL state, Adam's bias-correction step — change per call).

    Why: a step is ~100 kernel launches of 5-500 us; driven from Python the host needs ~1 ms per step,
    which is fully exposed whenever the caller reads the loss every step (main.py:106 accumulates it,
    bench.py's e2e arm `.item()`s it).  A replay costs the host ~20 us.

    Dropout: host-side seeds are frozen into the captured launches, so the library's device-resident
    seed salt (nr_set_seed_salt) is bumped inside the graph — every replay draws fresh masks, and the
    forward and backward kernels of one replay read the same value.

    The batch shape is fixed at capture; feed a ragged last batch through engine.train_step instead."""
    _live = None          # the instance whose salt the library currently points at

    def __init__(self, model, opt, example_batch, warmup=2):
        self.model, self.opt = model, opt
        # the update joins the graph when it has no host-side state: single GPU, or the symmetric-memory exchange
        self.tail_in_graph = bool(getattr(opt, 'capturable', False))
        self.static = [t.clone() for t in example_batch]
        dev = self.static[0].device
        if GraphedTrainStep._live is not None:
            raise RuntimeError('GraphedTrainStep: another instance still owns the library seed salt; close() it first')
        self.salt = torch.zeros(1, dtype=torch.int64, device=dev)
        L.call('nr_set_seed_salt', L.ptr(self.salt))
        GraphedTrainStep._live = self
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):          # warm-up off the capture stream (lazy kernel loading, allocator)
            for _ in range(warmup):
                self._fwd_bwd(tail=False)
            if self.tail_in_graph:
                # load the update's kernels before the capture without changing the optimiser state: one real
                # update, then everything it touched is put back (peers write into this rank's parameter copy, so
                # the restore is fenced by host barriers)
                keep = [t.clone() for t in (opt.flat_p, opt.m, opt.v, opt.step_dev)]
                opt.step_on_stream()
                torch.cuda.synchronize()
                if _dist_on():
                    dist.barrier()
                for t, k in zip((opt.flat_p, opt.m, opt.v, opt.step_dev), keep):
                    t.copy_(k)
                torch.cuda.synchronize()
                if _dist_on():
                    dist.barrier()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        n0 = L.lib.nr_launch_count()
        with torch.cuda.graph(self.graph):
            self._fwd_bwd()
        self.launches_per_replay = L.lib.nr_launch_count() - n0

    def _fwd_bwd(self, tail=True):
        L.call('nr_seed_salt_bump', L.ptr(self.salt), L.stream())
        self.opt.zero_grad()
        loss, score = self.model(*self.static)
        loss.backward(self.opt.one)
        self.loss, self.score, self.hits = loss.detach(), score.detach(), self.model.last_hits
        if tail and self.tail_in_graph:
            self.opt.step_on_stream()

    def __call__(self, history, history_mask, candidate, label):
        for s, t in zip(self.static, (history, history_mask, candidate, label)):
            if s.shape != t.shape:
                raise ValueError(f'GraphedTrainStep captured batch shape {tuple(s.shape)}, got {tuple(t.shape)}')
            s.copy_(t, non_blocking=True)
        self.graph.replay()
        if self.tail_in_graph:
            self.opt.step_count += 1
        else:
            self.opt.step()
        return self.loss, self.hits

    def close(self):
        """Unregister the device salt (the library keeps a raw pointer to it).  Idempotent; also runs from __del__ so
        that a step dropped without close() — an exception, a test — never leaves the library reading freed memory."""
        if GraphedTrainStep._live is self:
            GraphedTrainStep._live = None
            L.call('nr_set_seed_salt', None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# --------------------------------------------------------------------------------------
# eval path
# --------------------------------------------------------------------------------------
@torch.no_grad()
def build_news_cache(model, news_combined, batch=16384, shard=True):
    """src/main.py:185-198: news_encoder over EVERY row of news_combined (row 0 = the pad
    news included) -> device table (N+1, news_dim) fp32.  With a process group, rank r
    encodes the contiguous row range [r*per, (r+1)*per) and the shards are all-gathered
    in place."""
    was_training = model.training
    model.eval()
    rows = news_combined.shape[0]
    dev = news_combined.device
    H = model.args.news_dim
    W = dist.get_world_size() if (shard and _dist_on()) else 1
    r = dist.get_rank() if W > 1 else 0
    lo, hi, per = shard_rows(rows, W, r)
    table = torch.empty(per * W, H, dtype=torch.float32, device=dev)
    for b0 in range(lo, hi, batch):
        b1 = min(hi, b0 + batch)
        table[b0:b1] = model.news_encoder(news_combined[b0:b1])
    if hi < (r + 1) * per:
        table[hi:(r + 1) * per].zero_()
    if W > 1:
        # in place: rank r's slab already sits at its final offset of the output (NCCL in-place all-gather layout)
        dist.all_gather_into_tensor(table, table[r * per:(r + 1) * per])
    if was_training:
        model.train()
    return table[:rows]


@torch.no_grad()
def score_impressions(model, table, hist_idx, hist_mask, offsets, cand_idx, labels, batch=8192):
    """src/main.py:240-263 + src/dataset.py:64-74 on device.
    Impressions as CSR (offsets int64 (I+1), cand_idx int32, labels int8) with front-padded
    history indices hist_idx int32 (I,T) and hist_mask fp32 (I,T).
    Returns (scores fp32 flat, metrics fp64 (I,4), valid uint8 (I), sums fp64 (5):
    sum AUC, MRR, nDCG@5, nDCG@10 over scored impressions and their count)."""
    was_training = model.training
    model.eval()
    I, T = hist_idx.shape
    dev = table.device
    H = table.shape[1]
    total = cand_idx.numel()
    table = L.f32(table)
    user = torch.empty(I, H, dtype=torch.float32, device=dev)
    gathered = getattr(model.user_encoder, 'forward_gathered', None)
    for b0 in range(0, I, batch):
        b1 = min(I, b0 + batch)
        u = gathered(table, hist_idx[b0:b1], hist_mask[b0:b1]) if gathered else None
        if u is None:
            log_vecs = ops.gather_rows(table, hist_idx[b0:b1])             # dataset.py:67-68
            u = model.user_encoder(log_vecs, hist_mask[b0:b1])             # main.py:247
        user[b0:b1] = u
    scores = torch.empty(total, dtype=torch.float32, device=dev)
    metrics = torch.empty(I, 4, dtype=torch.float64, device=dev)
    valid = torch.empty(I, dtype=torch.uint8, device=dev)
    sums = torch.zeros(5, dtype=torch.float64, device=dev)
    offsets = offsets.contiguous().to(torch.int64)
    cand_idx = cand_idx.contiguous().to(torch.int32)
    labels = labels.contiguous().to(torch.int8)
    st = L.stream()
    L.call('nr_eval_score', L.ptr(table), table.shape[0], L.ptr(user), L.ptr(offsets), L.ptr(cand_idx),
           L.ptr(scores), I, total, H, st)                                  # main.py:253
    L.call('nr_eval_rank_metrics', L.ptr(scores), L.ptr(labels), L.ptr(offsets), L.ptr(metrics), L.ptr(valid),
           I, st)                                                           # main.py:250-258
    L.call('nr_eval_reduce_metrics', L.ptr(metrics), L.ptr(valid), L.ptr(sums), I, st)
    if was_training:
        model.train()
    return scores, metrics, valid, sums


def reduce_eval_sums(sums, n_impressions):
    """src/main.py:268-277: SUM over ranks of the 4 metric sums, the scored count and the
    total impression count.  Returns dict with both denominators."""
    t = torch.cat([sums, torch.tensor([float(n_impressions)], dtype=torch.float64, device=sums.device)])
    if _dist_on():
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    t = t.cpu()
    scored, total = float(t[4]), float(t[5])
    return {'sums': t[:4].numpy(), 'scored': int(scored), 'total': int(total),
            'mean_scored': (t[:4] / max(1.0, scored)).numpy(), 'mean_total': (t[:4] / max(1.0, total)).numpy()}
