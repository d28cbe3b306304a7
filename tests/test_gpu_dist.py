"""Multi-GPU numeric checks on real devices ("N-rank grads equal 1-rank grads on the concatenated
batch"; VERDICT r1 missing #7).  Two ranks over NCCL, spawned with torch.multiprocessing like src/main.py:309:

  * the REFERENCE's loop shape over the plugin — optim.Adam(model.parameters()) created BEFORE .cuda(), then
    DistributedDataParallel(model, device_ids=[rank]), backward, step (src/main.py:76-82,105-110): DDP's averaged
    gradient == the 1-rank gradient on the concatenated batch; parameters bit-identical across ranks after 3 steps
  * the engine's own path (FlatAdam + one all-reduce of the flat bucket, eager and graph-replayed): same two checks

Skipped with fewer than two GPUs; `gpurun --gpus 2` runs it."""
import os
import socket

import numpy as np
import pytest
import torch

import synth
from test_gpu_ops import T
from test_gpu_models import build_model

pytestmark = pytest.mark.gpu
TAGS = {'nrms': 'nrms_ulm0', 'naml': 'naml_ulm0_c11'}


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _batch(kind, B):
    """A batch of B samples in the layout of the golden case (same table sizes), deterministic."""
    k, kw = synth.MODEL_CASES[TAGS[kind]]
    if k == 'nrms':
        return synth.batch(777, B, synth.T_, synth.C_, kw['L'], kw['V'])
    F = 1 + int(kw['use_cat']) + int(kw['use_sub'])
    return synth.batch(777, B, synth.T_, synth.C_, kw['L'], kw['N'] + 1, F=F)


def _worker(rank, world, port, kind, prec, mode, out_dir):
    import torch.distributed as dist
    os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'] = '127.0.0.1', str(port)
    if mode.endswith('_nccl'):          # force the NCCL all-reduce + local Adam path
        os.environ['NR_FUSED_ALLREDUCE'] = '0'
        mode = mode[:-5]
    if mode.endswith('_early'):         # early side-stream exchange of the largest parameter even for this tiny vocabulary
        os.environ['NR_EXCHANGE_OVERLAP'] = 'always'
        mode = mode[:-6]
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
    try:
        from newsrecommendation_b200 import engine
        from newsrecommendation_b200.model import NRMS, NAML
        from test_gpu_models import ref_args
        k, kw, P, _ = synth.case_inputs(TAGS[kind])
        if k == 'nrms':
            a = ref_args(user_log_mask=False, num_words_title=kw['L'], precision=prec)
            model = NRMS.Model(a, P['news_encoder.embedding_matrix.weight'], 0, 0)
        else:
            a = ref_args(user_log_mask=False, num_words_title=kw['L'], use_category=True, use_subcategory=True,
                         freeze_embedding=True, precision=prec)
            model = NAML.Model(a, P['news_encoder.title_embeddings.weight'], synth.NCAT, synth.NSUB)
        model.load_state_dict({kk: T(v) for kk, v in P.items()})
        if rank == 1:       # rank 1 starts from DIFFERENT weights: the rank-0 broadcast has to repair that
            with torch.no_grad():
                for p in model.parameters():
                    if p.requires_grad:
                        p.add_(0.01)
        B = 8
        hist, mask, cand, label = [T(x) for x in _batch(kind, B)]
        sl = slice(rank * B // world, (rank + 1) * B // world)                  # sample i -> contiguous shards
        mine = [t[sl].cuda() for t in (hist, mask, cand, label)]
        grads = {}
        if mode == 'reference_loop':
            from torch.nn.parallel import DistributedDataParallel
            optimizer = torch.optim.Adam(model.parameters(), lr=3e-4)          # main.py:76 (before .cuda())
            model = model.cuda(rank)                                           # main.py:79
            ddp = DistributedDataParallel(model, device_ids=[rank])            # main.py:82 (broadcasts rank 0)
            ddp.train()
            for step in range(3):
                bz_loss, y_hat = ddp(*mine)                                    # main.py:105
                optimizer.zero_grad()                                          # main.py:108
                bz_loss.backward()                                             # main.py:109
                if step == 0:
                    grads = {n: p.grad.detach().cpu().clone() for n, p in model.named_parameters() if p.grad is not None}
                optimizer.step()                                               # main.py:110
        else:
            model = model.cuda(rank).train()
            opt = engine.FlatAdam(model, lr=3e-4)
            opt.broadcast_params()
            stepper = engine.GraphedTrainStep(model, opt, mine) if mode == 'engine_graph' else None
            for step in range(3):
                if stepper is None:
                    opt.zero_grad()
                    loss, _ = model(*mine)
                    loss.backward()
                    if step == 0:      # the bucket BEFORE the all-reduce is local; capture the reduced mean instead
                        g = opt.flat_g.clone()
                        dist.all_reduce(g)
                        g /= world
                        grads = {n: g[o:o + p.numel()].view(p.shape).cpu().clone()
                                 for (n, p), o in zip([(n, p) for n, p in model.named_parameters() if p.requires_grad], opt.offsets)}
                    opt.step()
                else:
                    stepper(*mine)
            if stepper is not None:
                stepper.close()
        torch.cuda.synchronize()
        sd = {n: p.detach().cpu().clone() for n, p in model.named_parameters() if p.requires_grad}
        torch.save({'grads': grads, 'params': sd}, os.path.join(out_dir, f'rank{rank}.pt'))
        dist.barrier(device_ids=[rank])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('kind,prec,mode', [('nrms', 'fp32', 'reference_loop'), ('naml', 'bf16', 'reference_loop'),
                                            ('nrms', 'bf16', 'reference_loop'), ('nrms', 'fp32', 'engine_eager'),
                                            ('nrms', 'bf16', 'engine_graph')])
def test_two_ranks_match_one_rank(kind, prec, mode, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs (gpurun --gpus 2)')
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), kind, prec, mode, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(os.path.join(tmp_path, 'rank0.pt'))
    r1 = torch.load(os.path.join(tmp_path, 'rank1.pt'))
    # parameters bit-identical across ranks after 3 steps (rank 1 started 0.01 away: broadcast + identical updates)
    for n in r0['params']:
        assert torch.equal(r0['params'][n], r1['params'][n]), n
    if mode == 'engine_graph':
        return
    # averaged 2-rank gradient of step 1 == 1-rank gradient on the concatenated batch (drop_rate 0)
    _, kw, P, _, model = build_model(TAGS[kind], precision=prec)
    model.train()
    hist, mask, cand, label = [T(x).cuda() for x in _batch(kind, 8)]
    loss, _ = model(hist, mask, cand, label)
    loss.backward()
    gmax = max(float(p.grad.abs().max()) for p in model.parameters() if p.grad is not None)
    tol = 2e-5 if prec == 'fp32' else 2e-2
    checked = 0
    for n, p in model.named_parameters():
        if p.grad is None:
            continue
        g1 = p.grad.detach().cpu().numpy()
        for r in (r0, r1):
            np.testing.assert_allclose(r['grads'][n].numpy(), g1, rtol=0, atol=tol * float(np.abs(g1).max()) + 1e-5 * gmax,
                                       err_msg=n)
        checked += 1
    assert checked >= 10


@pytest.mark.parametrize('mode', ['engine_eager', 'engine_graph', 'engine_eager_early', 'engine_graph_early'])
def test_fused_exchange_matches_nccl_allreduce(mode, tmp_path):
    """The one-kernel gradient exchange (multimem reduce-scatter + Adam + all-gather over symmetric memory,
    csrc/nr_collective.cu; `_early`: the largest parameter's slice exchanged on a side stream from inside the backward)
    against NCCL all-reduce + local Adam (DDP semantics, src/main.py:82,109-110): with two
    ranks a + b is order-free, so the parameters after 3 steps must agree bit for bit (eager) / to fp32 rounding of
    the differently ordered dropout-free graph step."""
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs (gpurun --gpus 2)')
    import torch.multiprocessing as mp
    res = {}
    base = mode.replace('_early', '')
    for tag in (mode, base + '_nccl'):
        d = tmp_path / tag
        d.mkdir()
        mp.spawn(_worker, args=(2, _free_port(), 'nrms', 'bf16', tag, str(d)), nprocs=2, join=True)
        res[tag] = torch.load(os.path.join(d, 'rank0.pt'))['params']
        assert all(torch.equal(res[tag][n], torch.load(os.path.join(d, 'rank1.pt'))['params'][n]) for n in res[tag])
    moved = 0
    for n in res[mode]:
        if n.endswith('W_K.bias') or n.endswith('att_fc2.bias'):
            continue      # analytically zero gradients (key shift / logit shift cancel): Adam turns rounding noise into +-lr steps
        a, b = res[mode][n], res[base + '_nccl'][n]
        # bf16 weight-gradient kernels reduce with atomics (run-to-run order noise ~1e-7 of the gradient); Adam's
        # first steps move every weight by ~lr, so the two runs agree to a small fraction of lr
        d = (a - b).abs()
        # elements whose gradient is at rounding-noise level take +-lr steps of either sign under Adam's normalisation:
        # allow a small fraction of such outliers, all bounded by the 3-step displacement 3 * lr
        assert float((d > 3e-5).float().mean()) < 5e-3, (n, float((d > 3e-5).float().mean()))
        assert float(d.max()) <= 3 * 3e-4 + 1e-6, (n, float(d.max()))
        moved += int((a - T(synth.case_inputs('nrms_ulm0')[2][n])).abs().max() > 1e-4) if n in synth.case_inputs('nrms_ulm0')[2] else 0
    assert moved >= 10
