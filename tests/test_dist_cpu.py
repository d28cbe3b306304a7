"""Host-side multi-rank logic on CPU: world_size-2 gloo process group (the GPU path uses the same code
over NCCL).  Covers the gradient mean (DDP semantics, src/main.py:82,109), the sharded news-vector cache
+ all-gather (main.py:185-198 re-designed), impression sharding i % world (prepare_data.py:57-58) and
the metric reduction (main.py:268-277)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _StubModel(torch.nn.Module):
    """news_encoder(row ids) -> deterministic vectors; stands in for the CUDA encoder on CPU."""

    class _A:
        news_dim = 8

    def __init__(self):
        super().__init__()
        self.args = self._A()

    def news_encoder(self, x):
        base = x[:, :1].float()
        return base * torch.arange(1, 9).float()[None, :] + 0.5


def _worker(rank, world, port, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from newsrecommendation_b200 import engine
    res = {}
    # 1. gradient mean
    g = torch.arange(10, dtype=torch.float32) * (rank + 1)
    scale = engine.allreduce_gradients(g)
    res['grad'] = (g * scale).numpy()
    # 2. sharded news cache + all-gather, including a row count that does not divide
    news = torch.arange(37, dtype=torch.int32).reshape(37, 1)
    table = engine.build_news_cache(_StubModel(), news, batch=5)
    res['table'] = table.numpy()
    res['range'] = engine.shard_rows(37, world, rank)
    # 3. impression sharding + metric reduction
    idx = list(engine.impression_shard(11, world, rank))
    sums = torch.tensor([float(sum(idx)), 1.0, 2.0, 3.0, float(len(idx))], dtype=torch.float64)
    r = engine.reduce_eval_sums(sums, len(idx))
    res['idx'] = idx
    res['red'] = (r['sums'], r['scored'], r['total'])
    out[rank] = res
    dist.barrier()
    dist.destroy_process_group()


def test_world2_gloo_host_logic():
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    r0, r1 = out[0], out[1]
    np.testing.assert_allclose(r0['grad'], np.arange(10) * 1.5)          # mean of 1x and 2x
    np.testing.assert_allclose(r1['grad'], r0['grad'])
    ref = np.arange(37, dtype=np.float32)[:, None] * np.arange(1, 9, dtype=np.float32)[None, :] + 0.5
    np.testing.assert_array_equal(r0['table'], ref)                       # every rank holds the full table
    np.testing.assert_array_equal(r1['table'], ref)
    assert r0['range'] == (0, 19, 19) and r1['range'] == (19, 37, 19)
    assert sorted(r0['idx'] + r1['idx']) == list(range(11)) and r0['idx'] == [0, 2, 4, 6, 8, 10]
    s0, scored0, total0 = r0['red']
    np.testing.assert_allclose(s0, [55.0, 2.0, 4.0, 6.0])
    assert scored0 == 11 and total0 == 11


def test_shard_rows_edges():
    from newsrecommendation_b200 import engine
    assert engine.shard_rows(0, 4, 2) == (0, 0, 0)
    assert engine.shard_rows(3, 4, 3) == (3, 3, 1)
    cover = []
    for r in range(8):
        lo, hi, per = engine.shard_rows(65239, 8, r)
        cover += list(range(lo, hi))[:1] + [hi]
        assert hi - lo <= per
    assert cover[0] == 0 and cover[-1] == 65239
