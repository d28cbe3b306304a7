"""Device-side batcher against the reference DatasetTrain / DatasetTest line mappers.

tests/golden/batcher.npz holds the outputs of the UNMODIFIED reference classes on
tests/golden/batcher_{train,test}.tsv (make_golden_batcher.py).  Integer work: everything bit exact."""
import os
import random

import numpy as np
import pytest
import torch

from oracle import nr_oracle as O

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def _gold():
    g = np.load(os.path.join(G, 'batcher.npz'))
    news_index = {'N%d' % i: i for i in range(1, int(g['N']) + 1)}
    return g, news_index


def test_oracle_line_mappers_match_reference():
    g, news_index = _gold()
    T = int(g['T'])
    for i, line in enumerate(open(os.path.join(G, 'batcher_train.tsv'))):
        h, m, c, l = O.train_line_mapper(line, news_index, g['news_combined'], T, int(g['label'][i]))
        assert np.array_equal(h, g['history'][i]) and np.array_equal(m, g['mask'][i])
        assert np.array_equal(c, g['candidate'][i])
    off = g['test_cand_off']
    for i, line in enumerate(open(os.path.join(G, 'batcher_test.tsv'))):
        h, m, c, lab = O.test_line_mapper(line, news_index, T)
        assert np.array_equal(h, g['test_hist'][i]) and np.array_equal(m, g['test_mask'][i])
        assert np.array_equal(c, g['test_cand'][off[i]:off[i + 1]])
        assert np.array_equal(lab, g['test_labels'][off[i]:off[i + 1]])


def test_index_builders_and_numpy_oracle(tmp_path):
    from newsrecommendation_b200 import batcher
    g, news_index = _gold()
    T, K = int(g['T']), int(g['K'])
    idx = batcher.TrainIndex.from_tsv(os.path.join(G, 'batcher_train.tsv'), news_index, K)
    assert len(idx) == g['label'].shape[0] and idx.npratio == K
    idx.save(tmp_path / 'tr.npz')
    idx = batcher.TrainIndex.load(tmp_path / 'tr.npz')
    h, m, c, l = O.batch_train_np(idx.hist_off, idx.hist_idx, idx.pos, idx.neg, np.arange(len(idx)), g['label'],
                                  g['news_combined'], T)
    assert np.array_equal(h, g['history']) and np.array_equal(m, g['mask']) and np.array_equal(c, g['candidate'])
    ev = batcher.EvalIndex.from_tsv(os.path.join(G, 'batcher_test.tsv'), news_index)
    ev.save(tmp_path / 'te.npz')
    ev = batcher.EvalIndex.load(tmp_path / 'te.npz')
    assert np.array_equal(ev.cand_off, g['test_cand_off']) and np.array_equal(ev.cand_idx, g['test_cand'])
    assert np.array_equal(ev.labels, g['test_labels'])
    for i in range(len(ev)):
        h, m = O.pad_history(ev.hist_idx[ev.hist_off[i]:ev.hist_off[i + 1]], T)
        assert np.array_equal(h, g['test_hist'][i]) and np.array_equal(m, g['test_mask'][i])
    with pytest.raises(ValueError):
        batcher.TrainIndex.from_tsv(os.path.join(G, 'batcher_train.tsv'), news_index, K + 1)


def test_batcher_refuses_cpu():
    from newsrecommendation_b200 import batcher
    g, news_index = _gold()
    idx = batcher.TrainIndex.from_tsv(os.path.join(G, 'batcher_train.tsv'), news_index, int(g['K']))
    with pytest.raises(RuntimeError):
        batcher.DeviceBatcher(idx, g['news_combined'], int(g['T']), device='cpu')


@pytest.mark.gpu
def test_device_batcher_matches_reference_golden():
    from newsrecommendation_b200 import batcher
    g, news_index = _gold()
    T, K = int(g['T']), int(g['K'])
    idx = batcher.TrainIndex.from_tsv(os.path.join(G, 'batcher_train.tsv'), news_index, K)
    b = batcher.DeviceBatcher(idx, g['news_combined'], T)
    h, m, c, l = b.batch(np.arange(len(idx)), g['label'])
    assert h.dtype == torch.int32 and m.dtype == torch.float32 and c.dtype == torch.int32 and l.dtype == torch.int64
    assert np.array_equal(h.cpu().numpy(), g['history']) and np.array_equal(m.cpu().numpy(), g['mask'])
    assert np.array_equal(c.cpu().numpy(), g['candidate'])
    # epoch(): file order, ragged last batch, labels = the reference's random.randint stream (main.py:284)
    random.seed(0)
    got = list(b.epoch(20, rng=random))
    assert [x[0].shape[0] for x in got] == [20, 20, 8]
    assert np.array_equal(torch.cat([x[3] for x in got]).cpu().numpy(), g['label'])
    assert np.array_equal(torch.cat([x[2] for x in got]).cpu().numpy(), g['candidate'])
    assert np.array_equal(torch.cat([x[0] for x in got]).cpu().numpy(), g['history'])
    # eval half
    ev = batcher.EvalIndex.from_tsv(os.path.join(G, 'batcher_test.tsv'), news_index).to('cuda')
    hi, hm = batcher.eval_history(ev, int(g['N']) + 1, T)
    assert np.array_equal(hi.cpu().numpy(), g['test_hist']) and np.array_equal(hm.cpu().numpy(), g['test_mask'])
    sel = torch.tensor([5, 0, 47, 5], dtype=torch.int64)
    hi, hm = batcher.eval_history(ev, int(g['N']) + 1, T, imp_ids=sel)
    assert np.array_equal(hi.cpu().numpy(), g['test_hist'][sel.numpy()])
    # empty batch
    h, m, c, l = b.batch(np.zeros(0, np.int64), np.zeros(0, np.int64))
    assert h.shape == (0, T, int(g['F']))


@pytest.mark.gpu
@pytest.mark.parametrize('I,B,T,K,F,N', [(20000, 4096, 50, 4, 30, 65239), (3000, 777, 7, 2, 3, 50), (500, 500, 64, 9, 33, 1000)])
def test_device_batcher_matches_numpy_oracle_at_size(I, B, T, K, F, N):
    from newsrecommendation_b200 import batcher
    rs = np.random.RandomState(I + B)
    lens = np.minimum(rs.lognormal(3.0, 1.0, size=I).astype(np.int64), 3 * T)
    lens[rs.rand(I) < 0.02] = 0
    off = np.zeros(I + 1, np.int64); off[1:] = np.cumsum(lens)
    idx = batcher.TrainIndex(off, rs.randint(0, N + 1, size=int(off[-1])).astype(np.int32),
                             rs.randint(0, N + 1, size=I).astype(np.int32), rs.randint(0, N + 1, size=(I, K)).astype(np.int32))
    news = rs.randint(0, 70000, size=(N + 1, F)).astype(np.int32)
    ids = rs.randint(0, I, size=B)
    lab = rs.randint(0, K + 1, size=B)
    h, m, c, l = batcher.DeviceBatcher(idx, news, T).batch(ids, lab)
    ho, mo, co, _ = O.batch_train_np(off, idx.hist_idx, idx.pos, idx.neg, ids, lab, news, T)
    assert np.array_equal(h.cpu().numpy(), ho) and np.array_equal(m.cpu().numpy(), mo) and np.array_equal(c.cpu().numpy(), co)
