"""Host code around the hot path against the reference: `mind_io` vs files and values the
UNMODIFIED reference `prepare_data.py` / `preprocess.py` / `parameters.py` produced (tests/golden/mind_tiny,
driver.json; make_golden_driver.py).  Byte / integer work: everything exact."""
import argparse
import filecmp
import json
import os
import shutil

import numpy as np
import pytest

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
TINY = os.path.join(G, 'mind_tiny')


def _gold():
    return json.load(open(os.path.join(G, 'driver.json')))


def test_prepare_data_byte_identical(tmp_path):
    from newsrecommendation_b200 import mind_io
    g = _gold()
    shutil.copy(os.path.join(TINY, 'behaviors.tsv'), tmp_path / 'behaviors.tsv')
    assert mind_io.prepare_training_data(str(tmp_path), 2, 4, 0) == g['n_train']
    assert mind_io.prepare_testing_data(str(tmp_path), 2) == g['n_test']
    for r in range(2):
        assert filecmp.cmp(tmp_path / f'behaviors_np4_{r}.tsv', os.path.join(TINY, f'ref_behaviors_np4_{r}.tsv'), shallow=False)
        assert filecmp.cmp(tmp_path / f'behaviors_{r}.tsv', os.path.join(TINY, f'ref_behaviors_{r}.tsv'), shallow=False)
    assert mind_io.count_lines(tmp_path / 'behaviors_0.tsv') + mind_io.count_lines(tmp_path / 'behaviors_1.tsv') == g['n_test']
    # another seed shuffles differently (the seed is honoured)
    mind_io.prepare_training_data(str(tmp_path), 2, 4, 1)
    assert not filecmp.cmp(tmp_path / 'behaviors_np4_0.tsv', os.path.join(TINY, 'ref_behaviors_np4_0.tsv'), shallow=False)


def test_read_news_and_doc_input_match_reference():
    from newsrecommendation_b200 import mind_io
    g = _gold()
    a = argparse.Namespace(use_category=True, use_subcategory=True)
    news, news_index, cd, sd = mind_io.read_news(os.path.join(TINY, 'news.tsv'), a, mode='train')
    assert news == g['news'] and news_index == g['news_index'] and cd == g['category_dict'] and sd == g['subcategory_dict']
    assert list(news) == sorted(g['news_index'], key=g['news_index'].get)      # insertion order = row order
    t, c, s = mind_io.get_doc_input(news, news_index, cd, sd, a)
    assert t.dtype == np.int32 and t.shape == (len(news) + 1, 1)
    assert t[:, 0].tolist() == g['title'] and c[:, 0].tolist() == g['category'] and s[:, 0].tolist() == g['subcategory']
    news_t, idx_t = mind_io.read_news(os.path.join(TINY, 'news.tsv'), a, mode='test')
    cd2 = {k: v for k, v in cd.items() if v != 2}
    _, c2, _ = mind_io.get_doc_input(news_t, idx_t, cd2, sd, a)
    assert c2[:, 0].tolist() == g['category_unseen'] and 0 in c2[1:, 0]
    comb = mind_io.combine_news_columns(t, c, s)
    assert comb.shape == (len(news) + 1, 3) and comb.flags['C_CONTIGUOUS'] and comb.dtype == np.int32
    a0 = argparse.Namespace(use_category=False, use_subcategory=False)
    _, _, cd0, sd0 = mind_io.read_news(os.path.join(TINY, 'news.tsv'), a0, mode='train')
    assert cd0 == {} and sd0 == {}
    t0, c0, s0 = mind_io.get_doc_input(news, news_index, cd0, sd0, a0)
    assert c0 is None and s0 is None and mind_io.combine_news_columns(t0, c0, s0).shape[1] == 1
    with pytest.raises(ValueError):
        mind_io.read_news(os.path.join(TINY, 'news.tsv'), a, mode='other')


def test_parameters_match_reference_defaults():
    from newsrecommendation_b200 import parameters
    g = _gold()
    d = vars(parameters.parse_args([]))
    for k, v in g['defaults'].items():
        assert d[k] == v, k
    line = ('--mode test --model NAML --batch_size 128 --user_log_mask True --load_ckpt_name epoch-5.pt '
            '--prepare False --nGPU 1 --enable_gpu True --use_category True --use_subcategory True '
            '--test_data_dir ../data/MINDsmall_dev --freeze_embedding True').split()
    d = vars(parameters.parse_args(line))
    for k, v in g['demo_test'].items():
        assert d[k] == v, k
    with pytest.raises(SystemExit):
        parameters.parse_args(['--use_category', 'maybe'])


def test_title_table_file_roundtrip(tmp_path):
    from newsrecommendation_b200 import mind_io
    full = mind_io.write_synthetic_mind_dir(str(tmp_path), n_news=12, n_impressions=5, L=3, D=8, seed=1)
    t = mind_io.read_news_embeddings(str(tmp_path))                     # gzip path (preprocess.py:227-232)
    assert t.dtype == np.float32 and np.array_equal(t, full) and not t[0].any()
    t = mind_io.read_news_embeddings(str(tmp_path), cache=True)         # writes the raw table
    raw = tmp_path / mind_io.TITLE_TABLE_RAW
    assert raw.exists() and os.path.getsize(raw) == mind_io.TABLE_HEADER_BYTES + full.nbytes
    m = mind_io.read_news_embeddings(str(tmp_path))                     # now memory-mapped
    assert isinstance(m, np.memmap) and np.array_equal(np.asarray(m), full)
    with open(raw, 'r+b') as f:
        f.truncate(os.path.getsize(raw) - 4)
    with pytest.raises(ValueError):
        mind_io.read_title_table(raw)
    (tmp_path / 'junk.nrtab').write_bytes(b'\x00' * 300)
    with pytest.raises(ValueError):
        mind_io.read_title_table(tmp_path / 'junk.nrtab')


def test_title_table_file_bf16_payload(tmp_path):
    """dtype code 1: bf16 payload, round-to-nearest-even on write (checked against torch's conversion, ties and
    signed values included), exact widening on read, half the file."""
    import torch
    from newsrecommendation_b200 import mind_io
    rs = np.random.RandomState(3)
    rows = mind_io._ROW_BLOCK + 37                               # more than one conversion block
    full = (rs.standard_normal((rows, 24)) * 0.4).astype(np.float32)
    full[0] = 0
    # exact ties between two bf16 neighbours (low half = 0x8000) with even and odd upper halves, both signs
    ties = np.array([0x3F808000, 0x3F818000, 0xBF808000, 0xBF818000, 0x3F80FFFF, 0x00000001, 0x7F7F7FFF],
                    dtype=np.uint32).view(np.float32)
    full[1, :ties.size] = ties
    path = tmp_path / 'title_embeddings.bpemb.nrtab'
    mind_io.write_title_table(path, full, dtype='bf16')
    assert os.path.getsize(path) == mind_io.TABLE_HEADER_BYTES + full.size * 2
    back = mind_io.read_title_table(path)
    want = torch.from_numpy(full).to(torch.bfloat16).float().numpy()
    assert back.dtype == np.float32 and back.shape == full.shape
    assert np.array_equal(back.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(mind_io.read_news_embeddings(str(tmp_path)), back)       # picked up like the fp32 file
    with pytest.raises(ValueError):
        mind_io.write_title_table(tmp_path / 'x.nrtab', full, dtype='fp16')
    bad = full.copy(); bad[2, 2] = np.inf
    with pytest.raises(ValueError):
        mind_io.write_title_table(tmp_path / 'y.nrtab', bad, dtype='bf16')
    # the driver's conversion mode writes what --table_dtype asks for
    from newsrecommendation_b200 import parameters
    assert parameters.parse_args([]).table_dtype == 'fp32'
    assert parameters.parse_args(['--table_dtype', 'bf16']).table_dtype == 'bf16'


def test_raw_table_dtype(tmp_path):
    from newsrecommendation_b200 import mind_io
    t = np.ones((3, 8), dtype=np.float32)
    mind_io.write_title_table(tmp_path / 'a.nrtab', t)
    mind_io.write_title_table(tmp_path / 'b.nrtab', t, dtype='bf16')
    (tmp_path / 'c.nrtab').write_bytes(b'not a table')
    assert mind_io.raw_table_dtype(tmp_path / 'a.nrtab') == 'fp32'
    assert mind_io.raw_table_dtype(tmp_path / 'b.nrtab') == 'bf16'
    assert mind_io.raw_table_dtype(tmp_path / 'c.nrtab') is None
    assert mind_io.raw_table_dtype(tmp_path / 'missing.nrtab') is None


def test_driver_rejects_ids_outside_their_table():
    """a title_token_ids / word_embeddings pair that does not match is an IndexError up front, as
    nn.Embedding would raise in the reference (src/model/NRMS.py:28), not an out-of-bounds device read."""
    import argparse
    import types
    import pytest
    import torch
    from newsrecommendation_b200 import main as drv
    enc = types.SimpleNamespace(embedding_matrix=types.SimpleNamespace(weight=torch.zeros(10, 4)))
    model = types.SimpleNamespace(news_encoder=enc)
    args = argparse.Namespace(model='NRMS', num_words_title=3)
    ok = np.array([[0, 1, 9], [2, 3, 4]], np.int32)
    drv._check_ids(model, ok, args)
    with pytest.raises(IndexError, match='title token ids'):
        drv._check_ids(model, np.array([[0, 1, 10]], np.int32), args)
    with pytest.raises(IndexError):
        drv._check_ids(model, np.array([[0, -1, 2]], np.int32), args)
