"""The `demo.sh train` -> `demo.sh test` flow through `newsrecommendation_b200.main` on a
synthetic MIND-format directory, checked against the CPU oracle:

* train: `--drop_rate 0` in fp32, the driver's weights after one epoch (device batcher + CUDA-graph step + flat
  Adam, ragged last batch eager) must equal an oracle loop (oracle forward, autograd, torch Adam as main.py:76)
  over the same lines and the same `random.randint` label draws (dataset.py:45);
* test: metrics printed by the driver vs the oracle restatement of main.py:185-277 on the saved checkpoint,
  with a dev directory whose news count differs from the train directory (F5)."""
import argparse
import os
import random

import numpy as np
import pytest
import torch

from oracle import nr_oracle as O

pytestmark = pytest.mark.gpu
L_TITLE, D = 20, 300


def _dirs(tmp_path):
    from newsrecommendation_b200 import mind_io
    tr, te = str(tmp_path / 'train'), str(tmp_path / 'dev')
    mind_io.write_synthetic_mind_dir(tr, n_news=220, n_impressions=60, L=L_TITLE, D=D, seed=2)
    mind_io.write_synthetic_mind_dir(te, n_news=180, n_impressions=120, L=L_TITLE, D=D, seed=3)
    return tr, te


def _common(tr, te, model_dir, precision):
    return ['--model', 'NAML', '--model_dir', model_dir, '--train_data_dir', tr, '--test_data_dir', te,
            '--use_category', 'True', '--use_subcategory', 'True', '--freeze_embedding', 'True',
            '--num_words_title', str(L_TITLE), '--precision', precision, '--nGPU', '1', '--enable_gpu', 'True',
            '--doc_sim_pairs', '20000', '--log_steps', '2']


def _oracle_train(tr, ckpt, lr, batch_size):
    """main.py:98-110 on the CPU: same files, same label draws, oracle forward + autograd + torch Adam."""
    from newsrecommendation_b200 import mind_io, batcher
    a = argparse.Namespace(use_category=True, use_subcategory=True)
    news, news_index, cd, sd = mind_io.read_news(os.path.join(tr, 'news.tsv'), a, 'train')
    comb = mind_io.combine_news_columns(*mind_io.get_doc_input(news, news_index, cd, sd, a))
    table = torch.from_numpy(np.asarray(mind_io.read_news_embeddings(tr)))
    P = {k: v.clone().requires_grad_(True) for k, v in ckpt.items()}
    P['news_encoder.title_embeddings.weight'] = table                    # frozen (demo.sh:12)
    opt = torch.optim.Adam([v for k, v in P.items() if v.requires_grad], lr=lr)
    idx = batcher.TrainIndex.from_tsv(os.path.join(tr, 'behaviors_np4_0.tsv'), news_index, 4)
    losses = []       # `random` continues from the caller's state (label draws, dataset.py:45)
    for s in range(0, len(idx), batch_size):
        e = min(len(idx), s + batch_size)
        label = np.array([random.randint(0, 4) for _ in range(e - s)], dtype=np.int64)
        h, m, c, l = O.batch_train_np(idx.hist_off, idx.hist_idx, idx.pos, idx.neg, np.arange(s, e), label, comb, 50)
        loss, _ = O.naml_forward(torch.from_numpy(h), torch.from_numpy(m), torch.from_numpy(c), torch.from_numpy(l),
                                 P, L_TITLE, D, True, True, False)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(loss.item())
    return {k: v.detach() for k, v in P.items()}, losses


def test_demo_train_then_test_fp32_vs_oracle(tmp_path):
    from newsrecommendation_b200 import main as drv, mind_io
    tr, te = _dirs(tmp_path)
    mdir = str(tmp_path / 'model')
    base = _common(tr, te, mdir, 'fp32')
    # initial weights: a 0-epoch run saves the freshly initialised model (epochs == start_epoch -> no steps)
    torch.manual_seed(0)
    from newsrecommendation_b200.model import NAML
    from newsrecommendation_b200 import parameters
    a0 = parameters.parse_args(base + ['--mode', 'train'])
    news, news_index, cd, sd = mind_io.read_news(os.path.join(tr, 'news.tsv'), a0, 'train')
    init = NAML.Model(a0, mind_io.read_news_embeddings(tr), len(cd), len(sd))
    os.makedirs(mdir, exist_ok=True)
    init_sd = {k: v.clone() for k, v in init.state_dict().items() if k != drv.TITLE_KEY}
    torch.save({'model_state_dict': init_sd, 'category_dict': cd, 'subcategory_dict': sd}, os.path.join(mdir, 'init.pt'))

    # ---- demo.sh train (drop 0 so that the oracle can follow) ----
    drv.main(base + ['--mode', 'train', '--epochs', '1', '--batch_size', '32', '--lr', '0.0003', '--user_log_mask', 'False',
                     '--prepare', 'True', '--drop_rate', '0', '--load_ckpt_name', 'init.pt', '--seed', '0'])
    torch.set_grad_enabled(True)
    ck = torch.load(os.path.join(mdir, 'epoch-1.pt'), map_location='cpu', weights_only=False)
    assert set(ck) == {'model_state_dict', 'category_dict', 'subcategory_dict'}             # main.py:120-127
    assert drv.TITLE_KEY not in ck['model_state_dict'] and ck['category_dict'] == cd
    assert list(ck['model_state_dict']) == list(init_sd)
    # the driver seeds `random` (main.py:284), prepare consumes draws, the batcher continues the stream: replay it
    random.seed(0)
    mind_io.prepare_training_data(tr, 1, 4, 0)
    st = random.getstate()
    n_lines = mind_io.count_lines(os.path.join(tr, 'behaviors_np4_0.tsv'))
    assert n_lines % 32 != 0                                                                # ragged last batch exercised

    random.setstate(st)
    Pw, losses = _oracle_train(tr, init_sd, 0.0003, 32)
    steps = len(losses)
    assert steps == (n_lines + 31) // 32
    # Adam moves every weight by at most lr per step; fp32 parity of the gradients (rel 1e-4) must keep the two
    # trajectories within a small fraction of that.  att_fc2.bias has an analytically ~0 gradient (the pooling
    # weights are normalised), so its Adam direction is rounding noise in both implementations: only bounded.
    for k, v in ck['model_state_dict'].items():
        d = np.abs(v.numpy() - Pw[k].numpy())
        moved = np.abs(Pw[k].numpy() - init_sd[k].numpy()).max()
        if k.endswith('att_fc2.bias'):
            assert d.max() <= 0.0003 * steps * 2, k
            continue
        assert d.max() <= 0.1 * 0.0003 * steps, (k, d.max(), moved)
        assert d.mean() <= 0.01 * 0.0003 * steps, (k, d.mean(), moved)
    assert max(np.abs(Pw[k].numpy() - init_sd[k].numpy()).max() for k in init_sd) > 0.5 * 0.0003 * steps   # it trained

    # ---- demo.sh test epoch-1.pt ----
    red, scores, metrics, valid = drv.main(base + ['--mode', 'test', '--batch_size', '128', '--user_log_mask', 'True',
                                                   '--load_ckpt_name', 'epoch-1.pt', '--prepare', 'True'])
    torch.set_grad_enabled(True)
    a = argparse.Namespace(use_category=True, use_subcategory=True)
    news_t, idx_t = mind_io.read_news(os.path.join(te, 'news.tsv'), a, 'test')
    comb = mind_io.combine_news_columns(*mind_io.get_doc_input(news_t, idx_t, ck['category_dict'], ck['subcategory_dict'], a))
    P = dict(ck['model_state_dict'])
    P[drv.TITLE_KEY] = torch.from_numpy(np.asarray(mind_io.read_news_embeddings(te)))
    table_o = O.build_news_cache(lambda x: O.naml_news_encoder(x, P, L_TITLE, D, True, True), comb)
    H, M, off, cand, lab = [], [], [0], [], []
    for line in open(os.path.join(te, 'behaviors_0.tsv')):
        h, m, c, l = O.test_line_mapper(line, idx_t, 50)
        H.append(h); M.append(m); cand.extend(c); lab.extend(l); off.append(len(cand))
    mets_o, scores_o = O.eval_impressions(lambda v, m: O.naml_user_encoder(v, m, P, True), table_o, np.stack(H), np.stack(M),
                                          np.asarray(off), np.asarray(cand, np.int32), np.asarray(lab, np.int8))
    summ = O.summarize_metrics(mets_o)
    assert red['total'] == len(H) == 120 and red['scored'] == summ['scored'] < red['total']
    np.testing.assert_allclose(red['mean_scored'], summ['mean_scored'], atol=1e-3)          # north_star tolerance
    np.testing.assert_allclose(scores.cpu().numpy(), scores_o, rtol=1e-4, atol=1e-5)


def test_demo_flow_bf16_graph_and_raw_table(tmp_path):
    """bench precision, dropout on, CUDA-graph step, the raw title-table file, unfrozen table key handling."""
    from newsrecommendation_b200 import main as drv, mind_io
    tr, te = _dirs(tmp_path)
    mdir = str(tmp_path / 'model')
    base = _common(tr, te, mdir, 'bf16')
    drv.main(base + ['--mode', 'read_embeddings'])
    assert os.path.exists(os.path.join(tr, mind_io.TITLE_TABLE_RAW)) and os.path.exists(os.path.join(te, mind_io.TITLE_TABLE_RAW))
    os.remove(os.path.join(tr, mind_io.TITLE_TABLE_GZ))                                     # only the raw table is left
    red, scores, metrics, valid = drv.main(base + ['--mode', 'train_test', '--epochs', '2', '--batch_size', '32', '--lr', '0.0003',
                                                   '--prepare', 'True', '--user_log_mask', 'False'])
    torch.set_grad_enabled(True)
    assert os.path.exists(os.path.join(mdir, 'epoch-2.pt'))
    assert red['total'] == 120 and 0 < red['scored'] < 120
    assert np.isfinite(red['mean_scored']).all() and (red['mean_scored'] > 0.05).all() and (red['mean_scored'] <= 1).all()
    assert torch.isfinite(scores).all()
    # a checkpoint from another news set that carries its table still loads (F5)
    ck = torch.load(os.path.join(mdir, 'epoch-2.pt'), map_location='cpu', weights_only=False)
    ck['model_state_dict'][drv.TITLE_KEY] = torch.zeros(5, L_TITLE * D)
    torch.save(ck, os.path.join(mdir, 'foreign.pt'))
    red2, *_ = drv.main(base + ['--mode', 'test', '--load_ckpt_name', 'foreign.pt', '--prepare', 'False', '--user_log_mask', 'False'])
    torch.set_grad_enabled(True)
    np.testing.assert_allclose(red2['mean_scored'], red['mean_scored'], atol=1e-12)


@pytest.mark.parametrize('model', ['NRMS', 'NAML'])
def test_driver_word_id_layouts(tmp_path, model):
    """`--model NRMS` behind the driver (F2/F3: token ids + word table read from title_token_ids.npy /
    word_embeddings.npy) and NAML over word ids (`--naml_word_ids True`, the layout commented out at NAML.py:40-45):
    train one epoch, test, and check the printed metrics against the oracle on the saved checkpoint."""
    from newsrecommendation_b200 import main as drv, mind_io, synthetic as S
    tr, te = _dirs(tmp_path)
    V = 3000
    for d, seed in ((tr, 11), (te, 12)):
        n_news = mind_io.count_lines(os.path.join(d, 'news.tsv'))
        np.save(os.path.join(d, 'title_token_ids.npy'), S.news_titles(n_news, L_TITLE, V, seed=seed))
        np.save(os.path.join(d, 'word_embeddings.npy'), S.word_table(V, D, seed=7))
    mdir = str(tmp_path / 'model')
    base = ['--model', model, '--model_dir', mdir, '--train_data_dir', tr, '--test_data_dir', te, '--num_words_title', str(L_TITLE),
            '--precision', 'fp32', '--nGPU', '1', '--doc_sim_pairs', '1000', '--log_steps', '100', '--prepare', 'True',
            '--use_category', 'True', '--use_subcategory', 'True', '--naml_word_ids', 'True', '--drop_rate', '0.2']
    red, scores, metrics, valid = drv.main(base + ['--mode', 'train_test', '--epochs', '1', '--batch_size', '32', '--lr', '0.0003',
                                                   '--user_log_mask', 'True'])
    torch.set_grad_enabled(True)
    ck = torch.load(os.path.join(mdir, 'epoch-1.pt'), map_location='cpu', weights_only=False)
    P = dict(ck['model_state_dict'])
    a = argparse.Namespace(use_category=model == 'NAML', use_subcategory=model == 'NAML')
    news_t, idx_t = mind_io.read_news(os.path.join(te, 'news.tsv'), a, 'test')
    _, cat, sub = mind_io.get_doc_input(news_t, idx_t, ck['category_dict'], ck['subcategory_dict'], a)
    tok = np.load(os.path.join(te, 'title_token_ids.npy'))
    if model == 'NRMS':
        assert 'news_encoder.embedding_matrix.weight' in P and P['news_encoder.embedding_matrix.weight'].shape == (V, D)
        comb = tok
        enc = lambda x: O.nrms_news_encoder(x, P, 20)
        uenc = lambda v, m: O.nrms_user_encoder(v, m, P, 20, True)
    else:
        comb = mind_io.combine_news_columns(tok, cat, sub)
        enc = lambda x: O.naml_news_encoder(x, P, L_TITLE, D, True, True, word_ids=True)
        uenc = lambda v, m: O.naml_user_encoder(v, m, P, True)
    table_o = O.build_news_cache(enc, comb)
    H, M, off, cand, lab = [], [], [0], [], []
    for line in open(os.path.join(te, 'behaviors_0.tsv')):
        h, m, c, l = O.test_line_mapper(line, idx_t, 50)
        H.append(h); M.append(m); cand.extend(c); lab.extend(l); off.append(len(cand))
    mets_o, scores_o = O.eval_impressions(uenc, table_o, np.stack(H), np.stack(M), np.asarray(off), np.asarray(cand, np.int32),
                                          np.asarray(lab, np.int8))
    summ = O.summarize_metrics(mets_o)
    assert red['scored'] == summ['scored']
    np.testing.assert_allclose(scores.cpu().numpy(), scores_o, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(red['mean_scored'], summ['mean_scored'], atol=1e-3)


def test_demo_flow_with_abstract_view(tmp_path):
    """`--use_abstract True` through the driver: second per-news table read from the raw table file, extra feature
    column, checkpoint without either frozen table, test-time metrics vs the oracle composition."""
    from newsrecommendation_b200 import main as drv, mind_io
    tr, te = _dirs(tmp_path)
    La = 50
    for d, seed in ((tr, 21), (te, 22)):
        n_rows = mind_io.count_lines(os.path.join(d, 'news.tsv')) + 1
        t = (np.random.RandomState(seed).standard_normal((n_rows, La * D)) * 0.4).astype(np.float32)
        t[0] = 0
        mind_io.write_title_table(os.path.join(d, mind_io.TITLE_TABLE_RAW.replace('title_', 'abstract_')), t)
    mdir = str(tmp_path / 'model')
    base = _common(tr, te, mdir, 'fp32') + ['--use_abstract', 'True', '--num_words_abstract', str(La), '--prepare', 'True']
    red, scores, metrics, valid = drv.main(base + ['--mode', 'train_test', '--epochs', '1', '--batch_size', '32', '--lr', '0.0003',
                                                   '--user_log_mask', 'True', '--drop_rate', '0.2'])
    torch.set_grad_enabled(True)
    ck = torch.load(os.path.join(mdir, 'epoch-1.pt'), map_location='cpu', weights_only=False)
    sd = ck['model_state_dict']
    assert drv.TITLE_KEY not in sd and drv.ABSTRACT_KEY not in sd and 'news_encoder.abstract_cnn.weight' in sd
    a = argparse.Namespace(use_category=True, use_subcategory=True)
    news_t, idx_t = mind_io.read_news(os.path.join(te, 'news.tsv'), a, 'test')
    title, cat, sub = mind_io.get_doc_input(news_t, idx_t, ck['category_dict'], ck['subcategory_dict'], a)
    comb = mind_io.combine_news_columns(title, title, cat, sub)
    P = dict(sd)
    P[drv.TITLE_KEY] = torch.from_numpy(np.asarray(mind_io.read_news_embeddings(te)))
    P[drv.ABSTRACT_KEY] = torch.from_numpy(np.asarray(mind_io.read_news_embeddings(te, field='abstract')))
    table_o = O.build_news_cache(lambda x: O.naml_news_encoder(x, P, L_TITLE, D, True, True, La=La), comb)
    H, M, off, cand, lab = [], [], [0], [], []
    for line in open(os.path.join(te, 'behaviors_0.tsv')):
        h, m, c, l = O.test_line_mapper(line, idx_t, 50)
        H.append(h); M.append(m); cand.extend(c); lab.extend(l); off.append(len(cand))
    mets_o, scores_o = O.eval_impressions(lambda v, m: O.naml_user_encoder(v, m, P, True), table_o, np.stack(H), np.stack(M),
                                          np.asarray(off), np.asarray(cand, np.int32), np.asarray(lab, np.int8))
    np.testing.assert_allclose(scores.cpu().numpy(), scores_o, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(red['mean_scored'], O.summarize_metrics(mets_o)['mean_scored'], atol=1e-3)
