"""CPU / eager-GPU comparators of bench.py: the REFERENCE's own modules (oracle/_ref, copied unmodified from
/root/reference/src by oracle/make_ref.py) driven exactly as src/main.py drives them.  Falls back to the oracle port
(oracle/nr_oracle.py) only when oracle/_ref is absent.  Nothing here imports the CUDA library: the process of
`bench.py --impl reference` maps no libnr_sm100.so.

  ref_train   src/main.py:105-110 verbatim (forward, zero_grad, backward, optim.Adam.step) on all host threads
  ref_eval    src/main.py:190-263 restated (news-vector cache, user encoder, np.dot, sklearn AUC + metrics.py)
  gpu_eager   the same reference modules .cuda(), fp32, TF32 off: the same-box comparator SURVEY 2.3 names
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
L_TITLE, T_HIST, K_NEG = 30, 50, 4


def _synth():
    from newsrecommendation_b200 import synthetic as S      # numpy only; the package object loads no library
    return S


def _ref():
    from oracle import make_ref
    return make_ref.load()


def build_reference_model(model_name, vocab, news_rows, device='cpu'):
    """Unmodified reference Model on synthetic tables.  Returns (model, news_combined, args)."""
    S = _synth()
    NRMS, NAML, _ = _ref()
    torch.manual_seed(0)
    if model_name == 'NRMS':
        args = S.default_args(num_words_title=L_TITLE, drop_rate=0.2)
        model = NRMS.Model(args, S.word_table(vocab))                       # NRMS.py:67
        news = S.news_titles(news_rows, L_TITLE, vocab)
    else:
        args = S.default_args(num_words_title=L_TITLE, drop_rate=0.2, use_category=True, use_subcategory=True,
                              freeze_embedding=True)
        model = NAML.Model(args, S.title_table(news_rows, L_TITLE), 17, 264)  # NAML.py:101
        news = S.naml_features(news_rows)
    return model.to(device), news, args


def _train_loop(model, batches, steps, lr, device, budget_s, sync=None):
    """src/main.py:76,98-110: optim.Adam over model.parameters(); per batch forward, zero_grad, backward, step."""
    opt = torch.optim.Adam(model.parameters(), lr=lr)                       # main.py:76
    model.train()
    loss_acc, acc_acc = 0.0, 0.0

    def one(i):
        nonlocal loss_acc, acc_acc
        history, history_mask, candidate, label = batches[i % len(batches)]
        bz_loss, y_hat = model(history, history_mask, candidate, label)     # main.py:105
        loss_acc = loss_acc + bz_loss.data.float()                          # main.py:106
        acc_acc = acc_acc + (torch.argmax(y_hat, dim=-1) == label).float().mean()   # utils.acc, main.py:107
        opt.zero_grad()                                                     # main.py:108
        bz_loss.backward()                                                  # main.py:109
        opt.step()                                                          # main.py:110

    one(0)                                                                  # warm-up
    if sync:
        sync()
    t0 = time.perf_counter()
    done = 0
    while done < steps and (time.perf_counter() - t0 < budget_s or done < 2):
        one(done + 1)
        done += 1
    if sync:
        sync()
    return done, time.perf_counter() - t0, float(loss_acc)


def ref_train(model_name, vocab, cpu_batch=32, budget_s=25.0, max_steps=6):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    S = _synth()
    if _ref() is None:
        return port_train(model_name, vocab, cpu_batch, budget_s, max_steps)
    model, news, _ = build_reference_model(model_name, vocab, 2000)
    batches = [[torch.from_numpy(x) for x in b] for b in S.train_batches(news, 2, cpu_batch)]
    done, dt, _ = _train_loop(model, batches, max_steps, 3e-4, 'cpu', budget_s)
    return {'value': done * cpu_batch / dt, 'unit': 'samples/s', 'cores': cores, 'kind': 'reference',
            'sample': f'{done} steps of src/main.py:105-110 (forward, zero_grad, backward, optim.Adam.step) on the '
                      f'unmodified reference {model_name} (oracle/_ref), batch {cpu_batch}, L={L_TITLE}, T={T_HIST}, '
                      f'npratio {K_NEG}, drop 0.2, V={vocab if model_name == "NRMS" else "-"}, 2000-news corpus, '
                      f'torch-CPU fp32, {cores} threads',
            'ms_per_step': dt / done * 1e3, 'batch': cpu_batch}


def ref_eval(model_name, vocab, impressions=1000, cache_news=8191, batch=128):
    """src/main.py:185-263 restated around the unmodified reference encoders and metrics.py: cache over `cache_news`+1
    news rows (the reference encodes every row, main.py:190-196), then `impressions` impressions whose indices fall in
    that table: user encoder in batches of `batch` (demo.sh:27), per-impression np.dot + roc_auc/mrr/ndcg@5/@10."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    S = _synth()
    ref = _ref()
    if ref is None:
        return None
    _, _, M = ref
    model, news, _ = build_reference_model(model_name, vocab, cache_news)
    model.eval()
    torch.set_grad_enabled(False)
    try:
        news_t = torch.from_numpy(news)
        t0 = time.perf_counter()
        news_scoring = []
        for b0 in range(0, news_t.shape[0], batch):                         # main.py:190-196
            news_vec = model.news_encoder(news_t[b0:b0 + batch])
            news_scoring.extend(news_vec.to(torch.device('cpu')).detach().numpy())
        news_scoring = np.array(news_scoring)                               # main.py:198
        t_cache = time.perf_counter() - t0
        hidx, hmask, off, cand, labels = S.eval_impressions(cache_news, impressions, seed=500)
        t0 = time.perf_counter()
        AUC, MRR, nDCG5, nDCG10 = [], [], [], []
        for b0 in range(0, impressions, batch):
            b1 = min(impressions, b0 + batch)
            log_vecs = torch.FloatTensor(news_scoring[hidx[b0:b1]])         # dataset.py:67-68 + collate main.py:212-216
            log_mask = torch.FloatTensor(hmask[b0:b1])
            user_vecs = model.user_encoder(log_vecs, log_mask).to(torch.device('cpu')).detach().numpy()   # main.py:247
            for i, user_vec in zip(range(b0, b1), user_vecs):
                label = labels[off[i]:off[i + 1]].astype(np.int64)
                if label.mean() == 0 or label.mean() == 1:                  # main.py:250-251
                    continue
                news_vec = news_scoring[cand[off[i]:off[i + 1]]]            # dataset.py:72
                score = np.dot(news_vec, user_vec)                          # main.py:253
                AUC.append(M.roc_auc_score(label, score)); MRR.append(M.mrr_score(label, score))
                nDCG5.append(M.ndcg_score(label, score, k=5)); nDCG10.append(M.ndcg_score(label, score, k=10))
        t_score = time.perf_counter() - t0
    finally:
        torch.set_grad_enabled(True)
    return {'value': impressions / t_score, 'unit': 'impressions/s', 'cores': cores, 'kind': 'reference',
            'news_per_s': news_scoring.shape[0] / t_cache,
            'metrics_mean_scored': [float(np.mean(x)) for x in (AUC, MRR, nDCG5, nDCG10)],
            'sample': f'{impressions} impressions through src/main.py:240-263 restated (reference {model_name} '
                      f'user_encoder in batches of {batch}, np.dot, sklearn roc_auc_score + metrics.py) after a '
                      f'{news_scoring.shape[0]}-news cache build (main.py:190-198, batch {batch}); torch-CPU fp32, '
                      f'{cores} threads'}


def gpu_eager(model_name, vocab, news_rows, batch, steps=10):
    """The reference modules run eagerly by torch on this GPU (cuBLAS/cuDNN/ATen kernels), fp32 with TF32 off: the
    same-box comparator of SURVEY 2.3.  Same step as ref_train; timed with CUDA events after 3 warm-up steps."""
    if _ref() is None or not torch.cuda.is_available():
        return None
    S = _synth()
    tf32 = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    try:
        model, news, _ = build_reference_model(model_name, vocab, news_rows, device='cuda')
        batches = [[torch.from_numpy(x).cuda() for x in b] for b in S.train_batches(news, 4, batch, seed=100)]
        opt = torch.optim.Adam(model.parameters(), lr=3e-4)
        model.train()

        def one(i):
            h, m, c, y = batches[i % 4]
            bz_loss, y_hat = model(h, m, c, y)
            opt.zero_grad()
            bz_loss.backward()
            opt.step()
            return bz_loss

        for i in range(3):
            one(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            loss = one(i)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out = {'value': batch / (ms * 1e-3), 'unit': 'samples/s', 'ms_per_step': ms, 'batch': batch,
               'final_loss': float(loss.item()),
               'what': f'unmodified reference {model_name} (oracle/_ref) .cuda(), eager torch {torch.__version__}, '
                       f'fp32, TF32 off, optim.Adam, batch {batch}; CUDA events over {steps} steps'}
        del model, opt, batches
        torch.cuda.empty_cache()
        return out
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = tf32


def port_train(model_name, vocab, cpu_batch, budget_s, max_steps):
    """Fallback when oracle/_ref is absent: the oracle port of the same step (kind 'port')."""
    from oracle import nr_oracle as O
    S = _synth()
    sys.path.insert(0, os.path.join(ROOT, 'tests', 'golden'))
    import synth
    cores = os.cpu_count() or 1
    B, p = cpu_batch, 0.2
    if model_name == 'NRMS':
        P = synth.nrms_weights(0, 8)
        P['news_encoder.embedding_matrix.weight'] = S.word_table(vocab)
        news = S.news_titles(2000, L_TITLE, vocab)
    else:
        P = synth.naml_weights(0, 2000, L_TITLE, True, True)
        news = S.naml_features(2000)
    P = {k: torch.from_numpy(v).clone() for k, v in P.items()}
    train = [k for k in P if not (model_name == 'NAML' and k.endswith('title_embeddings.weight'))]
    for k in train:
        P[k].requires_grad_(True)
    st = {k: (torch.zeros_like(P[k]), torch.zeros_like(P[k])) for k in train}
    batches = S.train_batches(news, 2, B)

    def step(i, s):
        h, m, c, y = [torch.from_numpy(x) for x in batches[i % 2]]
        keep = {}
        for tag, cnt in (('cand', B * (1 + K_NEG)), ('hist', B * T_HIST)):
            keep[tag + '1'] = torch.empty(cnt, L_TITLE, 300).bernoulli_(1 - p)
            if model_name == 'NRMS':
                keep[tag + '2'] = torch.empty(cnt, L_TITLE, 400).bernoulli_(1 - p)
        if model_name == 'NRMS':
            loss, _ = O.nrms_forward(h, m, c, y, P, 20, False, p, keep)
        else:
            loss, _ = O.naml_forward(h, m, c, y, P, L_TITLE, 300, True, True, False, p, keep)
        for k in train:
            P[k].grad = None
        loss.backward()
        with torch.no_grad():
            for k in train:
                g = P[k].grad if P[k].grad is not None else torch.zeros_like(P[k])
                O.adam_step(P[k], g, st[k][0], st[k][1], s, 3e-4)

    step(0, 1)
    t0 = time.perf_counter()
    done = 0
    while done < max_steps and (time.perf_counter() - t0 < budget_s or done < 2):
        step(done + 1, done + 2)
        done += 1
    dt = time.perf_counter() - t0
    return {'value': done * B / dt, 'unit': 'samples/s', 'cores': cores, 'kind': 'port',
            'sample': f'{done} train steps of batch {B} ({model_name}) through oracle/nr_oracle.py (oracle/_ref absent), '
                      f'torch-CPU, {cores} threads', 'ms_per_step': dt / done * 1e3, 'batch': B}
