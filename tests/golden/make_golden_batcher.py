"""Golden fixture for the device-side batcher: runs the UNMODIFIED reference DatasetTrain / DatasetTest
line mappers (src/dataset.py) on small synthetic behaviors files.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden_batcher.py
Writes batcher_train.tsv, batcher_test.tsv (the inputs) and batcher.npz (the reference's outputs)."""
import argparse
import os
import random
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = '/root/reference/src'
T, K, F, N = 50, 4, 32, 200


def synth_files():
    rs = np.random.RandomState(11)
    nid = lambda: ('N%d' % rs.randint(1, N + 1)) if rs.rand() > 0.05 else ('X%d' % rs.randint(1, 50))   # 5 % unknown ids
    lens = [0, 1, 49, 50, 51, 70, 120] + list(rs.randint(0, 90, size=41))
    tr, te = [], []
    for i, ln in enumerate(lens):
        hist = ' '.join(nid() for _ in range(ln))
        tr.append('\t'.join([str(i), 'U%d' % i, '11/11/2019 9:05:58 AM', hist, nid(), ' '.join(nid() for _ in range(K))]) + '\n')
        C = int(rs.randint(2, 40))
        labs = rs.randint(0, 2, size=C)
        if i % 9 == 0:
            labs[:] = 0              # all-negative impression (skipped by the metric loop, main.py:250-251)
        te.append('\t'.join([str(i), 'U%d' % i, '11/11/2019 9:05:58 AM', hist,
                             ' '.join('%s-%d' % (nid(), l) for l in labs)]) + '\n')
    return tr, te


def main():
    sys.path.insert(0, REF)
    from dataset import DatasetTrain, DatasetTest
    tr, te = synth_files()
    ptr, pte = os.path.join(HERE, 'batcher_train.tsv'), os.path.join(HERE, 'batcher_test.tsv')
    open(ptr, 'w').writelines(tr)
    open(pte, 'w').writelines(te)
    news_index = {'N%d' % i: i for i in range(1, N + 1)}
    news_combined = np.random.RandomState(12).randint(0, 30000, size=(N + 1, F)).astype(np.int32)
    news_combined[0] = 0
    args = argparse.Namespace(user_log_length=T, npratio=K)
    ds = DatasetTrain(ptr, news_index, news_combined, args)
    random.seed(0)                                   # main.py:284
    H, M, Cn, Lb = [], [], [], []
    for h, m, c, l in ds:                            # the reference's own iterator + line_mapper
        H.append(h); M.append(m); Cn.append(c); Lb.append(l)
    # DatasetTest returns news_scoring[...] rows: a table whose row i is [i] makes the outputs the indices
    scoring = np.arange(N + 1, dtype=np.float32)[:, None]
    dt = DatasetTest(pte, news_index, scoring, args)
    TH, TM, TC, TL, off = [], [], [], [], [0]
    for uf, lm, nf, lab in dt:
        TH.append(uf[:, 0].astype(np.int32)); TM.append(lm)
        TC.extend(nf[:, 0].astype(np.int32)); TL.extend(lab); off.append(len(TC))
    np.savez_compressed(os.path.join(HERE, 'batcher.npz'),
                        T=T, K=K, F=F, N=N, news_combined=news_combined,
                        history=np.stack(H), mask=np.stack(M), candidate=np.stack(Cn), label=np.asarray(Lb, np.int64),
                        test_hist=np.stack(TH), test_mask=np.stack(TM), test_cand_off=np.asarray(off, np.int64),
                        test_cand=np.asarray(TC, np.int32), test_labels=np.asarray(TL, np.int8))
    print('wrote', len(H), 'train lines,', len(TH), 'test lines')


if __name__ == '__main__':
    main()
