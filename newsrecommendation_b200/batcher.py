"""Device-side batcher: replaces the per-line Python of `DatasetTrain.line_mapper` /
`DatasetTest.line_mapper` (src/dataset.py:26-49, 64-74) and the `news_combined[...]` fancy indexing.

The behaviors files `prepare_data.py` writes (`behaviors_np{K}_{rank}.tsv`, `behaviors_{rank}.tsv`) are
tokenised ONCE into flat integer arrays (`TrainIndex`, `EvalIndex`; `.save()`/`.load()` keep them as one
uncompressed .npz next to the TSV), moved to the GPU, and every batch is assembled there by
`nr_batch_train` / `nr_batch_eval_history` (csrc/nr_batcher.cu): pad/truncate/mask, candidate ordering and
the feature gathers.  Parsing follows the reference line format exactly (tab-separated, field 3 history,
4 positives / `id-label` candidates, 5 negatives; unknown news ids -> index 0, dataset.py:14-15).
"""
import random as _random

import numpy as np
import torch

from . import _lib as L


def _nindex(news_index, nids):
    """DatasetTrain.trans_to_nindex (src/dataset.py:14-15)."""
    return [news_index[i] if i in news_index else 0 for i in nids]


class TrainIndex:
    """hist_off int64 (I+1), hist_idx int32, pos int32 (I), neg int32 (I,K) — one row per TSV line."""

    def __init__(self, hist_off, hist_idx, pos, neg):
        self.hist_off, self.hist_idx, self.pos, self.neg = hist_off, hist_idx, pos, neg

    def __len__(self):
        return int(self.pos.shape[0])

    @property
    def npratio(self):
        return int(self.neg.shape[1])

    @classmethod
    def from_tsv(cls, filename, news_index, npratio):
        off, hist, pos, neg = [0], [], [], []
        with open(filename) as f:
            for ln, line in enumerate(f):
                line = line.strip().split('\t')                      # dataset.py:34
                h = _nindex(news_index, line[3].split())             # :35, :39 (padding happens on the device)
                p = _nindex(news_index, line[4].split())             # :36, :42
                n = _nindex(news_index, line[5].split())             # :37, :43
                if len(p) != 1 or len(n) != npratio:
                    raise ValueError(f'{filename}:{ln + 1}: expected 1 positive and {npratio} negatives '
                                     f'(prepare_data.py:30-34), got {len(p)} and {len(n)}')
                hist.extend(h)
                off.append(len(hist))
                pos.append(p[0])
                neg.append(n)
        return cls(np.asarray(off, np.int64), np.asarray(hist, np.int32), np.asarray(pos, np.int32),
                   np.asarray(neg, np.int32).reshape(len(pos), npratio))

    def save(self, path):
        np.savez(path, kind='train', hist_off=self.hist_off, hist_idx=self.hist_idx, pos=self.pos, neg=self.neg)

    @classmethod
    def load(cls, path):
        d = np.load(path)
        return cls(d['hist_off'], d['hist_idx'], d['pos'], d['neg'])

    def to(self, device):
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
        return TrainIndex(t(self.hist_off), t(self.hist_idx), t(self.pos), t(self.neg))


class EvalIndex:
    """Histories as CSR + candidates as CSR (cand_off int64 (I+1), cand_idx int32, labels int8): the layout
    `engine.score_impressions` consumes."""

    def __init__(self, hist_off, hist_idx, cand_off, cand_idx, labels):
        self.hist_off, self.hist_idx = hist_off, hist_idx
        self.cand_off, self.cand_idx, self.labels = cand_off, cand_idx, labels

    def __len__(self):
        return int(self.cand_off.shape[0]) - 1

    @classmethod
    def from_tsv(cls, filename, news_index):
        hoff, hist, coff, cand, lab = [0], [], [0], [], []
        with open(filename) as f:
            for line in f:
                line = line.strip().split('\t')                                        # dataset.py:65
                hist.extend(_nindex(news_index, line[3].split()))                      # :66-67
                hoff.append(len(hist))
                imp = line[4].split()
                cand.extend(_nindex(news_index, [i.split('-')[0] for i in imp]))       # :70
                lab.extend(int(i.split('-')[1]) for i in imp)                          # :71
                coff.append(len(cand))
        return cls(np.asarray(hoff, np.int64), np.asarray(hist, np.int32), np.asarray(coff, np.int64),
                   np.asarray(cand, np.int32), np.asarray(lab, np.int8))

    def save(self, path):
        np.savez(path, kind='eval', hist_off=self.hist_off, hist_idx=self.hist_idx, cand_off=self.cand_off,
                 cand_idx=self.cand_idx, labels=self.labels)

    @classmethod
    def load(cls, path):
        d = np.load(path)
        return cls(d['hist_off'], d['hist_idx'], d['cand_off'], d['cand_idx'], d['labels'])

    def to(self, device):
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
        return EvalIndex(t(self.hist_off), t(self.hist_idx), t(self.cand_off), t(self.cand_idx), t(self.labels))


class DeviceBatcher:
    """Assembles `(history, history_mask, candidate, label)` — the four tensors `model.forward` takes
    (src/main.py:99-105) — on the GPU from a device-resident TrainIndex and news_combined table."""

    def __init__(self, index, news_combined, user_log_length, device='cuda'):
        self.device = torch.device(device)
        if self.device.type != 'cuda':
            raise RuntimeError('DeviceBatcher needs a CUDA device (there is no CPU path)')
        self.n = len(index)
        self.K = index.npratio
        self.T = int(user_log_length)
        self.idx = index.to(self.device)
        nc = torch.as_tensor(np.ascontiguousarray(news_combined)) if not torch.is_tensor(news_combined) else news_combined
        if nc.dtype != torch.int32 or nc.dim() != 2:
            raise TypeError('news_combined must be a 2-D int32 array (preprocess.py:52-54)')
        self.news = nc.to(self.device).contiguous()
        self.F = int(self.news.shape[1])

    def batch(self, sample_ids, label):
        """sample_ids int64 (B) rows of the index, label int64 (B) positive slots in [0, K]."""
        sample_ids = torch.as_tensor(sample_ids, dtype=torch.int64).to(self.device)
        label = torch.as_tensor(label, dtype=torch.int64).to(self.device)
        B = int(sample_ids.shape[0])
        if label.shape[0] != B:
            raise ValueError('sample_ids and label differ in length')
        hist = torch.empty(B, self.T, self.F, dtype=torch.int32, device=self.device)
        mask = torch.empty(B, self.T, dtype=torch.float32, device=self.device)
        cand = torch.empty(B, 1 + self.K, self.F, dtype=torch.int32, device=self.device)
        i = self.idx
        L.call('nr_batch_train', L.ptr(i.hist_off), L.ptr(i.hist_idx), L.ptr(i.pos), L.ptr(i.neg), L.ptr(sample_ids),
               L.ptr(label), L.ptr(self.news), self.news.shape[0], B, self.T, self.K, self.F, L.ptr(hist), L.ptr(mask),
               L.ptr(cand), L.stream())
        return hist, mask, cand, label

    def epoch(self, batch_size, rng=_random):
        """Batches in file order like the reference DataLoader over the IterableDataset (main.py:92, no
        drop_last: the last batch may be short); one `rng.randint(0, npratio)` per line, in line order
        (dataset.py:45), so a seeded `random` reproduces the reference's labels."""
        for s in range(0, self.n, batch_size):
            e = min(self.n, s + batch_size)
            label = torch.tensor([rng.randint(0, self.K) for _ in range(e - s)], dtype=torch.int64)
            yield self.batch(torch.arange(s, e, dtype=torch.int64), label)


def eval_history(index, news_rows, user_log_length, imp_ids=None):
    """Padded click indices + mask for impressions of a device-resident EvalIndex (dataset.py:64-68)."""
    dev = index.hist_off.device
    if dev.type != 'cuda':
        raise RuntimeError('eval_history needs a device-resident EvalIndex')
    B = len(index) if imp_ids is None else int(imp_ids.shape[0])
    T = int(user_log_length)
    out = torch.empty(B, T, dtype=torch.int32, device=dev)
    mask = torch.empty(B, T, dtype=torch.float32, device=dev)
    ids = None if imp_ids is None else torch.as_tensor(imp_ids, dtype=torch.int64).to(dev)
    L.call('nr_batch_eval_history', L.ptr(index.hist_off), L.ptr(index.hist_idx), L.ptr(ids) if ids is not None else None,
           int(news_rows), B, T, L.ptr(out), L.ptr(mask), L.stream())
    return out, mask
