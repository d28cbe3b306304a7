"""Host-side MIND file handling on either side of the hot path: what the driver
(`main.py` here) needs from the reference's `preprocess.py` / `prepare_data.py`, restated so the same data
directories and the same intermediate files work:

* `read_news`, `get_doc_input`      — `src/preprocess.py:16-74` (news.tsv -> indices, category maps, `news_combined` columns)
* `read_news_embeddings`            — `src/preprocess.py:227-232` (`title_embeddings.bpemb.npy.gz`), plus a raw
                                      table file (`*.nrtab`) that is memory-mapped instead of gunzipped: the
                                      2.3-5.8 GB gzip read is the reference's start-up cost
* `prepare_training_data`, `prepare_testing_data` — `src/prepare_data.py:14-66`; byte-identical output files for
                                      the same seed (same sequence of `random` draws), checked against files
                                      the reference wrote (tests/golden/prepare_*.tsv)

Pure host code: no device work happens here."""
import gzip
import os
import random
import struct

import numpy as np

TABLE_MAGIC = b'NRTAB1\x00\x00'
TABLE_HEADER_BYTES = 256            # payload starts 256-byte aligned (rows are copied to HBM with 128-bit loads)
TITLE_TABLE_GZ = 'title_embeddings.bpemb.npy.gz'
TITLE_TABLE_RAW = 'title_embeddings.bpemb.nrtab'


# ------------------------------------------------------------------------------------------------
# news.tsv
# ------------------------------------------------------------------------------------------------
def _next_index(d, key):
    """1-based index in order of first appearance (preprocess.py:8-13)."""
    if key not in d:
        d[key] = len(d) + 1


def read_news(news_path, args, mode='train'):
    """preprocess.py:16-45.  news: doc_id -> [category, subcategory]; news_index: doc_id -> 1-based row;
    the category maps are only built in train mode and only for enabled flags."""
    if mode not in ('train', 'test'):
        raise ValueError('Wrong mode!')
    news, news_index, category_dict, subcategory_dict = {}, {}, {}, {}
    with open(news_path, 'r', encoding='utf-8') as f:
        for ln, line in enumerate(f):
            cols = line.strip('\n').split('\t')
            if len(cols) != 8:
                raise ValueError(f'{news_path}:{ln + 1}: expected 8 tab-separated columns, got {len(cols)}')
            doc_id, category, subcategory = cols[0], cols[1], cols[2]
            _next_index(news_index, doc_id)
            if doc_id not in news:
                news[doc_id] = [category, subcategory]
            if mode == 'train':
                if args.use_category:
                    _next_index(category_dict, category)
                if args.use_subcategory:
                    _next_index(subcategory_dict, subcategory)
    if mode == 'train':
        return news, news_index, category_dict, subcategory_dict
    return news, news_index


def get_doc_input(news, news_index, category_dict, subcategory_dict, args):
    """preprocess.py:49-74: (N+1, 1) int32 columns; the 'title' column of the fork is the row index itself
    (F1), unseen categories map to 0, row 0 is the unknown-news row."""
    n = len(news) + 1
    rows = np.fromiter((news_index[d] for d in news), dtype=np.int64, count=len(news))
    title = np.zeros((n, 1), dtype='int32')
    title[rows, 0] = rows
    cat = sub = None
    if args.use_category:
        cat = np.zeros((n, 1), dtype='int32')
        cat[rows, 0] = np.fromiter((category_dict.get(news[d][0], 0) for d in news), dtype=np.int64, count=len(news))
    if args.use_subcategory:
        sub = np.zeros((n, 1), dtype='int32')
        sub[rows, 0] = np.fromiter((subcategory_dict.get(news[d][1], 0) for d in news), dtype=np.int64, count=len(news))
    return title, cat, sub


def combine_news_columns(*cols):
    """main.py:48,183: concatenate the present columns."""
    return np.ascontiguousarray(np.concatenate([c for c in cols if c is not None], axis=-1))


# ------------------------------------------------------------------------------------------------
# title-embedding table
# ------------------------------------------------------------------------------------------------
TABLE_DTYPES = {'fp32': 0, 'bf16': 1}     # header dtype codes
_ROW_BLOCK = 4096                         # rows converted per step (keeps the temporaries of a 5.8 GB table small)


def _bf16_bits(a):
    """fp32 -> bf16 bit patterns, round to nearest even (what the staging kernel's conversion does)."""
    u = np.ascontiguousarray(a, dtype='<f4').view(np.uint32)
    return ((u + (((u >> 16) & 1) + np.uint32(0x7FFF))) >> 16).astype('<u2')


def write_title_table(path, table, dtype='fp32'):
    """Raw table file: 256-byte header (magic, rows, cols, dtype code) + row-major little-endian payload.
    `cols` = L*D of the fork's per-news layout (NAML.py:104-107).  dtype 'fp32' is lossless; 'bf16' halves the
    file (1.2 GB instead of 2.35 GB for MIND-small) for the bf16 mode, whose staging kernel rounds every table
    row to bf16 anyway — it is NOT for the fp32 parity mode."""
    if dtype not in TABLE_DTYPES:
        raise ValueError(f'title table dtype must be one of {sorted(TABLE_DTYPES)}')
    if getattr(table, 'ndim', 0) != 2:
        raise ValueError('title table must be 2-D (N+1, L*D)')
    head = TABLE_MAGIC + struct.pack('<qqq', table.shape[0], table.shape[1], TABLE_DTYPES[dtype])
    tmp = str(path) + '.tmp'
    with open(tmp, 'wb') as f:
        f.write(head.ljust(TABLE_HEADER_BYTES, b'\x00'))
        for r0 in range(0, table.shape[0], _ROW_BLOCK):
            blk = np.ascontiguousarray(table[r0:r0 + _ROW_BLOCK], dtype='<f4')
            if dtype == 'bf16':
                if not np.isfinite(blk).all():
                    raise ValueError('title table holds non-finite values')
                blk = _bf16_bits(blk)
            blk.tofile(f)
    os.replace(tmp, path)


def read_title_table(path):
    """Reads a file written by `write_title_table`.  fp32 payloads are memory-mapped (read-only;
    `torch.from_numpy(np.asarray(...))` + `.cuda()` streams them to HBM without a host copy of a decompressed
    array); bf16 payloads are widened to an fp32 array block by block (exact: bf16 is the top half of fp32)."""
    with open(path, 'rb') as f:
        head = f.read(TABLE_HEADER_BYTES)
    if len(head) != TABLE_HEADER_BYTES or head[:8] != TABLE_MAGIC:
        raise ValueError(f'{path}: not a title table file')
    rows, cols, code = struct.unpack('<qqq', head[8:32])
    if code not in TABLE_DTYPES.values() or rows <= 0 or cols <= 0:
        raise ValueError(f'{path}: unsupported header (rows={rows}, cols={cols}, dtype code={code})')
    width = 4 if code == 0 else 2
    if os.path.getsize(path) != TABLE_HEADER_BYTES + rows * cols * width:
        raise ValueError(f'{path}: size does not match its header (truncated?)')
    if code == 0:
        return np.memmap(path, dtype='<f4', mode='r', offset=TABLE_HEADER_BYTES, shape=(rows, cols))
    bits = np.memmap(path, dtype='<u2', mode='r', offset=TABLE_HEADER_BYTES, shape=(rows, cols))
    out = np.empty((rows, cols), dtype=np.float32)
    for r0 in range(0, rows, _ROW_BLOCK):
        out[r0:r0 + _ROW_BLOCK] = (bits[r0:r0 + _ROW_BLOCK].astype(np.uint32) << 16).view(np.float32)
    return out


def raw_table_dtype(path):
    """'fp32' / 'bf16' payload of a raw table file, None when `path` is not one."""
    try:
        with open(path, 'rb') as f:
            head = f.read(32)
    except OSError:
        return None
    if len(head) < 32 or head[:8] != TABLE_MAGIC:
        return None
    code = struct.unpack('<q', head[24:32])[0]
    return {v: k for k, v in TABLE_DTYPES.items()}.get(code)


def news_embeddings_source(data_dir, field='title'):
    """Which file read_news_embeddings takes: 'raw' (the memory-mappable table, when at least as new as the .npy.gz)
    or 'gz'."""
    gz = os.path.join(data_dir, TITLE_TABLE_GZ.replace('title_', field + '_'))
    raw = os.path.join(data_dir, TITLE_TABLE_RAW.replace('title_', field + '_'))
    if os.path.exists(raw) and (not os.path.exists(gz) or os.path.getmtime(raw) >= os.path.getmtime(gz)):
        return 'raw'
    return 'gz'


def read_news_embeddings(data_dir, cache=False, field='title', cache_dtype='fp32'):
    """preprocess.py:227-232.  Prefers the raw table next to the .npy.gz when it is at least as new;
    `cache=True` writes it (as `cache_dtype`) after a gzip read so the next start-up skips the gunzip.  `field='abstract'` reads the
    second table of the abstract view (same format, `abstract_embeddings.bpemb.*`)."""
    gz = os.path.join(data_dir, TITLE_TABLE_GZ.replace('title_', field + '_'))
    raw = os.path.join(data_dir, TITLE_TABLE_RAW.replace('title_', field + '_'))
    if news_embeddings_source(data_dir, field) == 'raw':
        return read_title_table(raw)
    with gzip.GzipFile(gz, 'r') as f:
        table = np.load(f)
    if cache:
        write_title_table(raw, table, dtype=cache_dtype)
    return table


# ------------------------------------------------------------------------------------------------
# behaviors.tsv -> per-rank files
# ------------------------------------------------------------------------------------------------
def _draw_negatives(neg, k):
    """prepare_data.py:7-11: k negatives without replacement, the list repeated when it is too short."""
    pool = neg if k <= len(neg) else neg * (k // len(neg) + 1)
    return random.sample(pool, k)


def prepare_training_data(train_data_dir, nGPU, npratio, seed):
    """prepare_data.py:14-50: one line per positive of every impression that has both labels
    (iid, uid, time, history, pos, 'neg neg ...'), shuffled, dealt round-robin to `behaviors_np{K}_{rank}.tsv`.
    Uses the global `random` module in the reference's order (seed, one sample per positive, one shuffle)."""
    random.seed(seed)
    out = []
    with open(os.path.join(train_data_dir, 'behaviors.tsv'), 'r', encoding='utf-8') as f:
        for line in f:
            iid, uid, time, history, imp = line.strip().split('\t')
            pos, neg = [], []
            for item in imp.split(' '):
                nid, lab = item.split('-')
                if lab == '1':
                    pos.append(nid)
                elif lab == '0':
                    neg.append(nid)
            if not pos or not neg:
                continue
            for p in pos:
                out.append('\t'.join((iid, uid, time, history, p, ' '.join(_draw_negatives(neg, npratio)))) + '\n')
    random.shuffle(out)
    for r in range(nGPU):
        with open(os.path.join(train_data_dir, f'behaviors_np{npratio}_{r}.tsv'), 'w') as f:
            f.writelines(out[r::nGPU])
    return len(out)


def prepare_testing_data(test_data_dir, nGPU):
    """prepare_data.py:53-66: line i -> `behaviors_{i % nGPU}.tsv`, unchanged."""
    with open(os.path.join(test_data_dir, 'behaviors.tsv'), 'r', encoding='utf-8') as f:
        lines = f.readlines()
    for r in range(nGPU):
        with open(os.path.join(test_data_dir, f'behaviors_{r}.tsv'), 'w') as f:
            f.writelines(lines[r::nGPU])
    return len(lines)


def count_lines(path):
    n = 0
    with open(path, 'rb') as f:
        for chunk in iter(lambda: f.read(1 << 20), b''):
            n += chunk.count(b'\n')
    return n


# ------------------------------------------------------------------------------------------------
# synthetic MIND-format directory (datasets are unavailable offline; used by tests and the demo)
# ------------------------------------------------------------------------------------------------
def write_synthetic_mind_dir(data_dir, n_news=300, n_impressions=400, L=20, D=300, seed=0, n_cat=6, n_sub=15,
                             unknown_frac=0.03, table_format='gz'):
    """news.tsv (8 columns), behaviors.tsv (5 columns, `id-label` impressions) and the gzipped
    `(N+1, L*D)` title table in the layout `create_news_embeddings` writes (preprocess.py:112-168: row 0
    zeros, titles zero-padded to L tokens)."""
    os.makedirs(data_dir, exist_ok=True)
    rs = np.random.RandomState(seed)
    with open(os.path.join(data_dir, 'news.tsv'), 'w', encoding='utf-8') as f:
        for i in range(1, n_news + 1):
            f.write('\t'.join(['N%d' % i, 'cat%d' % rs.randint(n_cat), 'sub%d' % rs.randint(n_sub), 'title %d' % i,
                               'abstract %d' % i, 'https://x/%d' % i, '[]', '[]']) + '\n')
    table = np.zeros((n_news + 1, L, D), dtype=np.float32)
    n_tok = np.clip(np.rint(rs.normal(11.5, 4, size=n_news + 1)), 3, L).astype(int)
    for i in range(1, n_news + 1):
        table[i, :n_tok[i]] = rs.normal(0, 0.4, size=(n_tok[i], D))
    if table_format == 'gz':
        with gzip.GzipFile(os.path.join(data_dir, TITLE_TABLE_GZ), 'w', compresslevel=1) as f:
            np.save(f, table.reshape(n_news + 1, L * D))
    else:                                        # large synthetic sets: skip the (slow) gzip, write the raw table
        write_title_table(os.path.join(data_dir, TITLE_TABLE_RAW), table.reshape(n_news + 1, L * D))

    def nid():
        return 'X%d' % rs.randint(1, 50) if rs.rand() < unknown_frac else 'N%d' % rs.randint(1, n_news + 1)

    with open(os.path.join(data_dir, 'behaviors.tsv'), 'w', encoding='utf-8') as f:
        for i in range(n_impressions):
            hl = 0 if rs.rand() < 0.02 else int(min(80, rs.lognormal(3.0, 1.0)))
            C = int(np.clip(rs.lognormal(3.2, 0.8), 2, 300))
            labs = np.zeros(C, dtype=int)
            labs[rs.choice(C, size=min(C - 1, rs.randint(1, 4)), replace=False)] = 1
            if i % 57 == 56:
                labs[:] = 0                      # all-negative impression: dropped by prepare, skipped by the metrics
            f.write('\t'.join([str(i + 1), 'U%d' % rs.randint(1, 1000), '11/11/2019 9:05:58 AM',
                               ' '.join(nid() for _ in range(hl)),
                               ' '.join('%s-%d' % (nid(), l) for l in labs)]) + '\n')
    return table.reshape(n_news + 1, L * D)
