"""Driver for the `demo.sh train` / `demo.sh test` flow on the sm_100a path.

Same command line (`parameters.py`), same data directories and intermediate files, same log lines and the
same checkpoint dictionary as the reference's `src/main.py`, with its defects repaired:

* F3  `--model NRMS` constructs (the plugin accepts the extra positionals); NRMS needs token ids, which the
      fork's `get_doc_input` no longer produces (F2), so the driver reads `title_token_ids.npy` (N+1, L) int32
      and `word_embeddings.npy` (V, D) from the data directory for it.
* F4  `test()` reads the dev title table with the right arity and never touches `.cuda` on a CPU box
      (this path needs a CUDA device and says so).
* F5  checkpoints really drop the title table (it is an input file, not a weight, when frozen), and `test()`
      takes the table of the directory it scores, so train/dev news counts may differ.
* per-rank step counts are equalised (the reference can hang DDP at epoch end).

What runs where: per-line Python (`dataset.py`) -> `batcher.DeviceBatcher`; `optim.Adam` + DDP ->
`engine.FlatAdam` (one all-reduce + one fused Adam) with forward+backward replayed from a CUDA graph;
news cache (`main.py:185-198`) -> `engine.build_news_cache` (row shards + all-gather, table stays in HBM);
doc-sim loop (`main.py:201-208`) -> one batched device expression; impression loop + sklearn metrics
(`main.py:240-263`) -> `engine.score_impressions`; reductions (`main.py:268-277`) -> `engine.reduce_eval_sums`.

Launch: `python -m newsrecommendation_b200.main --mode train ...` spawns `--nGPU` processes itself
(`main.py:306-309`), or runs as one rank of a `torchrun` job when RANK/WORLD_SIZE are set.
"""
import logging
import os
import random
import sys
import time
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

from . import batcher, engine, mind_io
from .parameters import parse_args

TITLE_KEY = 'news_encoder.title_embeddings.weight'
ABSTRACT_KEY = 'news_encoder.abstract_embeddings.weight'


def setuplogger():
    """utils.py:20-27 (same format string)."""
    root = logging.getLogger()
    root.setLevel(logging.INFO)
    if not any(getattr(h, '_nr', False) for h in root.handlers):
        h = logging.StreamHandler(sys.stdout)
        h.setLevel(logging.INFO)
        h.setFormatter(logging.Formatter('[%(levelname)s %(asctime)s] %(message)s'))
        h._nr = True
        root.addHandler(h)


def _init_rank(rank, args):
    """main.py:22-34: (rank, world, distributed?)."""
    if not torch.cuda.is_available() or not args.enable_gpu:
        raise RuntimeError('the sm_100a path needs a CUDA device and --enable_gpu True (there is no CPU path)')
    distributed = rank is not None
    rank = rank or 0
    world = args.nGPU if distributed else 1
    torch.cuda.set_device(rank % torch.cuda.device_count())
    if distributed and not dist.is_initialized():
        setuplogger()
        dist.init_process_group('nccl', world_size=world, rank=rank, init_method='env://')
    if distributed:
        # spawned ranks do not inherit the parent's seeding (main.py:284-285 seeds only the launcher): without this
        # every rank would draw the same dropout masks and label positions and --seed would have no effect
        random.seed(args.seed + rank)
        torch.manual_seed(args.seed + rank)
    return rank, world, distributed


def _finish(distributed):
    if distributed:
        torch.cuda.synchronize()
        dist.barrier(device_ids=[torch.cuda.current_device()])
        dist.destroy_process_group()


def _build_model(data_dir, args, n_cat, n_sub):
    """main.py:63-64 / :169-170 (+ the abstract table when --use_abstract)."""
    kw = {}
    if args.model == 'NAML' and args.use_abstract:
        kw['abstract_embeddings_weight'] = _embedding_table(data_dir, args, 'abstract')
    return _plugin(args).Model(args, _embedding_table(data_dir, args), n_cat, n_sub, **kw)


def _plugin(args):
    """main.py:63 `importlib.import_module(f'model.{args.model}')` resolved to this package's plugin."""
    if args.model is None:
        raise SystemExit('--model {NAML,NRMS} is required')
    import importlib
    return importlib.import_module(f'newsrecommendation_b200.model.{args.model}')


def _news_inputs(data_dir, args, mode, category_dict=None, subcategory_dict=None):
    """main.py:38-48 / :180-183 -> (news_index, category_dict, subcategory_dict, news_combined int32 (N+1,F))."""
    news_path = os.path.join(data_dir, 'news.tsv')
    if mode == 'train':
        news, news_index, category_dict, subcategory_dict = mind_io.read_news(news_path, args, mode='train')
    else:
        news, news_index = mind_io.read_news(news_path, args, mode='test')
    title, cat, sub = mind_io.get_doc_input(news, news_index, category_dict, subcategory_dict, args)
    if args.model == 'NRMS' or args.naml_word_ids:
        tok = np.load(os.path.join(data_dir, 'title_token_ids.npy')).astype(np.int32)
        if tok.shape != (title.shape[0], args.num_words_title):
            raise ValueError(f'title_token_ids.npy has shape {tok.shape}, expected {(title.shape[0], args.num_words_title)}')
        title = tok
    if args.model == 'NRMS':
        cat = sub = None                      # NRMS is title-only (NRMS.py:79-95)
    abstract = None
    if args.model == 'NAML' and args.use_abstract:
        if args.naml_word_ids:
            abstract = np.load(os.path.join(data_dir, 'abstract_token_ids.npy')).astype(np.int32)
        else:
            abstract = title[:, :1].copy()    # fork layout: the row index again, into the abstract table
    return news_index, category_dict, subcategory_dict, mind_io.combine_news_columns(title, abstract, cat, sub)


def _embedding_table(data_dir, args, field='title'):
    if args.model == 'NRMS' or args.naml_word_ids:
        return np.load(os.path.join(data_dir, 'word_embeddings.npy'))
    table = mind_io.read_news_embeddings(data_dir, field=field)
    raw = os.path.join(data_dir, mind_io.TITLE_TABLE_RAW.replace('title_', field + '_'))
    if (args.precision == 'fp32' and mind_io.news_embeddings_source(data_dir, field) == 'raw'
            and mind_io.raw_table_dtype(raw) == 'bf16'):      # only when the bf16 file is the one actually read
        logging.warning(f'{raw} holds a bf16 payload: --precision fp32 will not match a run on the fp32 table')
    want = (args.num_words_title if field == 'title' else args.num_words_abstract) * args.word_embedding_dim
    if table.shape[1] != want:
        raise ValueError(f'{field} table has {table.shape[1]} columns, --num_words_{field} x --word_embedding_dim = {want}')
    if isinstance(table, np.memmap):          # read-only mapping: it is only ever copied (to HBM), never written
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            table = torch.from_numpy(table)
    return table


def _checkpoint_dict(model, args, category_dict, subcategory_dict):
    """main.py:120-127 with the filter that was meant (F5): the frozen title table stays out."""
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()
          if not (k in (TITLE_KEY, ABSTRACT_KEY) and args.freeze_embedding)}
    return {'model_state_dict': sd, 'category_dict': category_dict, 'subcategory_dict': subcategory_dict}


def _load_weights(model, state):
    for key in (TITLE_KEY, ABSTRACT_KEY):
        state = _keep_own_table(model, state, key)
    model.load_state_dict(state, strict=True)


def _keep_own_table(model, state, table_key):
    """Strict load of everything except a title table that is absent from the checkpoint or belongs to
    another news set (F5): the model keeps the table of the directory it was built from."""
    own = model.state_dict()
    state = dict(state)
    if table_key in own and (table_key not in state or state[table_key].shape != own[table_key].shape):
        state[table_key] = own[table_key]
    return state


def _check_ids(model, news_combined, args):
    """nn.Embedding raises IndexError for an id outside its table (what the reference does with a mismatched
    news.tsv / table pair).  The kernels read the padding row instead (include/nr_sm100.h: nr_rows_t.table_rows), so
    the driver rejects such inputs here, once per table, before anything runs."""
    enc = model.news_encoder
    nc = np.asarray(news_combined)

    def chk(cols, rows, what):
        if cols.size and (int(cols.min()) < 0 or int(cols.max()) >= rows):
            raise IndexError(f'{what}: ids span [{int(cols.min())}, {int(cols.max())}] but the table has {rows} rows')

    if args.model == 'NRMS':
        chk(nc[:, :args.num_words_title], enc.embedding_matrix.weight.shape[0], 'title token ids vs word_embeddings.npy')
        return
    col = args.num_words_title if enc.word_ids else 1
    chk(nc[:, :col], enc.title_embeddings.weight.shape[0], 'title ids vs the title table')
    if enc.use_abstract:
        w = enc.num_words_abstract if enc.word_ids else 1
        chk(nc[:, col:col + w], enc.abstract_embeddings.weight.shape[0], 'abstract ids vs the abstract table')
        col += w
    if enc.use_category:
        chk(nc[:, col], enc.category_emb.weight.shape[0], 'category ids vs category_emb')
        col += 1
    if enc.use_subcategory:
        chk(nc[:, col], enc.subcategory_emb.weight.shape[0], 'subcategory ids vs subcategory_emb')


def train(rank, args):
    rank, world, distributed = _init_rank(rank, args)
    dev = torch.device('cuda', torch.cuda.current_device())
    news_index, category_dict, subcategory_dict, news_combined = _news_inputs(args.train_data_dir, args, 'train')
    if rank == 0:
        logging.info('Initializing word embedding matrix...')
    model = _build_model(args.train_data_dir, args, len(category_dict), len(subcategory_dict))
    _check_ids(model, news_combined, args)
    if rank == 0:
        print("Model's state_dict:")
        for k, v in model.state_dict().items():
            print(k, '\t', v.size())
    if args.load_ckpt_name is not None:
        ckpt_path = os.path.join(args.model_dir, args.load_ckpt_name)
        if not os.path.exists(ckpt_path):
            raise FileNotFoundError(f'checkpoint {ckpt_path} does not exist')
        _load_weights(model, torch.load(ckpt_path, map_location='cpu', weights_only=False)['model_state_dict'])
        logging.info(f'Model loaded from {ckpt_path}.')
    model.cuda(dev).train()
    opt = engine.FlatAdam(model, lr=args.lr)
    opt.broadcast_params()                                         # DDP constructor (main.py:82)

    data_file_path = os.path.join(args.train_data_dir, f'behaviors_np{args.npratio}_{rank}.tsv')
    index = batcher.TrainIndex.from_tsv(data_file_path, news_index, args.npratio)
    loader = batcher.DeviceBatcher(index, news_combined, args.user_log_length, device=dev)
    n_batches = (len(index) + args.batch_size - 1) // args.batch_size
    if distributed:                                                # equal step counts on every rank
        t = torch.tensor([n_batches], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        n_batches = int(t.item())
    logging.info('Training...')
    try:
        _train_epochs(args, rank, dev, model, opt, loader, n_batches, category_dict, subcategory_dict)
    finally:
        if engine.GraphedTrainStep._live is not None:              # the library holds a raw pointer to its salt
            engine.GraphedTrainStep._live.close()
    _finish(distributed)
    return model


def _train_epochs(args, rank, dev, model, opt, loader, n_batches, category_dict, subcategory_dict):
    graphed = None
    for ep in range(args.start_epoch, args.epochs):
        loss_sum = torch.zeros((), device=dev)
        hit_sum = torch.zeros((), device=dev)
        seen = steps = 0
        t0 = time.time()
        for cnt, (history, history_mask, candidate, label) in enumerate(loader.epoch(args.batch_size)):
            if cnt >= n_batches:
                break
            B = int(label.shape[0])
            if args.graph and B == args.batch_size:
                if graphed is None:
                    graphed = engine.GraphedTrainStep(model, opt, (history, history_mask, candidate, label))
                bz_loss, hits = graphed(history, history_mask, candidate, label)
            else:                                                  # ragged last batch (no drop_last, main.py:92)
                bz_loss, hits = engine.train_step(model, opt, history, history_mask, candidate, label)
            loss_sum += bz_loss                                    # main.py:106-107, kept on the device
            hit_sum += hits / B
            seen += B
            steps += 1
            if cnt % args.log_steps == 0:
                logging.info('[{}][{}] Ed: {}, train_loss: {:.5f}, acc: {:.5f}'.format(
                    ep, rank, cnt * args.batch_size, loss_sum.item() / (cnt + 1), hit_sum.item() / (cnt + 1)))
            if rank == 0 and cnt != 0 and cnt % args.save_steps == 0:
                ckpt_path = os.path.join(args.model_dir, f'epoch-{ep + 1}-{cnt}.pt')
                torch.save(_checkpoint_dict(model, args, category_dict, subcategory_dict), ckpt_path)
                logging.info(f'Model saved to {ckpt_path}.')
        torch.cuda.synchronize()
        dt = time.time() - t0
        logging.info('Training finish.')
        logging.info('[{}][{}] epoch: {} samples in {:.2f} s ({:.0f} samples/s), train_loss: {:.5f}, acc: {:.5f}'.format(
            ep, rank, seen, dt, seen / max(dt, 1e-9), loss_sum.item() / max(1, steps), hit_sum.item() / max(1, steps)))
        if rank == 0:
            ckpt_path = os.path.join(args.model_dir, f'epoch-{ep + 1}.pt')
            torch.save(_checkpoint_dict(model, args, category_dict, subcategory_dict), ckpt_path)
            logging.info(f'Model saved to {ckpt_path}.')


@torch.no_grad()
def doc_similarity(table, pairs, seed):
    """main.py:199-208: mean cosine of `pairs` random (i, j), i != j, rows >= 1 — one device expression
    instead of a 10^6-iteration Python loop.  (The reference divides by the number of DRAWS, i == j draws
    contributing 0; kept.)"""
    n = table.shape[0]
    if n < 3 or pairs <= 0:
        return float('nan')
    g = torch.Generator(device=table.device)
    g.manual_seed(seed)
    acc = torch.zeros((), dtype=torch.float64, device=table.device)
    unit = table / table.norm(dim=1, keepdim=True).clamp_min(1e-30)
    for s in range(0, pairs, 1 << 18):
        m = min(1 << 18, pairs - s)
        i = torch.randint(1, n, (m,), generator=g, device=table.device)
        j = torch.randint(1, n, (m,), generator=g, device=table.device)
        cos = (unit[i] * unit[j]).sum(dim=1)
        acc += torch.where(i != j, cos, torch.zeros_like(cos)).double().sum()
    return float(acc.item() / pairs)


def test(rank, args):
    rank, world, distributed = _init_rank(rank, args)
    dev = torch.device('cuda', torch.cuda.current_device())
    ckpt_path = os.path.join(args.model_dir, args.load_ckpt_name) if args.load_ckpt_name is not None else None
    assert ckpt_path is not None and os.path.exists(ckpt_path), 'No checkpoint found.'          # main.py:162
    checkpoint = torch.load(ckpt_path, map_location='cpu', weights_only=False)
    category_dict, subcategory_dict = checkpoint['category_dict'], checkpoint['subcategory_dict']

    model = _build_model(args.test_data_dir, args, len(category_dict), len(subcategory_dict))
    _load_weights(model, checkpoint['model_state_dict'])
    logging.info(f'Model loaded from {ckpt_path}')
    model.cuda(dev).eval()
    torch.set_grad_enabled(False)

    news_index, _, _, news_combined = _news_inputs(args.test_data_dir, args, 'test', category_dict, subcategory_dict)
    _check_ids(model, news_combined, args)
    t0 = time.time()
    table = engine.build_news_cache(model, torch.from_numpy(news_combined).to(dev), batch=max(args.batch_size, 1024))
    torch.cuda.synchronize()
    logging.info('news scoring num: {} ({:.2f} s)'.format(table.shape[0], time.time() - t0))
    if rank == 0:
        logging.info(f'News doc-sim: {doc_similarity(table, args.doc_sim_pairs, args.seed)}')

    data_file_path = os.path.join(args.test_data_dir, f'behaviors_{rank}.tsv')
    index = batcher.EvalIndex.from_tsv(data_file_path, news_index).to(dev)
    n_local = len(index)
    hist, mask = batcher.eval_history(index, table.shape[0], args.user_log_length)
    t0 = time.time()
    scores, metrics, valid, sums = engine.score_impressions(model, table, hist, mask, index.cand_off, index.cand_idx,
                                                            index.labels, batch=max(args.batch_size, 1024))
    torch.cuda.synchronize()
    dt = time.time() - t0

    def print_metrics(who, cnt, x):                                 # main.py:229-230
        logging.info('[{}] {} samples: {}'.format(who, cnt, '\t'.join(['{:0.2f}'.format(i * 100) for i in x])))

    s = sums.cpu().numpy()
    print_metrics(rank, n_local, s[:4] / max(1.0, s[4]))            # running mean over scored impressions (main.py:265-266)
    logging.info('[{}] local_sample_num: {} ({:.0f} impressions/s)'.format(rank, n_local, n_local / max(dt, 1e-9)))
    red = engine.reduce_eval_sums(sums, n_local)
    if rank == 0:
        # main.py:274-277: the distributed branch divides by ALL impressions, the single-process branch by the
        # scored ones; both are reported by reduce_eval_sums, the line printed is the reference's.
        print_metrics('*', red['total'], red['mean_total'] if distributed else red['mean_scored'])
    _finish(distributed)
    return red, scores, metrics, valid


def _spawn_entry(rank, fn_name, args):
    globals()[fn_name](rank, args)


def _run(fn, args):
    if 'RANK' in os.environ and int(os.environ.get('WORLD_SIZE', '1')) > 1:       # one rank of a torchrun job
        args.nGPU = int(os.environ['WORLD_SIZE'])
        return fn(int(os.environ['RANK']), args)
    if args.nGPU == 1:
        return fn(None, args)
    torch.multiprocessing.spawn(_spawn_entry, nprocs=args.nGPU, args=(fn.__name__, args))    # main.py:309,331


def _check_split_files(data_dir, pattern, n):
    total = 0
    for i in range(n):
        p = os.path.join(data_dir, pattern.format(i))
        if not os.path.exists(p):
            logging.error(f'Splited data {p} for GPU {i} does not exist. Please set the parameter --prepare as True and rerun the code.')
            raise SystemExit(1)
        total += mind_io.count_lines(p)
    return total


def main(argv=None):
    setuplogger()
    args = parse_args(argv)
    for k in sorted(vars(args)):                                    # utils.dump_args
        logging.info(f'args[{k}]={getattr(args, k)}')
    random.seed(args.seed)                                          # main.py:284-285
    torch.manual_seed(args.seed)
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    os.environ.setdefault('MASTER_PORT', str(args.master_port))
    Path(args.model_dir).mkdir(parents=True, exist_ok=True)
    lead = int(os.environ.get('RANK', '0')) == 0
    n_ranks = int(os.environ['WORLD_SIZE']) if 'RANK' in os.environ and 'WORLD_SIZE' in os.environ else args.nGPU
    out = None

    if 'train' in args.mode:
        # under torchrun only rank 0 touches the split files; the others block in init_process_group until it joins
        if lead:
            if args.prepare:
                logging.info('Preparing training data...')
                total = mind_io.prepare_training_data(args.train_data_dir, n_ranks, args.npratio, args.seed)
            else:
                total = _check_split_files(args.train_data_dir, 'behaviors_np%d_{}.tsv' % args.npratio, n_ranks)
                logging.info('Skip training data preparation.')
            logging.info(f'{total} training samples, {total // args.batch_size // n_ranks} batches in total.')
        out = _run(train, args)

    if 'test' in args.mode:
        if args.mode == 'train_test' and args.load_ckpt_name is None:
            args.load_ckpt_name = f'epoch-{args.epochs}.pt'
        if lead:
            if args.prepare:
                logging.info('Preparing testing data...')
                total = mind_io.prepare_testing_data(args.test_data_dir, n_ranks)
            else:
                total = _check_split_files(args.test_data_dir, 'behaviors_{}.tsv', n_ranks)
                logging.info('Skip testing data preparation.')
            logging.info(f'{total} testing samples in total.')
        out = _run(test, args)

    if args.mode == 'create_embeddings':
        raise SystemExit('create_embeddings needs the BPEmb model (preprocess.py:112-168), which is an offline '
                         'preprocessing step outside this path; `read_embeddings` converts an existing '
                         'title_embeddings.bpemb.npy.gz into the raw table format')
    if args.mode == 'read_embeddings':
        for d in (args.train_data_dir, args.test_data_dir):
            if os.path.exists(os.path.join(d, mind_io.TITLE_TABLE_GZ)):
                t = mind_io.read_news_embeddings(d, cache=True, cache_dtype=args.table_dtype)
                logging.info(f'{d}: title table {t.shape}')
    return out


if __name__ == '__main__':
    main()
