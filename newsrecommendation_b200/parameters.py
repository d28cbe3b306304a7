"""Command-line flags of the driver: the reference's `src/parameters.py:5-62` flag set (same names,
types and defaults, booleans through `str2bool` like `src/utils.py:9-17`), so `demo.sh` command lines
work unchanged, plus the few switches the sm_100a path adds.

Unused-by-the-reference flags (`--num_words_abstract`, `--filter_num`, `--glove_embedding_path`) are
accepted and ignored, as there."""
import argparse
import os

MODES = ('train', 'test', 'train_test', 'create_embeddings', 'read_embeddings')


def str2bool(v):
    if isinstance(v, bool):
        return v
    s = str(v).lower()
    if s in ('yes', 'true', 't', 'y', '1'):
        return True
    if s in ('no', 'false', 'f', 'n', '0'):
        return False
    raise argparse.ArgumentTypeError('Boolean value expected.')


# (flag, type, default) in the order of src/parameters.py
_REFERENCE_FLAGS = (
    ('nGPU', int, 1), ('seed', int, 0), ('prepare', str2bool, True),
    ('train_data_dir', str, '../data/MINDsmall_train'), ('test_data_dir', str, '../data/MINDsmall_dev'),
    ('model_dir', str, '../model'), ('batch_size', int, 32), ('npratio', int, 4),
    ('enable_gpu', str2bool, True), ('filter_num', int, 3), ('log_steps', int, 100),
    ('epochs', int, 1), ('lr', float, 0.0001),
    ('num_words_title', int, 20), ('num_words_abstract', int, 50), ('user_log_length', int, 50),
    ('word_embedding_dim', int, 300), ('glove_embedding_path', str, '../data/glove.840B.300d.txt'),
    ('freeze_embedding', str2bool, False), ('news_dim', int, 400), ('news_query_vector_dim', int, 200),
    ('user_query_vector_dim', int, 200), ('num_attention_heads', int, 20), ('user_log_mask', str2bool, False),
    ('drop_rate', float, 0.2), ('save_steps', int, 10000), ('start_epoch', int, 0),
    ('load_ckpt_name', str, None), ('use_category', str2bool, False), ('use_subcategory', str2bool, False),
    ('category_emb_dim', int, 100),
)


def build_parser():
    ap = argparse.ArgumentParser(description='NRMS/NAML train/test on the sm_100a path (flags of src/parameters.py)')
    ap.add_argument('--mode', type=str, default='train', choices=list(MODES))
    ap.add_argument('--model', type=str, default=None, choices=['NAML', 'NRMS'])
    for name, typ, default in _REFERENCE_FLAGS:
        ap.add_argument('--' + name, type=typ, default=default)
    # additions of this implementation (none changes the meaning of a reference flag)
    ap.add_argument('--precision', type=str, default=os.environ.get('NR_PRECISION', 'bf16'), choices=['bf16', 'fp32'],
                    help='bf16 = tensor-core kernels (rel 1e-2), fp32 = parity kernels (rel 1e-5)')
    ap.add_argument('--naml_word_ids', type=str2bool, default=False,
                    help='NAML over token ids + word table (the layout commented out at NAML.py:40-45)')
    ap.add_argument('--use_abstract', type=str2bool, default=False,
                    help='NAML abstract view (BASELINE config 3 "+abstract"; uses --num_words_abstract and a second '
                         'per-news table abstract_embeddings.bpemb.npy.gz / .nrtab in the data directory)')
    ap.add_argument('--table_dtype', type=str, default='fp32', choices=['fp32', 'bf16'],
                    help='payload of the raw title table that --mode read_embeddings writes next to the .npy.gz '
                         '(bf16: half the file, for --precision bf16 only)')
    ap.add_argument('--graph', type=str2bool, default=True,
                    help='replay forward+backward of full batches from one CUDA graph')
    ap.add_argument('--doc_sim_pairs', type=int, default=1000000, help='random pairs of the doc-sim diagnostic (main.py:201-208)')
    ap.add_argument('--master_port', type=int, default=8888)
    return ap


def parse_args(argv=None):
    return build_parser().parse_args(argv)
