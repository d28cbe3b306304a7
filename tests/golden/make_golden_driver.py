"""Golden fixtures for the host code around the hot path: runs the UNMODIFIED reference
`prepare_data.py`, `preprocess.read_news/get_doc_input` and `parameters.parse_args` on a tiny synthetic
MIND-format directory.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden_driver.py
Writes tests/golden/mind_tiny/{news.tsv,behaviors.tsv} (inputs), ref_*.tsv (the split files the reference wrote)
and driver.json (read_news / get_doc_input outputs, parse_args defaults)."""
import argparse
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = '/root/reference/src'
OUT = os.path.join(HERE, 'mind_tiny')


def main():
    sys.path.insert(0, ROOT)
    from newsrecommendation_b200 import mind_io
    shutil.rmtree(OUT, ignore_errors=True)
    mind_io.write_synthetic_mind_dir(OUT, n_news=60, n_impressions=90, L=2, D=4, seed=5)
    os.remove(os.path.join(OUT, mind_io.TITLE_TABLE_GZ))           # not needed by the host-code fixtures
    # duplicate a news row and add an unseen-category row order quirk: first appearance wins (preprocess.py:8-13)
    with open(os.path.join(OUT, 'news.tsv'), 'a', encoding='utf-8') as f:
        f.write('\t'.join(['N7', 'catX', 'subX', 'dup title', 'a', 'u', '[]', '[]']) + '\n')
    sys.path.insert(0, REF)
    import prepare_data
    import preprocess
    n_tr = prepare_data.prepare_training_data(OUT, 2, 4, 0)
    n_te = prepare_data.prepare_testing_data(OUT, 2)
    for r in range(2):
        os.replace(os.path.join(OUT, f'behaviors_np4_{r}.tsv'), os.path.join(OUT, f'ref_behaviors_np4_{r}.tsv'))
        os.replace(os.path.join(OUT, f'behaviors_{r}.tsv'), os.path.join(OUT, f'ref_behaviors_{r}.tsv'))
    a = argparse.Namespace(use_category=True, use_subcategory=True)
    news, news_index, cd, sd = preprocess.read_news(os.path.join(OUT, 'news.tsv'), a, mode='train')
    t, c, s = preprocess.get_doc_input(news, news_index, cd, sd, a)
    # test mode with the TRAIN category maps minus one entry (unseen category -> 0, preprocess.py:68-71)
    cd2 = {k: v for k, v in cd.items() if v != 2}
    news_t, news_index_t = preprocess.read_news(os.path.join(OUT, 'news.tsv'), a, mode='test')
    t2, c2, s2 = preprocess.get_doc_input(news_t, news_index_t, cd2, sd, a)
    old = sys.argv
    sys.argv = ['main.py']
    import parameters
    defaults = vars(parameters.parse_args())
    sys.argv = ['main.py'] + ('--mode test --model NAML --batch_size 128 --user_log_mask True --load_ckpt_name epoch-5.pt '
                              '--prepare False --nGPU 1 --enable_gpu True --use_category True --use_subcategory True '
                              '--test_data_dir ../data/MINDsmall_dev --freeze_embedding True').split()
    demo_test = vars(parameters.parse_args())
    sys.argv = old
    json.dump({'n_train': n_tr, 'n_test': n_te, 'news': news, 'news_index': news_index, 'category_dict': cd,
               'subcategory_dict': sd, 'title': t[:, 0].tolist(), 'category': c[:, 0].tolist(),
               'subcategory': s[:, 0].tolist(), 'category_unseen': c2[:, 0].tolist(),
               'defaults': defaults, 'demo_test': demo_test},
              open(os.path.join(HERE, 'driver.json'), 'w'), indent=0, sort_keys=True)
    print('train lines', n_tr, 'test lines', n_te, 'news', len(news))


if __name__ == '__main__':
    main()
