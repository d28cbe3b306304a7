"""`demo.sh train` + `demo.sh test` on a synthetic MIND-format directory pair (datasets are unavailable offline):
writes train/dev directories, then runs the driver exactly as demo.sh does (NAML, category + subcategory views,
frozen title table, lr 3e-4, user_log_mask False for train / True for test).

    python -m newsrecommendation_b200.demo_synthetic --work /tmp/nr_demo --news 20000 --impressions 30000 --nGPU 1
"""
import argparse
import os
import time

from . import main as driver
from . import mind_io


def run(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument('--work', default='/tmp/nr_demo')
    ap.add_argument('--news', type=int, default=20000)
    ap.add_argument('--impressions', type=int, default=30000)
    ap.add_argument('--dev_impressions', type=int, default=20000)
    ap.add_argument('--title_len', type=int, default=20)
    ap.add_argument('--nGPU', type=int, default=1)
    ap.add_argument('--epochs', type=int, default=1)
    ap.add_argument('--batch_size', type=int, default=128)
    ap.add_argument('--precision', default='bf16')
    a = ap.parse_args(argv)
    tr, te, md = (os.path.join(a.work, d) for d in ('train', 'dev', 'model'))
    if int(os.environ.get('RANK', '0')) == 0 and not os.path.exists(os.path.join(te, 'behaviors.tsv')):
        t0 = time.time()
        mind_io.write_synthetic_mind_dir(tr, a.news, a.impressions, L=a.title_len, seed=0, n_cat=17, n_sub=264, table_format='raw')
        mind_io.write_synthetic_mind_dir(te, a.news * 9 // 10, a.dev_impressions, L=a.title_len, seed=1, n_cat=17, n_sub=264,
                                         table_format='raw')
        print(f'synthetic MIND directories written in {time.time() - t0:.1f} s', flush=True)
    common = ['--model', 'NAML', '--model_dir', md, '--nGPU', str(a.nGPU), '--enable_gpu', 'True', '--use_category', 'True',
              '--use_subcategory', 'True', '--freeze_embedding', 'True', '--num_words_title', str(a.title_len),
              '--precision', a.precision, '--prepare', 'True']
    driver.main(common + ['--mode', 'train', '--train_data_dir', tr, '--epochs', str(a.epochs), '--batch_size', str(a.batch_size),
                          '--lr', '0.0003', '--user_log_mask', 'False'])
    return driver.main(common + ['--mode', 'test', '--test_data_dir', te, '--batch_size', '128', '--user_log_mask', 'True',
                                 '--load_ckpt_name', f'epoch-{a.epochs}.pt'])


if __name__ == '__main__':
    run()
