"""One repetition of the eval leg of bench.py (config 4: 65,239-news cache + 73,152 impressions, NRMS, bf16) inside a
cudaProfilerStart/Stop range, for `ncu --profile-from-start off` launch lists."""
import sys, torch
sys.path.insert(0, '.')
from newsrecommendation_b200 import engine, synthetic as S
from newsrecommendation_b200.model import NRMS
args = S.default_args(num_words_title=30, drop_rate=0.2, precision='bf16', user_log_mask=True)
model = NRMS.Model(args, S.word_table(12508)).cuda().eval()
news = S.news_titles(65238, 30, 12508)
news_d = torch.from_numpy(news).cuda()
hidx, hmask, off, cand, labels = [torch.from_numpy(x).cuda() for x in S.eval_impressions(65238, 73152, seed=500)]
for i in range(2):
    table = engine.build_news_cache(model, news_d)
    res = engine.score_impressions(model, table, hidx, hmask, off, cand, labels)
    torch.cuda.synchronize()
    if i == 0:
        torch.cuda.profiler.start()
torch.cuda.profiler.stop()
print(engine.reduce_eval_sums(res[3], 73152)['mean_scored'])
