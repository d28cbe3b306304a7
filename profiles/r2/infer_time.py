import sys, torch, numpy as np
sys.path.insert(0, '.')
from newsrecommendation_b200 import _lib as L, synthetic as S
from newsrecommendation_b200.model import NRMS
args = S.default_args(num_words_title=30, drop_rate=0.2, precision='bf16')
model = NRMS.Model(args, S.word_table(12508)).cuda().eval()
news = torch.from_numpy(S.news_titles(65238, 30, 12508)).cuda()
for n in (7040, 16384, 65239):
    x = news[:n]
    with torch.no_grad():
        for _ in range(3): model.news_encoder(x)
        torch.cuda.synchronize()
        L.lib.nr_prof_enable(1)
        for _ in range(5): model.news_encoder(x)
        print(n, {k: (v[0], round(v[1] / 5, 4)) for k, v in L.prof_report().items()})
        L.lib.nr_prof_enable(0)
