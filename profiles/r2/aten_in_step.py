"""Which ATen kernels still run inside one eager train step?  (north_star: 'entirely through the hand-written kernels')"""
import sys, torch
sys.path.insert(0, '.')
from newsrecommendation_b200 import engine, synthetic as S
from newsrecommendation_b200.model import NRMS
args = S.default_args(num_words_title=30, drop_rate=0.2, precision='bf16')
model = NRMS.Model(args, S.word_table(12508)).cuda().train()
opt = engine.FlatAdam(model, lr=3e-4)
news = S.news_titles(65238, 30, 12508)
b = [torch.from_numpy(x).cuda() for x in S.train_batches(news, 1, 128)[0]]
for _ in range(3):
    engine.train_step(model, opt, *b)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
    engine.train_step(model, opt, *b)
    torch.cuda.synchronize()
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA and ('at::' in e.name or 'elementwise' in e.name or 'Memset' in e.name or 'Memcpy' in e.name):
        print(round(e.device_time, 1), e.name[:100])
print('---- cpu ops launching them')
for e in prof.key_averages(group_by_stack_n=4):
    if e.key in ('aten::fill_', 'aten::zero_', 'aten::zeros', 'aten::ones_like', 'aten::copy_', 'aten::add_', 'aten::mul', 'aten::div', 'aten::sum', 'aten::eq'):
        print(e.key, e.count, [s for s in e.stack[:4]])
