"""Which bias gradient is closer to the fp32 oracle: the one summed inside attn_bwd from fp32 fragments (compact
backward, default) or the ones-column of enc.dwqkv over the bf16-rounded dqkv (NR_DX_GATHER=0)?  Prints the max error
of W_Q/W_K/W_V.bias (and weights) of the title encoder against the oracle's autograd gradients, relative to max|grad|."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, 'tests'), os.path.join(ROOT, 'tests', 'golden')):
    sys.path.insert(0, p)
from oracle import nr_oracle as O
from test_gpu_ops import T
from test_gpu_models import build_model
from test_gpu_fused import _run

kind, kw, Pn, bt, model = build_model('nrms_ulm0', precision='bf16', drop_rate=0.0)
model.train()
os.environ.pop('NR_DX_GATHER', None)
_, _, g_c = _run(model, bt, seed=3)
os.environ['NR_DX_GATHER'] = '0'
_, _, g_a = _run(model, bt, seed=3)
P = {k: T(v).clone().requires_grad_(True) for k, v in Pn.items()}
hist, mask, cand, label = [T(a) for a in bt]
lo, so = O.nrms_forward(hist, mask, cand, label, P, 20, kw['user_log_mask'])
lo.backward()
for k in g_a:
    if 'news_encoder.multi_head' not in k:
        continue
    ref = P[k].grad.numpy()
    m = np.abs(ref).max()
    print('%-50s max|g| %.3e  compact err %.3e  all-rows err %.3e  (rel to max)' % (
        k, m, np.abs(g_c[k] - ref).max() / m, np.abs(g_a[k] - ref).max() / m))
