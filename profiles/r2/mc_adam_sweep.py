"""Size sweep of the fused reduce-scatter + Adam + all-gather kernel (nr_adam_step_dev with multicast pointers) on the
ranks of one box: is it bound by latency or by NVLink bandwidth?   torchrun --nproc-per-node N profiles/r2/mc_adam_sweep.py"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, '.')
import torch.distributed._symmetric_memory as symm_mem
from newsrecommendation_b200 import _lib as L
rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
torch.cuda.set_device(rank)
dist.init_process_group('nccl', device_id=torch.device('cuda', rank))
N = 16 << 20
p = symm_mem.empty(N, dtype=torch.float32, device='cuda'); g = symm_mem.empty(N, dtype=torch.float32, device='cuda')
hp, hg = symm_mem.rendezvous(p, dist.group.WORLD), symm_mem.rendezvous(g, dist.group.WORLD)
p.zero_(); g.fill_(1e-3)
m = torch.zeros(N, device='cuda'); v = torch.zeros(N, device='cuda'); step = torch.ones(1, dtype=torch.int64, device='cuda')
torch.cuda.synchronize(); dist.barrier()
for floats in (1 << 14, 1 << 16, 1 << 18, 600_000, 1 << 21, 1 << 23):
    lo, hi = rank * floats, (rank + 1) * floats
    if hi > N: break
    def run():
        hg.barrier(channel=0)
        L.call('nr_adam_step_dev', None, hg.multicast_ptr, L.ptr(p), hp.multicast_ptr, L.ptr(m), L.ptr(v), lo, hi, 3e-4, 0.9, 0.999, 1e-8,
               L.ptr(step), 1.0 / world, L.stream())
        hp.barrier(channel=1)
    for _ in range(3): run()
    torch.cuda.synchronize(); dist.barrier()
    L.lib.nr_prof_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): run()
    e1.record(); torch.cuda.synchronize()
    rep = L.prof_report(); L.lib.nr_prof_enable(0)
    if rank == 0:
        k = rep['adam_reduce_scatter_gather']
        print(f'slice {floats * 4 / 1e6:8.3f} MB  kernel {k[1] / k[0] * 1e3:7.1f} us   barrier+kernel+barrier {e0.elapsed_time(e1) / 20 * 1e3:7.1f} us', flush=True)
dist.barrier(); dist.destroy_process_group()
