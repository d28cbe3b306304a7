#!/bin/bash
# Final multi-GPU runs of round 2 on one 8-GPU box (torchrun exactly as the driver launches bench.py).
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
mkdir -p gpurun_out
timeout 400 $TR --nproc-per-node 8 --master-port 29601 bench.py --gpus 8 > gpurun_out/r2f_8gpu_default.json 2> gpurun_out/r2f_8gpu_default.err; echo default8 rc=$?
timeout 200 $TR --nproc-per-node 4 --master-port 29602 bench.py --gpus 4 --single --no-eval > gpurun_out/r2f_4gpu_nrms.json 2> /dev/null; echo nrms4 rc=$?
timeout 200 $TR --nproc-per-node 2 --master-port 29603 bench.py --gpus 2 --single --no-eval > gpurun_out/r2f_2gpu_nrms.json 2> /dev/null; echo nrms2 rc=$?
timeout 200 python bench.py --single --no-eval --no-cpu-baseline > gpurun_out/r2f_1gpu_nrms.json 2> /dev/null; echo nrms1 rc=$?
timeout 500 $TR --nproc-per-node 8 --master-port 29604 bench.py --gpus 8 --batch 1024 --vocab 70000 --news 160000 --eval-impressions 376000 > gpurun_out/r2f_8gpu_config5.json 2> gpurun_out/r2f_8gpu_config5.err; echo config5 rc=$?
