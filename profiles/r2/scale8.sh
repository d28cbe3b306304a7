#!/bin/bash
# 8-GPU runs of round 2 (one box, torchrun exactly as the driver launches bench.py).  Outputs: gpurun_out/r2_*.json
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
mkdir -p gpurun_out
timeout 400 $TR --master-port 29601 bench.py --gpus 8 > gpurun_out/r2_8gpu_default.json 2> gpurun_out/r2_8gpu_default.err; echo default rc=$?
NR_FUSED_ALLREDUCE=0 timeout 300 $TR --master-port 29602 bench.py --gpus 8 --single --no-eval > gpurun_out/r2_8gpu_nrms_nccl.json 2> gpurun_out/r2_8gpu_nrms_nccl.err; echo nccl rc=$?
timeout 500 $TR --master-port 29603 bench.py --gpus 8 --batch 1024 --vocab 70000 --news 160000 --eval-impressions 376000 > gpurun_out/r2_8gpu_config5.json 2> gpurun_out/r2_8gpu_config5.err; echo config5 rc=$?
NR_FUSED_ALLREDUCE=0 timeout 300 $TR --master-port 29604 bench.py --gpus 8 --single --no-eval --batch 1024 --vocab 70000 --news 160000 > gpurun_out/r2_8gpu_config5_nrms_nccl.json 2> gpurun_out/r2_8gpu_config5_nrms_nccl.err; echo config5-nccl rc=$?
tail -2 gpurun_out/r2_8gpu_default.err gpurun_out/r2_8gpu_config5.err
