#!/bin/bash
# Evidence after the compact title-encoder backward (same recipe as capture.sh, NRMS only: the NAML and eval kernels
# did not change): default bench line, ncu launch list of a short bench command, one-step `--set full` capture.
set -o pipefail
mkdir -p gpurun_out
python bench.py > gpurun_out/r2_bench_1gpu.json 2> gpurun_out/r2_bench_1gpu.err || exit 1
S="bench.py --single --no-cpu-baseline --no-eval --steps 2 --warmup 1"
python $S > gpurun_out/plain_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_nrms_launches.csv python $S > gpurun_out/ncu_launch.log 2>&1
P="bench.py --single --model NRMS --profile-step --no-eval"
python $P > gpurun_out/plain_step_NRMS.log 2>&1 && \
ncu --set full --clock-control none --profile-from-start off -f -o /tmp/r2_NRMS_step python $P > gpurun_out/ncu_step_NRMS.log 2>&1 && \
ncu -i /tmp/r2_NRMS_step.ncu-rep --page raw --csv > gpurun_out/r2_NRMS_step_raw.csv 2>/dev/null
ls -la gpurun_out | head -30
