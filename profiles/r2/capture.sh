#!/bin/bash
# Round-2 evidence on one B200: the default bench line, the ncu launch list of a short bench command and one-step
# `--set full` captures (NRMS, NAML, eval leg).  Each ncu run follows a plain run of the same command that exited 0.
set -o pipefail
mkdir -p gpurun_out
python bench.py > gpurun_out/r2_bench_1gpu.json 2> gpurun_out/r2_bench_1gpu.err || exit 1
python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2> gpurun_out/r2_ref.err || exit 2
S="bench.py --single --no-cpu-baseline --no-eval --steps 2 --warmup 1"
python $S > gpurun_out/plain_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_nrms_launches.csv python $S > gpurun_out/ncu_launch.log 2>&1
for M in NRMS NAML; do
  P="bench.py --single --model $M --profile-step --no-eval"
  python $P > gpurun_out/plain_step_$M.log 2>&1 && \
  ncu --set full --clock-control none --profile-from-start off -f -o /tmp/r2_${M}_step python $P > gpurun_out/ncu_step_$M.log 2>&1 && \
  ncu -i /tmp/r2_${M}_step.ncu-rep --page raw --csv > gpurun_out/r2_${M}_step_raw.csv 2>/dev/null
done
ls -la gpurun_out | head -30
