#!/bin/bash
# details page of the fused forward kernel (training variant) and the ncu launch list of the eval leg
mkdir -p gpurun_out
P="bench.py --single --profile-step --no-eval"
python $P > gpurun_out/plain_a.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:fused_enc -c 1 -f -o /tmp/r2_fused python $P > gpurun_out/ncu_a.log 2>&1 && \
ncu -i /tmp/r2_fused.ncu-rep --page details --csv > gpurun_out/r2_fused_fwd_details.csv 2>/dev/null && \
ncu -i /tmp/r2_fused.ncu-rep --page source --csv > gpurun_out/r2_fused_fwd_source.csv 2>/dev/null
E="profiles/r2/eval_only.py"
python $E > gpurun_out/plain_b.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2_eval_launches.csv python $E > gpurun_out/ncu_b.log 2>&1
ls -la gpurun_out | tail -8
