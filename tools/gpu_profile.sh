#!/bin/bash
# ncu evidence for the bench command (run under gpurun, 1 GPU). $1 = tag for the output names, $2 = kernel regex for the full capture
TAG=${1:-r1}; KREGEX=${2:-k_colbox}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-single --no-e2e"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KREGEX -s 8 -c 3 -o gpurun_out/prof_${TAG}_$KREGEX $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full capture rc=$?"
