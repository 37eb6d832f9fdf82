#!/bin/bash
# ncu evidence for the bench command (run under gpurun, 1 GPU).
#   $1 = tag for the output names, $2 = kernel regex for the full capture, $3 = launches to skip, $4 = launches to capture,
#   $5 = "nolist" to skip the launch-list pass
TAG=${1:-r1}; KREGEX=${2:-k_colbox}; SKIP=${3:-8}; COUNT=${4:-3}; LIST=${5:-list}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-single --no-e2e"
if [ "$LIST" = "list" ]; then
timeout 200 $CMD > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "launch list rc=$?"
fi
timeout 200 $CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:$KREGEX" -s $SKIP -c $COUNT -o gpurun_out/prof_${TAG} $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full capture rc=$?"
