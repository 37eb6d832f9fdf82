#!/bin/bash
# stream-priority A/B on the two-stream schedule (ORDERED and FAST)
mkdir -p gpurun_out
Q="--no-cpu --no-e2e --no-single --no-sweep --no-fast --windows 3 --steps 40"
for mode in ordered fast; do
for pr in "-1 0" "0 0" "0 -1" "-1 -1"; do
  set -- $pr
  BENCH_STREAM_PRIORITY=$1 RBV_FE_PRIORITY=$2 timeout 200 python bench.py $Q --reduce $mode > gpurun_out/s.json 2> gpurun_out/s.err; echo "$mode tracker=$1 fe=$2 rc=$? $(python -c "import json;d=json.load(open('gpurun_out/s.json'));print(round(d['value']), round(d['ms_per_step'],4), d['config']['window_ms'])" 2>/dev/null)"
done
done
