#!/bin/bash
# A/B of ORDERED forms on the two-stream schedule (RBV_TUNE_ORDERED 0 / 3 / 2)
mkdir -p gpurun_out
Q="--no-cpu --no-e2e --no-single --no-sweep --no-fast --windows 3 --steps 40"
for t in 0 3 2 0 3; do
  timeout 200 python bench.py $Q --tune 8=$t > gpurun_out/n_ov_$t.json 2> gpurun_out/n_ov_$t.err; echo "ORDERED=$t overlap2 rc=$? $(python -c "import json;d=json.load(open('gpurun_out/n_ov_$t.json'));print(round(d['value']), round(d['ms_per_step'],4), d['config']['window_ms'])" 2>/dev/null)"
done
