#!/bin/bash
# A/B of library variants on the two-stream schedule: tools/gpu_ab_variants.sh <lib names...> ("main" = the shipped library)
mkdir -p gpurun_out
Q="--no-cpu --no-e2e --no-single --no-sweep --no-fast --windows 3 --steps 40"
for rep in 1 2; do
for spec in "$@"; do
  v=${spec%%:*}; T=""; [ "$spec" != "$v" ] && T="--tune ${spec#*:}"
  if [ "$v" = main ]; then unset RBV_LIB; else export RBV_LIB=$PWD/rebvio_b200/lib_$v.so; fi
  timeout 200 python bench.py $Q $T > gpurun_out/q_$v.json 2> gpurun_out/q_$v.err; echo "$spec overlap2 rc=$? $(python -c "import json;d=json.load(open('gpurun_out/q_$v.json'));print(round(d['value']), round(d['ms_per_step'],4), d['config']['window_ms'])" 2>/dev/null)"
done
done
