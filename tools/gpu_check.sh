#!/bin/bash
# quick GPU check (under gpurun): ORDERED debug probe, the GPU suite (or a subset with `quick`), in-order and two-stream quick benches, kernel table
# usage: bash tools/gpu_check.sh <tag> [quick]
mkdir -p gpurun_out
export BENCH_TRACE=1
P=${1:-k}
for k in 6000 3491; do timeout 120 python tools/debug_ordered.py $k > gpurun_out/${P}_debug_$k.log 2>&1; echo "debug $k: differs=$(grep -c differs gpurun_out/${P}_debug_$k.log) $(grep -c 'bit-equal lane1: \[ True\] True' gpurun_out/${P}_debug_$k.log) bit-equal"; done
grep "mismatching" gpurun_out/${P}_debug_6000.log | cut -c1-120
[ "$2" = quick ] && PT="tests/test_gpu_parity.py -k teacher_forced" || PT=tests
timeout 1500 python -m pytest $PT -m gpu -q --timeout 400 -x > gpurun_out/${P}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${P}_pytest.log
tail -6 gpurun_out/${P}_pytest.log | cut -c1-200
Q="--no-cpu --no-e2e --no-single --no-sweep --no-fast --windows 2 --steps 20"
timeout 200 python bench.py $Q --no-overlap --kernel-table gpurun_out/${P}_ktable_inorder.csv > gpurun_out/${P}_q_inorder.json 2> gpurun_out/${P}_q_inorder.err; echo "in-order rc=$? $(python -c "import json;d=json.load(open('gpurun_out/${P}_q_inorder.json'));print(round(d['value']), round(d['ms_per_step'],4))" 2>/dev/null)"
timeout 200 python bench.py $Q > gpurun_out/${P}_q_ov2.json 2> gpurun_out/${P}_q_ov2.err; echo "overlap2 rc=$? $(python -c "import json;d=json.load(open('gpurun_out/${P}_q_ov2.json'));print(round(d['value']), round(d['ms_per_step'],4))" 2>/dev/null)"
head -24 gpurun_out/${P}_ktable_inorder.csv
