#!/bin/bash
# round-2 GPU pass A: parity suite, then the bench in both reduce modes (per-kernel tables written next to the lines)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/a_smi.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
timeout 900 python bench.py --reduce ordered --kernel-table gpurun_out/a_ktable_ordered.csv > gpurun_out/a_bench_ordered.json 2> gpurun_out/a_bench_ordered.err; echo "bench ordered rc=$?"
timeout 900 python bench.py --reduce fast --no-cpu --kernel-table gpurun_out/a_ktable_fast.csv > gpurun_out/a_bench_fast.json 2> gpurun_out/a_bench_fast.err; echo "bench fast rc=$?"
timeout 900 python bench.py --reduce ordered --no-overlap --no-cpu --no-e2e --no-single > gpurun_out/a_bench_ordered_inorder.json 2> gpurun_out/a_bench_ordered_inorder.err; echo "bench ordered in-order rc=$?"
cat gpurun_out/a_bench_ordered.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d.get('e2e',{}).get('value'), d.get('single_stream'), d['roofline']['kernel'], d['roofline']['frac'])"
cat gpurun_out/a_bench_fast.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d.get('e2e',{}).get('value'), d.get('single_stream'))"
