#!/bin/bash
# 2-GPU check of both bench arms, launched as the driver launches them
mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 60 --warmup 3 > gpurun_out/r_bench_${N}gpu.json 2> gpurun_out/r_bench_${N}gpu.err; echo "bench N=$N rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/r_bench_${N}gpu.json').read().strip().splitlines()[-1])
print('value', round(d['value']), 'n_gpus', d['n_gpus'], 'ms', round(d['ms_per_step'],4), 'win', d['config']['window_ms'], d['clocks'])
print('fast', d.get('fast_mode',{}).get('value'), 'e2e', d.get('e2e',{}).get('value'), 'host', d['config'].get('host_threads'))
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/r_ref_${N}gpu.json 2> gpurun_out/r_ref_${N}gpu.err; echo "reference N=$N rc=$? $(tail -1 gpurun_out/r_ref_${N}gpu.json | cut -c1-200)"
tail -3 gpurun_out/r_bench_${N}gpu.err
