#!/bin/bash
# round-2 GPU pass M: chain CTAs inside the evaluation launch
mkdir -p gpurun_out
P=${1:-m}
python tools/chain_timing.py 6000 2 2>&1 | tail -8
python tools/chain_timing.py 6000 64 2>&1 | tail -8
bash tools/gpu_run_k.sh $P $2
