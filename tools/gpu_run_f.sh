#!/bin/bash
# round-2 GPU pass F: compute-sanitizer on small cases (memcheck, racecheck, initcheck), then the parity suite
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
timeout 600 $CS --tool memcheck --print-limit 30 python tools/sanitize_case.py stages 4 > gpurun_out/f_mem_stages.log 2>&1; echo "memcheck stages rc=$?"; grep -c "Invalid\|out of bounds" gpurun_out/f_mem_stages.log; tail -4 gpurun_out/f_mem_stages.log
timeout 600 $CS --tool memcheck --print-limit 30 python tools/sanitize_case.py batch 4 > gpurun_out/f_mem_batch.log 2>&1; echo "memcheck batch rc=$?"; grep -c "Invalid\|out of bounds" gpurun_out/f_mem_batch.log; tail -4 gpurun_out/f_mem_batch.log
timeout 600 $CS --tool racecheck --print-limit 30 python tools/sanitize_case.py stages 3 > gpurun_out/f_race_stages.log 2>&1; echo "racecheck stages rc=$?"; grep -c "hazard\|Race" gpurun_out/f_race_stages.log; tail -4 gpurun_out/f_race_stages.log
timeout 600 $CS --tool initcheck --print-limit 30 python tools/sanitize_case.py stages 3 > gpurun_out/f_init_stages.log 2>&1; echo "initcheck stages rc=$?"; grep -c "Uninitialized" gpurun_out/f_init_stages.log; tail -4 gpurun_out/f_init_stages.log
timeout 300 python tools/sanitize_case.py batch 6 C2 > gpurun_out/f_plain_batch_c2.log 2>&1; echo "plain batch C2 64 lanes rc=$?"; tail -3 gpurun_out/f_plain_batch_c2.log
timeout 300 python tools/sanitize_case.py batch2 8 C2 > gpurun_out/f_plain_batch2_c2.log 2>&1; echo "plain batch2 C2 64 lanes rc=$?"; tail -3 gpurun_out/f_plain_batch2_c2.log
timeout 1200 python -m pytest tests -m gpu -q --timeout 400 > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest.log
tail -15 gpurun_out/f_pytest.log
