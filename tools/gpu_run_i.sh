#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/debug_ordered.py 6000 > gpurun_out/i_debug_6000.log 2>&1; echo "rc=$?"; tail -56 gpurun_out/i_debug_6000.log | cut -c1-200
