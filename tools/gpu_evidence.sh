#!/bin/bash
# evidence pass (under gpurun): suite, full bench, reference arm, ncu launch list + full captures, FAST-mode report
# usage: bash tools/gpu_evidence.sh <tag> [lean]   (lean: without the suite and the FAST-mode report)
mkdir -p gpurun_out
P=${1:-p}
[ "$2" = lean ] || timeout 1500 python -m pytest tests -m gpu -q --timeout 400 > gpurun_out/${P}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${P}_pytest.log; tail -3 gpurun_out/${P}_pytest.log
timeout 900 python bench.py --kernel-table gpurun_out/${P}_ktable.csv > gpurun_out/${P}_bench.json 2> gpurun_out/${P}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/${P}_bench.json'))
print('value', round(d['value']), 'ms', round(d['ms_per_step'],4), 'win', d['config']['window_ms'], 'launches', d['gpu_launches'], d['clocks'])
print('fast', d.get('fast_mode')); print('e2e', d.get('e2e')); 
print('single', d.get('single_stream',{}).get('latency_ms_p50'), d.get('single_stream_imu',{}).get('latency_ms_p50'))
r=d['roofline']; print('roofline', r['kernel'], round(r['frac'],3), round(r['alone']['frac'],3), r.get('latency_floor'))
r=d.get('roofline_streaming_kernel'); print('streaming', r and (r['kernel'], round(r['frac'],3), round(r['alone']['frac'],3), r['alone']['avg_launch_ms']))
print('cpu', d.get('cpu_baseline'))
for r in d.get('sweep', []): print(r['config'], r['lanes'], round(r['frames_per_s']), round(r['frame_frac'],3), round(r['scale_space']['frac'],3))
PY
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${P}_ref.json 2> gpurun_out/${P}_ref.err; echo "reference rc=$? $(cut -c1-300 gpurun_out/${P}_ref.json)"
CMD="python bench.py --steps 2 --warmup 3 --windows 1 --no-cpu --no-single --no-e2e --no-sweep --no-fast"
timeout 200 $CMD > gpurun_out/${P}_plain.log 2>&1 && timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 600 --csv --log-file gpurun_out/${P}_launches.csv $CMD > gpurun_out/${P}_ncu_list.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_boxpass|k_tryvel_ord4|k_field_band|k_detect_flags|k_directed|k_extrot_ord" -s 24 -c 14 -o gpurun_out/${P}_prof $CMD > gpurun_out/${P}_ncu_full.log 2>&1; echo "full capture rc=$?"
[ "$2" = lean ] || timeout 600 python tools/fast_mode_report.py 200 > gpurun_out/${P}_fast_mode_report.json 2> gpurun_out/${P}_fast_mode_report.err; echo "fast report rc=$?"; tail -2 gpurun_out/${P}_fast_mode_report.err 2>/dev/null; true
