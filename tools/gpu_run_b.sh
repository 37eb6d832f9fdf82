#!/bin/bash
# round-2 GPU pass B: chain-warp spread sweep (in-order kernel tables), parity suite, full bench, ncu captures of the big kernels
mkdir -p gpurun_out
Q="--no-cpu --no-e2e --no-single --no-sweep --no-fast --windows 2 --steps 20"
for rw in 1 2 4 8; do
  timeout 300 python bench.py $Q --tune 9=$rw --kernel-table gpurun_out/b_ktable_rw$rw.csv > gpurun_out/b_rw$rw.json 2> gpurun_out/b_rw$rw.err
  echo "rw=$rw rc=$? $(python -c "import json;d=json.load(open('gpurun_out/b_rw$rw.json'));print(d['value'], d['ms_per_step'])" 2>/dev/null) $(grep -E 'k_tryvel_ord4|k_extrot_ord' gpurun_out/b_ktable_rw$rw.csv | tr '\n' ' ')"
done
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -8 gpurun_out/b_pytest.log
timeout 1200 python bench.py --kernel-table gpurun_out/b_ktable.csv > gpurun_out/b_bench.json 2> gpurun_out/b_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/b_bench.json'))
print('value', d['value'], 'ms', d['ms_per_step'], 'win', d['config']['window_ms'])
print('fast', d.get('fast_mode')); print('e2e', d.get('e2e',{}).get('value')); print('single', d.get('single_stream',{}).get('latency_ms_p50'), d.get('single_stream_imu',{}).get('latency_ms_p50'))
print('roofline', d['roofline']['kernel'], d['roofline']['frac'], d['roofline']['alone']['frac'])
for r in d.get('sweep', []): print(r['config'], r['lanes'], round(r['frames_per_s']), round(r['frame_frac'],3), round(r['scale_space']['frac'],3))
print('cpu', d.get('cpu_baseline'))
"
C="python bench.py --steps 2 --warmup 1 --windows 1 --no-cpu --no-single --no-e2e --no-sweep --no-fast --no-overlap"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_field_band|k_detect_flags|k_boxpass|k_tryvel_ord4|k_directed|k_extrot_ord|k_detect_emit" -s 40 -c 14 -o gpurun_out/b_prof $C > gpurun_out/b_ncu_full.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out/b_prof.ncu-rep
