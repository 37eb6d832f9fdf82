#!/bin/bash
# round-2 GPU pass D: parity suite (k_boxpass2, reworked chain warps), in-order kernel tables old / new box pass, full bench, ncu of the changed kernels
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/d_pytest.log
tail -12 gpurun_out/d_pytest.log
Q="--no-cpu --no-e2e --no-single --no-sweep --no-fast --windows 2 --steps 20"
for t in "1=4" "1=2" "1=4,9=1"; do
  n=$(echo $t | tr '=,' '__')
  timeout 300 python bench.py $Q --tune $t --kernel-table gpurun_out/d_ktable_$n.csv > gpurun_out/d_q_$n.json 2> gpurun_out/d_q_$n.err
  echo "tune=$t rc=$? $(python -c "import json;d=json.load(open('gpurun_out/d_q_$n.json'));print(round(d['value']), round(d['ms_per_step'],4))" 2>/dev/null) $(grep -E 'k_tryvel_ord4|k_extrot_ord|k_boxpass' gpurun_out/d_ktable_$n.csv | tr '\n' ' ')"
done
timeout 1200 python bench.py --kernel-table gpurun_out/d_ktable.csv > gpurun_out/d_bench.json 2> gpurun_out/d_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/d_bench.json'))
print('value', d['value'], 'ms', d['ms_per_step'], 'win', d['config']['window_ms'])
print('fast', d.get('fast_mode')); print('e2e', d.get('e2e',{}).get('value')); print('single', d.get('single_stream',{}).get('latency_ms_p50'), d.get('single_stream_imu',{}).get('latency_ms_p50'))
print('roofline', d['roofline']['kernel'], d['roofline']['frac'], d['roofline']['alone']['frac'])
for r in d.get('sweep', []): print(r['config'], r['lanes'], round(r['frames_per_s']), round(r['frame_frac'],3), round(r['scale_space']['frac'],3))
"
C="python bench.py --steps 2 --warmup 1 --windows 1 --no-cpu --no-single --no-e2e --no-sweep --no-fast --no-overlap"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_boxpass2|k_tryvel_ord4|k_extrot_ord|k_field_band|k_detect_flags" -s 36 -c 12 -o gpurun_out/d_prof $C > gpurun_out/d_ncu_full.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out/d_prof.ncu-rep
