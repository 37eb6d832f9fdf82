#!/bin/bash
# round-2 GPU pass J: chain loop kept at one block per trip (unroller bug); failing subset first, then the suite, traced benches, full bench, FAST-mode report
mkdir -p gpurun_out
export BENCH_TRACE=1
timeout 120 python tools/debug_ordered.py 6000 > gpurun_out/j_debug_6000.log 2>&1; grep -c differs gpurun_out/j_debug_6000.log; grep "^chain" gpurun_out/j_debug_6000.log | cut -c1-150
T="tests/test_gpu_parity.py::test_cuda_vs_reference_golden_teacher_forced tests/test_gpu_parity.py::test_long_free_running_sequence_c1 tests/test_gpu_parity.py::test_large_frames_vs_oracle tests/test_imu_fusion.py::test_track_imu_equals_stage_calls_and_follows_reference"
timeout 500 python -m pytest $T -q --timeout 300 > gpurun_out/j_subset.log 2>&1; echo "== subset: $(tail -1 gpurun_out/j_subset.log)"; grep "^FAILED" gpurun_out/j_subset.log | cut -c1-110
timeout 1500 python -m pytest tests -m gpu -q --timeout 400 > gpurun_out/j_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j_pytest.log
tail -12 gpurun_out/j_pytest.log
Q="--no-cpu --no-e2e --no-single --no-sweep --no-fast --windows 2 --steps 20"
timeout 200 python bench.py $Q --no-overlap --kernel-table gpurun_out/j_ktable_inorder.csv > gpurun_out/j_q_inorder.json 2> gpurun_out/j_q_inorder.err; echo "in-order rc=$? $(python -c "import json;d=json.load(open('gpurun_out/j_q_inorder.json'));print(round(d['value']), round(d['ms_per_step'],4))" 2>/dev/null)"
timeout 200 python bench.py $Q > gpurun_out/j_q_ov2.json 2> gpurun_out/j_q_ov2.err; echo "overlap2 rc=$? $(python -c "import json;d=json.load(open('gpurun_out/j_q_ov2.json'));print(round(d['value']), round(d['ms_per_step'],4))" 2>/dev/null)"
tail -3 gpurun_out/j_q_ov2.err
cat gpurun_out/j_ktable_inorder.csv | head -26
timeout 200 python bench.py $Q --no-overlap --tune 1=4 --kernel-table gpurun_out/j_ktable_b2.csv > gpurun_out/j_q_b2.json 2> gpurun_out/j_q_b2.err; echo "boxpass2 in-order rc=$? $(grep k_boxpass2 gpurun_out/j_ktable_b2.csv)"
timeout 700 python bench.py --kernel-table gpurun_out/j_ktable_full.csv > gpurun_out/j_bench.json 2> gpurun_out/j_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/j_bench.json'))
print('value', d['value'], 'ms', d['ms_per_step'], 'win', d['config']['window_ms'])
print('fast', d.get('fast_mode')); print('e2e', d.get('e2e',{}).get('value')); print('single', d.get('single_stream',{}).get('latency_ms_p50'), d.get('single_stream_imu',{}).get('latency_ms_p50'))
print('roofline', d['roofline']['kernel'], d['roofline']['frac'], d['roofline']['alone']['frac'])
for r in d.get('sweep', []): print(r['config'], r['lanes'], round(r['frames_per_s']), round(r['frame_frac'],3), round(r['scale_space']['frac'],3))
"
timeout 600 python tools/fast_mode_report.py 200 > gpurun_out/j_fast_mode_report.json 2> gpurun_out/j_fast_mode_report.err; echo "fast report rc=$?"; tail -3 gpurun_out/j_fast_mode_report.err; python -c "
import json; d=json.load(open('gpurun_out/j_fast_mode_report.json')); print(json.dumps(d['summary'])); print({k:{m:(v['matcj_id_forward_mismatcj_per_frame']['mean'], v['pose_rel_err_after_200_frames']) for m,v in c.items()} for k,c in d['cases'].items()})"
