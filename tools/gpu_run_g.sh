#!/bin/bash
# round-2 GPU pass G: bisect the lane-dependent / wrong sums: same failing tests against library variants
mkdir -p gpurun_out
T="tests/test_gpu_parity.py::test_cuda_vs_reference_golden_teacher_forced tests/test_gpu_parity.py::test_long_free_running_sequence_c1 tests/test_gpu_parity.py::test_large_frames_vs_oracle tests/test_imu_fusion.py::test_track_imu_equals_stage_calls_and_follows_reference"
run() { # name, env...
  n=$1; shift
  env "$@" timeout 500 python -m pytest $T -q --timeout 300 > gpurun_out/g_$n.log 2>&1
  echo "== $n: $(tail -1 gpurun_out/g_$n.log)"; grep "^FAILED" gpurun_out/g_$n.log | cut -c1-110
}
run head X=1
run legacy_ordered RBV_TUNE_ENV=8=1
run v1_old_field RBV_LIB=$PWD/rebvio_b200/variants/librebvio_b200_v1.so
run v2_old_detect RBV_LIB=$PWD/rebvio_b200/variants/librebvio_b200_v2.so
run v3_old_both RBV_LIB=$PWD/rebvio_b200/variants/librebvio_b200_v3.so
run v3_legacy RBV_LIB=$PWD/rebvio_b200/variants/librebvio_b200_v3.so RBV_TUNE_ENV=8=1
