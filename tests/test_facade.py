"""the C++ pipeline façade `rebvio_b200::Rebvio` (rebvio_b200/cpp/rebvio_b200/rebvio.hpp — the reference's
Rebvio interface, rebvio.hpp:38-113, with its two worker loops) driven by the replay program rebvio_b200/cpp/rebvio_replay.cpp.

CPU: the program builds against the C ABI, its float32 IMU integrator (types::IntegratedImu, imu.hpp:33-150) is bit-identical
to the Python restatement that is pinned against the reference's integrator, and without a device the façade fails loudly.
GPU: the odometry log of a free-running sequence follows the UNMODIFIED reference pipeline (oracle/_ref, tests/imu_ref_run.py)
within 1e-4 absolute (metres / radians), equals the Python replay's log line for line, and does not depend on whether the two
worker loops run in lock step or concurrently."""
from __future__ import annotations

import os
import shutil
import subprocess

import numpy as np
import pytest

import common
from rebvio_b200 import capi, imu, replay, synth

SEQ, NF = "C5:320x240", 40
_EXE = {}


def _exe(tmp_path_factory):
    if "p" not in _EXE:
        if not shutil.which("g++"):
            pytest.skip("no g++")
        exe = str(tmp_path_factory.mktemp("facade") / "rebvio_replay")
        libdir = os.path.dirname(capi.LIB_PATH)
        cmd = ["g++", "-std=c++14", "-O2", "-ffp-contract=off", "-Wall", "-I" + os.path.join(common.ROOT, "include"),
               "-I" + os.path.join(common.ROOT, "rebvio_b200", "cpp"), os.path.join(common.ROOT, "rebvio_b200", "cpp", "rebvio_replay.cpp"),
               "-o", exe, "-L" + libdir, "-lrebvio_b200", "-Wl,-rpath," + libdir, "-lpthread"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        assert r.returncode == 0, r.stdout
        _EXE["p"] = exe
    return _EXE["p"]


def _sequence(path, n_frames, drop_imu_before_frame=None):
    spec = synth.config(SEQ, n_frames=n_frames)
    seq = synth.generate(spec)
    its, gyro, acc = seq.imu_ts_us, seq.gyro, seq.acc
    if drop_imu_before_frame is not None:   # leave a single IMU sample in one frame interval
        k = drop_imu_before_frame
        keep = ~((its > seq.ts_us[k - 1]) & (its <= seq.ts_us[k]))
        keep[np.nonzero(~keep)[0][0]] = True
        its, gyro, acc = its[keep], gyro[keep], acc[keep]
    replay.write_sequence(path, seq.frames, seq.ts_us, its, gyro, acc, spec.cam, spec.keylines_ref, spec.keylines_max)
    return spec, seq, (its, gyro, acc)


def test_cpp_integrated_imu_is_bit_identical_to_the_pinned_restatement(tmp_path, tmp_path_factory):
    exe = _exe(tmp_path_factory)
    spec, seq, (its, gyro, acc) = _sequence(str(tmp_path / "s.rbvseq"), 12, drop_imu_before_frame=7)
    r = subprocess.run([exe, str(tmp_path / "s.rbvseq"), str(tmp_path / "imu.txt"), "--imu-only"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stderr
    lines = open(tmp_path / "imu.txt").read().strip().split("\n")
    assert len(lines) == 12
    j, prev, compared, skipped = 0, None, 0, 0
    for k, line in enumerate(lines):
        ts = int(seq.ts_us[k])
        integ = imu.IntegratedImu()
        while j < len(its) and int(its[j]) <= ts:
            integ.add(its[j], gyro[j], acc[j]); j += 1
        f = line.split()
        assert int(f[0]) == ts
        if integ.n < 2:                     # imu.hpp:81 would divide by n - 1 = 0: both sides refuse
            assert f[1] == "none"
            with pytest.raises(ValueError):
                integ.get()
            skipped += 1
        else:
            rec = integ.get(frame_dt=np.float32(np.float32(ts - prev) / 1000000.0) if prev is not None else 0.0)
            got = np.array([int(x, 16) for x in f[1:]], np.uint32)
            assert np.array_equal(got, np.frombuffer(rec.tobytes(), np.uint32)), (k, got.view(np.float32), rec)
            compared += 1
        prev = ts
    assert compared == 11 and skipped == 1


def test_facade_fails_loudly_without_a_device(tmp_path, tmp_path_factory):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    exe = _exe(tmp_path_factory)
    _sequence(str(tmp_path / "s.rbvseq"), 3)
    r = subprocess.run([exe, str(tmp_path / "s.rbvseq"), str(tmp_path / "odo.txt")], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 1 and "rbv_create" in r.stderr          # no CPU fallback behind the façade
    assert open(tmp_path / "odo.txt").read() == ""


@pytest.mark.gpu
def test_facade_odometry_follows_reference_and_python_replay(tmp_path, tmp_path_factory):
    import io
    import imu_ref_run
    exe = _exe(tmp_path_factory)
    spec, seq, _ = _sequence(str(tmp_path / "s.rbvseq"), NF)
    logs = {}
    for mode, extra in (("lockstep", []), ("concurrent", ["--no-wait"])):
        p = str(tmp_path / f"odo_{mode}.txt")
        r = subprocess.run([exe, str(tmp_path / "s.rbvseq"), p] + extra, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        print(mode, r.stderr.strip())
        logs[mode] = open(p).read().strip().split("\n")
        assert f"{NF} edge images, {NF - 1} odometry lines" in r.stderr and "running 1" in r.stderr
    assert logs["lockstep"] == logs["concurrent"]                   # the hand-over between the two loops is deterministic
    # the Python replay makes the same device calls on the same records
    cam, det, mapc, core = capi.defaults()
    cam.fm, cam.cx, cam.cy, cam.rows, cam.cols = spec.cam.fm, spec.cam.cx, spec.cam.cy, spec.cam.rows, spec.cam.cols
    det.keylines_ref, det.keylines_max = spec.keylines_ref, spec.keylines_max
    out = io.StringIO()
    replay.replay(seq.frames, seq.ts_us, seq.imu_ts_us, seq.gyro, seq.acc, cam, det, mapc, core, out=out)
    assert logs["lockstep"] == out.getvalue().strip().split("\n")
    # the unmodified reference pipeline on the same frames and raw IMU samples
    recs = imu_ref_run.run_in_subprocess(SEQ, NF)
    worst = 0.0
    for line, ref in zip(logs["lockstep"], recs[1:]):
        f = line.split(" ")
        assert len(f) == 7 and int(f[0]) == ref["ts_us"] and all(len(x.split(".")[1]) == 6 for x in f[1:])
        v = np.array([float(x) for x in f[1:]])
        worst = max(worst, float(np.abs(v[:3] - ref["orientation"]).max()), float(np.abs(v[3:] - ref["position"]).max()))
    print("facade: worst absolute pose difference over %d frames: %.2e" % (len(recs) - 1, worst))
    assert worst < 1e-4


@pytest.mark.gpu
def test_facade_stops_like_the_reference(tmp_path, tmp_path_factory):
    """rebvio.cpp:247-252: fewer matches than global_min_matches_threshold ends the run; a flat sequence has none"""
    exe = _exe(tmp_path_factory)
    spec, seq, _ = _sequence(str(tmp_path / "s.rbvseq"), 6)
    flat = np.full_like(seq.frames, 90)
    replay.write_sequence(str(tmp_path / "flat.rbvseq"), flat, seq.ts_us, seq.imu_ts_us, seq.gyro, seq.acc, spec.cam, spec.keylines_ref, spec.keylines_max)
    r = subprocess.run([exe, str(tmp_path / "flat.rbvseq"), str(tmp_path / "odo.txt")], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert "stopping" in r.stderr and "running 0" in r.stderr
    assert len(open(tmp_path / "odo.txt").read().strip().split("\n")) == 1      # the frame that tripped the floor still reports (rebvio.cpp:270-286 runs before the loop test)
