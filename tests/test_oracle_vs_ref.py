"""CPU: the oracle against the UNMODIFIED reference compiled into oracle/_ref/librebvio_ref.so (make -C oracle ref),
run live on seeded inputs. Skipped where the reference library is not present (it is git-ignored; it is built by
__graft_entry__.build() wherever /root/reference exists and travels to the GPU box with the snapshot)."""
import numpy as np
import pytest

import common
import harness as H
from oracle import pyref

pytestmark = pytest.mark.skipif(not pyref.available(), reason="oracle/_ref/librebvio_ref.so not built")


def _pair(rows, cols, kref=12000, kmax=16000):
    cam, det, mapc, core = common.cfgs(0.6 * cols, cols / 2.0, rows / 2.0, rows, cols, kref, kmax)
    return pyref.RefLib(cam, det, core), pyref.OracleLib(cam, det, core, mapc)


@pytest.mark.parametrize("rows,cols", [(64, 96), (97, 131), (240, 320)])
def test_setup_divisors_integral_boxpass(rows, cols):
    ref, orc = _pair(rows, cols)
    wr, sr = ref.scale_setup()
    wo, so = orc.scale_setup()
    assert list(wr) == list(wo) == [7, 7, 7, 9, 9, 9]
    assert np.array_equal(H.bits(sr), H.bits(so))
    rng = np.random.RandomState(rows * 1000 + cols)
    img = (rng.randint(0, 256, (rows, cols)).astype(np.float32) * np.float32(3.0))
    assert np.array_equal(H.bits(ref.integral(img)), H.bits(orc.integral(img)))
    for s in range(2):
        for p in range(3):
            assert np.array_equal(H.bits(ref.divisors(s, p)), H.bits(orc.divisors(s, p)))
            assert np.array_equal(H.bits(ref.box_pass(s, p, img)), H.bits(orc.box_pass(s, p, img)))
    for a, b in zip(ref.scale_space(img), orc.scale_space(img)):
        assert np.array_equal(H.bits(a), H.bits(b))
    ref.close(); orc.close()


def test_full_chain_teacher_forced_320x240():
    spec, frames, R_rel = common.sequence("C5:320x240", 4)
    cam, det, mapc, core = common.cfgs_from_spec(spec)
    ref, orc = pyref.RefLib(cam, det, core), pyref.OracleLib(cam, det, core, mapc)
    T = H.run_stages(ref, frames, R_rel)
    G = H.run_stages(orc, frames, R_rel, teacher=T)
    rep = H.report(T, G)
    assert H.dirty(rep) == []
    common.assert_x_vs_ref(T, rep)
    ref.close(); orc.close()
    # and the other way round (fresh instances: the servo state is per detector): the reference driven with the
    # oracle's free-running state
    ref, orc = pyref.RefLib(cam, det, core), pyref.OracleLib(cam, det, core, mapc)
    T2 = H.run_stages(orc, frames, R_rel)
    G2 = H.run_stages(ref, frames, R_rel, teacher=T2)
    assert H.dirty(H.report(T2, G2)) == []
    ref.close(); orc.close()


@pytest.mark.parametrize("name,nf", [("C1", 3), ("C2", 4), ("C3", 2)])
def test_full_chain_teacher_forced_benchmark_sizes(name, nf):
    """the same pin at the sizes the GPU parity tests and the bench run on (BASELINE.json configs[0..3]: 640x480, 752x480,
    1280x720 with 20 k / 32 k keylines): unmodified reference free-running, oracle teacher-forced from its state, every
    stage bit-exact; X within the accuracy of the reference's own float32 LAPACK solve (tests/common.py)"""
    spec, frames, R_rel = common.sequence(name, nf)
    cam, det, mapc, core = common.cfgs_from_spec(spec)
    ref, orc = pyref.RefLib(cam, det, core), pyref.OracleLib(cam, det, core, mapc)
    T = H.run_stages(ref, frames, R_rel)
    G = H.run_stages(orc, frames, R_rel, teacher=T)
    rep = H.report(T, G)
    assert H.dirty(rep) == []
    common.assert_x_vs_ref(T, rep)
    assert T[1]["detect"]["n"] > 5000
    ref.close(); orc.close()


def test_try_vel_nonzero_velocity_and_residual_carry():
    """tryVel at a few velocities (incl. one that pushes keylines behind the camera / out of the image) with the
    residual vector carried from call to call as minimizeVel does (core.cpp:158,141)."""
    spec, frames, R_rel = common.sequence("C5:320x240", 4)
    cam, det, mapc, core = common.cfgs_from_spec(spec)
    ref, orc = pyref.RefLib(cam, det, core), pyref.OracleLib(cam, det, core, mapc)
    maps = []
    for impl in (ref, orc):
        a = impl.detect(frames[0]); b = impl.detect(frames[1])
        impl.build_distance_field(b)
        maps.append(a)
    n = ref.map_size(maps[0])
    res_r = np.zeros(n, np.float32); res_o = np.zeros(n, np.float32)
    for vel in ([0, 0, 0], [0.01, -0.02, 0.005], [0.3, 0.2, -0.9], [0, 0, -1.5], [-0.05, 0.05, 2.0]):
        for srmin in (1e3, 15.0):
            sr, JJr, JFr = ref.try_vel(maps[0], np.float32(vel), srmin, res_r)
            so, JJo, JFo = orc.try_vel(maps[1], np.float32(vel), srmin, res_o)
            assert H.bits(np.float32(sr)) == H.bits(np.float32(so))
            assert np.array_equal(H.bits(JJr), H.bits(JJo)) and np.array_equal(H.bits(JFr), H.bits(JFo))
            first = H.first_match_index(ref.map_get(maps[0]))
            assert np.array_equal(H.bits(res_r[first:]), H.bits(res_o[first:]))
            assert np.array_equal(ref.map_get(maps[0])["match_id_forward"], orc.map_get(maps[1])["match_id_forward"])
    ref.close(); orc.close()


def test_quantile_and_servo():
    spec, frames, _ = common.sequence("C5:320x240", 4)
    cam, det, mapc, core = common.cfgs_from_spec(spec)
    ref, orc = pyref.RefLib(cam, det, core), pyref.OracleLib(cam, det, core, mapc)
    mr, mo = ref.detect(frames[0]), orc.detect(frames[0])
    kl = ref.map_get(mr)
    rng = np.random.RandomState(3)
    kl["sigma_rho"] = rng.uniform(0.0, 25.0, len(kl)).astype(np.float32)
    ref.map_set(mr, kl); orc.map_set(mo, kl)
    for pct, bins in ((0.9, 100), (0.5, 10), (0.999, 1000), (1.0, 100)):
        assert H.bits(np.float32(ref.quantile(mr, pct, bins))) == H.bits(np.float32(orc.quantile(mo, pct, bins)))
    # servo: drive both detectors from the same (threshold, last count) states
    for thr, cnt in ((0.01, 0), (0.01, 16000), (0.4999, 50000), (0.0051, 0), (0.2, 12000)):
        ref.set_threshold(thr); ref.set_last_count(cnt)
        orc.set_threshold(thr); orc.set_last_count(cnt)
        ref.detect(frames[1]); orc.detect(frames[1])
        assert H.bits(np.float32(ref.get_threshold())) == H.bits(np.float32(orc.get_threshold()))
    ref.close(); orc.close()


def test_host_imu_integrator_and_odometry_line_vs_reference():
    """rebvio_b200/imu.py (host side of SURVEY 8(f2)/(f4)): the float32 IMU sample integrator against the reference's
    types::IntegratedImu as exercised by its own pipeline (R, gyro, cacc, dt per frame), the camera extrinsics against
    Camera::getRc2i / getTc2i, and the odometry log line format (rebvio.cpp:275-283)."""
    import imu_ref_run
    from rebvio_b200 import imu, synth
    name, nf = "C5:160x120", 14
    recs = imu_ref_run.run_in_subprocess(name, nf)
    seq = synth.generate(synth.config(name, n_frames=nf))
    assert np.array_equal(recs[0]["Rc2i"], imu.R_C2I) and np.array_equal(recs[0]["tc2i"], imu.T_C2I)
    j = 0
    worst = 0.0
    for k in range(nf):
        integ = imu.IntegratedImu()
        while j < len(seq.imu_ts_us) and seq.imu_ts_us[j] <= seq.ts_us[k]:
            integ.add(seq.imu_ts_us[j], seq.gyro[j], seq.acc[j]); j += 1
        if k == 0:
            continue
        rec = integ.get(frame_dt=np.float32(np.float32(int(seq.ts_us[k]) - int(seq.ts_us[k - 1])) / 1000000.0))[0]
        r = recs[k]
        assert r["estimated"]
        for mine, ref in ((rec["R"], r["imu_R"]), (rec["gyro"], r["imu_gyro"]), (rec["cacc"], r["imu_cacc"]),
                          (rec["dt_s"], r["imu_dt_s"]), (rec["frame_dt"], r["frame_dt"])):
            a, b = np.asarray(mine, np.float64).ravel(), np.asarray(ref, np.float64).ravel()
            worst = max(worst, float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-12)))
    assert worst <= 2e-7, worst       # float32 rounding of sin / cos (numpy vs glibc float overloads); measured 0 .. 6e-8
    assert imu.odometry_line(1403637145538319104 // 1000, [0.0123456789, -1.5, 2e-7], [1, -0.25, 3.14159265]) == \
        "1403637145538319 0.012346 -1.500000 0.000000 1.000000 -0.250000 3.141593"
