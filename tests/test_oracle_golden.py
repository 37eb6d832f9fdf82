"""CPU: the oracle (oracle/rebvio_oracle.c) against the committed golden vectors, which were produced by the
UNMODIFIED reference compiled from /root/reference (tools/make_golden.py). Bit-exact on every stage when each
stage is fed the reference's inputs (teacher forcing); X of extRotVel within the stated tolerance."""
import numpy as np
import pytest

import common
import harness as H
from oracle import pyref


@pytest.mark.parametrize("name", common.GOLDEN_CASES)
def test_oracle_matches_golden_teacher_forced(name):
    frames, R_rel, meta, recs, _ = common.load_golden(name)
    cam, det, mapc, core = common.cfgs_from_meta(meta)
    orc = pyref.OracleLib(cam, det, core, mapc)
    got = H.run_stages(orc, frames, R_rel, teacher=recs)
    rep = H.report(recs, got)
    assert H.dirty(rep) == []
    common.assert_x_vs_ref(recs, rep)
    orc.close()


@pytest.mark.parametrize("name", common.GOLDEN_CASES)
def test_oracle_free_running_until_first_solve(name):
    """Without teacher forcing everything up to (and including) Wx of the first extRotVel is still bit-exact;
    after that the float32-SVD vs FP64-Jacobi difference in X feeds back into the rotations."""
    frames, R_rel, meta, recs, _ = common.load_golden(name)
    cam, det, mapc, core = common.cfgs_from_meta(meta)
    orc = pyref.OracleLib(cam, det, core, mapc)
    got = H.run_stages(orc, frames, R_rel, n_frames=2)
    rep = H.report(recs[:2], got)
    for stage in ("detect", "field", "rotate1", "minimize", "forward", "extrot"):
        d = {k: v for k, v in rep[1][stage].items() if k != "X_rel"}
        assert d == {}, (stage, d)
    assert rep[0]["detect"] == {}
    orc.close()


@pytest.mark.parametrize("name", ["g160", "g131"])
def test_oracle_scale_space_golden(name):
    frames, _, meta, _, extra = common.load_golden(name)
    cam, det, mapc, core = common.cfgs_from_meta(meta)
    orc = pyref.OracleLib(cam, det, core, mapc)
    s0, s1, dog, mag = orc.scale_space(frames[0])
    for a, key in ((s0, "scale/s0"), (s1, "scale/s1"), (dog, "scale/dog"), (mag, "scale/mag")):
        assert np.array_equal(H.bits(a), H.bits(extra[key])), key
    orc.close()


def test_golden_hits_keyline_cap():
    """g131 exercises the keylines_max cut (edge_detector.cpp:113-119): every frame is capped and the detector mask
    holds no id beyond the cap."""
    _, _, meta, recs, _ = common.load_golden("g131")
    for r in recs:
        assert r["detect"]["n"] == meta["kmax"]
        assert r["detect"]["mask"].max() == meta["kmax"] - 1
