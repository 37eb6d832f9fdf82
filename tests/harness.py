"""Stage-by-stage parity harness shared by the CPU tests (oracle vs reference) and the GPU tests (CUDA vs
oracle / reference).

An "implementation" is any object with the RefLib method set (oracle/pyref.py): detect, build_distance_field,
rotate, minimize_vel_trace, forward_match, ext_rot_vel, directed_match, regularize, update_inverse_depth,
map_get / map_set / map_pixels / map_threshold / map_set_threshold / map_new, get/set_threshold, set_last_count.

`run_stages` replays the body of Rebvio::stateEstimationProcess (reference rebvio/src/rebvio.cpp:142-259) with a
minimal visual-only host fusion in place of the IMU filters (which are host code outside the hot path):
R comes from the synthetic ground truth instead of the gyro integral, and V / P_V / R0 come straight from
extRotVel's X and Wx (the reference's own path for the first frames, rebvio.cpp:194-200,229-232).

Teacher forcing: when `teacher` records are given, the implementation's maps and scalars are overwritten with
the teacher's state BEFORE every stage, so each stage is compared on identical inputs.
"""
from __future__ import annotations

import numpy as np

STAGES = ["detect", "field", "rotate1", "minimize", "forward", "extrot", "rotate2", "directed", "regularize", "kalman"]


def so3_exp(w):
    w = np.asarray(w, np.float64)
    th = np.linalg.norm(w)
    K = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    if th < 1e-12:
        return np.eye(3) + K
    return np.eye(3) + np.sin(th) / th * K + (1 - np.cos(th)) / th ** 2 * K @ K


def fusion(R_imu, Vg, Wx, X):
    """visual-only fusion (float32 outputs): R0 = exp(X[3:6]); V = R0 Vg + X[0:3]; P_V = inv(Wx)[0:3,0:3];
    Rgva = R with R^T <- R0 R^T (rebvio.cpp:193-200 without the gyro-bias filter)."""
    R0 = so3_exp(X[3:6])
    V = R0 @ np.asarray(Vg, np.float64) + np.asarray(X[0:3], np.float64)
    try:
        P = np.linalg.inv(np.asarray(Wx, np.float64))[:3, :3]
    except np.linalg.LinAlgError:
        P = np.eye(3) * 1e10
    Rt = R0 @ np.asarray(R_imu, np.float64).T
    return R0.astype(np.float32), V.astype(np.float32), P.astype(np.float32), Rt.T.astype(np.float32)


def snapshot(impl, m):
    return {"kl": impl.map_get(m).copy(), "px": impl.map_pixels(m).copy(), "thr": float(impl.map_threshold(m))}


def load(impl, m, snap):
    impl.map_set(m, snap["kl"], snap["px"])
    impl.map_set_threshold(m, snap["thr"])


def run_stages(impl, frames, R_rel, teacher=None, n_frames=None, want_field=True, iterations=5):
    """Returns a list (one per frame) of dicts stage -> outputs. frames: float32 images (0..765)."""
    recs = []
    old = None
    n_frames = n_frames or len(frames)
    for k in range(n_frames):
        rec = {}
        T = teacher[k] if teacher is not None else None
        if T is not None:
            impl.set_threshold(T["detect"]["thr_in"])
            impl.set_last_count(T["detect"]["count_in"])
        rec["detect"] = {"thr_in": float(impl.get_threshold()), "count_in": 0 if k == 0 else recs[-1]["detect"]["n"]}
        if T is not None:
            rec["detect"]["count_in"] = T["detect"]["count_in"]
        new = impl.detect(frames[k], 1000000 + 50000 * k)
        rec["detect"].update(new=snapshot(impl, new), n=impl.map_size(new), thr_out=float(impl.get_threshold()),
                             mask=impl.detector_mask().copy())
        if k == 0:
            recs.append(rec)
            old = new
            continue
        R = np.asarray(R_rel[k], np.float32)

        def force(stage_before, which=("old", "new")):
            if T is None:
                return
            if "old" in which:
                load(impl, old, T[stage_before]["old"])
            if "new" in which:
                load(impl, new, T[stage_before]["new"])

        # previous frame's end state of `old` (= last frame's `new` after kalman) is what the teacher had
        if T is not None:
            load(impl, old, teacher[k - 1]["kalman"]["new"] if k - 1 >= 1 else teacher[k - 1]["detect"]["new"])
            load(impl, new, T["detect"]["new"])
        impl.build_distance_field(new)
        rec["field"] = {}
        if want_field:
            fid, fdist = impl.field()
            rec["field"] = {"id": fid.copy(), "dist": np.where(fid >= 0, fdist, -1)}
        impl.rotate(old, R.T)
        rec["rotate1"] = {"old": snapshot(impl, old)}
        force("rotate1", ("old",))
        score, Vg, P_Vg, trace, residuals = impl.minimize_vel_trace(old, [0, 0, 0], iterations)
        rec["minimize"] = {"score": score, "Vg": Vg.copy(), "P_Vg": P_Vg.copy(), "trace": trace.copy(),
                           "residuals": residuals.copy(), "old": snapshot(impl, old)}
        force("minimize", ("old",))
        Vg_in = T["minimize"]["Vg"] if T is not None else Vg
        nf = impl.forward_match(old, new)
        rec["forward"] = {"n": nf, "new": snapshot(impl, new)}
        force("forward", ("new",))
        ok, Wx, X = impl.ext_rot_vel(new, Vg_in)
        rec["extrot"] = {"ok": ok, "Wx": Wx.copy(), "X": X.copy()}
        if T is not None:
            Wx, X = T["extrot"]["Wx"], T["extrot"]["X"]
        R0, V, P_V, Rgva = fusion(R, Vg_in, Wx, X)
        rec["extrot"].update(R0=R0, V=V, P_V=P_V, Rgva=Rgva)
        impl.rotate(old, R0)
        rec["rotate2"] = {"old": snapshot(impl, old)}
        force("rotate2", ("old",))
        n, kf = impl.directed_match(new, old, V, P_V, Rgva, 40.0)
        rec["directed"] = {"n": n, "kf": kf, "new": snapshot(impl, new), "probes": getattr(impl, "last_probes", 0)}
        force("directed", ("new",))
        nr = impl.regularize(new)
        rec["regularize"] = {"n": nr, "new": snapshot(impl, new)}
        force("regularize", ("new",))
        impl.update_inverse_depth(V)
        rec["kalman"] = {"new": snapshot(impl, new)}
        recs.append(rec)
        impl.map_free(old)
        old = new
    return recs


INT_FIELDS = ["id", "id_prev", "id_next", "match_id", "match_id_forward", "match_id_keyframe", "matches"]
FLOAT_FIELDS = ["pos", "pos_img", "match_pos_img", "gradient", "match_gradient", "gradient_norm", "match_gradient_norm",
                "rho", "sigma_rho"]


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def compare_keylines(a, b, float_exact=True, rtol=0.0, atol=0.0, fields=None):
    """Returns dict of mismatch counts; a,b structured arrays (KEYLINE_DTYPE)."""
    out = {"n": (len(a), len(b))}
    if len(a) != len(b):
        out["size_mismatch"] = 1
        return out
    for f in INT_FIELDS:
        if fields and f not in fields:
            continue
        bad = int(np.count_nonzero(a[f] != b[f]))
        if bad:
            out[f] = bad
    for f in FLOAT_FIELDS:
        if fields and f not in fields:
            continue
        if float_exact:
            bad = int(np.count_nonzero(bits(a[f]) != bits(b[f])))
        else:
            x, y = a[f].astype(np.float64), b[f].astype(np.float64)
            bad = int(np.count_nonzero(~(np.abs(x - y) <= atol + rtol * np.abs(y)) & ~(np.isnan(x) & np.isnan(y))))
        if bad:
            out[f] = bad
    return out


def clean(d):
    return {k: v for k, v in d.items() if k != "n"}


def first_match_index(kl):
    mf = kl["match_id_forward"]
    nz = np.nonzero(mf >= 0)[0]
    return int(nz[0]) if len(nz) else len(mf)


def rel_err(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    den = np.linalg.norm(b.ravel())
    return float(np.linalg.norm((a - b).ravel()) / den) if den > 0 else float(np.linalg.norm((a - b).ravel()))


def report(ref, got, float_exact=True):
    """Compare two run_stages() record lists. Returns list of per-frame dicts: stage -> {} (clean) or mismatch info."""
    out = []
    for k, (r, g) in enumerate(zip(ref, got)):
        f = {}
        d = clean(compare_keylines(r["detect"]["new"]["kl"], g["detect"]["new"]["kl"], float_exact))
        if r["detect"]["n"] != g["detect"]["n"]:
            d["count"] = (r["detect"]["n"], g["detect"]["n"])
        if not np.array_equal(r["detect"]["new"]["px"], g["detect"]["new"]["px"]):
            d["pixels"] = int(np.count_nonzero(r["detect"]["new"]["px"] != g["detect"]["new"]["px"])) if len(r["detect"]["new"]["px"]) == len(g["detect"]["new"]["px"]) else -1
        if not np.array_equal(r["detect"]["mask"], g["detect"]["mask"]):
            d["mask"] = int(np.count_nonzero(r["detect"]["mask"] != g["detect"]["mask"]))
        if bits(r["detect"]["new"]["thr"]) != bits(g["detect"]["new"]["thr"]):
            d["map_threshold"] = (r["detect"]["new"]["thr"], g["detect"]["new"]["thr"])
        f["detect"] = d
        if k == 0:
            out.append(f)
            continue
        d = {}
        if "id" in r["field"] and "id" in g["field"]:
            if not np.array_equal(r["field"]["id"], g["field"]["id"]):
                d["id"] = int(np.count_nonzero(r["field"]["id"] != g["field"]["id"]))
            if not np.array_equal(r["field"]["dist"], g["field"]["dist"]):
                d["dist"] = int(np.count_nonzero(r["field"]["dist"] != g["field"]["dist"]))
        f["field"] = d
        f["rotate1"] = clean(compare_keylines(r["rotate1"]["old"]["kl"], g["rotate1"]["old"]["kl"], float_exact))
        d = clean(compare_keylines(r["minimize"]["old"]["kl"], g["minimize"]["old"]["kl"], float_exact))
        tr_r, tr_g = r["minimize"]["trace"], g["minimize"]["trace"]
        first = first_match_index(r["minimize"]["old"]["kl"])
        nb = int(np.count_nonzero(bits(tr_r[:, :15]) != bits(tr_g[:, :15])))
        if nb:
            d["trace_bits"] = nb
            d["trace_rel"] = max(rel_err(tr_g[i, 3:12], tr_r[i, 3:12]) for i in range(len(tr_r)))
        ns = int(np.count_nonzero(bits(tr_r[:, 15]) != bits(tr_g[:, 15])))
        if ns:
            d["score_bits"] = ns
            d["score_rel"] = float(np.max(np.abs(tr_g[:, 15] - tr_r[:, 15]) / np.maximum(np.abs(tr_r[:, 15]), 1e-30)))
        if not np.array_equal(bits(r["minimize"]["Vg"]), bits(g["minimize"]["Vg"])):
            d["Vg_rel"] = rel_err(g["minimize"]["Vg"], r["minimize"]["Vg"])
        if not np.array_equal(bits(r["minimize"]["P_Vg"]), bits(g["minimize"]["P_Vg"])):
            d["P_Vg_rel"] = rel_err(g["minimize"]["P_Vg"], r["minimize"]["P_Vg"])
        rr, gr = r["minimize"]["residuals"][first:], g["minimize"]["residuals"][first:]
        nres = int(np.count_nonzero(bits(rr) != bits(gr))) if len(rr) == len(gr) else -1
        if nres:
            d["residuals"] = nres
        f["minimize"] = d
        d = clean(compare_keylines(r["forward"]["new"]["kl"], g["forward"]["new"]["kl"], float_exact))
        if r["forward"]["n"] != g["forward"]["n"]:
            d["count"] = (r["forward"]["n"], g["forward"]["n"])
        f["forward"] = d
        d = {}
        if not np.array_equal(bits(r["extrot"]["Wx"]), bits(g["extrot"]["Wx"])):
            d["Wx_bits"] = int(np.count_nonzero(bits(r["extrot"]["Wx"]) != bits(g["extrot"]["Wx"])))
            d["Wx_rel"] = rel_err(g["extrot"]["Wx"], r["extrot"]["Wx"])
        d["X_rel"] = rel_err(g["extrot"]["X"], r["extrot"]["X"])
        if r["extrot"]["ok"] != g["extrot"]["ok"]:
            d["ok"] = (r["extrot"]["ok"], g["extrot"]["ok"])
        f["extrot"] = d
        f["rotate2"] = clean(compare_keylines(r["rotate2"]["old"]["kl"], g["rotate2"]["old"]["kl"], float_exact))
        d = clean(compare_keylines(r["directed"]["new"]["kl"], g["directed"]["new"]["kl"], float_exact))
        if (r["directed"]["n"], r["directed"]["kf"]) != (g["directed"]["n"], g["directed"]["kf"]):
            d["count"] = ((r["directed"]["n"], r["directed"]["kf"]), (g["directed"]["n"], g["directed"]["kf"]))
        f["directed"] = d
        d = clean(compare_keylines(r["regularize"]["new"]["kl"], g["regularize"]["new"]["kl"], float_exact))
        if r["regularize"]["n"] != g["regularize"]["n"]:
            d["count"] = (r["regularize"]["n"], g["regularize"]["n"])
        f["regularize"] = d
        f["kalman"] = clean(compare_keylines(r["kalman"]["new"]["kl"], g["kalman"]["new"]["kl"], float_exact))
        out.append(f)
    return out


def dirty(rep, ignore=("X_rel",)):
    """stages with any mismatch, as [(frame, stage, info)]"""
    bad = []
    for k, f in enumerate(rep):
        for stage, d in f.items():
            dd = {a: b for a, b in d.items() if a not in ignore}
            if dd:
                bad.append((k, stage, dd))
    return bad


def run_lean(impl, frames, R_rel, snap_every=0, want_ids=False):
    """Free-running replay of the same loop as run_stages for LONG sequences: no per-stage map snapshots. Per frame it keeps
    the scalars every stage hands back (counts, Vg, P_Vg, score, Wx, X, V, R0), optionally the two match-index arrays the
    tracker decides (old map's match_id_forward after minimizeVel, new map's match_id after directedMatch), and the whole new map
    every `snap_every` frames and at the end. The pose is integrated like Rebvio::stateEstimationProcess does for its odometry
    (rebvio.cpp:263-273): R_k = R_{k-1} R0_k^T ... kept simple here as position += R_acc V, R_acc = R_acc R0^T (float64)."""
    recs, old = [], None
    Racc, pos = np.eye(3), np.zeros(3)
    for k in range(len(frames)):
        new = impl.detect(frames[k], 1000000 + 50000 * k)
        rec = {"n": impl.map_size(new)}
        if k == 0:
            recs.append(rec); old = new
            continue
        R = np.asarray(R_rel[k], np.float32)
        impl.build_distance_field(new)
        impl.rotate(old, R.T)
        score, Vg, P_Vg, trace, _ = impl.minimize_vel_trace(old, [0, 0, 0], 5)
        if want_ids:
            rec["fwd_ids"] = impl.map_get(old)["match_id_forward"].copy()
        rec.update(score=score, Vg=np.array(Vg), P_Vg=np.array(P_Vg), trace=np.array(trace))
        rec["nf"] = impl.forward_match(old, new)
        ok, Wx, X = impl.ext_rot_vel(new, Vg)
        R0, V, P_V, Rgva = fusion(R, Vg, Wx, X)
        rec.update(ok=ok, Wx=np.array(Wx), X=np.array(X), V=V, R0=R0)
        impl.rotate(old, R0)
        n, kf = impl.directed_match(new, old, V, P_V, Rgva, 40.0)
        if want_ids:
            rec["match_ids"] = impl.map_get(new)["match_id"].copy()
        rec.update(nm=n, kf=kf, nr=impl.regularize(new))
        impl.update_inverse_depth(V)
        pos = pos + Racc @ np.asarray(V, np.float64)
        Racc = Racc @ np.asarray(R0, np.float64).T
        rec.update(pos=pos.copy(), Racc=Racc.copy())
        if snap_every and (k % snap_every == 0 or k == len(frames) - 1):
            rec["new"] = snapshot(impl, new)
        recs.append(rec)
        impl.map_free(old)
        old = new
    return recs
