#!/usr/bin/env python
"""bench.py — frames/s of the REBVIO hot path (edge front end + edge tracker) on B200, against the HBM roofline and
next to the reference's own CPU implementation.

Workload (BASELINE.json `metric`: "frames/sec @752x480 (1 GPU p50 latency; batched streams at 1/2/4/8 GPU)"):
  * `value`  — batched replay (configs[3]): 64 independent EuRoC-shaped 752x480 sequences per GPU (the C2 generator with
    seeds 1000 + 64 g + i), one "step" = one frame of every sequence through detect + tracker (rbv_detect_device,
    rbv_track_g1, host fusion, rbv_track_g2). Frames resident in HBM before the timed region. Whole-job frames/s.
  * `e2e`    — the same through the public host-buffer API (BatchedPipeline.push(host frames) -> rbv_detect): every step
    copies that step's frames from pinned host memory to the device and reads counts / velocities / 6x6 systems back.
  * `single_stream` (N = 1) — configs[1]: one 752x480 stream, lanes = 1, p50 per-frame latency.
Multi-GPU: replicas only (one frame does not shard); ranks share nothing, `scaling` = "weak".
`--impl reference` times the UNMODIFIED reference (oracle/_ref/librebvio_ref_fast.so, compiled from /root/reference with
the reference's own optimisation flags) on the same workload with every host core; no GPU code on that arm.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np

METRIC = "frames/sec @752x480 (batched independent streams, whole job)"
LANES_PER_GPU = 64


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--lanes", type=int, default=LANES_PER_GPU)
    ap.add_argument("--config", default="C2", help="sequence shape of every lane (C2 = 752x480 EuRoC-shaped)")
    ap.add_argument("--reduce", default="ordered", choices=["fast", "ordered"])
    ap.add_argument("--no-single", action="store_true", help="skip the single-stream latency leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--undistort", action="store_true", help="e2e leg: raw camera frames, Camera::undistort (EuRoC cam0 model) on the device")
    ap.add_argument("--imu", action="store_true", help="run the reference's IMU filters on the device between the tracker halves "
                    "(rbv_track_imu: gyro-bias correction + scale-attitude-bias filter, SURVEY 8(f2)) instead of the visual-only fusion")
    ap.add_argument("--no-overlap", action="store_true", help="in-order schedule on one stream (detect, then the tracker of the same frame) "
                    "instead of the reference's two-thread schedule on two streams (detect of frame n+1 next to the tracker of n-1, n)")
    ap.add_argument("--overlap", type=int, default=2, choices=[1, 2], help="1 = host threshold servo, front end of frame n+1 next to the tracker "
                    "of (n-1, n); 2 = servo on the device, the front end runs one more frame ahead (BatchedPipeline overlap modes)")
    ap.add_argument("--tune", default="", help="implementation choices, e.g. '2=4,0=1' (rbv_set_tuning key=value pairs)")
    ap.add_argument("--kernel-table", default="", help="write the per-kernel CUDA-event table (one untimed step) here")
    return ap.parse_args()


def cfg_structs(spec):
    from rebvio_b200 import capi
    cam, det, mapc, core = capi.defaults()
    cam.fm, cam.cx, cam.cy, cam.rows, cam.cols = spec.cam.fm, spec.cam.cx, spec.cam.cy, spec.cam.rows, spec.cam.cols
    det.keylines_ref, det.keylines_max = spec.keylines_ref, spec.keylines_max
    return cam, det, mapc, core


class ClockSampler(threading.Thread):
    """samples SM clock and throttle reasons through NVML while the timed region runs"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def result(self):
        self.stop_flag = True
        self.join(timeout=1.0)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def cpu_reference_frames_per_s(spec, frames_u8, R_rel, n_frames, fast=True):
    """the UNMODIFIED reference on one core: detect + tracker stages per frame (same replay driver as the tests)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import pyref
    import harness as H
    cam, det, mapc, core = pyref.default_cfgs(cam=dict(fm=spec.cam.fm, cx=spec.cam.cx, cy=spec.cam.cy, rows=spec.cam.rows, cols=spec.cam.cols),
                                              det=dict(keylines_ref=spec.keylines_ref, keylines_max=spec.keylines_max))
    kind = "reference"
    try:
        impl = pyref.RefLib(cam, det, core, fast=fast)
    except (FileNotFoundError, OSError):
        impl = pyref.OracleLib(cam, det, core, mapc)   # the plain-C port when the compiled reference did not travel
        kind = "port"
    frames = [np.ascontiguousarray(f, np.float32) * np.float32(3.0) for f in frames_u8[:n_frames]]
    t0 = time.perf_counter()
    H.run_stages(impl, frames, R_rel[:n_frames], want_field=False)
    dt = time.perf_counter() - t0
    impl.close()
    return n_frames / dt, kind


# ---------------------------------------------------------------------------------------------------------------
def run_reference(args):
    """--impl reference: every host core runs the reference on its own sequence; one step = one frame per core."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import multiprocessing as mp
    from rebvio_b200 import synth
    spec = synth.config(args.config)
    cores = len(os.sched_getaffinity(0))
    nseq = min(8, cores)
    nf = args.warmup + args.steps + 1
    try:
        import torch
        dev = "cuda" if torch.cuda.is_available() else "cpu"
    except Exception:
        dev = "cpu"
    frames, R_rel = synth.batched_sequences(spec, [1000 + i for i in range(nseq)], nf, device=dev, n_textures=2)
    frames = frames.cpu().numpy()
    ctx = mp.get_context("spawn")
    conns, procs = [], []
    for w in range(cores):
        a, b = ctx.Pipe()
        p = ctx.Process(target=_ref_worker, args=(b, args.config, frames[:, w % nseq], R_rel[:, w % nseq], w), daemon=True)
        p.start()
        conns.append(a); procs.append(p)
    kinds = [c.recv() for c in conns]   # ready (frame 0 detected)
    for _ in range(args.warmup):
        for c in conns: c.send("step")
        for c in conns: c.recv()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        for c in conns: c.send("step")
        for c in conns: c.recv()
    dt = time.perf_counter() - t0
    for c in conns: c.send("quit")
    for p in procs: p.join(timeout=5)
    fps = cores * args.steps / dt
    line = {"metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "reference",
            "config": {"workload": f"{args.config}-shaped {spec.cam.cols}x{spec.cam.rows} independent sequences, one per host core "
                                   f"({cores} cores, {nseq} distinct sequences); step = one frame per core through the reference's detect + tracker stages"},
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": kinds[0],
                             "sample": f"{args.steps} frames on each of {cores} processes (reference -O3 build)"},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)
    return 0


def _ref_worker(conn, config, frames_u8, R_rel, w):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        os.sched_setaffinity(0, {sorted(os.sched_getaffinity(0))[w % len(os.sched_getaffinity(0))]})
    except Exception:
        pass
    from rebvio_b200 import synth
    from oracle import pyref
    import harness as H
    spec = synth.config(config)
    cam, det, mapc, core = pyref.default_cfgs(cam=dict(fm=spec.cam.fm, cx=spec.cam.cx, cy=spec.cam.cy, rows=spec.cam.rows, cols=spec.cam.cols),
                                              det=dict(keylines_ref=spec.keylines_ref, keylines_max=spec.keylines_max))
    try:
        impl, kind = pyref.RefLib(cam, det, core, fast=True), "reference"
    except (FileNotFoundError, OSError):
        impl, kind = pyref.OracleLib(cam, det, core, mapc), "port"
    f32 = lambda k: np.ascontiguousarray(frames_u8[k], np.float32) * np.float32(3.0)
    old = impl.detect(f32(0), 0)
    conn.send(kind)
    k = 1
    while True:
        if conn.recv() != "step":
            break
        img = f32(k)
        new = impl.detect(img, k)
        impl.build_distance_field(new)
        R = np.asarray(R_rel[k], np.float32)
        impl.rotate(old, R.T)
        s, Vg, P_Vg = impl.minimize_vel(old, [0, 0, 0])
        impl.forward_match(old, new)
        ok, Wx, X = impl.ext_rot_vel(new, Vg)
        R0, V, P_V, Rgva = H.fusion(R, Vg, Wx, X)
        impl.rotate(old, R0)
        impl.directed_match(new, old, V, P_V, Rgva, 40.0)
        impl.regularize(new)
        impl.update_inverse_depth(V)
        impl.map_free(old)
        old = new
        k += 1
        conn.send(1)


# ---------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from rebvio_b200 import capi, synth, roofline
    from rebvio_b200.pipeline import BatchedPipeline

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    spec = synth.config(args.config)
    cam, det, mapc, core = cfg_structs(spec)
    B, K, W = args.lanes, args.steps, args.warmup
    N = spec.cam.rows * spec.cam.cols
    lag = 0 if args.no_overlap else args.overlap   # overlapped schedule: push(k) enqueues detect(k) next to the tracker of the pair (k-lag-1, k-lag)
    first = 1 + lag                          # first push that has a pair to track
    nf = first + W + K
    seeds = [1000 + LANES_PER_GPU * rank + i for i in range(B)]
    frames_u8, R_rel = synth.batched_sequences(spec, seeds, nf, device=dev, n_textures=2)
    d_frames = frames_u8.to(torch.float32) * 3.0                     # rebvio.cpp:43 conversion, resident in HBM
    torch.cuda.synchronize()
    mode = capi.REDUCE_FAST if args.reduce == "fast" else capi.REDUCE_ORDERED
    stream = torch.cuda.Stream(device=dev, priority=int(os.environ.get("BENCH_STREAM_PRIORITY", "-1")))   # tracker stream above the library's front-end stream (see rbv_create)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    imu_cfg = True if args.imu else None

    def targs(k, lanes=slice(None)):
        """tracker arguments of frame k: the inter-frame rotation, or (--imu) an integrated-IMU record per lane built from it
        (R = R_rel, gravity-only compensated acceleration, 50 ms interval)"""
        R = np.asarray(R_rel[k, lanes], np.float32)
        if not args.imu:
            return dict(R_rel=R)
        a = np.zeros(R.shape[0], capi.IMU_IN)
        a["R"] = R.reshape(-1, 9); a["cacc"] = (0.0, -9.81, 0.0); a["dt_s"] = 1.0 / spec.fps; a["frame_dt"] = 1.0 / spec.fps
        return dict(imu_in=a)

    # ---- leg 1: device-resident inputs ----------------------------------------------------------------------
    pipe = BatchedPipeline(cam, det, mapc, core, B, device=local, reduce_mode=mode, cuda_stream=stream.cuda_stream, imu_cfg=imu_cfg,
                           overlap=lag)
    tune = [tuple(int(x) for x in kv.split("=")) for kv in args.tune.split(",") if kv]
    for key, val in tune:
        pipe.ctx.set_tuning(key, val)
    ptr = lambda k: d_frames[k].data_ptr()
    pargs = lambda k: targs(k - lag) if k >= first else {}     # arguments of the pair push(k) tracks
    for k in range(first):
        pipe.push(device_ptr=ptr(k))
    # warm-up; the first warm-up step runs with every kernel bracketed by CUDA events (in order, on one stream) to find the dominant kernel
    pipe.ctx.sync()                          # front ends enqueued ahead have drained: the profiled step runs alone
    pipe.ctx.profile_reset(); pipe.ctx.profile_set(-2)
    out = pipe.push(device_ptr=ptr(first), **pargs(first))
    pipe.ctx.sync()
    table = pipe.ctx.profile_read()
    pipe.ctx.profile_set(-1)
    for k in range(first + 1, first + W):
        out = pipe.push(device_ptr=ptr(k), **pargs(k))
    top = max(table, key=lambda n: table[n][0])
    if args.kernel_table and rank == 0:
        tot = sum(v[0] for v in table.values())
        with open(args.kernel_table, "w") as f:
            f.write(f"# per-kernel CUDA-event time of ONE step ({B} lanes x 1 frame, {spec.cam.cols}x{spec.cam.rows}); total {tot:.3f} ms\n")
            f.write("kernel,launches,total_ms,share\n")
            for n, v in sorted(table.items(), key=lambda kv: -kv[1][0]):
                f.write(f"{n},{v[2]},{v[0]:.4f},{v[0] / tot:.4f}\n")
    pipe.ctx.profile_reset(); pipe.ctx.profile_set(top)
    kl_sum, match_sum = 0.0, 0.0
    sampler = ClockSampler(local); sampler.start()
    launches0 = pipe.ctx.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for k in range(first + W, first + W + K):
        out = pipe.push(device_ptr=ptr(k), **pargs(k))
        kl_sum += float(out["keylines"].sum()); match_sum += float(out["matches"].sum())
    pipe.ctx.sync()                          # the front end enqueued ahead belongs to the window too: K detects + K tracker passes between the events
    ev1.record(stream)
    barrier()
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    clocks = sampler.result()
    gpu_launches = pipe.ctx.launch_count() - launches0
    prof = pipe.ctx.profile_read()
    pipe.ctx.profile_set(-1)
    frames_total = sum_over_ranks(B * K)
    value = frames_total / (ms * 1e-3)
    K_mean = kl_sum / (B * K)
    probes = pipe.ctx.probe_count(0) / max(1.0, float(out["keylines"][0]))
    top_ms, top_n, top_launches = prof[top]
    # planes per launch of the box-pass kernels: pass 0 = 1 (rowscan) / 1+1 (colbox x2 launches), passes 1,2 = 2
    planes = {"k_rowscan": 5.0 / 3.0, "k_rowscan4": 5.0 / 3.0, "k_colbox": 6.0 / 4.0, "k_boxpass": 11.0 / 6.0}.get(top, 1.0)
    words = B * N if top == "k_fill" else 0
    abytes = roofline.kernel_bytes(top, N, K_mean, B, planes=planes, P=probes, words=words)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    else:
        peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
    achieved = abytes / (top_ms / max(1, top_n) * 1e-3) / 1e9
    # DRAM traffic of the same kernel from the committed `ncu --set full` capture of this command (64 lanes, 752x480);
    # only quoted when the run has the capture's shape
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "r01_ncu_traffic.json")
    if os.path.exists(tpath) and B == LANES_PER_GPU and args.config == "C2":
        tj = json.load(open(tpath))
        if top in tj.get("bytes_per_launch", {}):
            traffic, traffic_src = float(tj["bytes_per_launch"][top]), "profiles/r01_ncu_traffic.json (" + tj.get("source", "") + ")"
    frame_gbs = roofline.frame_bytes(N, K_mean, probes) * value / max(1, world) / 1e9
    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"batched replay: {B} independent EuRoC-shaped {spec.cam.cols}x{spec.cam.rows} sequences per GPU "
                                   f"(C2 generator, seeds 1000+64*rank+i), step = 1 frame of every sequence through detect + tracker",
                       "lanes_per_gpu": B, "keylines_per_frame": K_mean, "matches_per_frame": match_sum / (B * K),
                       "schedule": (f"front end on its own stream up to {lag} frame(s) ahead of the tracker of the pair (n-1, n) on the tracker stream, as the "
                                    "reference's two threads (rebvio.cpp:56-90 / 92-293); every step = one detect + one tracker pass of all lanes"
                                    + ("; threshold servo on the device" if lag == 2 else "")) if lag
                                   else "in order on one stream",
                       "reduce_mode": args.reduce, "fusion": "imu filters on device (rbv_track_imu)" if args.imu else "visual-only (rbv_track)",
                       "parallelism": f"replicas x{world} (no collective)",
                       "l2": f"per-step working set {B} lanes x ~14 MB >> 126 MB L2 (inputs larger than L2, no flush needed)"},
            "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": abytes,
                         "avg_launch_ms": top_ms / max(1, top_n), "launches_timed": top_n,
                         "kernel_share_of_step": (top_ms / max(1, top_n)) * (top_launches / K) / (ms / K),
                         # with the two-stream schedule the events around a front-end kernel also see the tracker kernels that share
                         # the SMs with it; `alone` = the same kernel bracketed the same way in the in-order warm-up step (one stream,
                         # every kernel serialised) - the figure that compares with the ncu launch list
                         "concurrent": bool(lag),
                         "alone": {"avg_launch_ms": table[top][0] / max(1, table[top][1]),
                                   "achieved": abytes / (table[top][0] / max(1, table[top][1]) * 1e-3) / 1e9,
                                   "frac": abytes / (table[top][0] / max(1, table[top][1]) * 1e-3) / 1e9 / peak,
                                   "share_of_in_order_step": table[top][0] / sum(v[0] for v in table.values()),
                                   "in_order_step_kernel_sum_ms": sum(v[0] for v in table.values())},
                         "frame_level": {"achieved": frame_gbs, "frac": frame_gbs / peak,
                                         "bytes_per_frame": roofline.frame_bytes(N, K_mean, probes)}},
            "clocks": clocks, "gpu_launches": int(gpu_launches)}
    pipe.close()

    # ---- leg 2: end to end through the host-buffer API ---------------------------------------------------------
    # uint8 camera frames in pinned host memory -> rbv_detect_u8 (upload + x3 conversion on the device, rebvio.cpp:43)
    # -> tracker; the upload of frame k+1 is started after the front end of frame k (rbv_prefetch_frames) so that it
    # overlaps the tracker, like the reference's acquisition thread overlaps its estimation thread.
    if not args.no_e2e:
        h_frames = torch.empty((nf + 1, B, spec.cam.rows, spec.cam.cols), dtype=torch.uint8, pin_memory=True)
        h_frames[:nf].copy_(frames_u8)
        h_frames[nf].copy_(frames_u8[nf - 1])
        torch.cuda.synchronize()
        hnp = h_frames.numpy()
        pipe = BatchedPipeline(cam, det, mapc, core, B, device=local, reduce_mode=mode, cuda_stream=stream.cuda_stream, imu_cfg=imu_cfg,
                               overlap=lag)
        if args.undistort:   # rebvio/include/rebvio/camera.hpp:26-40
            fm = 0.5 * (458.654 + 457.296)
            pipe.ctx.set_undistort([fm, fm, 367.215, 248.375], [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0])
        for k in range(first):
            pipe.push(frames=hnp[k], next_frames=hnp[k + 1])
        for k in range(first, first + W):
            pipe.push(frames=hnp[k], next_frames=hnp[k + 1], **pargs(k))
        barrier()
        t0 = time.perf_counter()
        for k in range(first + W, first + W + K):
            pipe.push(frames=hnp[k], next_frames=hnp[k + 1], **pargs(k))
        pipe.ctx.sync()
        t1 = time.perf_counter()
        barrier()
        e2e_s = max_over_ranks(t1 - t0)
        # bytes per step: frames (1 byte per pixel) + thresholds + per-lane inputs of g1/g2 (64 floats each, twice);
        # outputs: counts, auto thresholds, LM state + 64 floats of lane outputs, match counters
        lm_bytes = 4 * (3 + 1 + 9 + 3 + 2 + 3 + 3 + 1 + 9 + 16 * 16)
        h2d = B * (N + 4 + 2 * 64 * 4)
        d2h = B * (8 + lm_bytes + 64 * 4 + 8 * 4)
        line["e2e"] = {"value": frames_total / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                       "ms_per_step": 1e3 * e2e_s / K,
                       "input": "uint8 frames in pinned host memory through rbv_detect_u8; next frame's upload overlapped (rbv_prefetch_frames)"
                                + ("; Camera::undistort on the device (rbv_set_undistort)" if args.undistort else "")}
        pipe.close()
        del h_frames, hnp

    # ---- leg 3: single-stream latency (configs[1]) on rank 0 -----------------------------------------------------
    if not args.no_single and rank == 0:
        nfs = min(nf, 1 + W + K)
        p1 = BatchedPipeline(cam, det, mapc, core, 1, device=local, reduce_mode=mode, cuda_stream=stream.cuda_stream, imu_cfg=imu_cfg)
        lat = []
        p1.push(device_ptr=d_frames[0, 0:1].data_ptr())
        for k in range(1, nfs):
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            p1.push(device_ptr=d_frames[k, 0:1].data_ptr(), **targs(k, slice(0, 1)))
            b.record(stream)
            torch.cuda.synchronize()
            if k > W:
                lat.append(a.elapsed_time(b))
        p50 = float(np.median(lat))
        line["single_stream"] = {"workload": "configs[1]: one 752x480 stream, lanes = 1", "latency_ms_p50": p50,
                                 "latency_ms_p90": float(np.percentile(lat, 90)), "frames_per_s": 1e3 / p50, "frames": len(lat)}
        p1.close()

    # ---- leg 4: the reference's CPU path on this box's host cores (rank 0, N = 1 only) -----------------------------
    if not args.no_cpu and rank == 0 and world == 1:
        ns = min(nf, 24)
        fps, kind = cpu_reference_frames_per_s(spec, frames_u8[:ns, 0].cpu().numpy(), R_rel[:ns, 0], ns)
        line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": 1, "kind": kind,
                                "sample": f"{ns} frames of lane 0's sequence, single thread, "
                                          + ("unmodified reference sources (-O3 -march=x86-64-v3)" if kind == "reference" else "plain-C oracle port")}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


_OUT = None


def emit(line):
    """the ONE JSON line, on the process's real stdout"""
    (_OUT or sys.stdout).write(json.dumps(line) + "\n")
    (_OUT or sys.stdout).flush()


if __name__ == "__main__":
    # libraries print to stdout as well (NCCL's "NCCL version ..." banner at the first collective): everything but the JSON line goes to stderr
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    a = parse()
    sys.exit(run_reference(a) if a.impl == "reference" else run_b200(a))
