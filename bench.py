#!/usr/bin/env python
"""bench.py — frames/s of the REBVIO hot path (edge front end + edge tracker) on B200, against the HBM roofline and
next to the reference's own CPU implementation.

Workload (BASELINE.json `metric`: "frames/sec @752x480 (1 GPU p50 latency; batched streams at 1/2/4/8 GPU)"):
  * `value`  — batched replay (configs[3]): 64 independent EuRoC-shaped 752x480 sequences per GPU (the C2 generator with
    seeds 1000 + 64 g + i), one "step" = one frame of every sequence through detect + tracker (rbv_detect_device,
    rbv_track_g1, host fusion, rbv_track_g2), in the library's shipped launch path (CUDA-graph replay) and its default,
    bit-exact reduce mode (RBV_REDUCE_ORDERED). Frames resident in HBM before the timed region. A window of --steps steps is
    timed with CUDA events R times back to back (config.windows); `value` is the MEDIAN window. Whole-job frames/s.
  * `fast_mode` — the same windows with RBV_REDUCE_FAST (FP64 tree sums: not bit-exact, see config.fast_mode_parity).
  * `e2e`    — the same through the public host-buffer API (BatchedPipeline.push(host frames) -> rbv_detect_u8): every step
    copies that step's frames from pinned host memory to the device and reads counts / velocities / 6x6 systems back.
  * `roofline` — the dominant kernel by CUDA-event time; its events are taken in windows of their own (never inside the timed
    one: events cannot bracket a kernel inside a replayed graph).
  * `roofline_streaming_kernel` — the same object for the dominant kernel that only streams (in ORDERED mode the dominant launch carries the
    lanes' serial float32 add chains, so its `roofline` also states that latency floor).
  * `config.schedule_check` — parity self-check of the timed configuration: --check-schedule N frames of all lanes through the in-order
    schedule and through the timed two-stream schedule, every tracker output of every pair and the newest maps compared byte for byte.
  * `single_stream`, `single_stream_imu` (N = 1) — configs[1]: one 752x480 stream, lanes = 1, p50 per-frame latency.
  * `sweep` (N = 1) — configs[2] and configs[4]: the 1280x720 dense scene and 320x240 ... 3840x2160, 3 timed steps each.
Multi-GPU: replicas only (one frame does not shard); ranks share nothing on the data path, the barrier and the
max-over-ranks of the timing go over gloo (rebvio_b200/replicas.py); `scaling` = "weak".
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
WINDOWS = 5          # timed windows of --steps steps each; the median is reported
SWEEP = [("C3", 32), ("C5:320x240", 64), ("C5:640x480", 64), ("C5:1280x720", 32), ("C5:1920x1080", 16), ("C5:2560x1440", 8),
         ("C5:3840x2160", 4)]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--lanes", type=int, default=LANES_PER_GPU)
    ap.add_argument("--config", default="C2", help="sequence shape of every lane (C2 = 752x480 EuRoC-shaped)")
    ap.add_argument("--reduce", default="ordered", choices=["fast", "ordered"], help="reduce mode of `value` and `e2e` (ordered = bit-exact default)")
    ap.add_argument("--windows", type=int, default=WINDOWS, help="timed windows of --steps steps, median reported")
    ap.add_argument("--no-single", action="store_true", help="skip the single-stream latency legs")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-fast", action="store_true", help="skip the RBV_REDUCE_FAST leg")
    ap.add_argument("--no-sweep", action="store_true", help="skip the configs[2] / configs[4] size sweep")
    ap.add_argument("--check-schedule", type=int, default=24, metavar="N", help="parity self-check of the benchmarked configuration: N frames of all lanes through the in-order schedule and through the timed two-stream schedule, every tracker output and the last map compared bit for bit (config.schedule_check; 0 = skip)")
    ap.add_argument("--undistort", action="store_true", help="e2e leg: raw camera frames, Camera::undistort (EuRoC cam0 model) on the device")
    ap.add_argument("--imu", action="store_true", help="run the reference's IMU filters on the device between the tracker halves "
                    "(rbv_track_imu: gyro-bias correction + scale-attitude-bias filter, SURVEY 8(f2)) instead of the visual-only fusion")
    ap.add_argument("--no-overlap", action="store_true", help="in-order schedule on one stream (detect, then the tracker of the same frame) "
                    "instead of the reference's two-thread schedule on two streams (detect of frame n+1 next to the tracker of n-1, n)")
    ap.add_argument("--overlap", type=int, default=2, choices=[1, 2], help="1 = host threshold servo, front end of frame n+1 next to the tracker "
                    "of (n-1, n); 2 = servo on the device, the front end runs one more frame ahead (BatchedPipeline overlap modes)")
    ap.add_argument("--tune", default="", help="implementation choices, e.g. '2=4,0=1' (rbv_set_tuning key=value pairs)")
    ap.add_argument("--kernel-table", default="", help="write the per-kernel CUDA-event table (one untimed in-order step) here")
    return ap.parse_args()


def trace(msg):
    """progress on stderr with BENCH_TRACE=1 (a hung GPU call otherwise leaves nothing behind)"""
    if os.environ.get("BENCH_TRACE"):
        sys.stderr.write(f"[bench {time.strftime('%H:%M:%S')}] {msg}\n")
        sys.stderr.flush()


def cfg_structs(spec):
    from rebvio_b200 import capi
    cam, det, mapc, core = capi.defaults()
    cam.fm, cam.cx, cam.cy, cam.rows, cam.cols = spec.cam.fm, spec.cam.cx, spec.cam.cy, spec.cam.rows, spec.cam.cols
    det.keylines_ref, det.keylines_max = spec.keylines_ref, spec.keylines_max
    return cam, det, mapc, core


def hbm_peak():
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        return float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """samples SM clock and throttle reasons through NVML (every 20 ms) while the timed region runs"""

    def __init__(self, index, cpu=None):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz, self.cpu = index, False, [], set(), None, cpu
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
        if self.cpu is not None:
            try:
                os.sched_setaffinity(threading.get_native_id(), {self.cpu})
            except Exception:
                pass
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
            time.sleep(0.02)

    def result(self):
        self.stop_flag = True
        self.join(timeout=1.0)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def pin_launch_thread(local, world):
    """Each rank's launch thread on a core of its own GPU's NUMA node (ranks whose GPUs share a node split that node's cores);
    the rest of the process (CUDA's helper threads, the clock sampler) on the remaining cores of the rank's slice.
    -> (launch core, sampler core, slice) or (None, None, None)"""
    try:
        import pynvml
        pynvml.nvmlInit()
        ncpu = os.cpu_count() or 1
        words = (ncpu + 63) // 64
        avail = os.sched_getaffinity(0)

        def node_cpus(i):
            mask = pynvml.nvmlDeviceGetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(i), words)
            cpus = sorted(c for c in range(ncpu) if (mask[c // 64] >> (c % 64)) & 1 and c in avail)
            return cpus or sorted(avail)

        mine = node_cpus(local)
        sharers = [r for r in range(world) if node_cpus(r) == mine]
        per = max(1, len(mine) // max(1, len(sharers)))
        pos = sharers.index(local) if local in sharers else 0
        sl = mine[pos * per:(pos + 1) * per] or mine
        os.sched_setaffinity(0, set(sl))                                   # process default (threads created from here on)
        os.sched_setaffinity(threading.get_native_id(), {sl[0]})          # the launch thread itself
        return sl[0], sl[-1], sl
    except Exception:
        return None, None, None


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
class PingPong:
    """virtual time over nf staged frames: 0, 1, .., nf-1, nf-2, .., 1, 0, 1, .. — a sequence played backwards is still a smooth
    camera motion, so any number of steps can be replayed from a bounded set of frames without a jump in the tracker's state.
    R(t) = rotation of the camera between the frames of steps t-1 and t (forwards R_rel[b], backwards R_rel[a]^T)."""

    def __init__(self, nf, R_rel):
        self.nf, self.period = nf, 2 * (nf - 1)
        self.R = np.asarray(R_rel, np.float32)
        self.Rt = np.ascontiguousarray(np.swapaxes(self.R, -1, -2))

    def frame(self, t):
        m = t % self.period
        return m if m < self.nf else self.period - m

    def rot(self, t, lanes=slice(None)):
        a, b = self.frame(t - 1), self.frame(t)
        return self.R[b, lanes] if b == a + 1 else self.Rt[a, lanes]


def run_b200(args):
    import torch
    import torch.distributed as dist
    from rebvio_b200 import capi, synth, roofline, replicas
    from rebvio_b200.pipeline import BatchedPipeline

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    torch.zeros(1, device=torch.device("cuda", local))   # CUDA context (and its helper threads) first: threads inherit their creator's mask
    launch_core, sampler_core, core_slice = pin_launch_thread(local, world)
    if world > 1:
        dist.init_process_group("gloo")          # harness only (barrier, max / sum of two scalars): the data path has no collective
    dev = torch.device("cuda", local)
    spec = synth.config(args.config)
    cam, det, mapc, core = cfg_structs(spec)
    B, K, W, RW = args.lanes, args.steps, args.warmup, max(1, args.windows)
    N = spec.cam.rows * spec.cam.cols
    lag = 0 if args.no_overlap else args.overlap   # overlapped schedule: push(t) enqueues detect(t) next to the tracker of the pair (t-lag-1, t-lag)
    first = 1 + lag                          # first push that has a pair to track
    nf = first + W + min(K, 60) + 1          # staged frames; longer runs replay them forwards and backwards (PingPong)
    seeds = replicas.sequence_seeds(rank, B)
    frames_u8, R_rel = synth.batched_sequences(spec, seeds, nf, device=dev, n_textures=2)
    d_frames = frames_u8.to(torch.float32) * 3.0                     # rebvio.cpp:43 conversion, resident in HBM
    torch.cuda.synchronize()
    pp = PingPong(nf, R_rel)
    mode = capi.REDUCE_FAST if args.reduce == "fast" else capi.REDUCE_ORDERED
    stream = torch.cuda.Stream(device=dev, priority=int(os.environ.get("BENCH_STREAM_PRIORITY", "-1")))   # tracker stream above the library's front-end stream (see rbv_create)
    peak, peak_src = hbm_peak()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    max_over_ranks = lambda x: replicas.reduce_scalar(x, "max")
    sum_over_ranks = lambda x: replicas.reduce_scalar(x, "sum")
    imu_cfg = True if args.imu else None
    tune = [tuple(int(x) for x in kv.split("=")) for kv in args.tune.split(",") if kv]

    def targs(t, lanes=slice(None), imu=args.imu):
        """tracker arguments of step t: the inter-frame rotation, or (--imu) an integrated-IMU record per lane built from it
        (R = R_rel, gravity-only compensated acceleration, 50 ms interval)"""
        R = np.ascontiguousarray(pp.rot(t, lanes), np.float32)
        if not imu:
            return dict(R_rel=R)
        a = np.zeros(R.shape[0], capi.IMU_IN)
        a["R"] = R.reshape(-1, 9); a["cacc"] = (0.0, -9.81, 0.0); a["dt_s"] = 1.0 / spec.fps; a["frame_dt"] = 1.0 / spec.fps
        return dict(imu_in=a)

    ptr = lambda t: d_frames[pp.frame(t)].data_ptr()
    pargs = lambda t: targs(t - lag) if t >= first else {}     # arguments of the pair push(t) tracks

    def device_leg(reduce_mode, windows, profile):
        """device-resident frames: warm-up, `windows` timed windows of K steps (graph replay, nothing but the two window events on the
        stream), then - profile only - event windows of their own for the dominant kernel"""
        pipe = BatchedPipeline(cam, det, mapc, core, B, device=local, reduce_mode=reduce_mode, cuda_stream=stream.cuda_stream, imu_cfg=imu_cfg,
                               overlap=lag)
        for key, val in tune:
            pipe.ctx.set_tuning(key, val)
        t = 0
        for t in range(first):
            pipe.push(device_ptr=ptr(t))
        t = first
        table = None
        trace(f"device leg reduce={reduce_mode}: {first} detect-only pushes done")
        if profile:   # one step with every kernel bracketed by CUDA events, in order on one stream: finds the dominant kernel (`alone`)
            pipe.ctx.sync()
            pipe.ctx.profile_reset(); pipe.ctx.profile_set(-2)
            pipe.push(device_ptr=ptr(t), **pargs(t)); t += 1
            pipe.ctx.sync()
            table = pipe.ctx.profile_read()
            pipe.ctx.profile_set(-1)
        trace("profiled step done")
        for _ in range(W):
            out = pipe.push(device_ptr=ptr(t), **pargs(t)); t += 1
        pipe.ctx.sync()
        trace("warm-up done")
        kl_sum = match_sum = 0.0
        kl_max = 0.0
        win_ms, launches = [], 0
        sampler = ClockSampler(local, sampler_core); sampler.start()
        for w in range(windows):
            l0 = pipe.ctx.launch_count()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            ev0.record(stream)
            for _ in range(K):
                out = pipe.push(device_ptr=ptr(t), **pargs(t)); t += 1
                kl_sum += float(out["keylines"].sum()); match_sum += float(out["matches"].sum())
                kl_max = max(kl_max, float(out["keylines"].max()))
            pipe.ctx.sync()                      # the front end enqueued ahead belongs to the window too: K detects + K tracker passes between the events
            ev1.record(stream)
            barrier()
            win_ms.append(max_over_ranks(ev0.elapsed_time(ev1)))
            launches = pipe.ctx.launch_count() - l0
            trace(f"window {w}: {win_ms[-1]:.2f} ms")
        clocks = sampler.result()
        res = {"win_ms": win_ms, "ms": float(np.median(win_ms)), "clocks": clocks, "launches": int(launches),
               "K_mean": kl_sum / (B * K * windows), "K_max": kl_max, "matches": match_sum / (B * K * windows),
               "probes": pipe.ctx.probe_count(0) / max(1.0, float(out["keylines"][0])), "table": table}
        if profile:   # the dominant kernel's events in the shipped two-stream schedule, outside the timed windows; then the same for the
            # dominant kernel that only streams (an ORDERED evaluation launch carries its lanes' serial float32 add chains: a latency floor)
            top = max(table, key=lambda n: table[n][0])
            stream_only = [n for n in table if n not in CHAIN_K]
            top_hbm = max(stream_only, key=lambda n: table[n][0]) if stream_only else top
            nprof = min(K, 10)
            res["top"], res["top_hbm"], res["prof_steps"], res["prof"] = top, top_hbm, nprof, {}
            for name in dict.fromkeys((top, top_hbm)):
                pipe.ctx.profile_reset(); pipe.ctx.profile_set(name)
                for _ in range(nprof):
                    pipe.push(device_ptr=ptr(t), **pargs(t)); t += 1
                pipe.ctx.sync()
                res["prof"][name] = pipe.ctx.profile_read()[name]
            pipe.ctx.profile_set(-1)
        pipe.close()
        return res

    def schedule_check(n):
        """the timed configuration against the in-order schedule on the same resident frames: outputs of every tracked pair and the
        newest map, byte for byte (in order vs the oracle / the compiled reference is what tests/ cover)"""
        n = min(n, nf)
        keys = ("Vg", "X", "Wx", "matches", "kf_matches", "ok") + (("imu",) if args.imu else ("V",))
        runs = []
        for lg in (0, lag):
            pipe = BatchedPipeline(cam, det, mapc, core, B, device=local, reduce_mode=mode, cuda_stream=stream.cuda_stream, imu_cfg=imu_cfg, overlap=lg)
            for key, val in tune:
                pipe.ctx.set_tuning(key, val)
            outs = []
            for t in range(n):
                o = pipe.push(device_ptr=d_frames[t].data_ptr(), **(targs(t - lg) if t >= 1 + lg else {}))
                if t >= 1 + lg:
                    outs.append(o)
            if lg:
                outs.extend(pipe.flush([targs(n - lg + i) for i in range(lg)]))
            pipe.ctx.sync()
            last = [pipe.newest_map().keylines(b)[0].tobytes() for b in range(0, B, max(1, B // 8))]
            runs.append(([{k: np.asarray(o[k]).tobytes() for k in keys} for o in outs], last))
            pipe.close()
        (a, la), (b, lb) = runs
        bad = [(i, k) for i, (x, y) in enumerate(zip(a, b)) for k in keys if x[k] != y[k]]
        return {"frames": n, "lanes": B, "pairs_compared": min(len(a), len(b)), "pairs_expected": n - 1, "mismatches": len(bad), "first_mismatch": bad[:1],
                "last_map_identical": la == lb and len(a) == len(b) == n - 1,
                "what": f"two-stream schedule (overlap {lag}) against in order on one stream, same frames: {', '.join(keys)} of every pair and the newest map of every 8th lane, byte for byte"}

    # ---- leg 1: device-resident inputs, default (bit-exact) reduce mode ---------------------------------------------
    main = device_leg(mode, RW, True)
    table, top = main["table"], main["top"]
    if args.kernel_table and rank == 0:
        tot = sum(v[0] for v in table.values())
        with open(args.kernel_table, "w") as f:
            f.write(f"# per-kernel CUDA-event time of ONE in-order step ({B} lanes x 1 frame, {spec.cam.cols}x{spec.cam.rows}, reduce {args.reduce}); total {tot:.3f} ms\n")
            f.write("kernel,launches,total_ms,share\n")
            for n, v in sorted(table.items(), key=lambda kv: -kv[1][0]):
                f.write(f"{n},{v[2]},{v[0]:.4f},{v[0] / tot:.4f}\n")
    ms = main["ms"]
    frames_total = sum_over_ranks(B * K)
    value = frames_total / (ms * 1e-3)
    K_mean, probes = main["K_mean"], main["probes"]
    def roof(name):
        """roofline object of one kernel: algorithmic bytes of a launch / its CUDA-event time, in the shipped schedule and alone"""
        k_ms, k_n, k_launches = main["prof"][name]
        # planes per launch of the box-pass kernels: pass 0 = 1 (rowscan) / 1+1 (colbox x2 launches), passes 1,2 = 2
        planes = {"k_rowscan": 5.0 / 3.0, "k_rowscan4": 5.0 / 3.0, "k_colbox": 6.0 / 4.0, "k_boxpass": 11.0 / 6.0, "k_boxpass2": 11.0 / 6.0}.get(name, 1.0)
        words = B * N if name == "k_fill" else 0
        abytes = roofline.kernel_bytes(name, N, K_mean, B, planes=planes, P=probes, words=words)
        achieved = abytes / (k_ms / max(1, k_n) * 1e-3) / 1e9
        # DRAM traffic of the same kernel from the committed `ncu --set full` capture of this command (64 lanes, 752x480);
        # only quoted when the run has the capture's shape
        traffic, traffic_src = None, None
        for tname in ("r02_ncu_traffic.json", "r01_ncu_traffic.json"):
            tpath = os.path.join(ROOT, "profiles", tname)
            if traffic is None and os.path.exists(tpath) and B == LANES_PER_GPU and args.config == "C2":
                tj = json.load(open(tpath))
                if name in tj.get("bytes_per_launch", {}):
                    traffic, traffic_src = float(tj["bytes_per_launch"][name]), f"profiles/{tname} (" + tj.get("source", "") + ")"
        alone_ms = table[name][0] / max(1, table[name][1])
        r = {"bound": "hbm", "kernel": name, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
             "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": abytes,
             "avg_launch_ms": k_ms / max(1, k_n), "launches_timed": k_n,
             "measured_in": f"{main['prof_steps']} steps of the same schedule right behind the timed windows (plain stream launches: "
                            "events cannot bracket a kernel inside a replayed graph)",
             "kernel_share_of_step": (k_ms / max(1, k_n)) * (k_launches / main["prof_steps"]) / (ms / K),
             # with the two-stream schedule the events around a kernel also see the other stream's kernels that share the SMs with it;
             # `alone` = the same kernel bracketed the same way in the in-order warm-up step (one stream, every kernel serialised) -
             # the figure that compares with the ncu launch list
             "concurrent": bool(lag),
             "alone": {"avg_launch_ms": alone_ms, "achieved": abytes / (alone_ms * 1e-3) / 1e9, "frac": abytes / (alone_ms * 1e-3) / 1e9 / peak,
                       "share_of_in_order_step": table[name][0] / sum(v[0] for v in table.values()),
                       "in_order_step_kernel_sum_ms": sum(v[0] for v in table.values())}}
        if name in CHAIN_K:
            # the launch carries one serial float32 add chain per lane (the reference's summation order): its floor is the longest lane's
            # keylines x one dependent FADD (4 cycles) at the SM clock, whatever the bandwidth
            kmax_lane = float(main["K_max"])
            floor_ms = kmax_lane * 4.0 / (main["clocks"]["sm_mhz"] * 1e3) if main["clocks"].get("sm_mhz") else None
            r["latency_floor"] = {"what": "one dependent float32 add per keyline and accumulator, in keyline order, one warp per lane (bit-exact sums)",
                                  "longest_lane_keylines": kmax_lane, "cycles_per_keyline": 4, "floor_ms": floor_ms,
                                  "floor_over_alone": (floor_ms / alone_ms) if floor_ms else None}
        return r

    top_hbm = main["top_hbm"]
    frame_gbs = roofline.frame_bytes(N, K_mean, probes) * value / max(1, world) / 1e9
    fast_parity = None
    fpath = os.path.join(ROOT, "profiles", "r02_fast_mode_report.json")
    if os.path.exists(fpath):
        fast_parity = json.load(open(fpath)).get("summary")
    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"batched replay: {B} independent EuRoC-shaped {spec.cam.cols}x{spec.cam.rows} sequences per GPU "
                                   f"(C2 generator, seeds 1000+64*rank+i), step = 1 frame of every sequence through detect + tracker",
                       "lanes_per_gpu": B, "keylines_per_frame": K_mean, "matches_per_frame": main["matches"],
                       "windows": RW, "window_ms": [round(x, 4) for x in main["win_ms"]],
                       "timing": f"median of {RW} back-to-back windows of {K} steps, CUDA events on the tracker stream, max over ranks per window; "
                                 "launch path = CUDA-graph replay (no per-kernel events inside the windows)",
                       "frames": f"{nf} frames per sequence staged in HBM, replayed forwards then backwards (continuous motion)",
                       "schedule": (f"front end on its own stream up to {lag} frame(s) ahead of the tracker of the pair (n-1, n) on the tracker stream, as the "
                                    "reference's two threads (rebvio.cpp:56-90 / 92-293); every step = one detect + one tracker pass of all lanes"
                                    + ("; threshold servo on the device" if lag == 2 else "")) if lag
                                   else "in order on one stream",
                       "reduce_mode": args.reduce + (" (float32 sums in keyline order: bit-exact against the reference)" if args.reduce == "ordered" else ""),
                       "fusion": "imu filters on device (rbv_track_imu)" if args.imu else "visual-only (rbv_track)",
                       "parallelism": f"replicas x{world} (no collective)",
                       "host_threads": {"launch_core": launch_core, "sampler_core": sampler_core, "slice": core_slice},
                       "l2": f"per-step working set {B} lanes x ~14 MB >> 126 MB L2 (inputs larger than L2, no flush needed)"},
            "roofline": dict(roof(top), frame_level={"achieved": frame_gbs, "frac": frame_gbs / peak,
                                                     "bytes_per_frame": roofline.frame_bytes(N, K_mean, probes)}),
            "clocks": main["clocks"], "gpu_launches": main["launches"]}
    if args.check_schedule and lag:
        line["config"]["schedule_check"] = schedule_check(args.check_schedule)
    if top_hbm != top:   # the dominant kernel that only streams, measured the same way
        line["roofline_streaming_kernel"] = roof(top_hbm)

    # ---- leg 1b: the other reduce mode ---------------------------------------------------------------------------------
    if not args.no_fast:
        other = capi.REDUCE_FAST if mode == capi.REDUCE_ORDERED else capi.REDUCE_ORDERED
        o = device_leg(other, min(RW, 3), False)
        key = "fast_mode" if other == capi.REDUCE_FAST else "ordered"
        line[key] = {"value": frames_total / (o["ms"] * 1e-3), "unit": "frames/s", "ms_per_step": o["ms"] / K, "windows": len(o["win_ms"]),
                     "reduce_mode": "fast (FP64 tree sums, deterministic, NOT bit-exact)" if other == capi.REDUCE_FAST else "ordered (bit-exact)"}
        if fast_parity is not None:
            line["config"]["fast_mode_parity"] = fast_parity

    # ---- leg 2: end to end through the host-buffer API ---------------------------------------------------------
    # uint8 camera frames in pinned host memory -> rbv_detect_u8 (upload + x3 conversion on the device, rebvio.cpp:43)
    # -> tracker; the upload of frame k+1 is started after the front end of frame k (rbv_prefetch_frames) so that it
    # overlaps the tracker, like the reference's acquisition thread overlaps its estimation thread.
    trace("device legs done")
    if not args.no_e2e:
        h_frames = torch.empty((nf, B, spec.cam.rows, spec.cam.cols), dtype=torch.uint8, pin_memory=True)
        h_frames.copy_(frames_u8)
        torch.cuda.synchronize()
        hnp = h_frames.numpy()
        hf = lambda t: hnp[pp.frame(t)]
        pipe = BatchedPipeline(cam, det, mapc, core, B, device=local, reduce_mode=mode, cuda_stream=stream.cuda_stream, imu_cfg=imu_cfg,
                               overlap=lag)
        for key, val in tune:
            pipe.ctx.set_tuning(key, val)
        if args.undistort:   # rebvio/include/rebvio/camera.hpp:26-40
            fm = 0.5 * (458.654 + 457.296)
            pipe.ctx.set_undistort([fm, fm, 367.215, 248.375], [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0])
        t = 0
        for t in range(first + W):
            pipe.push(frames=hf(t), next_frames=hf(t + 1), **pargs(t))
        t = first + W
        win_s = []
        for w in range(RW):
            barrier()
            t0 = time.perf_counter()
            for _ in range(K):
                pipe.push(frames=hf(t), next_frames=hf(t + 1), **pargs(t)); t += 1
            pipe.ctx.sync()
            t1 = time.perf_counter()
            barrier()
            win_s.append(max_over_ranks(t1 - t0))
        e2e_s = float(np.median(win_s))
        # bytes per step: frames (1 byte per pixel) + thresholds + per-lane inputs of g1/g2 (64 floats each, twice);
        # outputs: counts, auto thresholds, LM state + 64 floats of lane outputs, match counters
        lm_bytes = 4 * (3 + 1 + 9 + 3 + 2 + 3 + 3 + 1 + 9 + 16 * 16)
        h2d = B * (N + 4 + 2 * 64 * 4)
        d2h = B * (8 + lm_bytes + 64 * 4 + 8 * 4)
        line["e2e"] = {"value": frames_total / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                       "ms_per_step": 1e3 * e2e_s / K, "windows": RW, "window_ms": [round(1e3 * x, 4) for x in win_s],
                       "input": "uint8 frames in pinned host memory through rbv_detect_u8; next frame's upload overlapped (rbv_prefetch_frames); "
                                "wall clock around each window, median window"
                                + ("; Camera::undistort on the device (rbv_set_undistort)" if args.undistort else "")}
        pipe.close()
        del h_frames, hnp

    # ---- leg 3: single-stream latency (configs[1]) on rank 0 -----------------------------------------------------
    if not args.no_single and rank == 0:
        nfs = min(nf, 1 + W + K)
        for key, with_imu in (("single_stream", False), ("single_stream_imu", True)):
            p1 = BatchedPipeline(cam, det, mapc, core, 1, device=local, reduce_mode=mode, cuda_stream=stream.cuda_stream,
                                 imu_cfg=True if with_imu else None)
            lat = []
            p1.push(device_ptr=d_frames[0, 0:1].data_ptr())
            for k in range(1, nfs):
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                p1.push(device_ptr=d_frames[k, 0:1].data_ptr(), **targs(k, slice(0, 1), imu=with_imu))
                b.record(stream)
                torch.cuda.synchronize()
                if k > W:
                    lat.append(a.elapsed_time(b))
            p50 = float(np.median(lat))
            line[key] = {"workload": "configs[1]: one 752x480 stream, lanes = 1, " + ("IMU filters on the device (rbv_track_imu)" if with_imu
                                                                                      else "visual-only fusion on the host (rbv_track)"),
                         "latency_ms_p50": p50, "latency_ms_p90": float(np.percentile(lat, 90)), "frames_per_s": 1e3 / p50, "frames": len(lat),
                         "reduce_mode": args.reduce}
            p1.close()

    del d_frames, frames_u8
    torch.cuda.empty_cache()

    # ---- leg 4: size sweep (configs[2], configs[4]) on rank 0, N = 1 only -------------------------------------------
    if not args.no_sweep and rank == 0 and world == 1:
        line["sweep"] = run_sweep(torch, dev, local, stream, mode, peak)

    # ---- leg 5: the reference's CPU path on this box's host cores (rank 0, N = 1 only) -----------------------------
    if not args.no_cpu and rank == 0 and world == 1:
        ns = 24
        f_u8, R1 = synth.batched_sequences(spec, seeds[:1], ns, device=dev, n_textures=2)
        fps, kind = cpu_reference_frames_per_s(spec, f_u8[:, 0].cpu().numpy(), R1[:, 0], ns)
        line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": 1, "kind": kind,
                                "sample": f"{ns} frames of lane 0's sequence, single thread, "
                                          + ("unmodified reference sources (-O3 -march=x86-64-v3)" if kind == "reference" else "plain-C oracle port")}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


SCALE_K = ("k_boxpass", "k_boxpass2", "k_rowscan", "k_rowscan4", "k_colbox", "k_rowscan_lat", "k_colbox_lat")
DETECT_K = ("k_detect_flags", "k_detect_scan", "k_detect_emit", "k_detect_emit_wide", "k_join", "k_tune")
CHAIN_K = ("k_minimize_ord2", "k_minimize_ord4", "k_tryvel_ord2", "k_tryvel_ord4", "k_chain_tryvel", "k_extrot_ord", "k_chain_extrot", "k_tryvel_reduce", "k_extrot_reduce")   # launches that carry a serial add chain
TRYVEL_K = ("k_tryvel_fast", "k_tryvel_fast2", "k_tryvel_fast4", "k_tryvel_ord2", "k_tryvel_ord4", "k_chain_tryvel", "k_tryvel_eval", "k_tryvel_reduce")


def run_sweep(torch, dev, local, stream, mode, peak):
    """configs[2] (C3: 1280x720 dense-edge scene, ~20 k keylines) and configs[4] (320x240 ... 3840x2160): scale space, keyline detect
    and one tracker iteration per size as CUDA-event time of one in-order step -> algorithmic GB/s against the HBM roofline (bytes:
    scale space 48 N, detect 12 N + 104 K, one tryVel iteration 68 K, frame 64 N + (1260 + 4 P) K), plus frames/s over 3 timed steps."""
    from rebvio_b200 import synth, roofline
    from rebvio_b200.pipeline import BatchedPipeline
    rows = []
    for name, B in SWEEP:
        spec = synth.config(name)
        cam, det, mapc, core = cfg_structs(spec)
        N = spec.cam.rows * spec.cam.cols
        nf = 8
        frames_u8, R_rel = synth.batched_sequences(spec, [1000 + i for i in range(B)], nf, device=dev, n_textures=1)
        d = frames_u8.to(torch.float32) * 3.0
        del frames_u8
        p = BatchedPipeline(cam, det, mapc, core, B, device=local, reduce_mode=mode, cuda_stream=stream.cuda_stream)
        for k in range(3):
            p.push(device_ptr=d[k].data_ptr(), R_rel=R_rel[k] if k else None)
        p.ctx.profile_reset(); p.ctx.profile_set(-2)
        out = p.push(device_ptr=d[3].data_ptr(), R_rel=R_rel[3])
        t = p.ctx.profile_read(); p.ctx.profile_set(-1)
        p.push(device_ptr=d[4].data_ptr(), R_rel=R_rel[4])
        torch.cuda.synchronize()
        sampler = ClockSampler(local); sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ks = 0.0
        for k in range(5, 8):
            out = p.push(device_ptr=d[k].data_ptr(), R_rel=R_rel[k]); ks += float(out["keylines"].mean())
        e1.record(stream); torch.cuda.synchronize()
        clocks = sampler.result()
        ms = e0.elapsed_time(e1) / 3
        Km = ks / 3
        P = p.ctx.probe_count(0) / max(1.0, float(out["keylines"][0]))
        us = lambda names: 1e3 * sum(t[n][0] for n in names if n in t)
        sc, dt_, tv = us(SCALE_K), us(DETECT_K), us(TRYVEL_K) / 6.0
        g = lambda byts, u: (byts * B / (u * 1e-6) / 1e9) if u > 0 else 0.0
        gs, gd, gt = g(48.0 * N, sc), g(12.0 * N + 104.0 * Km, dt_), g(68.0 * Km, tv)
        gf = roofline.frame_bytes(N, Km, P) * B / (ms * 1e-3) / 1e9
        rows.append({"config": f"{name} {spec.cam.cols}x{spec.cam.rows}", "lanes": B, "keylines_per_frame": Km, "frames_per_s": B / (ms * 1e-3),
                     "ms_per_step": ms, "steps": 3, "schedule": "in order on one stream",
                     "scale_space": {"us": sc, "gbs": gs, "frac": gs / peak}, "detect": {"us": dt_, "gbs": gd, "frac": gd / peak},
                     "tryvel_iteration": {"us": tv, "gbs": gt, "frac": gt / peak}, "frame_gbs": gf, "frame_frac": gf / peak,
                     "clocks": {"sm_mhz": clocks["sm_mhz"], "reasons": clocks["reasons"]}})
        p.close(); del d
        torch.cuda.empty_cache()
    return rows


_OUT = None


def emit(line):
    """the ONE JSON line, on the process's real stdout"""
    (_OUT or sys.stdout).write(json.dumps(line) + "\n")
    (_OUT or sys.stdout).flush()


if __name__ == "__main__":
    # libraries print to stdout as well: everything but the JSON line goes to stderr
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    a = parse()
    sys.exit(run_reference(a) if a.impl == "reference" else run_b200(a))
