#!/usr/bin/env python
"""Kernel sweep (BASELINE.json configs[4] and configs[2]): scale space, keyline detect and one tracker iteration from 320x240 to
3840x2160 (+ the 1280x720 dense-edge scene C3), CUDA-event time per stage and algorithmic GB/s against the HBM roofline.
usage (GPU box): python tools/sweep.py > profiles/rNN_sweep.md"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from rebvio_b200 import synth, roofline
from rebvio_b200.pipeline import BatchedPipeline
import bench

CASES = [("C5:320x240", 64), ("C5:640x480", 64), ("C2", 64), ("C5:1280x720", 32), ("C3", 32), ("C5:1920x1080", 16),
         ("C5:2560x1440", 8), ("C5:3840x2160", 4), ("C2", 1), ("C3", 1), ("C5:3840x2160", 1)]
if len(sys.argv) > 1:   # e.g. 'C3:32,C2:1'
    CASES = [(a.rsplit(':', 1)[0], int(a.rsplit(':', 1)[1])) for a in sys.argv[1].split(',')]
peaks = os.path.join(ROOT, "MEASURED_PEAKS.json")
PEAK = float(json.load(open(peaks))["hbm_gbs"]) if os.path.exists(peaks) else 6650.0
SCALE = ("k_boxpass", "k_rowscan", "k_rowscan4", "k_colbox", "k_rowscan_lat", "k_colbox_lat")
DETECT = ("k_detect_flags", "k_detect_scan", "k_detect_emit", "k_detect_emit_wide", "k_join", "k_tune")
TRYVEL = ("k_tryvel_fast", "k_tryvel_fast2", "k_tryvel_fast4", "k_tryvel_ord2", "k_tryvel_ord4", "k_chain_tryvel")
dev = torch.device("cuda", 0)
print(f"# kernel sweep on {torch.cuda.get_device_name(0)}; HBM peak {PEAK:.0f} GB/s (MEASURED_PEAKS.json); algorithmic bytes: scale space 48 N, "
      "detect 12 N + 104 K (flags, emit, join), one tryVel iteration 68 K, frame 64 N + (1260 + 4 P) K; times = CUDA events of one step, "
      "frames/s = 6 timed steps\n")
print("| config | lanes | keylines/frame | frames/s | ms/step | scale space us (GB/s, frac) | detect us (GB/s, frac) | tryVel iter us (GB/s, frac) | field us | frame GB/s (frac) |")
print("|---|---|---|---|---|---|---|---|---|---|")
for name, B in CASES:
    spec = synth.config(name)
    cam, det, mapc, core = bench.cfg_structs(spec)
    N = spec.cam.rows * spec.cam.cols
    nf = 11
    frames_u8, R_rel = synth.batched_sequences(spec, [1000 + i for i in range(B)], nf, device=dev, n_textures=1)
    d = frames_u8.to(torch.float32) * 3.0
    del frames_u8
    stream = torch.cuda.Stream(device=dev)
    p = BatchedPipeline(cam, det, mapc, core, B, device=0, cuda_stream=stream.cuda_stream)
    for k in range(4):
        p.push(device_ptr=d[k].data_ptr(), R_rel=R_rel[k] if k else None)
    p.ctx.profile_reset(); p.ctx.profile_set(-2)
    out = p.push(device_ptr=d[4].data_ptr(), R_rel=R_rel[4])
    t = p.ctx.profile_read(); p.ctx.profile_set(-1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    ks = 0.0
    for k in range(5, 11):
        out = p.push(device_ptr=d[k].data_ptr(), R_rel=R_rel[k]); ks += float(out["keylines"].mean())
    e1.record(stream); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 6
    K = ks / 6
    P = p.ctx.probe_count(0) / max(1.0, float(out["keylines"][0]))
    us = lambda names: 1e3 * sum(t[n][0] for n in names if n in t)
    sc, dt = us(SCALE), us(DETECT)
    tv = us(TRYVEL) / 6.0
    fld = us(("k_field_band", "k_field_build", "k_fill", "k_pack_targets"))
    g = lambda byts, u: (byts * B / (u * 1e-6) / 1e9) if u > 0 else 0.0
    gs, gd, gt = g(48.0 * N, sc), g(12.0 * N + 104.0 * K, dt), g(68.0 * K, tv)
    fb = roofline.frame_bytes(N, K, P)
    gf = fb * B / (ms * 1e-3) / 1e9
    print(f"| {name} {spec.cam.cols}x{spec.cam.rows} | {B} | {K:.0f} | {B / (ms * 1e-3):.0f} | {ms:.3f} | {sc:.0f} ({gs:.0f}, {gs / PEAK:.2f}) | "
          f"{dt:.0f} ({gd:.0f}, {gd / PEAK:.2f}) | {tv:.1f} ({gt:.0f}, {gt / PEAK:.2f}) | {fld:.0f} | {gf:.0f} ({gf / PEAK:.2f}) |", flush=True)
    p.close(); del d; torch.cuda.empty_cache()
