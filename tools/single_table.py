"""per-kernel CUDA-event table of ONE single-stream frame (lanes = 1, 752x480) + wall/latency split"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from rebvio_b200 import capi, synth
from rebvio_b200.pipeline import BatchedPipeline
import bench
spec = synth.config("C2")
cam, det, mapc, core = bench.cfg_structs(spec)
dev = torch.device("cuda", 0)
frames_u8, R_rel = synth.batched_sequences(spec, [1000], 12, device=dev, n_textures=1)
d = frames_u8.to(torch.float32) * 3.0
stream = torch.cuda.Stream(device=dev)
p = BatchedPipeline(cam, det, mapc, core, 1, device=0, cuda_stream=stream.cuda_stream)
for kv in os.environ.get('RBV_TUNE', '').split(','):
    if kv:
        p.ctx.set_tuning(*[int(x) for x in kv.split('=')])
for k in range(6):
    p.push(device_ptr=d[k, 0:1].data_ptr(), R_rel=R_rel[k, 0:1] if k else None)
p.ctx.profile_reset(); p.ctx.profile_set(-2)
p.push(device_ptr=d[6, 0:1].data_ptr(), R_rel=R_rel[6, 0:1])
t = p.ctx.profile_read(); p.ctx.profile_set(-1)
tot = sum(v[0] for v in t.values())
print(f"# single stream, one frame: kernel sum {tot:.3f} ms, {sum(v[2] for v in t.values())} launches")
for n, v in sorted(t.items(), key=lambda kv: -kv[1][0]):
    print(f"{n},{v[2]},{v[0]*1e3:.1f} us")
lat = []
for k in range(7, 12):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    p.push(device_ptr=d[k, 0:1].data_ptr(), R_rel=R_rel[k, 0:1])
    torch.cuda.synchronize(); lat.append((time.perf_counter() - t0) * 1e3)
print("wall ms per frame:", ["%.3f" % x for x in lat])
