"""Per-kernel CUDA-event table of ONE steady-state step (after 6 warm-up frames) with the visual-only fusion (rbv_track) and
with the IMU filters on the device (rbv_track_imu), plus the host-clock step time of both. Usage: python tools/imu_step_table.py [lanes]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from rebvio_b200 import capi, synth
from rebvio_b200.pipeline import BatchedPipeline
import bench

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
spec = synth.config("C2")
cam, det, mapc, core = bench.cfg_structs(spec)
dev = torch.device("cuda", 0)
nf = 30
frames_u8, R_rel = synth.batched_sequences(spec, [1000 + i for i in range(B)], nf, device=dev, n_textures=2)
d = frames_u8.to(torch.float32) * 3.0
torch.cuda.synchronize()
for imu in (False, True):
    pipe = BatchedPipeline(cam, det, mapc, core, B, imu_cfg=True if imu else None)

    def targs(k):
        R = np.asarray(R_rel[k], np.float32)
        if not imu:
            return dict(R_rel=R)
        a = np.zeros(B, capi.IMU_IN)
        a["R"] = R.reshape(-1, 9); a["cacc"] = (0.0, -9.81, 0.0); a["dt_s"] = 0.05; a["frame_dt"] = 0.05
        return dict(imu_in=a)
    pipe.push(device_ptr=d[0].data_ptr())
    for k in range(1, 20):
        pipe.push(device_ptr=d[k].data_ptr(), **targs(k))
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for k in range(20, 28):
        out = pipe.push(device_ptr=d[k].data_ptr(), **targs(k))
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 8
    pipe.ctx.profile_reset(); pipe.ctx.profile_set(-2)
    out = pipe.push(device_ptr=d[28].data_ptr(), **targs(28))
    table = pipe.ctx.profile_read(); pipe.ctx.profile_set(-1)
    tot = sum(v[0] for v in table.values())
    print(f"# imu={imu} lanes={B}: host-clock step {dt * 1e3:.3f} ms; event table total {tot:.3f} ms; matches/lane {out['matches'].mean():.0f}"
          + (f"; K {out['imu']['K'][:3]}" if imu else ""))
    for n, v in sorted(table.items(), key=lambda kv: -kv[1][0])[:12]:
        print(f"{n},{v[2]},{v[0]:.4f}")
    pipe.close()
