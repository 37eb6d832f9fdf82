"""Wall-clock throughput of ONE 752x480 stream through the C++ Rebvio facade (rebvio_b200/cpp/rebvio_replay.cpp): lock step (drain after every
frame) against both worker loops free-running (detect of frame n+1 on the front-end stream next to the tracker of n-1, n). Run on the GPU box:
    python tools/facade_throughput.py"""
import subprocess, sys, os, time
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
from rebvio_b200 import synth, replay
spec = synth.config("C2", n_frames=120); seq = synth.generate(spec)
replay.write_sequence("/tmp/c2.rbvseq", seq.frames, seq.ts_us, seq.imu_ts_us, seq.gyro, seq.acc, spec.cam, spec.keylines_ref, spec.keylines_max)
exe="/tmp/rebvio_replay_gpu"
r=subprocess.run(["g++","-std=c++14","-O2","-ffp-contract=off","-Iinclude","-Irebvio_b200/cpp","rebvio_b200/cpp/rebvio_replay.cpp","-o",exe,"-Lrebvio_b200","-lrebvio_b200","-Wl,-rpath,"+os.path.join(os.getcwd(),"rebvio_b200"),"-lpthread"],capture_output=True,text=True); print(r.stderr)
for extra in ([], ["--no-wait"], ["--no-wait"], ["--no-wait"], ["--no-wait"]):
    r=subprocess.run([exe,"/tmp/c2.rbvseq","/tmp/odo.txt"]+extra,capture_output=True,text=True); print(extra, r.stderr.strip())
