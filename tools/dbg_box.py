import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import common, harness as H
from oracle import pyref
from rebvio_b200 import capi, host
rows, cols = 480, 752
cfg = common.cfgs(0.6 * cols, cols / 2.0, rows / 2.0, rows, cols, 100, 200)
orc = pyref.OracleLib(cfg[0], cfg[1], cfg[3], cfg[2])
conv = lambda cls, src: cls(*[getattr(src, f[0]) for f in cls._fields_])
rng = np.random.RandomState(rows + cols)
img = rng.randint(0, 256, (rows, cols)).astype(np.float32) * np.float32(3.0)
for tune in (1, 2, 1):
    ctx = host.Context(conv(capi.Camera, cfg[0]), conv(capi.DetectorCfg, cfg[1]), conv(capi.MapCfg, cfg[2]), conv(capi.CoreCfg, cfg[3]))
    ctx.set_tuning(capi.TUNE_BOX, tune)
    for s in range(2):
        for p in range(3):
            a = ctx.box_pass(s, p, img); b = orc.box_pass(s, p, img)
            bad = np.argwhere(H.bits(a) != H.bits(b))
            print("tune", tune, "s", s, "p", p, "ndiff", len(bad), "rows", np.unique(bad[:, 0])[:12] if len(bad) else "", "cols", (bad[:, 1].min(), bad[:, 1].max()) if len(bad) else "")
    ctx.close()
