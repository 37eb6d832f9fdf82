"""Flatten / restore tests/harness.run_stages records to an .npz (tests/golden/, written by tools/make_golden.py)."""
from __future__ import annotations

import json

import numpy as np


def _flatten(prefix, obj, out):
    if isinstance(obj, dict):
        for k, v in obj.items():
            _flatten(f"{prefix}/{k}" if prefix else str(k), v, out)
    else:
        out[prefix] = np.asarray(obj)


def save(path, frames_u8, R_rel, meta, recs, extra=None):
    out = {"frames": frames_u8, "R_rel": R_rel, "meta": np.frombuffer(json.dumps(meta).encode(), np.uint8)}
    for k, rec in enumerate(recs):
        _flatten(f"rec/{k}", rec, out)
    for k, v in (extra or {}).items():
        out[k] = v
    np.savez_compressed(path, **out)


def load(path):
    """-> frames_u8 [n,rows,cols], R_rel, meta dict, recs (list of nested dicts like run_stages returns), extra dict"""
    z = np.load(path)
    meta = json.loads(bytes(z["meta"]).decode())
    n = len(z["frames"])
    recs = [dict() for _ in range(n)]
    extra = {}
    for key in z.files:
        if key.startswith("rec/"):
            parts = key.split("/")
            d = recs[int(parts[1])]
            for p in parts[2:-1]:
                d = d.setdefault(p, {})
            v = z[key]
            d[parts[-1]] = v.item() if v.ndim == 0 else v
        elif key not in ("frames", "R_rel", "meta"):
            extra[key] = z[key]
    return z["frames"], z["R_rel"], meta, recs, extra
