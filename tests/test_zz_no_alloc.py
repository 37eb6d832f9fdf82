"""GPU (B200): rbv_box_pass_noalloc - the box pass that allocates nothing (static device scratch; no counterpart in the reference)
- through the C ABI against the existing entry point rbv_debug_box_pass on the same images:
bit-exact, the tolerance of every other box-pass test here, on two calls in a row that share the static scratch.

Collected last on purpose (file name): a failure here must not hide the other tests. Runs alone in
a fresh process as well as behind the other tests (it needs nothing they leave behind)."""
import ctypes as C

import numpy as np
import pytest

import common
import harness as H
from rebvio_b200 import capi, host

pytestmark = pytest.mark.gpu


def _noalloc(L, d_in, d_out, rows, cols, width):
    return L.rbv_box_pass_noalloc(C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()), rows, cols, width, None)


def test_box_pass_noalloc_equals_existing_entry_point_twice():
    import torch
    rows, cols = 240, 320
    cfg = common.cfgs(200.0, cols / 2.0, rows / 2.0, rows, cols, 3000, 4000)
    conv = lambda cls, src: cls(*[getattr(src, f[0]) for f in cls._fields_])
    ctx = host.Context(conv(capi.Camera, cfg[0]), conv(capi.DetectorCfg, cfg[1]), conv(capi.MapCfg, cfg[2]), conv(capi.CoreCfg, cfg[3]))
    L = ctx.L
    _, widths = ctx.setup()
    rng = np.random.default_rng(7)
    dev = torch.device("cuda:0")
    d_in = torch.empty((rows, cols), dtype=torch.float32, device=dev)
    d_out = torch.empty_like(d_in)
    torch.cuda.synchronize()
    # two calls in a row, different images and different widths: the second finds the first call's row prefix in the shared scratch
    for call, (scale, p) in enumerate([(0, 0), (1, 2)]):
        img = (rng.random((rows, cols)) * 765.0).astype(np.float32)
        want = ctx.box_pass(scale, p, img)                       # existing entry point (context scratch from cudaMalloc)
        d_in.copy_(torch.from_numpy(img))
        d_out.fill_(-1.0)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(dev)[0]
        capi.check(_noalloc(L, d_in, d_out, rows, cols, int(widths[scale * 3 + p])))
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info(dev)[0]
        got = d_out.cpu().numpy()
        assert np.array_equal(H.bits(got), H.bits(want)), ("call", call, "differs from rbv_debug_box_pass")
        if call == 1:   # (the first launch of a process may load the kernels' module: only the second call is held to "no new memory")
            assert free1 == free0, "device memory in use changed across rbv_box_pass_noalloc"
    ctx.close()


def test_box_pass_noalloc_rejects_what_does_not_fit():
    import torch
    L = capi.load()
    dev = torch.device("cuda:0")
    buf = torch.zeros(64, dtype=torch.float32, device=dev)
    # beyond the static scratch (1024 x 1024 pixels): refused before anything is launched, with a message that names the limit
    rc = _noalloc(L, buf, buf, 1025, 1024, 7)
    assert rc == capi.RBV_ERR_ARG and b"RBV_NOALLOC_MAX_PIXELS" in L.rbv_last_error()
    assert _noalloc(L, buf, buf, 8, 8, 4) == capi.RBV_ERR_ARG        # even width
    assert _noalloc(L, buf, buf, 8, 8, 17) == capi.RBV_ERR_ARG       # wider than the reference's divisor table
    assert L.rbv_box_pass_noalloc(None, C.c_void_p(buf.data_ptr()), 8, 8, 7, None) == capi.RBV_ERR_ARG
    torch.cuda.synchronize()
