"""lc_render_static — the allocation-free variant of lc_render (include/learn_comms.h; no counterpart in the reference, added
for callers that cannot call cudaMalloc).  Its per-pixel scratch lives in static
__device__ arrays (LC_RENDER_STATIC_MAX_PIXELS = 2^20 pixels) whose device addresses the library looks up with
cudaGetSymbolAddress.  Held against the existing entry point lc_render through the C ABI on two calls in a row (different
crowd states, different picture sizes: whatever the first call left in the shared scratch must not show in the second).
The two entry points launch the same kernels, so the pictures must be IDENTICAL — inside the <= 1 LSB per channel the
renderer's other tests (tests/test_render.py) allow.  This file is collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
MAX_PIXELS = 1 << 20
BACKGROUND = np.array([253, 246, 227, 255], np.uint8)   # render.cpp:365


def _sim(scene=9, wseed=12):
    import learn_comms_b200 as lcb
    from learn_comms_b200 import workloads as wl
    pos, goal, md, rng, walls = lcb.scene_full(scene, 0)
    m_cv, m_f = wl.weights(wseed)
    return lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, walls=walls, order="reference")


def _same(a, b, tag):
    d = np.abs(a.astype(np.int16) - b.astype(np.int16))
    assert d.max() <= 1, f"{tag}: a channel differs by {d.max()}"   # the tolerance of tests/test_render.py ...
    assert np.array_equal(a, b), f"{tag}: {int((d > 0).any(axis=2).sum())} pixels differ"   # ... and in fact none may differ


def test_render_static_equals_lc_render_on_two_calls_in_a_row():
    g = _sim()
    g.step(150)
    # call 1: a large picture fills most of the shared scratch
    want1 = g.render(0, cam_dist=0.6, width=1024, height=900)
    got1 = g.render_static(0, cam_dist=0.6, width=1024, height=900)
    _same(got1, want1, "first call")
    painted = (got1 != BACKGROUND).any(axis=2).sum()
    assert painted > 500, "the picture is empty: the test compares nothing"
    # call 2, straight after: another crowd state, another colour map, a SMALLER picture — stale keys of call 1 would show
    g.step(120)
    want2 = g.render(4, cam_dist=1.0, width=320, height=200)
    out = np.full((200, 320, 4), 7, np.uint8)
    got2 = g.render_static(4, cam_dist=1.0, width=320, height=200, out=out)
    assert got2 is out
    _same(got2, want2, "second call")
    assert not np.array_equal(got2, got1[:200, :320]), "the crowd did not move between the calls"
    # and the first request again: the static scratch keeps nothing of call 2 that matters
    g2 = _sim()
    g2.step(150)
    _same(g2.render_static(0, cam_dist=0.6, width=1024, height=900), want1, "a second handle sharing the scratch")


def test_render_static_allocates_nothing_and_takes_device_output():
    import torch
    import learn_comms_b200 as lcb
    g = _sim(3, 5)
    g.step(60)
    want = g.render(2, cam_dist=0.85, width=512, height=512)
    g.render_static(2, cam_dist=0.85, width=512, height=512)   # warm-up: kernels loaded, symbol addresses looked up
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    host = np.zeros((512, 512, 4), np.uint8)
    for _ in range(3):
        g.render_static(2, cam_dist=0.85, width=512, height=512, out=host)
    free1 = torch.cuda.mem_get_info()[0]
    assert free1 == free0, f"device memory moved by {free0 - free1} bytes across lc_render_static calls"
    _same(host, want, "host output")
    dev = torch.zeros((512, 512, 4), dtype=torch.uint8, device="cuda")   # the caller's buffer; the entry point allocates none
    rc = lcb.lib().lc_render_static(g.h, 2, C.c_float(0.85), 512, 512, C.c_void_p(dev.data_ptr()))
    assert rc == 0, lcb.lib().lc_last_error().decode()
    _same(dev.cpu().numpy(), want, "device output")


def test_render_static_rejects_sizes_beyond_its_scratch():
    import learn_comms_b200 as lcb
    g = _sim(3, 5)
    img = g.render_static(0, width=1024, height=1024)   # exactly the maximum
    _same(img, g.render(0, width=1024, height=1024), "maximum size")
    big = np.zeros((1024, 1025, 4), np.uint8)
    rc = lcb.lib().lc_render_static(g.h, 0, C.c_float(1.0), 1025, 1024, big.ctypes.data_as(C.c_void_p))
    assert rc == -5 and b"LC_RENDER_STATIC_MAX_PIXELS" in lcb.lib().lc_last_error()   # LC_ERR_CAPACITY, nothing written
    assert not big.any()
    with pytest.raises(lcb.LcError):
        g.render_static(0, width=0, height=10)
    assert 1024 * 1024 == MAX_PIXELS
