"""drt_trace_primary_noalloc (include/drt.h): the allocation-free variant of the primary-ray probe, whose device-side staging
lives in a static device array inside libdrt.so. It has no counterpart in the reference,
so its checker is the existing entry point drt_trace_primary: same kernel, same inputs, the hit records must be
the same BITS (ids are integers; t, u, v are compared as uint32 -- the tolerance of the primary-hit parity tests is zero).

Collected last on purpose (file name). Passes alone in a fresh process and after the others.
"""
import ctypes
import os

import numpy as np
import pytest

import oracles as O

pytestmark = pytest.mark.gpu

NOALLOC_MAX_PIXELS = 262144  # DRT_NOALLOC_MAX_PIXELS in include/drt.h


@pytest.fixture(scope="module")
def drt():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import dustraytracer_b200 as d
    return d


def _scene(drt, name):
    s = drt.Scene()
    s.loadGLTFmodel(O.scene_path(name))
    drt.BVHBuilder().buildIterative(s)
    return s


def _same_bits(a, b):
    return np.array_equal(np.frombuffer(a.tobytes(), np.uint32), np.frombuffer(b.tobytes(), np.uint32))


def test_header_states_the_limit():
    hdr = open(os.path.join(O.ROOT, "include", "drt.h")).read()
    assert "#define DRT_NOALLOC_MAX_PIXELS %d" % NOALLOC_MAX_PIXELS in hdr


@pytest.mark.parametrize("flags", [0, 3])  # production arithmetic; STRICT | HOST_ARG_ORDER
def test_noalloc_probe_equals_the_allocating_probe_on_two_calls_in_a_row(drt, flags):
    room, cornell = _scene(drt, "room"), _scene(drt, "cornell_box")
    cam = drt.Camera()
    r_ref, r = drt.Renderer(), drt.Renderer()
    for q in (r_ref, r):
        q.set_flags(flags)
    # two calls in a row on the shared scratch: a different scene, size and frame index each time, the second image SMALLER
    # than the first, so stale records of the first call would show if anything but this call's output were copied out
    calls = [(room, 512, 512, 1), (cornell, 200, 120, 7), (room, 203, 77, 2)]
    outs = []
    for scene, w, h, frame in calls:
        r_ref.ResizeBuffer(w, h)
        r.ResizeBuffer(w, h)
        want = r_ref.trace_primary(cam, scene, frame)
        got = r.trace_primary_noalloc(cam, scene, frame)
        assert got.shape == want.shape == (w * h,)
        assert _same_bits(got, want), "%dx%d frame %d: %d records differ" % (w, h, frame, int((got != want).sum()))
        assert (want["prim_id"] >= 0).any()
        outs.append(got)
    # and the same call repeated returns the same bits (the scratch carries nothing from one call into the next)
    again = r.trace_primary_noalloc(cam, room, 2)
    assert _same_bits(again, outs[-1])
    # the closest-ray counter advanced once per pixel per call, like the allocating probe's
    assert r.counters()["closest_rays"] == r_ref.counters()["closest_rays"] + 203 * 77


def test_noalloc_probe_rejects_sizes_beyond_the_static_scratch(drt):
    from dustraytracer_b200 import _lib
    scene = _scene(drt, "cube")
    cam = drt.Camera()
    r = drt.Renderer()
    r.ResizeBuffer(512, 512)  # exactly the maximum: accepted
    assert r.trace_primary_noalloc(cam, scene, 1).size == NOALLOC_MAX_PIXELS
    r.ResizeBuffer(513, 512)
    before = r.kernel_launches()
    with pytest.raises(drt.DrtError) as e:
        r.trace_primary_noalloc(cam, scene, 1)
    assert e.value.code == -1 and "DRT_NOALLOC_MAX_PIXELS" in str(e.value) and "513x512" in str(e.value)
    assert r.kernel_launches() == before  # rejected before any launch
    # too small an output buffer is rejected as well
    r.ResizeBuffer(64, 64)
    out = np.zeros(10, dtype=[("prim_id", np.int32), ("t", np.float32), ("u", np.float32), ("v", np.float32)])
    rc = _lib.lib().drt_trace_primary_noalloc(r._h, ctypes.byref(cam), scene._h, 1, out.ctypes.data, out.size)
    assert rc == -1
