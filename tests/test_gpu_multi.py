"""The multi-device renderer inside the library (drt_multi_*, NCCL bound at run time): a C++ host gets N GPUs without Python.
Runs with as many devices as the box has (2 under `gpurun --gpus 2`; with one device the n = 1 path and the error paths).
Checker: the single-device renderer (itself bit-exact against the oracle in test_gpu_parity.py)."""
import ctypes

import numpy as np
import pytest

import oracles as O

pytestmark = pytest.mark.gpu
STRICT = 1 | 2


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.fixture(scope="module")
def drt():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import dustraytracer_b200 as d
    return d


def ndev():
    import torch
    return torch.cuda.device_count()


def host_scene(drt, name):
    s = drt.Scene()
    s.loadGLTFmodel(O.scene_path(name))
    return s


def single(drt, name, W, H, st, batches, first=1, stride=1):
    s = host_scene(drt, name)
    drt.BVHBuilder().buildIterative(s)
    r = drt.Renderer(flags=STRICT)
    r.m_RendererSettings = st
    r.ResizeBuffer(W, H)
    r.set_frame_schedule(first, stride)
    for b in batches:
        r.Render(drt.Camera(), s, b)
    return r.readback(drt.DRT_ACCUM_RGB32F), r.counters()


def test_one_device_is_the_single_renderer(drt):
    W, H = 101, 67
    st = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=3, max_samples=100000)
    m = drt.MultiRenderer([0], flags=STRICT)
    m.m_RendererSettings = st
    m.set_scene(host_scene(drt, "room"))
    m.ResizeBuffer(W, H)
    m.Render(drt.Camera(), 3)
    m.Render(drt.Camera(), 2)
    want, c = single(drt, "room", W, H, st, [3, 2])
    assert np.array_equal(bits(m.readback(drt.DRT_ACCUM_RGB32F)), bits(want))
    assert m.counters()["closest_rays"] == c["closest_rays"] and m.getSampleCount() == 6 and m.nccl_reduces() == 0


def test_bad_device_lists_are_rejected(drt):
    with pytest.raises(drt.DrtError):
        drt.MultiRenderer([0, 0])
    with pytest.raises(drt.DrtError):
        drt.MultiRenderer([ndev() + 3])
    with pytest.raises(drt.DrtError):
        drt.MultiRenderer([])


@pytest.mark.parametrize("name", ["room", "animemaid"])
def test_tile_sharding_over_the_devices_reproduces_the_one_gpu_bits(drt, name):
    n = min(ndev(), 4)
    if n < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    W, H = 203, 117
    st = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=4, max_samples=100000)
    m = drt.MultiRenderer(list(range(n)), flags=STRICT)
    m.m_RendererSettings = st
    m.set_scene(host_scene(drt, name))
    m.set_sharding(tiles=True)
    m.ResizeBuffer(W, H)
    m.Render(drt.Camera(), 3)
    m.Render(drt.Camera(), 2)
    want, c = single(drt, name, W, H, st, [3, 2])
    got = m.readback(drt.DRT_ACCUM_RGB32F)
    assert np.array_equal(bits(got), bits(want))                     # x + 0 + ... + 0 = x: the single-GPU sums
    disp = m.readback(drt.DRT_RGBA32F)
    assert np.array_equal(bits(disp[:, :, :3]), bits(want / np.float32(5))) and np.all(disp[:, :, 3] == 1)
    mc = m.counters()
    assert (mc["closest_rays"], mc["shadow_rays"]) == (c["closest_rays"], c["shadow_rays"])
    assert m.nccl_reduces() == 2 and m.getSampleCount() == 6


def test_sample_range_sharding_equals_the_sum_of_its_ranks(drt):
    n = min(ndev(), 4)
    if n < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    W, H = 160, 90
    st = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=4, max_samples=100000)
    m = drt.MultiRenderer(list(range(n)), flags=STRICT)
    m.m_RendererSettings = st
    m.set_scene(host_scene(drt, "room"))
    m.ResizeBuffer(W, H)
    per = 3
    m.Render(drt.Camera(), per * n)   # device g: frame indices g+1, g+1+n, g+1+2n
    parts, rays = [], 0
    for g in range(n):
        a, c = single(drt, "room", W, H, st, [per], first=g + 1, stride=n)
        parts.append(a.astype(np.float64))
        rays += c["closest_rays"] + c["shadow_rays"]
    got = m.readback(drt.DRT_ACCUM_RGB32F)
    want = sum(parts)
    assert np.abs(got - want).max() <= 4e-6 * max(1.0, np.abs(want).max())   # fp32 reduce order inside NCCL; the terms are bit-exact
    if n == 2:
        assert np.array_equal(bits(got), bits((parts[0].astype(np.float32) + parts[1].astype(np.float32))))
    mc = m.counters()
    assert mc["closest_rays"] + mc["shadow_rays"] == rays
    assert m.getSampleCount() == per * n + 1
    # a frame of the same per * n samples from one device: same estimator, different summation order only
    one, _ = single(drt, "room", W, H, st, [per * n])
    assert np.abs(got - one).max() <= 1e-4
    # max_samples is honoured for the whole frame (Renderer.cu:82)
    m.m_RendererSettings = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=4, max_samples=per * n + 2)
    m.Render(drt.Camera(), 100)
    assert m.getSampleCount() == per * n + 2
