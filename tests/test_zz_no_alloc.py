"""rt_render_device_noalloc (include/rtc_b200.h): the megakernel render with its scratch in the library's file-scope
device array -- for callers that cannot call cudaMalloc.  No counterpart in the reference,
so the check is against the EXISTING entry point, rt_render_device, on the same scene, camera
and seeds, through the C ABI: within the tolerance the sample-range tests use (rtol 2e-5, atol 1e-5 on the float
radiance sums; bit-identical where both split the work the same way), on two calls in a row -- the array is shared
and keeps the last call's contents, so the second call runs over a dirty scratch.

Collected last on purpose.  Needs nothing from the other tests: it passes alone in
a fresh process and after the rest of the suite.  torch is only the caller's device memory here (inputs and outputs are
the caller's business; the entry point itself allocates nothing)."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL, ATOL = 2e-5, 1e-5  # tests/test_gpu_mega.py::test_sample_ranges_add_up


def _both(rtc, scene_id, width, spp, **kw):
    """(radiance of rt_render_device, radiance of rt_render_device_noalloc) for one frame, as numpy arrays."""
    import torch

    sess = rtc.Session(scene_id, width=width, spp=spp, **kw)
    cam = sess.camera
    scene = rtc.Scene(sess.flatten(), device=0)
    try:
        opts = rtc.make_opts(kernel="mega", rng="sample")
        dev = torch.device("cuda", 0)
        shape = (cam.img_height, cam.img_width, 3)
        want = torch.full(shape, -1.0, dtype=torch.float32, device=dev)
        got = torch.full(shape, -2.0, dtype=torch.float32, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        scene.render_device(cam, opts, want.data_ptr(), stream)
        torch.cuda.synchronize(dev)  # calls that overlap on the GPU would race on the shared array: serialise
        scene.render_device_noalloc(cam, opts, got.data_ptr(), stream)
        torch.cuda.synchronize(dev)
        return want.cpu().numpy(), got.cpu().numpy()
    finally:
        scene.close()


def test_no_alloc_render_matches_the_existing_entry_point_twice_in_a_row(rtc):
    # call 1: a small Book 1 frame -- all partial-sum frames fit, the work split is rt_render_device's own
    want, got = _both(rtc, 1, 96, 32)
    assert np.isfinite(want).mean() > 0.99 and want.max() > 1.0  # a lit frame, not the fill value
    assert np.allclose(got, want, rtol=RTOL, atol=ATOL, equal_nan=True)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))  # same split: the same bits
    # call 2, right after, over the scratch call 1 left behind: a 1000x1000 Cornell frame (12 MB per partial frame,
    # so only two of them fit the 32 MiB array and the trailing work items are merged: the sums are added in
    # another order, hence the tolerance)
    want, got = _both(rtc, 6, 1000, 16)
    assert want.shape == (1000, 1000, 3) and want.max() > 1.0
    assert np.allclose(got, want, rtol=RTOL, atol=ATOL, equal_nan=True)
    # call 3: back to a small frame of another scene (media, textures, 512-thread kernel): stale partial sums of the
    # larger frame must not leak into it
    want, got = _both(rtc, 7, 64, 16)
    assert np.allclose(got, want, rtol=RTOL, atol=ATOL, equal_nan=True)


def test_no_alloc_render_refuses_what_does_not_fit(rtc):
    import torch

    lib = rtc.lib.b200()
    sess = rtc.Session(7, width=3840, spp=8, aspect=16.0 / 9.0)  # 8.3 Mpixel: beyond RT_NOALLOC_MAX_PIXELS
    cam = sess.camera
    assert cam.img_width * cam.img_height > (32 << 20) // 24
    scene = rtc.Scene(sess.flatten(), device=0)
    try:
        out = torch.zeros(16, dtype=torch.float32, device="cuda:0")  # never written: the call must fail first
        opts = rtc.make_opts(kernel="mega", rng="sample")
        with pytest.raises(rtc.lib.RtError) as err:
            scene.render_device_noalloc(cam, opts, out.data_ptr())
        assert err.value.status == rtc.abi.RT_ERR_INVALID and "exceeds the maximum" in str(err.value)
        small = rtc.Session(7, width=32, spp=2).camera
        with pytest.raises(rtc.lib.RtError) as err:
            scene.render_device_noalloc(small, rtc.make_opts(kernel="exact", rng="sample"), out.data_ptr())
        assert err.value.status == rtc.abi.RT_ERR_UNSUPPORTED
        torch.cuda.synchronize()
        assert float(out.abs().sum()) == 0.0
        assert lib.rt_render_device_noalloc(None, C.byref(small), C.byref(opts), out.data_ptr(), None) == rtc.abi.RT_ERR_INVALID
    finally:
        scene.close()
