"""The benched kernel (RT_KERNEL_MEGA) against the REFERENCE'S OWN renders at real geometry and real sample counts.

tests/golden/frames.npz (tests/golden/make_golden.py frames) holds Camera_render of the unmodified reference
(/root/reference/src/raytracing.c:86-135, per-pixel seeds (17+j, 23+i)) for Book 1 final 400x225x256 spp, the Cornell
box 200x200x256 spp and Book 2 final 200x200x128 spp, depth 50, as 8-bit frames, and the same frames from independent
streams (the oracle, LIBM mode, per-(pixel, sample) seeds) -- the reference-vs-reference noise floor.

Stated tolerance, per channel, on the 8-bit tonemapped frames:
    RMSE(megakernel, reference)  <= 0.5 * RMSE(reference, independent streams)
    |mean(megakernel) - mean(reference)| <= 0.5 / 255
with the SAME seeds on both sides (RT_RNG_PIXEL against the reference's frame, RT_RNG_SAMPLE against the oracle's).
A systematic error on long paths, at chunk boundaries or in the trip bookkeeping would show as bias or as an RMSE at
the noise floor; agreement far inside the floor means most paths are literally the reference's paths.
"""
from __future__ import annotations

import numpy as np
import pytest

from conftest import GOLDEN

FRAME_SCENES = (1, 6, 7)


@pytest.fixture(scope="module")
def frames():
    return np.load(GOLDEN / "frames.npz")


def _stats(a, b):
    d = a.astype(np.float64) - b.astype(np.float64)
    return np.sqrt((d ** 2).mean(axis=(0, 1))), d.mean(axis=(0, 1)), float(np.count_nonzero(d)) / d.size


@pytest.mark.parametrize("scene_id", FRAME_SCENES)
def test_oracle_reproduces_the_reference_frame(rtc, oracle, frames, scene_id):
    """CPU: the pin at real size -- the oracle (LIBM, per-pixel streams) renders the reference's frame byte for byte
    (about half a minute for the three frames on 8 cores)."""
    width, height, spp, depth = (int(v) for v in frames[f"shape_{scene_id}"])
    sess = rtc.Session(scene_id, width=width, spp=spp, depth=depth)
    rgb8, _, _ = oracle.render(sess.flatten(), sess.camera, rtc.make_opts(kernel="exact", rng="pixel"),
                               oracle.MATH_LIBM)
    assert np.array_equal(rgb8, frames[f"ref_{scene_id}"])


@pytest.mark.gpu
@pytest.mark.parametrize("scene_id", FRAME_SCENES)
@pytest.mark.parametrize("rng", ["pixel", "sample"])
def test_megakernel_against_reference_frames(rtc, frames, scene_id, rng):
    width, height, spp, depth = (int(v) for v in frames[f"shape_{scene_id}"])
    ref, other = frames[f"ref_{scene_id}"], frames[f"other_{scene_id}"]
    floor, _, _ = _stats(ref, other)
    sess = rtc.Session(scene_id, width=width, spp=spp, depth=depth)
    assert (sess.height, sess.width) == (height, width)
    scene = rtc.Scene(sess.flatten())
    try:
        rgb8, _, st = scene.render(sess.camera, rtc.make_opts(kernel="mega", rng=rng))
    finally:
        scene.close()
    want = ref if rng == "pixel" else other  # same seeds on both sides
    rmse, bias, differing = _stats(rgb8, want)
    assert st["samples"] == width * height * spp
    assert (rmse <= 0.5 * floor).all(), (rmse, floor)
    assert (np.abs(bias) <= 0.5).all(), bias
    # and against the OTHER streams it must sit at the noise floor, not below or above it (an estimator check)
    cross, cross_bias, _ = _stats(rgb8, other if rng == "pixel" else ref)
    assert (cross <= 1.15 * floor + 0.2).all() and (cross >= 0.85 * floor - 0.2).all(), (cross, floor)
    assert (np.abs(cross_bias) <= 0.5).all(), cross_bias
