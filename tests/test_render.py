"""Headless DrawParticles: the CUDA point splat against its NumPy restatement (oracle/render_oracle.py)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import render_oracle as ro  # noqa: E402

from _oracle import PARTICLE  # noqa: E402


def _three():
    ps = np.zeros(3, dtype=PARTICLE)
    ps["px"] = [0.0, 10.9, 11.0]
    ps["py"] = [0.0, 0.0, 0.0]
    ps["mass"] = [3e10, 100.0, 0.0]      # core, ordinary, massless
    ps["radius"] = [4.0, 2.0, 1.0]
    return ps


def test_render_oracle_small_known_image():
    """Closed-form check of the restatement itself: colours, painter's order, integer truncation."""
    img = ro.render(_three(), 32, 16, (0.0, 0.0, 1.0))
    cx, cy = 16, 8                                      # world origin at the image centre
    assert tuple(img[cy, cx]) == ro.CC                  # core disc at the centre
    assert tuple(img[cy, cx + 5]) == ro.BG              # 5.5 px from the centre: outside r = 4
    # particle 1 at (int)10.9 = 10 and particle 2 at 11: the later (massless) one wins where they overlap
    assert tuple(img[cy, cx + 11]) == ro.EP
    assert tuple(img[cy, cx + 8]) == ro.NP_
    assert tuple(img[0, 0]) == ro.BG
    assert img[:, :, 3].min() == 255
    assert ro.MIN_GC_MASS == pytest.approx(4 * np.pi * 30 / 3 * 200 ** 3, rel=1e-6)


def test_fit_camera_oracle():
    ps = _three()
    ps["px"] = [-1000.0, 3000.0, 0.0]
    ps["py"] = [-500.0, 500.0, 0.0]
    tx, ty, zoom = ro.fit_camera(ps, 1280, 720)
    assert (tx, ty) == (1000.0, 0.0)
    assert zoom == pytest.approx(min(0.9 * 1280 / 4000, 0.9 * 720 / 1000))
    ps["px"] *= 0.01                                    # everything fits at zoom 1: never zoom in (ref main.c:215)
    ps["py"] *= 0.01
    assert ro.fit_camera(ps, 1280, 720)[2] == 1.0


@pytest.mark.gpu
@pytest.mark.parametrize("n,gal,size,min_radius", [(6000, 3, (320, 180), 0.0), (6000, 3, (640, 360), 40.0),
                                                    (20000, 2, (256, 256), 100.0)])
def test_cuda_splat_matches_restatement_bitwise(nb, oracle, n, gal, size, min_radius):
    arr, m = oracle.partition(nb.make_galaxies(n, gal))
    w, h = size
    with nb.SimPipeline(len(arr), m) as sim:
        sim.set_data(arr)
        sim.update(3, 0.01)
        state = sim.get_data()
        cam = sim.fit_camera(w, h)
        img = sim.render(w, h, cam, min_radius)
    ref_cam = ro.fit_camera(state, w, h)
    assert cam == tuple(float(v) for v in ref_cam)
    ref = ro.render(state, w, h, cam, min_radius)
    assert img.shape == ref.shape
    assert np.array_equal(img, ref), f"{np.count_nonzero((img != ref).any(axis=2))} pixels differ"
    assert (img[:, :, :3] != 22).any()                  # something was drawn


@pytest.mark.gpu
def test_cuda_splat_zoomed_in_and_offscreen(nb):
    ps = _three()
    with nb.SimPipeline(3, 2) as sim:
        sim.set_data(ps)
        for cam in [(0.0, 0.0, 1.0), (5.0, 1.0, 3.0), (1e6, 1e6, 1.0), (0.0, 0.0, 0.05)]:
            img = sim.render(64, 48, cam, 0.0)
            assert np.array_equal(img, ro.render(ps, 64, 48, cam, 0.0))


@pytest.mark.gpu
def test_write_ppm(nb, tmp_path):
    ps = _three()
    with nb.SimPipeline(3, 2) as sim:
        sim.set_data(ps)
        img = sim.render(32, 16, (0.0, 0.0, 1.0))
    path = tmp_path / "frame.ppm"
    nb.write_ppm(str(path), img)
    raw = path.read_bytes()
    assert raw.startswith(b"P6\n32 16\n255\n") and len(raw) == len(b"P6\n32 16\n255\n") + 32 * 16 * 3
