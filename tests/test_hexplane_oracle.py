"""CPU-only: pin oracle/hexplane_oracle.py against fixtures produced by the UNMODIFIED reference scene/hexplane.py
(tests/golden/hexplane_*.npz, written by tests/golden/make_golden_hexplane.py)."""
import glob
import os

import numpy as np
import pytest

import hexplane_oracle as hx

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = sorted(glob.glob(os.path.join(HERE, "golden", "hexplane_*.npz")))


def grids_of(g):
    return [[g["grid_%d_%d" % (l, ci)] for ci in range(6)] for l in range(int(g["L"]))]


def test_fixtures_present():
    assert len(FILES) >= 2


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(p) for p in FILES])
def test_forward_matches_reference(path):
    g = np.load(path)
    out, _ = hx.forward(g["pts"], g["time"], g["aabb"], grids_of(g))
    assert out.shape == g["out"].shape
    scale = np.abs(g["out"]).max()
    assert np.abs(out - g["out"]).max() <= 2e-6 * scale  # same fp32 expression tree; CPU/numpy may differ in FMA use


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(p) for p in FILES])
def test_backward_matches_reference_autograd(path):
    g = np.load(path)
    dgrids, dpts, dtime = hx.backward(g["pts"], g["time"], g["aabb"], grids_of(g), g["g_out"])
    for l in range(int(g["L"])):
        for ci in range(6):
            ref = g["g_grid_%d_%d" % (l, ci)]
            assert np.abs(dgrids[l][ci] - ref).max() <= 1e-5 * np.abs(ref).max() + 1e-7, (l, ci)
    assert np.abs(dpts - g["g_pts"]).max() <= 1e-5 * np.abs(g["g_pts"]).max()
    assert np.abs(dtime - g["g_time"]).max() <= 1e-5 * np.abs(g["g_time"]).max()
