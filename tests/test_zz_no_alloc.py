"""mmb_lens_apply_layered_no_alloc: the layered bulk-lens kernel with its device scratch (layer
table + staging of the points) in a static __device__ array -- nothing is allocated.  It must
give what mmb_lens_apply_layered gives, on two calls in a row (the array keeps the previous
call's contents), alone in a fresh process or after every other test (this file is collected
last on purpose)."""
import numpy as np
import pytest

from mayamatchmovesolver_b200 import lens
from mayamatchmovesolver_b200._capi import MmbError

pytestmark = pytest.mark.gpu

CAMS = np.array([[3.5, 3.6, 2.4, 1.0, 0.0, 0.0], [3.5, 3.6, 2.4, 1.0, 0.05, -0.03], [3.5, 3.6, 2.4, 1.1, 0.0, 0.02]])
LAYERS = [(0, None, CAMS[0]),
          (2, [0.05, 0.002, -0.001, 0.01, 0.0, 0.0, 10.0, 0.02], CAMS[1]),
          (3, [0.01, 0.005, -0.004, 0.002, 0.003, -0.002, 0.001, 0.001, -0.001, 0.0005, 2.0, 1.01, 0.99], CAMS[2])]
# the two entry points run the same kernel on the same inputs: the bulk-lens tests' own
# tolerances (test_lens_bulk_gpu.py TOL) hold with room to spare -- the results are identical
TOL = {0: 1e-13, 1: 1e-9}


@pytest.mark.parametrize("direction", [0, 1])
def test_no_alloc_matches_allocating_entry_point_twice(direction):
    rng = np.random.default_rng(40 + direction)
    # first call fills the whole scratch, the second is shorter and uses other layers: stale
    # points or a stale layer table from the first call would show
    first = rng.uniform(-0.45, 0.45, (lens.NO_ALLOC_MAX_POINTS, 2))
    second = rng.uniform(-0.45, 0.45, (1000, 2))
    for pts, layers in ((first, LAYERS), (second, LAYERS[1:2])):
        want = lens.apply_layered(layers, pts, direction)
        got = lens.apply_layered(layers, pts, direction, no_alloc=True)
        assert np.isfinite(got).all()
        assert np.abs(got - want).max() <= TOL[direction], np.abs(got - want).max()
        assert np.array_equal(got, want)


def test_no_alloc_limits():
    pts = np.zeros((lens.NO_ALLOC_MAX_POINTS + 1, 2))
    with pytest.raises(MmbError) as e:
        lens.apply_layered(LAYERS, pts, 0, no_alloc=True)
    assert "MMB_LENS_NO_ALLOC_MAX_POINTS" in str(e.value)
    assert lens.apply_layered(LAYERS, np.zeros((0, 2)), 0, no_alloc=True).shape == (0, 2)
    # still usable after a refused call
    one = np.array([[0.1, -0.2]])
    assert np.array_equal(lens.apply_layered(LAYERS, one, 1, no_alloc=True), lens.apply_layered(LAYERS, one, 1))
