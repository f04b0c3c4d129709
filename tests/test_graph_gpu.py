"""The unsharded global solve is ONE CUDA graph launch (conditional WHILE / IF nodes steered
from the device): no host round trip per LM iteration (north_star item 4).  The host-driven
loop (used by sharded solves) enqueues the same kernels, so both drivers must take the same
decisions and agree to rounding."""
import threading
import time

import numpy as np
import pytest

from mayamatchmovesolver_b200 import synth
from mayamatchmovesolver_b200.problem import Problem

pytestmark = pytest.mark.gpu

LENS = dict(degree2_distortion=0.05, degree2_u=0.002, degree2_v=-0.002, degree4_distortion=0.01)
SCENES = {
    "camera_bundles": dict(F=40, B=30, seed=5, fixed_bundles=4, period=40),
    "sparse_weighted": dict(F=20, B=14, seed=6, fixed_bundles=3, drop_fraction=0.15, weights=True, period=20),
    "bundles_only": dict(F=2, B=20, seed=1, solve_camera=False),
    "camera_only": dict(F=6, B=12, seed=2, solve_bundles=False),
    "lens_focal": dict(F=12, B=16, seed=4, fixed_bundles=16, solve_bundles=False, lens=LENS, solve_focal=True,
                       solve_lens=("degree2_distortion", "degree2_u", "degree2_v", "degree4_distortion"),
                       focal_init=33.0, period=12),
}


def _solve(sb, monkeypatch, graph, **opt):
    monkeypatch.setenv("MMB_LM_GRAPH", "1" if graph else "0")
    monkeypatch.setenv("MMB_LM_GRAPH_STRICT", "1")  # a graph that cannot be built is an error here
    with Problem(sb, **opt) as p:
        r = p.solve()
        stats = p.solve_stats()
        launches = p.launch_count()
    return r, stats, launches


@pytest.mark.parametrize("name", list(SCENES))
def test_graph_loop_equals_host_loop(name, monkeypatch):
    sb, _ = synth.make_scene(**SCENES[name])
    rg, sg, lg = _solve(sb, monkeypatch, True)
    rh, sh, lh = _solve(sb, monkeypatch, False)
    assert sg == (1, 1), sg                 # one graph launch, one synchronisation
    assert sh[1] == 0 and sh[0] > rh.iteration_attempt_num  # several reads per LM iteration
    assert (rg.reason_num, rg.iteration_num, rg.iteration_attempt_num) == (rh.reason_num, rh.iteration_num, rh.iteration_attempt_num)
    assert rg.iteration_jacobian_num == rh.iteration_jacobian_num and rg.error_is_better == rh.error_is_better
    # same kernels, same decisions, order-independent accumulation: the two drivers agree bit for bit
    # (the graph runs the Cholesky beside the SYRK, the host loop after it: same arithmetic per tile)
    assert np.array_equal(rg.x, rh.x)
    assert np.array_equal(rg.attr_values, rh.attr_values)
    assert np.array_equal(rg.error_dist, rh.error_dist)
    assert [rg.error_final, rg.error_avg, rg.error_min, rg.error_max] == [rh.error_final, rh.error_avg, rh.error_min, rh.error_max]
    # the device-side launch accounting reproduces the host's count (graph adds the k_cond kernels)
    assert lg >= lh
    assert rg.timer_jacobian_s > 0 and rg.timer_linear_s > 0 and rg.timer_function_s > 0


def test_graph_replay_on_same_handle(monkeypatch):
    """the instantiated graph is reused by a second solve of the same handle"""
    monkeypatch.setenv("MMB_LM_GRAPH", "1")
    monkeypatch.setenv("MMB_LM_GRAPH_STRICT", "1")
    sb, _ = synth.make_scene(**SCENES["camera_bundles"])
    with Problem(sb) as p:
        r1 = p.solve()
        r2 = p.solve()
        assert p.solve_stats() == (1, 1)
    np.testing.assert_allclose(r1.x, r2.x, rtol=1e-8, atol=1e-9)
    assert r1.iteration_num == r2.iteration_num


def test_iteration_cap_and_improper_input(monkeypatch):
    monkeypatch.setenv("MMB_LM_GRAPH", "1")
    monkeypatch.setenv("MMB_LM_GRAPH_STRICT", "1")
    sb, _ = synth.make_scene(**SCENES["camera_bundles"])
    with Problem(sb, iter_max=3) as p:
        r = p.solve()
    assert r.reason_num == 5 and r.iteration_num == 3        # maxfev reached (info 5)
    with Problem(sb, tau=0.0) as p:                            # factor = 0 -> "Improper input parameters"
        r = p.solve()
    assert r.reason_num == 0 and r.iteration_num == 1 and r.iteration_attempt_num == 0


def test_cancel_stops_device_loop(monkeypatch):
    """mmb_request_cancel from another thread (MComputation::isInterruptRequested analogue,
    adjust_solveFunc.cpp:321,567): the loop-head kernel polls the mapped flag."""
    monkeypatch.setenv("MMB_LM_GRAPH", "1")
    monkeypatch.setenv("MMB_LM_GRAPH_STRICT", "1")
    sb, _ = synth.make_scene(1000, 500, 3, noise_px=0.5)
    with Problem(sb, iter_max=100000, eps1=0.0, eps2=0.0, eps3=0.0) as p:
        t = threading.Timer(0.02, p.request_cancel)
        t0 = time.perf_counter()
        t.start()
        r = p.solve()
        dt = time.perf_counter() - t0
        t.join()
    assert r.user_interrupted and r.reason_num == -1
    assert dt < 5.0


def test_cholesky_beside_syrk_times_out_and_falls_back(monkeypatch):
    """The graph runs the Cholesky of the reduced system beside the SYRK that produces its tiles
    (per-tile ready flags).  If a flag never came the wait is bounded, the handle goes back to one
    kernel after the other and solves again: here the SYRK is told not to publish its flags."""
    sb, _ = synth.make_scene(**SCENES["camera_bundles"])
    ref, _, _ = _solve(sb, monkeypatch, True)
    monkeypatch.setenv("MMB_TEST_DROP_SYRK_FLAGS", "1")
    t0 = time.perf_counter()
    r, stats, _ = _solve(sb, monkeypatch, True)
    assert time.perf_counter() - t0 < 30.0
    assert stats == (1, 1)          # of the second attempt
    assert (r.reason_num, r.iteration_num) == (ref.reason_num, ref.iteration_num)
    assert np.array_equal(r.x, ref.x) and np.array_equal(r.error_dist, ref.error_dist)
