"""Sharded solves on the GPU.

On a one-GPU box the ranks are handles of one process on one device (the "local"
communicator backend: a reduction kernel over the ranks' HBM buffers, ordered by a host
barrier), so the sharded LM logic -- the same code that runs over NCCL -- is checked against
the unsharded solve.  With >= 2 GPUs the NCCL path runs as one process per GPU (torchrun).
"""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

from groups import run_ranks
from mayamatchmovesolver_b200 import synth
from mayamatchmovesolver_b200.distributed import SingleGroup, make_comm, solve_global_sharded, solve_per_frame_sharded
from mayamatchmovesolver_b200.problem import Problem

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpu_count():
    n = ctypes.c_int(0)
    try:
        ctypes.CDLL("libcudart.so").cudaGetDeviceCount(ctypes.byref(n))
    except OSError:
        return 0
    return n.value


def test_world1_shard_equals_unsharded():
    sb, _ = synth.make_scene(F=12, B=16, seed=9, fixed_bundles=3, period=12)
    with Problem(sb) as p:
        r1 = p.solve()
    r = solve_global_sharded(sb, SingleGroup())
    assert (r.reason_num, r.iteration_num) == (r1.reason_num, r1.iteration_num)
    # the assembly uses floating-point atomics: two runs agree to rounding, not bitwise
    np.testing.assert_allclose(r.attr_values, r1.attr_values, rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(r.error_dist, r1.error_dist, atol=1e-5)


GLOBAL_CASES = {
    "camera_bundles_24x20": dict(F=24, B=20, seed=21, fixed_bundles=4, period=24),
    "camera_bundles_40x30": dict(F=40, B=30, seed=5, fixed_bundles=4, period=40),
    "sparse_weighted": dict(F=20, B=14, seed=6, fixed_bundles=3, drop_fraction=0.15, weights=True, period=20),
    "camera_only": dict(F=9, B=12, seed=2, solve_bundles=False),
}


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("name", list(GLOBAL_CASES))
def test_global_sharded_matches_unsharded(name, world):
    sb, _ = synth.make_scene(**GLOBAL_CASES[name])
    with Problem(sb) as p:
        r1 = p.solve()
    outs = run_ranks(world, lambda g: solve_global_sharded(sb, g, local_backend=True))
    for r in outs:
        # whole-problem results are identical on every rank
        assert (r.reason_num, r.iteration_num, r.iteration_attempt_num, r.error_final) == \
               (outs[0].reason_num, outs[0].iteration_num, outs[0].iteration_attempt_num, outs[0].error_final)
        np.testing.assert_array_equal(r.attr_values, outs[0].attr_values)
    r = outs[0]
    print(name, world, "unsharded", r1, "sharded", r)
    # same LM decisions as the unsharded solve; values differ only by summation order
    assert (r.reason_num, r.iteration_num, r.iteration_attempt_num) == (r1.reason_num, r1.iteration_num, r1.iteration_attempt_num)
    scale = np.maximum(np.abs(r1.attr_values), 1.0)
    assert np.max(np.abs(r.attr_values - r1.attr_values) / scale) <= 1e-6
    assert np.max(np.abs(r.error_dist - r1.error_dist)) <= 1e-5
    assert abs(r.error_final - r1.error_final) <= 1e-6 * r1.error_final
    assert abs(r.error_avg - r1.error_avg) <= 1e-6 and abs(r.error_max - r1.error_max) <= 1e-5


def test_session_communicator_is_reused_across_solves():
    """mmb_comm_create once, mmb_problem_use_comm per handle (the handle does not own it)"""
    sb, _ = synth.make_scene(**GLOBAL_CASES["camera_bundles_24x20"])
    with Problem(sb) as p:
        r1 = p.solve()

    def body(g):
        comm = make_comm(g, local_backend=True)
        a = solve_global_sharded(sb, g, comm=comm)
        b = solve_global_sharded(sb, g, comm=comm)
        comm.close()
        return a, b
    for a, b in run_ranks(2, body):
        for r in (a, b):
            assert (r.reason_num, r.iteration_num, r.iteration_attempt_num) == (r1.reason_num, r1.iteration_num, r1.iteration_attempt_num)
            scale = np.maximum(np.abs(r1.attr_values), 1.0)
            assert np.max(np.abs(r.attr_values - r1.attr_values) / scale) <= 1e-6


@pytest.mark.parametrize("world", [2, 4])
def test_per_frame_sharded_matches_unsharded(world):
    # cfg2 shape, smaller: independent per-frame solves, no collective
    sb, _ = synth.make_scene(F=30, B=16, seed=22, solve_bundles=False, period=30)
    with Problem(sb, frame_solve_mode=1) as p:
        q1, per1 = p.solve_per_frame()
    outs = run_ranks(world, lambda g: solve_per_frame_sharded(sb, g))
    for rp, per in outs:
        np.testing.assert_array_equal(rp.attr_values, q1.attr_values)
        np.testing.assert_array_equal(rp.error_dist, q1.error_dist)
        assert [x.iteration_num for x in per] == [x.iteration_num for x in per1]
        assert [x.reason_num for x in per] == [x.reason_num for x in per1]
        assert rp.iteration_num == q1.iteration_num and rp.reason_num == q1.reason_num
        assert abs(rp.error_final - q1.error_final) <= 1e-12 * q1.error_final
        assert abs(rp.error_avg - q1.error_avg) <= 1e-12 * q1.error_avg
        assert rp.error_max == q1.error_max and rp.error_min == q1.error_min


def test_more_ranks_than_frames_is_refused():
    from mayamatchmovesolver_b200._capi import MmbError
    sb, _ = synth.make_scene(F=2, B=8, seed=1, solve_camera=False)
    with pytest.raises(MmbError):
        Problem(sb, shard=(2, 3))


@pytest.mark.skipif(_gpu_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_nccl_two_gpus():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29641", os.path.join(ROOT, "tests", "mp_worker.py"), "gpu"]
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600,
                         env=dict(os.environ, PYTHONPATH=ROOT))
    assert out.returncode == 0, out.stdout[-4000:]
    assert "rank 0 ok" in out.stdout and "rank 1 ok" in out.stdout
