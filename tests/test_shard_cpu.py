"""Host logic of the multi-GPU path on CPU: frame ranges, the
local -> unsharded parameter map, merging of shard vectors and of per-frame results.
The two-process case runs over torch.distributed with the gloo backend (world size 2)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from mayamatchmovesolver_b200 import _capi, synth
from mayamatchmovesolver_b200.distributed import merge_shard_vectors
from mayamatchmovesolver_b200.problem import aggregate_results, shard_frames, shard_param_map

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("F,world", [(10, 1), (10, 3), (240, 8), (7, 7), (1000, 8), (5, 8)])
def test_shard_frames_partition(F, world):
    ranges = [shard_frames(F, world, r) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == F
    for (a0, a1), (b0, b1) in zip(ranges, ranges[1:]):
        assert a1 == b0
    sizes = [b - a for a, b in ranges]
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def test_shard_frames_bad_arguments():
    lib = _capi.load()
    f0, f1 = C.c_int32(), C.c_int32()
    assert lib.mmb_shard_frames(10, 0, 0, C.byref(f0), C.byref(f1)) == _capi.MMB_ERR_INVALID_ARGUMENT
    assert lib.mmb_shard_frames(10, 2, 2, C.byref(f0), C.byref(f1)) == _capi.MMB_ERR_INVALID_ARGUMENT
    assert lib.mmb_shard_frames(10, 2, 1, None, None) == _capi.MMB_ERR_INVALID_ARGUMENT


@pytest.mark.parametrize("world", [1, 2, 3, 5])
def test_param_map_covers_reference_order(world):
    # camera tx..rz animated (attr-major, frame-minor: adjust_relationships.cpp:223-337), bundles static
    F, B = 11, 4
    sb, _ = synth.make_scene(F, B, seed=3)
    n = 6 * F + 3 * B
    maps = [shard_param_map(sb, world, r) for r in range(world)]
    seen = np.zeros(n, np.int32)
    for r, m in enumerate(maps):
        f0, f1 = shard_frames(F, world, r)
        assert len(m) == 6 * (f1 - f0) + 3 * B
        # animated attr a, local frame f -> a*F + f0 + f
        for a in range(6):
            np.testing.assert_array_equal(m[a * (f1 - f0):(a + 1) * (f1 - f0)], a * F + np.arange(f0, f1))
        np.testing.assert_array_equal(m[6 * (f1 - f0):], 6 * F + np.arange(3 * B))
        seen[m] += 1
    assert (seen[:6 * F] == 1).all() and (seen[6 * F:] == world).all()
    merged = merge_shard_vectors(n, maps, [m.astype(np.float64) for m in maps])
    np.testing.assert_array_equal(merged, np.arange(n))


def _result(**kw):
    r = _capi.Result()
    for k, v in kw.items():
        setattr(r, k, v)
    return r


def test_aggregate_results_add_divide():
    # SolverResult::add / divide, adjust_results.h:89-117
    rs = [_result(success=1, reason_num=2, error_final=4.0, error_avg=1.0, error_min=0.5, error_max=2.0,
                  iteration_num=5, iteration_function_num=6, iteration_jacobian_num=12, error_is_better=1),
          _result(success=1, reason_num=1, error_final=2.0, error_avg=3.0, error_min=0.25, error_max=1.0,
                  iteration_num=3, iteration_function_num=4, iteration_jacobian_num=6, error_is_better=1),
          _result(success=0, reason_num=5, error_final=3.0, error_avg=2.0, error_min=1.0, error_max=7.0,
                  iteration_num=2, iteration_function_num=3, iteration_jacobian_num=6, error_is_better=0)]
    a = aggregate_results(rs)
    assert a.success == 0 and a.reason_num == 5 and a.error_is_better == 0
    assert a.error_final == pytest.approx(3.0) and a.error_avg == pytest.approx(2.0)
    assert a.error_min == 0.25 and a.error_max == 7.0
    assert (a.iteration_num, a.iteration_function_num, a.iteration_jacobian_num) == (10, 13, 24)
    assert a.reason_str.startswith(b"Number of calls to fcn")
    one = aggregate_results(rs[:1])  # count == 1: divide() leaves the sums alone
    assert one.error_final == 4.0 and one.error_avg == 1.0


def test_two_process_gloo_host_path():
    """world_size 2 over gloo: both ranks shard the same scene, exchange their shard vectors
    and per-frame records through TorchGroup and arrive at the same unsharded result."""
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29631", PYTHONPATH=ROOT)
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "mp_worker.py"), "cpu"],
                              env=dict(env, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
    assert "rank 0 ok" in outs[0] and "rank 1 ok" in outs[1], outs
