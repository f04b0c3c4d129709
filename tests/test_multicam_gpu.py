"""cfg5 shape at test size: several witness cameras seeing the same bundles (per-frame
eliminated block of 6*C animated attributes), unsharded and frame-sharded, with each rank
holding only its own frames (mmb_problem_create_shard_local)."""
import numpy as np
import pytest

from groups import run_ranks
from mayamatchmovesolver_b200 import synth
from mayamatchmovesolver_b200.distributed import solve_global_sharded
from mayamatchmovesolver_b200.problem import Problem, shard_frames
from oraclelib import OracleProblem

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("adt", [0, 1])
def test_multicam_residuals_and_jacobian_match_oracle(adt):
    sb, _ = synth.make_multicam_scene(3, 6, 10, 1)
    o = OracleProblem(sb, auto_diff_type=adt)
    with Problem(sb, auto_diff_type=adt) as p:
        assert (p.n, p.nobs, p.m) == (o.n, o.nobs, o.m) == (3 * 6 * 6 + 30, 180, 360)
        e2m, p2a = p.index_maps()
        oe2m, op2a = o.index_maps()
        assert np.array_equal(e2m, oe2m) and np.array_equal(p2a, op2a)
        x = p.initial_parameters()
        f, d = p.eval_residuals(x)
        fo, do = o.eval_residuals(x)
        assert np.abs(f - fo).max() <= 1e-9 * np.abs(fo).max()
        J, Jo = p.eval_jacobian_dense(x), o.eval_jacobian_dense(x)
        assert np.abs(J - Jo).max() <= 1e-9 * np.abs(Jo).max()


def test_multicam_solve_matches_oracle():
    sb, _ = synth.make_multicam_scene(3, 8, 12, 2, fixed_bundles=4)
    o = OracleProblem(sb)
    ro = o.solve()
    with Problem(sb) as p:
        r = p.solve()
    print(r, ro)
    assert (r.reason_num, r.iteration_num, r.iteration_attempt_num) == (ro.reason_num, ro.iteration_num, ro.iteration_attempt_num)
    scale = np.maximum(np.abs(ro.attr_values), 1.0)
    assert np.max(np.abs(r.attr_values - ro.attr_values) / scale) <= 1e-6
    assert np.max(np.abs(r.error_dist - ro.error_dist)) <= 1e-5


@pytest.mark.parametrize("world", [2, 3])
def test_multicam_local_slices_match_unsharded(world):
    C, F, B, seed = 4, 18, 16, 3
    sb, _ = synth.make_multicam_scene(C, F, B, seed, fixed_bundles=4)
    with Problem(sb) as p:
        r1 = p.solve()

    def body(g):
        f0, f1 = shard_frames(F, g.world, g.rank)
        mine, _ = synth.make_multicam_scene(C, F, B, seed, frame_range=(f0, f1), fixed_bundles=4)
        return solve_global_sharded(mine, g, local_backend=True, local_slice=(f0, F))
    outs = run_ranks(world, body)
    r = outs[0]
    print(world, r1, r)
    assert (r.reason_num, r.iteration_num, r.iteration_attempt_num) == (r1.reason_num, r1.iteration_num, r1.iteration_attempt_num)
    scale = np.maximum(np.abs(r1.attr_values), 1.0)
    assert np.max(np.abs(r.attr_values - r1.attr_values) / scale) <= 1e-6
    assert np.max(np.abs(r.error_dist - r1.error_dist)) <= 1e-5
