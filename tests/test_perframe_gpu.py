"""frameSolveMode = 1: one launch solves every frame; parity against the oracle
run frame by frame the way solve_v1 does (adjust_base.cpp:1430-1484) and its
results summed like CommandResult::add / divide (adjust_results.h:89-117)."""
import numpy as np
import pytest

from mayamatchmovesolver_b200 import Problem
from mayamatchmovesolver_b200.synth import make_scene

from oraclelib import OracleProblem

pytestmark = pytest.mark.gpu

OPTS = dict(iter_max=100, tau=1.0, eps1=1e-6, eps2=1e-6, eps3=1e-6, delta=1e-4, auto_diff_type=0,
            auto_param_scale=0, accept_only_better=1, image_width=2048.0)


def oracle_per_frame(sb, **opts):
    """The reference's per-frame loop, one 1-frame oracle problem at a time."""
    F = sb.F
    per, attrs = [], []
    for f in range(F):
        one = sb.slice_frames([f])
        r = OracleProblem(one, faithful=True, **opts).solve()
        per.append(r)
        attrs.append(np.asarray(r.attr_values))
    agg = dict(success=min(int(r.success) for r in per),
               error_final=sum(r.error_final for r in per) / F,
               error_avg=sum(r.error_avg for r in per) / F,
               error_min=min(r.error_min for r in per), error_max=max(r.error_max for r in per),
               iterations=sum(r.iteration_num for r in per), reason=per[-1].reason_num)
    return agg, per, np.stack(attrs, 1)  # attrs[a, f]


@pytest.mark.parametrize("case", ["camera6_24x20", "camera6_weights_drop", "camera_focal_central", "camera6_lmdif",
                                  "camera6_lmdif_maxfev", "camera3_translate"])
def test_per_frame_parity(case):
    opts = dict(OPTS)
    if case == "camera6_24x20":
        sb, _ = make_scene(24, 20, 5, solve_bundles=False, period=24.0)
    elif case == "camera6_weights_drop":
        sb, _ = make_scene(16, 30, 6, solve_bundles=False, weights=True, drop_fraction=0.2, period=16.0)
    elif case == "camera3_translate":
        # three animated attributes per frame (the other register-resident block size)
        sb, info = make_scene(12, 20, 9, solve_camera=False, solve_bundles=False, period=12.0)
        rng = np.random.default_rng(3)
        for k in ("tx", "ty", "tz"):
            ref = info["cam_tfm"][k]
            for i in range(sb.F):
                sb.set_value(ref, sb.get_value(ref, i) + rng.normal(0, 0.05), i)
            sb.add_attr(ref)
    elif case.startswith("camera6_lmdif"):
        # solverType 1 = cminpack_lmdif: relative forward steps, 6 function calls per Jacobian
        # count towards `iterations` (maxfev); the second case stops on it (reason 5)
        sb, _ = make_scene(10, 20, 8, solve_bundles=False, period=10.0)
        opts["solver_type"] = 1
        opts["iter_max"] = 400 if case == "camera6_lmdif" else 20
    else:
        sb, _ = make_scene(8, 24, 7, solve_bundles=False, period=8.0)
        opts["auto_diff_type"] = 1
    agg, per, attrs = oracle_per_frame(sb, **opts)
    with Problem(sb, frame_solve_mode=1, **opts) as p:
        r, frames = p.solve_per_frame()
        r2 = p.solve()  # mmb_solve routes by options.frame_solve_mode
    A = attrs.shape[0]
    got = r.attr_values.reshape(A, sb.F)  # parameters are attr-major, one per frame
    for f in range(sb.F):
        assert frames[f].reason_num == per[f].reason_num, (f, frames[f], per[f])
        assert frames[f].iteration_num == per[f].iteration_num, (f, frames[f], per[f])
        assert frames[f].iteration_function_num == per[f].iteration_function_num
        assert frames[f].iteration_jacobian_num == per[f].iteration_jacobian_num
        np.testing.assert_allclose(frames[f].error_final, per[f].error_final, rtol=1e-6)
        np.testing.assert_allclose(frames[f].error_avg, per[f].error_avg, rtol=1e-6)
    np.testing.assert_allclose(got, attrs, rtol=1e-6, atol=1e-8)
    assert int(r.success) == agg["success"]
    assert r.iteration_num == agg["iterations"]
    assert r.reason_num == agg["reason"]
    np.testing.assert_allclose(r.error_final, agg["error_final"], rtol=1e-6)
    np.testing.assert_allclose(r.error_avg, agg["error_avg"], rtol=1e-6)
    np.testing.assert_allclose(r.error_min, agg["error_min"], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(r.error_max, agg["error_max"], rtol=1e-6)
    np.testing.assert_allclose(r2.attr_values, r.attr_values, rtol=0, atol=0)


def test_per_frame_refuses_static_attrs():
    from mayamatchmovesolver_b200._capi import MmbError
    sb, _ = make_scene(4, 20, 8, solve_bundles=True, fixed_bundles=3, period=4.0)
    with Problem(sb, frame_solve_mode=1, **OPTS) as p:
        with pytest.raises(MmbError):
            p.solve_per_frame()


def test_per_frame_cfg2_full_size():
    """BASELINE cfg2 (240 frames x 50 bundles, camera only): every frame converges and the
    batched result is independent of how many frames share the launch."""
    sb, _ = make_scene(240, 50, 2, solve_bundles=False)
    with Problem(sb, frame_solve_mode=1, **OPTS) as p:
        r, frames = p.solve_per_frame()
    assert r.success and all(fr.reason_num in (1, 2, 3, 4) for fr in frames)
    assert r.error_avg < 1.0
    sub = sb.slice_frames(np.arange(100, 110))
    with Problem(sub, frame_solve_mode=1, **OPTS) as p:
        r10, _ = p.solve_per_frame()
    np.testing.assert_array_equal(r10.attr_values.reshape(6, 10), r.attr_values.reshape(6, 240)[:, 100:110])
