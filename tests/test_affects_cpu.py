"""CPU: the marker <-> attribute relationship (mmSolverAffects,
src/mmSolver/cmd/MMSolverAffectsCmd.cpp; findMarkerToAttributeRelationship,
adjust_relationships.cpp:369-466) through the C-ABI's host-only mmb_affects, and the
used / unused split of solve_v1 (adjust_base.cpp:574-610, 1322-1395) in the product's host
code (driven through the emulated thread bodies) against the oracle's literal restatement."""
import ctypes as C

import numpy as np
import pytest

from mayamatchmovesolver_b200 import cmd
from mayamatchmovesolver_b200.problem import affects
from mayamatchmovesolver_b200.scene import SceneBuilder
from oraclelib import EmuProblem, OracleProblem, oracle


def _scene():
    """camera A (tx solved) sees bundles 0, 1; camera B (nothing solved) sees bundles 1, 2;
    bundle 0 tx, bundle 1 ty solved; bundle 2 has no solved attribute -> marker (B, 2) is
    affected by nothing; an attribute of a transform nothing hangs under affects no marker."""
    sb = SceneBuilder(frames=[1, 2, 3])
    ta = sb.add_transform(name="camA_tfm", tx=[0.0, 0.1, 0.2], tz=6.0)
    tb = sb.add_transform(name="camB_tfm", tx=2.0, tz=6.5, ry=8.0)
    ca = sb.add_camera(ta, name="camA")
    cb = sb.add_camera(tb, name="camB")
    grp = sb.add_transform(name="grp", tz=-12.0)
    bt = [sb.add_transform(parent=grp, name="bnd%d_tfm" % i, tx=-1.0 + i, ty=0.2 * i) for i in range(3)]
    bn = [sb.add_bundle(t, name="bnd%d" % i) for i, t in enumerate(bt)]
    lonely = sb.add_transform(name="lonely", tx=3.0)
    rng = np.random.default_rng(3)
    for name, c, b in (("mA0", ca, 0), ("mA1", ca, 1), ("mB1", cb, 1), ("mB2", cb, 2)):
        sb.add_marker(c, bn[b], rng.uniform(-0.2, 0.2, 3), rng.uniform(-0.1, 0.1, 3), name=name)
    sb.add_attr(ta["tx"])
    sb.add_attr(bt[0]["tx"])
    sb.add_attr(lonely["tx"])
    sb.add_attr(bt[1]["ty"])
    return sb


def test_affects_table_and_used_flags():
    table, mu, au = affects(_scene())
    #            camA.tx bnd0.tx lonely.tx bnd1.ty
    assert table.astype(int).tolist() == [[1, 1, 0, 0],   # mA0
                                          [1, 0, 0, 1],   # mA1
                                          [0, 0, 0, 1],   # mB1
                                          [0, 0, 0, 0]]   # mB2
    assert mu.tolist() == [True, True, True, False] and au.tolist() == [True, True, False, True]


def test_mmsolveraffects_command_strings():
    sb = _scene()
    out = cmd.mmSolverAffects(sb, mode="returnString")
    assert "marker_affects_attribute=mA0#camA_tfm.tx#1" in out and "marker_affects_attribute=mB2#bnd1_tfm.ty#0" in out
    assert len(out) == 16
    assert cmd.mmSolverAffects(sb, mode="addAttrsToMarkers") is None and sb.affects[("mB1", "bnd1_tfm.ty")] is True
    with pytest.raises(ValueError):
        cmd.mmSolverAffects(sb, mode="nope")


@pytest.mark.parametrize("rum,rua", [(1, 1), (1, 0), (0, 1), (0, 0)])
def test_unused_objects_are_removed_like_the_reference(rum, rua):
    sb = _scene()
    o = OracleProblem(sb, remove_unused_markers=rum, remove_unused_attributes=rua)
    e = EmuProblem(sb, remove_unused_markers=rum, remove_unused_attributes=rua)
    # 3 animated (camA.tx) + statics; markers x 3 frames
    assert o.n == 3 + (2 if rua else 3) and o.nobs == 3 * (3 if rum else 4)
    assert (e.n, e.nobs) == (o.n, o.nobs)
    mu = np.zeros(4, np.uint8); au = np.zeros(4, np.uint8)
    u8 = C.POINTER(C.c_uint8)
    oracle().orc_used_objects(o.h, mu.ctypes.data_as(u8), au.ctypes.data_as(u8))
    assert mu.tolist() == ([1, 1, 1, 0] if rum else [1, 1, 1, 1]) and au.tolist() == ([1, 1, 0, 1] if rua else [1, 1, 1, 1])
    x = o.initial_parameters() + 0.01
    fo, do = o.eval_residuals(x)
    fe, de = e.eval_residuals(x)
    assert np.abs(fo - fe).max() <= 1e-9 * np.abs(fo).max()
    Jo = o.eval_jacobian_dense(x)
    Je, _ = e.eval_jacobian_dense(x)
    assert np.abs(Jo - Je).max() <= 1e-9 * np.abs(Jo).max() and np.array_equal(Jo != 0, Je != 0)
