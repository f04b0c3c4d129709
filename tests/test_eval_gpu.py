"""GPU parity: residuals (solveFunc iflag=1) and FD Jacobian (iflag=2) of the
CUDA path, called through the C-ABI, against the CPU oracle on the same seeded
inputs.  Tolerances: residuals 1e-9 relative to the largest residual;
Jacobian entries 1e-9 relative to the largest entry of the matrix (FD
quotients amplify rounding by 1/delta = 1e4).
"""
import numpy as np
import pytest

from mayamatchmovesolver_b200 import synth
from mayamatchmovesolver_b200.problem import Problem, default_options
from oraclelib import OracleProblem, blocks_to_dense

pytestmark = pytest.mark.gpu

CASES = [
    dict(F=7, B=5, seed=11, drop_fraction=0.2, weights=True),
    dict(F=1, B=20, seed=1, solve_camera=False),
    dict(F=24, B=10, seed=2, solve_bundles=False),
    dict(F=30, B=40, seed=3),
    dict(F=70, B=9, seed=12, drop_fraction=0.3, weights=True),   # sparse visibility in the rows kernel (F >= 32)
    dict(F=12, B=16, seed=4, lens=dict(degree2_distortion=0.05, degree2_u=0.002, degree2_v=-0.002, degree4_distortion=0.01),
         solve_focal=True, solve_lens=("degree2_distortion", "degree2_u", "degree2_v", "degree4_distortion"), focal_init=33.0),
    # the other lens models in the loop, coefficients and focal length solved (in_frame: the
    # LDPK inverse only converges inside the image)
    dict(F=6, B=12, seed=6, in_frame=True, lens_model="3de_classic",
         lens=dict(distortion=0.06, anamorphic_squeeze=1.02, curvature_x=0.01, curvature_y=-0.005, quartic_distortion=0.01),
         solve_focal=True, solve_lens=("distortion", "curvature_x", "quartic_distortion")),
    dict(F=6, B=12, seed=7, in_frame=True, lens_model="3de_radial_std_deg4",
         lens=dict(degree2_distortion=0.05, degree2_u=0.002, degree4_distortion=0.01, cylindric_direction=20.0, cylindric_bending=0.05),
         solve_lens=("degree2_distortion", "cylindric_bending")),
    dict(F=6, B=12, seed=8, in_frame=True, lens_model="3de_anamorphic_std_deg4",
         lens=dict(degree2_cx02=0.02, degree2_cy02=0.01, degree2_cx22=-0.01, degree4_cx04=0.01, degree4_cy44=0.002,
                   lens_rotation=3.0, squeeze_x=1.02, squeeze_y=0.99),
         solve_focal=True, solve_lens=("degree2_cx02", "degree2_cy02", "lens_rotation", "squeeze_x")),
    dict(F=6, B=12, seed=9, in_frame=True, lens_model="3de_anamorphic_std_deg4_rescaled",
         lens=dict(degree2_cx02=0.02, degree2_cy02=0.01, degree4_cx24=0.004, lens_rotation=-2.0, squeeze_x=1.01,
                   squeeze_y=1.0, rescale=1.05),
         solve_lens=("degree2_cx02", "rescale")),
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("adt", [0, 1])
@pytest.mark.parametrize("path", ["generic", "tiled"])
def test_residuals_and_jacobian_match_oracle(case, adt, path, monkeypatch):
    # both Jacobian kernels: the generic per-observation one and the register-tiled fast
    # path (which only engages for the shapes it supports; others fall back to generic)
    monkeypatch.setenv("MMB_JACOBIAN_PATH", path)
    sb, _ = synth.make_scene(**case)
    o = OracleProblem(sb, auto_diff_type=adt)
    with Problem(sb, auto_diff_type=adt) as p:
        assert (p.n, p.nobs, p.m) == (o.n, o.nobs, o.m)
        e2m, p2a = p.index_maps()
        oe2m, op2a = o.index_maps()
        assert np.array_equal(e2m, oe2m) and np.array_equal(p2a, op2a)  # bit-exact indexing
        assert np.array_equal(p.marker_weights(), o.marker_weights())
        x = p.initial_parameters()
        assert np.array_equal(x, o.initial_parameters())
        rng = np.random.default_rng(5)
        for trial in range(2):
            xt = x + (rng.normal(0, 1e-2, x.shape) if trial else 0.0)
            f, d = p.eval_residuals(xt)
            fo, do = o.eval_residuals(xt)
            scale = np.abs(fo).max()
            assert np.abs(f - fo).max() <= 1e-9 * scale
            assert np.abs(d - do).max() <= 1e-9 * np.abs(do).max()
            J = p.eval_jacobian_dense(xt)
            Jo = o.eval_jacobian_dense(xt)
            assert np.array_equal(J != 0, Jo != 0) or np.abs(J - Jo).max() <= 1e-9 * np.abs(Jo).max()
            assert np.abs(J - Jo).max() <= 1e-9 * np.abs(Jo).max()
            vals, cols = p.eval_jacobian_blocks(xt)
            assert np.array_equal(blocks_to_dense(vals, cols, p.n), J)
            # values-only call: the kernel instance the LM loop and bench use (no column ids)
            # (different schedule and FMA contraction: equal to rounding, amplified by 1/delta)
            fast = p.eval_jacobian_blocks(xt, want_cols=False)
            assert np.abs(fast - vals).max() <= 1e-9 * np.abs(Jo).max()


def test_full_size_properties():
    """cfg3 shape (1000 x 500): size-independent checks."""
    sb, info = synth.make_scene(1000, 500, 3, noise_px=0.0, init_noise_t=0.0, init_noise_r=0.0, init_noise_b=0.0)
    with Problem(sb) as p:
        assert p.nobs == 500000 and p.n == 7500
        x = p.initial_parameters()
        f, d = p.eval_residuals(x)
        # noise-free markers at the truth: zero reprojection error
        assert np.abs(d).max() < 1e-7
        vals, cols = p.eval_jacobian_blocks(x)
        assert vals.shape == (9, 1000000)
        # every observation touches its frame's 6 camera params and its bundle's 3
        e2m, p2a = p.index_maps()
        fr = e2m[:, 1]
        assert np.array_equal(cols[:, :6], (np.arange(6)[None, :] * 1000 + fr[:, None]))
        assert np.array_equal(cols[:, 6:], 6000 + 3 * e2m[:, :1] + np.arange(3)[None, :])
        assert np.isfinite(vals).all() and np.abs(vals).max() > 1.0


BOUNDED_CASES = [
    dict(F=7, B=9, seed=11, drop_fraction=0.2, weights=True),                  # generic kernel, sparse visibility
    dict(F=40, B=30, seed=3),                                                  # F >= 32: the rows kernel
    dict(F=36, B=18, seed=4, in_frame=True,                                    # lens in the loop: the lens-tile kernel
         lens=dict(degree2_distortion=0.05, degree2_u=0.002, degree2_v=-0.002, degree4_distortion=0.01),
         solve_focal=True, solve_lens=("degree2_distortion", "degree4_distortion"), focal_init=33.0),
]


@pytest.mark.parametrize("case", BOUNDED_CASES)
@pytest.mark.parametrize("adt", [0, 1])
@pytest.mark.parametrize("path", ["generic", "tiled"])
def test_bounded_offset_scaled_attrs_match_oracle(case, adt, path, monkeypatch):
    """`-attr node.attr min max offset scale` (arg_flags_solve_object.cpp:288-318): every branch of
    parameterBoundFromInternalToExternal (adjust_base.cpp:194-258) on the device -- both bounds,
    upper only, lower only, offset / scale -- and FD steps whose sign calculateParameterDelta flips
    (adjust_solveFunc.cpp:148-180), in every Jacobian kernel."""
    from kat_scenes import apply_bounds
    monkeypatch.setenv("MMB_JACOBIAN_PATH", path)
    sb, info = synth.make_scene(**case)
    apply_bounds(sb, info)
    o = OracleProblem(sb, auto_diff_type=adt)
    with Problem(sb, auto_diff_type=adt) as p:
        x = p.initial_parameters()
        assert np.array_equal(x, o.initial_parameters())
        _, p2a = o.index_maps()
        flips = ((x + 1e-4) > np.array(sb.attr_max)[p2a[:, 0]]).sum()
        assert flips > 0  # the scene does exercise the sign flip
        rng = np.random.default_rng(7)
        for trial in range(2):
            xt = x + (rng.normal(0, 1e-2, x.shape) if trial else 0.0)
            f, d = p.eval_residuals(xt)
            fo, do = o.eval_residuals(xt)
            assert np.abs(f - fo).max() <= 1e-9 * np.abs(fo).max()
            assert np.abs(d - do).max() <= 1e-9 * np.abs(do).max()
            J = p.eval_jacobian_dense(xt)
            Jo = o.eval_jacobian_dense(xt)
            assert np.abs(J - Jo).max() <= 1e-9 * np.abs(Jo).max()
            fast = p.eval_jacobian_blocks(xt, want_cols=False)
            vals, cols = p.eval_jacobian_blocks(xt)
            assert np.abs(fast - vals).max() <= 1e-9 * np.abs(Jo).max()


def test_lens_index_literal_matches_oracle():
    """options.lens_index_literal = 1: the reference's literal `markerIndex + frameIndex` read of
    the marker-frame lens table (adjust_measureErrors.cpp:463) -- with two cameras of which only
    one has a lens, the literal index picks another marker's lens for most observations."""
    sb = _two_camera_lens_scene()
    for literal in (0, 1):
        o = OracleProblem(sb, lens_index_literal=literal)
        with Problem(sb, lens_index_literal=literal) as p:
            x = p.initial_parameters()
            f, d = p.eval_residuals(x)
            fo, do = o.eval_residuals(x)
            assert np.abs(f - fo).max() <= 1e-9 * np.abs(fo).max()
            J = p.eval_jacobian_dense(x)
            Jo = o.eval_jacobian_dense(x)
            assert np.abs(J - Jo).max() <= 1e-9 * np.abs(Jo).max()
            if literal == 0:
                f0 = f
    assert np.abs(f - f0).max() > 1e-3  # the two indexings do differ on this scene


def _two_camera_lens_scene():
    from mayamatchmovesolver_b200.scene import SceneBuilder
    rng = np.random.default_rng(21)
    F, B = 6, 5
    sb = SceneBuilder(frames=np.arange(1, F + 1))
    lens = sb.add_lens("3de_radial_std_deg4", camera_params=(3.5, 3.6, 2.4, 1.0, 0.0, 0.0), degree2_distortion=0.08,
                       degree4_distortion=0.02)
    cams = []
    for c in range(2):
        t = sb.add_transform(tx=np.linspace(-1, 1, F) + c, ty=0.3 * c, tz=np.linspace(5, 6, F), ry=np.linspace(-3, 3, F))
        cams.append((t, sb.add_camera(t, lens=lens if c == 0 else None)))
    pts = np.stack([rng.uniform(-3, 3, B), rng.uniform(-2, 2, B), rng.uniform(-20, -12, B)], 1)
    bt = [sb.add_transform(tx=p[0], ty=p[1], tz=p[2]) for p in pts]
    bn = [sb.add_bundle(t) for t in bt]
    for c in range(2):
        for b in range(B):
            sb.add_marker(cams[c][1], bn[b], rng.uniform(-0.3, 0.3, F), rng.uniform(-0.2, 0.2, F))
    for t, _ in cams:
        sb.add_attr(t["tx"]); sb.add_attr(t["ry"])
    sb.add_attr(lens["degree2_distortion"])
    for t in bt:
        sb.add_attr(t["tx"]); sb.add_attr(t["tz"])
    return sb
