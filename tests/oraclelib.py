"""ctypes loader for the CPU oracle (oracle/liboracle.so), the reference's own
lens code (oracle/_ref/libmmlens_ref.so) and the test-only host emulation of
the CUDA thread bodies (tests/emu/libemu.so).  TEST INFRASTRUCTURE ONLY.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from mayamatchmovesolver_b200 import _capi
from mayamatchmovesolver_b200._capi import Options, Result, SceneDesc, c_f64p, c_i32p

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")


def _make():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR], stdout=subprocess.DEVNULL)


_oracle = None
_ref = None
_emu = None


def oracle():
    global _oracle
    if _oracle is None:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".h", ".cpp"))]
        if not os.path.exists(path) or os.path.getmtime(path) < max(os.path.getmtime(s) for s in srcs):
            _make()
        lib = C.CDLL(path)
        lib.orc_create.argtypes = [C.POINTER(SceneDesc), C.POINTER(Options), C.POINTER(C.c_void_p)]
        lib.orc_destroy.argtypes = [C.c_void_p]
        lib.orc_sizes.argtypes = [C.c_void_p, c_i32p, c_i32p, c_i32p]
        lib.orc_index_maps.argtypes = [C.c_void_p, c_i32p, c_i32p]
        lib.orc_weights.argtypes = [C.c_void_p, c_f64p]
        lib.orc_used_objects.argtypes = [C.c_void_p, C.POINTER(C.c_uint8), C.POINTER(C.c_uint8)]
        lib.orc_set_faithful.argtypes = [C.c_void_p, C.c_int]
        lib.orc_set_central_literal.argtypes = [C.c_void_p, C.c_int]
        lib.orc_initial_parameters.argtypes = [C.c_void_p, c_f64p]
        lib.orc_eval_residuals.argtypes = [C.c_void_p, c_f64p, c_f64p, c_f64p]
        lib.orc_eval_jacobian_dense.argtypes = [C.c_void_p, c_f64p, c_f64p, c_f64p, C.c_int32, C.c_int32]
        lib.orc_solve.argtypes = [C.c_void_p, C.POINTER(Result), c_f64p, c_f64p, c_f64p, C.c_int32]
        lib.orc_reproject.argtypes = [c_f64p, C.c_int, c_f64p, c_f64p, c_f64p]
        lib.orc_projection_matrix.argtypes = [c_f64p, c_f64p]
        lib.orc_local_matrix.argtypes = [c_f64p, C.c_int, c_f64p]
        lib.orc_lens_apply.argtypes = [C.c_int, c_f64p, c_f64p, C.c_int, C.c_int, c_f64p, c_f64p]
        lib.orc_bound_int_to_ext.argtypes = [C.c_double] * 5
        lib.orc_bound_int_to_ext.restype = C.c_double
        lib.orc_bound_ext_to_int.argtypes = [C.c_double] * 5
        lib.orc_bound_ext_to_int.restype = C.c_double
        _oracle = lib
    return _oracle


def ref_lens():
    """The reference's own mmlens, or None when it was never built (no /root/reference
    and no prebuilt oracle/_ref)."""
    global _ref
    if _ref is None:
        path = os.path.join(ORACLE_DIR, "_ref", "libmmlens_ref.so")
        if not os.path.exists(path) and os.path.isdir("/root/reference"):
            _make()
        if not os.path.exists(path):
            return None
        lib = C.CDLL(path)
        lib.ref_lens_apply.argtypes = [C.c_int, c_f64p, c_f64p, C.c_int, C.c_int, c_f64p, c_f64p]
        _ref = lib
    return _ref


def emu():
    global _emu
    if _emu is None:
        d = os.path.join(ROOT, "tests", "emu")
        path = os.path.join(d, "libemu.so")
        csrc = os.path.join(ROOT, "mayamatchmovesolver_b200", "csrc")
        deps = [os.path.join(d, "emu.cpp")] + [os.path.join(csrc, f) for f in ("problem.cpp", "problem.h", "eval_bodies.cuh", "device_math.cuh")]
        if not os.path.exists(path) or os.path.getmtime(path) < max(os.path.getmtime(s) for s in deps):
            subprocess.check_call(["g++", "-O1", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-o", path,
                                   os.path.join(d, "emu.cpp")])
        lib = C.CDLL(path)
        lib.emu_create.argtypes = [C.POINTER(SceneDesc), C.POINTER(Options), C.POINTER(C.c_void_p)]
        lib.emu_destroy.argtypes = [C.c_void_p]
        lib.emu_sizes.argtypes = [C.c_void_p, c_i32p, c_i32p, c_i32p]
        lib.emu_residuals.argtypes = [C.c_void_p, c_f64p, c_f64p, c_f64p]
        lib.emu_jacobian.argtypes = [C.c_void_p, c_f64p, c_f64p, c_f64p, c_f64p, c_i32p]
        lib.emu_extra.argtypes = [C.c_void_p]
        lib.emu_extra_jacobian.argtypes = [C.c_void_p, c_f64p, c_f64p, c_i32p]
        _emu = lib
    return _emu


def host_options(**overrides):
    """Reference defaults without needing the CUDA library (mirrors mmb_options_default)."""
    o = Options()
    o.iter_max = 100; o.tau = 1.0; o.eps1 = o.eps2 = o.eps3 = 1e-6; o.delta = 1e-4
    o.auto_diff_type = 0; o.auto_param_scale = 1; o.robust_loss_type = 0; o.robust_loss_scale = 1.0
    o.scene_graph_mode = 2; o.solver_type = 2; o.time_eval_mode = 0; o.accept_only_better = 1
    o.image_width = 2048.0; o.frame_solve_mode = 0; o.remove_unused_markers = 1
    o.remove_unused_attributes = 1; o.lens_index_literal = 0; o.device = 0
    for k, v in overrides.items():
        setattr(o, k, v)
    return o


def _p(a, t=c_f64p):
    return a.ctypes.data_as(t)


class OracleProblem:
    def __init__(self, scene, options=None, faithful=False, **ov):
        self.lib = oracle()
        self.options = options if options is not None else host_options(**ov)
        desc, self._keep = scene.desc()
        self.h = C.c_void_p()
        st = self.lib.orc_create(C.byref(desc), C.byref(self.options), C.byref(self.h))
        assert st == 0, st
        n, nobs, m = C.c_int32(), C.c_int32(), C.c_int32()
        self.lib.orc_sizes(self.h, C.byref(n), C.byref(nobs), C.byref(m))
        self.n, self.nobs, self.m = n.value, nobs.value, m.value
        self.lib.orc_set_faithful(self.h, 1 if faithful else 0)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.orc_destroy(self.h)
            self.h = None

    def index_maps(self):
        e2m = np.zeros((self.nobs, 2), np.int32)
        p2a = np.zeros((self.n, 2), np.int32)
        self.lib.orc_index_maps(self.h, _p(e2m, c_i32p), _p(p2a, c_i32p))
        return e2m, p2a

    def marker_weights(self):
        w = np.zeros(self.nobs)
        self.lib.orc_weights(self.h, _p(w))
        return w

    def initial_parameters(self):
        x = np.zeros(self.n)
        self.lib.orc_initial_parameters(self.h, _p(x))
        return x

    def eval_residuals(self, x):
        x = np.ascontiguousarray(x, np.float64)
        f = np.zeros(self.m)
        d = np.zeros(self.nobs)
        self.lib.orc_eval_residuals(self.h, _p(x), _p(f), _p(d))
        return f, d

    def eval_jacobian_dense(self, x, structural_skip=True):
        x = np.ascontiguousarray(x, np.float64)
        f, _ = self.eval_residuals(x)
        ld = max(self.m, self.n)
        J = np.zeros((self.n, ld))
        self.lib.orc_eval_jacobian_dense(self.h, _p(x), _p(f), _p(J), ld, 1 if structural_skip else 0)
        return J[:, :self.m].T.copy()

    def solve_sparse(self, threads=0, max_outer=0):
        """block-sparse CPU solve (oracle/oracle_sparse.h): feasible at full cfg3 / cfg4 size"""
        from mayamatchmovesolver_b200.problem import SolveResult
        r = Result()
        x = np.zeros(self.n); av = np.zeros(self.n); d = np.zeros(self.nobs)
        self.lib.orc_solve_sparse.argtypes = [C.c_void_p, C.POINTER(Result), c_f64p, c_f64p, c_f64p, C.c_int32, C.c_int32]
        st = self.lib.orc_solve_sparse(self.h, C.byref(r), _p(x), _p(av), _p(d), int(threads), int(max_outer))
        res = SolveResult(r, x, av, d)
        res.status = st
        return res

    def solve(self, structural_skip=True):
        from mayamatchmovesolver_b200.problem import SolveResult
        r = Result()
        x = np.zeros(self.n); av = np.zeros(self.n); d = np.zeros(self.nobs)
        st = self.lib.orc_solve(self.h, C.byref(r), _p(x), _p(av), _p(d), 1 if structural_skip else 0)
        res = SolveResult(r, x, av, d)
        res.status = st
        return res


class EmuProblem:
    def __init__(self, scene, options=None, **ov):
        self.lib = emu()
        self.options = options if options is not None else host_options(**ov)
        desc, self._keep = scene.desc()
        self.h = C.c_void_p()
        st = self.lib.emu_create(C.byref(desc), C.byref(self.options), C.byref(self.h))
        assert st == 0, st
        n, nobs, k = C.c_int32(), C.c_int32(), C.c_int32()
        self.lib.emu_sizes(self.h, C.byref(n), C.byref(nobs), C.byref(k))
        self.n, self.nobs, self.kmax = n.value, nobs.value, k.value
        self.n_extra = self.lib.emu_extra(self.h)
        self.m = 2 * self.nobs + self.n_extra

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.emu_destroy(self.h)
            self.h = None

    def eval_residuals(self, x):
        x = np.ascontiguousarray(x, np.float64)
        f = np.zeros(self.m); d = np.zeros(self.nobs)
        self.lib.emu_residuals(self.h, _p(x), _p(f), _p(d))
        return f, d

    def eval_jacobian_dense(self, x):
        x = np.ascontiguousarray(x, np.float64)
        f = np.zeros(self.m); d = np.zeros(self.nobs)
        vals = np.zeros((self.kmax, 2 * self.nobs)); cols = np.zeros((self.nobs, self.kmax), np.int32)
        self.lib.emu_jacobian(self.h, _p(x), _p(f), _p(d), _p(vals), _p(cols, c_i32p))
        J = blocks_to_dense(vals, cols, self.n)
        if self.n_extra:
            jext = np.zeros(self.n_extra); pid = np.zeros(self.n_extra, np.int32)
            self.lib.emu_extra_jacobian(self.h, _p(x), _p(jext), _p(pid, c_i32p))
            rows = np.zeros((self.n_extra, self.n))
            rows[np.arange(self.n_extra), pid] = jext
            J = np.vstack([J, rows])
        return J, f


def blocks_to_dense(vals, cols, n):
    k, m = vals.shape
    nobs = m // 2
    J = np.zeros((m, n))
    for c in range(k):
        pid = cols[:, c]
        ok = pid >= 0
        idx = np.nonzero(ok)[0]
        J[2 * idx, pid[ok]] = vals[c, 2 * idx]
        J[2 * idx + 1, pid[ok]] = vals[c, 2 * idx + 1]
    return J


def oracle_mmsolver(scene, scene_graph_mode=2, **flags):
    """The command surface (cmd.mmSolver's flag handling) driving the CPU oracle instead of the
    CUDA library: -> (SolveResult, {plug: solved value}).  Used to pin the oracle to the
    reference's own Maya-hosted tests (tests/kat_scenes.py)."""
    from mayamatchmovesolver_b200 import cmd
    opts, rest = cmd._options_from_kwargs(flags)
    objects = {k: v for k, v in rest.items() if k in cmd._OBJECT_FLAGS or k in cmd._STIFF_FLAGS}
    opts["scene_graph_mode"] = scene_graph_mode
    sb = cmd.select_scene(scene, **objects) if objects else scene
    o = OracleProblem(sb, **opts)
    r = o.solve()
    _, p2a = o.index_maps()
    values = {}
    plugs = [e[0] for e in objects.get("attr", [])]
    for i in range(o.n):
        a, f = int(p2a[i, 0]), int(p2a[i, 1])
        if a < len(plugs) and f <= 0:
            values[plugs[a]] = float(r.attr_values[i])
    return r, values
