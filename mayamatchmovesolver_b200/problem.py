"""Python handle over the C-ABI (`mmb_problem`): what the reference's Python
API reaches through `maya.cmds.mmSolver` (solverstep.py:505-738), minus Maya.
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import Options, Result, c_f64p, c_i32p, check


def default_options(**overrides):
    lib = _capi.load()
    o = Options()
    lib.mmb_options_default(C.byref(o))
    for k, v in overrides.items():
        if not hasattr(o, k):
            raise TypeError("unknown option %r" % k)
        setattr(o, k, v)
    return o


def shard_frames(num_frames, world, rank):
    lib = _capi.load()
    f0, f1 = C.c_int32(), C.c_int32()
    check(lib.mmb_shard_frames(int(num_frames), int(world), int(rank), C.byref(f0), C.byref(f1)), "mmb_shard_frames")
    return f0.value, f1.value


def shard_param_map(scene, world, rank):
    """unsharded parameter id of every parameter of rank's shard (host only, no GPU)"""
    lib = _capi.load()
    desc, keep = scene.desc()
    n = C.c_int32()
    check(lib.mmb_shard_param_map(C.byref(desc), int(world), int(rank), C.byref(n), None), "mmb_shard_param_map")
    out = np.zeros(n.value, np.int32)
    check(lib.mmb_shard_param_map(C.byref(desc), int(world), int(rank), C.byref(n), out.ctypes.data_as(c_i32p)),
          "mmb_shard_param_map")
    return out


def local_slice_param_map(scene, frame_begin, frames_global):
    """the same map for a scene that already is frames [frame_begin, frame_begin + F) of a
    frames_global-frame scene (mmb_problem_create_shard_local; shard.cpp shard_param_map)"""
    out, gbase = [], 0
    for ref in scene.arrays()["attr_ref"]:
        if ref < 0:   # animated: one parameter per frame, frame-minor
            out.append(gbase + frame_begin + np.arange(scene.F, dtype=np.int64))
            gbase += frames_global
        else:
            out.append(np.array([gbase], dtype=np.int64))
            gbase += 1
    return np.concatenate(out).astype(np.int32) if out else np.zeros(0, np.int32)


def comm_unique_id(local_backend=False):
    """128 bytes: an ncclUniqueId (or the id of an in-process group for single-GPU tests)"""
    lib = _capi.load()
    buf = (C.c_uint8 * _capi.MMB_COMM_ID_BYTES)()
    check(lib.mmb_comm_unique_id(buf, 1 if local_backend else 0), "mmb_comm_unique_id")
    return bytes(buf)


def fp64_peaks(device=0):
    """-> (DFMA TFLOP/s, DMMA TFLOP/s) measured on `device`"""
    lib = _capi.load()
    a, b = C.c_double(), C.c_double()
    check(lib.mmb_bench_fp64_peaks(int(device), C.byref(a), C.byref(b)), "mmb_bench_fp64_peaks")
    return a.value, b.value


def stream_floor(read_bytes, write_bytes, reps=30, device=0):
    """-> median ms of a plain streaming kernel of that size (mmb_bench_stream_floor)"""
    lib = _capi.load()
    ms = C.c_double()
    check(lib.mmb_bench_stream_floor(int(device), int(read_bytes), int(write_bytes), int(reps), C.byref(ms)),
          "mmb_bench_stream_floor")
    return ms.value


class Comm:
    """A communicator that outlives Problem handles (mmb_comm_create: collective over the
    ranks of `comm_id`; NCCL set-up and warm-up are paid once)."""

    def __init__(self, comm_id, rank, world, device=0):
        self._lib = _capi.load()
        self._h = C.c_void_p()
        buf = (C.c_uint8 * _capi.MMB_COMM_ID_BYTES).from_buffer_copy(bytes(comm_id))
        check(self._lib.mmb_comm_create(buf, int(rank), int(world), int(device), C.byref(self._h)), "mmb_comm_create")
        self.rank, self.world, self.device = int(rank), int(world), int(device)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.mmb_comm_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def aggregate_results(results):
    """CommandResult::add / divide over per-frame mmb_result records"""
    lib = _capi.load()
    arr = (Result * len(results))(*results)
    out = Result()
    check(lib.mmb_aggregate_results(arr, len(results), C.byref(out)), "mmb_aggregate_results")
    return out


class SolveResult:
    """Field names follow SolverResult / ErrorMetricsResult (adjust_results.h:119-286)."""

    def __init__(self, r, x, attr_values, error_dist):
        self.success = bool(r.success)
        self.reason_num = r.reason_num
        self.reason_string = (r.reason_str or b"").decode()
        self.error_final = r.error_final
        self.error_avg, self.error_min, self.error_max = r.error_avg, r.error_min, r.error_max
        self.iteration_num = r.iteration_num
        self.iteration_function_num = r.iteration_function_num
        self.iteration_jacobian_num = r.iteration_jacobian_num
        self.iteration_attempt_num = r.iteration_attempt_num
        self.user_interrupted = bool(r.user_interrupted)
        self.error_is_better = bool(r.error_is_better)
        self.timer_solve_s = r.timer_solve_s
        self.timer_jacobian_s = r.timer_jacobian_s
        self.timer_function_s = r.timer_function_s
        self.timer_linear_s = r.timer_linear_s
        self.x = x
        self.attr_values = attr_values
        self.error_dist = error_dist
        self.raw = r

    def __repr__(self):
        return ("SolveResult(success=%s reason=%d nfev=%d err_final=%.6g avg=%.6g max=%.6g)" %
                (self.success, self.reason_num, self.iteration_num, self.error_final, self.error_avg,
                 self.error_max))


def _copy_result(r):
    o = Result()
    C.memmove(C.byref(o), C.byref(r), C.sizeof(Result))
    return o


class PerFrameResults:
    """The per-frame records of mmb_solve_per_frame as a read-only sequence of SolveResult;
    entries are built when asked for (a large batch has tens of thousands of frames)."""

    def __init__(self, raw):
        self._raw = raw

    def __len__(self):
        return len(self._raw)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError(i)
        return SolveResult(_copy_result(self._raw[i]), None, None, None)

    def __iter__(self):
        return (self[i] for i in range(len(self)))


class Problem:
    def __init__(self, scene, options=None, shard=None, local_slice=None, **opt_overrides):
        """shard = (rank, world): the handle covers this rank's frame range of `scene`
        (mmb_problem_create_shard); parameters and observations are then numbered locally.
        local_slice = (frame_begin, frames_global): `scene` already IS this rank's frame range
        of a larger scene (mmb_problem_create_shard_local)."""
        self._lib = _capi.load()
        self.options = options if options is not None else default_options(**opt_overrides)
        desc, keep = scene.desc()
        self.num_frames = int(desc.num_frames)
        self.frame_begin, self.shard = 0, None
        self._h = C.c_void_p()
        if local_slice is not None:
            f0, fg = int(local_slice[0]), int(local_slice[1])
            check(self._lib.mmb_problem_create_shard_local(C.byref(desc), C.byref(self.options), f0, fg,
                                                           C.byref(self._h)), "mmb_problem_create_shard_local")
            self.shard = shard
            self.frame_begin = f0
            self.param_map = local_slice_param_map(scene, f0, fg)
        elif shard is None:
            check(self._lib.mmb_problem_create(C.byref(desc), C.byref(self.options), C.byref(self._h)),
                  "mmb_problem_create")
        else:
            rank, world = int(shard[0]), int(shard[1])
            f0, f1 = shard_frames(self.num_frames, world, rank)
            check(self._lib.mmb_problem_create_shard(C.byref(desc), C.byref(self.options), rank, world,
                                                     C.byref(self._h)), "mmb_problem_create_shard")
            self.shard = (rank, world)
            self.frame_begin, self.num_frames = f0, f1 - f0
            self.param_map = shard_param_map(scene, world, rank)
        n, nobs, m = C.c_int32(), C.c_int32(), C.c_int32()
        check(self._lib.mmb_problem_sizes(self._h, C.byref(n), C.byref(nobs), C.byref(m)), "mmb_problem_sizes")
        self.n, self.nobs, self.m = n.value, nobs.value, m.value

    def close(self):
        if getattr(self, "_h", None):
            self._lib.mmb_problem_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def index_maps(self):
        e2m = np.zeros((self.nobs, 2), np.int32)
        p2a = np.zeros((self.n, 2), np.int32)
        check(self._lib.mmb_get_index_maps(self._h, e2m.ctypes.data_as(c_i32p), p2a.ctypes.data_as(c_i32p)),
              "mmb_get_index_maps")
        return e2m, p2a

    def used_objects(self):
        """-> (used_marker_index, used_attr_index): position in the scene's marker / attr lists of
        every marker / attribute that takes part (removeUnusedMarkers / removeUnusedAttributes,
        adjust_base.cpp:1322-1395); index_maps() numbers these used lists"""
        nm, na = C.c_int32(), C.c_int32()
        check(self._lib.mmb_problem_used_objects(self._h, C.byref(nm), None, C.byref(na), None), "mmb_problem_used_objects")
        mi = np.zeros(nm.value, np.int32)
        ai = np.zeros(na.value, np.int32)
        check(self._lib.mmb_problem_used_objects(self._h, C.byref(nm), mi.ctypes.data_as(c_i32p), C.byref(na),
                                                 ai.ctypes.data_as(c_i32p)), "mmb_problem_used_objects")
        return mi, ai

    def marker_weights(self):
        w = np.zeros(self.nobs)
        check(self._lib.mmb_get_marker_weights(self._h, w.ctypes.data_as(c_f64p)), "mmb_get_marker_weights")
        return w

    def initial_parameters(self):
        x = np.zeros(self.n)
        check(self._lib.mmb_get_initial_parameters(self._h, x.ctypes.data_as(c_f64p)), "mmb_get_initial_parameters")
        return x

    def eval_residuals(self, x=None):
        """x = None: the scene at its own attribute values (no parameter applied)"""
        x = None if x is None else np.ascontiguousarray(x, dtype=np.float64)
        fvec = np.zeros(self.m)
        dist = np.zeros(self.nobs)
        check(self._lib.mmb_eval_residuals(self._h, None if x is None else x.ctypes.data_as(c_f64p), fvec.ctypes.data_as(c_f64p),
                                           dist.ctypes.data_as(c_f64p)), "mmb_eval_residuals")
        return fvec, dist

    def eval_jacobian_dense(self, x):
        """column-major fjac as cminpack receives it, returned as (m, n) array"""
        x = np.ascontiguousarray(x, dtype=np.float64)
        ld = max(self.m, 1)
        fjac = np.zeros((self.n, ld))
        check(self._lib.mmb_eval_jacobian_dense(self._h, x.ctypes.data_as(c_f64p), fjac.ctypes.data_as(c_f64p), ld),
              "mmb_eval_jacobian_dense")
        return fjac.T

    def jacobian_kmax(self):
        k = C.c_int32()
        check(self._lib.mmb_eval_jacobian_blocks(self._h, None, None, None, C.byref(k)), "mmb_eval_jacobian_blocks")
        return k.value

    def eval_jacobian_blocks(self, x, want_cols=True):
        x = np.ascontiguousarray(x, dtype=np.float64)
        k = self.jacobian_kmax()
        vals = np.zeros((k, 2 * self.nobs))   # the marker rows; stiffness / smoothness rows: eval_jacobian_dense
        cols = np.zeros((self.nobs, k), np.int32) if want_cols else None
        kk = C.c_int32()
        check(self._lib.mmb_eval_jacobian_blocks(self._h, x.ctypes.data_as(c_f64p), vals.ctypes.data_as(c_f64p),
                                                 cols.ctypes.data_as(c_i32p) if want_cols else None, C.byref(kk)),
              "mmb_eval_jacobian_blocks")
        return (vals, cols) if want_cols else vals

    def solve(self):
        r = Result()
        x = np.zeros(self.n)
        av = np.zeros(self.n)
        dist = np.zeros(self.nobs)
        check(self._lib.mmb_solve(self._h, C.byref(r), x.ctypes.data_as(c_f64p), av.ctypes.data_as(c_f64p),
                                  dist.ctypes.data_as(c_f64p)), "mmb_solve")
        return SolveResult(r, x, av, dist)

    def last_residuals(self):
        """fvec of the accepted point of the last solve (errorList, adjust_results.h:298-311)"""
        fvec = np.zeros(self.m)
        check(self._lib.mmb_get_last_residuals(self._h, fvec.ctypes.data_as(c_f64p)), "mmb_get_last_residuals")
        return fvec

    def solve_per_frame(self):
        """frameSolveMode=1: -> (aggregate SolveResult, [per-frame SolveResult])"""
        r = Result()
        F = self.num_frames
        per = (Result * F)()
        x = np.zeros(self.n)
        av = np.zeros(self.n)
        dist = np.zeros(self.nobs)
        check(self._lib.mmb_solve_per_frame(self._h, C.byref(r), per, x.ctypes.data_as(c_f64p),
                                            av.ctypes.data_as(c_f64p), dist.ctypes.data_as(c_f64p)),
              "mmb_solve_per_frame")
        return SolveResult(r, x, av, dist), PerFrameResults(per)

    def attach_comm(self, comm_id, rank, world):
        """collective over the ranks of `comm_id` (bytes from comm_unique_id): mmb_solve on this
        handle becomes the sharded global solve"""
        buf = (C.c_uint8 * _capi.MMB_COMM_ID_BYTES).from_buffer_copy(bytes(comm_id))
        check(self._lib.mmb_problem_attach_comm(self._h, buf, int(rank), int(world)), "mmb_problem_attach_comm")

    def use_comm(self, comm):
        """lend a long-lived Comm to this handle (it must outlive the handle's solves)"""
        check(self._lib.mmb_problem_use_comm(self._h, comm._h), "mmb_problem_use_comm")
        self._comm_keepalive = comm

    def request_cancel(self):
        self._lib.mmb_request_cancel(self._h)

    def bench_eval(self, reps=10, flush_l2=True):
        """-> (ms per whole evaluation, ms of the residual/Jacobian kernel alone, launches)"""
        ms = C.c_double()
        msj = C.c_double()
        launches = C.c_int32()
        check(self._lib.mmb_bench_eval(self._h, reps, 1 if flush_l2 else 0, C.byref(ms), C.byref(msj),
                                       C.byref(launches)), "mmb_bench_eval")
        return ms.value, msj.value, launches.value

    def launch_count(self):
        return int(self._lib.mmb_launch_count(self._h))

    def solve_stats(self):
        """-> (host synchronisations, CUDA graph launches) of the last solve on this handle"""
        a, b = C.c_int64(), C.c_int64()
        check(self._lib.mmb_solve_stats(self._h, C.byref(a), C.byref(b)), "mmb_solve_stats")
        return a.value, b.value


def affects(scene):
    """findMarkerToAttributeRelationship (adjust_relationships.cpp:369-466) of a scene:
    -> (marker_affects_attr [M, A] bool, marker_used [M] bool, attr_used [A] bool).  Host only."""
    lib = _capi.load()
    desc, keep = scene.desc()
    M, A = desc.num_markers, desc.num_attrs
    table = np.zeros((M, max(A, 1)), np.uint8)
    mu = np.zeros(M, np.uint8)
    au = np.zeros(max(A, 1), np.uint8)
    u8 = C.POINTER(C.c_uint8)
    check(lib.mmb_affects(C.byref(desc), table.ctypes.data_as(u8) if A else None, mu.ctypes.data_as(u8),
                          au.ctypes.data_as(u8)), "mmb_affects")
    return table[:, :A].astype(bool), mu.astype(bool), au[:A].astype(bool)
