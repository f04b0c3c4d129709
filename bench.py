#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 mmSolver solve hot path.

Metric (BASELINE.json): marker-frame residual+FD-Jacobian evals/s.
One *eval* = one observation (marker x frame) producing its 2 residuals and its
whole block-Jacobian row pair (all 1+k perturbed re-evaluations count as that
one eval).

A *step* = one residual+FD-Jacobian evaluation of the whole workload scene
(setParameters + entity variants + the residual/Jacobian kernel), inputs
resident in HBM.  L2 is flushed (256 MiB memset, outside the event brackets)
before every step.  `value` = evals of all ranks / max-over-ranks device time.
`e2e` = the same metric through the public C-ABI solve call (mmb_problem_create
+ mmb_solve + destroy) with HOST buffers: evals done by the solve / wall time.
Extra keys report LM iterations/s and solved frames/s of the global solve and
of the per-frame (batched) solve.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = "cfg3: global camera+bundle solve, 1000 frames x 500 bundles per GPU (n=7500, 500000 observations per GPU)"
CPU_SAMPLE = dict(F=40, B=40, seed=3)   # bounded sample of the same workload for the CPU legs


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            return json.load(fh), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self._halt = threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [t.strip() for t in out.strip().split(",")]
                if len(parts) >= 6:
                    self.samples.append(parts)
            except Exception:
                pass
            self._halt.wait(0.1)

    def stop(self):
        self._halt.set()
        self.join(timeout=5)
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        reasons = set()
        for s in self.samples:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.samples)}


def cpu_reference_leg(steps=1, warmup=0):
    """The reference's own CPU cost structure for this path, restated (oracle/):
    one whole-scene evaluation per Jacobian column, dense column-major J, dense
    pivoted-QR lmder, single thread (the reference has no threads under
    src/mmSolver/adjust).  Bounded sample of the workload; returns evals/s."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oraclelib import OracleProblem
    from mayamatchmovesolver_b200 import synth
    sb, _ = synth.make_scene(fixed_bundles=3, **CPU_SAMPLE)
    times, evals, iters = [], 0, 0
    for i in range(warmup + steps):
        o = OracleProblem(sb, faithful=True, iter_max=12)
        t0 = time.perf_counter()
        r = o.solve(structural_skip=False)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
            evals += o.nobs * r.iteration_attempt_num
            iters += r.iteration_attempt_num
    total = sum(times)
    sample = ("scaled cfg3 (F=%d, B=%d: n=%d, m=%d), oracle/ faithful dense path, %d LM iterations per solve"
              % (CPU_SAMPLE["F"], CPU_SAMPLE["B"], o.n, o.m, iters // max(1, steps)))
    return dict(value=evals / total, seconds=total, lm_iters_per_s=iters / total,
                frames_per_s=CPU_SAMPLE["F"] * steps / total, sample=sample, ms_per_step=1e3 * total / steps)


def cpu_block_sparse_leg(max_outer=2):
    """BASELINE.md "B2": the block-sparse CPU restatement (oracle/oracle_sparse.h: the same
    evaluation arithmetic and MINPACK control law, Jacobian columns only where they live,
    Schur complement instead of the dense QR) on FULL cfg3, all host cores (OpenMP), bounded
    to `max_outer` LM iterations.  Not the reference's cost structure -- what a CPU can do
    with the structure this library exploits."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oraclelib import OracleProblem
    from mayamatchmovesolver_b200 import synth
    sb, _ = synth.make_scene(1000, 500, 200, noise_px=0.5)
    o = OracleProblem(sb, iter_max=20)
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    r = o.solve_sparse(threads=cores, max_outer=max_outer)
    dt = time.perf_counter() - t0
    return dict(value=o.nobs * r.iteration_attempt_num / dt, unit="evals/s", cores=cores, kind="port", seconds=dt,
                lm_iters_per_s=r.iteration_attempt_num / dt,
                sample="full cfg3 (F=1000, B=500: n=7500, m=1000000), oracle/ block-sparse path, first %d LM iterations"
                       % r.iteration_attempt_num)


def cpu_block_sparse_steps(steps, warmup):
    """`steps` timed steps of the block-sparse CPU restatement on FULL cfg3, all host cores: one
    step = ONE LM iteration (residuals + FD block Jacobian + normal equations + lmpar trials +
    trial evaluation) of the cfg3 solve from its start point.  -> per-step seconds, evals."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oraclelib import OracleProblem
    from mayamatchmovesolver_b200 import synth
    sb, _ = synth.make_scene(1000, 500, 200, noise_px=0.5)
    cores = os.cpu_count() or 1
    o = OracleProblem(sb, iter_max=20)
    times = []
    budget_s = 150.0   # the whole arm must end within a few minutes whatever --steps says
    i = 0
    while i < warmup + steps:
        t0 = time.perf_counter()
        r = o.solve_sparse(threads=cores, max_outer=1)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
            if sum(times) + dt > budget_s:
                break
        i += 1
    return times, o.nobs * max(1, r.iteration_attempt_num), cores, o


def run_reference(args):
    """The reference arm: the reference's CPU algorithm for this path (oracle/ restatement; the
    plugin itself needs Maya + Rust and cannot be built here) on the box's host cores, on the
    GPU arm's config (full cfg3).  The reference's own dense cost structure cannot hold cfg3
    (J = 60 GB, twice), so the same-config number is the block-sparse restatement on all cores;
    the faithful dense single-thread path is timed on a scaled cfg3 and reported beside it."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    times, evals_per_step, cores, o = cpu_block_sparse_steps(max(1, args.steps), min(args.warmup, 1))
    steps = len(times)   # the steps that really ran (bounded by the time budget)
    total = sum(times)
    value = evals_per_step * steps / total
    sample = ("full cfg3 (F=1000, B=500: n=%d, m=%d), oracle/ block-sparse restatement, OpenMP on %d cores; "
              "one step = one LM iteration from the start point" % (o.n, o.m, cores))
    dense = cpu_reference_leg(steps=1)
    line = {
        "impl": "reference", "metric": "marker-frame residual+FD-Jacobian evals/s", "value": value,
        "unit": "evals/s", "n_gpus": args.gpus, "steps": steps, "warmup": min(args.warmup, 1),
        "ms_per_step": 1e3 * total / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": value, "unit": "evals/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "lm_iters_per_s": steps / total, "solved_frames_per_s": None,
        "faithful_dense_path": {"what": "the reference's own cost structure (one whole-scene evaluation per Jacobian column, "
                                        "dense J, dense pivoted-QR lmder, 1 thread); infeasible at full cfg3 (dense J = 60 GB)",
                                "value": dense["value"], "unit": "evals/s", "cores": 1, "sample": dense["sample"],
                                "seconds": dense["seconds"]},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def run_own(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from mayamatchmovesolver_b200 import synth
    from mayamatchmovesolver_b200.distributed import SingleGroup, TorchGroup
    from mayamatchmovesolver_b200.distributed import make_comm
    from mayamatchmovesolver_b200.problem import Problem

    group = TorchGroup() if dist is not None else SingleGroup()
    # Weak scaling: ONE scene of 1000 x N frames, frame-range shards of 1000 frames per GPU
    # (SURVEY.md 8e).  Every rank builds the same scene from the same seed.
    F1, B = 1000, 500
    F = F1 * world

    def rank_scene(seed, **kw):
        """this rank's 1000 frames of the 1000 x N-frame scene (same bundles on every rank, the
        camera path continues across ranks): no rank generates or ingests another rank's frames"""
        return synth.make_scene(F1, B, seed, frame_offset=rank * F1, **kw)[0]

    def rank_problem(scene, **kw):
        if world > 1:
            return Problem(scene, device=local, local_slice=(rank * F1, F), **kw)   # mmb_problem_create_shard_local
        return Problem(scene, device=local, **kw)

    sb = rank_scene(3)
    p = rank_problem(sb)
    k = p.jacobian_kmax()
    bytes_per_obs = 48 + 16 * k  # marker 24 B + residuals 16 B + distance 8 B + 2*k*8 B block J

    def barrier():
        if dist is not None:
            dist.barrier()

    def allmax(vals):
        if dist is None:
            return [float(v) for v in vals]
        import torch
        t = torch.tensor(list(vals), device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.cpu()]

    def allsum(vals):
        if dist is None:
            return [float(v) for v in vals]
        import torch
        t = torch.tensor(list(vals), device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return [float(v) for v in t.cpu()]

    sampler = ClockSampler(local)
    for _ in range(args.warmup):
        p.bench_eval(1, True)
    barrier()
    sampler.start()
    launches = 0
    tot_ms = jac_ms = 0.0
    t_wall0 = time.perf_counter()
    for _ in range(args.steps):
        ms, msj, nl = p.bench_eval(1, True)
        tot_ms += ms
        jac_ms += msj
        launches += nl  # kernels inside the step's event bracket (the kernel-only bracket is separate)
    t_wall = time.perf_counter() - t_wall0
    barrier()
    clocks = sampler.stop()
    tot_ms, jac_ms = allmax([tot_ms, jac_ms])
    launches, evals_per_step = allsum([launches, p.nobs])
    value = evals_per_step * args.steps / (tot_ms * 1e-3)
    ms_per_step = tot_ms / args.steps
    nobs_rank, n_rank = p.nobs, p.n
    p.close()

    # ---- end to end through the public solve call, host buffers ----------
    # N = 1: mmb_problem_create + mmb_solve + destroy (one CUDA graph launch per solve).
    # N > 1: the sharded global solve of the N x 1000-frame scene: create_shard_local (every rank
    # holds only its 1000 frames) + use_comm (NCCL) + mmb_solve with one all-reduce of the packed
    # lower triangle of the reduced system per lmpar trial + destroy.
    E2E_REPS = 3
    comm = make_comm(group, device=local)  # NCCL communicator of the session, made once

    def e2e_once(seed):
        sb2 = rank_scene(seed, noise_px=0.5)
        arrays = sb2.arrays()
        h2d = sum(a.nbytes for a in arrays.values())
        barrier()
        t0 = time.perf_counter()
        q = rank_problem(sb2, iter_max=20)
        with q:
            if world > 1:
                q.use_comm(comm)
            r = q.solve()
            d2h = 8 * (2 * q.n + q.nobs) + 160
            nobs = q.nobs
            stats = q.solve_stats()
            solve_launches = q.launch_count()
        dt = time.perf_counter() - t0
        return r, dt, h2d, d2h, nobs, stats, solve_launches
    e2e_once(100)  # warm-up (module load, memory pool, NCCL communicator set-up)
    e2e_dt = e2e_solve = 0.0
    e2e_iters = 0
    for rep in range(E2E_REPS):
        r, dt, h2d, d2h, nobs, stats, solve_launches = e2e_once(200 + rep)
        dt, ts = allmax([dt, r.timer_solve_s])
        e2e_dt += dt
        e2e_solve += ts
        e2e_iters += r.iteration_attempt_num
    (nobs_all,) = allsum([nobs])
    e2e_value = nobs_all * e2e_iters / e2e_dt
    lm_iters_per_s = e2e_iters / e2e_solve
    frames_per_s = F * E2E_REPS / e2e_dt
    if comm is not None:
        comm.close()

    # ---- strong scaling: the FIXED cfg3 problem (1000 frames x 500 bundles) over the N ranks,
    # 1000 / N frames each; the replicated Cholesky of the 1500-dim reduced system and the
    # all-reduce are the serial (Amdahl) part
    strong = None
    if world > 1 and F1 % world == 0:
        Fs = F1 // world
        def strong_once(seed):
            sbs, _ = synth.make_scene(Fs, B, seed, noise_px=0.5, frame_offset=rank * Fs)
            barrier()
            t0 = time.perf_counter()
            with Problem(sbs, device=local, iter_max=20, local_slice=(rank * Fs, F1)) as q:
                q.use_comm(comm_strong)
                r = q.solve()
            return time.perf_counter() - t0, r
        comm_strong = make_comm(group, device=local)
        strong_once(100)
        ts = []
        for rep in range(3):
            dt, rs = strong_once(200 + rep)
            (dt,) = allmax([dt])
            ts.append(dt)
        comm_strong.close()
        strong = {"workload": "fixed cfg3: 1000 frames x 500 bundles split over %d GPUs (%d frames each)" % (world, Fs),
                  "seconds": sum(ts) / len(ts), "lm_outer_iterations": rs.iteration_attempt_num, "reason_num": rs.reason_num,
                  "device_linear_s": rs.timer_linear_s, "device_jacobian_s": rs.timer_jacobian_s}

    # ---- per-frame mode (cfg2): 240 x N frames x 50 bundles, frames sharded, no collective
    def perframe_once(seed):
        sbp, _ = synth.make_scene(240, 50, seed, solve_bundles=False, frame_offset=240 * rank)
        barrier()
        t0 = time.perf_counter()
        q = (Problem(sbp, device=local, frame_solve_mode=1, local_slice=(240 * rank, 240 * world)) if world > 1
             else Problem(sbp, device=local, frame_solve_mode=1))
        with q:
            agg, per = q.solve_per_frame()
        return time.perf_counter() - t0, agg
    perframe_once(2)
    pf_dt, pf_agg = perframe_once(2)
    (pf_dt,) = allmax([pf_dt])

    # the same mode at a size that fills the GPU: 24 000 frames per GPU, every rank generating and
    # holding only its own frames (mmb_problem_create_shard_local: no rank ingests another's frames)
    PF_BIG = 24000
    def perframe_big_once():
        f0 = rank * PF_BIG
        sbp, _ = synth.make_scene(PF_BIG, 50, 2, solve_bundles=False, frame_offset=f0)
        barrier()
        t0 = time.perf_counter()
        q = (Problem(sbp, device=local, frame_solve_mode=1, local_slice=(f0, PF_BIG * world)) if world > 1
             else Problem(sbp, device=local, frame_solve_mode=1))
        with q:
            agg, per = q.solve_per_frame()
        return time.perf_counter() - t0, agg
    perframe_big_once()
    pfb_dt, pfb_agg = perframe_big_once()
    (pfb_dt,) = allmax([pfb_dt])

    # ---- cfg4 evaluation step: lens distortion (LDPK fixed-point inverse) inside every
    # residual / FD evaluation -> FP64-bound, reported against the measured DFMA ceiling
    cfg4 = None
    if world == 1:
        from mayamatchmovesolver_b200.problem import fp64_peaks
        lens = dict(degree2_distortion=0.05, degree2_u=0.002, degree2_v=-0.002, degree4_distortion=0.01,
                    degree4_u=0.0, degree4_v=0.0)
        sb4, _ = synth.make_scene(F1, B, 4, lens=lens, solve_focal=True, solve_lens=tuple(lens), focal_init=33.0,
                                  in_frame=True)
        with Problem(sb4, device=local) as p4:
            k4 = p4.jacobian_kmax()
            for _ in range(3):
                p4.bench_eval(1, True)
            t4 = [p4.bench_eval(1, True) for _ in range(20)]
            step4 = sum(t[0] for t in t4) / len(t4)
            jac4 = sum(t[1] for t in t4) / len(t4)
            cfg4 = {"workload": "cfg4: cfg3 + focal + 6 radial-decentered deg-4 lens coefficients (n=%d), lens in the loop" % p4.n,
                    "evals_per_s": p4.nobs / (step4 * 1e-3), "ms_per_step": step4, "jacobian_kernel_ms": jac4,
                    "local_params_per_obs_k": k4, "bound": "fp64",
                    "algorithmic_GBps": p4.nobs * (48 + 16 * k4) / (jac4 * 1e-3) / 1e9}
        dfma, dmma = fp64_peaks(local)
        cfg4["fp64_peaks_measured_tflops"] = {"dfma": dfma, "dmma_m8n8k4": dmma}
        # the whole cfg4 solve (north_star's target shape) through the public call, host buffers
        t4s = []
        for _ in range(3):
            t0 = time.perf_counter()
            with Problem(sb4, device=local, iter_max=30) as q4:
                r4 = q4.solve()
            t4s.append(time.perf_counter() - t0)
        cfg4["solve"] = {"what": "mmb_problem_create + mmb_solve (iter_max=30) + destroy, host buffers, best of 3",
                         "seconds": min(t4s), "lm_outer_iterations": r4.iteration_attempt_num, "reason_num": r4.reason_num,
                         "error_avg_px": r4.error_avg, "device_seconds": r4.timer_solve_s,
                         "device_jacobian_s": r4.timer_jacobian_s, "device_linear_s": r4.timer_linear_s}

    # ---- the headline kernel on a launch the L2 cannot absorb: cfg3 x 16 (16 000 frames x 500
    # bundles = 8e6 observations, 1.5 GB of block Jacobian per launch)
    hbm_resident = None
    if world == 1:
        FB = 16000
        sbb, _ = synth.make_scene(FB, B, 3)
        with Problem(sbb, device=local) as pb:
            kb = pb.jacobian_kmax()
            for _ in range(3):
                pb.bench_eval(1, True)
            tb = [pb.bench_eval(1, True) for _ in range(10)]
            stepb = sum(t[0] for t in tb) / len(tb)
            jacb = sum(t[1] for t in tb) / len(tb)
            bytes_b = pb.nobs * (48 + 16 * kb)
            hbm_resident = {"workload": "cfg3 x 16: %d frames x %d bundles, %d observations, %.2f GB per launch" % (FB, B, pb.nobs, bytes_b / 1e9),
                            "kernel_ms": jacb, "step_ms": stepb, "algorithmic_bytes_per_launch": bytes_b,
                            "achieved": bytes_b / (jacb * 1e-3) / 1e9, "unit": "GB/s",
                            "whole_step_GBps": bytes_b / (stepb * 1e-3) / 1e9,
                            "evals_per_s": pb.nobs / (stepb * 1e-3), "traffic": None}
        del sbb

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    peaks, peaks_src = _peaks()
    algo_bytes = nobs_rank * bytes_per_obs
    achieved = algo_bytes / (jac_ms / args.steps * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic_k_jacobian.json")
    if os.path.exists(tp):
        with open(tp) as fh:
            traffic = json.load(fh).get("dram_bytes_per_launch")
    line = {
        "metric": "marker-frame residual+FD-Jacobian evals/s", "value": value, "unit": "evals/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames": F, "frames_per_gpu": F1, "bundles": B, "params_per_gpu": n_rank,
                   "observations_per_gpu": nobs_rank,
                   "local_params_per_obs_k": k, "l2": "flushed before every step (256 MiB memset outside the event brackets)",
                   "sharding": "one scene of 1000 x N frames; each rank generates and holds only its own 1000 frames (mmb_problem_create_shard_local); no data-path collective in the evaluation step"},
        "clocks": clocks,
        "gpu_launches": int(launches),
        "wall_s_timed_region": t_wall,
        "e2e": {"value": e2e_value, "unit": "evals/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "what": ("mmb_problem_create + mmb_solve (iter_max=20) + destroy with host buffers, mean of %d solves; "
                         "evals = observations x Jacobian evaluations" % E2E_REPS) +
                        ("; sharded global solve (mmb_problem_create_shard_local: each rank holds only its 1000 frames), "
                         "one NCCL all-reduce of the packed lower triangle of the reduced system per lmpar trial, "
                         "session communicator (mmb_comm_create) made once outside the timed region" if world > 1 else ""),
                "seconds": e2e_dt / E2E_REPS, "lm_outer_iterations": e2e_iters / E2E_REPS, "nfev": r.iteration_num,
                "reason_num": r.reason_num, "error_avg_px": r.error_avg,
                "host_syncs_per_solve": stats[0], "graph_launches_per_solve": stats[1], "kernel_launches_per_solve": solve_launches},
        "lm_iters_per_s": lm_iters_per_s,
        "solved_frames_per_s": frames_per_s,
        "per_frame_mode": {"workload": "cfg2: %d frames x 50 bundles, one LM solve per frame (n=6, m=100 each), one launch" % (240 * world),
                           "solved_frames_per_s": 240 * world / pf_dt, "seconds": pf_dt,
                           "error_avg_px": pf_agg.error_avg, "what": "create + mmb_solve_per_frame + destroy, host buffers"},
        "per_frame_mode_large": {"workload": "cfg2 shape at %d frames per GPU (%d frames in all) x 50 bundles, each rank holds only its own frames"
                                             % (PF_BIG, PF_BIG * world),
                                 "solved_frames_per_s": PF_BIG * world / pfb_dt, "seconds": pfb_dt, "error_avg_px": pfb_agg.error_avg,
                                 "what": "create (slice-local) + mmb_solve_per_frame + destroy, host buffers, max over ranks"},
        "roofline": {"bound": "hbm", "kernel": "k_jacobian_rows", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                     "frac": achieved / peaks["hbm_gbs"], "traffic": traffic, "algorithmic_bytes_per_launch": algo_bytes,
                     "peak_source": peaks_src, "kernel_ms": jac_ms / args.steps,
                     "whole_step": {"what": "variants + Jacobian kernel (the timed step): algorithmic bytes / step time",
                                    "ms": ms_per_step, "achieved": algo_bytes / (ms_per_step * 1e-3) / 1e9,
                                    "frac": algo_bytes / (ms_per_step * 1e-3) / 1e9 / peaks["hbm_gbs"]},
                     "note": "cfg3's 96 MB launch fits the 126 MB L2: its stores are L2-absorbed (DRAM counter `traffic` < "
                             "algorithmic bytes), so `frac` is algorithmic bandwidth against the HBM copy peak, not DRAM "
                             "utilisation; `hbm_resident` is the same kernel on a launch the L2 cannot hold"},
    }
    if hbm_resident is not None:
        hbm_resident["frac"] = hbm_resident["achieved"] / peaks["hbm_gbs"]
        hbm_resident["whole_step_frac"] = hbm_resident["whole_step_GBps"] / peaks["hbm_gbs"]
        tp2 = os.path.join(ROOT, "profiles", "traffic_k_jacobian_hbm_resident.json")
        if os.path.exists(tp2):
            with open(tp2) as fh:
                hbm_resident["traffic"] = json.load(fh).get("dram_bytes_per_launch")
        line["roofline"]["hbm_resident"] = hbm_resident
    # the same bytes through a plain streaming kernel (12 MB read, 84 MB written, L2 flushed):
    # the share of the 2-GiB copy peak a transfer of THIS size can reach at all
    if rank == 0:
        from mayamatchmovesolver_b200.problem import stream_floor
        read_b = 24 * nobs_rank
        floor_ms = stream_floor(read_b, algo_bytes - read_b, 30, local)
        line["roofline"]["same_size_stream"] = {
            "what": "plain kernel reading %d B and writing %d B, median of 30, same L2 flush" % (read_b, algo_bytes - read_b),
            "ms": floor_ms, "GBps": algo_bytes / (floor_ms * 1e-3) / 1e9, "frac_of_peak": algo_bytes / (floor_ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
            "kernel_vs_stream": floor_ms / (jac_ms / args.steps)}
    if cfg4 is not None:
        line["cfg4_eval"] = cfg4
    if strong is not None:
        line["strong_scaling_cfg3"] = strong
    if world == 1 and not args.no_cpu:
        leg = cpu_reference_leg(steps=1)
        line["cpu_baseline"] = {"value": leg["value"], "unit": "evals/s", "cores": 1, "kind": "port",
                                "sample": leg["sample"], "seconds": leg["seconds"]}
        line["cpu_baseline_block_sparse"] = cpu_block_sparse_leg()
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_own(args)


if __name__ == "__main__":
    main()
