"""Bulk lens kernel throughput (device-resident, CUDA events) for every model and direction,
with the reference's compiled mmlens on one host core beside it."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, "tests")
sys.path.insert(0, ".")
from mayamatchmovesolver_b200 import lens  # noqa: E402
from oraclelib import _p, ref_lens  # noqa: E402

P = {1: [0.1, 1.0, 0.01, 0.02, 0.05], 2: [0.1, 0.02, -0.01, 0.05, 0.01, -0.02, 10.0, 0.1],
     3: [0.02, 0.01, -0.015, 0.005, 0.03, -0.01, 0.002, 0.004, -0.003, 0.001, 5.0, 1.02, 0.98],
     4: [0.02, 0.01, -0.015, 0.005, 0.03, -0.01, 0.002, 0.004, -0.003, 0.001, 5.0, 1.02, 0.98, 1.1]}
CAM = np.array([3.5, 3.6, 2.4, 1.0, 0.05, -0.03])
N = 1 << 25   # 33.5 M points: 512 MB in + 512 MB out, far above the 126 MB L2
out = []
ref = ref_lens()
for model in (1, 2, 3, 4):
    co = np.zeros(14)
    co[:len(P[model])] = P[model]
    for direction in (0, 1):
        ms = lens.bench(model, co, direction, N, reps=10, camera_params=CAM)
        row = {"model": model, "direction": "undistort" if direction == 0 else "distort", "points": N, "ms": ms,
               "points_per_s": N / ms * 1e3, "GBps": 32.0 * N / ms / 1e6}
        if ref is not None:
            n = 1 << 20
            pts = np.random.default_rng(0).uniform(-0.5, 0.5, (n, 2))
            o = np.zeros_like(pts)
            t = time.perf_counter()
            ref.ref_lens_apply(model, _p(co), _p(CAM), direction, n, _p(pts), _p(o))
            row["cpu_ref_points_per_s"] = n / (time.perf_counter() - t)
        t = time.perf_counter()
        pts = np.random.default_rng(0).uniform(-0.5, 0.5, (1 << 23, 2))
        lens.apply(model, co, pts, direction, CAM)
        t = time.perf_counter()
        lens.apply(model, co, pts, direction, CAM)
        row["host_buffers_points_per_s"] = (1 << 23) / (time.perf_counter() - t)
        out.append(row)
        print(json.dumps(row), flush=True)
json.dump(out, open("gpurun_out/lens_bench.json", "w"), indent=1)
