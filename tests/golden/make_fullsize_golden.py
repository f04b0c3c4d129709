"""Full-size golden vectors from the block-sparse CPU oracle (oracle/oracle_sparse.h).

cfg3 (1000 frames x 500 bundles, camera + bundles, n = 7488 with 4 bundles held to fix the
gauge) and cfg4 (the same plus focal length and six lens coefficients, n = 7495) are solved
on the CPU with the reference's evaluation arithmetic and MINPACK control law; the solved
attribute values, the result record and a sample of the per-marker deviations are stored in
tests/golden/fullsize_<cfg>.npz.  tests/test_fullsize_gpu.py compares the GPU solve of the
same seeded scene with them.  Takes a few minutes of CPU:

    python tests/golden/make_fullsize_golden.py [cfg3] [cfg4]
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from mayamatchmovesolver_b200 import synth  # noqa: E402
from oraclelib import OracleProblem  # noqa: E402

LENS = dict(degree2_distortion=0.05, degree2_u=0.002, degree2_v=-0.002, degree4_distortion=0.01,
            degree4_u=0.0, degree4_v=0.0)
SAMPLE_STRIDE = 97


def scene(name):
    if name == "cfg3":
        return synth.make_scene(1000, 500, 3, fixed_bundles=4)
    if name == "cfg4":
        return synth.make_scene(1000, 500, 4, lens=LENS, solve_focal=True, solve_lens=tuple(LENS), focal_init=33.0,
                                in_frame=True, fixed_bundles=4)
    raise KeyError(name)


def main():
    for name in (sys.argv[1:] or ["cfg3", "cfg4"]):
        sb, _ = scene(name)
        o = OracleProblem(sb)
        t0 = time.time()
        r = o.solve_sparse(threads=0)
        dt = time.time() - t0
        print(name, r, "njev", r.iteration_attempt_num, "%.1f s" % dt, flush=True)
        np.savez_compressed(os.path.join(HERE, "fullsize_%s.npz" % name),
                            attr_values=r.attr_values, x=r.x,
                            dist_sample=r.error_dist[::SAMPLE_STRIDE], sample_stride=SAMPLE_STRIDE,
                            summary=np.array([r.reason_num, r.iteration_num, r.iteration_attempt_num,
                                              r.iteration_function_num, r.iteration_jacobian_num, int(r.error_is_better)]),
                            errors=np.array([r.error_final, r.error_avg, r.error_min, r.error_max]),
                            cpu_seconds=dt)


if __name__ == "__main__":
    main()
