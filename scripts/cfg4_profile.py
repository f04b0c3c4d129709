"""cfg4: cfg3 + 3DE radial-decentered deg-4 lens + focal length solved (lens distortion inside
every residual / FD evaluation).  Prints the evaluation-step timing and one solve."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mayamatchmovesolver_b200 import synth  # noqa: E402
from mayamatchmovesolver_b200.problem import Problem  # noqa: E402

F = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
B = int(sys.argv[2]) if len(sys.argv) > 2 else 500
it = int(sys.argv[3]) if len(sys.argv) > 3 else 30
LENS = dict(degree2_distortion=0.05, degree2_u=0.002, degree2_v=-0.002, degree4_distortion=0.01, degree4_u=0.0, degree4_v=0.0)
sb, _ = synth.make_scene(F, B, 4, noise_px=0.5, lens=LENS, solve_focal=True, solve_lens=tuple(LENS), in_frame=True, focal_init=33.0)
with Problem(sb, iter_max=it) as p:
    k = p.jacobian_kmax()
    for _ in range(3):
        p.bench_eval(1, True)
    ms = [p.bench_eval(1, True) for _ in range(20)]
    step = sum(m[0] for m in ms) / len(ms)
    jac = sum(m[1] for m in ms) / len(ms)
    print("cfg4 F=%d B=%d n=%d nobs=%d k=%d: eval step %.3f ms, Jacobian kernel %.3f ms -> %.3e evals/s, %.1f GB/s algorithmic (48+16k B/obs)"
          % (F, B, p.n, p.nobs, k, step, jac, p.nobs / (step * 1e-3), p.nobs * (48 + 16 * k) / (jac * 1e-3) / 1e9))
for rep in range(2):
    t0 = time.perf_counter()
    with Problem(sb, iter_max=it) as p:
        r = p.solve()
        st = p.solve_stats()
    dt = time.perf_counter() - t0
    print("create+solve+destroy %.4f s (solve %.4f: jac %.4f func %.4f linear %.4f) %r LM iterations %d host syncs %d"
          % (dt, r.timer_solve_s, r.timer_jacobian_s, r.timer_function_s, r.timer_linear_s, r, r.iteration_attempt_num, st[0]))
