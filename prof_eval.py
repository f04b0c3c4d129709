import sys; sys.path.insert(0,'.')
from mayamatchmovesolver_b200 import synth
from mayamatchmovesolver_b200.problem import Problem
sb,_ = synth.make_scene(1000,500,3)
p = Problem(sb)
print(p.bench_eval(5, True))
