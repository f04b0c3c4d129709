"""mayamatchmovesolver_b200 -- B200-native solve hot path of mmSolver.

Host layer (Python, ctypes, no PyTorch) over libmmsolver_b200.so: the
hand-written sm_100a CUDA kernels + C++ host code behind the C-ABI of
include/mmsolver_b200.h.  There is no CPU fallback: compute calls fail loudly
without the built library or without a CUDA device.
"""
from . import _capi, cmd, lens, loadmarker, scene, solver, solveresult, synth  # noqa: F401
from .cmd import mmSolver  # noqa: F401
from .problem import Problem, SolveResult, default_options  # noqa: F401
from .scene import SceneBuilder  # noqa: F401

__all__ = ["Problem", "SolveResult", "SceneBuilder", "default_options", "mmSolver"]
