"""Critical-path timeline of the reduced-system Cholesky (k_chol_tiles) on cfg3."""
import ctypes as C
import sys

sys.path.insert(0, ".")
import numpy as np  # noqa: E402

from mayamatchmovesolver_b200 import _capi, synth  # noqa: E402
from mayamatchmovesolver_b200.problem import Problem  # noqa: E402

sb, _ = synth.make_scene(1000, 500, 3)
with Problem(sb, iter_max=4) as p:
    r = p.solve()
    buf = (C.c_uint64 * 768)()
    _capi.check(_capi.load().mmb_debug_chol_trace(buf, 768), "trace")
t = np.array(buf[:], dtype=np.float64).reshape(64, 2, 6)
nb = 24
print("mean per-column period %.2f us" % (np.mean(t[1:nb, 0, 5] - t[0:nb - 1, 0, 5]) / 1e3))
j = slice(2, nb - 1)
print("diag: flag->smem %.2f, smem->upd %.2f, upd->potrf %.2f, potrf->pub %.2f"
      % tuple(np.mean(t[j, 0, b] - t[j, 0, a]) / 1e3 for a, b in ((0, 1), (1, 2), (2, 3), (3, 5))))
print("sub : dflag - diag pub %.2f, dflag->trsm %.2f, trsm->pub %.2f; next diag flag - sub pub %.2f"
      % (np.mean(t[j, 1, 4] - t[j, 0, 5]) / 1e3, np.mean(t[j, 1, 3] - t[j, 1, 4]) / 1e3, np.mean(t[j, 1, 5] - t[j, 1, 3]) / 1e3,
         np.mean(t[3:nb, 0, 0] - t[2:nb - 1, 1, 5]) / 1e3))
