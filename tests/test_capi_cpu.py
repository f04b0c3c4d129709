"""CPU: the C-ABI library loads, exports every symbol include/mmsolver_b200.h declares, and
refuses to compute without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from mayamatchmovesolver_b200 import _capi, build, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()  # nvcc cross-compiles sm_100a without a GPU
    return _capi.load()


def test_header_symbols_all_exported_and_bound(lib):
    hdr = open(os.path.join(ROOT, "include", "mmsolver_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(mmb_[a-z_0-9]+)\s*\(", hdr))
    bound = {name for name, _, _ in _capi.SYMBOLS}
    assert declared == bound, (declared ^ bound)
    for name in declared:
        assert hasattr(lib, name)


def test_struct_layouts_match_header_sizes(lib):
    # sizes computed from the header with the host C compiler
    import subprocess
    import tempfile
    src = '#include <stdio.h>\n#include "mmsolver_b200.h"\nint main(){printf("%zu %zu %zu\\n", sizeof(mmb_scene_desc), sizeof(mmb_options), sizeof(mmb_result));return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "s.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        sizes = [int(v) for v in subprocess.check_output([exe]).split()]
    assert sizes == [C.sizeof(_capi.SceneDesc), C.sizeof(_capi.Options), C.sizeof(_capi.Result)]


def test_options_default_are_reference_defaults(lib):
    o = _capi.Options()
    lib.mmb_options_default(C.byref(o))
    # adjust_defines.h:92-152 / arg_flags_solve_info.h
    assert (o.iter_max, o.tau, o.eps1, o.eps2, o.eps3, o.delta) == (100, 1.0, 1e-6, 1e-6, 1e-6, 1e-4)
    assert (o.auto_diff_type, o.auto_param_scale, o.solver_type, o.accept_only_better) == (0, 1, 2, 1)
    assert o.image_width == 2048.0 and o.frame_solve_mode == 0


def test_no_gpu_means_error_not_fallback(lib):
    from conftest import HAS_GPU
    if HAS_GPU:
        pytest.skip("GPU present")
    sb, _ = synth.make_scene(F=2, B=3, seed=0)
    desc, keep = sb.desc()
    o = _capi.Options()
    lib.mmb_options_default(C.byref(o))
    h = C.c_void_p()
    st = lib.mmb_problem_create(C.byref(desc), C.byref(o), C.byref(h))
    assert st == _capi.MMB_ERR_NO_DEVICE and not h.value
    assert b"no CPU path" in lib.mmb_status_string(st) or b"no CPU path" in lib.mmb_last_error()


def test_product_does_not_reference_oracle():
    pkg = os.path.join(ROOT, "mayamatchmovesolver_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle/" not in txt and "liboracle" not in txt and "oraclelib" not in txt, f
