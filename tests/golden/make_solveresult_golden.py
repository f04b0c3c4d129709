"""Generates tests/golden/solveresult_golden.json with the REFERENCE's own parser.

Runs python/mmSolver/_api/solveresult.py of /root/reference (its Maya imports stubbed: the
functions used here are pure string handling) on a sample of mmSolver command strings and
stores what it returns.  tests/test_cmd_cpu.py checks mayamatchmovesolver_b200.solveresult
against this file.  Run in the build container only (the reference is not on the GPU box):

    python tests/golden/make_solveresult_golden.py
"""
import importlib.util
import json
import logging
import os
import sys
import types

REF = "/root/reference/python"
HERE = os.path.dirname(os.path.abspath(__file__))

SAMPLE = [
    "numberOfParameters=42", "numberOfErrors=120", "numberOfMarkerErrors=120",
    "numberOfAttrStiffnessErrors=0", "numberOfAttrSmoothnessErrors=0",
    "success=1", "reason_num=2", "error_final=8.13955", "error_final_average=0.51381",
    "error_final_maximum=1.67977", "error_final_minimum=0.043739", "iteration_num=17",
    "iteration_function_num=17", "iteration_jacobian_num=672", "user_interrupted=0",
    "timer_solve=0.0513", "timer_function=0.0014", "timer_jacobian=0.0038", "timer_parameter=0",
    "timer_error=0", "ticks_solve=51300000", "ticks_function=1400000", "ticks_jacobian=3800000",
    "ticks_parameter=0", "ticks_error=0",
    "error_per_marker_per_frame=mkr_A#1#0.25", "error_per_marker_per_frame=mkr_A#2#0.5",
    "error_per_marker_per_frame=mkr_B#1#0.75", "error_per_marker_per_frame=mkr_B#2#1e-05",
    "error_per_frame=1#0.5", "error_per_frame=2#0.250005",
    "solve_parameter_list=1.5#-2.25#3e-07#", "solve_error_list=0.1#0.2#0.3#0.4#",
]


def load_reference_parser():
    maya = types.ModuleType("maya")
    maya.cmds = types.ModuleType("maya.cmds")
    sys.modules["maya"] = maya
    sys.modules["maya.cmds"] = maya.cmds
    pkg = types.ModuleType("mmSolver")
    pkg.__path__ = [os.path.join(REF, "mmSolver")]
    sys.modules["mmSolver"] = pkg
    logger = types.ModuleType("mmSolver.logger")
    logger.get_logger = lambda *a, **k: logging.getLogger("mmSolver")
    sys.modules["mmSolver.logger"] = logger
    pkg.logger = logger
    utils = types.ModuleType("mmSolver.utils")
    utils.__path__ = [os.path.join(REF, "mmSolver", "utils")]
    sys.modules["mmSolver.utils"] = utils
    pkg.utils = utils
    spec = importlib.util.spec_from_file_location("mmSolver.utils.python_compat",
                                                  os.path.join(REF, "mmSolver", "utils", "python_compat.py"))
    compat = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(compat)
    sys.modules["mmSolver.utils.python_compat"] = compat
    utils.python_compat = compat
    spec = importlib.util.spec_from_file_location("ref_solveresult",
                                                  os.path.join(REF, "mmSolver", "_api", "solveresult.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ref = load_reference_parser()
    data = ref.parse_command_result(SAMPLE)
    out = {
        "sample": SAMPLE,
        "parsed_keys": sorted(data.keys()),
        "solver_stats": ref._string_get_solver_stats(data),
        "error_stats": ref._string_get_error_stats(data),
        "timer_stats": ref._string_get_timer_stats(data),
        "print_stats": ref._string_get_print_stats(data),
        "error_per_frame": {repr(k): v for k, v in ref._string_get_error_per_frame(data).items()},
        "error_per_marker_per_frame": {m: {repr(k): v for k, v in d.items()}
                                       for m, d in ref._string_get_error_per_marker_per_frame(data).items()},
    }
    path = os.path.join(HERE, "solveresult_golden.json")
    with open(path, "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)
    print("wrote", path)


if __name__ == "__main__":
    main()
