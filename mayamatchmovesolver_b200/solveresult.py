"""Reading the `key=value` strings of the mmSolver command.

Mirrors python/mmSolver/_api/solveresult.py:41-260 of the reference (the part that
does not touch Maya nodes): the command returns a list of strings `key=value`, list values
are joined with '#'.  tests/golden/solveresult_golden.json holds what the reference's own
parser makes of a sample; this module must agree with it.
"""
import collections
import math

KEY_VALUE_SEP_CHAR = "="
SPLIT_SEP_CHAR = "#"


def parse_command_result(cmd_result):
    """list of 'key=value' strings -> {key: [value | [values...], ...]}"""
    data = collections.defaultdict(list)
    for res in cmd_result:
        key, _, value = res.partition(KEY_VALUE_SEP_CHAR)
        if not key:
            continue
        if SPLIT_SEP_CHAR in value:
            value = value.split(SPLIT_SEP_CHAR)
        data[key].append(value)
    return data


def _convert(typ, value, index=0):
    """value = the list stored under a key (or one '#'-split entry of it)"""
    if value is None or len(value) == 0:
        return typ()
    v = value[index]
    if not isinstance(v, (str, float, int, bool)):
        return typ()
    if typ is int:
        return int(float(v))
    if typ is bool:
        return bool(int(v))
    out = typ(v)
    if typ is float and (math.isinf(out) or math.isnan(out)):
        return -1.0
    return out


def _table(data, name_keys):
    return {name: _convert(typ, data.get(key)) for name, key, typ in name_keys}


def solver_stats(data):
    return _table(data, [
        ("success", "success", bool), ("stop_message", "reason_string", str), ("stop_id", "reason_num", int),
        ("iteration_total_calls", "iteration_num", int), ("iteration_function_calls", "iteration_function_num", int),
        ("iteration_jacobian_calls", "iteration_jacobian_num", int), ("attempts", "iteration_attempt_num", int),
        ("user_interrupted", "user_interrupted", bool)])


def error_stats(data):
    return _table(data, [
        ("initial", "error_initial", float), ("maximum", "error_maximum", float), ("final", "error_final", float),
        ("final_average", "error_final_average", float), ("final_maximum", "error_final_maximum", float),
        ("final_minimum", "error_final_minimum", float), ("jt", "error_jt", float), ("dp", "error_dp", float)])


def timer_stats(data):
    return _table(data, [
        ("solve_seconds", "timer_solve", float), ("function_seconds", "timer_function", float),
        ("jacobian_seconds", "timer_jacobian", float), ("parameter_seconds", "timer_parameter", float),
        ("error_seconds", "timer_error", float), ("solve_ticks", "ticks_solve", int),
        ("function_ticks", "ticks_function", int), ("jacobian_ticks", "ticks_jacobian", int),
        ("parameter_ticks", "ticks_parameter", int), ("error_ticks", "ticks_error", int)])


def print_stats(data):
    return _table(data, [("number_of_parameters", "numberOfParameters", int),
                         ("number_of_errors", "numberOfErrors", int)])


def error_per_frame(data):
    out = {}
    for value in data.get("error_per_frame") or []:
        out[_convert(float, value, 0)] = _convert(float, value, 1)
    return out


def error_per_marker_per_frame(data):
    out = collections.defaultdict(dict)
    for value in data.get("error_per_marker_per_frame") or []:
        out[_convert(str, value, 0)][_convert(float, value, 1)] = _convert(float, value, 2)
    return out


class SolveResultStrings:
    """What the reference's SolveResult object exposes from the command strings
    (solveresult.py:402-470)."""

    def __init__(self, cmd_result):
        self.raw = list(cmd_result)
        data = parse_command_result(cmd_result)
        self.solver_stats = solver_stats(data)
        self.error_stats = error_stats(data)
        self.timer_stats = timer_stats(data)
        self.print_stats = print_stats(data)
        self.per_frame_error = error_per_frame(data)
        self.per_marker_per_frame_error = error_per_marker_per_frame(data)
        vals = data.get("solve_parameter_list") or [[]]
        self.solve_parameter_list = [float(v) for v in vals[0] if v != ""] if isinstance(vals[0], list) else []
        vals = data.get("solve_error_list") or [[]]
        self.solve_error_list = [float(v) for v in vals[0] if v != ""] if isinstance(vals[0], list) else []

    def get_success(self):
        return self.solver_stats["success"]

    def get_final_error(self):
        return self.error_stats["final"]

    def get_frame_error_list(self):
        return dict(self.per_frame_error)

    def get_marker_error_list(self, marker_node=None):
        if marker_node is None:
            return {k: dict(v) for k, v in self.per_marker_per_frame_error.items()}
        return dict(self.per_marker_per_frame_error.get(marker_node, {}))
