// Levenberg-Marquardt driver on the device (lm_global.cu).
#pragma once
#include "device_problem.h"

namespace mmb {
mmb_status lm_solve(DeviceProblem *p, mmb_result *res, double *x_out, double *attr_out, double *dist_out);
mmb_status lm_solve_per_frame(DeviceProblem *p, mmb_result *res, double *x_out, double *attr_out, double *dist_out,
                              mmb_result *per_frame_out);
void lm_workspace_destroy(LmWorkspace *w);
// sets the flag the device-resident loop polls (pinned, mapped) and the host loop's flag
void lm_request_cancel(DeviceProblem *p);
// debug: critical-path timeline of the last k_chol_tiles launch (lm_kernels.cuh g_chol_trace)
mmb_status lm_chol_trace(unsigned long long *out, int count);
}  // namespace mmb
