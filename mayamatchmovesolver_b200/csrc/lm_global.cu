#include "lm.h"
namespace mmb {
struct LmWorkspace { int dummy; };
void lm_workspace_destroy(LmWorkspace *w) { delete w; }
mmb_status lm_solve(DeviceProblem *, mmb_result *, double *, double *, double *) {
    set_last_error("lm_solve: not built yet");
    return MMB_ERR_UNSUPPORTED;
}
}  // namespace mmb
