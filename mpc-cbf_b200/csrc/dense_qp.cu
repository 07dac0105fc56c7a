// placeholder until the ADMM kernel lands
#include <string>
#include "../../include/mpccbf.h"
#include "kernels.h"
namespace mpccbf {
int dense_qp_solve_host(::mpccbf_handle*, int32_t, int32_t, int32_t, const double*, const double*, const double*,
                        const double*, const double*, const double*, const double*, double*, double*, int32_t*, int32_t*) {
    set_error("dense ADMM kernel not built yet");
    return MPCCBF_ERR_UNSUPPORTED;
}
}
