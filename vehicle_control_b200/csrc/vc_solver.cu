// placeholder -- replaced by the batched interior-point NMPC solver
#include "vc_handle.h"
namespace vc {
int launch_nmpc_solve(vc_handle*, int, const double*, const double*, const double*, const double*, int, double*,
                      double*, double*, int*, int*, double*) { set_error("nmpc solver not built"); return -9; }
int launch_nmpc_eval(vc_handle*, int, const double*, const double*, const double*, const double*, int,
                     const double*, const double*, double*) { set_error("nmpc solver not built"); return -9; }
}
