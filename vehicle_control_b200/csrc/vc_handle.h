// Internal handle shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include "../../include/vc_api.h"

struct vc_handle {
    int device = 0;
    cudaStream_t stream = nullptr;
    vc_model_params params{};
    vc_mpc_config cfg{};
    // track tables (device)
    double* d_kc = nullptr;   // curvature spline coefficients [n_k]
    int n_k = 0;
    double k_x0 = 0.0, k_step = 0.05;
    double* d_xc = nullptr;   // x(u) spline coefficients [n_xy]
    double* d_yc = nullptr;
    int n_xy = 0;
    double length = 0.0;
    // track atlas (device): several curvature tables, one per track id (per-instance tracks)
    double* d_atlas = nullptr;     // concatenated curvature spline coefficients
    void* d_atlas_desc = nullptr;  // [n_atlas] descriptors {x0, dx, offset, n}
    int n_atlas = 0;
    // tracks built on the device (vc_build_tracks): waypoint / centre-line spline / curvature-table storage
    double* d_bt = nullptr;        // [wx | wy | cx | cy] each [sum n_wp], then the raw curvature tables [sum n_k]
    std::vector<int> bt_wp_off, bt_n_wp, bt_k_off, bt_n_k;   // per built track (host copies)
    std::vector<double> bt_length;
    long long bt_total_wp = 0, bt_total_k = 0;
    int sm_count = 148;
    int max_smem_optin = 0;
    long long launches = 0;
    // scratch for the host-buffer solve wrapper
    void* d_scratch = nullptr;
    size_t scratch_bytes = 0;
    // NMPC solver workspace
    void* d_ws = nullptr;
    size_t ws_bytes = 0;
    int* d_active = nullptr;   // [2] counters + 2 x [active_cap] compacted instance lists
    int active_cap = 0;
    int* h_active = nullptr;
    // host waits of the solver loop: 0 = spin (cudaStreamSynchronize, lowest latency), 1 = sleep on a blocking-sync event
    // (vc_set_blocking_sync: for several handles driven by more host threads than there are cores to spin on)
    int blocking_sync = 0;
    cudaEvent_t sync_event = nullptr;
    int last_passes = 0;
    // optional per-kernel timing of the solver loop (vc_set_profiling)
    int profiling = 0;
    std::vector<cudaEvent_t> prof_events;
    double prof_ms[4] = {0, 0, 0, 0};   // k_eval_st, k_eval_pm, k_step, launches of each
    long long prof_launches[3] = {0, 0, 0};
    long long prof_instance_passes = 0;   // sum over profiled passes of the number of unfinished instances
};

namespace vc {
struct TrackDesc { double x0, dx; int off, n; };   // one curvature table of the atlas: uniform sites x0 + i dx, coefficients at off
void set_error(const std::string& msg);
int check_cuda(cudaError_t e, const char* what);
// wait until everything enqueued on the handle's stream has finished (spinning or sleeping, see vc_handle::blocking_sync)
inline cudaError_t stream_wait(vc_handle* h) {
    if (!h->blocking_sync) return cudaStreamSynchronize(h->stream);
    if (!h->sync_event) {
        cudaError_t e = cudaEventCreateWithFlags(&h->sync_event, cudaEventBlockingSync | cudaEventDisableTiming);
        if (e != cudaSuccess) return e;
    }
    cudaError_t e = cudaEventRecord(h->sync_event, h->stream);
    if (e != cudaSuccess) return e;
    return cudaEventSynchronize(h->sync_event);
}
// launchers implemented in vc_solver.cu
int launch_nmpc_solve(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                      const double* d_obst, int n_obs, double* d_X, double* d_U, double* d_obj, int* d_status,
                      int* d_iters, double* d_kkt);
int launch_nmpc_solve_static(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                             const double* d_obst, int n_obs, double* d_X, double* d_U, double* d_obj, int* d_status,
                             int* d_iters, double* d_kkt);
int launch_nmpc_eval(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                     const double* d_obst, int n_obs, const double* d_X, const double* d_U, double* d_out);
int debug_read(vc_handle* h, int b, int array, double* h_out, int n);
}  // namespace vc
