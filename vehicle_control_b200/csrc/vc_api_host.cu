// C-ABI entry points of the NMPC solve (include/vc_api.h): device-pointer call, evaluation-only
// call, and the host-buffer wrapper used for end-to-end timing.
#include "vc_handle.h"

using namespace vc;

extern "C" {

int vc_nmpc_solve(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                  const double* d_obst, int n_obs, double* d_X, double* d_U, double* d_obj, int* d_status,
                  int* d_iters, double* d_kkt) {
    if (!h) { set_error("vc_nmpc_solve: null handle"); return -1; }
    if (B <= 0) return 0;
    return launch_nmpc_solve(h, B, d_x0, d_kappa, d_ds, d_obst, n_obs, d_X, d_U, d_obj, d_status, d_iters, d_kkt);
}

int vc_nmpc_solve_static(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                         const double* d_obst, int n_obs, double* d_X, double* d_U, double* d_obj, int* d_status,
                         int* d_iters, double* d_kkt) {
    if (!h) { set_error("vc_nmpc_solve_static: null handle"); return -1; }
    if (B <= 0) return 0;
    return launch_nmpc_solve_static(h, B, d_x0, d_kappa, d_ds, d_obst, n_obs, d_X, d_U, d_obj, d_status, d_iters, d_kkt);
}

int vc_nmpc_eval(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                 const double* d_obst, int n_obs, const double* d_X, const double* d_U, double* d_out) {
    if (!h) { set_error("vc_nmpc_eval: null handle"); return -1; }
    if (B <= 0) return 0;
    return launch_nmpc_eval(h, B, d_x0, d_kappa, d_ds, d_obst, n_obs, d_X, d_U, d_out);
}

int vc_nmpc_solve_host(vc_handle* h, int B, const double* h_x0, const double* h_kappa, const double* h_ds,
                       const double* h_obst, int n_obs, double* h_X, double* h_U, double* h_obj, int* h_status,
                       int* h_iters, double* h_kkt) {
    if (!h) { set_error("vc_nmpc_solve_host: null handle"); return -1; }
    if (B <= 0) return 0;
    const int H = h->cfg.N + h->cfg.M;
    const size_t nB = (size_t)B;
    // layout of the scratch arena (doubles): x0 | kappa | ds | obst | X | U | obj | kkt | status,iters (int)
    const size_t o_x0 = 0, o_k = o_x0 + nB * 8, o_ds = o_k + nB * H, o_ob = o_ds + nB * H,
                 o_X = o_ob + nB * (size_t)n_obs * 3, o_U = o_X + nB * 8 * H, o_obj = o_U + nB * 2 * H,
                 o_kkt = o_obj + nB, o_int = o_kkt + nB, total = o_int + nB;   // 2 int32 per instance = 1 double
    const size_t bytes = total * sizeof(double);
    cudaSetDevice(h->device);
    if (bytes > h->scratch_bytes) {
        cudaFree(h->d_scratch);
        h->d_scratch = nullptr; h->scratch_bytes = 0;
        int rc = check_cuda(cudaMalloc(&h->d_scratch, bytes), "cudaMalloc(scratch)");
        if (rc) return rc;
        h->scratch_bytes = bytes;
    }
    double* d = (double*)h->d_scratch;
    int* d_status = (int*)(d + o_int);
    int* d_iters = d_status + nB;
    cudaStream_t st = h->stream;
    cudaMemcpyAsync(d + o_x0, h_x0, nB * 8 * sizeof(double), cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(d + o_k, h_kappa, nB * H * sizeof(double), cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(d + o_ds, h_ds, nB * H * sizeof(double), cudaMemcpyHostToDevice, st);
    if (h_obst && n_obs > 0)
        cudaMemcpyAsync(d + o_ob, h_obst, nB * n_obs * 3 * sizeof(double), cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(d + o_X, h_X, nB * 8 * H * sizeof(double), cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(d + o_U, h_U, nB * 2 * H * sizeof(double), cudaMemcpyHostToDevice, st);
    int rc = launch_nmpc_solve(h, B, d + o_x0, d + o_k, d + o_ds, (h_obst && n_obs > 0) ? d + o_ob : nullptr, n_obs,
                               d + o_X, d + o_U, d + o_obj, d_status, d_iters, d + o_kkt);
    if (rc) return rc;
    cudaMemcpyAsync(h_X, d + o_X, nB * 8 * H * sizeof(double), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(h_U, d + o_U, nB * 2 * H * sizeof(double), cudaMemcpyDeviceToHost, st);
    if (h_obj) cudaMemcpyAsync(h_obj, d + o_obj, nB * sizeof(double), cudaMemcpyDeviceToHost, st);
    if (h_kkt) cudaMemcpyAsync(h_kkt, d + o_kkt, nB * sizeof(double), cudaMemcpyDeviceToHost, st);
    if (h_status) cudaMemcpyAsync(h_status, d_status, nB * sizeof(int), cudaMemcpyDeviceToHost, st);
    if (h_iters) cudaMemcpyAsync(h_iters, d_iters, nB * sizeof(int), cudaMemcpyDeviceToHost, st);
    return check_cuda(stream_wait(h), "vc_nmpc_solve_host");
}

// CascadedMPC.command() for a batch, host buffers in and out: _init_horizon on the device (needs vc_set_track), the
// solve, and the copies, all inside the call.  h_X / h_U: previous prediction in (warm start), solution out.
int vc_nmpc_command_host(vc_handle* h, int B, const double* h_x0, const double* h_obst, int n_obs, double* h_X, double* h_U,
                         double* h_obj, int* h_status, int* h_iters, double* h_kkt) {
    if (!h) { set_error("vc_nmpc_command_host: null handle"); return -1; }
    if (B <= 0) return 0;
    if (!h_x0 || !h_X || !h_U) { set_error("vc_nmpc_command_host: null buffer"); return -1; }
    const int H = h->cfg.N + h->cfg.M;
    const size_t nB = (size_t)B;
    const size_t o_x0 = 0, o_k = o_x0 + nB * 8, o_ds = o_k + nB * H, o_ob = o_ds + nB * H,
                 o_X = o_ob + nB * (size_t)n_obs * 3, o_U = o_X + nB * 8 * H, o_obj = o_U + nB * 2 * H,
                 o_kkt = o_obj + nB, o_int = o_kkt + nB, total = o_int + nB;
    const size_t bytes = total * sizeof(double);
    cudaSetDevice(h->device);
    if (bytes > h->scratch_bytes) {
        cudaFree(h->d_scratch);
        h->d_scratch = nullptr; h->scratch_bytes = 0;
        int rc = check_cuda(cudaMalloc(&h->d_scratch, bytes), "cudaMalloc(scratch)");
        if (rc) return rc;
        h->scratch_bytes = bytes;
    }
    double* d = (double*)h->d_scratch;
    int* d_status = (int*)(d + o_int);
    int* d_iters = d_status + nB;
    cudaStream_t st = h->stream;
    const bool obst = h_obst && n_obs > 0;
    cudaMemcpyAsync(d + o_x0, h_x0, nB * 8 * sizeof(double), cudaMemcpyHostToDevice, st);
    if (obst) cudaMemcpyAsync(d + o_ob, h_obst, nB * n_obs * 3 * sizeof(double), cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(d + o_X, h_X, nB * 8 * H * sizeof(double), cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(d + o_U, h_U, nB * 2 * H * sizeof(double), cudaMemcpyHostToDevice, st);
    int rc = vc_horizon_params(h, B, d + o_x0, d + o_X, d + o_ds, d + o_k);
    if (rc) return rc;
    rc = launch_nmpc_solve(h, B, d + o_x0, d + o_k, d + o_ds, obst ? d + o_ob : nullptr, n_obs, d + o_X, d + o_U, d + o_obj,
                           d_status, d_iters, d + o_kkt);
    if (rc) return rc;
    cudaMemcpyAsync(h_X, d + o_X, nB * 8 * H * sizeof(double), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(h_U, d + o_U, nB * 2 * H * sizeof(double), cudaMemcpyDeviceToHost, st);
    if (h_obj) cudaMemcpyAsync(h_obj, d + o_obj, nB * sizeof(double), cudaMemcpyDeviceToHost, st);
    if (h_kkt) cudaMemcpyAsync(h_kkt, d + o_kkt, nB * sizeof(double), cudaMemcpyDeviceToHost, st);
    if (h_status) cudaMemcpyAsync(h_status, d_status, nB * sizeof(int), cudaMemcpyDeviceToHost, st);
    if (h_iters) cudaMemcpyAsync(h_iters, d_iters, nB * sizeof(int), cudaMemcpyDeviceToHost, st);
    return check_cuda(stream_wait(h), "vc_nmpc_command_host");
}

int vc_nmpc_debug_read(vc_handle* h, int b, int array, double* h_out, int n) {
    if (!h || !h_out) { set_error("vc_nmpc_debug_read: null argument"); return -1; }
    return debug_read(h, b, array, h_out, n);
}

int vc_set_blocking_sync(vc_handle* h, int on) {
    if (!h) { set_error("vc_set_blocking_sync: null handle"); return -1; }
    h->blocking_sync = on ? 1 : 0;
    return 0;
}

long long vc_workspace_bytes(vc_handle* h) { return h ? (long long)h->ws_bytes : 0; }

int vc_set_profiling(vc_handle* h, int on) {
    if (!h) { set_error("vc_set_profiling: null handle"); return -1; }
    h->profiling = on;
    for (int i = 0; i < 4; ++i) h->prof_ms[i] = 0.0;
    for (int i = 0; i < 3; ++i) h->prof_launches[i] = 0;
    h->prof_instance_passes = 0;
    return 0;
}

int vc_get_profile(vc_handle* h, double* out_ms3, long long* out_launches3, int* out_last_passes) {
    if (!h) { set_error("vc_get_profile: null handle"); return -1; }
    for (int i = 0; i < 3; ++i) {
        if (out_ms3) out_ms3[i] = h->prof_ms[i];
        if (out_launches3) out_launches3[i] = h->prof_launches[i];
    }
    if (out_last_passes) *out_last_passes = h->last_passes;
    return 0;
}

long long vc_get_profile_instance_passes(vc_handle* h) { return h ? h->prof_instance_passes : -1; }

}  // extern "C"
