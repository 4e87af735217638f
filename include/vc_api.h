/* vc_api.h -- C-ABI of the B200 batched NMPC engine (libvc_b200.so).
 *
 * The reference (neverorfrog/vehicle-control) has no FFI of its own: the hot path sits behind
 * Python classes that call CasADi/IPOPT.  Each entry point below replaces one of those call
 * sites; the reference-side binding is a ctypes stub.
 *
 * Conventions
 *   - every pointer argument named d_* is a DEVICE pointer to contiguous row-major memory owned by
 *     the caller (fp64 unless stated int32); h_* are host pointers, copied during the call.
 *   - all functions return 0 on success, <0 on error; vc_last_error() gives a thread-local message.
 *   - work is enqueued on the handle's CUDA stream.  The rollout / track / plant entry points are asynchronous;
 *     vc_nmpc_solve is NOT: its interior-point loop reads the number of unfinished instances back once per pass
 *     (a 4-byte copy + stream synchronisation) to size the next launches, so it returns with the results complete.
 *     A handle is bound to one device and is not re-entrant; handles are independent of each other, so several batches
 *     can be in flight on one GPU, one handle (on its own stream) and one host thread each.
 *   - state order  SingleTrack: Ux Uy r delta s ey epsi t   (models/dynamic_car.py:246-247)
 *                  PointMass  : V s ey epsi t              (models/dynamic_point_mass.py:155-156)
 *     input order  SingleTrack: Fx w ; PointMass: Fx Fy     (dynamic_car.py:208-209, dynamic_point_mass.py:120-121)
 */
#ifndef VC_API_H
#define VC_API_H

#ifdef __cplusplus
extern "C" {
#endif

/* config/models/dynamic_car.yaml, flattened (consumed by models/dynamic_car.py:49-191) */
typedef struct vc_model_params {
    double dt;                 /* plant sampling time                      dynamic_car.yaml:1  */
    double l, m, Izz, a, b, h, eps, Peng;
    double Xdf, Xdr, Xbf, Xbr; /* drive / brake distribution                                   */
    double Caf, Car;           /* cornering stiffness                                          */
    double Cd, muf, mur, Av2, Frr;
    double g_eff;              /* 9.88*cos(theta)*cos(phi)                 dynamic_car.py:61,98 */
} vc_model_params;

/* config/controllers/{singletrack,cascaded}.yaml (consumed by controllers/mpc/cascaded_mpc.py) */
typedef struct vc_mpc_config {
    int N;                     /* horizon      (SingleTrack stages)        cascaded_mpc.py:42   */
    int M;                     /* horizon_pm   (PointMass stages)          cascaded_mpc.py:48   */
    int obstacles;             /* 0/1                                      cascaded_mpc.py:173  */
    int max_iter;              /* IPM iteration cap                                             */
    double mpc_dt, ds_pm;
    /* cost_weights */
    double w_time, w_speed, w_ey, w_epsi, w_w, w_Fx, w_Fy, w_dev_st, w_dev_pm, w_boundary, w_slip,
        w_obstacles, w_switch_F;
    /* input / state constraints */
    double w_min, w_max, Ux_min, max_speed, delta_min, delta_max, ey_min, ey_max;
    double V_min, ey_min_pm, ey_max_pm;
    /* interior-point options (IPOPT defaults used by the reference: cascaded_mpc.py:55-65) */
    double tol;                /* 1e-8                                                          */
    double acceptable_tol;     /* 1e-6                                                          */
    int acceptable_iter;       /* 15                                                            */
    int obj_scaling;           /* 1 = gradient-based objective scaling (max gradient 100)       */
    double mu_init;            /* 0.1                                                           */
    double bound_relax;        /* 1e-8                                                          */
    int warm_rungs;            /* >0: the first iteration starts this many rungs down the barrier ladder (warm starts) */
    int max_soc;               /* IPOPT max_soc: second-order correction steps per iteration (default 4)             */
    int watchdog_trigger;      /* IPOPT watchdog_shortened_iter_trigger (default 10; 0 = watchdog off)                */
    int watchdog_trials;       /* IPOPT watchdog_trial_iter_max (default 3)                                           */
    /* where IPOPT would enter its restoration phase (no acceptable step along a direction) the engine damps the step
     * (larger delta_w) and re-solves, at most max_damp times, then re-initialises the iterate by a rollout (twice) */
    int max_damp;              /* damped re-solves of one iteration's direction (8)                                   */
    int wide_max;              /* passes with at most this many unfinished instances run one BLOCK per instance
                                  (latency) instead of one warp (throughput); 0 = never (296 = two per SM)           */
} vc_mpc_config;

typedef struct vc_handle vc_handle;

/* solver exit status written to d_status[b] by vc_nmpc_solve */
enum { VC_SOLVED = 0, VC_ACCEPTABLE = 1, VC_MAX_ITER = 2, VC_EVAL_ERROR = 3, VC_LINESEARCH_FAIL = 4 };

const char* vc_last_error(void);
int vc_version(void);

/* replaces: DynamicCar/DynamicPointMass/CascadedMPC construction (racing.py:43-63).  `stream`
 * is a cudaStream_t (NULL = legacy default stream). */
int vc_create(const vc_model_params* h_params, const vc_mpc_config* h_cfg, int device, void* stream,
              vc_handle** out);
int vc_destroy(vc_handle* h);
int vc_set_config(vc_handle* h, const vc_mpc_config* h_cfg);
/* fills the solver options of *cfg (tol .. wide_max, max_iter) with the defaults: IPOPT's as the reference leaves them
 * (cascaded_mpc.py:55-65) and this engine's policies; the controller part (N .. ey_max_pm) is left untouched */
int vc_default_config(vc_mpc_config* h_cfg);

/* replaces: Track._precompute_curvatures / Track.k (environment/track.py:156-167) and the x/y
 * splines behind rel2glob (track.py:102-129,206-252).  Cubic not-a-knot B-spline COEFFICIENTS built
 * on the host over uniformly spaced sites (curvature: k_x0 + i*k_step, i<n_k; x/y: integer waypoint
 * index, i<n_xy, evaluated at s/length*n_xy); the knot vectors are arithmetic and not passed.
 * Copied to the device here (synchronous).  h_x_coefs/h_y_coefs may be NULL if rel2glob is unused. */
int vc_set_track(vc_handle* h, const double* h_k_coefs, int n_k, double k_x0, double k_step,
                 const double* h_x_coefs, const double* h_y_coefs, int n_xy, double length);

/* replaces: Track.k(s) for arrays (track.py:166). */
int vc_track_curvature(vc_handle* h, int n, const double* d_s, double* d_kappa);

/* replaces: Track.rel2glob (track.py:102-107) / CascadedMPC.get_state_prediction (cascaded_mpc.py:340-352).
 * d_out[n,3] = (x, y, psi). */
int vc_rel2glob(vc_handle* h, int n, const double* d_s, const double* d_ey, const double* d_epsi, double* d_out);

/* replaces: CascadedMPC._init_horizon (cascaded_mpc.py:316-338): per-stage ds and curvature from
 * the previous prediction.  d_Xprev[B,8,H] (only row Ux of the first N columns is read), d_x0[B,8]. */
int vc_horizon_params(vc_handle* h, int B, const double* d_x0, const double* d_Xprev, double* d_ds,
                      double* d_kappa);

/* replaces: car.spatial_transition applied N times (dynamic_car.py:188-191) and CasADi's AD of it.
 * d_X[B,N+1,8]; d_AB[B,N,8,10] = d x_{n+1} / d (x_n,u_n), may be NULL. */
int vc_rollout_st(vc_handle* h, int B, int N, const double* d_x0, const double* d_U, const double* d_kappa,
                  const double* d_ds, double* d_X, double* d_AB);

/* replaces: point_mass.spatial_transition applied M times (dynamic_point_mass.py:100-103).
 * d_X[B,M+1,5]; d_AB[B,M,5,7] may be NULL. */
int vc_rollout_pm(vc_handle* h, int B, int M, const double* d_x0, const double* d_U, const double* d_kappa,
                  const double* d_ds, double* d_X, double* d_AB);

/* replaces: RacingCar.drive (racing_car.py:34-46): kappa = k(s); x <- RK4_t(x,u,kappa,dt) in place. */
int vc_plant_step(vc_handle* h, int B, double* d_x, const double* d_u, double dt);

/* Per-instance tracks (the reference builds ONE Track per simulation, track.py:84-167): an atlas of
 * the curvature splines of several tracks (Track.k of each, track.py:156-167).  h_coefs = the tables concatenated,
 * h_offsets[n_tracks+1] their extents, h_x0 / h_step the uniform sites of each table (host pointers, copied once).
 * vc_horizon_params_tracks / vc_plant_step_tracks = vc_horizon_params / vc_plant_step with the table of
 * d_track_id[b] (int32, device) for instance b; same arithmetic, bit for bit. */
int vc_set_track_atlas(vc_handle* h, int n_tracks, const double* h_coefs, const int* h_offsets, const double* h_x0,
                       const double* h_step);
int vc_horizon_params_tracks(vc_handle* h, int B, const int* d_track_id, const double* d_x0, const double* d_Xprev,
                             double* d_ds, double* d_kappa);
int vc_plant_step_tracks(vc_handle* h, int B, const int* d_track_id, double* d_x, const double* d_u, double dt);

/* replaces: Track.__init__ for a batch of GENERATED tracks, on the device: Track._construct_path
 * (track.py:254-296: corner polygon -> waypoints every `resolution` m -> centred moving average of half-width `smoothing`,
 * numpy's summation order), Track._construct_spline (:206-252: not-a-knot cubic B-spline through the waypoints, track
 * length by the trapezoid rule), Track.get_curvature (:109-119) sampled every 0.05 m and Track._precompute_curvatures
 * (:156-167: cubic B-spline through that table = Track.k).  h_corners[T][max_corners][2] (host), h_n_corners[T],
 * h_resolution[T], h_smoothing[T].  The built curvature splines REPLACE the handle's track atlas: track t is then
 * `track_id == t` in vc_horizon_params_tracks / vc_plant_step_tracks.  Synchronises (two counts are read back to size the
 * tables).  vc_built_track_info: length / number of waypoints / curvature-table size of built track t.
 * vc_built_track_read copies a table of track t to the host: what = 0 Track.k coefficients, 1 / 2 smoothed waypoints x / y,
 * 3 / 4 coefficients of x(u) / y(u), 5 the curvature table.  vc_built_track_rel2glob: (s, ey) -> (cx, cy) on track t, the
 * obstacle placement of Track._construct_obstacles (:131-138); d_xy[n,2]. */
int vc_build_tracks(vc_handle* h, int T, int max_corners, const double* h_corners, const int* h_n_corners,
                    const double* h_resolution, const int* h_smoothing);
int vc_built_track_info(vc_handle* h, int t, double* length, int* n_waypoints, int* n_k);
int vc_built_track_read(vc_handle* h, int t, int what, double* h_out, int n);
int vc_built_track_rel2glob(vc_handle* h, int t, int n, const double* d_s, const double* d_ey, double* d_xy);

/* replaces: the tyre helper Functions car.Fx_f/Fx_r/Fy_f/Fy_r/alpha_f/alpha_r (dynamic_car.py:79-142)
 * for arrays.  d_out[B,6] = Fx_f Fx_r Fy_f Fy_r alpha_f alpha_r. */
int vc_tyre_forces(vc_handle* h, int B, const double* d_x, const double* d_u, double* d_out);

/* replaces: CascadedMPC.command -> opti.solve() (cascaded_mpc.py:306-314).  One NMPC instance per
 * batch entry: primal-dual interior point with a per-instance Riccati KKT factorisation.
 *   d_x0[B,8]; d_kappa[B,H]; d_ds[B,H]; d_obst[B,n_obs,3] = (s,ey,radius) or NULL;
 *   d_X[B,8,H], d_U[B,2,H]: warm start in, solution out (layout of state_prediction / action_prediction);
 *   d_obj[B]; d_status[B] int32; d_iters[B] int32; d_kkt[B] final scaled NLP error. */
int vc_nmpc_solve(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                  const double* d_obst, int n_obs, double* d_X, double* d_U, double* d_obj, int* d_status,
                  int* d_iters, double* d_kkt);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (a tight serving loop, a harness that forbids allocation).  Same arguments, kernels and results as vc_nmpc_solve, but the
 * call allocates NOTHING: the solver workspace and the lists of unfinished instances live in a file-scope
 * `static __device__` array of VC_STATIC_SCRATCH_BYTES (16 MiB) inside the library, which every GPU that loads the library
 * holds until the process exits.  Limit: B * ((596 * H + 112) * 8 + 8) + 264 <= VC_STATIC_SCRATCH_BYTES with H = N + M, i.e.
 * at most 63 instances at H = 55 (cascaded 20 + 35), 99 at H = 35, 70 at H = 50; a larger request is rejected with an error
 * (vc_last_error) and nothing is launched.  All handles and all calls on one GPU share the array: two calls that overlap on
 * the GPU (different streams or handles) race -- serialise them.  The array keeps the previous call's contents; the call
 * re-initialises every record it uses.  Like vc_nmpc_solve it reads one int back per pass (it is not capturable). */
#define VC_STATIC_SCRATCH_BYTES (16u << 20)
int vc_nmpc_solve_static(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                         const double* d_obst, int n_obs, double* d_X, double* d_U, double* d_obj, int* d_status,
                         int* d_iters, double* d_kkt);

/* Objective / constraint residual of a given trajectory (no solve): diagnostics + parity tests.
 * d_out[B,3] = objective, max equality violation, min inequality margin. */
int vc_nmpc_eval(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                 const double* d_obst, int n_obs, const double* d_X, const double* d_U, double* d_out);

/* Host-buffer convenience wrapper of the same call (the e2e path): copies inputs H2D, solves,
 * copies X,U,obj,status,iters back; synchronises the stream. */
int vc_nmpc_solve_host(vc_handle* h, int B, const double* h_x0, const double* h_kappa, const double* h_ds,
                       const double* h_obst, int n_obs, double* h_X, double* h_U, double* h_obj,
                       int* h_status, int* h_iters, double* h_kkt);

/* replaces: CascadedMPC.command(state) for a batch, host buffers in and out (cascaded_mpc.py:306-338): _init_horizon on the
 * device (vc_set_track must have been called), the solve, and every host<->device copy inside the call.  h_x0[B,8];
 * h_obst[B,n_obs,3] or NULL; h_X[B,8,H] / h_U[B,2,H]: previous prediction in (the warm start), solution out.  Synchronises. */
int vc_nmpc_command_host(vc_handle* h, int B, const double* h_x0, const double* h_obst, int n_obs, double* h_X, double* h_U,
                         double* h_obj, int* h_status, int* h_iters, double* h_kkt);

/* How the host waits inside vc_nmpc_solve / vc_nmpc_*_host (no reference counterpart: the reference's solve is host code).
 * 0 (default): spin on the stream (lowest latency: a single vehicle's tick).  1: sleep on a blocking-sync event -- for
 * several handles in flight, each driven by its own host thread (vehicle_control_b200/pipeline.py), when there are fewer
 * cores than waiting threads (8 ranks x 4 lanes on one box). */
int vc_set_blocking_sync(vc_handle* h, int on);

/* Measurement hooks (no reference counterpart; the reference times command() with time.time(), racing.py:231-233):
 * with profiling on, vc_nmpc_solve brackets every kernel of its iteration loop with CUDA events on the handle's
 * stream.  vc_get_profile: accumulated ms and launch counts of k_eval_st, k_eval_pm, k_step since
 * vc_set_profiling, and the number of interior-point passes of the last solve. */
int vc_set_profiling(vc_handle* h, int on);
int vc_get_profile(vc_handle* h, double* out_ms3, long long* out_launches3, int* out_last_passes);
/* sum over the profiled passes of the number of unfinished instances each pass processed (units of the roofline) */
long long vc_get_profile_instance_passes(vc_handle* h);

/* Diagnostics: copies n doubles of solver-workspace array `array` (enum WsArray in csrc/vc_solver.cu; -1 = scalar
 * block) of instance b of the last vc_nmpc_solve to the host.  Synchronises. */
int vc_nmpc_debug_read(vc_handle* h, int b, int array, double* h_out, int n);

/* bytes of device workspace currently held by the handle for vc_nmpc_solve */
long long vc_workspace_bytes(vc_handle* h);

/* number of kernels this handle has launched so far (bench bookkeeping) */
long long vc_launch_count(vc_handle* h);

#ifdef __cplusplus
}
#endif
#endif /* VC_API_H */
