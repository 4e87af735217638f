// Batched dynamics kernels for sm_100a: SingleTrack RK4 rollout (+ exact Jacobians), PointMass Euler
// rollout (+ Jacobians), track-spline lookups / horizon parameters, closed-loop plant step.
// All arithmetic fp64; none of it is a dense contraction, so no tensor cores.
#include "vc_handle.h"
#include "vc_model.cuh"
#include "vc_jac.cuh"
#include <cstdlib>
#ifndef VC_JAC_K_DEFAULT
#define VC_JAC_K_DEFAULT 0
#endif

namespace vc {

// ------------------------------------------------------------------------------------------------
// K1a: SingleTrack spatial rollout, states only.  One thread per trajectory.
// replaces car.spatial_transition chained N times (dynamic_car.py:188-191)
__global__ void __launch_bounds__(128)
k_rollout_st(Params p, int B, int N, const double* __restrict__ x0, const double* __restrict__ U,
             const double* __restrict__ kappa, const double* __restrict__ ds, double* __restrict__ X) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double x[8], xn[8], u[2];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = x0[(size_t)b * 8 + i];
    double* Xb = X + (size_t)b * (N + 1) * 8;
#pragma unroll
    for (int i = 0; i < 8; ++i) Xb[i] = x[i];
    for (int n = 0; n < N; ++n) {
        const size_t bn = (size_t)b * N + n;
        u[0] = U[bn * 2]; u[1] = U[bn * 2 + 1];
        st_rk4<true>(p, x, u, kappa[bn], ds[bn], xn);
#pragma unroll
        for (int i = 0; i < 8; ++i) { x[i] = xn[i]; Xb[(n + 1) * 8 + i] = xn[i]; }
    }
}

// K1b: SingleTrack spatial rollout with [A|B] = d x+ / d(x,u) (8x10 per step).
// Eight lanes per trajectory; lane j carries the tangent along active input j
// (Ux Uy r delta ey epsi | Fx w); columns s and t of A are unit vectors (f does not depend on s,t).
// The 8x10 block is staged in shared memory so each trajectory-step is written as one 640-byte run.
#define K1_TPB 128
#define K1_GROUPS (K1_TPB / 8)
__global__ void __launch_bounds__(K1_TPB)
k_rollout_st_jac(Params p, int B, int N, const double* __restrict__ x0, const double* __restrict__ U,
                 const double* __restrict__ kappa, const double* __restrict__ ds, double* __restrict__ X,
                 double* __restrict__ AB) {
    __shared__ double tile[K1_GROUPS][80];
    const int g = threadIdx.x >> 3, j = threadIdx.x & 7;
    const int b = blockIdx.x * K1_GROUPS + g;
    const bool live = b < B;
    const int bb = live ? b : B - 1;
    // column of [A|B] owned by this lane
    const int col = j < 4 ? j : (j < 6 ? j + 1 : j + 2);   // 0 1 2 3 5 6 8 9
    double xv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) xv[i] = x0[(size_t)bb * 8 + i];
    double* Xb = X + (size_t)bb * (N + 1) * 8;
    if (live) Xb[j] = xv[j];
    for (int n = 0; n < N; ++n) {
        const size_t bn = (size_t)bb * N + n;
        Dual1 x[8], u[2], xn[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = Dual1{xv[i], (col == i) ? 1.0 : 0.0};
        u[0] = Dual1{U[bn * 2], col == 8 ? 1.0 : 0.0};
        u[1] = Dual1{U[bn * 2 + 1], col == 9 ? 1.0 : 0.0};
        st_rk4<true>(p, x, u, kappa[bn], ds[bn], xn);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            xv[i] = xn[i].v;
            tile[g][i * 10 + col] = xn[i].d;
        }
        if (j == 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { tile[g][i * 10 + 4] = (i == 4) ? 1.0 : 0.0; tile[g][i * 10 + 7] = (i == 7) ? 1.0 : 0.0; }
        }
        __syncwarp();
        if (live) {
            Xb[(n + 1) * 8 + j] = xv[j];
            double* out = AB + bn * 80;
#pragma unroll
            for (int q = 0; q < 10; ++q) out[q * 8 + j] = tile[g][q * 8 + j];
        }
        __syncwarp();
    }
}

// K1c: the same with K Jacobian columns per lane (DualK<K>): 8/K lanes per trajectory, so the state values and every
// transcendental are evaluated 8/K times per step instead of 8 times.
template <int K>
__global__ void __launch_bounds__(K1_TPB)
k_rollout_st_jac_k(Params p, int B, int N, const double* __restrict__ x0, const double* __restrict__ U,
                   const double* __restrict__ kappa, const double* __restrict__ ds, double* __restrict__ X,
                   double* __restrict__ AB) {
    constexpr int L = 8 / K, GROUPS = K1_TPB / L;
    __shared__ double tile[GROUPS][80];
    const int g = threadIdx.x / L, j = threadIdx.x % L;
    const int b = blockIdx.x * GROUPS + g;
    const bool live = b < B;
    const int bb = live ? b : B - 1;
    int cols[K];   // columns of [A|B] owned by this lane, out of 0 1 2 3 5 6 8 9
#pragma unroll
    for (int c = 0; c < K; ++c) { const int a = j * K + c; cols[c] = a < 4 ? a : (a < 6 ? a + 1 : a + 2); }
    double xv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) xv[i] = x0[(size_t)bb * 8 + i];
    double* Xb = X + (size_t)bb * (N + 1) * 8;
    if (live)
        for (int i = j; i < 8; i += L) Xb[i] = xv[i];
    for (int n = 0; n < N; ++n) {
        const size_t bn = (size_t)bb * N + n;
        DualK<K> x[8], u[2], xn[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            x[i].v = xv[i];
#pragma unroll
            for (int c = 0; c < K; ++c) x[i].d[c] = cols[c] == i ? 1.0 : 0.0;
        }
        u[0].v = U[bn * 2]; u[1].v = U[bn * 2 + 1];
#pragma unroll
        for (int c = 0; c < K; ++c) { u[0].d[c] = cols[c] == 8 ? 1.0 : 0.0; u[1].d[c] = cols[c] == 9 ? 1.0 : 0.0; }
        st_rk4<true>(p, x, u, kappa[bn], ds[bn], xn);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            xv[i] = xn[i].v;
#pragma unroll
            for (int c = 0; c < K; ++c) tile[g][i * 10 + cols[c]] = xn[i].d[c];
        }
        if (j == 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { tile[g][i * 10 + 4] = (i == 4) ? 1.0 : 0.0; tile[g][i * 10 + 7] = (i == 7) ? 1.0 : 0.0; }
        }
        __syncwarp();
        if (live) {
            for (int i = j; i < 8; i += L) Xb[(n + 1) * 8 + i] = xv[i];
            double* out = AB + bn * 80;
#pragma unroll
            for (int q = 0; q < 80 / L; ++q) out[q * L + j] = tile[g][q * L + j];
        }
        __syncwarp();
    }
}

// K1d: the same outputs from the hand-derived Jacobian of the right-hand side (vc_jac.cuh) chained through the four RK4
// stages.  One thread per trajectory: the 39 non-zero partials of one right-hand-side evaluation live in registers, the
// sensitivity D = d(active states)/d(x0 active, u) (6 x 8) and the weighted sum of the stage Jacobians (7 x 8) live in
// shared memory, element-major with the thread index fastest (stride KA_LD: conflict-free for the owner thread and for
// the warp-cooperative read-out).  After the fourth stage the warp writes each trajectory's 8x10 block as one 640-byte
// run.  About a third of the arithmetic of the dual-number kernels: no value or transcendental is evaluated twice.
#define KA_TPB 64
constexpr int KA_LD = KA_TPB + 1;
constexpr int KA_ROWS = 48 + 56;   // D, acc
__global__ void __launch_bounds__(KA_TPB)
k_rollout_st_jac_an(Params p, int B, int N, const double* __restrict__ x0, const double* __restrict__ U,
                    const double* __restrict__ kappa, const double* __restrict__ ds, double* __restrict__ X,
                    double* __restrict__ AB) {
    extern __shared__ double ka_sm[];
    const int tid = threadIdx.x, lane = tid & 31, wbase = tid & ~31;
    double* sD = ka_sm + tid;                       // element e of this thread: sD[e * KA_LD]
    double* sA = ka_sm + 48 * KA_LD + tid;
    const int b = blockIdx.x * KA_TPB + tid;
    const bool live = b < B;
    const int bb = live ? b : B - 1;
    double x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = x0[(size_t)bb * 8 + i];
    double* Xb = X + (size_t)bb * (N + 1) * 8;
    if (live) {
#pragma unroll
        for (int i = 0; i < 8; ++i) Xb[i] = x[i];
    }
    const size_t blk0 = (size_t)blockIdx.x * KA_TPB + wbase;   // first trajectory of this warp
    const int nt = (int)(blk0 >= (size_t)B ? 0 : ((size_t)B - blk0 < 32 ? (size_t)B - blk0 : 32));   // live trajectories of the warp
    size_t bn = (size_t)bb * N;
    double u0 = N > 0 ? U[bn * 2] : 0.0, u1 = N > 0 ? U[bn * 2 + 1] : 0.0, kap = N > 0 ? kappa[bn] : 0.0, h = N > 0 ? ds[bn] : 0.0;
    for (int n = 0; n < N; ++n, ++bn) {
        StCtl ct;
        st_ctl(p, u0, ct);
        const double w = u1, kp = kap, hh = h;
        if (n + 1 < N) {   // next step's inputs travel while this step computes
            u0 = U[(bn + 1) * 2]; u1 = U[(bn + 1) * 2 + 1]; kap = kappa[bn + 1]; h = ds[bn + 1];
        }
        double xs[8], f[8], facc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) xs[i] = x[i];
#pragma unroll 1
        for (int s = 0; s < 4; ++s) {
            StRhsJac J;
            st_rhs_spatial_jac(p, ct, xs, w, kp, f, J);
            const double wgt = (s == 0 || s == 3) ? 1.0 : 2.0;
            const double cs = s == 2 ? hh : 0.5 * hh;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (s == 0) facc[i] = f[i]; else facc[i] = facc[i] + wgt * f[i];
                xs[i] = x[i] + cs * f[i];
            }
            if (s == 0) {   // D = [I | 0]: K = J
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    double k[7];
                    st_jac_unitcol(J, c, k);
#pragma unroll
                    for (int i = 0; i < 7; ++i) sA[(i * 8 + c) * KA_LD] = k[i];
#pragma unroll
                    for (int i = 0; i < 6; ++i) sD[(c * 6 + i) * KA_LD] = (c == i ? 1.0 : 0.0) + cs * k[i];
                }
            } else {
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    double d[6], k[7];
#pragma unroll
                    for (int i = 0; i < 6; ++i) d[i] = sD[(c * 6 + i) * KA_LD];
                    st_jac_col(J, d, c < 6 ? 0 : c - 5, k);
#pragma unroll
                    for (int i = 0; i < 7; ++i) sA[(i * 8 + c) * KA_LD] += wgt * k[i];
                    if (s < 3) {
#pragma unroll
                        for (int i = 0; i < 6; ++i) sD[(c * 6 + i) * KA_LD] = (c == i ? 1.0 : 0.0) + cs * k[i];
                    }
                }
            }
        }
        const double h6 = hh * (1.0 / 6.0);
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = x[i] + h6 * facc[i];
        if (live) {
#pragma unroll
            for (int i = 0; i < 8; ++i) Xb[(n + 1) * 8 + i] = x[i];
        }
        __syncwarp();
        // read-out: the warp writes each of its trajectories' 8x10 block as one 640-byte run (acc rows are in output order:
        // entry (row, col) of the block = acc row i * 8 + c, or a unit-vector entry for row s / columns s, t)
        int o_src[3];
        double o_unit[3];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const int o = lane + 32 * q, row = o / 10, col = o - row * 10;
            const int i = row < 4 ? row : (row == IS ? -1 : row - 1);                    // Ux Uy r delta | ey epsi t -> 0..6
            const int c = col < 4 ? col : ((col == IS || col == IT) ? -1 : (col < IT ? col - 1 : col - 2));   // -> 0..7
            o_src[q] = (o < 80 && i >= 0 && c >= 0) ? (i * 8 + c) * KA_LD : -1;
            o_unit[q] = row == col ? 1.0 : 0.0;
        }
        const double* a = ka_sm + 48 * KA_LD + wbase;
        double* out = AB + ((size_t)blk0 * N + n) * 80 + lane;
#pragma unroll 4
        for (int t = 0; t < nt; ++t) {
            const double ht = __shfl_sync(0xffffffffu, h6, t);
#pragma unroll
            for (int q = 0; q < 3; ++q)
                if (lane + 32 * q < 80)
                    out[(size_t)t * N * 80 + 32 * q] = o_src[q] < 0 ? o_unit[q] : o_unit[q] + ht * a[o_src[q] + t];
        }
        __syncwarp();
    }
}

// K2: PointMass spatial Euler rollout, optional [A|B] (5x7 per step).  One thread per trajectory.
// replaces point_mass.spatial_transition (dynamic_point_mass.py:93-103)
__global__ void __launch_bounds__(128)
k_rollout_pm(Params p, int B, int M, const double* __restrict__ x0, const double* __restrict__ U,
             const double* __restrict__ kappa, const double* __restrict__ ds, double* __restrict__ X,
             double* __restrict__ AB) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double x[5], xn[5], u[2];
#pragma unroll
    for (int i = 0; i < 5; ++i) x[i] = x0[(size_t)b * 5 + i];
    double* Xb = X + (size_t)b * (M + 1) * 5;
#pragma unroll
    for (int i = 0; i < 5; ++i) Xb[i] = x[i];
    for (int n = 0; n < M; ++n) {
        const size_t bn = (size_t)b * M + n;
        u[0] = U[bn * 2]; u[1] = U[bn * 2 + 1];
        const double kp = kappa[bn], h = ds[bn];
        pm_euler<true>(p, x, u, kp, h, xn);
        if (AB) {
            double* out = AB + bn * 35;
            for (int c = 0; c < 7; ++c) {
                Dual1 xd[5], ud[2], xo[5];
#pragma unroll
                for (int i = 0; i < 5; ++i) xd[i] = Dual1{x[i], c == i ? 1.0 : 0.0};
                ud[0] = Dual1{u[0], c == 5 ? 1.0 : 0.0};
                ud[1] = Dual1{u[1], c == 6 ? 1.0 : 0.0};
                pm_euler<true>(p, xd, ud, kp, h, xo);
#pragma unroll
                for (int i = 0; i < 5; ++i) out[i * 7 + c] = xo[i].d;
            }
        }
#pragma unroll
        for (int i = 0; i < 5; ++i) { x[i] = xn[i]; Xb[(n + 1) * 5 + i] = xn[i]; }
    }
}

// K3: horizon parameters (cascaded_mpc.py:316-338).  Persistent CTAs stage the curvature-spline
// coefficient table in shared memory once, then sweep instances.
__global__ void __launch_bounds__(256)
k_horizon_params(int B, int N, int M, double mpc_dt, double ds_pm, const double* __restrict__ kc, UniformKnots kt,
                 int use_smem, const double* __restrict__ x0, const double* __restrict__ Xprev,
                 double* __restrict__ ds, double* __restrict__ kappa) {
    extern __shared__ double s_c[];
    const double* c = kc;
    if (use_smem) {
        for (int i = threadIdx.x; i < kt.n; i += blockDim.x) s_c[i] = kc[i];
        __syncthreads();
        c = s_c;
    }
    const int H = N + M;
    for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < B; b += gridDim.x * blockDim.x) {
        const double* ux = Xprev + (size_t)b * 8 * H;   // row 0 (Ux) of state_prediction[8,H]
        double* dsb = ds + (size_t)b * H;
        double* kb = kappa + (size_t)b * H;
        const double s0 = x0[(size_t)b * 8 + IS];
        double cum = 0.0, d0 = 0.0, dlast = 0.0, slast = s0;
        for (int n = 0; n < N; ++n) {
            const double d = mpc_dt * ux[n];
            if (n == 0) d0 = d;
            cum += d;
            const double s = cum - d0 + s0;
            dsb[n] = d;
            kb[n] = bspline3(c, kt, s, nullptr);
            dlast = d; slast = s;
        }
        double cum_pm = 0.0;
        for (int m = 0; m < M; ++m) {
            cum_pm += ds_pm;
            const double s = cum_pm - dlast + slast;
            dsb[N + m] = ds_pm;
            kb[N + m] = bspline3(c, kt, s, nullptr);
        }
    }
}

__global__ void k_curvature(int n, const double* __restrict__ kc, UniformKnots kt, const double* __restrict__ s,
                            double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = bspline3(kc, kt, s[i], nullptr);
}

// rel2glob (track.py:102-129): orientation from the spline tangent at fmod(s, L)
__global__ void k_rel2glob(int n, const double* __restrict__ xc, const double* __restrict__ yc, UniformKnots kt,
                           double length, const double* __restrict__ s, const double* __restrict__ ey,
                           const double* __restrict__ epsi, double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double scale = (double)kt.n / length;
    const double sv = s[i];
    const double sm = fmod(sv, length);
    double dx, dy, tmp;
    bspline3(xc, kt, sm * scale, &dx);
    bspline3(yc, kt, sm * scale, &dy);
    const double mag = sqrt(dx * dx + dy * dy);
    const double th = atan2(dy / mag, dx / mag);
    const double x = bspline3(xc, kt, sv * scale, &tmp);
    const double y = bspline3(yc, kt, sv * scale, &tmp);
    double sn, cs;
    sincos(th, &sn, &cs);
    double psi = th + epsi[i];
    const double pi = 3.14159265358979323846;
    if (psi < -pi) psi = 2 * pi + psi; else if (psi > pi) psi = psi - 2 * pi;   // common_utils.py:29-37
    out[(size_t)i * 3 + 0] = x - sn * ey[i];
    out[(size_t)i * 3 + 1] = y + cs * ey[i];
    out[(size_t)i * 3 + 2] = psi;
}

// K5: closed-loop plant step (racing_car.py:34-46): kappa = k(s) (no fmod), temporal RK4.
__global__ void __launch_bounds__(128)
k_plant_step(Params p, int B, const double* __restrict__ kc, UniformKnots kt, double* __restrict__ x,
             const double* __restrict__ u, double dt) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double xv[8], xn[8], uv[2];
#pragma unroll
    for (int i = 0; i < 8; ++i) xv[i] = x[(size_t)b * 8 + i];
    uv[0] = u[(size_t)b * 2]; uv[1] = u[(size_t)b * 2 + 1];
    const double kp = bspline3(kc, kt, xv[IS], nullptr);
    st_rk4<false>(p, xv, uv, kp, dt, xn);
#pragma unroll
    for (int i = 0; i < 8; ++i) x[(size_t)b * 8 + i] = xn[i];
}

// ---- per-instance tracks: a track atlas holds the curvature tables of several tracks; every
// instance carries a track id.  Same arithmetic as k_horizon_params / k_plant_step, the table chosen per instance.

__global__ void __launch_bounds__(256)
k_horizon_params_atlas(int B, int N, int M, double mpc_dt, double ds_pm, const double* __restrict__ atlas,
                       const TrackDesc* __restrict__ desc, int n_tracks, const int* __restrict__ track_id,
                       const double* __restrict__ x0, const double* __restrict__ Xprev, double* __restrict__ ds,
                       double* __restrict__ kappa) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int H = N + M;
    const int tid = min(max(track_id[b], 0), n_tracks - 1);
    const TrackDesc td = desc[tid];
    const UniformKnots kt{td.x0, td.dx, td.n};
    const double* c = atlas + td.off;
    const double* ux = Xprev + (size_t)b * 8 * H;
    double* dsb = ds + (size_t)b * H;
    double* kb = kappa + (size_t)b * H;
    const double s0 = x0[(size_t)b * 8 + IS];
    double cum = 0.0, d0 = 0.0, dlast = 0.0, slast = s0;
    for (int n = 0; n < N; ++n) {
        const double d = mpc_dt * ux[n];
        if (n == 0) d0 = d;
        cum += d;
        const double s = cum - d0 + s0;
        dsb[n] = d;
        kb[n] = bspline3(c, kt, s, nullptr);
        dlast = d; slast = s;
    }
    double cum_pm = 0.0;
    for (int m = 0; m < M; ++m) {
        cum_pm += ds_pm;
        const double s = cum_pm - dlast + slast;
        dsb[N + m] = ds_pm;
        kb[N + m] = bspline3(c, kt, s, nullptr);
    }
}

__global__ void __launch_bounds__(128)
k_plant_step_atlas(Params p, int B, const double* __restrict__ atlas, const TrackDesc* __restrict__ desc, int n_tracks,
                   const int* __restrict__ track_id, double* __restrict__ x, const double* __restrict__ u, double dt) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int tid = min(max(track_id[b], 0), n_tracks - 1);
    const TrackDesc td = desc[tid];
    double xv[8], xn[8], uv[2];
#pragma unroll
    for (int i = 0; i < 8; ++i) xv[i] = x[(size_t)b * 8 + i];
    uv[0] = u[(size_t)b * 2]; uv[1] = u[(size_t)b * 2 + 1];
    const double kp = bspline3(atlas + td.off, UniformKnots{td.x0, td.dx, td.n}, xv[IS], nullptr);
    st_rk4<false>(p, xv, uv, kp, dt, xn);
#pragma unroll
    for (int i = 0; i < 8; ++i) x[(size_t)b * 8 + i] = xn[i];
}

__global__ void k_tyre_forces(Params p, int B, const double* __restrict__ x, const double* __restrict__ u,
                              double* __restrict__ out) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    Tyres<double> ty;
    tyres(p, x[(size_t)b * 8 + IUX], x[(size_t)b * 8 + IUY], x[(size_t)b * 8 + IR], x[(size_t)b * 8 + IDELTA],
          u[(size_t)b * 2], ty);
    double* o = out + (size_t)b * 6;
    o[0] = ty.fxf; o[1] = ty.fxr; o[2] = ty.fyf; o[3] = ty.fyr; o[4] = atan(ty.tan_f); o[5] = atan(ty.tan_r);   // alpha_f, alpha_r (dynamic_car.py:111-115)
}

}  // namespace vc

// =================================================================================================
// C-ABI (include/vc_api.h)
// =================================================================================================
namespace vc {
static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    set_error(std::string(what) + ": " + cudaGetErrorString(e));
    return -2;
}
static int bad(const char* msg) { set_error(msg); return -1; }
static UniformKnots kknots(const vc_handle* h) { return UniformKnots{h->k_x0, h->k_step, h->n_k}; }
static UniformKnots xyknots(const vc_handle* h) { return UniformKnots{0.0, 1.0, h->n_xy}; }
}  // namespace vc

using namespace vc;

#define VC_REQUIRE(c, msg) do { if (!(c)) return bad(msg); } while (0)
#define VC_LAUNCH_CHECK(h, what) do { (h)->launches++; int rc_ = check_cuda(cudaGetLastError(), what); if (rc_) return rc_; } while (0)

extern "C" {

const char* vc_last_error(void) { return g_err.c_str(); }
int vc_version(void) { return 100; }

int vc_create(const vc_model_params* hp, const vc_mpc_config* hc, int device, void* stream, vc_handle** out) {
    VC_REQUIRE(hp && out, "vc_create: null argument");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) {
        set_error("vc_create: no CUDA device (this library has no CPU fallback)");
        return -3;
    }
    VC_REQUIRE(device >= 0 && device < ndev, "vc_create: bad device index");
    int rc = check_cuda(cudaSetDevice(device), "cudaSetDevice");
    if (rc) return rc;
    vc_handle* h = new vc_handle();
    h->device = device;
    h->stream = (cudaStream_t)stream;
    h->params = *hp;
    if (hc) h->cfg = *hc;
    cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device);
    cudaDeviceGetAttribute(&h->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    *out = h;
    return 0;
}

int vc_destroy(vc_handle* h) {
    if (!h) return 0;
    cudaSetDevice(h->device);
    cudaFree(h->d_kc); cudaFree(h->d_xc); cudaFree(h->d_yc); cudaFree(h->d_scratch);
    cudaFree(h->d_atlas); cudaFree(h->d_atlas_desc); cudaFree(h->d_bt);
    cudaFree(h->d_ws); cudaFree(h->d_active);
    if (h->h_active) cudaFreeHost(h->h_active);
    if (h->sync_event) cudaEventDestroy(h->sync_event);
    delete h;
    return 0;
}

int vc_default_config(vc_mpc_config* c) {
    VC_REQUIRE(c, "vc_default_config: null argument");
    // IPOPT 3.14 defaults, which the reference leaves untouched (cascaded_mpc.py:55-65)
    c->max_iter = 3000; c->tol = 1e-8; c->acceptable_tol = 1e-6; c->acceptable_iter = 15; c->obj_scaling = 1;
    c->mu_init = 0.1; c->bound_relax = 1e-8; c->max_soc = 4; c->warm_rungs = 0;
    c->watchdog_trigger = 10; c->watchdog_trials = 3;
    // engine policies
    c->max_damp = 8; c->wide_max = 296;
    return 0;
}

int vc_set_config(vc_handle* h, const vc_mpc_config* hc) {
    VC_REQUIRE(h && hc, "vc_set_config: null argument");
    VC_REQUIRE(hc->N >= 1 && hc->M >= 0, "vc_set_config: need horizon N >= 1, M >= 0");
    VC_REQUIRE(hc->max_iter >= 1 && hc->tol > 0 && hc->mu_init > 0 && hc->max_damp >= 0 && hc->watchdog_trials >= 0,
               "vc_set_config: solver options not set (start from vc_default_config)");
    h->cfg = *hc;
    return 0;
}

int vc_set_track(vc_handle* h, const double* kc, int n_k, double k_x0, double k_step, const double* xc,
                 const double* yc, int n_xy, double length) {
    VC_REQUIRE(h && kc && n_k >= 8 && k_step > 0, "vc_set_track: bad curvature table");
    cudaSetDevice(h->device);
    cudaFree(h->d_kc); cudaFree(h->d_xc); cudaFree(h->d_yc);
    h->d_kc = h->d_xc = h->d_yc = nullptr;
    int rc = check_cuda(cudaMalloc(&h->d_kc, sizeof(double) * n_k), "cudaMalloc(track)");
    if (rc) return rc;
    rc = check_cuda(cudaMemcpy(h->d_kc, kc, sizeof(double) * n_k, cudaMemcpyHostToDevice), "cudaMemcpy(track)");
    if (rc) return rc;
    h->n_k = n_k; h->k_x0 = k_x0; h->k_step = k_step; h->length = length; h->n_xy = 0;
    if (xc && yc && n_xy >= 8) {
        rc = check_cuda(cudaMalloc(&h->d_xc, sizeof(double) * n_xy), "cudaMalloc(track xy)");
        if (rc) return rc;
        rc = check_cuda(cudaMalloc(&h->d_yc, sizeof(double) * n_xy), "cudaMalloc(track xy)");
        if (rc) return rc;
        cudaMemcpy(h->d_xc, xc, sizeof(double) * n_xy, cudaMemcpyHostToDevice);
        rc = check_cuda(cudaMemcpy(h->d_yc, yc, sizeof(double) * n_xy, cudaMemcpyHostToDevice), "cudaMemcpy(track xy)");
        if (rc) return rc;
        h->n_xy = n_xy;
    }
    return 0;
}

int vc_track_curvature(vc_handle* h, int n, const double* d_s, double* d_kappa) {
    VC_REQUIRE(h && h->d_kc, "vc_track_curvature: track not set");
    if (n <= 0) return 0;
    k_curvature<<<(n + 255) / 256, 256, 0, h->stream>>>(n, h->d_kc, kknots(h), d_s, d_kappa);
    VC_LAUNCH_CHECK(h, "k_curvature");
    return 0;
}

int vc_rel2glob(vc_handle* h, int n, const double* d_s, const double* d_ey, const double* d_epsi, double* d_out) {
    VC_REQUIRE(h && h->d_xc && h->d_yc, "vc_rel2glob: x/y spline tables not set");
    if (n <= 0) return 0;
    k_rel2glob<<<(n + 255) / 256, 256, 0, h->stream>>>(n, h->d_xc, h->d_yc, xyknots(h), h->length, d_s, d_ey,
                                                        d_epsi, d_out);
    VC_LAUNCH_CHECK(h, "k_rel2glob");
    return 0;
}

int vc_horizon_params(vc_handle* h, int B, const double* d_x0, const double* d_Xprev, double* d_ds, double* d_kappa) {
    VC_REQUIRE(h && h->d_kc, "vc_horizon_params: track not set");
    VC_REQUIRE(h->cfg.N >= 1, "vc_horizon_params: config not set");
    if (B <= 0) return 0;
    const size_t smem = sizeof(double) * (size_t)h->n_k;
    // stage the table in shared memory when it fits and there is enough work to amortise the fill
    int use_smem = (smem <= (size_t)h->max_smem_optin && B >= 4 * h->sm_count) ? 1 : 0;
    int grid = use_smem ? h->sm_count : (B + 255) / 256;
    if (use_smem) {
        int rc = check_cuda(cudaFuncSetAttribute(k_horizon_params, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                            "cudaFuncSetAttribute(k_horizon_params)");
        if (rc) return rc;
    }
    k_horizon_params<<<grid, 256, use_smem ? smem : 0, h->stream>>>(B, h->cfg.N, h->cfg.M, h->cfg.mpc_dt, h->cfg.ds_pm,
                                                                  h->d_kc, kknots(h), use_smem, d_x0, d_Xprev, d_ds,
                                                                  d_kappa);
    VC_LAUNCH_CHECK(h, "k_horizon_params");
    return 0;
}

int vc_rollout_st(vc_handle* h, int B, int N, const double* d_x0, const double* d_U, const double* d_kappa,
                  const double* d_ds, double* d_X, double* d_AB) {
    VC_REQUIRE(h && N >= 0, "vc_rollout_st: bad argument");
    if (B <= 0) return 0;
    if (d_AB) {
        static const int jac_k = getenv("VC_JAC_K") ? atoi(getenv("VC_JAC_K")) : VC_JAC_K_DEFAULT;   // Jacobian columns per lane
        if (jac_k == 0) {   // hand-derived Jacobian (default)
            constexpr size_t smem = sizeof(double) * KA_ROWS * KA_LD;
            int rc = check_cuda(cudaFuncSetAttribute(k_rollout_st_jac_an, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                                "cudaFuncSetAttribute(k_rollout_st_jac_an)");
            if (rc) return rc;
            k_rollout_st_jac_an<<<(B + KA_TPB - 1) / KA_TPB, KA_TPB, smem, h->stream>>>(h->params, B, N, d_x0, d_U, d_kappa, d_ds, d_X, d_AB);
        } else if (jac_k == 2) {
            constexpr int G = K1_TPB / 4;
            k_rollout_st_jac_k<2><<<(B + G - 1) / G, K1_TPB, 0, h->stream>>>(h->params, B, N, d_x0, d_U, d_kappa, d_ds, d_X, d_AB);
        } else if (jac_k == 4) {
            constexpr int G = K1_TPB / 2;
            k_rollout_st_jac_k<4><<<(B + G - 1) / G, K1_TPB, 0, h->stream>>>(h->params, B, N, d_x0, d_U, d_kappa, d_ds, d_X, d_AB);
        } else {
            k_rollout_st_jac<<<(B + K1_GROUPS - 1) / K1_GROUPS, K1_TPB, 0, h->stream>>>(h->params, B, N, d_x0, d_U, d_kappa,
                                                                                     d_ds, d_X, d_AB);
        }
    } else {
        k_rollout_st<<<(B + 127) / 128, 128, 0, h->stream>>>(h->params, B, N, d_x0, d_U, d_kappa, d_ds, d_X);
    }
    VC_LAUNCH_CHECK(h, "k_rollout_st");
    return 0;
}

int vc_rollout_pm(vc_handle* h, int B, int M, const double* d_x0, const double* d_U, const double* d_kappa,
                  const double* d_ds, double* d_X, double* d_AB) {
    VC_REQUIRE(h && M >= 0, "vc_rollout_pm: bad argument");
    if (B <= 0) return 0;
    k_rollout_pm<<<(B + 127) / 128, 128, 0, h->stream>>>(h->params, B, M, d_x0, d_U, d_kappa, d_ds, d_X, d_AB);
    VC_LAUNCH_CHECK(h, "k_rollout_pm");
    return 0;
}

int vc_plant_step(vc_handle* h, int B, double* d_x, const double* d_u, double dt) {
    VC_REQUIRE(h && h->d_kc, "vc_plant_step: track not set");
    if (B <= 0) return 0;
    k_plant_step<<<(B + 127) / 128, 128, 0, h->stream>>>(h->params, B, h->d_kc, kknots(h), d_x, d_u, dt);
    VC_LAUNCH_CHECK(h, "k_plant_step");
    return 0;
}

int vc_set_track_atlas(vc_handle* h, int n_tracks, const double* h_coefs, const int* h_offsets, const double* h_x0,
                       const double* h_step) {
    VC_REQUIRE(h && n_tracks >= 1 && h_coefs && h_offsets && h_x0 && h_step, "vc_set_track_atlas: bad argument");
    cudaSetDevice(h->device);
    cudaFree(h->d_atlas); cudaFree(h->d_atlas_desc);
    h->d_atlas = nullptr; h->d_atlas_desc = nullptr; h->n_atlas = 0;
    std::vector<TrackDesc> desc(n_tracks);
    for (int t = 0; t < n_tracks; ++t) {
        const int n = h_offsets[t + 1] - h_offsets[t];
        VC_REQUIRE(n >= 8 && h_step[t] > 0, "vc_set_track_atlas: bad curvature table");
        desc[t] = TrackDesc{h_x0[t], h_step[t], h_offsets[t], n};
    }
    const size_t total = (size_t)h_offsets[n_tracks];
    int rc = check_cuda(cudaMalloc(&h->d_atlas, sizeof(double) * total), "cudaMalloc(track atlas)");
    if (rc) return rc;
    rc = check_cuda(cudaMalloc(&h->d_atlas_desc, sizeof(TrackDesc) * n_tracks), "cudaMalloc(track atlas)");
    if (rc) return rc;
    cudaMemcpy(h->d_atlas, h_coefs, sizeof(double) * total, cudaMemcpyHostToDevice);
    rc = check_cuda(cudaMemcpy(h->d_atlas_desc, desc.data(), sizeof(TrackDesc) * n_tracks, cudaMemcpyHostToDevice),
                    "cudaMemcpy(track atlas)");
    if (rc) return rc;
    h->n_atlas = n_tracks;
    return 0;
}

int vc_horizon_params_tracks(vc_handle* h, int B, const int* d_track_id, const double* d_x0, const double* d_Xprev,
                             double* d_ds, double* d_kappa) {
    VC_REQUIRE(h && h->n_atlas > 0 && d_track_id, "vc_horizon_params_tracks: track atlas not set");
    VC_REQUIRE(h->cfg.N >= 1, "vc_horizon_params_tracks: config not set");
    if (B <= 0) return 0;
    k_horizon_params_atlas<<<(B + 255) / 256, 256, 0, h->stream>>>(B, h->cfg.N, h->cfg.M, h->cfg.mpc_dt, h->cfg.ds_pm,
                                                                 h->d_atlas, (const TrackDesc*)h->d_atlas_desc,
                                                                 h->n_atlas, d_track_id, d_x0, d_Xprev, d_ds, d_kappa);
    VC_LAUNCH_CHECK(h, "k_horizon_params_atlas");
    return 0;
}

int vc_plant_step_tracks(vc_handle* h, int B, const int* d_track_id, double* d_x, const double* d_u, double dt) {
    VC_REQUIRE(h && h->n_atlas > 0 && d_track_id, "vc_plant_step_tracks: track atlas not set");
    if (B <= 0) return 0;
    k_plant_step_atlas<<<(B + 127) / 128, 128, 0, h->stream>>>(h->params, B, h->d_atlas,
                                                            (const TrackDesc*)h->d_atlas_desc, h->n_atlas, d_track_id,
                                                            d_x, d_u, dt);
    VC_LAUNCH_CHECK(h, "k_plant_step_atlas");
    return 0;
}

int vc_tyre_forces(vc_handle* h, int B, const double* d_x, const double* d_u, double* d_out) {
    VC_REQUIRE(h, "vc_tyre_forces: null handle");
    if (B <= 0) return 0;
    k_tyre_forces<<<(B + 127) / 128, 128, 0, h->stream>>>(h->params, B, d_x, d_u, d_out);
    VC_LAUNCH_CHECK(h, "k_tyre_forces");
    return 0;
}

long long vc_launch_count(vc_handle* h) { return h ? h->launches : 0; }

}  // extern "C"
