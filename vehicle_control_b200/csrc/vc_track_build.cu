// Track CONSTRUCTION on the device for a batch of generated tracks ("track segments" as an instance
// dimension).  Restates, per track:
//   Track._construct_path        vehicle_control/environment/track.py:254-296  corner polygon -> waypoints every `resolution`
//                                                                        m -> centred moving average of half-width `smoothing`
//   Track._construct_spline      :206-252  not-a-knot cubic B-spline through the waypoints (sites 0..n-1), track length =
//                                          trapezoid of |d(x,y)/du| over the sites
//   Track.get_curvature          :109-119  |x' y'' - x'' y'| / (x'^2 + y'^2)^1.5 of the length-parametrised spline
//   Track._precompute_curvatures :156-167  table of the curvature every 0.05 m, interpolated again by a cubic B-spline = Track.k
//   Track._construct_obstacles   :131-138  (s, ey) -> (cx, cy) through rel2glob
// The result is installed as the handle's track atlas (vc_horizon_params_tracks / vc_plant_step_tracks index it by
// track id), so a batch can carry one generated track per instance without a round trip to the host.
//
// Parallelisation: waypoints, smoothing windows and curvature samples are one thread per point; each interpolation system
// (collocation matrix of the not-a-knot B-spline: tridiagonal but for one extra entry in the second and the second-last row)
// is eliminated by one thread -- a few thousand sequential steps, setup work, not the hot path.
#include <cmath>
#include <vector>
#include "vc_handle.h"
#include "vc_model.cuh"

namespace vc {

constexpr double CURV_STEP = 0.05;     // track.py:157
constexpr int MAX_CORNERS = 64;

struct BuildIn { const double* corners; const int* n_corners; const double* resolution; const int* smoothing; int max_corners; };

// number of waypoints of a track: sum over the polygon's edges of int(distance / resolution)  (track.py:262-269)
__global__ void k_tb_count(int T, BuildIn in, int* __restrict__ n_wp) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    const double* c = in.corners + (size_t)t * in.max_corners * 2;
    int n = 0;
    for (int i = 0; i + 1 < in.n_corners[t]; ++i) {
        const double dx = c[2 * (i + 1)] - c[2 * i], dy = c[2 * (i + 1) + 1] - c[2 * i + 1];
        n += (int)(sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy))) / in.resolution[t]);
    }
    n_wp[t] = n;
}

// raw waypoints: np.linspace(start, end, n, endpoint=False) of every edge = start + j * ((end - start) / n), no contraction
__global__ void k_tb_waypoints(int T, BuildIn in, const int* __restrict__ wp_off, const int* __restrict__ n_wp,
                               double* __restrict__ rx, double* __restrict__ ry) {
    const int t = blockIdx.y;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T || j >= n_wp[t]) return;
    const double* c = in.corners + (size_t)t * in.max_corners * 2;
    int first = 0;
    for (int i = 0; i + 1 < in.n_corners[t]; ++i) {
        const double dx = c[2 * (i + 1)] - c[2 * i], dy = c[2 * (i + 1) + 1] - c[2 * i + 1];
        const int cnt = (int)(sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy))) / in.resolution[t]);
        if (j < first + cnt) {
            const double q = (double)(j - first);
            rx[wp_off[t] + j] = __dadd_rn(__dmul_rn(q, dx / (double)cnt), c[2 * i]);
            ry[wp_off[t] + j] = __dadd_rn(__dmul_rn(q, dy / (double)cnt), c[2 * i + 1]);
            return;
        }
        first += cnt;
    }
}

// numpy's pairwise summation (np.add.reduce of a contiguous double array), bit for bit: the reference smooths with
// np.mean over a Python-list window, and the closed-loop golden runs resolve the last bits of the waypoints on the curves
__device__ double np_block_sum(const double* a, int n) {   // numpy's unrolled block, n <= 128
    if (n < 8) {
        double r = 0.0;
        for (int i = 0; i < n; ++i) r = __dadd_rn(r, a[i]);
        return r;
    }
    double r0 = a[0], r1 = a[1], r2 = a[2], r3 = a[3], r4 = a[4], r5 = a[5], r6 = a[6], r7 = a[7];
    int i = 8;
    for (; i < n - (n % 8); i += 8) {
        r0 = __dadd_rn(r0, a[i]); r1 = __dadd_rn(r1, a[i + 1]); r2 = __dadd_rn(r2, a[i + 2]); r3 = __dadd_rn(r3, a[i + 3]);
        r4 = __dadd_rn(r4, a[i + 4]); r5 = __dadd_rn(r5, a[i + 5]); r6 = __dadd_rn(r6, a[i + 6]); r7 = __dadd_rn(r7, a[i + 7]);
    }
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r0, r1), __dadd_rn(r2, r3)), __dadd_rn(__dadd_rn(r4, r5), __dadd_rn(r6, r7)));
    for (; i < n; ++i) res = __dadd_rn(res, a[i]);
    return res;
}
__device__ double np_pairwise_sum(const double* a, int n) {
    // numpy splits n > 128 into (n2, n - n2) with n2 = n / 2 rounded down to a multiple of 8 and adds left + right: the
    // split tree is walked with an explicit stack (no device recursion: its stack cannot be sized statically)
    const double* sa[28];
    int sn[28];
    double sv[28];
    int sst[28];
    int top = 0;
    sa[0] = a; sn[0] = n; sst[0] = 0;
    double ret = 0.0;
    while (top >= 0) {
        if (sst[top] == 0) {
            if (sn[top] <= 128) { ret = np_block_sum(sa[top], sn[top]); --top; continue; }
            int n2 = sn[top] / 2;
            n2 -= n2 % 8;
            sst[top] = 1;
            sa[top + 1] = sa[top]; sn[top + 1] = n2; sst[top + 1] = 0;
            ++top;
        } else if (sst[top] == 1) {          // the left half returned `ret`
            int n2 = sn[top] / 2;
            n2 -= n2 % 8;
            sv[top] = ret; sst[top] = 2;
            sa[top + 1] = sa[top] + n2; sn[top + 1] = sn[top] - n2; sst[top + 1] = 0;
            ++top;
        } else {                             // the right half returned `ret`
            ret = __dadd_rn(sv[top], ret);
            --top;
        }
    }
    return ret;
}

// centred moving average of half-width w away from the two ends (track.py:272-287)
__global__ void k_tb_smooth(int T, const int* __restrict__ smoothing, const int* __restrict__ wp_off, const int* __restrict__ n_wp,
                            const double* __restrict__ rx, const double* __restrict__ ry, double* __restrict__ wx,
                            double* __restrict__ wy) {
    const int t = blockIdx.y;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T || j >= n_wp[t]) return;
    const int w = smoothing[t], n = n_wp[t], o = wp_off[t];
    if (j < w || j > n - w - 1) { wx[o + j] = rx[o + j]; wy[o + j] = ry[o + j]; return; }
    const int m = 2 * w + 1;
    wx[o + j] = np_pairwise_sum(rx + o + j - w, m) / (double)m;
    wy[o + j] = np_pairwise_sum(ry + o + j - w, m) / (double)m;
}

// the four cubic B-spline basis functions that are non-zero on span k, at x (Cox - de Boor): N[r] = B_{k-3+r}(x)
__device__ void basis4(const UniformKnots& t, int k, double x, double* N) {
    double left[4], right[4];
    N[0] = 1.0;
    for (int j = 1; j <= 3; ++j) {
        left[j] = x - t(k + 1 - j);
        right[j] = t(k + j) - x;
        double saved = 0.0;
        for (int r = 0; r < j; ++r) {
            const double tmp = N[r] / (right[r + 1] + left[j - r]);
            N[r] = saved + right[r + 1] * tmp;
            saved = left[j - r] * tmp;
        }
        N[j] = saved;
    }
}

// Not-a-knot cubic B-spline interpolation of y at the uniform sites x0 + i dx (scipy make_interp_spline(k=3) /
// CasADi interpolant("bspline")): solves the collocation system A c = y, A[i][j] = B_j(x_i).  Row 0 / n-1 are unit rows,
// rows 2..n-3 are tridiagonal (B_{i-1..i+1}), row 1 has columns 0..3, row n-2 has columns n-4..n-1: a banded elimination
// with two sub- and two super-diagonals, no pivoting (the rows are diagonally dominant).  band = scratch [n][5].
__device__ void notaknot_fit(int n, double x0, double dx, const double* __restrict__ y, double* __restrict__ c,
                             double* __restrict__ band) {
    const UniformKnots t{x0, dx, n};
    for (int i = 0; i < n; ++i) {
        double* r = band + (size_t)i * 5;        // r[2 + (j - i)] = A[i][j]
        r[0] = r[1] = r[2] = r[3] = r[4] = 0.0;
        const double x = x0 + i * dx;
        int k = i + 2;
        if (k < 3) k = 3;
        if (i <= 1) k = 3;
        if (i >= n - 2) k = n - 1;
        double N[4];
        basis4(t, k, i == n - 1 ? t(n) : x, N);
        for (int q = 0; q < 4; ++q) {
            const int j = k - 3 + q, d = j - i;
            if (d >= -2 && d <= 2) r[2 + d] = N[q];
        }
        c[i] = y[i];
    }
    for (int i = 0; i < n; ++i) {                // forward elimination
        const double* ri = band + (size_t)i * 5;
        const double piv = 1.0 / ri[2];
        for (int m = 1; m <= 2 && i + m < n; ++m) {
            double* rm = band + (size_t)(i + m) * 5;
            const double f = rm[2 - m] * piv;
            if (f == 0.0) continue;
            rm[2 - m] = 0.0;
            for (int d = 1; d <= 2; ++d)
                if (2 - m + d <= 4) rm[2 - m + d] -= f * ri[2 + d];
            c[i + m] -= f * c[i];
        }
    }
    for (int i = n - 1; i >= 0; --i) {           // back substitution
        const double* ri = band + (size_t)i * 5;
        double v = c[i];
        if (i + 1 < n) v -= ri[3] * c[i + 1];
        if (i + 2 < n) v -= ri[4] * c[i + 2];
        c[i] = v / ri[2];
    }
}

// value, first and second derivative of the cubic B-spline with coefficients c at x (de Boor on the derivative splines)
__device__ void bspline3_d2(const double* __restrict__ c, const UniformKnots& t, double x, double& v, double& d1, double& d2) {
    const int n = t.n;
    double jf = floor((x - t.x0) / t.dx);
    jf = jf < 0.0 ? 0.0 : (jf > (double)n ? (double)n : jf);
    int k = (int)jf + 2;
    if (k < 3) k = 3;
    if (k > n - 1) k = n - 1;
    while (k > 3 && x < t(k)) --k;
    while (k < n - 1 && x >= t(k + 1)) ++k;
    v = bspline3(c, t, x, &d1);
    // first-derivative spline (degree 2): q_j = 3 (c_j - c_{j-1}) / (t_{j+3} - t_j), j = k-2..k
    double q[3];
    for (int r = 0; r < 3; ++r) {
        const int j = k - 2 + r;
        q[r] = 3.0 * (c[j] - c[j - 1]) / (t(j + 3) - t(j));
    }
    // second-derivative spline (degree 1): p_j = 2 (q_j - q_{j-1}) / (t_{j+2} - t_j), j = k-1..k
    const double p0 = 2.0 * (q[1] - q[0]) / (t(k + 1) - t(k - 1));
    const double p1 = 2.0 * (q[2] - q[1]) / (t(k + 2) - t(k));
    const double a = (x - t(k)) / (t(k + 1) - t(k));
    d2 = (1.0 - a) * p0 + a * p1;
}

// one thread per (track, coordinate): centre-line spline x(u), y(u) through the smoothed waypoints, sites u = 0..n-1
__global__ void k_tb_fit_xy(int T, const int* __restrict__ wp_off, const int* __restrict__ n_wp, const double* __restrict__ wx,
                            const double* __restrict__ wy, double* __restrict__ cx, double* __restrict__ cy,
                            double* __restrict__ band) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * T) return;
    const int t = i >> 1, o = wp_off[t], n = n_wp[t];
    const double* y = (i & 1) ? wy + o : wx + o;
    double* c = (i & 1) ? cy + o : cx + o;
    // band scratch: [2][total_wp][5]; wp_off[T] = total
    notaknot_fit(n, 0.0, 1.0, y, c, band + ((size_t)(i & 1) * wp_off[T] + o) * 5);
}

// track length: trapezoid over the integer sites of |d(x,y)/du| (track.py:231-241); one warp per track
__global__ void k_tb_length(int T, const int* __restrict__ wp_off, const int* __restrict__ n_wp, const double* __restrict__ cx,
                            const double* __restrict__ cy, double* __restrict__ speed, double* __restrict__ length) {
    const int t = blockIdx.x, lane = threadIdx.x;
    if (t >= T) return;
    const int o = wp_off[t], n = n_wp[t];
    const UniformKnots kt{0.0, 1.0, n};
    for (int j = lane; j < n; j += 32) {
        double dx, dy;
        bspline3(cx + o, kt, (double)j, &dx);
        bspline3(cy + o, kt, (double)j, &dy);
        speed[o + j] = sqrt(dx * dx + dy * dy);
    }
    __syncwarp();
    if (lane == 0) {   // np.trapezoid: sum(d * (y[1:] + y[:-1]) / 2): pairwise sum of the products, as numpy adds them
        double* prod = speed;   // in place: prod[j] = (y[j+1] + y[j]) / 2
        for (int j = 0; j + 1 < n; ++j) prod[o + j] = __dmul_rn(1.0, __dadd_rn(speed[o + j + 1], speed[o + j])) / 2.0;
        length[t] = np_pairwise_sum(prod + o, n - 1);
    }
}

// curvature table: kappa(s_j), s_j = j * 0.05, of the spline parametrised by arc length estimate s = u * L / n
__global__ void k_tb_curvature(int T, const int* __restrict__ wp_off, const int* __restrict__ n_wp, const int* __restrict__ k_off,
                               const int* __restrict__ n_k, const double* __restrict__ length, const double* __restrict__ cx,
                               const double* __restrict__ cy, double* __restrict__ table) {
    const int t = blockIdx.y;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T || j >= n_k[t]) return;
    const int o = wp_off[t], n = n_wp[t];
    const UniformKnots kt{0.0, 1.0, n};
    const double L = length[t], ups = (double)n / L;
    const double s = fmod(j * CURV_STEP, L);
    double v, dx, ddx, dy, ddy;
    bspline3_d2(cx + o, kt, s * ups, v, dx, ddx);
    bspline3_d2(cy + o, kt, s * ups, v, dy, ddy);
    dx *= ups; dy *= ups; ddx *= ups * ups; ddy *= ups * ups;
    table[k_off[t] + j] = fabs(dx * ddy - ddx * dy) / pow(dx * dx + dy * dy, 1.5);
}

// one thread per track: Track.k = not-a-knot cubic B-spline through the curvature table (sites j * 0.05)
__global__ void k_tb_fit_k(int T, const int* __restrict__ k_off, const int* __restrict__ n_k, const double* __restrict__ table,
                           double* __restrict__ atlas, double* __restrict__ band) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    notaknot_fit(n_k[t], 0.0, CURV_STEP, table + k_off[t], atlas + k_off[t], band + (size_t)k_off[t] * 5);
}

// (s, ey) -> (cx, cy) on built track t (Track._construct_obstacles, track.py:131-138 = rel2glob with epsi = 0)
__global__ void k_tb_rel2glob(int n, const double* __restrict__ cx, const double* __restrict__ cy, int n_wp, double length,
                              const double* __restrict__ s, const double* __restrict__ ey, double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const UniformKnots kt{0.0, 1.0, n_wp};
    const double scale = (double)n_wp / length;
    const double sm = fmod(s[i], length);
    double dx, dy, tmp;
    bspline3(cx, kt, sm * scale, &dx);
    bspline3(cy, kt, sm * scale, &dy);
    const double mag = sqrt(dx * dx + dy * dy);
    const double th = atan2(dy / mag, dx / mag);
    double sn, cs;
    sincos(th, &sn, &cs);
    out[(size_t)i * 2] = bspline3(cx, kt, s[i] * scale, &tmp) - sn * ey[i];
    out[(size_t)i * 2 + 1] = bspline3(cy, kt, s[i] * scale, &tmp) + cs * ey[i];
}

}  // namespace vc

using namespace vc;

#define TB_REQUIRE(c, msg) do { if (!(c)) { set_error(msg); return -1; } } while (0)
#define TB_CUDA(e, what) do { int rc_ = check_cuda((e), what); if (rc_) return rc_; } while (0)

extern "C" {

int vc_build_tracks(vc_handle* h, int T, int max_corners, const double* h_corners, const int* h_n_corners,
                    const double* h_resolution, const int* h_smoothing) {
    TB_REQUIRE(h && T >= 1 && h_corners && h_n_corners && h_resolution && h_smoothing, "vc_build_tracks: bad argument");
    TB_REQUIRE(max_corners >= 2 && max_corners <= MAX_CORNERS, "vc_build_tracks: max_corners must be in 2..64");
    for (int t = 0; t < T; ++t) {
        TB_REQUIRE(h_n_corners[t] >= 2 && h_n_corners[t] <= max_corners, "vc_build_tracks: bad corner count");
        TB_REQUIRE(h_resolution[t] > 0 && h_smoothing[t] >= 0, "vc_build_tracks: bad resolution / smoothing");
    }
    cudaSetDevice(h->device);
    cudaStream_t st = h->stream;
    // ---- inputs
    double *d_corners = nullptr, *d_res = nullptr;
    int *d_nc = nullptr, *d_sm = nullptr, *d_cnt = nullptr;   // d_cnt: [n_wp T | wp_off T+1 | n_k T | k_off T+1]
    TB_CUDA(cudaMalloc(&d_corners, sizeof(double) * (size_t)T * max_corners * 2), "cudaMalloc(track corners)");
    TB_CUDA(cudaMalloc(&d_res, sizeof(double) * T), "cudaMalloc(track resolution)");
    TB_CUDA(cudaMalloc(&d_nc, sizeof(int) * T), "cudaMalloc(track corners)");
    TB_CUDA(cudaMalloc(&d_sm, sizeof(int) * T), "cudaMalloc(track smoothing)");
    TB_CUDA(cudaMalloc(&d_cnt, sizeof(int) * (4 * (size_t)T + 2)), "cudaMalloc(track counts)");
    struct Guard { double *a, *b; int *c, *d, *e; double* f = nullptr; double* g = nullptr;
                   ~Guard() { cudaFree(a); cudaFree(b); cudaFree(c); cudaFree(d); cudaFree(e); cudaFree(f); cudaFree(g); } } guard{d_corners, d_res, d_nc, d_sm, d_cnt};
    cudaMemcpyAsync(d_corners, h_corners, sizeof(double) * (size_t)T * max_corners * 2, cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(d_res, h_resolution, sizeof(double) * T, cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(d_nc, h_n_corners, sizeof(int) * T, cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(d_sm, h_smoothing, sizeof(int) * T, cudaMemcpyHostToDevice, st);
    int *d_nwp = d_cnt, *d_wpoff = d_cnt + T, *d_nk = d_cnt + 2 * T + 1, *d_koff = d_cnt + 3 * T + 1;
    const BuildIn in{d_corners, d_nc, d_res, d_sm, max_corners};
    k_tb_count<<<(T + 127) / 128, 128, 0, st>>>(T, in, d_nwp);
    h->launches++;
    std::vector<int> n_wp(T), wp_off(T + 1, 0);
    TB_CUDA(cudaMemcpyAsync(n_wp.data(), d_nwp, sizeof(int) * T, cudaMemcpyDeviceToHost, st), "vc_build_tracks");
    TB_CUDA(cudaStreamSynchronize(st), "vc_build_tracks: waypoint count");
    int max_wp = 0;
    for (int t = 0; t < T; ++t) {
        TB_REQUIRE(n_wp[t] >= 8 && n_wp[t] > 2 * h_smoothing[t] + 1, "vc_build_tracks: too few waypoints for the smoothing window");
        wp_off[t + 1] = wp_off[t] + n_wp[t];
        max_wp = n_wp[t] > max_wp ? n_wp[t] : max_wp;
    }
    const size_t W = (size_t)wp_off[T];
    cudaMemcpyAsync(d_wpoff, wp_off.data(), sizeof(int) * (T + 1), cudaMemcpyHostToDevice, st);
    // ---- waypoints, smoothing, centre-line splines, length
    double* d_tmp = nullptr;     // [rx | ry | speed | length T] + band scratch [2][W][5]
    TB_CUDA(cudaMalloc(&d_tmp, sizeof(double) * (3 * W + T + 10 * W)), "cudaMalloc(track build scratch)");
    guard.f = d_tmp;
    double *rx = d_tmp, *ry = d_tmp + W, *speed = d_tmp + 2 * W, *d_len = d_tmp + 3 * W, *band = d_tmp + 3 * W + T;
    double* d_xy = nullptr;      // [wx | wy | cx | cy]
    TB_CUDA(cudaMalloc(&d_xy, sizeof(double) * 4 * W), "cudaMalloc(track splines)");
    guard.g = d_xy;
    const dim3 gw((max_wp + 127) / 128, T);
    k_tb_waypoints<<<gw, 128, 0, st>>>(T, in, d_wpoff, d_nwp, rx, ry);
    k_tb_smooth<<<gw, 128, 0, st>>>(T, d_sm, d_wpoff, d_nwp, rx, ry, d_xy, d_xy + W);
    k_tb_fit_xy<<<(2 * T + 31) / 32, 32, 0, st>>>(T, d_wpoff, d_nwp, d_xy, d_xy + W, d_xy + 2 * W, d_xy + 3 * W, band);
    k_tb_length<<<T, 32, 0, st>>>(T, d_wpoff, d_nwp, d_xy + 2 * W, d_xy + 3 * W, speed, d_len);
    h->launches += 4;
    std::vector<double> length(T);
    TB_CUDA(cudaMemcpyAsync(length.data(), d_len, sizeof(double) * T, cudaMemcpyDeviceToHost, st), "vc_build_tracks");
    TB_CUDA(cudaStreamSynchronize(st), "vc_build_tracks: centre-line splines");
    // ---- curvature tables and their splines (the atlas)
    std::vector<int> n_k(T), k_off(T + 1, 0);
    int max_k = 0;
    for (int t = 0; t < T; ++t) {
        TB_REQUIRE(length[t] == length[t] && length[t] > 1.0, "vc_build_tracks: degenerate track");
        n_k[t] = (int)ceil((length[t] - 0.1) / CURV_STEP);     // len(np.arange(0, L - 0.1, 0.05))
        k_off[t + 1] = k_off[t] + n_k[t];
        max_k = n_k[t] > max_k ? n_k[t] : max_k;
    }
    const size_t K = (size_t)k_off[T];
    cudaMemcpyAsync(d_nk, n_k.data(), sizeof(int) * T, cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(d_koff, k_off.data(), sizeof(int) * (T + 1), cudaMemcpyHostToDevice, st);
    double *d_bt = nullptr, *d_atlas = nullptr, *d_band2 = nullptr;
    TrackDesc* d_desc = nullptr;
    TB_CUDA(cudaMalloc(&d_bt, sizeof(double) * (4 * W + K)), "cudaMalloc(built tracks)");
    TB_CUDA(cudaMalloc(&d_atlas, sizeof(double) * K), "cudaMalloc(track atlas)");
    TB_CUDA(cudaMalloc(&d_desc, sizeof(TrackDesc) * T), "cudaMalloc(track atlas)");
    TB_CUDA(cudaMalloc(&d_band2, sizeof(double) * 5 * K), "cudaMalloc(track build scratch)");
    cudaMemcpyAsync(d_bt, d_xy, sizeof(double) * 4 * W, cudaMemcpyDeviceToDevice, st);
    double* table = d_bt + 4 * W;
    k_tb_curvature<<<dim3((max_k + 127) / 128, T), 128, 0, st>>>(T, d_wpoff, d_nwp, d_koff, d_nk, d_len, d_bt + 2 * W, d_bt + 3 * W, table);
    k_tb_fit_k<<<(T + 31) / 32, 32, 0, st>>>(T, d_koff, d_nk, table, d_atlas, d_band2);
    h->launches += 2;
    std::vector<TrackDesc> desc(T);
    for (int t = 0; t < T; ++t) desc[t] = TrackDesc{0.0, CURV_STEP, k_off[t], n_k[t]};
    cudaMemcpyAsync(d_desc, desc.data(), sizeof(TrackDesc) * T, cudaMemcpyHostToDevice, st);
    int rc = check_cuda(cudaStreamSynchronize(st), "vc_build_tracks: curvature splines");
    cudaFree(d_band2);
    if (rc) { cudaFree(d_bt); cudaFree(d_atlas); cudaFree(d_desc); return rc; }
    // ---- install: the built tracks become the handle's atlas
    cudaFree(h->d_atlas); cudaFree(h->d_atlas_desc); cudaFree(h->d_bt);
    h->d_atlas = d_atlas; h->d_atlas_desc = d_desc; h->n_atlas = T; h->d_bt = d_bt;
    h->bt_wp_off.assign(wp_off.begin(), wp_off.end()); h->bt_n_wp = n_wp; h->bt_k_off.assign(k_off.begin(), k_off.end());
    h->bt_n_k = n_k; h->bt_length = length; h->bt_total_wp = (long long)W; h->bt_total_k = (long long)K;
    return check_cuda(cudaGetLastError(), "vc_build_tracks");
}

int vc_built_track_info(vc_handle* h, int t, double* length, int* n_waypoints, int* n_k) {
    TB_REQUIRE(h && h->d_bt && t >= 0 && t < (int)h->bt_n_wp.size(), "vc_built_track_info: no such built track");
    if (length) *length = h->bt_length[t];
    if (n_waypoints) *n_waypoints = h->bt_n_wp[t];
    if (n_k) *n_k = h->bt_n_k[t];
    return 0;
}

int vc_built_track_read(vc_handle* h, int t, int what, double* h_out, int n) {
    TB_REQUIRE(h && h->d_bt && h_out && t >= 0 && t < (int)h->bt_n_wp.size(), "vc_built_track_read: no such built track");
    TB_REQUIRE(what >= 0 && what <= 5, "vc_built_track_read: what must be 0..5");
    const double* src;
    int len;
    const size_t W = (size_t)h->bt_total_wp;
    if (what == 0) { src = h->d_atlas + h->bt_k_off[t]; len = h->bt_n_k[t]; }                 // Track.k coefficients
    else if (what == 5) { src = h->d_bt + 4 * W + h->bt_k_off[t]; len = h->bt_n_k[t]; }     // raw curvature table
    else { src = h->d_bt + (size_t)(what - 1) * W + h->bt_wp_off[t]; len = h->bt_n_wp[t]; }  // 1 wx, 2 wy, 3 x coefs, 4 y coefs
    TB_REQUIRE(n >= 0 && n <= len, "vc_built_track_read: n exceeds the table");
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    return check_cuda(cudaMemcpy(h_out, src, sizeof(double) * n, cudaMemcpyDeviceToHost), "vc_built_track_read");
}

int vc_built_track_rel2glob(vc_handle* h, int t, int n, const double* d_s, const double* d_ey, double* d_xy) {
    TB_REQUIRE(h && h->d_bt && t >= 0 && t < (int)h->bt_n_wp.size() && d_s && d_ey && d_xy, "vc_built_track_rel2glob: bad argument");
    if (n <= 0) return 0;
    const size_t W = (size_t)h->bt_total_wp;
    k_tb_rel2glob<<<(n + 127) / 128, 128, 0, h->stream>>>(n, h->d_bt + 2 * W + h->bt_wp_off[t], h->d_bt + 3 * W + h->bt_wp_off[t],
                                                          h->bt_n_wp[t], h->bt_length[t], d_s, d_ey, d_xy);
    h->launches++;
    return check_cuda(cudaGetLastError(), "k_tb_rel2glob");
}

}  // extern "C"
