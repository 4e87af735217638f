// Vehicle models of the hot path, templated on the scalar type (double / Dual1 / HDual).
//
// Reference formulas restated (not copied) from:
//   SingleTrack tyres/loads/ODE : vehicle_control/models/dynamic_car.py:78-163
//   SingleTrack spatial ODE     : vehicle_control/models/dynamic_car.py:170-187
//   PointMass ODE + Euler       : vehicle_control/models/dynamic_point_mass.py:81-103
//   RK4 / Euler one-step maps   : vehicle_control/utils/integrators.py:15-37
// State order (dynamic_car.py:246-247): Ux Uy r delta s ey epsi t ; input Fx w.
// PointMass (dynamic_point_mass.py:155-156): V s ey epsi t ; input Fx Fy.
#pragma once
#include "vc_ad.cuh"
#include "../../include/vc_api.h"

namespace vc {

typedef vc_model_params Params;

enum { IUX = 0, IUY = 1, IR = 2, IDELTA = 3, IS = 4, IEY = 5, IEPSI = 6, IT = 7 };
enum { PV = 0, PS = 1, PEY = 2, PEPSI = 3, PT = 4 };

template <class T> struct Tyres {
    T fxf, fxr;        // longitudinal force split          dynamic_car.py:78-86
    T fzf, fzr;        // normal loads                       :98-102
    T fymax_f, fymax_r;  // friction-ellipse lateral limit   :107-108
    T tan_f, tan_r;    // tan of the slip angles             :111-115 (alpha itself is never needed, see tyres())
    T sd, cd;          // sin / cos of the steering angle (shared with the chassis equations)
    T fyf, fyr;        // modified Fiala lateral forces      :121-142
};

// drive/brake split: tanh blend between Xd and Xb distributions
template <class T> VC_HD void force_split(const Params& p, const T& Fx, T& fxf, T& fxr) {
    const T th = ad_tanh(2.0 * (Fx / 1000.0 + 0.5));
    const T Xf = (p.Xdf - p.Xbf) / 2 * th + (p.Xdf + p.Xbf) / 2;
    const T Xr = (p.Xdr - p.Xbr) / 2 * th + (p.Xdr + p.Xbr) / 2;   // tanh(-z) = -tanh(z)
    fxf = Fx * Xf;
    fxr = Fx * Xr;
}

template <class T> VC_HD void normal_loads(const Params& p, const T& Ux, const T& Fx, T& fzf, T& fzr) {
    const T base = p.g_eff + p.Av2 * (Ux * Ux);
    const T shift = (p.h / p.l) * Fx;
    fzf = (p.b / p.l) * p.m * base - shift;
    fzr = (p.a / p.l) * p.m * base + shift;
}

// modified Fiala brush model; branch chosen on values (CasADi if_else semantics).  The reference tests
// |alpha| <= alpha_mod = atan(3 Fymax eps / C); tan is monotone on (-pi/2, pi/2), so the same branch is taken by
// |tan alpha| <= 3 Fymax eps / C, and sign(alpha) = sign(tan alpha): no inverse trigonometry on the device.
template <class T>
VC_HD T fiala(double C, double eps, const T& ta, const T& fymax) {
    const double tmod = 3.0 * val(fymax) * eps / C;
    if (fabs(val(ta)) <= tmod) {
        const T inv = recip(fymax);
        return -C * ta + (C * C / 3.0) * (ad_abs(ta) * ta) * inv
               - (C * C * C / 27.0) * (ta * ta * ta) * (inv * inv);
    }
    const double sg = val(ta) > 0.0 ? 1.0 : (val(ta) < 0.0 ? -1.0 : 0.0);
    return -C * (1 - 2 * eps + eps * eps) * ta - (3 * eps * eps - 2 * eps * eps * eps) * sg * fymax;
}

// Slip angles (dynamic_car.py:111-115): alpha_f = atan((Uy + a r)/Ux) - delta, alpha_r = atan((Uy - b r)/Ux).  Every use
// of them downstream is through tan(alpha) (tyre force, slip cost) or cos(alpha) (friction-limit rows), so they are
// carried as tangents: tan(alpha_r) = z_r, tan(alpha_f) = (z_f cos(delta) - sin(delta)) / (cos(delta) + z_f sin(delta))
// -- the same functions of the state as the reference's atan/tan chain (to rounding), without fp64 atan/tan calls.
template <class T>
VC_HD void tyres(const Params& p, const T& Ux, const T& Uy, const T& r, const T& delta, const T& Fx,
                 Tyres<T>& o) {
    force_split(p, Fx, o.fxf, o.fxr);
    normal_loads(p, Ux, Fx, o.fzf, o.fzr);
    const T lf = p.muf * o.fzf, lr = p.mur * o.fzr;
    const T qf = 0.99 * o.fxf, qr = 0.99 * o.fxr;
    o.fymax_f = ad_sqrt(lf * lf - qf * qf);
    o.fymax_r = ad_sqrt(lr * lr - qr * qr);
    const T iux = recip(Ux);
    const T zf = (Uy + p.a * r) * iux;
    ad_sincos(delta, o.sd, o.cd);
    o.tan_f = (zf * o.cd - o.sd) * recip(o.cd + zf * o.sd);
    o.tan_r = (Uy - p.b * r) * iux;
    o.fyf = fiala(p.Caf, p.eps, o.tan_f, o.fymax_f);
    o.fyr = fiala(p.Car, p.eps, o.tan_r, o.fymax_r);
}

// cos(alpha) from tan(alpha), |alpha| < pi/2
template <class T> VC_HD T cos_from_tan(const T& t) { return recip(ad_sqrt(1.0 + t * t)); }

// SingleTrack right-hand side. SPATIAL=false: d/dt (dynamic_car.py:153-160); true: d/ds (:170-187).
template <bool SPATIAL, class T>
VC_HD void st_rhs(const Params& p, const T* x, const T* u, double kappa, T* f) {
    const T &Ux = x[IUX], &Uy = x[IUY], &r = x[IR], &delta = x[IDELTA], &ey = x[IEY], &epsi = x[IEPSI];
    const T &Fx = u[0], &w = u[1];
    Tyres<T> ty;
    tyres(p, Ux, Uy, r, delta, Fx, ty);
    T se, ce;
    const T &sd = ty.sd, &cd = ty.cd;
    ad_sincos(epsi, se, ce);
    const T Fd = p.Frr + p.Cd * (Ux * Ux);
    const T lat = ty.fyf * cd + ty.fxf * sd;
    const T Ux_dot = (ty.fxf * cd - ty.fyf * sd + ty.fxr - Fd) / p.m + r * Uy;
    const T Uy_dot = (lat + ty.fyr) / p.m - r * Ux;
    const T r_dot = (p.a * lat - p.b * ty.fyr) / p.Izz;
    const T s_dot = (Ux * ce - Uy * se) * recip(1.0 - kappa * ey);
    const T ey_dot = Ux * se + Uy * ce;
    const T epsi_dot = r - kappa * s_dot;
    if (SPATIAL) {
        const T is = recip(s_dot);
        f[IUX] = Ux_dot * is; f[IUY] = Uy_dot * is; f[IR] = r_dot * is; f[IDELTA] = w * is;
        f[IS] = cst<T>(1.0); f[IEY] = ey_dot * is; f[IEPSI] = epsi_dot * is; f[IT] = is;
    } else {
        f[IUX] = Ux_dot; f[IUY] = Uy_dot; f[IR] = r_dot; f[IDELTA] = w;
        f[IS] = s_dot; f[IEY] = ey_dot; f[IEPSI] = epsi_dot; f[IT] = cst<T>(1.0);
    }
}

// classical RK4 with the reference's operation order: x + h*(1/6)*(k1 + 2 k2 + 2 k3 + k4).
// The four stages run as a rolled loop so that the (large, fully inlined) right-hand side is emitted once per
// instantiation: the hyper-dual kernels are instruction-cache bound otherwise.
template <bool SPATIAL, class T>
VC_HD void st_rk4(const Params& p, const T* x, const T* u, double kappa, double h, T* xn) {
    T acc[8], xs[8], f[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) xs[i] = x[i];
#pragma unroll 1
    for (int s = 0; s < 4; ++s) {
        st_rhs<SPATIAL>(p, xs, u, kappa, f);
        const double wgt = (s == 0 || s == 3) ? 1.0 : 2.0;
        const double c = s == 2 ? h : 0.5 * h;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (s == 0) acc[i] = f[i]; else acc[i] = acc[i] + wgt * f[i];
            xs[i] = x[i] + c * f[i];
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) xn[i] = x[i] + (h * (1.0 / 6.0)) * acc[i];
}

// PointMass right-hand side; SPATIAL as above (dynamic_point_mass.py:81-99)
template <bool SPATIAL, class T>
VC_HD void pm_rhs(const Params& p, const T* x, const T* u, double kappa, T* f) {
    const T &V = x[PV], &ey = x[PEY], &epsi = x[PEPSI];
    const T &Fx = u[0], &Fy = u[1];
    T se, ce;
    ad_sincos(epsi, se, ce);
    const T Fd = p.Frr + p.Cd * (V * V);
    const T V_dot = (Fx - Fd) / p.m;
    const T s_dot = (V * ce) * recip(1.0 - kappa * ey);
    const T ey_dot = V * se;
    const T epsi_dot = Fy * recip(p.m * V) - kappa * s_dot;
    if (SPATIAL) {
        const T is = recip(s_dot);
        f[PV] = V_dot * is; f[PS] = cst<T>(1.0); f[PEY] = ey_dot * is; f[PEPSI] = epsi_dot * is; f[PT] = is;
    } else {
        f[PV] = V_dot; f[PS] = s_dot; f[PEY] = ey_dot; f[PEPSI] = epsi_dot; f[PT] = cst<T>(1.0);
    }
}

template <bool SPATIAL, class T>
VC_HD void pm_euler(const Params& p, const T* x, const T* u, double kappa, double h, T* xn) {
    T f[5];
    pm_rhs<SPATIAL>(p, x, u, kappa, f);
#pragma unroll
    for (int i = 0; i < 5; ++i) xn[i] = x[i] + h * f[i];
}

// Cubic B-spline evaluation (de Boor) on a coefficient table exported by the host Track.
// Restates CasADi interpolant("bspline") = not-a-knot cubic through uniformly spaced sites with
// polynomial extrapolation (track.py:156-167, 228-229).  Sites are x0 + i*dx, i = 0..n-1, so the
// knot vector is arithmetic: 4x first site, sites 2..n-3, 4x last site; only coefficients are stored.
struct UniformKnots {
    double x0, dx; int n;
    VC_HD double operator()(int i) const {
        return i < 4 ? x0 : (i >= n ? x0 + (n - 1) * dx : x0 + (i - 2) * dx);
    }
};

// returns S(x); if deriv != nullptr also S'(x)
template <class CoefPtr>
VC_HD double bspline3(CoefPtr c, const UniformKnots& t, double x, double* deriv) {
    const int n = t.n;
    double jf = floor((x - t.x0) / t.dx);
    jf = jf < 0.0 ? 0.0 : (jf > (double)n ? (double)n : jf);
    int k = (int)jf + 2;                        // span guess: t(k) <= x < t(k+1)
    if (k < 3) k = 3;
    if (k > n - 1) k = n - 1;
    while (k > 3 && x < t(k)) --k;
    while (k < n - 1 && x >= t(k + 1)) ++k;
    const double tm2 = t(k - 2), tm1 = t(k - 1), t0 = t(k), t1 = t(k + 1), t2 = t(k + 2), t3 = t(k + 3);
    const double e0 = c[k - 3], e1 = c[k - 2], e2 = c[k - 1], e3 = c[k];
    const double a1 = (x - tm2) / (t1 - tm2), a2 = (x - tm1) / (t2 - tm1), a3 = (x - t0) / (t3 - t0);
    const double f1 = (1 - a1) * e0 + a1 * e1;
    const double f2 = (1 - a2) * e1 + a2 * e2;
    const double f3 = (1 - a3) * e2 + a3 * e3;
    const double b2 = (x - tm1) / (t1 - tm1), b3 = (x - t0) / (t2 - t0);
    const double g2 = (1 - b2) * f1 + b2 * f2;
    const double g3 = (1 - b3) * f2 + b3 * f3;
    const double c3 = (x - t0) / (t1 - t0);
    if (deriv) *deriv = 3.0 * (g3 - g2) / (t1 - t0);
    return (1 - c3) * g2 + c3 * g3;
}

}  // namespace vc
