// ORACLE (test infrastructure only -- never linked into or loaded by the product path).
//
// Host restatement of the reference's vehicle models and NMPC stage functions, generic over the scalar type
// (double for values, Jet<N> for exact first and second derivatives).  Written in the reference's own form --
// slip ANGLES through atan, tyre forces through tan(alpha), alpha_mod through atan -- i.e. not the tangent form
// the CUDA kernels use.  Follows, formula by formula:
//   SingleTrack tyres / loads / ODE   vehicle_control/models/dynamic_car.py:78-163, spatial ODE :170-187
//   PointMass ODE                     vehicle_control/models/dynamic_point_mass.py:81-103
//   RK4 / Euler one-step maps         vehicle_control/utils/integrators.py:15-37
//   NLP stage cost / constraints      vehicle_control/controllers/mpc/cascaded_mpc.py:91-304
// Parity status: pinned through tests/golden (plant steps of the 44 saved reference laps, IPOPT's saved first
// actions); see tests/test_oracle_cpp.py.
#pragma once
#include "jet.hpp"
#include "../../include/vc_api.h"   // parameter / configuration structs only (the public interface)

namespace vco {

typedef vc_model_params Par;
typedef vc_mpc_config Cfg;

// state order ST: Ux Uy r delta s ey epsi t | Fx w ; PM: V s ey epsi t | Fx Fy
enum { UX = 0, UY, RR, DELTA, SS, EY, EPSI, TT, FX, WW };
enum { PV = 0, PS, PEY, PEPSI, PT, PFX, PFY };

template <class T> struct Tyre { T fxf, fxr, fzf, fzr, fymax_f, fymax_r, alf, alr, amod_f, amod_r, fyf, fyr; };

template <class T> T fiala(double C, double eps, const T& alpha, const T& amod, const T& fymax) {   // dynamic_car.py:121-128
    const T ta = tan(alpha);
    if (std::fabs(val(alpha)) <= val(amod))
        return -C * ta + (C * C) * (fabs(ta) * ta) / (3.0 * fymax) - (C * C * C) * (ta * ta * ta) / (27.0 * (fymax * fymax));
    const double sg = val(alpha) > 0 ? 1.0 : (val(alpha) < 0 ? -1.0 : 0.0);
    return -C * (1 - 2 * eps + eps * eps) * ta - fymax * ((3 * eps * eps - 2 * eps * eps * eps) * sg);
}

template <class T> Tyre<T> tyre(const Par& p, const T& Ux, const T& Uy, const T& r, const T& delta, const T& Fx) {
    Tyre<T> o;
    const T Xf = (p.Xdf - p.Xbf) / 2 * tanh(2.0 * (Fx / 1000.0 + 0.5)) + (p.Xdf + p.Xbf) / 2;        // :78
    const T Xr = (p.Xbr - p.Xdr) / 2 * tanh(-2.0 * (Fx / 1000.0 + 0.5)) + (p.Xdr + p.Xbr) / 2;       // :83
    o.fxf = Fx * Xf; o.fxr = Fx * Xr;                                                              // :80,85
    o.fzf = (p.b / p.l) * p.m * (p.g_eff + p.Av2 * (Ux * Ux)) - p.h * Fx / p.l;                      // :98
    o.fzr = (p.a / p.l) * p.m * (p.g_eff + p.Av2 * (Ux * Ux)) + p.h * Fx / p.l;                      // :101
    const T mf = p.muf * o.fzf, mr = p.mur * o.fzr, qf = 0.99 * o.fxf, qr = 0.99 * o.fxr;
    o.fymax_f = sqrt(mf * mf - qf * qf);                                                            // :107
    o.fymax_r = sqrt(mr * mr - qr * qr);                                                            // :108
    o.alf = atan((Uy + p.a * r) / Ux) - delta;                                                      // :111
    o.alr = atan((Uy - p.b * r) / Ux);                                                              // :114
    o.amod_f = atan(3.0 * o.fymax_f * p.eps / p.Caf);                                               // :119
    o.amod_r = atan(3.0 * o.fymax_r * p.eps / p.Car);                                               // :132
    o.fyf = fiala(p.Caf, p.eps, o.alf, o.amod_f, o.fymax_f);
    o.fyr = fiala(p.Car, p.eps, o.alr, o.amod_r, o.fymax_r);
    return o;
}

// temporal right-hand side (dynamic_car.py:144-163); f[8]
template <class T> void st_dot(const Par& p, const T* x, const T* u, double kappa, T* f) {
    const T &Ux = x[UX], &Uy = x[UY], &r = x[RR], &delta = x[DELTA], &ey = x[EY], &epsi = x[EPSI];
    const Tyre<T> ty = tyre(p, Ux, Uy, r, delta, u[0]);
    const T Fd = p.Frr + p.Cd * (Ux * Ux);
    const T cd = cos(delta), sd = sin(delta);
    f[UX] = (ty.fxf * cd - ty.fyf * sd + ty.fxr - Fd) / p.m + r * Uy;
    f[UY] = (ty.fyf * cd + ty.fxf * sd + ty.fyr) / p.m - r * Ux;
    f[RR] = (p.a * (ty.fyf * cd + ty.fxf * sd) - p.b * ty.fyr) / p.Izz;
    f[DELTA] = u[1];
    f[SS] = (Ux * cos(epsi) - Uy * sin(epsi)) / (1.0 - kappa * ey);
    f[EY] = Ux * sin(epsi) + Uy * cos(epsi);
    f[EPSI] = r - kappa * f[SS];
    f[TT] = T(1.0);
}
template <class T> void st_prime(const Par& p, const T* x, const T* u, double kappa, T* f) {   // :170-187
    T d[8];
    st_dot(p, x, u, kappa, d);
    const T sd = d[SS];
    f[UX] = d[UX] / sd; f[UY] = d[UY] / sd; f[RR] = d[RR] / sd; f[DELTA] = d[DELTA] / sd;
    f[SS] = T(1.0); f[EY] = d[EY] / sd; f[EPSI] = d[EPSI] / sd; f[TT] = 1.0 / sd;
}
// integrators.py:30-34: x + h*(1/6)*(k1 + 2 k2 + 2 k3 + k4)
template <bool SPATIAL, class T> void st_rk4(const Par& p, const T* x, const T* u, double kappa, double h, T* xn) {
    T k1[8], k2[8], k3[8], k4[8], xs[8];
    auto f = [&](const T* xx, T* out) { if (SPATIAL) st_prime(p, xx, u, kappa, out); else st_dot(p, xx, u, kappa, out); };
    f(x, k1);
    for (int i = 0; i < 8; ++i) xs[i] = x[i] + (0.5 * h) * k1[i];
    f(xs, k2);
    for (int i = 0; i < 8; ++i) xs[i] = x[i] + (0.5 * h) * k2[i];
    f(xs, k3);
    for (int i = 0; i < 8; ++i) xs[i] = x[i] + h * k3[i];
    f(xs, k4);
    for (int i = 0; i < 8; ++i) xn[i] = x[i] + (h * (1.0 / 6.0)) * (k1[i] + 2.0 * k2[i] + 2.0 * k3[i] + k4[i]);
}

// PointMass spatial Euler step (dynamic_point_mass.py:81-103, integrators.py:17-23)
template <class T> void pm_step(const Par& p, const T* x, const T* u, double kappa, double h, T* xn) {
    const T &V = x[PV], &ey = x[PEY], &epsi = x[PEPSI];
    const T Fd = p.Frr + p.Cd * (V * V);
    const T V_dot = (u[0] - Fd) / p.m;
    const T s_dot = (V * cos(epsi)) / (1.0 - kappa * ey);
    const T ey_dot = V * sin(epsi);
    const T epsi_dot = u[1] / (p.m * V) - kappa * s_dot;
    xn[PV] = x[PV] + h * (V_dot / s_dot);
    xn[PS] = x[PS] + h;
    xn[PEY] = x[PEY] + h * (ey_dot / s_dot);
    xn[PEPSI] = x[PEPSI] + h * (epsi_dot / s_dot);
    xn[PT] = x[PT] + h * (1.0 / s_dot);
}

inline double relax_lo(double b, double rl) { return b - rl * std::fmax(1.0, std::fabs(b)); }
inline double relax_hi(double b, double rl) { return b + rl * std::fmax(1.0, std::fabs(b)); }

// obstacle term W ds / (dist - (radius + 0.1)) summed over the obstacles (cascaded_mpc.py:173-177, :233-237).  Inside
// the pole the reference's term is negative and unbounded: not a valid point (NaN -> the line search backtracks).
template <class T> T obstacle_cost(const double* obst, int n_obs, double W, double ds, const T& s, const T& ey) {
    T acc(0.0);
    for (int o = 0; o < n_obs; ++o) {
        const T a = s - obst[3 * o], b = ey - obst[3 * o + 1];
        const T gap = sqrt(a * a + b * b) - (obst[3 * o + 2] + 0.1);
        if (!(val(gap) > 0.0)) return T(NAN);
        acc = acc + (W * ds) / gap;
    }
    return acc;
}

// SingleTrack stage n < N: cost terms that depend on this stage's variables only (cascaded_mpc.py:130-165,173-177),
// the 10 inequality rows c(y) >= 0 (:91-128, emission order) and -- for the switching stage -- T = Fy_f + Fy_r
template <class T>
void st_stage(const Par& p, const Cfg& c, const T* y, double ds, const double* obst, int n_obs, bool terminal,
              T& cost, T* ineq, T* fy_sum) {
    const T &Ux = y[UX], &Uy = y[UY], &r = y[RR], &delta = y[DELTA], &ey = y[EY], &epsi = y[EPSI], &Fx = y[FX], &w = y[WW];
    const Tyre<T> ty = tyre(p, Ux, Uy, r, delta, Fx);
    const double rl = c.bound_relax;
    cost = (c.w_dev_st * ds) * (ey * ey) + c.w_w * (w * w);
    if (val(ey) < c.ey_min) cost = cost + (c.w_boundary * ds) * ((ey - c.ey_min) * (ey - c.ey_min));
    if (val(ey) > c.ey_max) cost = cost + (c.w_boundary * ds) * ((ey - c.ey_max) * (ey - c.ey_max));
    {
        const T taf = fabs(tan(ty.alf)), tmf = tan(ty.amod_f);
        if (val(taf) >= val(tmf)) cost = cost + c.w_slip * ((taf - tmf) * (taf - tmf));
        const T tar = fabs(tan(ty.alr)), tmr = tan(ty.amod_r);
        if (val(tar) >= val(tmr)) cost = cost + c.w_slip * ((tar - tmr) * (tar - tmr));
    }
    if (c.obstacles && n_obs > 0) cost = cost + obstacle_cost(obst, n_obs, c.w_obstacles, ds, y[SS], ey);
    if (terminal) {   // :279-304, SingleTrack as the final model
        if (val(Ux) >= c.max_speed) cost = cost + c.w_speed * ((Ux - c.max_speed) * (Ux - c.max_speed));
        cost = cost + c.w_time * y[TT] + c.w_ey * (ey * ey) + c.w_epsi * (epsi * epsi);
    }
    const T bf = p.muf * ty.fzf * cos(ty.alf), br = p.mur * ty.fzr * cos(ty.alr);
    ineq[0] = Ux - relax_lo(c.Ux_min, rl);
    ineq[1] = delta - relax_lo(c.delta_min, rl);
    ineq[2] = relax_hi(c.delta_max, rl) - delta;
    ineq[3] = p.Peng / Ux - Fx + rl;
    ineq[4] = w - relax_lo(c.w_min, rl);
    ineq[5] = relax_hi(c.w_max, rl) - w;
    ineq[6] = bf - ty.fxf + rl;
    ineq[7] = ty.fxf + bf + rl;
    ineq[8] = br - ty.fxr + rl;
    ineq[9] = ty.fxr + br + rl;
    if (fy_sum) *fy_sum = ty.fyf + ty.fyr;
}

// switching map SingleTrack column N-1 -> PointMass column N (cascaded_mpc.py:256-277), PointMass state order
template <class T> void st_switch(const T* y, T* xpm) {
    xpm[PV] = sqrt(y[UX] * y[UX] + y[UY] * y[UY]);
    xpm[PS] = y[SS];
    xpm[PEY] = y[EY];
    xpm[PEPSI] = atan(y[UY] / y[UX]) + y[EPSI];
    xpm[PT] = y[TT];
}

// PointMass stage m in [N, H): own-variable cost terms (:204-226,233-237), 2 inequality rows (:181-192)
template <class T>
void pm_stage(const Par& p, const Cfg& c, const T* y, double ds, const double* obst, int n_obs, bool terminal, T& cost, T* ineq) {
    const T &V = y[PV], &ey = y[PEY], &epsi = y[PEPSI];
    const double rl = c.bound_relax;
    cost = (c.w_dev_pm * ds) * (ey * ey);
    if (val(ey) < c.ey_min_pm) cost = cost + (c.w_boundary * ds) * ((ey - c.ey_min_pm) * (ey - c.ey_min_pm));
    if (val(ey) > c.ey_max_pm) cost = cost + (c.w_boundary * ds) * ((ey - c.ey_max_pm) * (ey - c.ey_max_pm));
    if (c.obstacles && n_obs > 0) cost = cost + obstacle_cost(obst, n_obs, c.w_obstacles, ds, y[PS], ey);
    if (terminal) {
        if (val(V) >= c.max_speed) cost = cost + c.w_speed * ((V - c.max_speed) * (V - c.max_speed));
        cost = cost + c.w_time * y[PT] + c.w_ey * (ey * ey) + c.w_epsi * (epsi * epsi);
    }
    ineq[0] = V - relax_lo(c.V_min, rl);
    ineq[1] = p.Peng / V - y[PFX] + rl;
}

}  // namespace vco
