// Stage functions of the NMPC transcription, templated on the scalar type (double / Dual1 / HDual).
//
// Restates (does not copy) the NLP that the reference builds with CasADi Opti:
//   variables / parameters   vehicle_control/controllers/mpc/cascaded_mpc.py:68-89
//   SingleTrack constraints  :91-128      SingleTrack stage cost   :130-179
//   PointMass constraints    :181-202     PointMass stage cost     :204-239
//   switching cost           :241-254     switching constraints    :256-277
//   terminal cost            :279-304
// The slew-rate terms (:167-171, :228-231), the obstacle terms (:173-177, :233-237), the terminal time
// weight and everything that touches s or t are assembled by the step kernel (vc_solver.cu), because they are
// quadratic / closed form; the functions here are the ones that need automatic differentiation.
//
// Stage variable order y = (x, u):  SingleTrack  Ux Uy r delta s ey epsi t | Fx w      (10)
//                                   PointMass    V s ey epsi t | Fx Fy                  (7)
#pragma once
#include "vc_model.cuh"

namespace vc {

typedef vc_mpc_config Cfg;

enum { YFX = 8, YW = 9 };                    // SingleTrack stage variable indices of the controls
enum { QV = 0, QS = 1, QEY = 2, QEPSI = 3, QT = 4, QFX = 5, QFY = 6 };   // PointMass stage variables
enum { NI_ST = 10, NI_PM = 2 };              // inequality rows per stage (Opti splits bounded() in two)

// IPOPT bound_relax_factor applied to constant bounds: b -/+ relax*max(1,|b|)
VC_HD double relax_lo(double b, double rl) { return b - rl * fmax(1.0, fabs(b)); }
VC_HD double relax_hi(double b, double rl) { return b + rl * fmax(1.0, fabs(b)); }

template <class T> VC_HD T ad_cos(const T& x) { T s, c; ad_sincos(x, s, c); return c; }
template <class T> VC_HD T sq(const T& x) { return x * x; }

// ---- SingleTrack stage: cost (without slew / obstacles / time) and the 10 inequality rows c(y) >= 0.
// `terminal` adds the terminal cost when this SingleTrack stage is the last column (M == 0).
template <class T>
VC_HD void st_cost_ineq(const Params& p, const Cfg& c, const T* y, double ds, bool terminal, T& cost, T* ineq) {
    const T &Ux = y[IUX], &Uy = y[IUY], &r = y[IR], &delta = y[IDELTA], &ey = y[IEY], &epsi = y[IEPSI];
    const T &Fx = y[YFX], &w = y[YW];
    Tyres<T> ty;
    tyres(p, Ux, Uy, r, delta, Fx, ty);
    const double rl = c.bound_relax;
    // ---- cost :130-165
    cost = (c.w_dev_st * ds) * sq(ey) + c.w_w * sq(w);
    if (val(ey) < c.ey_min) cost = cost + (c.w_boundary * ds) * sq(ey - c.ey_min);
    if (val(ey) > c.ey_max) cost = cost + (c.w_boundary * ds) * sq(ey - c.ey_max);
    {   // tan(alphamod) = 3 Fymax eps / C_alpha   (dynamic_car.py:119,132)
        const T taf = ad_abs(ty.tan_f), tmf = (3.0 * p.eps / p.Caf) * ty.fymax_f;
        if (val(taf) >= val(tmf)) cost = cost + c.w_slip * sq(taf - tmf);
        const T tar = ad_abs(ty.tan_r), tmr = (3.0 * p.eps / p.Car) * ty.fymax_r;
        if (val(tar) >= val(tmr)) cost = cost + c.w_slip * sq(tar - tmr);
    }
    if (terminal) {   // :279-304 with the SingleTrack as final model
        if (val(Ux) >= c.max_speed) cost = cost + c.w_speed * sq(Ux - c.max_speed);
        cost = cost + c.w_ey * sq(ey) + c.w_epsi * sq(epsi);
    }
    // ---- inequalities :91-128 (emission order)
    const T bf = p.muf * ty.fzf * cos_from_tan(ty.tan_f);
    const T br = p.mur * ty.fzr * cos_from_tan(ty.tan_r);
    ineq[0] = Ux - relax_lo(c.Ux_min, rl);
    ineq[1] = delta - relax_lo(c.delta_min, rl);
    ineq[2] = relax_hi(c.delta_max, rl) - delta;
    ineq[3] = p.Peng * recip(Ux) - Fx + rl;
    ineq[4] = w - relax_lo(c.w_min, rl);
    ineq[5] = relax_hi(c.w_max, rl) - w;
    ineq[6] = bf - ty.fxf + rl;
    ineq[7] = ty.fxf + bf + rl;
    ineq[8] = br - ty.fxr + rl;
    ineq[9] = ty.fxr + br + rl;
}

// ---- switching map SingleTrack column N-1 -> PointMass column N (:256-277), PointMass state order,
// and the lateral force the first PointMass Fy is tied to by the switching cost (:241-254).
template <class T>
VC_HD void st_switch(const Params& p, const T* y, T* xpm, T& fy_sum) {
    const T &Ux = y[IUX], &Uy = y[IUY];
    xpm[PV] = ad_sqrt(Ux * Ux + Uy * Uy);
    xpm[PS] = y[IS];
    xpm[PEY] = y[IEY];
    xpm[PEPSI] = ad_atan(Uy * recip(Ux)) + y[IEPSI];
    xpm[PT] = y[IT];
    Tyres<T> ty;
    tyres(p, Ux, Uy, y[IR], y[IDELTA], y[YFX], ty);
    fy_sum = ty.fyf + ty.fyr;
}

// ---- PointMass stage: cost (without slew / obstacles / time) and the 2 inequality rows.
template <class T>
VC_HD void pm_cost_ineq(const Params& p, const Cfg& c, const T* y, double ds, bool terminal, T& cost, T* ineq) {
    const T &V = y[QV], &ey = y[QEY], &epsi = y[QEPSI], &Fx = y[QFX];
    const double rl = c.bound_relax;
    cost = (c.w_dev_pm * ds) * sq(ey);
    if (val(ey) < c.ey_min_pm) cost = cost + (c.w_boundary * ds) * sq(ey - c.ey_min_pm);
    if (val(ey) > c.ey_max_pm) cost = cost + (c.w_boundary * ds) * sq(ey - c.ey_max_pm);
    if (terminal) {
        if (val(V) >= c.max_speed) cost = cost + c.w_speed * sq(V - c.max_speed);
        cost = cost + c.w_ey * sq(ey) + c.w_epsi * sq(epsi);
    }
    ineq[0] = V - relax_lo(c.V_min, rl);
    ineq[1] = p.Peng * recip(V) - Fx + rl;
}

// ---- obstacle term of one stage (closed form): W ds / (dist - (radius + 0.1)), summed over obstacles.
// Returns the value; g = d/d(s,ey); h = (ss, s-ey, ey-ey) second derivatives.  Points inside an obstacle's pole
// (dist <= radius + 0.1), where the reference's term turns negative and unbounded, evaluate to NaN so that the line
// search treats them like IPOPT treats an evaluation error (backtrack); solutions outside are unaffected.
VC_HD double obstacle_cost(const double* obst, int n_obs, double W, double ds, double s, double ey, double* g,
                           double* h) {
    double v = 0.0;
    if (g) { g[0] = g[1] = 0.0; }
    if (h) { h[0] = h[1] = h[2] = 0.0; }
    for (int o = 0; o < n_obs; ++o) {
        const double a = s - obst[3 * o], b = ey - obst[3 * o + 1], rad = obst[3 * o + 2] + 0.1;
        const double d = sqrt(a * a + b * b);
        if (!(d > rad)) return NAN;   // inside the pole the reciprocal term is negative and unbounded: not a valid point
        const double q = 1.0 / (d - rad);
        const double k = W * ds;
        v += k * q;
        if (g) {
            const double dq = -k * q * q;          // d/dd
            const double id = 1.0 / d;
            g[0] += dq * a * id;
            g[1] += dq * b * id;
            if (h) {
                const double ddq = 2.0 * k * q * q * q;   // d2/dd2
                const double id3 = id * id * id;
                h[0] += ddq * a * a * id * id + dq * (id - a * a * id3);
                h[1] += ddq * a * b * id * id + dq * (-a * b * id3);
                h[2] += ddq * b * b * id * id + dq * (id - b * b * id3);
            }
        }
    }
    return v;
}

}  // namespace vc
