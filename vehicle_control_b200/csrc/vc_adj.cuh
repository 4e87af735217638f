// Hand-written adjoints (reverse mode) of the SingleTrack stage functions, templated on the scalar type.
//
// The exact Lagrangian Hessian IPOPT gets from CasADi ("expand": True, cascaded_mpc.py:65) is assembled here
// forward-over-reverse: the reverse sweep below yields the gradient of the scalar stage Lagrangian
//     l(y) = sf cost(y) - z . c(y) - lam . Phi(y)
// and, run in Dual1 arithmetic seeded along one active variable a, also d/da of that gradient = column a of the
// Hessian.  Eight threads (one per direction) replace the 36 hyper-dual pair evaluations: the value part is recomputed
// 8 instead of 36 times and every operation carries 2 instead of 4 components.
//
// Forward formulas: vc_model.cuh / vc_nlp.cuh (same operation order, so values agree bit for bit with the templates
// there); reference: vehicle_control/models/dynamic_car.py:78-187, utils/integrators.py:26-37,
// controllers/mpc/cascaded_mpc.py:91-179,241-277.
#pragma once
#include "vc_nlp.cuh"

namespace vc {

// ---- forward intermediates of tyres() that its adjoint needs
template <class T> struct TyreTape {
    T th, Xf, Xr, fxf, fxr, fzf, fzr, lf, lr, qf, qr, fymax_f, fymax_r, iux, zf, sd, cd, iden, num, tan_f, tan_r, fyf, fyr;
    bool in_f, in_r;   // Fiala branch taken (cubic / saturated)
};
// adjoints of the tyre outputs (inputs of the reverse sweep)
template <class T> struct TyreAdj { T fxf, fxr, fzf, fzr, fymax_f, fymax_r, tan_f, tan_r, sd, cd, fyf, fyr; };
// adjoints of the model inputs (outputs of the reverse sweep); accumulated into
template <class T> struct InAdj { T Ux, Uy, r, delta, ey, epsi, Fx, w; };

template <class T> VC_HD T zero() { return cst<T>(0.0); }

template <class T>
VC_HD void tyres_fwd(const Params& p, const T& Ux, const T& Uy, const T& r, const T& delta, const T& Fx, TyreTape<T>& o) {
    o.th = ad_tanh(2.0 * (Fx / 1000.0 + 0.5));
    o.Xf = (p.Xdf - p.Xbf) / 2 * o.th + (p.Xdf + p.Xbf) / 2;
    o.Xr = (p.Xdr - p.Xbr) / 2 * o.th + (p.Xdr + p.Xbr) / 2;
    o.fxf = Fx * o.Xf;
    o.fxr = Fx * o.Xr;
    const T base = p.g_eff + p.Av2 * (Ux * Ux);
    const T shift = (p.h / p.l) * Fx;
    o.fzf = (p.b / p.l) * p.m * base - shift;
    o.fzr = (p.a / p.l) * p.m * base + shift;
    o.lf = p.muf * o.fzf; o.lr = p.mur * o.fzr;
    o.qf = 0.99 * o.fxf; o.qr = 0.99 * o.fxr;
    o.fymax_f = ad_sqrt(o.lf * o.lf - o.qf * o.qf);
    o.fymax_r = ad_sqrt(o.lr * o.lr - o.qr * o.qr);
    o.iux = recip(Ux);
    o.zf = (Uy + p.a * r) * o.iux;
    ad_sincos(delta, o.sd, o.cd);
    o.num = o.zf * o.cd - o.sd;
    o.iden = recip(o.cd + o.zf * o.sd);
    o.tan_f = o.num * o.iden;
    o.tan_r = (Uy - p.b * r) * o.iux;
    o.in_f = fabs(val(o.tan_f)) <= 3.0 * val(o.fymax_f) * p.eps / p.Caf;
    o.in_r = fabs(val(o.tan_r)) <= 3.0 * val(o.fymax_r) * p.eps / p.Car;
    o.fyf = fiala(p.Caf, p.eps, o.tan_f, o.fymax_f);
    o.fyr = fiala(p.Car, p.eps, o.tan_r, o.fymax_r);
}

// adjoint of fy = fiala(C, eps, ta, fymax): accumulates into ta_b, fymax_b
template <class T>
VC_HD void fiala_rev(double C, double eps, bool inner, const T& ta, const T& fymax, const T& fy_b, T& ta_b, T& fymax_b) {
    if (inner) {
        const T inv = recip(fymax);
        const T at = ad_abs(ta);
        // fy = -C ta + (C^2/3) |ta| ta inv - (C^3/27) ta^3 inv^2
        const T dta = (2.0 * C * C / 3.0) * (at * inv) - (C * C * C / 9.0) * ((ta * ta) * (inv * inv)) - C;
        const T dinv = (C * C / 3.0) * (at * ta) - (2.0 * C * C * C / 27.0) * ((ta * ta * ta) * inv);
        ta_b = ta_b + fy_b * dta;
        fymax_b = fymax_b - (fy_b * dinv) * (inv * inv);
    } else {
        const double sg = val(ta) > 0.0 ? 1.0 : (val(ta) < 0.0 ? -1.0 : 0.0);
        ta_b = ta_b - (C * (1 - 2 * eps + eps * eps)) * fy_b;
        fymax_b = fymax_b - ((3 * eps * eps - 2 * eps * eps * eps) * sg) * fy_b;
    }
}

// adjoint of tyres(): consumes the output adjoints `b`, accumulates into the input adjoints `in`
template <class T>
VC_HD void tyres_rev(const Params& p, const T& Ux, const T& Uy, const T& r, const T& Fx, const TyreTape<T>& t, TyreAdj<T> b,
                     InAdj<T>& in) {
    fiala_rev(p.Car, p.eps, t.in_r, t.tan_r, t.fymax_r, b.fyr, b.tan_r, b.fymax_r);
    fiala_rev(p.Caf, p.eps, t.in_f, t.tan_f, t.fymax_f, b.fyf, b.tan_f, b.fymax_f);
    // tan_r = (Uy - b r) iux
    T iux_b = b.tan_r * (Uy - p.b * r);
    in.Uy = in.Uy + b.tan_r * t.iux;
    in.r = in.r - p.b * (b.tan_r * t.iux);
    // tan_f = num * iden, num = zf cd - sd, iden = 1 / (cd + zf sd)
    const T num_b = b.tan_f * t.iden;
    const T den_b = -((b.tan_f * t.num) * (t.iden * t.iden));
    const T zf_b = num_b * t.cd + den_b * t.sd;
    const T cd_b = b.cd + num_b * t.zf + den_b;
    const T sd_b = b.sd - num_b + den_b * t.zf;
    in.delta = in.delta + sd_b * t.cd - cd_b * t.sd;
    // zf = (Uy + a r) iux
    iux_b = iux_b + zf_b * (Uy + p.a * r);
    in.Uy = in.Uy + zf_b * t.iux;
    in.r = in.r + p.a * (zf_b * t.iux);
    in.Ux = in.Ux - iux_b * (t.iux * t.iux);
    // fymax = sqrt(l^2 - q^2)
    const T hf = b.fymax_f * recip(t.fymax_f), hr = b.fymax_r * recip(t.fymax_r);
    const T fzf_b = b.fzf + p.muf * (hf * t.lf), fzr_b = b.fzr + p.mur * (hr * t.lr);
    const T fxf_b = b.fxf - 0.99 * (hf * t.qf), fxr_b = b.fxr - 0.99 * (hr * t.qr);
    // fz = (b|a)/l m base -+ shift, base = g + Av2 Ux^2, shift = h/l Fx
    const T base_b = ((p.b / p.l) * p.m) * fzf_b + ((p.a / p.l) * p.m) * fzr_b;
    in.Fx = in.Fx + (p.h / p.l) * (fzr_b - fzf_b);
    in.Ux = in.Ux + (2.0 * p.Av2) * (base_b * Ux);
    // fxf = Fx Xf, fxr = Fx Xr, X* = c1 th + c0, th = tanh(2 (Fx/1000 + 0.5))
    const T th_b = ((p.Xdf - p.Xbf) / 2) * (fxf_b * Fx) + ((p.Xdr - p.Xbr) / 2) * (fxr_b * Fx);
    in.Fx = in.Fx + fxf_b * t.Xf + fxr_b * t.Xr + (2.0 / 1000.0) * (th_b * (1.0 - t.th * t.th));
}

// ---- spatial right-hand side: forward with tape, and adjoint
template <class T> struct RhsTape { TyreTape<T> ty; T se, ce, lat, Uxd, Uyd, rd, ik, nm, sdot, eyd, epd, is; };

template <class T>
VC_HD void st_rhs_fwd(const Params& p, const T* x, const T* u, double kappa, RhsTape<T>& t, T* f) {
    const T &Ux = x[IUX], &Uy = x[IUY], &r = x[IR], &ey = x[IEY];
    tyres_fwd(p, Ux, Uy, r, x[IDELTA], u[0], t.ty);
    ad_sincos(x[IEPSI], t.se, t.ce);
    const T Fd = p.Frr + p.Cd * (Ux * Ux);
    t.lat = t.ty.fyf * t.ty.cd + t.ty.fxf * t.ty.sd;
    t.Uxd = (t.ty.fxf * t.ty.cd - t.ty.fyf * t.ty.sd + t.ty.fxr - Fd) / p.m + r * Uy;
    t.Uyd = (t.lat + t.ty.fyr) / p.m - r * Ux;
    t.rd = (p.a * t.lat - p.b * t.ty.fyr) / p.Izz;
    t.ik = recip(1.0 - kappa * ey);
    t.nm = Ux * t.ce - Uy * t.se;
    t.sdot = t.nm * t.ik;
    t.eyd = Ux * t.se + Uy * t.ce;
    t.epd = r - kappa * t.sdot;
    t.is = recip(t.sdot);
    f[IUX] = t.Uxd * t.is; f[IUY] = t.Uyd * t.is; f[IR] = t.rd * t.is; f[IDELTA] = u[1] * t.is;
    f[IS] = cst<T>(1.0); f[IEY] = t.eyd * t.is; f[IEPSI] = t.epd * t.is; f[IT] = t.is;
}

// adjoint of the spatial rhs: fb[8] = adjoint of f; accumulates into `in`; `extra` = additional adjoints of the tyre
// outputs at this point (the stage cost / inequality rows share the tyre evaluation of the first RK4 stage)
template <class T>
VC_HD void st_rhs_rev(const Params& p, const T* x, const T* u, double kappa, const RhsTape<T>& t, const T* fb,
                      const TyreAdj<T>* extra, InAdj<T>& in) {
    const T &Ux = x[IUX], &Uy = x[IUY], &r = x[IR];
    const T is_b = fb[IUX] * t.Uxd + fb[IUY] * t.Uyd + fb[IR] * t.rd + fb[IDELTA] * u[1] + fb[IEY] * t.eyd + fb[IEPSI] * t.epd + fb[IT];
    const T Uxd_b = fb[IUX] * t.is, Uyd_b = fb[IUY] * t.is, rd_b = fb[IR] * t.is, eyd_b = fb[IEY] * t.is, epd_b = fb[IEPSI] * t.is;
    in.w = in.w + fb[IDELTA] * t.is;
    const T sdot_b = -(is_b * (t.is * t.is)) - kappa * epd_b;
    in.r = in.r + epd_b;
    const T nm_b = sdot_b * t.ik, ik_b = sdot_b * t.nm;
    const T se_b = eyd_b * Ux - nm_b * Uy, ce_b = eyd_b * Uy + nm_b * Ux;
    in.Ux = in.Ux + eyd_b * t.se + nm_b * t.ce;
    in.Uy = in.Uy + eyd_b * t.ce - nm_b * t.se;
    in.ey = in.ey + kappa * (ik_b * (t.ik * t.ik));
    in.epsi = in.epsi + se_b * t.ce - ce_b * t.se;
    T lat_b = (p.a / p.Izz) * rd_b + Uyd_b / p.m;
    TyreAdj<T> b;
    if (extra) b = *extra;
    else { b.fxf = b.fxr = b.fzf = b.fzr = b.fymax_f = b.fymax_r = b.tan_f = b.tan_r = b.sd = b.cd = b.fyf = b.fyr = zero<T>(); }
    b.fyr = b.fyr - (p.b / p.Izz) * rd_b + Uyd_b / p.m;
    in.r = in.r - Uyd_b * Ux + Uxd_b * Uy;
    in.Ux = in.Ux - Uyd_b * r;
    in.Uy = in.Uy + Uxd_b * r;
    const T tm = Uxd_b / p.m;
    b.fxf = b.fxf + tm * t.ty.cd + lat_b * t.ty.sd;
    b.fyf = b.fyf - tm * t.ty.sd + lat_b * t.ty.cd;
    b.cd = b.cd + tm * t.ty.fxf + lat_b * t.ty.fyf;
    b.sd = b.sd - tm * t.ty.fyf + lat_b * t.ty.fxf;
    b.fxr = b.fxr + tm;
    in.Ux = in.Ux - (2.0 * p.Cd) * (tm * Ux);
    tyres_rev(p, Ux, Uy, r, u[0], t.ty, b, in);
}

// ---- stage cost + inequality rows: forward values (from a tyre tape at y) and the adjoint seeds they send to the
// tyre outputs / directly to the inputs.  cost_b = adjoint of the cost (sf), cb[j] = adjoint of row j (-z_j).
template <class T>
VC_HD void st_cost_ineq_fwd(const Params& p, const Cfg& c, const T* y, const TyreTape<T>& ty, double ds, bool terminal,
                            T& cost, T* ineq) {
    const T &Ux = y[IUX], &delta = y[IDELTA], &ey = y[IEY], &epsi = y[IEPSI], &Fx = y[YFX], &w = y[YW];
    const double rl = c.bound_relax;
    cost = (c.w_dev_st * ds) * sq(ey) + c.w_w * sq(w);
    if (val(ey) < c.ey_min) cost = cost + (c.w_boundary * ds) * sq(ey - c.ey_min);
    if (val(ey) > c.ey_max) cost = cost + (c.w_boundary * ds) * sq(ey - c.ey_max);
    {
        const T taf = ad_abs(ty.tan_f), tmf = (3.0 * p.eps / p.Caf) * ty.fymax_f;
        if (val(taf) >= val(tmf)) cost = cost + c.w_slip * sq(taf - tmf);
        const T tar = ad_abs(ty.tan_r), tmr = (3.0 * p.eps / p.Car) * ty.fymax_r;
        if (val(tar) >= val(tmr)) cost = cost + c.w_slip * sq(tar - tmr);
    }
    if (terminal) {
        if (val(Ux) >= c.max_speed) cost = cost + c.w_speed * sq(Ux - c.max_speed);
        cost = cost + c.w_ey * sq(ey) + c.w_epsi * sq(epsi);
    }
    const T bf = p.muf * ty.fzf * cos_from_tan(ty.tan_f);
    const T br = p.mur * ty.fzr * cos_from_tan(ty.tan_r);
    ineq[0] = Ux - relax_lo(c.Ux_min, rl);
    ineq[1] = delta - relax_lo(c.delta_min, rl);
    ineq[2] = relax_hi(c.delta_max, rl) - delta;
    ineq[3] = p.Peng * ty.iux - Fx + rl;
    ineq[4] = w - relax_lo(c.w_min, rl);
    ineq[5] = relax_hi(c.w_max, rl) - w;
    ineq[6] = bf - ty.fxf + rl;
    ineq[7] = ty.fxf + bf + rl;
    ineq[8] = br - ty.fxr + rl;
    ineq[9] = ty.fxr + br + rl;
}

template <class T>
VC_HD void st_cost_ineq_rev(const Params& p, const Cfg& c, const T* y, const TyreTape<T>& ty, double ds, bool terminal,
                            double cost_b, const double* cb, TyreAdj<T>& b, InAdj<T>& in) {
    const T &Ux = y[IUX], &ey = y[IEY], &epsi = y[IEPSI], &w = y[YW];
    b.fxf = b.fxr = b.fzf = b.fzr = b.fymax_f = b.fymax_r = b.tan_f = b.tan_r = b.sd = b.cd = b.fyf = b.fyr = zero<T>();
    // cost
    T dey = (2.0 * c.w_dev_st * ds) * ey;
    if (val(ey) < c.ey_min) dey = dey + (2.0 * c.w_boundary * ds) * (ey - c.ey_min);
    if (val(ey) > c.ey_max) dey = dey + (2.0 * c.w_boundary * ds) * (ey - c.ey_max);
    if (terminal) {
        dey = dey + (2.0 * c.w_ey) * ey;
        in.epsi = in.epsi + (cost_b * 2.0 * c.w_epsi) * epsi;
        if (val(Ux) >= c.max_speed) in.Ux = in.Ux + (cost_b * 2.0 * c.w_speed) * (Ux - c.max_speed);
    }
    in.ey = in.ey + cost_b * dey;
    in.w = in.w + (cost_b * 2.0 * c.w_w) * w;
    {
        const double k3f = 3.0 * p.eps / p.Caf, k3r = 3.0 * p.eps / p.Car;
        const T taf = ad_abs(ty.tan_f), tmf = k3f * ty.fymax_f;
        if (val(taf) >= val(tmf)) {
            const T e_b = (cost_b * 2.0 * c.w_slip) * (taf - tmf);
            b.tan_f = b.tan_f + (val(ty.tan_f) < 0.0 ? -e_b : e_b);
            b.fymax_f = b.fymax_f - k3f * e_b;
        }
        const T tar = ad_abs(ty.tan_r), tmr = k3r * ty.fymax_r;
        if (val(tar) >= val(tmr)) {
            const T e_b = (cost_b * 2.0 * c.w_slip) * (tar - tmr);
            b.tan_r = b.tan_r + (val(ty.tan_r) < 0.0 ? -e_b : e_b);
            b.fymax_r = b.fymax_r - k3r * e_b;
        }
    }
    // inequality rows (adjoint of row j = cb[j])
    in.Ux = in.Ux + cb[0] - (cb[3] * p.Peng) * (ty.iux * ty.iux);
    in.delta = in.delta + (cb[1] - cb[2]);
    in.Fx = in.Fx - cb[3];
    in.w = in.w + (cb[4] - cb[5]);
    const double bf_b = cb[6] + cb[7], br_b = cb[8] + cb[9];
    b.fxf = b.fxf + (cb[7] - cb[6]);
    b.fxr = b.fxr + (cb[9] - cb[8]);
    const T cf = cos_from_tan(ty.tan_f), cr = cos_from_tan(ty.tan_r);
    b.fzf = b.fzf + (bf_b * p.muf) * cf;
    b.fzr = b.fzr + (br_b * p.mur) * cr;
    // d cos(atan t) / dt = -t cos^3
    b.tan_f = b.tan_f - (bf_b * p.muf) * (ty.fzf * (ty.tan_f * (cf * cf * cf)));
    b.tan_r = b.tan_r - (br_b * p.mur) * (ty.fzr * (ty.tan_r * (cr * cr * cr)));
}

}  // namespace vc

namespace vc {

// ---- one SingleTrack stage, forward-over-reverse along ONE direction (T = Dual1 seeded on an active variable;
// T = double gives values and the plain gradient).
//   mode 0: interior stage (RK4 transition to the next SingleTrack stage, multipliers lam[8])
//   mode 1: last SingleTrack stage of a pure SingleTrack controller (no transition; terminal cost)
//   mode 2: switching stage (transition = st_switch into the PointMass, multipliers lam[5]); T = Fy_f + Fy_r enters the
//           Lagrangian with the CONSTANT weight kT = kt_w (fy_next - T), kt_w = -2 sf c1: the curvature part of the
//           switching cost sf c1 (Fy_N - T)^2; its Gauss-Newton part goes through the lifted variable of the Riccati sweep
// Outputs: cost, ineq[10], xn[8] (mode 0) or xn[0..4] = PointMass state and xn[5] = T (mode 2), all with their
// directional derivative; g = gradient of  l = sf cost - z.c - lam.Phi (+ kT T)  w.r.t. the 8 active variables, whose
// dual part is the Hessian column of the direction.
template <class T>
VC_HD void st_stage_for(const Params& p, const Cfg& c, const T* y, double kap, double ds, int mode, double sf, const double* z,
                        const double* lam, double kt_w, double fy_next, T& cost, T* ineq, T* xn, InAdj<T>& g, double* kT_out = nullptr) {
    T xs[4][6];        // RK4 stage points, active state components Ux Uy r delta ey epsi
    T x[8], f[8], acc[8];
    RhsTape<T> tp;
    g.Ux = g.Uy = g.r = g.delta = g.ey = g.epsi = g.Fx = g.w = zero<T>();
    const int map6[6] = {IUX, IUY, IR, IDELTA, IEY, IEPSI};
    double cb[NI_ST];   // adjoint of inequality row j in the Lagrangian: -z_j
#pragma unroll
    for (int j = 0; j < NI_ST; ++j) cb[j] = -z[j];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = y[i];
    if (mode != 0) {
        // no RK4: cost / inequality rows (+ switching map) at y
        tyres_fwd(p, y[IUX], y[IUY], y[IR], y[IDELTA], y[YFX], tp.ty);
        st_cost_ineq_fwd(p, c, y, tp.ty, ds, mode == 1, cost, ineq);
        TyreAdj<T> b;
        st_cost_ineq_rev(p, c, y, tp.ty, ds, mode == 1, sf, cb, b, g);
        if (mode == 2) {
            const T &Ux = y[IUX], &Uy = y[IUY];
            const T V = ad_sqrt(Ux * Ux + Uy * Uy);
            const T q = Uy * tp.ty.iux;
            xn[PV] = V; xn[PS] = y[IS]; xn[PEY] = y[IEY]; xn[PEPSI] = ad_atan(q) + y[IEPSI]; xn[PT] = y[IT];
            xn[5] = tp.ty.fyf + tp.ty.fyr;
            // adjoints: -lam . xp + kT T
            const T iV = recip(V);
            g.Ux = g.Ux - lam[PV] * (Ux * iV);
            g.Uy = g.Uy - lam[PV] * (Uy * iV);
            g.ey = g.ey - lam[PEY];
            g.epsi = g.epsi - lam[PEPSI];
            const T dq = recip(1.0 + q * q);
            const T q_b = -lam[PEPSI] * dq;
            g.Uy = g.Uy + q_b * tp.ty.iux;
            g.Ux = g.Ux - (q_b * Uy) * (tp.ty.iux * tp.ty.iux);
            const double kT = kt_w * (fy_next - val(xn[5]));
            if (kT_out) *kT_out = kT;
            b.fyf = b.fyf + kT; b.fyr = b.fyr + kT;
        }
        tyres_rev(p, y[IUX], y[IUY], y[IR], y[YFX], tp.ty, b, g);
        return;
    }
    // ---- forward RK4 (integrators.py:30-34 operation order, as st_rk4)
#pragma unroll 1
    for (int s = 0; s < 4; ++s) {
#pragma unroll
        for (int i = 0; i < 6; ++i) xs[s][i] = x[map6[i]];
        st_rhs_fwd(p, x, y + 8, kap, tp, f);
        if (s == 0) st_cost_ineq_fwd(p, c, y, tp.ty, ds, false, cost, ineq);
        const double wgt = (s == 0 || s == 3) ? 1.0 : 2.0;
        const double cc = s == 2 ? ds : 0.5 * ds;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (s == 0) acc[i] = f[i]; else acc[i] = acc[i] + wgt * f[i];
            x[i] = y[i] + cc * f[i];
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) xn[i] = y[i] + (ds * (1.0 / 6.0)) * acc[i];
    // ---- reverse sweep: adjoint of Phi is -lam (constant); k_b of stage s = w_s h/6 (-lam) + c_s xs_{s+1}_b
    T nb[6];           // adjoint of the NEXT stage point (active state components)
#pragma unroll
    for (int i = 0; i < 6; ++i) nb[i] = zero<T>();
#pragma unroll 1
    for (int s = 3; s >= 0; --s) {
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = y[i];       // s, t are not read by the rhs
#pragma unroll
        for (int i = 0; i < 6; ++i) x[map6[i]] = xs[s][i];
        st_rhs_fwd(p, x, y + 8, kap, tp, f);
        const double wgt = ((s == 0 || s == 3) ? 1.0 : 2.0) * (ds * (1.0 / 6.0));
        const double cn = s == 3 ? 0.0 : (s == 2 ? ds : 0.5 * ds);   // x_{s+1} = y + cn k_s
        T fb[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) fb[i] = cst<T>(-wgt * lam[i]);
#pragma unroll
        for (int i = 0; i < 6; ++i) fb[map6[i]] = fb[map6[i]] + cn * nb[i];
        InAdj<T> st;
        st.Ux = st.Uy = st.r = st.delta = st.ey = st.epsi = st.Fx = st.w = zero<T>();
        if (s == 0) {
            TyreAdj<T> b;
            st_cost_ineq_rev(p, c, y, tp.ty, ds, false, sf, cb, b, st);
            st_rhs_rev(p, x, y + 8, kap, tp, fb, &b, st);
        } else {
            st_rhs_rev<T>(p, x, y + 8, kap, tp, fb, nullptr, st);
        }
        nb[0] = st.Ux; nb[1] = st.Uy; nb[2] = st.r; nb[3] = st.delta; nb[4] = st.ey; nb[5] = st.epsi;
        g.Ux = g.Ux + st.Ux; g.Uy = g.Uy + st.Uy; g.r = g.r + st.r; g.delta = g.delta + st.delta;
        g.ey = g.ey + st.ey; g.epsi = g.epsi + st.epsi; g.Fx = g.Fx + st.Fx; g.w = g.w + st.w;
    }
    // direct dependence Phi = y + ...
    g.Ux = g.Ux - lam[IUX]; g.Uy = g.Uy - lam[IUY]; g.r = g.r - lam[IR]; g.delta = g.delta - lam[IDELTA];
    g.ey = g.ey - lam[IEY]; g.epsi = g.epsi - lam[IEPSI];
}

}  // namespace vc
