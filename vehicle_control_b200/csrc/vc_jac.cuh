// Hand-derived Jacobian of the SPATIAL SingleTrack right-hand side and the RK4 chain rule on it.
//
// Replaces what CasADi's AD does for `ca.jacobian(car.spatial_transition(x, u, k, ds), [x, u])`
// (vehicle_control/models/dynamic_car.py:78-191, utils/integrators.py:26-37).  The forward-mode dual numbers of
// vc_ad.cuh recompute the value part and every transcendental once per group of Jacobian columns; here the partial
// derivatives of one right-hand-side evaluation are formed once (39 structurally non-zero entries) and the four RK4
// stages are chained with sparse row products: about a third of the arithmetic of the dual-number kernel.
//
// Variables the right-hand side depends on ("active"): v = (Ux Uy r delta ey epsi | Fx w); s and t never enter f, and
// f[s] = 1, so columns s, t of A are unit vectors and row s is a unit vector.
#pragma once
#include "vc_model.cuh"

namespace vc {

// d f_spatial / d v in factored form.  With f_i = g_i / s_dot:  d f_i = g_i,v / s_dot - f_i (s_dot,v / s_dot), so every row
// is its "temporal" part (sparse) minus f_i times ONE common row sv = s_dot,v / s_dot (columns Ux Uy ey epsi).
struct StRhsJac {
    double g[3][5];        // rows Ux', Uy', r': temporal partials / s_dot, columns Ux Uy r delta Fx
    double sv[4];          // s_dot,v / s_dot, columns Ux Uy ey epsi
    double fU, fV, fR, fD, fE, ris, is;   // f of rows Ux Uy r delta ey, r / s_dot, 1 / s_dot
    double eU, eV, eP;     // row ey': temporal partials / s_dot, columns Ux Uy epsi
};

// partial derivatives of the modified Fiala force (same branches as fiala() in vc_model.cuh)
VC_HD void fiala_partials(double C, double eps, double ta, double fymax, double inv, double& fy, double& d_ta, double& d_fm) {
    const double tmod = 3.0 * fymax * eps / C;   // (C is a parameter: the quotient is formed as in fiala())
    if (fabs(ta) <= tmod) {
        const double at = fabs(ta);
        fy = -C * ta + (C * C / 3.0) * (at * ta) * inv - (C * C * C / 27.0) * (ta * ta * ta) * (inv * inv);
        d_ta = (2.0 * C * C / 3.0) * (at * inv) - (C * C * C / 9.0) * ((ta * ta) * (inv * inv)) - C;
        const double d_inv = (C * C / 3.0) * (at * ta) - (2.0 * C * C * C / 27.0) * ((ta * ta * ta) * inv);
        d_fm = -d_inv * (inv * inv);
    } else {
        const double sg = ta > 0.0 ? 1.0 : (ta < 0.0 ? -1.0 : 0.0);
        fy = -C * (1 - 2 * eps + eps * eps) * ta - (3 * eps * eps - 2 * eps * eps * eps) * sg * fymax;
        d_ta = -C * (1 - 2 * eps + eps * eps);
        d_fm = -(3 * eps * eps - 2 * eps * eps * eps) * sg;
    }
}

// what depends on the control Fx only (the same for the four RK4 stages of a step): drive/brake split and its derivative
struct StCtl { double fxf, fxr, fxf_F, fxr_F, shift; };
VC_HD void st_ctl(const Params& p, double Fx, StCtl& c) {
    const double th = tanh(2.0 * (Fx / 1000.0 + 0.5));
    const double dth = (1.0 - th * th) * (2.0 / 1000.0);
    const double cf1 = (p.Xdf - p.Xbf) / 2, cr1 = (p.Xdr - p.Xbr) / 2;
    const double Xf = cf1 * th + (p.Xdf + p.Xbf) / 2;
    const double Xr = cr1 * th + (p.Xdr + p.Xbr) / 2;
    c.fxf = Fx * Xf; c.fxr = Fx * Xr;
    c.fxf_F = Xf + Fx * (cf1 * dth); c.fxr_F = Xr + Fx * (cr1 * dth);
    c.shift = (p.h / p.l) * Fx;
}

// f = d x / d s (the operations of st_rhs<true, double>; divisions by the constants m, Izz are multiplications by their
// reciprocals) and its Jacobian
VC_HD void st_rhs_spatial_jac(const Params& p, const StCtl& ct, const double* x, double w, double kappa, double* f, StRhsJac& J) {
    const double Ux = x[IUX], Uy = x[IUY], r = x[IR], delta = x[IDELTA], ey = x[IEY], epsi = x[IEPSI];
    // ---- tyres (vc_model.cuh: force_split, normal_loads, tyres)
    const double fxf = ct.fxf, fxr = ct.fxr, fxf_F = ct.fxf_F, fxr_F = ct.fxr_F, shift = ct.shift;
    const double base = p.g_eff + p.Av2 * (Ux * Ux);
    const double fzf = (p.b / p.l) * p.m * base - shift;
    const double fzr = (p.a / p.l) * p.m * base + shift;
    const double base_U = 2.0 * p.Av2 * Ux;
    const double lf = p.muf * fzf, lr = p.mur * fzr;
    const double qf = 0.99 * fxf, qr = 0.99 * fxr;
    const double fmf = sqrt(lf * lf - qf * qf);
    const double fmr = sqrt(lr * lr - qr * qr);
    const double ifmf = 1.0 / fmf, ifmr = 1.0 / fmr;
    // d fymax = (l dl - q dq) / fymax
    const double fmf_U = lf * (p.muf * (p.b / p.l) * p.m * base_U) * ifmf;
    const double fmr_U = lr * (p.mur * (p.a / p.l) * p.m * base_U) * ifmr;
    const double fmf_F = (lf * (-p.muf * (p.h / p.l)) - qf * (0.99 * fxf_F)) * ifmf;
    const double fmr_F = (lr * (p.mur * (p.h / p.l)) - qr * (0.99 * fxr_F)) * ifmr;
    const double iux = 1.0 / Ux;
    const double zf = (Uy + p.a * r) * iux;
    double sd, cd;
#if defined(__CUDA_ARCH__)
    sincos(delta, &sd, &cd);
#else
    sd = sin(delta); cd = cos(delta);
#endif
    const double iden = 1.0 / (cd + zf * sd);
    const double tan_f = (zf * cd - sd) * iden;
    const double tan_r = (Uy - p.b * r) * iux;
    double fyf, fyr, df_ta, df_fm, dr_ta, dr_fm;
    fiala_partials(p.Caf, p.eps, tan_f, fmf, ifmf, fyf, df_ta, df_fm);
    fiala_partials(p.Car, p.eps, tan_r, fmr, ifmr, fyr, dr_ta, dr_fm);
    // d tan_f / d zf = 1 / den^2 ; d tan_f / d delta = -(1 + tan_f^2)
    const double tf_z = df_ta * (iden * iden);
    const double fyf_U = tf_z * (-zf * iux) + df_fm * fmf_U;
    const double fyf_V = tf_z * iux;
    const double fyf_r = tf_z * (p.a * iux);
    const double fyf_d = -df_ta * (1.0 + tan_f * tan_f);
    const double fyf_F = df_fm * fmf_F;
    const double fyr_U = dr_ta * (-tan_r * iux) + dr_fm * fmr_U;
    const double fyr_V = dr_ta * iux;
    const double fyr_r = dr_ta * (-p.b * iux);
    const double fyr_F = dr_fm * fmr_F;
    // ---- chassis (st_rhs)
    double se, ce;
#if defined(__CUDA_ARCH__)
    sincos(epsi, &se, &ce);
#else
    se = sin(epsi); ce = cos(epsi);
#endif
    const double Fd = p.Frr + p.Cd * (Ux * Ux);
    const double lat = fyf * cd + fxf * sd;
    const double im = 1.0 / p.m, iI = 1.0 / p.Izz;
    const double Ux_dot = (fxf * cd - fyf * sd + fxr - Fd) * im + r * Uy;
    const double Uy_dot = (lat + fyr) * im - r * Ux;
    const double r_dot = (p.a * lat - p.b * fyr) * iI;
    const double q = 1.0 / (1.0 - kappa * ey);
    const double s_dot = (Ux * ce - Uy * se) * q;
    const double ey_dot = Ux * se + Uy * ce;
    const double epsi_dot = r - kappa * s_dot;
    const double is = 1.0 / s_dot;
    f[IUX] = Ux_dot * is; f[IUY] = Uy_dot * is; f[IR] = r_dot * is; f[IDELTA] = w * is;
    f[IS] = 1.0; f[IEY] = ey_dot * is; f[IEPSI] = epsi_dot * is; f[IT] = is;
    // ---- temporal partials g_v of (Ux_dot, Uy_dot, r_dot), columns Ux Uy r delta Fx
    const double lat_U = fyf_U * cd, lat_V = fyf_V * cd, lat_r = fyf_r * cd;
    const double lat_d = fyf_d * cd - fyf * sd + fxf * cd;
    const double lat_F = fyf_F * cd + fxf_F * sd;
    const double a_U = (-fyf_U * sd - 2.0 * p.Cd * Ux) * im;
    const double a_V = (-fyf_V * sd) * im + r;
    const double a_r = (-fyf_r * sd) * im + Uy;
    const double a_d = (-fxf * sd - fyf_d * sd - fyf * cd) * im;
    const double a_F = (fxf_F * cd - fyf_F * sd + fxr_F) * im;
    const double b_U = (lat_U + fyr_U) * im - r;
    const double b_V = (lat_V + fyr_V) * im;
    const double b_r = (lat_r + fyr_r) * im - Ux;
    const double b_d = lat_d * im;
    const double b_F = (lat_F + fyr_F) * im;
    const double c_U = (p.a * lat_U - p.b * fyr_U) * iI;
    const double c_V = (p.a * lat_V - p.b * fyr_V) * iI;
    const double c_r = (p.a * lat_r - p.b * fyr_r) * iI;
    const double c_d = (p.a * lat_d) * iI;
    const double c_F = (p.a * lat_F - p.b * fyr_F) * iI;
    // s_dot partials (columns Ux Uy ey epsi), pre-multiplied by 1/s_dot:  sv = s_v / s_dot
    J.sv[0] = (ce * q) * is; J.sv[1] = (-se * q) * is; J.sv[2] = kappa * q; J.sv[3] = (-ey_dot * q) * is;
    J.g[0][0] = a_U * is; J.g[0][1] = a_V * is; J.g[0][2] = a_r * is; J.g[0][3] = a_d * is; J.g[0][4] = a_F * is;
    J.g[1][0] = b_U * is; J.g[1][1] = b_V * is; J.g[1][2] = b_r * is; J.g[1][3] = b_d * is; J.g[1][4] = b_F * is;
    J.g[2][0] = c_U * is; J.g[2][1] = c_V * is; J.g[2][2] = c_r * is; J.g[2][3] = c_d * is; J.g[2][4] = c_F * is;
    J.fU = f[IUX]; J.fV = f[IUY]; J.fR = f[IR]; J.fD = f[IDELTA]; J.fE = f[IEY]; J.ris = r * is; J.is = is;
    // ey_dot: /Ux = se, /Uy = ce, /epsi = Ux ce - Uy se
    J.eU = se * is; J.eV = ce * is; J.eP = (Ux * ce - Uy * se) * is;
}

// K[:, c] = J * [D[:, c]; e_u(c)] for one column c of the sensitivity D = d(active states)/d v0 (rows Ux Uy r delta ey
// epsi).  `d` = that column (6 values), cu = 0: state column, 1: column Fx, 2: column w.  k: rows Ux Uy r delta ey epsi t.
VC_HD void st_jac_col(const StRhsJac& J, const double* d, int cu, double* k) {
    const double t = J.sv[0] * d[0] + J.sv[1] * d[1] + J.sv[2] * d[4] + J.sv[3] * d[5];
    k[0] = J.g[0][0] * d[0] + J.g[0][1] * d[1] + J.g[0][2] * d[2] + J.g[0][3] * d[3] - J.fU * t;
    k[1] = J.g[1][0] * d[0] + J.g[1][1] * d[1] + J.g[1][2] * d[2] + J.g[1][3] * d[3] - J.fV * t;
    k[2] = J.g[2][0] * d[0] + J.g[2][1] * d[1] + J.g[2][2] * d[2] + J.g[2][3] * d[3] - J.fR * t;
    k[3] = -J.fD * t;
    k[4] = J.eU * d[0] + J.eV * d[1] + J.eP * d[5] - J.fE * t;
    k[5] = J.is * d[2] - J.ris * t;
    k[6] = -J.is * t;
    if (cu == 1) { k[0] += J.g[0][4]; k[1] += J.g[1][4]; k[2] += J.g[2][4]; }
    if (cu == 2) k[3] += J.is;
}

// column c of J itself (the product with D = [I | 0] of the first RK4 stage): c = Ux Uy r delta ey epsi Fx w
VC_HD void st_jac_unitcol(const StRhsJac& J, int c, double* k) {
    const int ms[8] = {0, 1, -1, -1, 2, 3, -1, -1};   // position of column c in sv
    const int mg[8] = {0, 1, 2, 3, -1, -1, 4, -1};    // ... in the rows g
    const bool ht = ms[c] >= 0, hg = mg[c] >= 0;
    const double t = ht ? J.sv[ms[c]] : 0.0;
    k[0] = (hg ? J.g[0][mg[c]] : 0.0) - (ht ? J.fU * t : 0.0);
    k[1] = (hg ? J.g[1][mg[c]] : 0.0) - (ht ? J.fV * t : 0.0);
    k[2] = (hg ? J.g[2][mg[c]] : 0.0) - (ht ? J.fR * t : 0.0);
    k[3] = (c == 7 ? J.is : 0.0) - (ht ? J.fD * t : 0.0);
    k[4] = (c == 0 ? J.eU : (c == 1 ? J.eV : (c == 5 ? J.eP : 0.0))) - (ht ? J.fE * t : 0.0);
    k[5] = (c == 2 ? J.is : 0.0) - (ht ? J.ris * t : 0.0);
    k[6] = ht ? -J.is * t : 0.0;
}

// One RK4 step with its Jacobian, plain doubles, one caller per trajectory (host check and reference for the kernel).
// AB[8][10] row-major: d x+ / d (Ux Uy r delta s ey epsi t | Fx w).
VC_HD void st_rk4_jac(const Params& p, const double* x, const double* u, double kappa, double h, double* xn, double* AB) {
    // active column c of v0 -> column of AB; active row i -> state row
    const int colmap[8] = {0, 1, 2, 3, 5, 6, 8, 9};
    const int rowmap[7] = {IUX, IUY, IR, IDELTA, IEY, IEPSI, IT};
    double D[8][6], acc[8][7], xs[8], f[8], facc[8];
    for (int c = 0; c < 8; ++c)
        for (int i = 0; i < 6; ++i) D[c][i] = (c == i) ? 1.0 : 0.0;
    for (int i = 0; i < 8; ++i) xs[i] = x[i];
    StCtl ct;
    st_ctl(p, u[0], ct);
    for (int s = 0; s < 4; ++s) {
        StRhsJac J;
        st_rhs_spatial_jac(p, ct, xs, u[1], kappa, f, J);
        const double wgt = (s == 0 || s == 3) ? 1.0 : 2.0;
        const double cs = s == 2 ? h : 0.5 * h;
        for (int i = 0; i < 8; ++i) {
            if (s == 0) facc[i] = f[i]; else facc[i] = facc[i] + wgt * f[i];
            xs[i] = x[i] + cs * f[i];
        }
        for (int c = 0; c < 8; ++c) {
            double k[7];
            if (s == 0) st_jac_unitcol(J, c, k); else st_jac_col(J, D[c], c < 6 ? 0 : c - 5, k);
            for (int i = 0; i < 7; ++i) acc[c][i] = s == 0 ? k[i] : acc[c][i] + wgt * k[i];
            for (int i = 0; i < 6; ++i) D[c][i] = ((c == i) ? 1.0 : 0.0) + cs * k[i];
        }
    }
    for (int i = 0; i < 8; ++i) xn[i] = x[i] + (h * (1.0 / 6.0)) * facc[i];
    for (int i = 0; i < 80; ++i) AB[i] = 0.0;
    AB[IS * 10 + IS] = 1.0; AB[IT * 10 + IT] = 1.0;
    for (int c = 0; c < 8; ++c)
        for (int i = 0; i < 7; ++i)
            AB[rowmap[i] * 10 + colmap[c]] = ((rowmap[i] == colmap[c]) ? 1.0 : 0.0) + (h * (1.0 / 6.0)) * acc[c][i];
}

}  // namespace vc
