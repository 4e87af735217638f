// Host-side check of the hand-written adjoints (vehicle_control_b200/csrc/vc_adj.cuh, compiled with g++): the
// forward-over-reverse gradient and Hessian columns of the SingleTrack stage Lagrangian against the hyper-dual (pair per
// evaluation) templates of vc_model.cuh / vc_nlp.cuh that the first-generation k_eval_st used.  Prints the largest
// relative deviations; exit code 0 when all are below 1e-10.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include "../../vehicle_control_b200/csrc/vc_adj.cuh"

using namespace vc;

static double urand(double a, double b) { return a + (b - a) * (rand() / (double)RAND_MAX); }

int main() {
    Params p{};
    p.dt = 0.05; p.l = 3; p.m = 1700; p.Izz = 3000; p.a = 1.1; p.b = 1.4; p.h = 0.55; p.eps = 0.9; p.Peng = 170000;
    p.Xdf = 1; p.Xdr = 0; p.Xbf = 0.78; p.Xbr = 0.22; p.Caf = 234000; p.Car = 390000; p.Cd = 0.4243; p.muf = 0.9; p.mur = 0.95;
    p.Av2 = 0.0; p.Frr = 220; p.g_eff = 9.88;
    Cfg c{};
    c.N = 20; c.M = 35; c.w_time = 5; c.w_speed = 1; c.w_ey = 1; c.w_epsi = 2; c.w_w = 4; c.w_Fx = 1e-6; c.w_Fy = 1e-6;
    c.w_dev_st = 1e-3; c.w_dev_pm = 1e-3; c.w_boundary = 10; c.w_slip = 1500; c.w_obstacles = 10; c.w_switch_F = 1e-5;
    c.w_min = -0.4; c.w_max = 0.4; c.Ux_min = 3; c.max_speed = 18; c.delta_min = -0.45; c.delta_max = 0.45; c.ey_min = -3; c.ey_max = 3;
    c.V_min = 3; c.ey_min_pm = -3; c.ey_max_pm = 3; c.bound_relax = 1e-8;
    const int act[8] = {IUX, IUY, IR, IDELTA, IEY, IEPSI, YFX, YW};
    double worst_g = 0, worst_h = 0, worst_v = 0, worst_j = 0;
    srand(7);
    int cases = 0;
    for (int trial = 0; trial < 600; ++trial) {
        const int mode = trial % 3;
        double yv[10] = {urand(4, 25), urand(-1.5, 1.5), urand(-0.8, 0.8), urand(-0.3, 0.3), urand(0, 300), urand(-3.5, 3.5),
                         urand(-0.3, 0.3), urand(0, 20), urand(-6000, 5000), urand(-0.4, 0.4)};
        if (trial % 7 == 0) { yv[IUY] *= 3; yv[IR] *= 2; }          // saturated tyres / slip penalty active
        if (trial % 11 == 0) yv[IUX] = urand(17.5, 19);             // terminal speed penalty
        const double kap = urand(0, 0.05), ds = urand(0.1, 0.7), sf = urand(0.01, 1.0), kt_w = urand(-1e-4, 0), fy_next = urand(-8000, 8000);
        double z[10], lam[8];
        for (int j = 0; j < 10; ++j) z[j] = urand(0, 2);
        for (int i = 0; i < 8; ++i) lam[i] = urand(-3, 3);
        // friction limit must be real: |Fx_f| < mu Fz / 0.99
        {
            double ytest[10]; for (int i = 0; i < 10; ++i) ytest[i] = yv[i];
            Tyres<double> ty; tyres(p, ytest[IUX], ytest[IUY], ytest[IR], ytest[IDELTA], ytest[YFX], ty);
            if (!(ty.fymax_f == ty.fymax_f) || !(ty.fymax_r == ty.fymax_r)) continue;
        }
        ++cases;
        // ---- reference: hyper-dual pairs
        double Href[8][8], gref[8], vals[19], Jref[18][8];
        for (int ia = 0; ia < 8; ++ia)
            for (int ib = 0; ib <= ia; ++ib) {
                HDual y[10];
                for (int i = 0; i < 10; ++i) y[i] = HDual{yv[i], i == act[ia] ? 1.0 : 0.0, i == act[ib] ? 1.0 : 0.0, 0.0};
                HDual cost, ineq[10];
                st_cost_ineq(p, c, y, ds, mode == 1, cost, ineq);
                double hab = sf * cost.ab, ga = sf * cost.a;
                for (int j = 0; j < 10; ++j) { hab -= z[j] * ineq[j].ab; ga -= z[j] * ineq[j].a; if (ia == ib) Jref[j][ia] = ineq[j].a; }
                vals[0] = cost.v;
                for (int j = 0; j < 10; ++j) vals[1 + j] = ineq[j].v;
                if (mode == 0) {
                    HDual xn[8];
                    st_rk4<true>(p, y, y + 8, kap, ds, xn);
                    for (int i = 0; i < 8; ++i) { hab -= lam[i] * xn[i].ab; ga -= lam[i] * xn[i].a; vals[11 + i] = xn[i].v; if (ia == ib) Jref[10 + i][ia] = xn[i].a; }
                } else if (mode == 2) {
                    HDual xp[5], T;
                    st_switch(p, y, xp, T);
                    for (int i = 0; i < 5; ++i) { hab -= lam[i] * xp[i].ab; ga -= lam[i] * xp[i].a; vals[11 + i] = xp[i].v; if (ia == ib) Jref[10 + i][ia] = xp[i].a; }
                    const double kT = kt_w * (fy_next - T.v);
                    hab += kT * T.ab; ga += kT * T.a; vals[16] = T.v; if (ia == ib) Jref[15][ia] = T.a;
                }
                Href[ia][ib] = Href[ib][ia] = hab;
                if (ia == ib) gref[ia] = ga;
            }
        // ---- forward over reverse, one direction at a time
        for (int ia = 0; ia < 8; ++ia) {
            Dual1 y[10], cost, ineq[10], xn[8];
            for (int i = 0; i < 10; ++i) y[i] = Dual1{yv[i], i == act[ia] ? 1.0 : 0.0};
            InAdj<Dual1> g;
            st_stage_for<Dual1>(p, c, y, kap, ds, mode, sf, z, lam, kt_w, fy_next, cost, ineq, xn, g);
            const Dual1 gg[8] = {g.Ux, g.Uy, g.r, g.delta, g.ey, g.epsi, g.Fx, g.w};
            for (int ib = 0; ib < 8; ++ib) {
                worst_g = fmax(worst_g, fabs(gg[ib].v - gref[ib]) / fmax(1.0, fabs(gref[ib])));
                const double e = fabs(gg[ib].d - Href[ib][ia]) / fmax(1.0, fabs(Href[ib][ia]));
                if (e > 1e-9) printf("trial %d mode %d H[%d][%d] = %.12e ref %.12e\n", trial, mode, ib, ia, gg[ib].d, Href[ib][ia]);
                worst_h = fmax(worst_h, e);
            }
            worst_v = fmax(worst_v, fabs(cost.v - vals[0]) / fmax(1.0, fabs(vals[0])));
            for (int j = 0; j < 10; ++j) {
                worst_v = fmax(worst_v, fabs(ineq[j].v - vals[1 + j]) / fmax(1.0, fabs(vals[1 + j])));
                worst_j = fmax(worst_j, fabs(ineq[j].d - Jref[j][ia]) / fmax(1.0, fabs(Jref[j][ia])));
            }
            const int nout = mode == 0 ? 8 : (mode == 2 ? 6 : 0);
            for (int i = 0; i < nout; ++i) {
                worst_v = fmax(worst_v, fabs(xn[i].v - vals[11 + i]) / fmax(1.0, fabs(vals[11 + i])));
                worst_j = fmax(worst_j, fabs(xn[i].d - Jref[10 + i][ia]) / fmax(1.0, fabs(Jref[10 + i][ia])));
            }
        }
    }
    printf("cases %d  max rel dev: values %.3e  jacobians %.3e  gradient %.3e  hessian %.3e\n", cases, worst_v, worst_j, worst_g, worst_h);
    return (worst_v < 1e-13 && worst_j < 1e-11 && worst_g < 1e-10 && worst_h < 1e-10) ? 0 : 1;
}
