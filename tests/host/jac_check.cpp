// Host-side check of the hand-derived SingleTrack RK4 Jacobian (vehicle_control_b200/csrc/vc_jac.cuh, compiled with g++)
// against the forward-mode dual numbers of vc_ad.cuh / vc_model.cuh run column by column (what the first-generation
// rollout kernels did).  Prints the largest relative deviations; exit code 0 when values agree to 1e-13 and Jacobian
// entries to 1e-11 (relative to the largest entry of the block).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include "../../vehicle_control_b200/csrc/vc_jac.cuh"

using namespace vc;

static double urand(double a, double b) { return a + (b - a) * (rand() / (double)RAND_MAX); }

int main() {
    Params p{};
    p.dt = 0.05; p.l = 3; p.m = 1700; p.Izz = 3000; p.a = 1.1; p.b = 1.4; p.h = 0.55; p.eps = 0.9; p.Peng = 170000;
    p.Xdf = 1; p.Xdr = 0; p.Xbf = 0.78; p.Xbr = 0.22; p.Caf = 234000; p.Car = 390000; p.Cd = 0.4243; p.muf = 0.9; p.mur = 0.95;
    p.Av2 = 0.0; p.Frr = 220; p.g_eff = 9.88;
    double worst_v = 0, worst_j = 0;
    int cases = 0, saturated = 0;
    srand(11);
    for (int trial = 0; trial < 2000; ++trial) {
        double x[8] = {urand(4, 25), urand(-1.5, 1.5), urand(-0.8, 0.8), urand(-0.3, 0.3), urand(0, 300), urand(-3.5, 3.5),
                       urand(-0.3, 0.3), urand(0, 20)};
        double u[2] = {urand(-6000, 5000), urand(-0.4, 0.4)};
        if (trial % 5 == 0) { x[IUY] *= 3; x[IR] *= 2; }   // saturated tyres
        if (trial % 13 == 0) p.Av2 = 0.35; else p.Av2 = 0.0;   // aerodynamic downforce term (zero in the shipped configs)
        const double kap = urand(-0.05, 0.05), h = urand(0.1, 0.7);
        {
            Tyres<double> ty; tyres(p, x[IUX], x[IUY], x[IR], x[IDELTA], u[0], ty);
            if (!(ty.fymax_f == ty.fymax_f) || !(ty.fymax_r == ty.fymax_r)) continue;
            if (fabs(ty.tan_r) > 3.0 * ty.fymax_r * p.eps / p.Car) ++saturated;
        }
        double xn[8], AB[80];
        st_rk4_jac(p, x, u, kap, h, xn, AB);
        if (!(xn[0] == xn[0])) continue;
        ++cases;
        double ref[80], xr[8], scale = 1.0;
        for (int c = 0; c < 10; ++c) {
            Dual1 xd[8], ud[2], xo[8];
            for (int i = 0; i < 8; ++i) xd[i] = Dual1{x[i], c == i ? 1.0 : 0.0};
            ud[0] = Dual1{u[0], c == 8 ? 1.0 : 0.0}; ud[1] = Dual1{u[1], c == 9 ? 1.0 : 0.0};
            st_rk4<true>(p, xd, ud, kap, h, xo);
            for (int i = 0; i < 8; ++i) { ref[i * 10 + c] = xo[i].d; xr[i] = xo[i].v; scale = fmax(scale, fabs(xo[i].d)); }
        }
        for (int i = 0; i < 8; ++i) worst_v = fmax(worst_v, fabs(xn[i] - xr[i]) / fmax(1.0, fabs(xr[i])));
        for (int i = 0; i < 80; ++i) {
            const double e = fabs(AB[i] - ref[i]) / scale;
            if (e > 1e-9) printf("trial %d AB[%d][%d] = %.12e ref %.12e\n", trial, i / 10, i % 10, AB[i], ref[i]);
            worst_j = fmax(worst_j, e);
        }
    }
    printf("cases %d (rear tyre saturated in %d)  max rel dev: values %.3e  jacobians %.3e\n", cases, saturated, worst_v, worst_j);
    return (cases > 1500 && worst_v < 1e-13 && worst_j < 1e-11) ? 0 : 1;
}
