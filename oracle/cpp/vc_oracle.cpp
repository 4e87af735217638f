// ORACLE (test infrastructure only -- never linked into or loaded by the product path; only tests/, smoke() and
// bench.py's cpu_baseline / --impl reference legs may use it).
//
// C++17 fp64 restatement of the reference's NMPC solve for the CPU: CascadedMPC.command -> opti.solve()
// (vehicle_control/controllers/mpc/cascaded_mpc.py:306-314), i.e. the NLP of cascaded_mpc.py:68-304 with exact
// derivatives (jet.hpp, the role of CasADi's AD) and IPOPT's published algorithm (Waechter & Biegler, Math. Prog. 106,
// 2006; IPOPT 3.14 defaults as the reference leaves them, cascaded_mpc.py:55-65): monotone barrier, fraction to the
// boundary, filter line search with second-order correction, delta_w inertia correction, gradient-based objective
// scaling, bound_relax_factor 1e-8, acceptable_tol 1e-6 / 15.  The KKT system is solved by a block-tridiagonal
// LDL^T (forward block elimination over the stages, Bunch-Kaufman pivoting inside the <= 18x18 blocks, inertia from
// the pivots) -- the role of MA27 in the reference, and a different factorisation from the product's Riccati sweep.
// IPOPT's restoration phase is NOT restated: where IPOPT would restore, the step is damped (larger delta_w) and, if that
// fails, the iterate is re-initialised by a dynamically feasible rollout (same substitute as oracle/ipm.py).
//
// Parity status: pinned by IPOPT's saved first actions of the reference's runs (tests/golden/nmpc_oracle.npz, <= 1e-6)
// and cross-checked against the independent Python oracle (oracle/nlp.py + oracle/ipm.py: torch autodiff + dense
// LDL^T) on the same instances (tests/test_oracle_cpp.py).
//
// Build: oracle/cpp/Makefile (g++ -O3 -fopenmp -shared) -> oracle/_ref/libvc_oracle.so
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "model.hpp"

namespace vco {

enum { ST_SOLVED = 0, ST_ACCEPTABLE = 1, ST_MAX_ITER = 2, ST_EVAL_ERROR = 3, ST_LINESEARCH_FAIL = 4 };

struct Options {          // extern "C" mirror: vco_options
    int max_iter;         // IPOPT default 3000
    int max_soc;          // IPOPT default 4
    int max_restorations; // re-initialisations by rollout where IPOPT would enter restoration (substitute)
    int verbose;
    double obj_scale;     // > 0: fixed objective scaling (else gradient based at the initial point)
    int watchdog;         // IPOPT watchdog_shortened_iter_trigger (default 10; 0 = off)
};

// IPOPT defaults (the reference sets none of them)
static constexpr double KAPPA_EPS = 10.0, KAPPA_MU = 0.2, THETA_MU = 1.5, TAU_MIN = 0.99, SLACK_PUSH = 1e-3,
                        KAPPA_SIGMA = 1e10, S_MAX = 100.0, GAMMA_THETA = 1e-5, GAMMA_PHI = 1e-8, ETA_PHI = 1e-8,
                        LS_DELTA = 1.0, S_THETA = 1.1, S_PHI = 2.3, MAX_GRADIENT = 100.0, KAPPA_SOC = 0.99;
static constexpr int MAX_BACKTRACK = 12;

// ------------------------------------------------------------------------------------------------ small dense LDL^T
// Bunch-Kaufman (partial pivoting, 1x1 and 2x2 pivots): P A P^T = L D L^T for a symmetric indefinite matrix, n <= NB.
constexpr int NB = 18;
struct Ldl {
    int n = 0;
    double a[NB][NB];
    int perm[NB], ptype[NB];
    int pos = 0, neg = 0, zero = 0;
    void factor(const double* A, int n_, int lda) {
        n = n_; pos = neg = zero = 0;
        for (int i = 0; i < n; ++i) { perm[i] = i; ptype[i] = 0; for (int j = 0; j < n; ++j) a[i][j] = A[i * lda + j]; }
        const double alpha = (1.0 + std::sqrt(17.0)) / 8.0;
        int k = 0;
        while (k < n) {
            const double absakk = std::fabs(a[k][k]);
            int imax = k; double colmax = 0.0;
            for (int i = k + 1; i < n; ++i) if (std::fabs(a[i][k]) > colmax) { colmax = std::fabs(a[i][k]); imax = i; }
            int kstep = 1, kp = k;
            if (!(std::fmax(absakk, colmax) > 0.0)) {   // zero column (or NaN): singular pivot
                ptype[k] = 1; ++zero; a[k][k] = 0.0; ++k; continue;
            }
            if (absakk < alpha * colmax) {
                double rowmax = 0.0;
                for (int j = k; j < n; ++j) if (j != imax) rowmax = std::fmax(rowmax, std::fabs(a[imax][j]));
                if (absakk >= alpha * colmax * (colmax / rowmax)) kp = k;
                else if (std::fabs(a[imax][imax]) >= alpha * rowmax) kp = imax;
                else { kp = imax; kstep = 2; }
            }
            const int kk = k + kstep - 1;
            if (kp != kk) {
                for (int j = 0; j < n; ++j) std::swap(a[kk][j], a[kp][j]);   // rows (incl. the finished columns of L)
                for (int i = k; i < n; ++i) std::swap(a[i][kk], a[i][kp]);   // columns of the trailing matrix
                std::swap(perm[kk], perm[kp]);
            }
            if (kstep == 1) {
                const double d = a[k][k];
                ptype[k] = 1;
                if (d > 0) ++pos; else if (d < 0) ++neg; else ++zero;
                double l[NB];
                for (int i = k + 1; i < n; ++i) l[i] = a[i][k] / d;
                for (int i = k + 1; i < n; ++i)
                    for (int j = k + 1; j < n; ++j) a[i][j] -= l[i] * a[j][k];
                for (int i = k + 1; i < n; ++i) { a[i][k] = l[i]; a[k][i] = 0.0; }
                k += 1;
            } else {
                const double d11 = a[k][k], d21 = a[k + 1][k], d22 = a[k + 1][k + 1];
                const double det = d11 * d22 - d21 * d21;
                ptype[k] = 2; ptype[k + 1] = 0;
                if (det < 0) { ++pos; ++neg; }
                else if (det > 0) { if (d11 > 0) pos += 2; else neg += 2; }
                else { ++zero; if (d11 + d22 > 0) ++pos; else ++neg; }
                double l1[NB], l2[NB];
                for (int i = k + 2; i < n; ++i) {
                    l1[i] = (a[i][k] * d22 - a[i][k + 1] * d21) / det;
                    l2[i] = (-a[i][k] * d21 + a[i][k + 1] * d11) / det;
                }
                for (int i = k + 2; i < n; ++i)
                    for (int j = k + 2; j < n; ++j) a[i][j] -= l1[i] * a[j][k] + l2[i] * a[j][k + 1];
                for (int i = k + 2; i < n; ++i) { a[i][k] = l1[i]; a[i][k + 1] = l2[i]; a[k][i] = a[k + 1][i] = 0.0; }
                k += 2;
            }
        }
    }
    void solve(const double* b, double* x) const {
        double y[NB];
        for (int i = 0; i < n; ++i) y[i] = b[perm[i]];
        for (int k = 0; k < n;) {   // L y' = y
            const int st = ptype[k] == 2 ? 2 : 1;
            for (int c = k; c < k + st; ++c)
                for (int i = k + st; i < n; ++i) y[i] -= a[i][c] * y[c];
            k += st;
        }
        for (int k = 0; k < n;) {   // D
            if (ptype[k] == 2) {
                const double d11 = a[k][k], d21 = a[k + 1][k], d22 = a[k + 1][k + 1], det = d11 * d22 - d21 * d21;
                const double y0 = y[k], y1 = y[k + 1];
                y[k] = (y0 * d22 - y1 * d21) / det; y[k + 1] = (-y0 * d21 + y1 * d11) / det;
                k += 2;
            } else { y[k] = a[k][k] != 0.0 ? y[k] / a[k][k] : 0.0; k += 1; }
        }
        for (int k = n - 1; k >= 0;) {   // L^T
            const int st = (k > 0 && ptype[k] == 0) ? 2 : 1;
            const int k0 = k - st + 1;
            for (int c = k0; c <= k; ++c)
                for (int i = k + 1; i < n; ++i) y[c] -= a[i][c] * y[i];
            k = k0 - 1;
        }
        for (int i = 0; i < n; ++i) x[perm[i]] = y[i];
    }
};

// ------------------------------------------------------------------------------------------------ problem / iterate
struct StageIt { double y[10], lam[8], s[10], z[10]; };
struct StageEv {
    double f;            // objective share of this stage (own terms + the coupling to the next stage's controls)
    double gf[10];       // d(objective)/d y_k, unscaled, all contributions
    double c[10];        // inequality values
    double Jc[10][10];   // [row][var]
    double g[8];         // residual of the equality that defines x_k
    double AB[8][10];    // d Phi_k / d y_k  (rows: state k+1)
    double W[10][10];    // Hessian of the Lagrangian, block (y_k, y_k)
    double Wx[2][10];    // cross block (u_{k+1}, y_k)
    double dlag[10];     // gradient of the Lagrangian w.r.t. y_k (dual infeasibility)
};

struct Problem {
    Par p; Cfg c;
    int N, M, H, n_obs;
    const double *x0, *kappa, *ds, *obst;
    int nx(int k) const { return k < N ? 8 : 5; }
    int nv(int k) const { return nx(k) + 2; }
    int ni(int k) const { return k < N ? 10 : 2; }
    void slew(int k, double& c0, double& c1) const {   // cascaded_mpc.py:167-171, :228-231, :241-254
        c0 = c1 = 0.0;
        if (k < 0 || k >= H - 1) return;
        if (k < N - 1) c0 = c.w_Fx / ds[k];
        else if (k == N - 1) c0 = c1 = c.w_switch_F / ds[k];
        else { c0 = c.w_Fx / ds[k]; c1 = c.w_Fy / ds[k]; }
    }
};

// values at a point: objective, equality residuals g[k][..], inequality values c[k][..]
static bool eval_values(const Problem& P, const std::vector<StageIt>& it, double& f, std::vector<StageEv>& ev) {
    const int N = P.N, H = P.H;
    f = 0.0;
    for (int k = 0; k < H; ++k) {
        const double* y = it[k].y;
        StageEv& e = ev[k];
        double cost, T = 0.0, c0, c1;
        P.slew(k, c0, c1);
        if (k < N) {
            st_stage<double>(P.p, P.c, y, P.ds[k], P.obst, P.n_obs, k == H - 1, cost, e.c, (k == N - 1 && P.M > 0) ? &T : nullptr);
            if (k == 0) for (int i = 0; i < 8; ++i) e.g[i] = y[i] - P.x0[i];
            if (k < N - 1) {
                double xn[8];
                st_rk4<true, double>(P.p, y, y + 8, P.kappa[k], P.ds[k], xn);
                for (int i = 0; i < 8; ++i) ev[k + 1].g[i] = it[k + 1].y[i] - xn[i];
                cost += c0 * (it[k + 1].y[FX] - y[FX]) * (it[k + 1].y[FX] - y[FX]);
            } else if (P.M > 0) {
                double xp[5];
                st_switch<double>(y, xp);
                for (int i = 0; i < 5; ++i) ev[k + 1].g[i] = it[k + 1].y[i] - xp[i];
                const double e0 = it[k + 1].y[PFX] - y[FX], e1 = it[k + 1].y[PFY] - T;
                cost += c0 * e0 * e0 + c1 * e1 * e1;
            }
        } else {
            pm_stage<double>(P.p, P.c, y, P.ds[k], P.obst, P.n_obs, k == H - 1, cost, e.c);
            if (k == 0) for (int i = 0; i < 5; ++i) e.g[i] = y[i] - P.x0[i];
            if (k < H - 1) {
                double xn[5];
                pm_step<double>(P.p, y, y + 5, P.kappa[k], P.ds[k], xn);
                for (int i = 0; i < 5; ++i) ev[k + 1].g[i] = it[k + 1].y[i] - xn[i];
                const double e0 = it[k + 1].y[PFX] - y[PFX], e1 = it[k + 1].y[PFY] - y[PFY];
                cost += c0 * e0 * e0 + c1 * e1 * e1;
            }
        }
        e.f = cost;
        f += cost;
    }
    if (!std::isfinite(f)) return false;
    for (int k = 0; k < H; ++k) {
        for (int i = 0; i < P.nx(k); ++i) if (!std::isfinite(ev[k].g[i])) return false;
        for (int j = 0; j < P.ni(k); ++j) if (!std::isfinite(ev[k].c[j])) return false;
    }
    return true;
}

// values + exact derivatives: objective gradient, constraint Jacobians, Hessian of
//   L = sf f + sum_k lam_k . g_k - sum_k z_k . c_k
template <int NV> static void take_hess(const Jet<NV>& q, double w, double W[10][10]) {
    for (int i = 0; i < NV; ++i) for (int j = 0; j < NV; ++j) W[i][j] += w * q.hess(i, j);
}
static bool eval_all(const Problem& P, const std::vector<StageIt>& it, double sf, double& f, std::vector<StageEv>& ev) {
    const int N = P.N, H = P.H;
    f = 0.0;
    for (int k = 0; k < H; ++k) { std::memset(ev[k].gf, 0, sizeof ev[k].gf); std::memset(ev[k].W, 0, sizeof ev[k].W); std::memset(ev[k].Wx, 0, sizeof ev[k].Wx); std::memset(ev[k].Jc, 0, sizeof ev[k].Jc); std::memset(ev[k].AB, 0, sizeof ev[k].AB); }
    for (int k = 0; k < H; ++k) {
        StageEv& e = ev[k];
        const double* yv = it[k].y;
        double c0, c1;
        P.slew(k, c0, c1);
        if (k < N) {
            typedef Jet<10> J;
            J y[10], cost, ineq[10], T;
            for (int i = 0; i < 10; ++i) y[i] = J::var(yv[i], i);
            const bool sw = k == N - 1 && P.M > 0;
            st_stage<J>(P.p, P.c, y, P.ds[k], P.obst, P.n_obs, k == H - 1, cost, ineq, sw ? &T : nullptr);
            if (k == 0) for (int i = 0; i < 8; ++i) e.g[i] = yv[i] - P.x0[i];
            if (k < N - 1) {
                J xn[8];
                st_rk4<true, J>(P.p, y, y + 8, P.kappa[k], P.ds[k], xn);
                const double* ln = it[k + 1].lam;
                for (int i = 0; i < 8; ++i) {
                    ev[k + 1].g[i] = it[k + 1].y[i] - xn[i].v;
                    for (int a = 0; a < 10; ++a) e.AB[i][a] = xn[i].g[a];
                    take_hess<10>(xn[i], -ln[i], e.W);
                }
                const double e0 = it[k + 1].y[FX] - yv[FX];
                cost.v += c0 * e0 * e0;
                e.gf[FX] += -2.0 * c0 * e0; ev[k + 1].gf[FX] += 2.0 * c0 * e0;
                e.W[FX][FX] += sf * 2.0 * c0; ev[k + 1].W[FX][FX] += sf * 2.0 * c0; e.Wx[0][FX] += -sf * 2.0 * c0;
            } else if (sw) {
                J xp[5];
                st_switch<J>(y, xp);
                const double* ln = it[k + 1].lam;
                for (int i = 0; i < 5; ++i) {
                    ev[k + 1].g[i] = it[k + 1].y[i] - xp[i].v;
                    for (int a = 0; a < 10; ++a) e.AB[i][a] = xp[i].g[a];
                    take_hess<10>(xp[i], -ln[i], e.W);
                }
                const double e0 = it[k + 1].y[PFX] - yv[FX];
                cost.v += c0 * e0 * e0;
                e.gf[FX] += -2.0 * c0 * e0; ev[k + 1].gf[PFX] += 2.0 * c0 * e0;
                e.W[FX][FX] += sf * 2.0 * c0; ev[k + 1].W[PFX][PFX] += sf * 2.0 * c0; e.Wx[0][FX] += -sf * 2.0 * c0;
                const J E = it[k + 1].y[PFY] - T;          // Fy_N - (Fy_f + Fy_r)(y_{N-1})
                cost = cost + c1 * (E * E);
                ev[k + 1].gf[PFY] += 2.0 * c1 * E.v;
                ev[k + 1].W[PFY][PFY] += sf * 2.0 * c1;
                for (int a = 0; a < 10; ++a) e.Wx[1][a] += sf * 2.0 * c1 * E.g[a];
            }
            e.f = cost.v;
            for (int a = 0; a < 10; ++a) e.gf[a] += cost.g[a];
            take_hess<10>(cost, sf, e.W);
            for (int j = 0; j < 10; ++j) {
                e.c[j] = ineq[j].v;
                for (int a = 0; a < 10; ++a) e.Jc[j][a] = ineq[j].g[a];
                take_hess<10>(ineq[j], -it[k].z[j], e.W);
            }
        } else {
            typedef Jet<7> J;
            J y[7], cost, ineq[2];
            for (int i = 0; i < 7; ++i) y[i] = J::var(yv[i], i);
            pm_stage<J>(P.p, P.c, y, P.ds[k], P.obst, P.n_obs, k == H - 1, cost, ineq);
            if (k == 0) for (int i = 0; i < 5; ++i) e.g[i] = yv[i] - P.x0[i];
            if (k < H - 1) {
                J xn[5];
                pm_step<J>(P.p, y, y + 5, P.kappa[k], P.ds[k], xn);
                const double* ln = it[k + 1].lam;
                for (int i = 0; i < 5; ++i) {
                    ev[k + 1].g[i] = it[k + 1].y[i] - xn[i].v;
                    for (int a = 0; a < 7; ++a) e.AB[i][a] = xn[i].g[a];
                    take_hess<7>(xn[i], -ln[i], e.W);
                }
                const double e0 = it[k + 1].y[PFX] - yv[PFX], e1 = it[k + 1].y[PFY] - yv[PFY];
                cost.v += c0 * e0 * e0 + c1 * e1 * e1;
                e.gf[PFX] += -2.0 * c0 * e0; ev[k + 1].gf[PFX] += 2.0 * c0 * e0;
                e.gf[PFY] += -2.0 * c1 * e1; ev[k + 1].gf[PFY] += 2.0 * c1 * e1;
                e.W[PFX][PFX] += sf * 2.0 * c0; ev[k + 1].W[PFX][PFX] += sf * 2.0 * c0; e.Wx[0][PFX] += -sf * 2.0 * c0;
                e.W[PFY][PFY] += sf * 2.0 * c1; ev[k + 1].W[PFY][PFY] += sf * 2.0 * c1; e.Wx[1][PFY] += -sf * 2.0 * c1;
            }
            e.f = cost.v;
            for (int a = 0; a < 7; ++a) e.gf[a] += cost.g[a];
            take_hess<7>(cost, sf, e.W);
            for (int j = 0; j < 2; ++j) {
                e.c[j] = ineq[j].v;
                for (int a = 0; a < 7; ++a) e.Jc[j][a] = ineq[j].g[a];
                take_hess<7>(ineq[j], -it[k].z[j], e.W);
            }
        }
        f += e.f;
    }
    // gradient of the Lagrangian: sf grad f + Jg^T lam - Jc^T z
    bool ok = std::isfinite(f);
    for (int k = 0; k < H; ++k) {
        StageEv& e = ev[k];
        const int nv = P.nv(k), nx = P.nx(k);
        for (int a = 0; a < nv; ++a) {
            double d = sf * e.gf[a];
            if (a < nx) d += it[k].lam[a];
            if (k < H - 1) for (int i = 0; i < P.nx(k + 1); ++i) d -= e.AB[i][a] * it[k + 1].lam[i];
            for (int j = 0; j < P.ni(k); ++j) d -= e.Jc[j][a] * it[k].z[j];
            e.dlag[a] = d;
            ok = ok && std::isfinite(d);
        }
        for (int i = 0; i < nx; ++i) ok = ok && std::isfinite(e.g[i]);
        for (int j = 0; j < P.ni(k); ++j) ok = ok && std::isfinite(e.c[j]);
    }
    return ok;
}

// ------------------------------------------------------------------------------------------------ KKT system
// unknowns per stage: z_k = [dlam_k (nx) ; dy_k (nv)];  K is block tridiagonal in this ordering:
//   D_k = [[0, (I 0)], [(I 0)^T, W_k + Jc^T Sigma Jc + delta I]],   E_k = [[0, -AB_{k-1}], [0, X_k]]  (block row k, col k-1)
struct Kkt {
    int H = 0;
    std::vector<Ldl> S;
    std::vector<std::vector<double>> F;   // F_k = S_{k-1}^{-1} E_k^T  (nb_{k-1} x nb_k), k >= 1
    std::vector<int> nb;
    int pos = 0, neg = 0, zero = 0;

    static void build_E(const Problem& P, const std::vector<StageEv>& ev, int k, double* E /* nb_k x nb_{k-1} */) {
        const int nxk = P.nx(k), nvk = P.nv(k), nxp = P.nx(k - 1), nvp = P.nv(k - 1), nbp = nxp + nvp;
        std::fill(E, E + (nxk + nvk) * nbp, 0.0);
        for (int i = 0; i < nxk; ++i) for (int a = 0; a < nvp; ++a) E[i * nbp + nxp + a] = -ev[k - 1].AB[i][a];
        for (int j = 0; j < 2; ++j) for (int a = 0; a < nvp; ++a) E[(nxk + nxk + j) * nbp + nxp + a] = ev[k - 1].Wx[j][a];
    }
    bool factor(const Problem& P, const std::vector<StageIt>& it, const std::vector<StageEv>& ev, double delta) {
        H = P.H; S.resize(H); F.resize(H); nb.resize(H);
        pos = neg = zero = 0;
        double D[NB * NB], E[NB * NB], Et[NB * NB];
        for (int k = 0; k < H; ++k) {
            const int nx = P.nx(k), nv = P.nv(k), n = nx + nv;
            nb[k] = n;
            std::fill(D, D + n * n, 0.0);
            for (int i = 0; i < nx; ++i) { D[i * n + nx + i] = 1.0; D[(nx + i) * n + i] = 1.0; }
            for (int a = 0; a < nv; ++a)
                for (int b = 0; b < nv; ++b) {
                    double w = ev[k].W[a][b];
                    for (int j = 0; j < P.ni(k); ++j) w += ev[k].Jc[j][a] * (it[k].z[j] / it[k].s[j]) * ev[k].Jc[j][b];
                    D[(nx + a) * n + nx + b] = w + (a == b ? delta : 0.0);
                }
            if (k > 0) {
                const int np = nb[k - 1];
                build_E(P, ev, k, E);
                F[k].assign((size_t)np * n, 0.0);
                // F_k = S_{k-1}^{-1} E_k^T, column by column
                double col[NB], sol[NB];
                for (int c = 0; c < n; ++c) {
                    for (int r = 0; r < np; ++r) col[r] = E[c * np + r];
                    S[k - 1].solve(col, sol);
                    for (int r = 0; r < np; ++r) F[k][(size_t)r * n + c] = sol[r];
                }
                for (int i = 0; i < n; ++i)
                    for (int j = 0; j < n; ++j) {
                        double a = 0.0;
                        for (int r = 0; r < np; ++r) a += E[i * np + r] * F[k][(size_t)r * n + j];
                        Et[i * n + j] = a;
                    }
                for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) D[i * n + j] -= 0.5 * (Et[i * n + j] + Et[j * n + i]);
            }
            S[k].factor(D, n, n);
            pos += S[k].pos; neg += S[k].neg; zero += S[k].zero;
        }
        int me = 0, nvar = 0;
        for (int k = 0; k < H; ++k) { me += P.nx(k); nvar += P.nv(k); }
        return zero == 0 && neg == me && pos == nvar;
    }
    // rhs / sol: per stage [nx + nv]
    void solve(std::vector<std::vector<double>>& r) const {
        double t[NB];
        for (int k = 1; k < H; ++k) {   // r'_k = r_k - F_k^T r'_{k-1}
            const int n = nb[k], np = nb[k - 1];
            for (int j = 0; j < n; ++j) {
                double a = 0.0;
                for (int q = 0; q < np; ++q) a += F[k][(size_t)q * n + j] * r[k - 1][q];
                r[k][j] -= a;
            }
        }
        for (int k = H - 1; k >= 0; --k) {
            S[k].solve(r[k].data(), t);
            if (k < H - 1) {
                const int n1 = nb[k + 1], n = nb[k];
                for (int q = 0; q < n; ++q) {
                    double a = 0.0;
                    for (int j = 0; j < n1; ++j) a += F[k + 1][(size_t)q * n1 + j] * r[k + 1][j];
                    t[q] -= a;
                }
            }
            for (int q = 0; q < nb[k]; ++q) r[k][q] = t[q];
        }
    }
};

// ------------------------------------------------------------------------------------------------ repair rollout
// restoration substitute (same as oracle/nlp.py NLP.repair): hold the speed (Fx = drag), no steering rate, PointMass Fy
// follows the centre-line curvature; states re-generated by rolling the models out from x0
static void repair(const Problem& P, std::vector<StageIt>& it) {
    const int N = P.N, H = P.H;
    if (N > 0) for (int i = 0; i < 8; ++i) it[0].y[i] = P.x0[i];
    else for (int i = 0; i < 5; ++i) it[0].y[i] = P.x0[i];
    for (int k = 0; k < N; ++k) {
        double* y = it[k].y;
        y[FX] = P.p.Frr + P.p.Cd * y[UX] * y[UX]; y[WW] = 0.0;
        if (k < N - 1) st_rk4<true, double>(P.p, y, y + 8, P.kappa[k], P.ds[k], it[k + 1].y);
    }
    if (P.M > 0) {
        if (N > 0) st_switch<double>(it[N - 1].y, it[N].y);
        for (int k = N; k < H; ++k) {
            double* y = it[k].y;
            y[PFX] = P.p.Frr + P.p.Cd * y[PV] * y[PV]; y[PFY] = P.p.m * y[PV] * y[PV] * P.kappa[k];
            if (k < H - 1) pm_step<double>(P.p, y, y + 5, P.kappa[k], P.ds[k], it[k + 1].y);
        }
    }
    if (P.c.obstacles && P.n_obs > 0)
        for (int k = 1; k < H; ++k) {
            double* y = it[k].y;
            const int si = k < N ? (int)SS : (int)PS, ei = k < N ? (int)EY : (int)PEY;
            for (int o = 0; o < P.n_obs; ++o) {
                const double R = P.obst[3 * o + 2] + 0.1 + 0.5, a = y[si] - P.obst[3 * o], e = y[ei] - P.obst[3 * o + 1];
                if (a * a + e * e < R * R) y[ei] = P.obst[3 * o + 1] + (e > 1e-9 ? 1.0 : -1.0) * std::sqrt(R * R - a * a);   // |e| <= 1e-9: dead centre
            }
        }
}

// ------------------------------------------------------------------------------------------------ interior point
struct Result { int status, iters; double obj, err, mu, sf; };

static Result solve_once(const Problem& P, std::vector<StageIt>& it, const Options& opt, double sf_fixed) {
    const int H = P.H;
    const double tol = P.c.tol, acc_tol = P.c.acceptable_tol;
    std::vector<StageEv> ev(H), evt(H);
    std::vector<StageIt> tr(H);
    double f;
    for (int k = 0; k < H; ++k) {
        for (int i = 0; i < 8; ++i) it[k].lam[i] = 0.0;
        for (int j = 0; j < 10; ++j) { it[k].s[j] = 1.0; it[k].z[j] = 0.0; }
    }
    double sf = 1.0;
    bool ok = eval_all(P, it, 1.0, f, ev);
    if (!ok) return Result{ST_EVAL_ERROR, 0, f, INFINITY, P.c.mu_init, sf};
    if (sf_fixed > 0) sf = sf_fixed;
    else if (P.c.obj_scaling) {
        double g0 = 0.0;
        for (int k = 0; k < H; ++k) for (int a = 0; a < P.nv(k); ++a) g0 = std::fmax(g0, std::fabs(ev[k].gf[a]));
        if (g0 > MAX_GRADIENT) sf = MAX_GRADIENT / g0;
    }
    int me = 0, mi = 0;
    for (int k = 0; k < H; ++k) { me += P.nx(k); mi += P.ni(k); }
    double mu = P.c.mu_init;
    for (int k = 0; k < H; ++k)
        for (int j = 0; j < P.ni(k); ++j) {
            it[k].s[j] = std::fmax(ev[k].c[j], SLACK_PUSH * std::fmax(1.0, std::fabs(ev[k].c[j])));
            it[k].z[j] = mu / it[k].s[j];
        }
    std::vector<std::pair<double, double>> filt;
    double delta_last = 0.0, theta0 = -1.0, err0 = INFINITY;
    int acc_count = 0, status = ST_MAX_ITER, iter = 0;
    Kkt kkt;
    std::vector<std::vector<double>> rhs(H), dir(H), gsoc(H), csoc(H);
    std::vector<StageIt> d1(H), d2(H);   // directions: y, lam, s, z fields reused as dy, dlam, ds, dz
    auto errors = [&](double mu_) {
        double rd = 0.0, prim = 0.0, cm = 0.0, sl = 0.0, sz = 0.0;
        for (int k = 0; k < H; ++k) {
            for (int a = 0; a < P.nv(k); ++a) rd = std::fmax(rd, std::fabs(ev[k].dlag[a]));
            for (int i = 0; i < P.nx(k); ++i) { prim = std::fmax(prim, std::fabs(ev[k].g[i])); sl += std::fabs(it[k].lam[i]); }
            for (int j = 0; j < P.ni(k); ++j) {
                prim = std::fmax(prim, std::fabs(ev[k].c[j] - it[k].s[j]));
                sz += std::fabs(it[k].z[j]);
                cm = std::fmax(cm, std::fabs(it[k].s[j] * it[k].z[j] - mu_));
            }
        }
        const double sd = std::fmax(S_MAX, (sl + sz) / std::max(1, me + mi)) / S_MAX;
        const double sc = std::fmax(S_MAX, sz / std::max(1, mi)) / S_MAX;
        return std::fmax(std::fmax(rd / sd, prim), cm / sc);
    };
    // direction for given equality / (c - s) residuals (nullptr: the current ones); returns a_p, a_d, fills d
    auto direction = [&](const std::vector<std::vector<double>>* gs, const std::vector<std::vector<double>>* cs, std::vector<StageIt>& d,
                         double tau, double& a_p, double& a_d) {
        for (int k = 0; k < H; ++k) {
            const int nx = P.nx(k), nv = P.nv(k), ni = P.ni(k);
            rhs[k].assign(nx + nv, 0.0);
            for (int i = 0; i < nx; ++i) rhs[k][i] = -(gs ? (*gs)[k][i] : ev[k].g[i]);
            for (int a = 0; a < nv; ++a) {
                // -(sf gf + Jg^T lam - Jc^T (mu/s - sigma rc)) = -(dlag + Jc^T z - Jc^T (mu/s - sigma rc))
                double v = ev[k].dlag[a];
                for (int j = 0; j < ni; ++j) {
                    const double s = it[k].s[j], z = it[k].z[j], rc = cs ? (*cs)[k][j] : ev[k].c[j] - s;
                    v += ev[k].Jc[j][a] * (z - (mu / s - (z / s) * rc));
                }
                rhs[k][nx + a] = -v;
            }
        }
        kkt.solve(rhs);
        a_p = 1.0; a_d = 1.0;
        for (int k = 0; k < H; ++k) {
            const int nx = P.nx(k), nv = P.nv(k), ni = P.ni(k);
            for (int i = 0; i < nx; ++i) d[k].lam[i] = rhs[k][i];
            for (int a = 0; a < nv; ++a) d[k].y[a] = rhs[k][nx + a];
            for (int j = 0; j < ni; ++j) {
                const double s = it[k].s[j], z = it[k].z[j], rc = cs ? (*cs)[k][j] : ev[k].c[j] - s;
                double dsl = rc;
                for (int a = 0; a < nv; ++a) dsl += ev[k].Jc[j][a] * d[k].y[a];
                const double dz = mu / s - z - (z / s) * dsl;
                d[k].s[j] = dsl; d[k].z[j] = dz;
                if (dsl < 0) a_p = std::fmin(a_p, -tau * s / dsl);
                if (dz < 0) a_d = std::fmin(a_d, -tau * z / dz);
            }
        }
    };
    auto make_trial = [&](const std::vector<StageIt>& d, double alpha) {
        for (int k = 0; k < H; ++k) {
            for (int a = 0; a < P.nv(k); ++a) tr[k].y[a] = it[k].y[a] + alpha * d[k].y[a];
            for (int j = 0; j < P.ni(k); ++j) tr[k].s[j] = it[k].s[j] + alpha * d[k].s[j];
        }
    };
    auto measures = [&](const std::vector<StageIt>& pt, const std::vector<StageEv>& e, double fv, double& theta, double& phi) {
        theta = 0.0; double lns = 0.0;
        for (int k = 0; k < H; ++k) {
            for (int i = 0; i < P.nx(k); ++i) theta += std::fabs(e[k].g[i]);
            for (int j = 0; j < P.ni(k); ++j) { theta += std::fabs(e[k].c[j] - pt[k].s[j]); lns += std::log(pt[k].s[j]); }
        }
        phi = sf * fv - mu * lns;
    };

    // IPOPT's watchdog (IpBacktrackingLineSearch: StartWatchDog / StopWatchDog): after `watchdog` consecutive shortened
    // steps the full step is taken for up to 3 iterations without asking the filter, the acceptance test being made
    // against the stored reference point; if no acceptable point turns up, back to the stored iterate and backtrack
    bool in_wd = false, skip_first = false;
    int wd_short = 0, wd_trial = 0;
    std::vector<StageIt> it_wd;
    double theta_wd = 0, phi_wd = 0, dphi_wd = 0, atest_wd = 0;
    for (iter = 0; iter <= opt.max_iter; ++iter) {
        ok = eval_all(P, it, sf, f, ev);
        if (!ok) { status = ST_EVAL_ERROR; break; }
        err0 = errors(0.0);
        if (opt.verbose) {
            double th, ph; measures(it, ev, f, th, ph);
            std::fprintf(stderr, "it %3d f %.10f E0 %.3e mu %.3e theta %.3e\n", iter, f, err0, mu, th);
        }
        if (err0 <= tol) { status = ST_SOLVED; break; }
        acc_count = err0 <= acc_tol ? acc_count + 1 : 0;
        if (acc_count >= P.c.acceptable_iter) { status = ST_ACCEPTABLE; break; }
        if (iter == opt.max_iter) break;
        bool changed = false;
        while (mu > tol / 11.0 * (1 + 1e-12) && errors(mu) <= KAPPA_EPS * mu) {
            mu = std::fmax(tol / 11.0, std::fmin(KAPPA_MU * mu, std::pow(mu, THETA_MU)));
            changed = true;
        }
        if (changed) filt.clear();
        const double tau = std::fmax(TAU_MIN, 1.0 - mu);
        double theta, phi;
        measures(it, ev, f, theta, phi);
        if (theta0 < 0) theta0 = theta;
        const double theta_min = 1e-4 * std::fmax(1.0, theta0), theta_max = 1e4 * std::fmax(1.0, theta0);
        const bool first_ever = delta_last == 0.0;
        double trial = 0.0, alpha = 0.0, a_d = 1.0;
        bool accepted = false, failed = false;
        const std::vector<StageIt>* used = &d1;
        auto acceptable_ref = [&](double a_test, double tht, double pht, double dphi, bool& augment, double theta, double phi) {
            augment = false;
            if (tht > theta_max) return false;
            for (auto& fe : filt) if (tht >= fe.first && pht >= fe.second) return false;
            const bool switching = theta <= theta_min && dphi < 0 && a_test * std::pow(-dphi, S_PHI) > LS_DELTA * std::pow(theta, S_THETA);
            if (switching) { if (pht <= phi + ETA_PHI * a_test * dphi) return true; }
            else if (tht <= (1 - GAMMA_THETA) * theta || pht <= phi - GAMMA_PHI * theta) { augment = true; return true; }
            return std::fabs(pht - phi) <= 1e-13 * std::fmax(1.0, std::fabs(phi)) && tht <= theta + 1e-13;
        };
        auto acceptable = [&](double a_test, double tht, double pht, double dphi, bool& augment) {
            return acceptable_ref(a_test, tht, pht, dphi, augment, theta, phi);
        };
        bool wd_restored = false;
        int n_steps = 0;
        while (!accepted && !failed) {
            while (!kkt.factor(P, it, ev, trial)) {   // delta_w ladder
                if (trial == 0.0) trial = first_ever ? 1e-4 : std::fmax(1e-20, delta_last / 3.0);
                else trial *= first_ever ? 100.0 : 8.0;
                if (trial > 1e40) { failed = true; break; }
            }
            if (failed) break;
            double a_p;
            direction(nullptr, nullptr, d1, tau, a_p, a_d);
            double dphi = 0.0;
            for (int k = 0; k < H; ++k) {
                for (int a = 0; a < P.nv(k); ++a) dphi += sf * ev[k].gf[a] * d1[k].y[a];
                for (int j = 0; j < P.ni(k); ++j) dphi -= mu * d1[k].s[j] / it[k].s[j];
            }
            alpha = a_p;
            if (!in_wd && opt.watchdog > 0 && wd_short >= opt.watchdog && !skip_first && trial == 0.0) {
                in_wd = true; it_wd = it; theta_wd = theta; phi_wd = phi; dphi_wd = dphi; atest_wd = a_p; wd_trial = 0;
            }
            if (in_wd) {
                make_trial(d1, a_p);
                double ft, tht = 0, pht = 0;
                const bool okv = eval_values(P, tr, ft, evt);
                bool augment = false, acc = false;
                if (okv) { measures(tr, evt, ft, tht, pht); acc = acceptable_ref(atest_wd, tht, pht, dphi_wd, augment, theta_wd, phi_wd); }
                if (acc) {
                    in_wd = false; accepted = true; used = &d1; n_steps = 0;
                    if (augment) filt.emplace_back((1 - GAMMA_THETA) * theta_wd, phi_wd - GAMMA_PHI * theta_wd);
                } else if (!okv || ++wd_trial > 3) {
                    it = it_wd; in_wd = false; wd_short = 0; skip_first = true; wd_restored = true;
                } else { accepted = true; used = &d1; n_steps = 0; }   // full step taken on trust
                break;
            }
            if (skip_first) alpha *= 0.5;
            for (int bt = skip_first ? 1 : 0; bt < MAX_BACKTRACK && !accepted; ++bt, alpha *= 0.5) {
                n_steps = bt;
                make_trial(d1, alpha);
                double ft, tht = 0, pht = 0;
                const bool okv = eval_values(P, tr, ft, evt);
                bool augment = false;
                if (okv) {
                    measures(tr, evt, ft, tht, pht);
                    if (acceptable(alpha, tht, pht, dphi, augment)) { accepted = true; used = &d1; }
                }
                if (opt.verbose > 1)
                    std::fprintf(stderr, "        bt %d alpha %.3e ok %d f %.8f theta %.4e phi %.8e (ref theta %.4e phi %.8e dphi %.3e) acc %d\n", bt, alpha,
                                 (int)okv, ft, tht, pht, theta, phi, dphi, (int)accepted);
                // second-order correction (IPOPT Alg. A-5.5..5.9): first trial point rejected, infeasibility not reduced
                if (!accepted && bt == 0 && okv && tht >= theta && opt.max_soc > 0) {
                    for (int k = 0; k < H; ++k) {
                        gsoc[k].assign(8, 0.0); csoc[k].assign(10, 0.0);
                        for (int i = 0; i < P.nx(k); ++i) gsoc[k][i] = alpha * ev[k].g[i] + evt[k].g[i];
                        for (int j = 0; j < P.ni(k); ++j) csoc[k][j] = alpha * (ev[k].c[j] - it[k].s[j]) + (evt[k].c[j] - tr[k].s[j]);
                    }
                    double theta_soc_old = tht;   // IPOPT compares every correction with the previous TRIAL point's violation
                    for (int ps = 0; ps < opt.max_soc; ++ps) {
                        double a_p2, a_d2, ft2, tht2, pht2;
                        direction(&gsoc, &csoc, d2, tau, a_p2, a_d2);
                        make_trial(d2, a_p2);
                        const bool ok2 = eval_values(P, tr, ft2, evt);
                        if (!ok2) break;
                        measures(tr, evt, ft2, tht2, pht2);
                        if (acceptable(alpha, tht2, pht2, dphi, augment)) {
                            accepted = true; used = &d2; alpha = a_p2; a_d = a_d2; n_steps = 0;
                            break;
                        }
                        if (tht2 > KAPPA_SOC * theta_soc_old) break;
                        theta_soc_old = tht2;
                        for (int k = 0; k < H; ++k) {
                            for (int i = 0; i < P.nx(k); ++i) gsoc[k][i] = a_p2 * gsoc[k][i] + evt[k].g[i];
                            for (int j = 0; j < P.ni(k); ++j) csoc[k][j] = a_p2 * csoc[k][j] + (evt[k].c[j] - tr[k].s[j]);
                        }
                    }
                }
                if (accepted) {
                    if (augment) filt.emplace_back((1 - GAMMA_THETA) * theta, phi - GAMMA_PHI * theta);
                    break;
                }
            }
            if (accepted) break;
            // IPOPT would enter restoration here: damp the step instead (larger delta_w) and retry
            trial = std::fmax(trial, 1e-4) * 100.0;
            if (trial > 1e12) failed = true;
        }
        if (wd_restored) { --iter; continue; }   // back at the stored iterate: the regular line search follows
        skip_first = false;
        if (failed || !accepted) { status = ST_LINESEARCH_FAIL; break; }
        if (trial > 0) delta_last = trial;
        wd_short = n_steps == 0 ? 0 : wd_short + 1;
        if (opt.verbose) std::fprintf(stderr, "      alpha %.3e a_d %.3e delta_w %.2e soc %d\n", alpha, a_d, trial, used == &d2);
        const std::vector<StageIt>& d = *used;
        for (int k = 0; k < H; ++k) {
            for (int a = 0; a < P.nv(k); ++a) it[k].y[a] += alpha * d[k].y[a];
            for (int i = 0; i < P.nx(k); ++i) it[k].lam[i] += alpha * d[k].lam[i];
            for (int j = 0; j < P.ni(k); ++j) {
                it[k].s[j] += alpha * d[k].s[j];
                double z = it[k].z[j] + a_d * d[k].z[j];
                z = std::fmin(std::fmax(z, mu / (KAPPA_SIGMA * it[k].s[j])), KAPPA_SIGMA * mu / it[k].s[j]);
                it[k].z[j] = z;
            }
        }
    }
    if (iter > opt.max_iter) iter = opt.max_iter;
    double fv;
    eval_values(P, it, fv, evt);
    return Result{status, iter, fv, err0, mu, sf};
}

static Result solve(const Problem& P, std::vector<StageIt>& it, const Options& opt) {
    Result r = solve_once(P, it, opt, opt.obj_scale);
    int restarts = 0;
    while (r.status == ST_EVAL_ERROR && r.iters == 0 && restarts < opt.max_restorations) {
        ++restarts;   // invalid starting point: the repaired point sets the objective scaling
        repair(P, it);
        r = solve_once(P, it, opt, opt.obj_scale);
    }
    while (r.status == ST_LINESEARCH_FAIL && restarts < opt.max_restorations) {
        ++restarts;
        repair(P, it);
        const int before = r.iters + 1;
        r = solve_once(P, it, opt, r.sf);
        r.iters += before;
    }
    return r;
}

}  // namespace vco

// ------------------------------------------------------------------------------------------------ C interface
extern "C" {

typedef struct vco_options { int max_iter, max_soc, max_restorations, verbose; double obj_scale; int watchdog; } vco_options;

int vco_version(void) { return 2; }
int vco_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// Same array conventions as vc_nmpc_solve (include/vc_api.h), host pointers: x0[B,8], kappa[B,H], ds[B,H],
// obst[B,n_obs,3] or NULL, X[B,8,H] / U[B,2,H] warm start in, solution out; obj/status/iters/kkt[B].
int vco_nmpc_solve(const vc_model_params* par, const vc_mpc_config* cfg, const vco_options* o, int B, const double* x0,
                   const double* kappa, const double* ds, const double* obst, int n_obs, double* X, double* U, double* obj,
                   int* status, int* iters, double* kkt, int n_threads) {
    using namespace vco;
    const int N = cfg->N, M = cfg->M, H = N + M;
    if (H < 1 || H > 4096) return -1;
    Options opt{o ? o->max_iter : 3000, o ? o->max_soc : 4, o ? o->max_restorations : 2, o ? o->verbose : 0, o ? o->obj_scale : 0.0,
                o ? o->watchdog : 10};
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#pragma omp parallel for schedule(dynamic, 1)
#endif
    for (int b = 0; b < B; ++b) {
        Problem P;
        P.p = *par; P.c = *cfg; P.N = N; P.M = M; P.H = H;
        P.n_obs = (cfg->obstacles && obst) ? n_obs : 0;
        P.x0 = x0 + (size_t)b * 8; P.kappa = kappa + (size_t)b * H; P.ds = ds + (size_t)b * H;
        P.obst = P.n_obs ? obst + (size_t)b * n_obs * 3 : nullptr;
        std::vector<StageIt> it(H);
        double* Xb = X + (size_t)b * 8 * H; double* Ub = U + (size_t)b * 2 * H;
        for (int k = 0; k < H; ++k) {
            const int nx = P.nx(k);
            std::memset(&it[k], 0, sizeof(StageIt));
            for (int i = 0; i < nx; ++i) it[k].y[i] = Xb[i * H + k];
            it[k].y[nx] = Ub[k]; it[k].y[nx + 1] = Ub[H + k];
        }
        const Result r = solve(P, it, opt);
        for (int k = 0; k < H; ++k) {
            const int nx = P.nx(k);
            for (int i = 0; i < nx; ++i) Xb[i * H + k] = it[k].y[i];
            Ub[k] = it[k].y[nx]; Ub[H + k] = it[k].y[nx + 1];
        }
        if (obj) obj[b] = r.obj;
        if (status) status[b] = r.status;
        if (iters) iters[b] = r.iters;
        if (kkt) kkt[b] = r.err;
    }
    return 0;
}

// objective / max equality violation / min inequality margin of given trajectories (mirror of vc_nmpc_eval)
int vco_nmpc_eval(const vc_model_params* par, const vc_mpc_config* cfg, int B, const double* x0, const double* kappa,
                  const double* ds, const double* obst, int n_obs, const double* X, const double* U, double* out) {
    using namespace vco;
    const int N = cfg->N, M = cfg->M, H = N + M;
    for (int b = 0; b < B; ++b) {
        Problem P;
        P.p = *par; P.c = *cfg; P.N = N; P.M = M; P.H = H;
        P.n_obs = (cfg->obstacles && obst) ? n_obs : 0;
        P.x0 = x0 + (size_t)b * 8; P.kappa = kappa + (size_t)b * H; P.ds = ds + (size_t)b * H;
        P.obst = P.n_obs ? obst + (size_t)b * n_obs * 3 : nullptr;
        std::vector<StageIt> it(H);
        std::vector<StageEv> ev(H);
        const double* Xb = X + (size_t)b * 8 * H; const double* Ub = U + (size_t)b * 2 * H;
        for (int k = 0; k < H; ++k) {
            const int nx = P.nx(k);
            for (int i = 0; i < nx; ++i) it[k].y[i] = Xb[i * H + k];
            it[k].y[nx] = Ub[k]; it[k].y[nx + 1] = Ub[H + k];
        }
        double f, gmax = 0.0, cmin = INFINITY;
        eval_values(P, it, f, ev);
        for (int k = 0; k < H; ++k) {
            for (int i = 0; i < P.nx(k); ++i) gmax = std::fmax(gmax, std::fabs(ev[k].g[i]));
            for (int j = 0; j < P.ni(k); ++j) cmin = std::fmin(cmin, ev[k].c[j]);
        }
        out[3 * b] = f; out[3 * b + 1] = gmax; out[3 * b + 2] = cmin;
    }
    return 0;
}

// SingleTrack spatial RK4 rollout + [A|B] Jacobians (mirror of vc_rollout_st): X[B,N+1,8], AB[B,N,8,10] or NULL
int vco_rollout_st(const vc_model_params* par, int B, int N, const double* x0, const double* U, const double* kappa,
                   const double* ds, double* X, double* AB, int n_threads) {
    using namespace vco;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#pragma omp parallel for schedule(static)
#endif
    for (int b = 0; b < B; ++b) {
        double* Xb = X + (size_t)b * (N + 1) * 8;
        for (int i = 0; i < 8; ++i) Xb[i] = x0[(size_t)b * 8 + i];
        for (int n = 0; n < N; ++n) {
            const double* u = U + ((size_t)b * N + n) * 2;
            const double k = kappa[(size_t)b * N + n], h = ds[(size_t)b * N + n];
            if (AB) {
                typedef Jet<10, false> J;   // first-order jets: value + the 10 columns of [A|B]
                J y[10], xn[8];
                for (int i = 0; i < 8; ++i) y[i] = J::var(Xb[n * 8 + i], i);
                y[8] = J::var(u[0], 8); y[9] = J::var(u[1], 9);
                st_rk4<true, J>(*par, y, y + 8, k, h, xn);
                double* ab = AB + ((size_t)b * N + n) * 80;
                for (int i = 0; i < 8; ++i) { Xb[(n + 1) * 8 + i] = xn[i].v; for (int a = 0; a < 10; ++a) ab[i * 10 + a] = xn[i].g[a]; }
            } else {
                st_rk4<true, double>(*par, Xb + n * 8, u, k, h, Xb + (n + 1) * 8);
            }
        }
    }
    return 0;
}

// temporal RK4 plant step (mirror of vc_plant_step with kappa given): x[B,8] in place
int vco_plant_step(const vc_model_params* par, int B, double* x, const double* u, const double* kappa, double dt) {
    using namespace vco;
    for (int b = 0; b < B; ++b) {
        double xn[8];
        st_rk4<false, double>(*par, x + (size_t)b * 8, u + (size_t)b * 2, kappa[b], dt, xn);
        for (int i = 0; i < 8; ++i) x[(size_t)b * 8 + i] = xn[i];
    }
    return 0;
}

}  // extern "C"
