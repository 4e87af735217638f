// Batched NMPC solve for sm_100a: primal-dual interior point (IPOPT's published algorithm: monotone barrier,
// filter line search, inertia correction) with a per-instance Riccati factorisation of the KKT system.
//
// Replaces CascadedMPC.command -> opti.solve() (vehicle_control/controllers/mpc/cascaded_mpc.py:306-314), i.e.
// CasADi's AD (exact Lagrangian Hessian, "expand": True, :65) + IPOPT (:54-66) + the HSL MA27 LDL^T.
//
// Kernels per interior-point iteration (host-driven loop, all unfinished instances in lock step; finished ones cost nothing):
//   k_eval_st              eight threads per (instance, SingleTrack stage): forward in Dual1 arithmetic + the hand-written
//                          adjoint sweep of vc_adj.cuh -> Jacobian columns, gradient, exact Hessian column (forward over reverse)
//   k_eval_pm              one thread per (instance, PointMass stage, derivative pair): hyper-dual evaluation
//   k_step_a               one warp per instance: error / barrier update (warp-shuffle reductions), block tridiagonal KKT
//                          solve by a Riccati recursion staged in shared memory (delta_w inertia ladder), directions
//   k_step_b               one warp per instance: fraction to the boundary, filter line search with lanes over stages,
//                          second-order correction, watchdog (k_step_*_wide: one CTA per instance for the late passes)
//   k_update               one warp per (instance, stage): the accepted primal-dual step
// IPOPT's restoration phase is not restated: where it would start, the step is damped, then the iterate is re-initialised
// by a dynamically feasible rollout, then the instance ends with VC_LINESEARCH_FAIL.
// All fp64; per-stage matrices are <= 11x11, so there is nothing for tensor cores here.
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <mutex>
#include <vector>
#include "vc_handle.h"
#include "vc_adj.cuh"

namespace vc {

#define FULL_MASK 0xffffffffu

// ------------------------------------------------------------------------------------------------
// per-instance workspace (global memory): one contiguous slab per instance made of H stage RECORDS of
// WS_PER_STAGE doubles each, followed by the scalar block.  Inside a record the arrays the Riccati sweep streams
// come first and are contiguous (backward sweep: [0, WS_BWD); forward sweep: AB and [WS_FWD0, WS_FWD1)), so that a
// warp stages the next stage's block in shared memory with 16-byte cp.async copies while it works on the current one.
enum WsArray {
    W_HC = 0,    // [36] condensed Hessian block, packed lower triangle over the ACTIVE stage variables
    W_AB,        // [80] d(next state)/d(active vars): [i][a] stride 8; switch stage: row 5 = grad T, [48] = T
    W_RD,        // [10] dual infeasibility  (eval: active entries; step: s, t, slew, obstacle parts)
    W_V1,        // [10] Jc^T (1/s)
    W_V2,        // [10] Jc^T (z c / s)
    W_OBS,       // [8]  obstacle term: value, d/ds, d/dey, ss, s-ey, ey-ey; [6],[7] = slew weights tying u_k to u_{k-1}
    W_AUX,       // [10] Riccati pre-pass: direction of s at this stage; cost-to-go gradient of t of the NEXT stage;
                 //      [2..9] second-order correction: Jc^T sigma (c_soc - (c - s)) per active variable (soc_residual)
    W_GRES,      // [8]  equality residual of the row block that defines x_k
    W_KM,        // [24] feedback gains K[2][10], kff[2]
    W_PV,        // [10] Riccati cost-to-go gradient
    W_PM,        // [100] Riccati cost-to-go Hessian
    W_Y,         // [10] stage variables (x,u); PointMass stages use 7
    W_LAM,       // [8]  multiplier of the equality that defines x_k (k=0: initial state)
    W_S,         // [10] slacks of the stage inequalities
    W_Z,         // [10] their multipliers
    W_JC,        // [80] inequality Jacobian [j][a] stride 8
    W_GF,        // [10] scaled objective gradient
    W_CRES,      // [10] inequality values c(y_k)
    W_FST,       // [1]  stage cost (unscaled; AD part only)
    W_PAD,       // [1]  keeps every record 16-byte aligned
    W_DY,        // [10] primal direction
    W_DLAM,      // [8]
    W_DSL,       // [10]
    W_DZ,        // [10]
    W_GT,        // [8]   equality residual at the last trial point
    W_CT,        // [10]  inequality values at the last trial point
    W_GS,        // [8]   second-order-correction residuals (equalities)
    W_CS,        // [10]  second-order-correction residuals (c - s rows)
    W_DY2,       // [10]  corrected direction
    W_DLAM2,     // [8]
    W_DSL2,      // [10]
    W_DZ2,       // [10]
    W_BAK,       // [38]  watchdog: copy of the iterate (y lam s z) at the stored reference point
    W_COUNT
};
struct WsTable { int width[W_COUNT + 1]; int off[W_COUNT + 1]; };
constexpr WsTable make_ws_table() {
    WsTable t{};
    const int w[W_COUNT] = {36, 80, 10, 10, 10, 8, 10, 8, 24, 10, 100, 10, 8, 10, 10, 80, 10, 10, 1, 1,
                            10, 8, 10, 10, 8, 10, 8, 10, 10, 8, 10, 10, 38};
    int o = 0;
    for (int i = 0; i < W_COUNT; ++i) { t.width[i] = w[i]; t.off[i] = o; o += w[i]; }
    t.width[W_COUNT] = 0; t.off[W_COUNT] = o;
    return t;
}
constexpr WsTable WS_TABLE = make_ws_table();
__device__ __constant__ int c_ws_off[W_COUNT + 1] = {
    WS_TABLE.off[0], WS_TABLE.off[1], WS_TABLE.off[2], WS_TABLE.off[3], WS_TABLE.off[4], WS_TABLE.off[5], WS_TABLE.off[6],
    WS_TABLE.off[7], WS_TABLE.off[8], WS_TABLE.off[9], WS_TABLE.off[10], WS_TABLE.off[11], WS_TABLE.off[12],
    WS_TABLE.off[13], WS_TABLE.off[14], WS_TABLE.off[15], WS_TABLE.off[16], WS_TABLE.off[17], WS_TABLE.off[18],
    WS_TABLE.off[19], WS_TABLE.off[20], WS_TABLE.off[21], WS_TABLE.off[22], WS_TABLE.off[23], WS_TABLE.off[24],
    WS_TABLE.off[25], WS_TABLE.off[26], WS_TABLE.off[27], WS_TABLE.off[28], WS_TABLE.off[29], WS_TABLE.off[30],
    WS_TABLE.off[31], WS_TABLE.off[32], WS_TABLE.off[33]};
constexpr int ws_width(int a) { return WS_TABLE.width[a]; }
constexpr int ws_offset(int a) { return WS_TABLE.off[a]; }   // offset inside a stage record
constexpr int WS_PER_STAGE = ws_offset(W_COUNT);
constexpr int WS_BWD = ws_offset(W_KM);       // [0, WS_BWD): what the backward sweep reads of a stage
constexpr int WS_FWD0 = ws_offset(W_GRES);    // forward sweep: AB block and [WS_FWD0, WS_FWD1)
constexpr int WS_FWD1 = ws_offset(W_Y);
static_assert(WS_PER_STAGE % 2 == 0 && WS_BWD % 2 == 0 && WS_FWD0 % 2 == 0 && WS_FWD1 % 2 == 0 &&
              ws_offset(W_AB) % 2 == 0 && ws_offset(W_RD) % 2 == 0, "16-byte alignment of the staged blocks");
static_assert(W_COUNT == 33, "update make_ws_table / c_ws_off");
constexpr int O_IT = ws_offset(W_Y), O_DIR = ws_offset(W_DY), O_DIR2 = ws_offset(W_DY2), O_BAK = ws_offset(W_BAK);
constexpr int O_GS = ws_offset(W_GS), O_GRES_C = ws_offset(W_GRES), O_JC = ws_offset(W_JC), O_CS = ws_offset(W_CS), O_CRES = ws_offset(W_CRES),
              O_S = ws_offset(W_S), O_Z = ws_offset(W_Z), O_KM = ws_offset(W_KM), O_PM = ws_offset(W_PM), O_PV = ws_offset(W_PV);
static_assert(ws_offset(W_LAM) == O_IT + 10 && ws_offset(W_S) == O_IT + 18 && ws_offset(W_Z) == O_IT + 28 &&
              ws_offset(W_DLAM) == O_DIR + 10 && ws_offset(W_DSL) == O_DIR + 18 && ws_offset(W_DZ) == O_DIR + 28 &&
              ws_offset(W_DLAM2) == O_DIR2 + 10 && ws_offset(W_DSL2) == O_DIR2 + 18 && ws_offset(W_DZ2) == O_DIR2 + 28,
              "iterate / direction blocks must share one layout");
constexpr int NSCAL = 112;
#ifdef VC_TIMING
#define VC_TICK(slot) do { if (rank_for_tick == 0) ws.scal(b)[84 + (slot)] = (double)clock64(); } while (0)
#else
#define VC_TICK(slot) do { } while (0)
#endif
// scalar block of an instance
enum Scal { S_MU = 0, S_SF, S_DELTA_LAST, S_THETA0, S_NFILT, S_ACC, S_MODE, S_RESTARTS, S_OBJ, S_ERR, S_ITERS,
            S_STATUS, S_DBG_ALPHA, S_DBG_AP, S_DBG_THETA, S_DBG_DPHI, S_FILT = 16, S_NREST = 48, S_DBG_NAN = 50, S_DBG_THT0, S_DBG_PHT0, S_DBG_PHI, S_DBG_SOC, S_DBG_THT2, S_DBG_AP2, S_SHORT /* theta[16] */, S_FILP = 32 /* phi[16] */ };
constexpr int FILT_MAX = 16;
// hand-over block between the two halves of an iteration (k_step_a -> k_step_b)
enum ScalAB { S_PHASE = 64, S_THETA, S_PHI, S_TRIAL, S_AP, S_AD, S_DPHI, S_MU_NEW, S_NFILT_NEW, S_THETA0_NEW, S_ACC_NEW,
              S_NSHORT_NEW, S_OBJ_NEW, S_ERR_NEW, S_DAMP_TRIAL, S_NDAMP, S_UPD_DIR, S_UPD_ALPHA, S_UPD_AD };
// watchdog state (IPOPT IpBacktrackingLineSearch: StartWatchDog / StopWatchDog)
enum ScalWd { S_WD_IN = 96, S_WD_SHORT, S_WD_TRIAL, S_WD_THETA, S_WD_PHI, S_WD_DPHI, S_WD_ATEST, S_SKIP_FIRST };
enum Phase { PH_NONE = 0, PH_SEARCH = 1, PH_FAILED = 3, PH_BADINIT = 4 };
enum Mode { MODE_INIT = 0, MODE_STEP = 1, MODE_DONE = 2, MODE_REINIT = 3 };

struct Ws {
    double* base;
    int H;
    size_t stride;   // doubles per instance
    __host__ __device__ double* rec(int b, int k) const { return base + (size_t)b * stride + (size_t)k * WS_PER_STAGE; }
    __host__ __device__ double* at(int b, int a, int k) const {
#if defined(__CUDA_ARCH__)
        return rec(b, k) + c_ws_off[a];
#else
        return rec(b, k) + ws_offset(a);
#endif
    }
    __host__ __device__ double* scal(int b) const { return base + (size_t)b * stride + (size_t)WS_PER_STAGE * H; }
};

struct Prob {   // device-side view of one vc_nmpc_solve call
    Params p;
    Cfg c;
    int B, N, M, H, n_obs;
    int watchdog_trigger, watchdog_trials;   // IPOPT watchdog_shortened_iter_trigger (10) / watchdog_trial_iter_max (3)
    double mu_start;   // barrier parameter of the first iteration (mu_init advanced by `warm_rungs` rungs of the ladder)
    int max_soc;       // second-order correction steps per iteration (IPOPT max_soc = 4)
    int max_damp;      // damped re-solves of one iteration's direction before the iterate is restored / given up
    const double *x0, *kappa, *ds, *obst;
    double *X, *U, *obj, *kkt;
    int *status, *iters;
    // compacted list of unfinished instances of this pass (written by the previous k_step) and its length
    const int* act_list;
    const int* act_count;
};

// active (differentiated) stage variables
__device__ __constant__ int c_act_st[8] = {IUX, IUY, IR, IDELTA, IEY, IEPSI, YFX, YW};
__device__ __constant__ int c_act_pm[5] = {QV, QEY, QEPSI, QFX, QFY};

__device__ __forceinline__ void pair_from_index(int q, int& a, int& b) {   // q -> (a >= b), packed lower triangle
    a = 0;
    while ((a + 1) * (a + 2) / 2 <= q) ++a;
    b = q - a * (a + 1) / 2;
}

// ------------------------------------------------------------------------------------------------
// k_prepare: user layout X[B,8,H], U[B,2,H] (state_prediction / action_prediction) -> stage-major workspace
__global__ void k_prepare(Prob pr, Ws ws, int* list0, int* count0) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= pr.B * pr.H) return;
    const int b = t / pr.H, k = t % pr.H, H = pr.H;
    if (k == 0) list0[b] = b;
    if (t == 0) *count0 = pr.B;
    const double* Xb = pr.X + (size_t)b * 8 * H;
    const double* Ub = pr.U + (size_t)b * 2 * H;
    double* y = ws.at(b, W_Y, k);
    if (k < pr.N) {
        for (int i = 0; i < 8; ++i) y[i] = Xb[i * H + k];
        y[8] = Ub[k]; y[9] = Ub[H + k];
    } else {
        for (int i = 0; i < 5; ++i) y[i] = Xb[i * H + k];
        y[5] = Ub[k]; y[6] = Ub[H + k]; y[7] = y[8] = y[9] = 0.0;
    }
    double *lam = ws.at(b, W_LAM, k), *s = ws.at(b, W_S, k), *z = ws.at(b, W_Z, k);
    for (int i = 0; i < 8; ++i) lam[i] = 0.0;
    for (int i = 0; i < 10; ++i) { s[i] = 1.0; z[i] = 0.0; }
    if (k == 0) {
        double* sc = ws.scal(b);
        for (int i = 0; i < NSCAL; ++i) sc[i] = 0.0;
        sc[S_MU] = pr.mu_start; sc[S_SF] = 1.0; sc[S_THETA0] = -1.0; sc[S_MODE] = MODE_INIT; sc[S_STATUS] = -1.0;
        // outputs are defined whatever happens later (an instance that is never finalised reads as "iteration limit")
        if (pr.status) pr.status[b] = VC_MAX_ITER;
        if (pr.iters) pr.iters[b] = 0;
        if (pr.obj) pr.obj[b] = NAN;
        if (pr.kkt) pr.kkt[b] = INFINITY;
    }
}

// ------------------------------------------------------------------------------------------------
// k_eval_st: eight threads per (instance, SingleTrack stage), one per active variable a.  Each runs the stage functions
// forward in Dual1 arithmetic seeded along a (column a of the dynamics / inequality Jacobians) and the hand-written
// adjoint sweep of vc_adj.cuh in the same arithmetic (gradient of the stage Lagrangian + column a of its exact Hessian);
// the Gauss-Newton part Jc^T (z/s) Jc is assembled from the eight Jacobian columns with warp shuffles.
// 192 threads x 2 CTAs per SM (170 registers): the kernel is ~158 KB of straight-line code that every warp runs once, so
// instruction fetch is its top stall (ncu: "no instruction"); warps of one CTA start together and share the fetches:
// 11.8 ms per full pass against 14.2 (128 x 4), 16.6 (64 x 8), 12.1 (256 x 2), 12.0 (384 x 1), 12.4 (512 x 1)
#ifndef VC_EVAL_MINBLOCKS
#define VC_EVAL_MINBLOCKS 2
#endif
#ifndef VC_EVAL_TPB
#define VC_EVAL_TPB 192
#endif
constexpr int EVAL_ST_LANES = 8;
__global__ void __launch_bounds__(VC_EVAL_TPB, VC_EVAL_MINBLOCKS) k_eval_st(Prob pr, Ws ws) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int N = pr.N;
    const bool live = t < (long long)(*pr.act_count) * N * EVAL_ST_LANES;
    const unsigned mask = __ballot_sync(FULL_MASK, live);   // groups of 8 lanes enter or leave together
    if (!live) return;
    const int ia = (int)(t % EVAL_ST_LANES), k = (int)((t / EVAL_ST_LANES) % N), b = pr.act_list[(int)(t / ((long long)EVAL_ST_LANES * N))];
    const int va = c_act_st[ia];
    const double* yv = ws.at(b, W_Y, k);
    Dual1 y[10];
#pragma unroll
    for (int i = 0; i < 10; ++i) y[i] = Dual1{yv[i], i == va ? 1.0 : 0.0};
    const double sf = ws.scal(b)[S_SF];
    const double kap = pr.kappa[(size_t)b * pr.H + k], ds = pr.ds[(size_t)b * pr.H + k];
    const bool last_st = k == N - 1;
    const int mode = !last_st ? 0 : (pr.M > 0 ? 2 : 1);
    double z[NI_ST], sg[NI_ST], lam[8];
    {
        const double *zz = ws.at(b, W_Z, k), *ss = ws.at(b, W_S, k);
#pragma unroll
        for (int j = 0; j < NI_ST; ++j) { z[j] = zz[j]; sg[j] = 1.0 / ss[j]; }
    }
    const double* yn = mode != 1 ? ws.at(b, W_Y, k + 1) : nullptr;
    if (mode != 1) {
        const double* ln = ws.at(b, W_LAM, k + 1);
#pragma unroll
        for (int i = 0; i < 8; ++i) lam[i] = (mode == 0 || i < 5) ? ln[i] : 0.0;
    }
    Dual1 cost, ineq[NI_ST], xn[8];
    InAdj<Dual1> g;
    double kT = 0.0;
    st_stage_for<Dual1>(pr.p, pr.c, y, kap, ds, mode, sf, z, lam, mode == 2 ? -2.0 * (sf * pr.c.w_switch_F / ds) : 0.0,
                        mode == 2 ? yn[QFY] : 0.0, cost, ineq, xn, g, &kT);
    // ---- column a of the Jacobians, gradients
    double* ab = ws.at(b, W_AB, k);
    double* jc = ws.at(b, W_JC, k);
    double v1 = 0.0, v2 = 0.0, ga = sf * cost.d;
#pragma unroll
    for (int j = 0; j < NI_ST; ++j) {
        jc[j * 8 + ia] = ineq[j].d;
        v1 += ineq[j].d * sg[j];
        v2 += ineq[j].d * z[j] * ineq[j].v * sg[j];
    }
    if (mode == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) ab[i * 8 + ia] = xn[i].d;
    } else if (mode == 2) {
#pragma unroll
        for (int i = 0; i < 6; ++i) ab[i * 8 + ia] = xn[i].d;   // rows 0..4: PointMass state, row 5: grad T
        ga += kT * xn[5].d;
    }
    const Dual1 gg[8] = {g.Ux, g.Uy, g.r, g.delta, g.ey, g.epsi, g.Fx, g.w};
    double rda = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) if (i == ia) rda = gg[i].v;
    if (va < 8) rda += ws.at(b, W_LAM, k)[va];
    ws.at(b, W_RD, k)[va] = rda;
    ws.at(b, W_GF, k)[va] = ga;
    ws.at(b, W_V1, k)[va] = v1;
    ws.at(b, W_V2, k)[va] = v2;
    // ---- Hessian column a: exact Lagrangian part from the adjoint sweep + Gauss-Newton part of the barrier term
    double hcol[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) hcol[i] = gg[i].d;
    const int base = (threadIdx.x & 31) & ~(EVAL_ST_LANES - 1);
    __syncwarp(mask);
#pragma unroll
    for (int j = 0; j < NI_ST; ++j) {
        const double w = z[j] * sg[j] * ineq[j].d;
#pragma unroll
        for (int i = 0; i < 8; ++i) hcol[i] += w * __shfl_sync(mask, ineq[j].d, base + i);
    }
    double* hc = ws.at(b, W_HC, k);
#pragma unroll
    for (int i = 0; i < 8; ++i)
        if (i >= ia) hc[i * (i + 1) / 2 + ia] = hcol[i];
    if (ia == 0) {
        double* cres = ws.at(b, W_CRES, k);
#pragma unroll
        for (int j = 0; j < NI_ST; ++j) cres[j] = ineq[j].v;
        ws.at(b, W_FST, k)[0] = cost.v;
        if (mode != 1) {
            double* gres = ws.at(b, W_GRES, k + 1);
            const int nn = mode == 0 ? 8 : 5;
            for (int i = 0; i < nn; ++i) gres[i] = yn[i] - xn[i].v;
            if (mode == 2) ab[6 * 8] = xn[5].v;
        }
        if (k == 0) {
            double* g0 = ws.at(b, W_GRES, 0);
            for (int i = 0; i < 8; ++i) g0[i] = yv[i] - pr.x0[(size_t)b * 8 + i];
        }
    }
}

// k_eval_pm: same for the PointMass stages (5 active variables -> 15 pairs).
__global__ void __launch_bounds__(128) k_eval_pm(Prob pr, Ws ws) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int M = pr.M, N = pr.N, H = pr.H;
    if (t >= (long long)(*pr.act_count) * M * 15) return;
    const int q = (int)(t % 15), k = N + (int)((t / 15) % M), b = pr.act_list[(int)(t / (15LL * M))];
    const double* sc = ws.scal(b);
    int ia, ib;
    pair_from_index(q, ia, ib);
    const int va = c_act_pm[ia], vb = c_act_pm[ib];
    const bool diag = ia == ib;
    const double* yv = ws.at(b, W_Y, k);
    HDual y[7];
#pragma unroll
    for (int i = 0; i < 7; ++i) y[i] = HDual{yv[i], i == va ? 1.0 : 0.0, i == vb ? 1.0 : 0.0, 0.0};
    const double sf = sc[S_SF];
    const double kap = pr.kappa[(size_t)b * H + k], ds = pr.ds[(size_t)b * H + k];
    HDual cost, ineq[NI_PM];
    pm_cost_ineq(pr.p, pr.c, y, ds, k == H - 1, cost, ineq);
    const double *z = ws.at(b, W_Z, k), *s = ws.at(b, W_S, k);
    double hab = sf * cost.ab, ga = sf * cost.a, rda = ga, v1 = 0.0, v2 = 0.0;
    double* jc = ws.at(b, W_JC, k);
#pragma unroll
    for (int j = 0; j < NI_PM; ++j) {
        const double zj = z[j], is = 1.0 / s[j];
        hab += -zj * ineq[j].ab + zj * is * ineq[j].a * ineq[j].b;
        rda -= zj * ineq[j].a;
        v1 += ineq[j].a * is;
        v2 += ineq[j].a * zj * ineq[j].v * is;
        if (diag) jc[j * 8 + ia] = ineq[j].a;
    }
    if (k < H - 1) {
        HDual xn[5];
        pm_euler<true>(pr.p, y, y + 5, kap, ds, xn);
        const double* lam = ws.at(b, W_LAM, k + 1);
        const double* yn = ws.at(b, W_Y, k + 1);
        double* gres = ws.at(b, W_GRES, k + 1);
        double* ab = ws.at(b, W_AB, k);
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            hab -= lam[i] * xn[i].ab;
            rda -= lam[i] * xn[i].a;
            if (diag) ab[i * 8 + ia] = xn[i].a;
            if (q == 0) gres[i] = yn[i] - xn[i].v;
        }
    }
    if (diag && va < 5) rda += ws.at(b, W_LAM, k)[va];
    ws.at(b, W_HC, k)[q] = hab;
    if (diag) {
        ws.at(b, W_RD, k)[va] = rda;
        ws.at(b, W_GF, k)[va] = ga;
        ws.at(b, W_V1, k)[va] = v1;
        ws.at(b, W_V2, k)[va] = v2;
    }
    if (q == 0) {
        double* cres = ws.at(b, W_CRES, k);
        cres[0] = ineq[0].v; cres[1] = ineq[1].v;
        ws.at(b, W_FST, k)[0] = cost.v;
        if (N == 0 && k == 0) {   // pure PointMass controller (not used by the shipped configs)
            double* g0 = ws.at(b, W_GRES, 0);
            for (int i = 0; i < 5; ++i) g0[i] = yv[i] - pr.x0[(size_t)b * 8 + i];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// warp helpers
#define FULL 0xffffffffu
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(FULL, v, o));
    return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(FULL, v, o));
    return v;
}
// NaN-propagating max (fmax drops NaNs; evaluation errors must not be hidden)
__device__ __forceinline__ double nmax(double a, double b) { return (a != a || b != b) ? NAN : fmax(a, b); }
__device__ __forceinline__ double warp_nmax(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = nmax(v, __shfl_xor_sync(FULL, v, o));
    return v;
}

// stage geometry ---------------------------------------------------------------------------------
struct Dims { int nx, np, n, nu_off; bool st; };
__device__ __forceinline__ Dims stage_dims(int N, int k) {
    Dims d;
    d.st = k < N;
    d.nx = d.st ? 8 : 5;
    d.np = k == 0 ? 0 : (d.st ? 1 : 2);
    d.n = d.nx + d.np;
    d.nu_off = d.st ? 8 : 5;   // index of the first control inside the stage variable vector y
    return d;
}
// weights of the control-continuity terms between stage k and k+1 (unscaled): (c0 on Fx, c1 on Fy)
//   cascaded_mpc.py:167-171 (SingleTrack), :228-231 (PointMass), :241-254 (switching; c1 ties Fy_N to Fy_f+Fy_r)
__device__ __forceinline__ void slew_weights(const Prob& pr, const double* dsb, int k, double& c0, double& c1) {
    c0 = c1 = 0.0;
    if (k < 0 || k >= pr.H - 1) return;
    const double ids = 1.0 / dsb[k];
    if (k < pr.N - 1) c0 = pr.c.w_Fx * ids;
    else if (k == pr.N - 1) { c0 = pr.c.w_switch_F * ids; c1 = c0; }
    else { c0 = pr.c.w_Fx * ids; c1 = pr.c.w_Fy * ids; }
}
// previous-stage value that control j of stage k+1 is tied to
__device__ __forceinline__ double slew_prev(const Prob& pr, const Ws& ws, int b, int k, int j, const double* yk) {
    if (k == pr.N - 1 && pr.M > 0 && j == 1) return ws.at(b, W_AB, k)[6 * 8];   // T = Fy_f + Fy_r (from k_eval_st)
    return yk[(k < pr.N ? 8 : 5) + j];
}

// Values of one stage at a trial point (plain doubles): objective share, constraint violation share,
// log-barrier share.  y = stage variables, yn = next stage's (nullptr at the last stage), sl = trial slacks.
struct StageVal { double f, theta, lns, gmax, cmin; };
__device__ __noinline__ StageVal stage_trial(const Prob& pr, int b, int k, const double* y, const double* yn, const double* sl,
                                double* og0 = nullptr, double* ogn = nullptr, double* oc = nullptr) {
    StageVal o{0.0, 0.0, 0.0, 0.0, INFINITY};
    const int H = pr.H, N = pr.N;
    const double kap = pr.kappa[(size_t)b * H + k];
    const double* dsb = pr.ds + (size_t)b * H;
    const double ds = dsb[k];
    double c0, c1;
    slew_weights(pr, dsb, k, c0, c1);
    if (k < N) {
        double cost, ineq[NI_ST];
        st_cost_ineq(pr.p, pr.c, y, ds, k == N - 1 && pr.M == 0, cost, ineq);
        o.f = cost;
#pragma unroll
        for (int j = 0; j < NI_ST; ++j) { o.theta += fabs(ineq[j] - sl[j]); o.lns += log(sl[j]); o.cmin = fmin(o.cmin, ineq[j]); if (oc) oc[j] = ineq[j]; }
        if (k == 0)
            for (int i = 0; i < 8; ++i) { const double r = y[i] - pr.x0[(size_t)b * 8 + i], e = fabs(r); o.theta += e; o.gmax = nmax(o.gmax, e); if (og0) og0[i] = r; }
        if (k < N - 1) {
            double xn[8];
            st_rk4<true>(pr.p, y, y + 8, kap, ds, xn);
#pragma unroll
            for (int i = 0; i < 8; ++i) { const double r = yn[i] - xn[i], e = fabs(r); o.theta += e; o.gmax = nmax(o.gmax, e); if (ogn) ogn[i] = r; }
            o.f += c0 * sq(yn[8] - y[8]);
        } else if (pr.M > 0) {
            double xp[5], T;
            st_switch(pr.p, y, xp, T);
#pragma unroll
            for (int i = 0; i < 5; ++i) { const double r = yn[i] - xp[i], e = fabs(r); o.theta += e; o.gmax = nmax(o.gmax, e); if (ogn) ogn[i] = r; }
            o.f += c0 * (sq(yn[QFX] - y[8]) + sq(yn[QFY] - T));
        }
        if (pr.c.obstacles && pr.n_obs > 0)
            o.f += obstacle_cost(pr.obst + (size_t)b * pr.n_obs * 3, pr.n_obs, pr.c.w_obstacles, ds, y[IS], y[IEY],
                                 nullptr, nullptr);
        if (k == H - 1) o.f += pr.c.w_time * y[IT];
    } else {
        double cost, ineq[NI_PM];
        pm_cost_ineq(pr.p, pr.c, y, ds, k == H - 1, cost, ineq);
        o.f = cost;
        for (int j = 0; j < NI_PM; ++j) { o.theta += fabs(ineq[j] - sl[j]); o.lns += log(sl[j]); o.cmin = fmin(o.cmin, ineq[j]); if (oc) oc[j] = ineq[j]; }
        if (k < H - 1) {
            double xn[5];
            pm_euler<true>(pr.p, y, y + 5, kap, ds, xn);
#pragma unroll
            for (int i = 0; i < 5; ++i) { const double r = yn[i] - xn[i], e = fabs(r); o.theta += e; o.gmax = nmax(o.gmax, e); if (ogn) ogn[i] = r; }
            o.f += c0 * sq(yn[QFX] - y[QFX]) + c1 * sq(yn[QFY] - y[QFY]);
        }
        if (pr.c.obstacles && pr.n_obs > 0)
            o.f += obstacle_cost(pr.obst + (size_t)b * pr.n_obs * 3, pr.n_obs, pr.c.w_obstacles, ds, y[QS], y[QEY],
                                 nullptr, nullptr);
        if (k == H - 1) o.f += pr.c.w_time * y[QT];
    }
    return o;
}

// ------------------------------------------------------------------------------------------------
// k_step building blocks (one warp per instance; `lane` strides over stages or matrix entries)
#ifndef VC_STEP_WARPS
#define VC_STEP_WARPS 4
#endif
#ifndef VC_STEP_MINBLOCKS
#define VC_STEP_MINBLOCKS 4
#endif
constexpr int STEP_WARPS = VC_STEP_WARPS;
constexpr int RING_DOUBLES = 768;
struct __align__(16) WarpSm {
    double P[100];     // cost-to-go Hessian of the NEXT stage (stride 10)
    double M[120];     // [A~ B~] of the lifted transition (stride 12)
    double PMm[120];   // P * M
    double G[144];     // stage Hessian + M^T P M (stride 12), order [x | p | u]
    double p[12], r[12], g[12], x[12], tv[12], K[24];
    double ring[RING_DOUBLES];   // stage blocks staged by cp.async ahead of the Riccati sweeps
};

struct Measures { double rdmax, prim, sumlam, sumz, theta, f, lns, gfmax; };

// IPOPT options the reference leaves at their defaults (cascaded_mpc.py:55-65)
struct IpmConst {
    static constexpr double kappa_eps = 10.0, kappa_mu = 0.2, theta_mu = 1.5, tau_min = 0.99, slack_push = 1e-3,
        kappa_sigma = 1e10, s_max = 100.0, gamma_theta = 1e-5, gamma_phi = 1e-8, eta_phi = 1e-8, delta = 1.0,
        s_theta = 1.1, s_phi = 2.3, max_gradient = 100.0;
};

// completes the gradients that k_eval left open (s, t, obstacle and control-continuity parts) and reduces
// the error measures over the warp
__device__ __noinline__ Measures finish_gradients(const Prob& pr, const Ws& ws, int b, int lane) {
    const int H = pr.H, N = pr.N;
    const double* sc = ws.scal(b);
    const double sf = sc[S_SF];
    const double* dsb = pr.ds + (size_t)b * H;
    Measures m{0, 0, 0, 0, 0, 0, 0, 0};
    for (int k = lane; k < H; k += 32) {
        const Dims d = stage_dims(N, k);
        const double* y = ws.at(b, W_Y, k);
        const int si = d.st ? IS : QS, ti = d.st ? IT : QT, eyi = d.st ? IEY : QEY;
        double og[2] = {0, 0}, oh[3] = {0, 0, 0}, ov = 0.0;
        if (pr.c.obstacles && pr.n_obs > 0)
            ov = obstacle_cost(pr.obst + (size_t)b * pr.n_obs * 3, pr.n_obs, pr.c.w_obstacles, dsb[k], y[si], y[eyi], og, oh);
        double* ob = ws.at(b, W_OBS, k);
        ob[0] = ov; ob[1] = og[0]; ob[2] = og[1]; ob[3] = oh[0]; ob[4] = oh[1]; ob[5] = oh[2];
        double p0, p1, c0, c1, g0 = 0.0, g1 = 0.0, fsl = 0.0;
        slew_weights(pr, dsb, k - 1, p0, p1);
        ob[6] = p0; ob[7] = p1;
        slew_weights(pr, dsb, k, c0, c1);
        const double u0 = y[d.nu_off], u1 = y[d.nu_off + 1];
        if (k > 0) {
            const double* yp = ws.at(b, W_Y, k - 1);
            g0 += 2.0 * p0 * (u0 - slew_prev(pr, ws, b, k - 1, 0, yp));
            g1 += 2.0 * p1 * (u1 - slew_prev(pr, ws, b, k - 1, 1, yp));
        }
        double *rd = ws.at(b, W_RD, k), *gf = ws.at(b, W_GF, k), *v1 = ws.at(b, W_V1, k), *v2 = ws.at(b, W_V2, k);
        const double* lam = ws.at(b, W_LAM, k);
        double lam_s_next = 0.0, lam_t_next = 0.0;
        if (k < H - 1) {
            const Dims dn = stage_dims(N, k + 1);
            const double* yn = ws.at(b, W_Y, k + 1);
            const double e0 = yn[dn.nu_off] - u0;
            g0 -= 2.0 * c0 * e0;
            fsl += c0 * e0 * e0;
            if (k == N - 1) {   // switching: Fy_N is tied to T(y_{N-1}); that gradient was taken by k_eval_st
                const double e1 = yn[QFY] - slew_prev(pr, ws, b, k, 1, y);
                fsl += c1 * e1 * e1;
            } else {
                const double e1 = yn[dn.nu_off + 1] - u1;
                g1 -= 2.0 * c1 * e1;
                fsl += c1 * e1 * e1;
            }
            const double* ln = ws.at(b, W_LAM, k + 1);
            lam_s_next = ln[dn.st ? IS : QS];
            lam_t_next = ln[dn.st ? IT : QT];
        }
        const double wt = (k == H - 1) ? sf * pr.c.w_time : 0.0;
        rd[d.nu_off] += sf * g0; gf[d.nu_off] += sf * g0;
        rd[d.nu_off + 1] += sf * g1; gf[d.nu_off + 1] += sf * g1;
        rd[si] = lam[si] - lam_s_next + sf * og[0]; gf[si] = sf * og[0];
        rd[ti] = lam[ti] - lam_t_next + wt; gf[ti] = wt;
        rd[eyi] += sf * og[1]; gf[eyi] += sf * og[1];
        v1[si] = v1[ti] = v2[si] = v2[ti] = 0.0;
        for (int i = 0; i < d.nx + 2; ++i) { m.rdmax = nmax(m.rdmax, fabs(rd[i])); m.gfmax = nmax(m.gfmax, fabs(gf[i])); }
        const double* gr = ws.at(b, W_GRES, k);
        for (int i = 0; i < d.nx; ++i) { m.prim = nmax(m.prim, fabs(gr[i])); m.theta += fabs(gr[i]); m.sumlam += fabs(lam[i]); }
        const double *cr = ws.at(b, W_CRES, k), *s = ws.at(b, W_S, k), *z = ws.at(b, W_Z, k);
        for (int j = 0; j < (d.st ? NI_ST : NI_PM); ++j) {
            const double r = cr[j] - s[j];
            m.prim = nmax(m.prim, fabs(r)); m.theta += fabs(r); m.sumz += fabs(z[j]); m.lns += log(s[j]);
        }
        m.f += ws.at(b, W_FST, k)[0] + ov + fsl + (k == H - 1 ? pr.c.w_time * y[ti] : 0.0);
    }
    m.rdmax = warp_nmax(m.rdmax); m.gfmax = warp_nmax(m.gfmax); m.prim = warp_nmax(m.prim);
    m.sumlam = warp_sum(m.sumlam); m.sumz = warp_sum(m.sumz); m.theta = warp_sum(m.theta);
    m.f = warp_sum(m.f); m.lns = warp_sum(m.lns);
    return m;
}

__device__ double compl_error(const Prob& pr, const Ws& ws, int b, int lane, double mu) {
    double e = 0.0;
    for (int k = lane; k < pr.H; k += 32) {
        const double *s = ws.at(b, W_S, k), *z = ws.at(b, W_Z, k);
        for (int j = 0; j < (k < pr.N ? NI_ST : NI_PM); ++j) e = fmax(e, fabs(s[j] * z[j] - mu));
    }
    return warp_max(e);
}

// slack / multiplier initialisation (IPOPT warm_start_*_bound_push 1e-3; oracle/ipm.py)
__device__ void init_slacks(const Prob& pr, const Ws& ws, int b, int lane, double mu) {
    for (int k = lane; k < pr.H; k += 32) {
        const double* cr = ws.at(b, W_CRES, k);
        double *s = ws.at(b, W_S, k), *z = ws.at(b, W_Z, k), *lam = ws.at(b, W_LAM, k);
        for (int j = 0; j < (k < pr.N ? NI_ST : NI_PM); ++j) {
            s[j] = fmax(cr[j], IpmConst::slack_push * fmax(1.0, fabs(cr[j])));
            z[j] = mu / s[j];
        }
        for (int i = 0; i < 8; ++i) lam[i] = 0.0;
    }
}

// writes the result of a finished instance back in the reference's layout
__device__ void finalize(const Prob& pr, const Ws& ws, int b, int lane, int status, double obj, double err) {
    const int H = pr.H;
    double* Xb = pr.X + (size_t)b * 8 * H;
    double* Ub = pr.U + (size_t)b * 2 * H;
    for (int k = lane; k < H; k += 32) {
        const double* y = ws.at(b, W_Y, k);
        if (k < pr.N) {
            for (int i = 0; i < 8; ++i) Xb[i * H + k] = y[i];
            Ub[k] = y[8]; Ub[H + k] = y[9];
        } else {
            for (int i = 0; i < 5; ++i) Xb[i * H + k] = y[i];   // rows 5..7 of PointMass columns are free variables
            Ub[k] = y[5]; Ub[H + k] = y[6];
        }
    }
    if (lane == 0) {
        double* sc = ws.scal(b);
        sc[S_MODE] = MODE_DONE; sc[S_STATUS] = status; sc[S_OBJ] = obj; sc[S_ERR] = err;
        if (pr.obj) pr.obj[b] = obj;
        if (pr.kkt) pr.kkt[b] = err;
        if (pr.status) pr.status[b] = status;
        if (pr.iters) pr.iters[b] = (int)sc[S_ITERS];
    }
}

// position of stage variable i of stage k inside the Riccati ordering [x | p | u]
__device__ __forceinline__ int gpos(const Dims& d, int i) { return i < d.nu_off ? i : d.n + (i - d.nu_off); }

// ---- cp.async staging of stage blocks (16-byte copies, L2 -> shared memory, bypassing L1)
__device__ __forceinline__ void cp_async16(double* smem, const double* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// offsets inside a stage record / inside the staged copies
constexpr int O_HC = ws_offset(W_HC), O_AB = ws_offset(W_AB), O_RD = ws_offset(W_RD), O_V1 = ws_offset(W_V1),
              O_V2 = ws_offset(W_V2), O_OBS = ws_offset(W_OBS), O_AUX = ws_offset(W_AUX), O_GRES = ws_offset(W_GRES);
constexpr int BWD_DEPTH = 4, BWD_SLOT = WS_BWD + 2;                 // backward ring: 4 slots of the [0, WS_BWD) block
// forward ring: 4 slots of  AB | GRES KM PV PM[0..70)  (the reduced cost-to-go Hessian is at most 7 x 7, stride 10)
constexpr int FWD_PM_USED = 70;
constexpr int FWD_TAIL = ws_offset(W_PM) + FWD_PM_USED - WS_FWD0;
constexpr int FWD_DEPTH = 4, FWD_SLOT = 80 + FWD_TAIL;
constexpr int F_AB = 0, F_GRES = 80, F_KM = F_GRES + 8, F_PV = F_KM + 24, F_PM = F_PV + 10;
static_assert(BWD_DEPTH * BWD_SLOT <= RING_DOUBLES && FWD_DEPTH * FWD_SLOT <= RING_DOUBLES, "ring too small");
static_assert(BWD_SLOT % 2 == 0 && FWD_SLOT % 2 == 0 && FWD_TAIL % 2 == 0, "slot alignment");
static_assert(ws_offset(W_KM) == WS_FWD0 + 8 && ws_offset(W_PV) == WS_FWD0 + 32 && ws_offset(W_PM) == WS_FWD0 + 42, "forward slot layout");
// right-hand-side-only backward sweep: 3 slots of  [AB .. GRES]  |  [KM PV PM[0..70)]  of the stage record
constexpr int RHS_A = WS_BWD - ws_offset(W_AB), RHS_B = 24 + 10 + FWD_PM_USED, RHS_DEPTH = 3, RHS_SLOT = RHS_A + RHS_B;
static_assert(RHS_DEPTH * RHS_SLOT <= RING_DOUBLES && RHS_A % 2 == 0 && RHS_B % 2 == 0, "rhs ring");

// A solve step is run by a GROUP of NT threads per instance: one warp (NT = 32, four instances per block) while the
// batch is large, four warps (NT = 128, named barrier) once few instances are left and latency is what counts.
// Every entry of every matrix is computed by the same arithmetic in both cases, so results do not depend on NT.
template <int NT> __device__ __forceinline__ void gsync() {
    if (NT == 32) __syncwarp();
    else asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
}

// ---- reduced stage state of the Riccati sweeps.  The arclength s and the time t are cyclic coordinates of the
// linearised problem: s_{k+1} = s_k + ds_k has no control behind it, so its direction is the running sum of the
// equality residuals and only shifts the linear term of ey (obstacle cross term); t appears in no Hessian block, so its
// multiplier direction is a running sum of dual residuals that enters the other rows through d t_{k+1} / d y_k.  Both
// are handled by prefix sums over the stages (lanes over stages) and the matrix recursion runs on
//   SingleTrack  (Ux Uy r delta ey epsi | lifted u_{k-1})   6 + np        PointMass  (V ey epsi | lifted)   3 + np
struct RDims { int nxr, np, n, nuo; bool st; };
__device__ __forceinline__ RDims rdims(int N, int k) {
    RDims d;
    d.st = k < N;
    d.nxr = d.st ? 6 : 3;
    d.np = k == 0 ? 0 : (d.st ? 1 : 2);
    d.n = d.nxr + d.np;
    d.nuo = d.st ? 8 : 5;   // index of the first control inside the stage variable vector y
    return d;
}
// position of ACTIVE variable a (index into c_act_st / c_act_pm) in the Riccati ordering [x^ | p | u]; the first nxr
// active variables are the reduced state in order, the last two the controls
__device__ __forceinline__ int apos(const RDims& d, int a) { return a < d.nxr ? a : d.n + (a - d.nxr); }
// index inside y / inside a full state vector of ACTIVE variable a (arithmetic form of c_act_st / c_act_pm)
__device__ __forceinline__ int yidx(bool st, int a) { return st ? a + (a >= 4) + (a >= 6) : a + (a >= 1) + (a >= 3); }

// ---- per stage KIND index tables (shared memory, built once per block): where every entry of the stage matrices comes
// from inside the staged stage block, so that a stage is assembled by straight table-driven copies
enum { NKIND = 5, G_MAX = 81, M_MAX = 63 };
enum { GF_EY = 1, GF_P0 = 2, GF_M0 = 4, GF_P1 = 8, GF_M1 = 16, GF_DELTA = 32 };
struct KindTab {
    short gdst[G_MAX], gsrc[G_MAX];   // G entry: destination (stride 12); offset of its Hessian entry in the block or -1
    short mdst[M_MAX], msrc[M_MAX];   // M entry: destination (stride 12); offset inside the AB block, -1 = 0, -2 = 1
    unsigned char gflag[G_MAX];       // closed-form additions (obstacle ey-ey, control continuity, delta_w)
    short ng, nm, n, nn, nxr, np, dnn, dnxr, st, nst, has_next, trow;
};
__device__ __forceinline__ int stage_kind(int k, int N, int H) {
    return k == 0 ? 0 : (k == H - 1 ? 4 : (k == N - 1 ? 2 : (k < N - 1 ? 1 : 3)));
}
__device__ void build_kind_tables(const Prob& pr, KindTab* tabs, int tid, int nthreads) {
    const int N = pr.N, H = pr.H;
    const int rep[NKIND] = {0, 1, N - 1, N, H - 1};   // a stage of each kind (kinds that do not occur are never looked up)
    for (int kd = 0; kd < NKIND; ++kd) {
        const int k = rep[kd];
        if (k < 0 || k >= H || stage_kind(k, N, H) != kd) continue;
        KindTab& T = tabs[kd];
        const RDims d = rdims(N, k);
        const bool has_next = k < H - 1;
        const RDims dn = rdims(N, has_next ? k + 1 : k);
        const int n = d.n, nn = n + 2;
        if (tid == 0) {
            T.ng = nn * nn; T.nm = has_next ? dn.n * nn : 0; T.n = n; T.nn = nn; T.nxr = d.nxr; T.np = d.np;
            T.dnn = dn.n; T.dnxr = dn.nxr; T.st = d.st; T.nst = dn.st; T.has_next = has_next;
            T.trow = (k + 1 < N ? IT : QT);
        }
        for (int e = tid; e < nn * nn; e += nthreads) {
            const int i = e / nn, j = e - i * nn;
            // active variable behind a position: reduced state, or control; lifted positions have none
            const int ai = i < d.nxr ? i : (i >= n ? d.nxr + (i - n) : -1), aj = j < d.nxr ? j : (j >= n ? d.nxr + (j - n) : -1);
            int src = -1;
            if (ai >= 0 && aj >= 0) { const int hi = ai > aj ? ai : aj, lo = ai > aj ? aj : ai; src = O_HC + hi * (hi + 1) / 2 + lo; }
            int fl = 0;
            const int eyp = d.st ? 4 : 1;
            if (i == eyp && j == eyp) fl |= GF_EY;
            for (int q = 0; q < d.np; ++q) {
                const int pu = n + q, pp = d.nxr + q;
                if ((i == pu && j == pu) || (i == pp && j == pp)) fl |= q == 0 ? GF_P0 : GF_P1;
                if ((i == pu && j == pp) || (i == pp && j == pu)) fl |= q == 0 ? GF_M0 : GF_M1;
            }
            if (i == j && (i < d.nxr || i >= n)) fl |= GF_DELTA;
            T.gdst[e] = (short)(i * 12 + j); T.gsrc[e] = (short)src; T.gflag[e] = (unsigned char)fl;
        }
        if (has_next)
            for (int e = tid; e < dn.n * nn; e += nthreads) {
                const int l = e / nn, j = e - l * nn;
                const int aj = j < d.nxr ? j : (j >= n ? d.nxr + (j - n) : -1);
                int src = -1;
                if (l < dn.nxr) { if (aj >= 0) src = yidx(dn.st, l) * 8 + aj; }
                else if (l == dn.nxr) { if (j == n) src = -2; }                         // p0 <- Fx_k
                else if (!d.st) { if (j == n + 1) src = -2; }                           // p1 <- Fy_k
                else if (aj >= 0) src = 5 * 8 + aj;                                    // p1 <- T(y_{N-1}) = Fy_f + Fy_r
                T.mdst[e] = (short)(l * 12 + j); T.msrc[e] = (short)src;
            }
    }
}

__device__ __forceinline__ void stage_fetch_bwd(const double* rec0, WarpSm& sm, int k, int rank, int nt) {
    if (k >= 0) {
        const double* rec = rec0 + k * WS_PER_STAGE;
        double* dst = sm.ring + (k % BWD_DEPTH) * BWD_SLOT;
        for (int c = rank; c < WS_BWD / 2; c += nt) cp_async16(dst + 2 * c, rec + 2 * c);
    }
    cp_async_commit();
}
__device__ __forceinline__ void stage_fetch_rhs(const double* rec0, WarpSm& sm, int k, int rank, int nt) {
    if (k >= 0) {
        const double* rec = rec0 + k * WS_PER_STAGE;
        double* dst = sm.ring + (k % RHS_DEPTH) * RHS_SLOT;
        for (int c = rank; c < (RHS_A + RHS_B) / 2; c += nt) {
            const bool a = c < RHS_A / 2;
            cp_async16(dst + 2 * c, rec + (a ? O_AB + 2 * c : O_KM + 2 * (c - RHS_A / 2)));
        }
    }
    cp_async_commit();
}
__device__ __forceinline__ void stage_fetch_fwd(const double* rec0, WarpSm& sm, int k, int H, int rank, int nt) {
    if (k < H) {
        const double* rec = rec0 + k * WS_PER_STAGE;
        double* dst = sm.ring + (k % FWD_DEPTH) * FWD_SLOT;
        for (int c = rank; c < 40 + FWD_TAIL / 2; c += nt) {
            const bool ab = c < 40;
            cp_async16(dst + (ab ? F_AB + 2 * c : F_GRES + 2 * (c - 40)), rec + (ab ? O_AB + 2 * c : WS_FWD0 + 2 * (c - 40)));
        }
    }
    cp_async_commit();
}

// ---- stage products of the backward sweep: PMm = P M, tv = P r + p, G += M^T PMm, g += M^T tv
// generic shapes: four interleaved accumulators per lane, entry indices by reciprocal multiplication
__device__ __noinline__ void stage_products_generic(WarpSm& sm, int lane, int dnn, int nn) {
    const int cnt = dnn * nn, rcp_nn = 65536 / nn + 1;
    {   // PMm = P M : up to 4 entries per lane, accumulated side by side
        int ii[4], jj[4];
        double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int e = min(lane + 32 * t, cnt - 1);
            ii[t] = (e * rcp_nn) >> 16; jj[t] = e - ii[t] * nn;
        }
        for (int l = 0; l < dnn; ++l) {
#pragma unroll
            for (int t = 0; t < 4; ++t) acc[t] += sm.P[ii[t] * 10 + l] * sm.M[l * 12 + jj[t]];
        }
#pragma unroll
        for (int t = 0; t < 4; ++t)
            if (lane + 32 * t < cnt) sm.PMm[ii[t] * 12 + jj[t]] = acc[t];
    }
    if (lane < dnn) {                                 // tv = P r + p
        double a = sm.p[lane];
        for (int l = 0; l < dnn; ++l) a += sm.P[lane * 10 + l] * sm.r[l];
        sm.tv[lane] = a;
    }
    __syncwarp();
    {   // G += M^T (P M)
        const int cnt2 = nn * nn;
        int ii[4], jj[4];
        double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int e = min(lane + 32 * t, cnt2 - 1);
            ii[t] = (e * rcp_nn) >> 16; jj[t] = e - ii[t] * nn;
        }
        for (int l = 0; l < dnn; ++l) {
#pragma unroll
            for (int t = 0; t < 4; ++t) acc[t] += sm.M[l * 12 + ii[t]] * sm.PMm[l * 12 + jj[t]];
        }
#pragma unroll
        for (int t = 0; t < 4; ++t)
            if (lane + 32 * t < cnt2) sm.G[ii[t] * 12 + jj[t]] += acc[t];
    }
    if (lane < nn) {                                   // g += M^T tv
        double a = 0.0;
        for (int l = 0; l < dnn; ++l) a += sm.M[l * 12 + lane] * sm.tv[l];
        sm.g[lane] += a;
    }
    __syncwarp();
}

// compile-time shapes: every lane owns a 2x2 block of the result (one shared-memory load per multiply-add instead of
// two, 16-byte loads along the rows); G is symmetric, so only its upper block triangle is computed and mirrored
template <int DN, int NN>
__device__ __forceinline__ void stage_products(WarpSm& sm, int lane) {
    constexpr int RB = (DN + 1) / 2, CB = (NN + 1) / 2;
    static_assert(RB * CB <= 32 && CB * (CB + 1) / 2 <= 32, "one block per lane");
    if (lane < RB * CB) {
        const int bi = lane / CB, bj = lane - bi * CB, i0 = 2 * bi, j0 = 2 * bj;
        double a00 = 0.0, a01 = 0.0, a10 = 0.0, a11 = 0.0;
#pragma unroll
        for (int l = 0; l < DN; ++l) {
            const double p0 = sm.P[i0 * 10 + l], p1 = sm.P[(i0 + 1) * 10 + l];
            const double2 m = *reinterpret_cast<const double2*>(&sm.M[l * 12 + j0]);
            a00 += p0 * m.x; a01 += p0 * m.y; a10 += p1 * m.x; a11 += p1 * m.y;
        }
        sm.PMm[i0 * 12 + j0] = a00;
        if (j0 + 1 < NN) sm.PMm[i0 * 12 + j0 + 1] = a01;
        if (i0 + 1 < DN) {
            sm.PMm[(i0 + 1) * 12 + j0] = a10;
            if (j0 + 1 < NN) sm.PMm[(i0 + 1) * 12 + j0 + 1] = a11;
        }
    }
    if (lane >= 32 - DN) {                                 // tv = P r + p on the idle lanes
        const int i = lane - (32 - DN);
        double a = sm.p[i];
#pragma unroll
        for (int l = 0; l < DN; ++l) a += sm.P[i * 10 + l] * sm.r[l];
        sm.tv[i] = a;
    }
    __syncwarp();
    if (lane < CB * (CB + 1) / 2) {                        // upper block triangle of G += M^T (P M)
        int bj = (int)((sqrtf(8.0f * lane + 1.0f) - 1.0f) * 0.5f);
        if ((bj + 1) * (bj + 2) / 2 <= lane) ++bj;
        if (bj * (bj + 1) / 2 > lane) --bj;
        const int bi = lane - bj * (bj + 1) / 2;           // bi <= bj
        const int i0 = 2 * bi, j0 = 2 * bj;
        double a00 = 0.0, a01 = 0.0, a10 = 0.0, a11 = 0.0;
#pragma unroll
        for (int l = 0; l < DN; ++l) {
            const double2 mi = *reinterpret_cast<const double2*>(&sm.M[l * 12 + i0]);
            const double2 q = *reinterpret_cast<const double2*>(&sm.PMm[l * 12 + j0]);
            a00 += mi.x * q.x; a01 += mi.x * q.y; a10 += mi.y * q.x; a11 += mi.y * q.y;
        }
        const bool i1 = i0 + 1 < NN, j1 = j0 + 1 < NN;
        sm.G[i0 * 12 + j0] += a00;
        if (j1) sm.G[i0 * 12 + j0 + 1] += a01;
        if (i1) sm.G[(i0 + 1) * 12 + j0] += a10;
        if (i1 && j1) sm.G[(i0 + 1) * 12 + j0 + 1] += a11;
        if (bi != bj) {                                    // mirror block
            sm.G[j0 * 12 + i0] += a00;
            if (j1) sm.G[(j0 + 1) * 12 + i0] += a01;
            if (i1) sm.G[j0 * 12 + i0 + 1] += a10;
            if (i1 && j1) sm.G[(j0 + 1) * 12 + i0 + 1] += a11;
        }
    } else if (lane >= 32 - NN) {                          // g += M^T tv on the idle lanes
        const int c = lane - (32 - NN);
        double a = 0.0;
#pragma unroll
        for (int l = 0; l < DN; ++l) a += sm.M[l * 12 + c] * sm.tv[l];
        sm.g[c] += a;
    }
    __syncwarp();
}

// the same products with one entry per thread (block-wide groups).  `blocked` = the warp path would use the register-
// blocked template for this shape: there only the upper block triangle of G is computed and mirrored, which this
// reproduces entry for entry (2x2 blocks: (i,j) with i/2 > j/2 takes the value computed for (j,i)).
template <int NT>
__device__ __forceinline__ void stage_products_wide(WarpSm& sm, int rank, int dnn, int nn, bool blocked) {
    for (int e = rank; e < dnn * nn; e += NT) {                 // PMm = P M
        const int i = e / nn, j = e - i * nn;
        double acc = 0.0;
        for (int l = 0; l < dnn; ++l) acc += sm.P[i * 10 + l] * sm.M[l * 12 + j];
        sm.PMm[i * 12 + j] = acc;
    }
    if (rank >= NT - 16 && rank - (NT - 16) < dnn) {            // tv = P r + p
        const int i = rank - (NT - 16);
        double a = sm.p[i];
        for (int l = 0; l < dnn; ++l) a += sm.P[i * 10 + l] * sm.r[l];
        sm.tv[i] = a;
    }
    gsync<NT>();
    for (int e = rank; e < nn * nn; e += NT) {                  // G += M^T (P M)
        const int i = e / nn, j = e - i * nn;
        const bool mirror = blocked && (i >> 1) > (j >> 1);
        const int ii = mirror ? j : i, jj = mirror ? i : j;
        double acc = 0.0;
        for (int l = 0; l < dnn; ++l) acc += sm.M[l * 12 + ii] * sm.PMm[l * 12 + jj];
        sm.G[i * 12 + j] += acc;
    }
    if (rank >= NT - 16 && rank - (NT - 16) < nn) {             // g += M^T tv
        const int c = rank - (NT - 16);
        double a = 0.0;
        for (int l = 0; l < dnn; ++l) a += sm.M[l * 12 + c] * sm.tv[l];
        sm.g[c] += a;
    }
    gsync<NT>();
}

// warp prefix / suffix sums (inclusive) used by the s / t passes
__device__ __forceinline__ double warp_scan_up(double v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double t = __shfl_up_sync(FULL, v, o); if (lane >= o) v += t; }
    return v;
}
__device__ __forceinline__ double warp_scan_down(double v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double t = __shfl_down_sync(FULL, v, o); if (lane + o < 32) v += t; }
    return v;
}

// Backward Riccati sweep + forward substitution of the condensed KKT system
//   [Hc + delta I  Jg^T] [dw  ]   [-(rd - mu v1 + v2)]
//   [Jg            0   ] [dlam] = [-g                ]
// Returns false when some reduced Hessian block is not positive definite (wrong inertia for this delta).
// The per-stage blocks are streamed through a shared-memory ring with cp.async, several stages ahead of the math;
// every stage matrix is assembled by table-driven copies out of the staged block (KindTab).
template <int NT, bool RHS>
__device__ __noinline__ bool riccati(const Prob& pr, const Ws& ws, WarpSm& sm, const KindTab* tabs, int b, int rank, double mu, double delta, bool soc) {
    const int H = pr.H, N = pr.N;
    // second-order correction: same matrix, residuals replaced by the accumulated c_soc (IPOPT 2.4 / Alg. A-5.7)
    const int o_gres = soc ? O_GS : O_GRES_C, o_dy = soc ? O_DIR2 : O_DIR, o_dlam = o_dy + 10;
    double* const rec0 = ws.rec(b, 0);
    const double sf = ws.scal(b)[S_SF];
    bool ok = true;
    // ---- pre-pass (first warp, lanes over stages): direction of s = running sum of its equality residuals;
    // cost-to-go gradient of t = running sum (from the end) of its dual residuals
    if (rank < 32) {
        double carry = 0.0;
        for (int base = 0; base < H; base += 32) {
            const int k = base + rank;
            double* rec = rec0 + k * WS_PER_STAGE;
            double v = k < H ? -rec[o_gres + (k < N ? IS : QS)] : 0.0;
            v = warp_scan_up(v, rank) + carry;
            if (k < H) { rec[o_dy + (k < N ? IS : QS)] = v; rec[O_AUX] = v; }
            carry = __shfl_sync(FULL, v, 31);
        }
        carry = 0.0;
        for (int base = ((H - 1) / 32) * 32; base >= 0; base -= 32) {
            const int k = base + rank;
            double* rec = rec0 + k * WS_PER_STAGE;
            double v = k < H ? rec[O_RD + (k < N ? IT : QT)] : 0.0;
            v = warp_scan_down(v, rank) + carry;
            if (k < H) {
                rec[o_dlam + (k < N ? IT : QT)] = -v;
                if (k >= 1) rec[O_AUX + 1 - WS_PER_STAGE] = v;
                if (k == H - 1) rec[O_AUX + 1] = 0.0;
            }
            carry = __shfl_sync(FULL, v, 0);
        }
    }
    gsync<NT>();
    if (RHS) {
        for (int j = 1; j <= RHS_DEPTH; ++j) stage_fetch_rhs(rec0, sm, H - j, rank, NT);
        cp_async_wait<RHS_DEPTH - 1>();
    } else {
        for (int j = 1; j <= BWD_DEPTH; ++j) stage_fetch_bwd(rec0, sm, H - j, rank, NT);
        cp_async_wait<BWD_DEPTH - 1>();
    }
    gsync<NT>();
    for (int k = H - 1; k >= 0; --k) {
        const KindTab& T = tabs[stage_kind(k, N, H)];
        const int n = T.n, nn = T.nn, nxr = T.nxr;
        // staged block of this stage; in the right-hand-side sweep the slot starts at AB and carries the stored factors
        const double* sb = RHS ? sm.ring + (k % RHS_DEPTH) * RHS_SLOT - O_AB : sm.ring + (k % BWD_DEPTH) * BWD_SLOT;
        const double* sf_k = sm.ring + (k % RHS_DEPTH) * RHS_SLOT + RHS_A;   // RHS: KM | PV | PM of this stage
        double* rec = rec0 + k * WS_PER_STAGE;
        {   // ---- assemble G (stage Hessian over [x^ | p | u] + closed-form parts), M, the linear term g
            const double* ob = sb + O_OBS;
            const double cc0 = 2.0 * sf * ob[6], cc1 = 2.0 * sf * ob[7], eyey = sf * ob[5];
            for (int e = rank; !RHS && e < T.ng; e += NT) {
                const int src = T.gsrc[e], fl = T.gflag[e];
                double v = src >= 0 ? sb[src] : 0.0;
                if (fl) {
                    if (fl & GF_EY) v += eyey;
                    if (fl & GF_P0) v += cc0;
                    if (fl & GF_M0) v -= cc0;
                    if (fl & GF_P1) v += cc1;
                    if (fl & GF_M1) v -= cc1;
                    if (fl & GF_DELTA) v += delta;
                }
                sm.G[T.gdst[e]] = v;
            }
            for (int e = rank; e < T.nm; e += NT) {
                const int src = T.msrc[e];
                sm.M[T.mdst[e]] = src >= 0 ? sb[O_AB + src] : (src == -2 ? 1.0 : 0.0);
            }
            // linear term of active variable a on lane a of the LAST warp of the group
            const int a = rank - (NT - 32);
            if (a >= 0 && a < nxr + 2) {
                const int yi = yidx(T.st != 0, a);
                double f = sb[O_RD + yi] - mu * sb[O_V1 + yi] + sb[O_V2 + yi];
                if (soc) f += sb[O_AUX + 2 + a];   // + Jc^T sigma (c_soc - (c - s)), prepared by soc_residual
                if (yi == (T.st ? IEY : QEY)) f += sf * ob[4] * sb[O_AUX];            // obstacle cross term (s, ey) x ds
                if (T.has_next) f += sb[O_AB + T.trow * 8 + a] * sb[O_AUX + 1];       // d t_{k+1} / d a x cost-to-go of t
                sm.g[a < nxr ? a : n + (a - nxr)] = f;
            } else if (a >= nxr + 2 && a < n + 2) sm.g[a - 2] = 0.0;                  // lifted variables: no linear term
        }
        gsync<NT>();
        if (RHS) {
            // ---- right-hand-side-only sweep on the stored factors: g~ = g + M^T (P r + p), kff = -R^-1 g~_u, p = g~_x + K^T g~_u
            if (T.has_next) {
                const int dnn = T.dnn;
                const int q = rank - (NT - 16);
                if (q >= 0 && q < dnn) {
                    double a = sm.p[q];
                    for (int l = 0; l < dnn; ++l) a += sm.P[q * 10 + l] * sm.r[l];
                    sm.tv[q] = a;
                }
                gsync<NT>();
                if (rank < nn) {
                    double a = 0.0;
                    for (int l = 0; l < dnn; ++l) a += sm.M[l * 12 + rank] * sm.tv[l];
                    sm.g[rank] += a;
                }
                gsync<NT>();
            }
            const double gu0 = sm.g[n], gu1 = sm.g[n + 1];
            const double i00 = sf_k[22], i01 = sf_k[23], i11 = sf_k[9];
            if (rank == 31) { rec[O_KM + 20] = -(i00 * gu0 + i01 * gu1); rec[O_KM + 21] = -(i01 * gu0 + i11 * gu1); }
            const int q = rank - (NT - 16);
            if (q >= 0 && q < n) {
                const double v = sm.g[q] + sf_k[q] * gu0 + sf_k[10 + q] * gu1;
                sm.p[q] = v;
                rec[O_PV + q] = v;
            }
            for (int e = rank; e < n * 10; e += NT) sm.P[e] = sf_k[34 + e];   // this stage's cost-to-go Hessian, for stage k - 1
            const int q2 = rank - (NT - 32);
            if (q2 >= 0 && q2 < 12) sm.r[q2] = q2 < nxr ? -(soc ? rec[o_gres + yidx(T.st != 0, q2)] : sb[O_GRES + yidx(T.st != 0, q2)]) : 0.0;
            cp_async_wait<RHS_DEPTH - 2>();
            gsync<NT>();
            stage_fetch_rhs(rec0, sm, k - RHS_DEPTH, rank, NT);
            continue;
        }
        if (T.has_next) {
            // G += M^T P M, g += M^T (P r + p): register-blocked for the stage shapes of the shipped controllers
            const int dnn = T.dnn;
            const bool blocked = (dnn == 7 && (nn == 9 || nn == 8)) || (dnn == 5 && (nn == 7 || nn == 9));
            if (NT != 32) stage_products_wide<NT>(sm, rank, dnn, nn, blocked);
            else if (dnn == 7 && nn == 9) stage_products<7, 9>(sm, rank);
            else if (dnn == 5 && nn == 7) stage_products<5, 7>(sm, rank);
            else if (dnn == 5 && nn == 9) stage_products<5, 9>(sm, rank);
            else if (dnn == 7 && nn == 8) stage_products<7, 8>(sm, rank);
            else stage_products_generic(sm, rank, dnn, nn);
        }
        // ---- reduced Hessian of the two controls, feedback gains, cost-to-go of this stage (one phase: every thread
        // derives the gain columns it needs from G itself)
        const double ra = sm.G[n * 12 + n], rb = 0.5 * (sm.G[n * 12 + n + 1] + sm.G[(n + 1) * 12 + n]),
                     rdd = sm.G[(n + 1) * 12 + n + 1];
        const double det = ra * rdd - rb * rb;
        if (!(ra > 0.0) || !(det > 0.0) || !isfinite(det)) { ok = false; break; }
        const double idet = 1.0 / det;
        const double i00 = rdd * idet, i01 = -rb * idet, i11 = ra * idet;
        const double gu0 = sm.g[n], gu1 = sm.g[n + 1];
        const double f0 = -(i00 * gu0 + i01 * gu1), f1 = -(i01 * gu0 + i11 * gu1);   // feed-forward
        double* km = rec + O_KM;
        if (rank < n) {
            const double g0 = sm.G[n * 12 + rank], g1 = sm.G[(n + 1) * 12 + rank];
            km[rank] = -(i00 * g0 + i01 * g1); km[10 + rank] = -(i01 * g0 + i11 * g1);
        }
        if (rank == 31) { km[20] = f0; km[21] = f1; km[22] = i00; km[23] = i01; km[9] = i11; }   // n <= 8: slot 9 is free
        double* pm = rec + O_PM;
        double* pv = rec + O_PV;
        const int rcp_n = 65536 / n + 1;
        for (int e = rank; e < n * n; e += NT) {               // P = Gxx + Gxu K   (symmetrised)
            const int i = (e * rcp_n) >> 16, j = e - i * n;
            const double gi0 = sm.G[n * 12 + i], gi1 = sm.G[(n + 1) * 12 + i], gj0 = sm.G[n * 12 + j], gj1 = sm.G[(n + 1) * 12 + j];
            const double ki0 = -(i00 * gi0 + i01 * gi1), ki1 = -(i01 * gi0 + i11 * gi1);
            const double kj0 = -(i00 * gj0 + i01 * gj1), kj1 = -(i01 * gj0 + i11 * gj1);
            const double a = sm.G[i * 12 + j] + sm.G[i * 12 + n] * kj0 + sm.G[i * 12 + n + 1] * kj1;
            const double c = sm.G[j * 12 + i] + sm.G[j * 12 + n] * ki0 + sm.G[j * 12 + n + 1] * ki1;
            const double v = 0.5 * (a + c);
            sm.P[i * 10 + j] = v;
            pm[i * 10 + j] = v;
        }
        {
            const int q = rank - (NT - 16);                    // last half-warp of the group
            if (q >= 0 && q < n) {
                const double v = sm.g[q] + sm.G[q * 12 + n] * f0 + sm.G[q * 12 + n + 1] * f1;
                sm.p[q] = v;
                pv[q] = v;
            }
            // the equality residual of THIS stage's state is the offset of the transition into it (used at stage k-1)
            const int q2 = rank - (NT - 32);
            if (q2 >= 0 && q2 < 12) sm.r[q2] = q2 < nxr ? -(soc ? rec[o_gres + yidx(T.st != 0, q2)] : sb[O_GRES + yidx(T.st != 0, q2)]) : 0.0;
        }
        // the slot of stage k is free after the barrier: queue stage k - BWD_DEPTH; stage k - 1 must have landed
        cp_async_wait<BWD_DEPTH - 2>();
        gsync<NT>();
        stage_fetch_bwd(rec0, sm, k - BWD_DEPTH, rank, NT);
    }
    cp_async_wait<0>();
    gsync<NT>();
    if (!ok) return false;
    // ---- forward substitution (stage blocks AB | GRES KM PV PM streamed ahead); x ping-pongs between two buffers
    for (int j = 0; j < FWD_DEPTH; ++j) stage_fetch_fwd(rec0, sm, j, H, rank, NT);
    cp_async_wait<FWD_DEPTH - 2>();          // stages 0 and 1
    gsync<NT>();
    double* xc = sm.x;
    double* xn = sm.tv;
    {
        const KindTab& T0 = tabs[stage_kind(0, N, H)];
        if (rank < 12) xc[rank] = rank < T0.nxr ? -(soc ? rec0[o_gres + yidx(T0.st != 0, rank)] : sm.ring[F_GRES + yidx(T0.st != 0, rank)]) : 0.0;
    }
    gsync<NT>();
    for (int k = 0; k < H; ++k) {
        const KindTab& T = tabs[stage_kind(k, N, H)];
        const int n = T.n, nxr = T.nxr;
        const double* fb = sm.ring + (k % FWD_DEPTH) * FWD_SLOT;
        double* rec = rec0 + k * WS_PER_STAGE;
        // controls u = K x + kff: every thread that needs them computes them (two short dot products, no barrier)
        double u0, u1;
        {
            const double* km = fb + F_KM;
            u0 = km[20]; u1 = km[21];
            for (int l = 0; l < n; ++l) { const double xl = xc[l]; u0 += km[l] * xl; u1 += km[10 + l] * xl; }
        }
        double* dy = rec + o_dy;
        const bool st = T.st != 0;
        if (rank < nxr) dy[yidx(st, rank)] = xc[rank];
        if (rank == 30) dy[st ? 8 : 5] = u0;
        if (rank == 31) dy[st ? 9 : 6] = u1;
        {
            const int q = rank - (NT - 16);                    // multiplier direction on the last half-warp
            if (q >= 0 && q < nxr) {
                const double* pm = fb + F_PM;
                double a = fb[F_PV + q];
                for (int l = 0; l < n; ++l) a += pm[q * 10 + l] * xc[l];
                rec[o_dlam + yidx(st, q)] = -a;
            }
        }
        if (T.has_next) {
            const bool nst = T.nst != 0;
            const int dnxr = T.dnxr, dnn = T.dnn;
            const double* fn = sm.ring + ((k + 1) % FWD_DEPTH) * FWD_SLOT;
            if (rank < dnn) {
                double a;
                if (rank < dnxr || (rank == dnxr + 1 && st && !nst)) {
                    // next reduced state component (or, at the switch, the lifted T): row of [A B] times (x^, u)
                    const int row = rank < dnxr ? yidx(nst, rank) : 5;
                    a = rank < dnxr ? -(soc ? rec[WS_PER_STAGE + o_gres + row] : fn[F_GRES + row]) : 0.0;
                    const double* abr = fb + F_AB + row * 8;
                    for (int l = 0; l < nxr; ++l) a += abr[l] * xc[l];
                    a += abr[nxr] * u0 + abr[nxr + 1] * u1;
                } else a = rank == dnxr ? u0 : u1;             // lifted copies of the controls
                xn[rank] = a;
            }
        }
        cp_async_wait<FWD_DEPTH - 3>();      // stage k + 2 has landed (queued two stages ago)
        gsync<NT>();
        stage_fetch_fwd(rec0, sm, k + FWD_DEPTH, H, rank, NT);   // into the slot of stage k, free after the barrier
        double* t = xc; xc = xn; xn = t;
    }
    cp_async_wait<0>();
    gsync<NT>();
    // ---- post-pass (first warp, lanes over stages): direction of t = running sum of its linearised increments;
    // multiplier direction of s = running sum (from the end) of its stationarity row
    if (rank < 32) {
        double carry = 0.0;
        for (int base = 0; base < H; base += 32) {
            const int k = base + rank;
            double v = 0.0;
            if (k < H) {
                const double* rec = rec0 + k * WS_PER_STAGE;
                v = -rec[o_gres + (k < N ? IT : QT)];
                if (k >= 1) {
                    const bool pst = k - 1 < N;
                    const double* ab = rec - WS_PER_STAGE + O_AB + (k < N ? IT : QT) * 8;
                    const double* dyp = rec - WS_PER_STAGE + o_dy;
                    for (int a = 0; a < (pst ? 8 : 5); ++a) v += ab[a] * dyp[yidx(pst, a)];
                }
            }
            v = warp_scan_up(v, rank) + carry;
            if (k < H) rec0[k * WS_PER_STAGE + o_dy + (k < N ? IT : QT)] = v;
            carry = __shfl_sync(FULL, v, 31);
        }
        carry = 0.0;
        for (int base = ((H - 1) / 32) * 32; base >= 0; base -= 32) {
            const int k = base + rank;
            double v = 0.0;
            if (k < H) {
                const double* rec = rec0 + k * WS_PER_STAGE;
                const double* ob = rec + O_OBS;
                const double* dyk = rec + o_dy;
                v = rec[O_RD + (k < N ? IS : QS)] + sf * (ob[3] * dyk[k < N ? IS : QS] + ob[4] * dyk[k < N ? IEY : QEY]);
            }
            v = warp_scan_down(v, rank) + carry;
            if (k < H) rec0[k * WS_PER_STAGE + o_dlam + (k < N ? IS : QS)] = -v;
            carry = __shfl_sync(FULL, v, 0);
        }
    }
    gsync<NT>();
    return true;
}

// ---- helper threads of a block-wide group: warp 0 runs the (scalar) interior-point logic of its instance and calls
// the others in for the factorisations.  Commands go through shared memory, hand-over by the named barrier.
constexpr int WIDE_NT = 128;
enum GroupOp { OP_EXIT = 0, OP_RICCATI = 1, OP_RICCATI_RHS = 2 };
struct GroupCmd { int op, soc; double mu, delta; };
struct Group { int rank, nt; GroupCmd* cmd; const KindTab* tabs; };   // cmd == nullptr for a warp group

__device__ __forceinline__ bool riccati_call(const Prob& pr, const Ws& ws, WarpSm& sm, int b, const Group& g, double mu, double delta, bool soc) {
    if (g.nt == 32) return riccati<32, false>(pr, ws, sm, g.tabs, b, g.rank, mu, delta, soc);
    if (g.rank == 0) { g.cmd->op = OP_RICCATI; g.cmd->soc = soc; g.cmd->mu = mu; g.cmd->delta = delta; }
    gsync<WIDE_NT>();   // releases the helpers (group_helpers)
    return riccati<WIDE_NT, false>(pr, ws, sm, g.tabs, b, g.rank, mu, delta, soc);
}
// new right-hand side (second-order correction) on the factors the last riccati_call of this iterate stored
__device__ __forceinline__ bool riccati_rhs_call(const Prob& pr, const Ws& ws, WarpSm& sm, int b, const Group& g, double mu, bool soc) {
    if (g.nt == 32) return riccati<32, true>(pr, ws, sm, g.tabs, b, g.rank, mu, 0.0, soc);
    if (g.rank == 0) { g.cmd->op = OP_RICCATI_RHS; g.cmd->soc = soc; g.cmd->mu = mu; g.cmd->delta = 0.0; }
    gsync<WIDE_NT>();
    return riccati<WIDE_NT, true>(pr, ws, sm, g.tabs, b, g.rank, mu, 0.0, soc);
}
__device__ __forceinline__ void group_helpers(const Prob& pr, const Ws& ws, WarpSm& sm, int b, const Group& g) {
    for (;;) {
        gsync<WIDE_NT>();
        const GroupCmd c = *g.cmd;
        if (c.op == OP_EXIT) return;
        if (c.op == OP_RICCATI_RHS) riccati<WIDE_NT, true>(pr, ws, sm, g.tabs, b, g.rank, c.mu, c.delta, c.soc != 0);
        else riccati<WIDE_NT, false>(pr, ws, sm, g.tabs, b, g.rank, c.mu, c.delta, c.soc != 0);
    }
}
__device__ __forceinline__ void group_release(const Group& g) {   // called by warp 0 when its instance's step is finished
    if (g.nt == 32) return;
    if (g.rank == 0) g.cmd->op = OP_EXIT;
    gsync<WIDE_NT>();
}

// slack / multiplier directions of a computed primal direction, fraction-to-the-boundary step sizes and the
// directional derivative of the barrier objective (warp reductions)
__device__ __noinline__ void slack_directions(const Prob& pr, const Ws& ws, int b, int lane, double mu, double tau, bool soc,
                                 double& a_p, double& a_d, double& dphi) {
    const int H = pr.H, N = pr.N, M = pr.M;
    a_p = 1.0; a_d = 1.0; dphi = 0.0;
    // lanes over (stage, inequality row): N*10 SingleTrack rows, then M*2 PointMass rows
    const int n_rows = N * NI_ST + M * NI_PM;
    for (int e = lane; e < n_rows; e += 32) {
        const bool st = e < N * NI_ST;
        const int k = st ? e / NI_ST : N + (e - N * NI_ST) / NI_PM;
        const int j = st ? e - k * NI_ST : (e - N * NI_ST) - (k - N) * NI_PM;
        const double *jc = ws.at(b, W_JC, k) + j * 8, *dy = ws.at(b, soc ? W_DY2 : W_DY, k);
        const double sj = ws.at(b, W_S, k)[j], zj = ws.at(b, W_Z, k)[j];
        double a = soc ? ws.at(b, W_CS, k)[j] : ws.at(b, W_CRES, k)[j] - sj;
        if (st) {
            double jr[8], dv[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) { jr[q] = jc[q]; dv[q] = dy[c_act_st[q]]; }
#pragma unroll
            for (int q = 0; q < 8; ++q) a += jr[q] * dv[q];
        } else {
            double jr[5], dv[5];
#pragma unroll
            for (int q = 0; q < 5; ++q) { jr[q] = jc[q]; dv[q] = dy[c_act_pm[q]]; }
#pragma unroll
            for (int q = 0; q < 5; ++q) a += jr[q] * dv[q];
        }
        const double dzj = mu / sj - zj - (zj / sj) * a;
        ws.at(b, soc ? W_DSL2 : W_DSL, k)[j] = a;
        ws.at(b, soc ? W_DZ2 : W_DZ, k)[j] = dzj;
        if (a < 0.0) a_p = fmin(a_p, -tau * sj / a);
        if (dzj < 0.0) a_d = fmin(a_d, -tau * zj / dzj);
        dphi -= mu * a / sj;
    }
    // objective part of the directional derivative: lanes over (stage, variable)
    for (int e = lane; e < H * 10; e += 32) {
        const int k = e / 10, i = e - k * 10;
        if (i < (k < N ? 10 : 7)) dphi += ws.at(b, W_GF, k)[i] * ws.at(b, soc ? W_DY2 : W_DY, k)[i];
    }
    a_p = warp_min(a_p); a_d = warp_min(a_d); dphi = warp_sum(dphi);
}

// evaluates the trial point (y + alpha dy, s + alpha ds) with lanes over stages; optionally keeps the constraint
// values for a second-order correction.  Returns false on an evaluation error (NaN/Inf).
__device__ __noinline__ bool trial_point(const Prob& pr, const Ws& ws, int b, int lane, double alpha, bool soc, bool store,
                            double& ft, double& tht, double& lnt) {
    const int H = pr.H, N = pr.N;
    ft = 0.0; tht = 0.0; lnt = 0.0;
    const int a_dy = soc ? W_DY2 : W_DY, a_ds = soc ? W_DSL2 : W_DSL;
    for (int k = lane; k < H; k += 32) {
        double yt[10], yn[10], st_[10];
        const int nv = k < N ? 10 : 7;
        const double *y = ws.at(b, W_Y, k), *dy = ws.at(b, a_dy, k);
        for (int i = 0; i < nv; ++i) yt[i] = y[i] + alpha * dy[i];
        if (k < H - 1) {
            const double *y1 = ws.at(b, W_Y, k + 1), *dy1 = ws.at(b, a_dy, k + 1);
            for (int i = 0; i < (k + 1 < N ? 10 : 7); ++i) yn[i] = y1[i] + alpha * dy1[i];
        }
        const double *s = ws.at(b, W_S, k), *dsl = ws.at(b, a_ds, k);
        for (int j = 0; j < (k < N ? NI_ST : NI_PM); ++j) st_[j] = s[j] + alpha * dsl[j];
        const StageVal v = stage_trial(pr, b, k, yt, yn, st_, store ? ws.at(b, W_GT, 0) : nullptr,
                                       store && k < H - 1 ? ws.at(b, W_GT, k + 1) : nullptr, store ? ws.at(b, W_CT, k) : nullptr);
        if (store)   // c(x_trial) - s_trial
            for (int j = 0; j < (k < N ? NI_ST : NI_PM); ++j) ws.at(b, W_CT, k)[j] -= st_[j];
        ft += v.f; tht += v.theta; lnt += v.lns;
    }
    ft = warp_sum(ft); tht = warp_sum(tht); lnt = warp_sum(lnt);
    __syncwarp();
    return isfinite(ft) && isfinite(tht) && isfinite(lnt);
}

// c_soc <- alpha * (first ? c(x_k) : c_soc) + c(trial)      for the equality rows and the (c - s) rows
__device__ __noinline__ void soc_residual(const Prob& pr, const Ws& ws, int b, int lane, double alpha, bool /*soc*/, bool first) {
    for (int k = lane; k < pr.H; k += 32) {
        const bool st = k < pr.N;
        double *gs = ws.at(b, W_GS, k), *cs = ws.at(b, W_CS, k);
        const double *gt = ws.at(b, W_GT, k), *ct = ws.at(b, W_CT, k), *gr = ws.at(b, W_GRES, k);
        const double *cr = ws.at(b, W_CRES, k), *s = ws.at(b, W_S, k);
        for (int i = 0; i < (st ? 8 : 5); ++i) gs[i] = alpha * (first ? gr[i] : gs[i]) + gt[i];
        for (int j = 0; j < (st ? NI_ST : NI_PM); ++j) cs[j] = alpha * (first ? cr[j] - s[j] : cs[j]) + ct[j];
        // linear-term correction of the corrected system, per active variable: Jc^T sigma (c_soc - (c - s))
        const double *jc = ws.at(b, W_JC, k), *z = ws.at(b, W_Z, k);
        double v3[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int j = 0; j < (st ? NI_ST : NI_PM); ++j) {
            const double w = (z[j] / s[j]) * (cs[j] - (cr[j] - s[j]));
            for (int a = 0; a < (st ? 8 : 5); ++a) v3[a] += jc[j * 8 + a] * w;
        }
        double* aux = ws.at(b, W_AUX, k);
        for (int a = 0; a < 8; ++a) aux[2 + a] = v3[a];
    }
    __syncwarp();
}

// restoration substitute (oracle/nlp.py NLP.repair): guess-independent, dynamically feasible re-initialisation
// -- hold the speed (Fx = drag), no steering rate, PointMass Fy follows the centre-line curvature -- with the
// states re-generated by rolling the models out from x0
__device__ __noinline__ void repair_rollout(const Prob& pr, const Ws& ws, int b, bool neutral = true) {
    const int H = pr.H, N = pr.N;
    const double* kap = pr.kappa + (size_t)b * H;
    const double* dsb = pr.ds + (size_t)b * H;
    double* y = ws.at(b, W_Y, 0);
    for (int i = 0; i < 8; ++i) y[i] = pr.x0[(size_t)b * 8 + i];
    for (int k = 0; k < N; ++k) {
        y = ws.at(b, W_Y, k);
        if (neutral) { y[YFX] = pr.p.Frr + pr.p.Cd * y[IUX] * y[IUX]; y[YW] = 0.0; }
        if (k < N - 1) st_rk4<true>(pr.p, y, y + 8, kap[k], dsb[k], ws.at(b, W_Y, k + 1));
    }
    if (pr.M > 0) {
        double T;
        st_switch(pr.p, ws.at(b, W_Y, N - 1), ws.at(b, W_Y, N), T);
        for (int k = N; k < H; ++k) {
            y = ws.at(b, W_Y, k);
            if (neutral) { y[QFX] = pr.p.Frr + pr.p.Cd * y[QV] * y[QV]; y[QFY] = pr.p.m * y[QV] * y[QV] * kap[k]; }
            if (k < H - 1) pm_euler<true>(pr.p, y, y + 5, kap[k], dsb[k], ws.at(b, W_Y, k + 1));
        }
    }
    // the obstacle term has a pole at dist = radius + 0.1 and is negative inside it: move guessed positions that
    // fall inside an obstacle (plus a margin) sideways out of it (towards the side they are on; dead centre -> right)
    if (pr.c.obstacles && pr.n_obs > 0) {
        const double* ob = pr.obst + (size_t)b * pr.n_obs * 3;
        for (int k = 1; k < H; ++k) {
            y = ws.at(b, W_Y, k);
            const int si = k < N ? IS : QS, eyi = k < N ? IEY : QEY;
            for (int o = 0; o < pr.n_obs; ++o) {
                const double R = ob[3 * o + 2] + 0.1 + 0.5, a = y[si] - ob[3 * o], e = y[eyi] - ob[3 * o + 1];
                if (a * a + e * e < R * R) y[eyi] = ob[3 * o + 1] + (e > 1e-9 ? 1.0 : -1.0) * sqrt(R * R - a * a);   // |e| <= 1e-9: dead centre
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// one interior-point iteration of an unfinished instance, in two halves (k_step_a: test, barrier update, factorisation,
// directions; k_step_b: line search).  `lane` = lane of the warp that runs the scalar logic; g = its group.
__device__ void step_part_a(const Prob& pr, const Ws& ws, WarpSm& sm, int b, int lane, const Group& g) {
    double* sc = ws.scal(b);
    const int mode = (int)sc[S_MODE];
    if (mode == MODE_DONE) return;
    const int N = pr.N, M = pr.M;
    const Cfg& c = pr.c;
    if (lane == 0) sc[S_PHASE] = PH_NONE;
    const int rank_for_tick = lane; (void)rank_for_tick;
    VC_TICK(0);
    const Measures m = finish_gradients(pr, ws, b, lane);
    VC_TICK(1);
    const bool finite = isfinite(m.f) && isfinite(m.theta) && isfinite(m.rdmax) && isfinite(m.prim);
    double mu = sc[S_MU];
    const double sf = sc[S_SF];
    if (mode == MODE_INIT || mode == MODE_REINIT) {
        if (!finite) {   // invalid starting point (NaN / inside an obstacle pole): restoration substitute, else give up
            const int restarts = (int)sc[S_RESTARTS];
            if (restarts >= 2) { finalize(pr, ws, b, lane, VC_EVAL_ERROR, m.f, INFINITY); return; }
            if (lane == 0) sc[S_PHASE] = PH_BADINIT;   // the rollout runs in k_step_b
            return;
        }
        if (mode == MODE_INIT && c.obj_scaling && m.gfmax > IpmConst::max_gradient && lane == 0)
            sc[S_SF] = IpmConst::max_gradient / m.gfmax;   // gradient-based objective scaling at the initial point
        init_slacks(pr, ws, b, lane, mu);
        if (lane == 0) { sc[S_MODE] = MODE_STEP; sc[S_NFILT] = 0; sc[S_THETA0] = -1.0; sc[S_ACC] = 0; }
        return;
    }
    // ---- convergence test (IPOPT overall error E_0) and barrier update
    const int me = 8 + 8 * (N - 1) + (M > 0 ? 5 * M : 0), mi = NI_ST * N + NI_PM * M;
    const double sd = fmax(IpmConst::s_max, (m.sumlam + m.sumz) / (me + mi)) / IpmConst::s_max;
    const double scs = fmax(IpmConst::s_max, m.sumz / mi) / IpmConst::s_max;
    const double e_dp = fmax(m.rdmax / sd, m.prim);
    const double err0 = fmax(e_dp, compl_error(pr, ws, b, lane, 0.0) / scs);
    const double obj = m.f;
    int iters = (int)sc[S_ITERS];
    if (!finite) { finalize(pr, ws, b, lane, VC_EVAL_ERROR, obj, INFINITY); return; }
    if (err0 <= c.tol) { finalize(pr, ws, b, lane, VC_SOLVED, obj, err0); return; }
    int acc = (err0 <= c.acceptable_tol) ? (int)sc[S_ACC] + 1 : 0;
    if (acc >= c.acceptable_iter) { finalize(pr, ws, b, lane, VC_ACCEPTABLE, obj, err0); return; }
    if (iters >= c.max_iter) { finalize(pr, ws, b, lane, VC_MAX_ITER, obj, err0); return; }
    int nfilt = (int)sc[S_NFILT];
    const double mu_floor = c.tol / 11.0;
    while (mu > mu_floor * (1 + 1e-12) &&
           fmax(e_dp, compl_error(pr, ws, b, lane, mu) / scs) <= IpmConst::kappa_eps * mu) {
        mu = fmax(mu_floor, fmin(IpmConst::kappa_mu * mu, pow(mu, IpmConst::theta_mu)));
        nfilt = 0;
    }
    const double tau = fmax(IpmConst::tau_min, 1.0 - mu);
    const double theta = m.theta, phi = sf * m.f - mu * m.lns;
    double theta0 = sc[S_THETA0];
    if (theta0 < 0.0) theta0 = theta;
    const double theta_min = 1e-4 * fmax(1.0, theta0), theta_max = 1e4 * fmax(1.0, theta0);
    const double delta_last = sc[S_DELTA_LAST];
    const bool first_ever = delta_last == 0.0;
    double trial = sc[S_DAMP_TRIAL], a_p, a_d, dphi;   // > 0: damped re-solve of this iterate's direction (see k_step_b)
    bool failed = false;
    VC_TICK(2);
    // ---- inertia-correcting factorisation (IPOPT delta_w ladder) and the slack / multiplier directions
    while (!riccati_call(pr, ws, sm, b, g, mu, trial, false)) {
        if (trial == 0.0) trial = first_ever ? 1e-4 : fmax(1e-20, delta_last / 3.0);
        else trial *= first_ever ? 100.0 : 8.0;
        if (trial > 1e40) { failed = true; break; }
    }
    VC_TICK(4);
    if (!failed) slack_directions(pr, ws, b, lane, mu, tau, false, a_p, a_d, dphi);
    VC_TICK(5);
    if (lane == 0) {
        sc[S_PHASE] = failed ? PH_FAILED : PH_SEARCH;
        sc[S_MU_NEW] = mu; sc[S_THETA] = theta; sc[S_PHI] = phi; sc[S_THETA0_NEW] = theta0; sc[S_NFILT_NEW] = nfilt;
        sc[S_TRIAL] = trial; sc[S_ACC_NEW] = acc; sc[S_OBJ_NEW] = obj; sc[S_ERR_NEW] = err0;
        if (!failed) { sc[S_AP] = a_p; sc[S_AD] = a_d; sc[S_DPHI] = dphi; }
    }
    (void)theta_min; (void)theta_max; (void)iters;
}

// results of the backtracking trial points alpha = a_p 2^-bt evaluated ahead of time by the other warps of a
// k_step_b_wide block (nullptr: evaluate them one after the other)
struct TrialCache { double ok, f, theta, lns; };

__device__ void step_part_b(const Prob& pr, const Ws& ws, WarpSm& sm, int b, int lane, const Group& g, const TrialCache* cache = nullptr) {
    double* sc = ws.scal(b);
    const int mode = (int)sc[S_MODE];
    if (mode == MODE_DONE) return;
    const int H = pr.H, N = pr.N;
    const Cfg& c = pr.c;
    const int phase = (int)sc[S_PHASE];
    if (phase == PH_NONE) return;
    if (phase == PH_BADINIT) {   // invalid starting point (NaN / inside an obstacle pole): restoration substitute
        // (an invalid FIRST point has no gradient to scale by: the repaired point then sets the objective scaling)
        if (lane == 0) { repair_rollout(pr, ws, b); sc[S_RESTARTS] += 1.0; }
        return;
    }
    const double mu = sc[S_MU_NEW], sf = sc[S_SF];
    const int iters = (int)sc[S_ITERS];
    const double obj = sc[S_OBJ_NEW], err0 = sc[S_ERR_NEW];
    const int acc = (int)sc[S_ACC_NEW];
    int nfilt = (int)sc[S_NFILT_NEW];
    const double theta = sc[S_THETA], phi = sc[S_PHI], theta0 = sc[S_THETA0_NEW];
    const double tau = fmax(IpmConst::tau_min, 1.0 - mu);
    const double theta_min = 1e-4 * fmax(1.0, theta0), theta_max = 1e4 * fmax(1.0, theta0);
    const double delta_last = sc[S_DELTA_LAST];
    const bool first_ever = delta_last == 0.0;
    double trial = sc[S_TRIAL], alpha = 0.0, a_d = sc[S_AD], dbg_ap = 0.0, dbg_dphi = 0.0, dbg_tht0 = 0.0, dbg_pht0 = 0.0;
    int dbg_nan = 0;
    double dbg_tht2 = 0.0, dbg_ap2 = 0.0;
    bool accepted = false, failed = phase == PH_FAILED;
    int n_damp = 0;
    double* filt_t = sc + S_FILT;
    double* filt_p = sc + S_FILP;
    bool use_soc = false;
    // acceptance of a trial point by the filter (IPOPT Alg. A, steps A-5.3/5.4) relative to a reference point (the
    // current iterate, or the stored one while the watchdog runs); all lanes hold the same values
    auto acceptable_ref = [&](double a_test, double tht, double pht, double dphi, bool& augment, double theta, double phi) -> bool {
        augment = false;
        if (tht > theta_max) return false;
        for (int i = 0; i < nfilt; ++i)
            if (tht >= filt_t[i] && pht >= filt_p[i]) return false;
        const bool switching = theta <= theta_min && dphi < 0.0 &&
                               a_test * pow(-dphi, IpmConst::s_phi) > IpmConst::delta * pow(theta, IpmConst::s_theta);
        if (switching) {
            if (pht <= phi + IpmConst::eta_phi * a_test * dphi) return true;
        } else if (tht <= (1 - IpmConst::gamma_theta) * theta || pht <= phi - IpmConst::gamma_phi * theta) {
            augment = true;
            return true;
        }
        // steps at the rounding floor: both measures unchanged
        return fabs(pht - phi) <= 1e-13 * fmax(1.0, fabs(phi)) && tht <= theta + 1e-13;
    };
    auto acceptable = [&](double a_test, double tht, double pht, double dphi, bool& augment) -> bool {
        return acceptable_ref(a_test, tht, pht, dphi, augment, theta, phi);
    };
    // ---- IPOPT's watchdog (IpBacktrackingLineSearch StartWatchDog / StopWatchDog): after `watchdog_trigger` consecutive
    // shortened steps the full step is taken for up to `watchdog_trials` iterations without asking the filter, the
    // acceptance test being made against the stored reference point; if no acceptable point turns up, the stored iterate
    // is restored and the regular backtracking follows (first trial point skipped)
    bool in_wd = sc[S_WD_IN] != 0.0;
    const bool skip_first = sc[S_SKIP_FIRST] != 0.0;
    int wd_short = (int)sc[S_WD_SHORT];
    int n_steps = 0;
    bool forced = false;
    if (!failed && !in_wd && pr.watchdog_trigger > 0 && wd_short >= pr.watchdog_trigger && !skip_first && sc[S_DAMP_TRIAL] == 0.0) {
        in_wd = true;
        for (int e = lane; e < H * 38; e += 32) {           // store the reference iterate
            const int k = e / 38, i = e - k * 38;
            double* r = ws.rec(b, k);
            r[O_BAK + i] = r[O_IT + i];
        }
        if (lane == 0) {
            sc[S_WD_IN] = 1.0; sc[S_WD_TRIAL] = 0.0; sc[S_WD_THETA] = theta; sc[S_WD_PHI] = phi; sc[S_WD_DPHI] = sc[S_DPHI];
            sc[S_WD_ATEST] = sc[S_AP];
        }
        __syncwarp();
    }
    if (!failed && in_wd) {
        const double a_p = sc[S_AP];
        double ft, tht, lnt;
        bool okv;
        if (cache) { okv = cache[0].ok != 0.0; ft = cache[0].f; tht = cache[0].theta; lnt = cache[0].lns; }
        else okv = trial_point(pr, ws, b, lane, a_p, false, true, ft, tht, lnt);
        bool augment = false;
        const double th_ref = sc[S_WD_THETA], ph_ref = sc[S_WD_PHI];
        const bool acc_wd = okv && acceptable_ref(sc[S_WD_ATEST], tht, sf * ft - mu * lnt, sc[S_WD_DPHI], augment, th_ref, ph_ref);
        const int wd_trial = (int)sc[S_WD_TRIAL] + 1;
        __syncwarp();
        if (acc_wd) {
            if (augment && nfilt < FILT_MAX) {
                if (lane == 0) { filt_t[nfilt] = (1 - IpmConst::gamma_theta) * th_ref; filt_p[nfilt] = ph_ref - IpmConst::gamma_phi * th_ref; }
                ++nfilt;
            }
            if (lane == 0) sc[S_WD_IN] = 0.0;
            accepted = true; alpha = a_p; dbg_ap = a_p;
        } else if (!okv || wd_trial > pr.watchdog_trials) {
            for (int e = lane; e < H * 38; e += 32) {       // back to the stored iterate; the next pass backtracks from it
                const int k = e / 38, i = e - k * 38;
                double* r = ws.rec(b, k);
                r[O_IT + i] = r[O_BAK + i];
            }
            if (lane == 0) { sc[S_WD_IN] = 0.0; sc[S_WD_SHORT] = 0.0; sc[S_SKIP_FIRST] = 1.0; sc[S_MU] = mu; sc[S_NFILT] = nfilt; }
            return;
        } else {
            if (lane == 0) sc[S_WD_TRIAL] = wd_trial;
            accepted = true; forced = true; alpha = a_p; dbg_ap = a_p;   // full step taken on trust
        }
    }
    while (!accepted && !failed) {   // one direction per pass (computed by k_step_a)
        const double a_p = sc[S_AP], dphi = sc[S_DPHI];
        dbg_ap = a_p; dbg_dphi = dphi;
        // ---- backtracking filter line search, lanes over stages
        alpha = skip_first ? 0.5 * a_p : a_p;
        for (int bt = skip_first ? 1 : 0; bt < 12 && !accepted; ++bt, alpha *= 0.5) {
            n_steps = bt;
            double ft, tht, lnt;
            bool okv;
            if (cache) { okv = cache[bt].ok != 0.0; ft = cache[bt].f; tht = cache[bt].theta; lnt = cache[bt].lns; }
            else okv = trial_point(pr, ws, b, lane, alpha, false, bt == 0, ft, tht, lnt);
            if (!okv) ++dbg_nan;
            if (bt == 0) { dbg_tht0 = tht; dbg_pht0 = sf * ft - mu * lnt; }
            bool augment = false;
            if (okv && acceptable(alpha, tht, sf * ft - mu * lnt, dphi, augment)) accepted = true;
            // ---- second-order correction (IPOPT Alg. A-5.5 .. A-5.9): first trial step rejected and infeasibility
            // not reduced -> re-solve with c_soc = alpha c(x) + c(x + alpha d), at most 4 times
            if (!accepted && bt == 0 && okv && tht >= theta && pr.max_soc > 0) {
                soc_residual(pr, ws, b, lane, alpha, false, true);
                double theta_soc_old = tht;   // IPOPT compares every correction with the previous TRIAL point's violation
                for (int ps = 0; ps < pr.max_soc; ++ps) {
                    if (!riccati_rhs_call(pr, ws, sm, b, g, mu, true)) break;
                    double a_p2, a_d2, dphi2, ft2, tht2, lnt2;
                    slack_directions(pr, ws, b, lane, mu, tau, true, a_p2, a_d2, dphi2);
                    const bool ok2 = trial_point(pr, ws, b, lane, a_p2, true, true, ft2, tht2, lnt2);
                    if (ps == 0) { dbg_tht2 = ok2 ? tht2 : -1.0; dbg_ap2 = a_p2; }
                    if (ok2 && acceptable(alpha, tht2, sf * ft2 - mu * lnt2, dphi, augment)) {
                        accepted = true; use_soc = true; alpha = a_p2; a_d = a_d2; n_steps = 0;
                        break;
                    }
                    if (!ok2 || tht2 > 0.99 * theta_soc_old) break;
                    theta_soc_old = tht2;
                    soc_residual(pr, ws, b, lane, a_p2, true, false);
                }
            }
            if (accepted) {
                if (augment && nfilt < FILT_MAX) {
                    if (lane == 0) { filt_t[nfilt] = (1 - IpmConst::gamma_theta) * theta; filt_p[nfilt] = phi - IpmConst::gamma_phi * theta; }
                    ++nfilt;
                }
                break;
            }
        }
        __syncwarp();
        if (accepted) break;
        // no acceptable step along this direction: damp it (larger delta_w) and retry; IPOPT would restore here.
        // The re-solve happens in the NEXT pass (k_step_a starts its delta_w ladder at S_DAMP_TRIAL; the iterate is
        // unchanged, so the evaluation kernels reproduce the same derivatives): a pass never holds the whole batch
        // for more than one factorisation + one line search of its slowest instance
        trial = fmax(trial, 1e-4) * 100.0;
        n_damp = (int)sc[S_NDAMP] + 1;
        if (trial > 1e12 || n_damp > pr.max_damp) { failed = true; break; }
        if (lane == 0) { sc[S_DAMP_TRIAL] = trial; sc[S_NDAMP] = n_damp; sc[S_ITERS] = iters + 1; }   // a pass counts against max_iter
        return;
    }
    if (lane == 0) { sc[S_DAMP_TRIAL] = 0.0; sc[S_NDAMP] = 0.0; }
    if (failed) {
        const int restarts = (int)sc[S_RESTARTS];
        if (restarts >= 2) { finalize(pr, ws, b, lane, VC_LINESEARCH_FAIL, obj, err0); return; }
        __syncwarp();
        if (lane == 0) {
            repair_rollout(pr, ws, b);
            sc[S_RESTARTS] = restarts + 1; sc[S_MODE] = MODE_REINIT; sc[S_MU] = c.mu_init; sc[S_DELTA_LAST] = 0.0;
            sc[S_ITERS] = iters + 1;
        }
        for (int k = lane; k < H; k += 32) {   // neutral multipliers / slacks for the re-evaluation
            double *lam = ws.at(b, W_LAM, k), *s = ws.at(b, W_S, k), *z = ws.at(b, W_Z, k);
            for (int i = 0; i < 8; ++i) lam[i] = 0.0;
            for (int j = 0; j < 10; ++j) { s[j] = 1.0; z[j] = 0.0; }
        }
        return;
    }
    // ---- accept: the primal-dual update itself is streaming work (38 doubles per stage read twice, written once) and
    // runs in k_update at full occupancy; here only its parameters are handed over
    if (lane == 0) { sc[S_UPD_DIR] = use_soc ? 2.0 : 1.0; sc[S_UPD_ALPHA] = alpha; sc[S_UPD_AD] = a_d; }
    (void)forced;
    if (lane == 0) {
        sc[S_WD_SHORT] = n_steps == 0 ? 0.0 : wd_short + 1.0;
        sc[S_SKIP_FIRST] = 0.0;
        sc[S_MU] = mu; sc[S_NFILT] = nfilt; sc[S_THETA0] = theta0; sc[S_ACC] = acc; sc[S_ITERS] = iters + 1;
        if (trial > 0.0) sc[S_DELTA_LAST] = trial;
        sc[S_OBJ] = obj; sc[S_ERR] = err0;
        sc[S_DBG_THT2] = dbg_tht2; sc[S_DBG_AP2] = dbg_ap2;
        sc[S_DBG_NAN] = dbg_nan; sc[S_DBG_THT0] = dbg_tht0; sc[S_DBG_PHT0] = dbg_pht0; sc[S_DBG_PHI] = phi; sc[S_DBG_SOC] = use_soc;
        sc[S_DBG_ALPHA] = alpha; sc[S_DBG_AP] = dbg_ap; sc[S_DBG_THETA] = theta; sc[S_DBG_DPHI] = dbg_dphi;
    }
}

// the two halves of an iteration as separate kernels: the factorisation half needs half the registers of the
// line-search half (no model code), so it runs at twice the occupancy
#ifndef VC_STEPA_MINBLOCKS
#define VC_STEPA_MINBLOCKS 4
#endif
__global__ void __launch_bounds__(32 * STEP_WARPS, VC_STEPA_MINBLOCKS) k_step_a(Prob pr, Ws ws) {
    extern __shared__ __align__(16) unsigned char step_smem[];   // WarpSm[STEP_WARPS] | KindTab[NKIND] (> 48 KB at 8 warps)
    WarpSm* s_all = reinterpret_cast<WarpSm*>(step_smem);
    KindTab* s_tabs = reinterpret_cast<KindTab*>(step_smem + sizeof(WarpSm) * STEP_WARPS);
    build_kind_tables(pr, s_tabs, threadIdx.x, blockDim.x);
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int i = blockIdx.x * STEP_WARPS + wid;
    if (i >= *pr.act_count) return;
    step_part_a(pr, ws, s_all[wid], pr.act_list[i], lane, Group{lane, 32, nullptr, s_tabs});
}
__global__ void __launch_bounds__(32 * STEP_WARPS, VC_STEP_MINBLOCKS) k_step_b(Prob pr, Ws ws, int* next_list, int* next_count) {
    extern __shared__ __align__(16) unsigned char step_smem[];
    WarpSm* s_all = reinterpret_cast<WarpSm*>(step_smem);
    KindTab* s_tabs = reinterpret_cast<KindTab*>(step_smem + sizeof(WarpSm) * STEP_WARPS);
    build_kind_tables(pr, s_tabs, threadIdx.x, blockDim.x);
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int i = blockIdx.x * STEP_WARPS + wid;
    if (i >= *pr.act_count) return;
    const int b = pr.act_list[i];
    if (lane == 0) ws.scal(b)[S_UPD_DIR] = 0.0;
    __syncwarp();
    if (ws.scal(b)[S_MODE] != (double)MODE_DONE) step_part_b(pr, ws, s_all[wid], b, lane, Group{lane, 32, nullptr, s_tabs});
    __syncwarp();
    if (lane == 0 && ws.scal(b)[S_MODE] != (double)MODE_DONE) next_list[atomicAdd(next_count, 1)] = b;
}

// Late passes hold few instances on an otherwise idle GPU: what counts then is the latency of ONE iteration.  A whole
// block serves one instance: warp 0 runs the unchanged scalar logic, four warps share the Riccati sweeps (one matrix
// entry per thread), and in the line-search half warp w evaluates the trial point alpha = a_p 2^-w speculatively
// before warp 0 applies the unchanged acceptance logic to the cached values (same first acceptable alpha).  Every
// number is produced by the same arithmetic as in the warp-per-instance kernels: results are identical.
__global__ void __launch_bounds__(WIDE_NT, 4) k_step_a_wide(Prob pr, Ws ws) {
    __shared__ WarpSm s_one;
    __shared__ GroupCmd s_cmd;
    __shared__ KindTab s_tabs[NKIND];
    if ((int)blockIdx.x >= *pr.act_count) return;
    build_kind_tables(pr, s_tabs, threadIdx.x, blockDim.x);
    __syncthreads();
    const int b = pr.act_list[blockIdx.x];
    const Group g{(int)threadIdx.x, WIDE_NT, &s_cmd, s_tabs};
    if (threadIdx.x < 32) {
        step_part_a(pr, ws, s_one, b, threadIdx.x, g);
        group_release(g);
    } else {
        group_helpers(pr, ws, s_one, b, g);
    }
}
constexpr int LS_TRIALS = 12;
__global__ void __launch_bounds__(32 * LS_TRIALS, 1) k_step_b_wide(Prob pr, Ws ws, int* next_list, int* next_count) {
    __shared__ WarpSm s_one;
    __shared__ GroupCmd s_cmd;
    __shared__ KindTab s_tabs[NKIND];
    __shared__ TrialCache s_trial[LS_TRIALS];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if ((int)blockIdx.x >= *pr.act_count) return;
    build_kind_tables(pr, s_tabs, threadIdx.x, blockDim.x);
    const int b = pr.act_list[blockIdx.x];
    const Group g{(int)threadIdx.x, WIDE_NT, &s_cmd, s_tabs};
    const double* sc = ws.scal(b);
    const bool live = sc[S_MODE] != (double)MODE_DONE;
    const bool search = live && (int)sc[S_PHASE] == PH_SEARCH;
    if (search) {
        double alpha = sc[S_AP];
        for (int w = 0; w < wid; ++w) alpha *= 0.5;
        double ft, tht, lnt;
        const bool okv = trial_point(pr, ws, b, lane, alpha, false, wid == 0, ft, tht, lnt);
        if (lane == 0) s_trial[wid] = TrialCache{okv ? 1.0 : 0.0, ft, tht, lnt};
    }
    __syncthreads();
    if (threadIdx.x >= WIDE_NT) return;                 // the second-order-correction sweeps are shared by four warps
    if (wid != 0) { group_helpers(pr, ws, s_one, b, g); return; }
    if (lane == 0) ws.scal(b)[S_UPD_DIR] = 0.0;
    __syncwarp();
    if (live) step_part_b(pr, ws, s_one, b, lane, g, search ? s_trial : nullptr);
    group_release(g);
    __syncwarp();
    if (lane == 0 && ws.scal(b)[S_MODE] != (double)MODE_DONE) next_list[atomicAdd(next_count, 1)] = b;
}

// Accepted step of every instance of the pass: one warp per (instance, stage).  Lane i < 28 owns element i of the
// contiguous [y 10 | lam 8 | s 10] row, lanes 18..27 also update the multiplier z of the slack they hold, with the
// kappa_sigma safeguard of IPOPT (same arithmetic as the in-kernel update of the fused k_step).
__global__ void __launch_bounds__(128) k_update(Prob pr, Ws ws) {
    const int lane = threadIdx.x & 31;
    const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int H = pr.H;
    const int ii = (int)(w / H), k = (int)(w - (long long)ii * H);
    if (ii >= *pr.act_count) return;
    const int b = pr.act_list[ii];
    const double* sc = ws.scal(b);
    const int dir = (int)sc[S_UPD_DIR];
    if (dir == 0) return;
    const double alpha = sc[S_UPD_ALPHA], a_d = sc[S_UPD_AD], mu = sc[S_MU];
    const int odir = dir == 2 ? O_DIR2 : O_DIR;
    const bool st = k < pr.N;
    const int i = lane, j = lane - 18;
    const int lim = i < 10 ? (st ? 10 : 7) : (i < 18 ? (st ? 18 : 15) : 18 + (st ? NI_ST : NI_PM));
    if (i >= lim) return;
    double* r = ws.rec(b, k);
    const double v = r[O_IT + i] + alpha * r[odir + i];
    r[O_IT + i] = v;
    if (j >= 0) {
        double zj = r[O_IT + 28 + j] + a_d * r[odir + 28 + j];
        zj = fmin(fmax(zj, mu / (IpmConst::kappa_sigma * v)), IpmConst::kappa_sigma * mu / v);
        r[O_IT + 28 + j] = zj;
    }
}

// instances still unfinished when the pass bound is reached: written back as they are, status VC_MAX_ITER
__global__ void __launch_bounds__(128) k_finalize_survivors(Prob pr, Ws ws) {
    const int lane = threadIdx.x & 31;
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (i >= *pr.act_count) return;
    const int b = pr.act_list[i];
    const double* sc = ws.scal(b);
    if (sc[S_MODE] != (double)MODE_DONE) finalize(pr, ws, b, lane, VC_MAX_ITER, sc[S_OBJ], sc[S_ERR]);
}

// k_nmpc_eval: objective / max equality violation / min inequality margin of given trajectories (no solve)
__global__ void __launch_bounds__(128) k_nmpc_eval(Prob pr, const double* X, const double* U, double* out) {
    const int lane = threadIdx.x & 31;
    const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (b >= pr.B) return;
    const int H = pr.H, N = pr.N;
    const double* Xb = X + (size_t)b * 8 * H;
    const double* Ub = U + (size_t)b * 2 * H;
    double f = 0.0, gmax = 0.0, cmin = INFINITY;
    for (int k = lane; k < H; k += 32) {
        double y[10], yn[10], sl[10];
        for (int kk = k; kk <= k + 1 && kk < H; ++kk) {
            double* d = kk == k ? y : yn;
            const int nx = kk < N ? 8 : 5;
            for (int i = 0; i < nx; ++i) d[i] = Xb[i * H + kk];
            d[nx] = Ub[kk]; d[nx + 1] = Ub[H + kk];
        }
        for (int j = 0; j < 10; ++j) sl[j] = 1.0;
        const StageVal v = stage_trial(pr, b, k, y, yn, sl);
        f += v.f; gmax = nmax(gmax, v.gmax); cmin = fmin(cmin, v.cmin);
    }
    f = warp_sum(f); gmax = warp_nmax(gmax); cmin = warp_min(cmin);
    if (lane == 0) { out[(size_t)b * 3] = f; out[(size_t)b * 3 + 1] = gmax; out[(size_t)b * 3 + 2] = cmin; }
}

// ------------------------------------------------------------------------------------------------
// host side
constexpr size_t STEP_SMEM = sizeof(WarpSm) * STEP_WARPS + sizeof(KindTab) * NKIND;
static int step_kernels_smem_optin() {   // dynamic shared memory beyond 48 KB needs the opt-in, once per device context
    int rc = check_cuda(cudaFuncSetAttribute(k_step_a, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)STEP_SMEM), "cudaFuncSetAttribute(k_step_a)");
    if (rc) return rc;
    return check_cuda(cudaFuncSetAttribute(k_step_b, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)STEP_SMEM), "cudaFuncSetAttribute(k_step_b)");
}
static int fill_prob(vc_handle* h, Prob& pr, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                     const double* d_obst, int n_obs) {
    if (h->cfg.N < 1 || h->cfg.M < 0) { set_error("nmpc: config not set (need horizon N >= 1, M >= 0)"); return -1; }
    if (h->cfg.N + h->cfg.M > 4096) { set_error("nmpc: horizon too long"); return -1; }
    if (!d_x0 || !d_kappa || !d_ds) { set_error("nmpc: null input pointer"); return -1; }
    if (h->cfg.obstacles && n_obs > 0 && !d_obst) { set_error("nmpc: obstacles enabled but d_obst is null"); return -1; }
    pr.p = h->params; pr.c = h->cfg;
    pr.B = B; pr.N = h->cfg.N; pr.M = h->cfg.M; pr.H = pr.N + pr.M;
    pr.n_obs = (h->cfg.obstacles && d_obst) ? n_obs : 0;
    pr.max_damp = h->cfg.max_damp; pr.max_soc = h->cfg.max_soc;
    pr.watchdog_trigger = h->cfg.watchdog_trigger; pr.watchdog_trials = h->cfg.watchdog_trials;
    {   // warm-started solves may enter the monotone barrier ladder a few rungs down (option, default off)
        double mu0 = h->cfg.mu_init;
        const double mu_floor = h->cfg.tol / 11.0;
        for (int i = 0; i < h->cfg.warm_rungs; ++i) mu0 = fmax(mu_floor, fmin(IpmConst::kappa_mu * mu0, pow(mu0, IpmConst::theta_mu)));
        pr.mu_start = mu0;
    }
    pr.x0 = d_x0; pr.kappa = d_kappa; pr.ds = d_ds; pr.obst = d_obst;
    pr.X = pr.U = pr.obj = pr.kkt = nullptr; pr.status = pr.iters = nullptr;
    pr.act_list = pr.act_count = nullptr;
    return 0;
}

int launch_nmpc_eval(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                     const double* d_obst, int n_obs, const double* d_X, const double* d_U, double* d_out) {
    Prob pr;
    int rc = fill_prob(h, pr, B, d_x0, d_kappa, d_ds, d_obst, n_obs);
    if (rc) return rc;
    cudaSetDevice(h->device);
    const long long threads = (long long)B * 32;
    k_nmpc_eval<<<(unsigned)((threads + 127) / 128), 128, 0, h->stream>>>(pr, d_X, d_U, d_out);
    h->launches++;
    return check_cuda(cudaGetLastError(), "k_nmpc_eval");
}

int launch_nmpc_solve(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                      const double* d_obst, int n_obs, double* d_X, double* d_U, double* d_obj, int* d_status,
                      int* d_iters, double* d_kkt) {
    Prob pr;
    int rc = fill_prob(h, pr, B, d_x0, d_kappa, d_ds, d_obst, n_obs);
    if (rc) return rc;
    if (!d_X || !d_U) { set_error("nmpc: null X/U pointer"); return -1; }
    pr.X = d_X; pr.U = d_U; pr.obj = d_obj; pr.kkt = d_kkt; pr.status = d_status; pr.iters = d_iters;
    cudaSetDevice(h->device);
    rc = step_kernels_smem_optin();
    if (rc) return rc;
    Ws ws;
    ws.H = pr.H;
    ws.stride = (size_t)WS_PER_STAGE * pr.H + NSCAL;
    const size_t bytes = ws.stride * sizeof(double) * (size_t)B;
    if (bytes > h->ws_bytes) {
        cudaFree(h->d_ws);
        h->d_ws = nullptr; h->ws_bytes = 0;
        rc = check_cuda(cudaMalloc(&h->d_ws, bytes), "cudaMalloc(nmpc workspace)");
        if (rc) return rc;
        h->ws_bytes = bytes;
    }
    if (B > h->active_cap) {   // two ping-pong index lists [B] + two counters
        cudaFree(h->d_active);
        h->d_active = nullptr; h->active_cap = 0;
        rc = check_cuda(cudaMalloc(&h->d_active, sizeof(int) * (2 * (size_t)B + 2)), "cudaMalloc(active lists)");
        if (rc) return rc;
        h->active_cap = B;
    }
    if (!h->h_active) {
        rc = check_cuda(cudaMallocHost(&h->h_active, sizeof(int)), "cudaMallocHost(active)");
        if (rc) return rc;
    }
    int* lists[2] = {h->d_active + 2, h->d_active + 2 + h->active_cap};
    int* counts[2] = {h->d_active, h->d_active + 1};
    ws.base = (double*)h->d_ws;
    cudaStream_t st = h->stream;
    k_prepare<<<(B * pr.H + 127) / 128, 128, 0, st>>>(pr, ws, lists[0], counts[0]);
    h->launches++;
    int cur = 0, n_act = B;
    // passes that do not count as iterations: initialisation (<= 3 with re-initialisations), the final test, one per
    // watchdog restore (at most one in watchdog_trigger + 1 iterations); survivors of the bound are finalised below
    const int max_pass = h->cfg.max_iter + h->cfg.max_iter / 8 + 32;
    const bool debug = getenv("VC_DEBUG") != nullptr;
    int pass = 0;
    const bool prof = h->profiling != 0;
    const int wide_max = h->cfg.wide_max;
    const bool trace = prof && getenv("VC_PASS_TRACE") != nullptr;
    size_t ev = 0;
    auto mark = [&]() {
        if (!prof) return;
        if (ev >= h->prof_events.size()) { cudaEvent_t e; cudaEventCreate(&e); h->prof_events.push_back(e); }
        cudaEventRecord(h->prof_events[ev++], st);
    };
    for (; pass < max_pass; ++pass) {
        // grids are sized by the last count read back (an upper bound); the kernels read the exact count on the device
        const long long t_st = (long long)n_act * pr.N * EVAL_ST_LANES, t_pm = (long long)n_act * pr.M * 15;
        pr.act_list = lists[cur]; pr.act_count = counts[cur];
        if (prof) h->prof_instance_passes += n_act;
        mark();
        k_eval_st<<<(unsigned)((t_st + VC_EVAL_TPB - 1) / VC_EVAL_TPB), VC_EVAL_TPB, 0, st>>>(pr, ws);
        mark();
        if (pr.M > 0) k_eval_pm<<<(unsigned)((t_pm + 127) / 128), 128, 0, st>>>(pr, ws);
        cudaMemsetAsync(counts[1 - cur], 0, sizeof(int), st);
        mark();
        if (n_act <= wide_max) {
            k_step_a_wide<<<n_act, WIDE_NT, 0, st>>>(pr, ws);
            if (trace) mark();
            k_step_b_wide<<<n_act, 32 * LS_TRIALS, 0, st>>>(pr, ws, lists[1 - cur], counts[1 - cur]);
        } else {
            k_step_a<<<(n_act + STEP_WARPS - 1) / STEP_WARPS, 32 * STEP_WARPS, STEP_SMEM, st>>>(pr, ws);
            if (trace) mark();
            k_step_b<<<(n_act + STEP_WARPS - 1) / STEP_WARPS, 32 * STEP_WARPS, STEP_SMEM, st>>>(pr, ws, lists[1 - cur], counts[1 - cur]);
        }
        k_update<<<(unsigned)(((long long)n_act * pr.H * 32 + 127) / 128), 128, 0, st>>>(pr, ws);
        mark();
        cur = 1 - cur;
        h->launches += pr.M > 0 ? 5 : 4;
        if (debug) {   // VC_DEBUG=1: trace instance 0
            double sc[NSCAL];
            cudaMemcpyAsync(sc, ws.scal(0), sizeof(sc), cudaMemcpyDeviceToHost, st);
            cudaStreamSynchronize(st);
            fprintf(stderr, "[vc] pass %3d mode %d it %3d mu %.3e sf %.4g obj %.10f err %.3e delta %.2e nfilt %d restarts %d status %d | alpha %.3e a_p %.3e theta %.3e dphi %.3e nan %d tht0 %.3e pht0 %.6e phi %.6e soc %d tht2 %.3e ap2 %.3e\n",
                    pass, (int)sc[S_MODE], (int)sc[S_ITERS], sc[S_MU], sc[S_SF], sc[S_OBJ], sc[S_ERR], sc[S_DELTA_LAST],
                    (int)sc[S_NFILT], (int)sc[S_RESTARTS], (int)sc[S_STATUS], sc[S_DBG_ALPHA], sc[S_DBG_AP], sc[S_DBG_THETA], sc[S_DBG_DPHI], (int)sc[S_DBG_NAN], sc[S_DBG_THT0], sc[S_DBG_PHT0], sc[S_DBG_PHI], (int)sc[S_DBG_SOC], sc[S_DBG_THT2], sc[S_DBG_AP2]);
        }
        if (pass >= 3) {   // no instance can finish before its third pass (init, >= 1 step, convergence test)
            cudaMemcpyAsync(h->h_active, counts[cur], sizeof(int), cudaMemcpyDeviceToHost, st);
            rc = check_cuda(stream_wait(h), "nmpc iteration");
            if (rc) return rc;
            n_act = *h->h_active;
            if (n_act == 0) break;
        }
    }
    if (n_act > 0) {   // pass bound reached with instances left
        pr.act_list = lists[cur]; pr.act_count = counts[cur];
        k_finalize_survivors<<<(n_act * 32 + 127) / 128, 128, 0, st>>>(pr, ws);
        h->launches++;
    }
    h->last_passes = pass < max_pass ? pass + 1 : max_pass;
    if (prof) {
        cudaStreamSynchronize(st);
        const size_t per = trace ? 5 : 4;   // VC_PASS_TRACE=1 adds an event between the two step kernels
        for (size_t i = 0; i + per - 1 < ev; i += per) {
            float a, b2, c2, c3 = 0.f;
            cudaEventElapsedTime(&a, h->prof_events[i], h->prof_events[i + 1]);
            cudaEventElapsedTime(&b2, h->prof_events[i + 1], h->prof_events[i + 2]);
            cudaEventElapsedTime(&c2, h->prof_events[i + 2], h->prof_events[i + per - 1]);
            if (trace) cudaEventElapsedTime(&c3, h->prof_events[i + 2], h->prof_events[i + 3]);
            h->prof_ms[0] += a; h->prof_ms[1] += b2; h->prof_ms[2] += c2;
            if (trace) fprintf(stderr, "[vc] pass %3zu eval_st %.3f eval_pm %.3f step %.3f ms (step_a %.3f)\n", i / per, a, b2, c2, c3);
            h->prof_launches[0]++; h->prof_launches[1] += pr.M > 0; h->prof_launches[2]++;
        }
    }
    return check_cuda(cudaGetLastError(), "nmpc solve");
}


// ---- vc_nmpc_solve_static: the same solve with NO allocation (include/vc_api.h) -------------------------------------------
// Callers that may not call cudaMalloc (a tight serving loop, a harness that forbids it) get the solver workspace and the
// two active lists from this file-scope array instead of the handle's cudaMalloc'ed slab.  Every GPU that loads the library
// holds the 16 MiB until exit; every call shares them (calls that overlap on one GPU race), and they keep the last call's
// contents: k_prepare re-initialises every instance record it uses and the pass loop clears the list counters.
constexpr size_t STATIC_SCRATCH_BYTES = (size_t)VC_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];
constexpr int STATIC_MAX_DEVICES = 64;
static unsigned char* g_static_scratch_addr[STATIC_MAX_DEVICES] = {};   // device address of the array, per device
static std::mutex g_static_mutex;

// In host code `g_static_scratch` is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress,
// once per device, with that device current.
static int static_scratch_address(int device, unsigned char** out) {
    if (device < 0 || device >= STATIC_MAX_DEVICES) { set_error("vc_nmpc_solve_static: device index out of range"); return -1; }
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_scratch_addr[device]) {
        int rc = check_cuda(cudaSetDevice(device), "cudaSetDevice(static scratch)");
        if (rc) return rc;
        void* ptr = nullptr;
        rc = check_cuda(cudaGetSymbolAddress(&ptr, g_static_scratch), "cudaGetSymbolAddress(static scratch)");
        if (rc) return rc;
        if (!ptr) { set_error("vc_nmpc_solve_static: null symbol address"); return -1; }
        g_static_scratch_addr[device] = (unsigned char*)ptr;
    }
    *out = g_static_scratch_addr[device];
    return 0;
}

int launch_nmpc_solve_static(vc_handle* h, int B, const double* d_x0, const double* d_kappa, const double* d_ds,
                             const double* d_obst, int n_obs, double* d_X, double* d_U, double* d_obj, int* d_status,
                             int* d_iters, double* d_kkt) {
    Prob pr;
    int rc = fill_prob(h, pr, B, d_x0, d_kappa, d_ds, d_obst, n_obs);
    if (rc) return rc;
    if (!d_X || !d_U) { set_error("nmpc: null X/U pointer"); return -1; }
    pr.X = d_X; pr.U = d_U; pr.obj = d_obj; pr.kkt = d_kkt; pr.status = d_status; pr.iters = d_iters;
    Ws ws;
    ws.H = pr.H;
    ws.stride = (size_t)WS_PER_STAGE * pr.H + NSCAL;
    const size_t ws_bytes = (ws.stride * sizeof(double) * (size_t)B + 255) & ~(size_t)255;
    const size_t list_bytes = sizeof(int) * (2 * (size_t)B + 2);
    if (ws_bytes + list_bytes > STATIC_SCRATCH_BYTES) {
        char msg[256];
        snprintf(msg, sizeof(msg), "vc_nmpc_solve_static: B = %d instances at horizon %d need %zu bytes of scratch, the static "
                 "array holds %zu (at most %zu instances at this horizon); use vc_nmpc_solve", B, pr.H, ws_bytes + list_bytes,
                 STATIC_SCRATCH_BYTES, (STATIC_SCRATCH_BYTES - 256 - 8) / (ws.stride * sizeof(double) + 2 * sizeof(int)));
        set_error(msg);
        return -1;
    }
    unsigned char* scratch = nullptr;
    rc = static_scratch_address(h->device, &scratch);
    if (rc) return rc;
    cudaSetDevice(h->device);
    rc = step_kernels_smem_optin();
    if (rc) return rc;
    ws.base = (double*)scratch;
    int* act = (int*)(scratch + ws_bytes);
    int* counts[2] = {act, act + 1};
    int* lists[2] = {act + 2, act + 2 + B};
    cudaStream_t st = h->stream;
    k_prepare<<<(B * pr.H + 127) / 128, 128, 0, st>>>(pr, ws, lists[0], counts[0]);
    h->launches++;
    int cur = 0, n_act = B, pass = 0;
    const int max_pass = h->cfg.max_iter + h->cfg.max_iter / 8 + 32;
    const int wide_max = h->cfg.wide_max;
    for (; pass < max_pass; ++pass) {
        const long long t_st = (long long)n_act * pr.N * EVAL_ST_LANES, t_pm = (long long)n_act * pr.M * 15;
        pr.act_list = lists[cur]; pr.act_count = counts[cur];
        k_eval_st<<<(unsigned)((t_st + VC_EVAL_TPB - 1) / VC_EVAL_TPB), VC_EVAL_TPB, 0, st>>>(pr, ws);
        if (pr.M > 0) k_eval_pm<<<(unsigned)((t_pm + 127) / 128), 128, 0, st>>>(pr, ws);
        cudaMemsetAsync(counts[1 - cur], 0, sizeof(int), st);
        if (n_act <= wide_max) {
            k_step_a_wide<<<n_act, WIDE_NT, 0, st>>>(pr, ws);
            k_step_b_wide<<<n_act, 32 * LS_TRIALS, 0, st>>>(pr, ws, lists[1 - cur], counts[1 - cur]);
        } else {
            k_step_a<<<(n_act + STEP_WARPS - 1) / STEP_WARPS, 32 * STEP_WARPS, STEP_SMEM, st>>>(pr, ws);
            k_step_b<<<(n_act + STEP_WARPS - 1) / STEP_WARPS, 32 * STEP_WARPS, STEP_SMEM, st>>>(pr, ws, lists[1 - cur], counts[1 - cur]);
        }
        k_update<<<(unsigned)(((long long)n_act * pr.H * 32 + 127) / 128), 128, 0, st>>>(pr, ws);
        cur = 1 - cur;
        h->launches += pr.M > 0 ? 5 : 4;
        if (pass >= 3) {   // survivors of this pass, read into a stack variable (no pinned buffer to allocate)
            int left = 0;
            cudaMemcpyAsync(&left, counts[cur], sizeof(int), cudaMemcpyDeviceToHost, st);
            rc = check_cuda(cudaStreamSynchronize(st), "nmpc iteration (static scratch)");
            if (rc) return rc;
            n_act = left;
            if (n_act == 0) break;
        }
    }
    if (n_act > 0) {
        pr.act_list = lists[cur]; pr.act_count = counts[cur];
        k_finalize_survivors<<<(n_act * 32 + 127) / 128, 128, 0, st>>>(pr, ws);
        h->launches++;
    }
    h->last_passes = pass < max_pass ? pass + 1 : max_pass;
    return check_cuda(cudaGetLastError(), "nmpc solve (static scratch)");
}


int debug_read(vc_handle* h, int b, int array, double* h_out, int n) {
    if (!h->d_ws) { set_error("debug_read: no workspace"); return -1; }
    const int H = h->cfg.N + h->cfg.M;
    Ws ws; ws.H = H; ws.stride = (size_t)WS_PER_STAGE * H + NSCAL; ws.base = (double*)h->d_ws;
    cudaStreamSynchronize(h->stream);
    if (array < 0) return check_cuda(cudaMemcpy(h_out, ws.scal(b), sizeof(double) * n, cudaMemcpyDeviceToHost), "debug_read");
    if (array >= W_COUNT) { set_error("debug_read: bad array"); return -1; }
    const int w = ws_width(array);   // returned as [H][width], gathered from the stage records
    for (int k = 0; k < H && k * w < n; ++k) {
        const int cnt = (n - k * w) < w ? (n - k * w) : w;
        int rc = check_cuda(cudaMemcpy(h_out + (size_t)k * w, ws.rec(b, k) + ws_offset(array), sizeof(double) * cnt,
                                       cudaMemcpyDeviceToHost), "debug_read");
        if (rc) return rc;
    }
    return 0;
}
}  // namespace vc
