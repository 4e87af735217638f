"""ORACLE (test infrastructure only -- never imported by the product path).

CPU restatement of the solver the reference calls: IPOPT 3.14 (bundled with CasADi 3.6.7, pinned in
the reference's pixi.lock:35,123-125; call site cascaded_mpc.py:54-66,308) with the options the
reference sets -- everything default except warm_start_init_point=yes: tol 1e-8, monotone barrier
update mu+ = max(tol/11, min(0.2 mu, mu^1.5)) from mu_init 0.1, fraction-to-boundary tau =
max(0.99, 1-mu), filter line search, inertia correction of the KKT matrix by delta_w, gradient-based
objective scaling (nlp_scaling_max_gradient 100), bound_relax_factor 1e-8, acceptable_tol 1e-6 / 15
iterations.  IPOPT is a third-party dependency absent from /root/reference and from this image, so
this file restates its PUBLISHED algorithm (Waechter & Biegler, Math. Prog. 106, 2006) on a dense
KKT system with scipy's symmetric-indefinite LDL^T.  It does not implement the restoration phase, the
second-order correction or the watchdog (oracle/cpp has the last two): they change the iterate path, and
on a few "branch" ticks of the non-convex NLP also which local solution it ends in.

Parity status: PINNED by the reference's cold-start known answers (tests/test_oracle_nmpc.py).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import scipy.linalg as sla


@dataclass
class IpmOptions:
    tol: float = 1e-8
    max_iter: int = 300
    mu_init: float = 0.1
    acceptable_tol: float = 1e-6
    acceptable_iter: int = 15
    obj_scaling: bool = True
    max_gradient: float = 100.0
    kappa_eps: float = 10.0
    kappa_mu: float = 0.2
    theta_mu: float = 1.5
    tau_min: float = 0.99
    slack_push: float = 1e-3
    kappa_sigma: float = 1e10
    s_max: float = 100.0
    gamma_theta: float = 1e-5
    gamma_phi: float = 1e-8
    eta_phi: float = 1e-8
    delta: float = 1.0
    s_theta: float = 1.1
    s_phi: float = 2.3
    max_backtrack: int = 12
    max_restorations: int = 2
    verbose: bool = False
    obj_scale: float = 0.0     # >0: use this objective scaling instead of computing it at w0


@dataclass
class IpmResult:
    w: np.ndarray
    obj: float
    status: str
    iters: int
    kkt_error: float
    mu: float
    lam_g: np.ndarray
    z: np.ndarray
    obj_scale: float


def _inertia(d):
    """Inertia of the block-diagonal D of an LDL^T factorisation."""
    n = d.shape[0]
    pos = neg = zero = 0
    i = 0
    while i < n:
        if i + 1 < n and d[i + 1, i] != 0.0:
            ev = np.linalg.eigvalsh(d[i:i + 2, i:i + 2])
            for e in ev:
                pos += e > 0
                neg += e < 0
                zero += e == 0
            i += 2
        else:
            e = d[i, i]
            pos += e > 0
            neg += e < 0
            zero += e == 0
            i += 1
    return pos, neg, zero


def solve(nlp, w0, opt: IpmOptions = IpmOptions()) -> IpmResult:
    """IPM with the restoration substitute: if the filter line search finds no acceptable step, the
    iterate is re-initialised by NLP.repair (dynamic rollout) and the barrier loop restarts, keeping the
    objective scaling of the ORIGINAL initial point (that scaling decides where IPOPT's answer sits)."""
    res = _solve_once(nlp, w0, opt)
    restarts = 0
    while res.status == "eval_error" and res.iters == 0 and restarts < opt.max_restorations and hasattr(nlp, "repair"):
        restarts += 1          # invalid starting point: re-initialise, the repaired point sets the objective scaling
        res = _solve_once(nlp, nlp.repair(res.w), opt)
    while res.status == "linesearch_fail" and restarts < opt.max_restorations and hasattr(nlp, "repair"):
        restarts += 1
        o2 = IpmOptions(**{**opt.__dict__, "obj_scale": res.obj_scale})
        r2 = _solve_once(nlp, nlp.repair(res.w), o2)
        r2.iters += res.iters + 1
        res = r2
    return res


def _solve_once(nlp, w0, opt: IpmOptions) -> IpmResult:
    w = np.array(w0, dtype=float)
    n = w.size
    # objective scaling from the gradient at the initial guess
    sf = 1.0
    if opt.obj_scale > 0:
        sf = opt.obj_scale
    elif opt.obj_scaling:
        g0 = np.abs(nlp.grad_f(w)).max()
        if g0 > opt.max_gradient:
            sf = opt.max_gradient / g0
    f0, g0v, c0 = nlp.values(w)
    if not (np.isfinite(f0) and np.isfinite(g0v).all() and np.isfinite(c0).all()):
        return IpmResult(w, f0, "eval_error", 0, np.inf, opt.mu_init, np.zeros(g0v.size), np.zeros(c0.size), sf)
    me, mi = g0v.size, c0.size
    mu = opt.mu_init
    s = np.maximum(c0, opt.slack_push * np.maximum(1.0, np.abs(c0)))
    z = mu / s
    lam = np.zeros(me)
    filt = []
    delta_last = 0.0
    acc_count = 0
    status = "max_iter"
    it = 0
    err0 = np.inf
    theta0 = None

    def errors(gf, Jg, Jc, gv, cv, mu_):
        rd = sf * gf + Jg.T @ lam - Jc.T @ z
        sd = max(opt.s_max, (np.abs(lam).sum() + np.abs(z).sum()) / max(1, me + mi)) / opt.s_max
        sc = max(opt.s_max, np.abs(z).sum() / max(1, mi)) / opt.s_max
        prim = max(np.abs(gv).max(), np.abs(cv - s).max())
        return max(np.abs(rd).max() / sd, prim, np.abs(s * z - mu_).max() / sc)

    for it in range(opt.max_iter + 1):
        fv, gf, gv, Jg, cv, Jc, W = nlp.eval_all(w, lam, z, sf)
        err0 = errors(gf, Jg, Jc, gv, cv, 0.0)
        if opt.verbose:
            print(f"it {it:3d} f {fv:.10f} E0 {err0:.3e} mu {mu:.3e} theta {np.abs(gv).sum() + np.abs(cv - s).sum():.3e}")
        if err0 <= opt.tol:
            status = "solved"
            break
        acc_count = acc_count + 1 if err0 <= opt.acceptable_tol else 0
        if acc_count >= opt.acceptable_iter:
            status = "acceptable"
            break
        if it == opt.max_iter:
            break
        # barrier update (possibly several rungs at once)
        changed = False
        while mu > opt.tol / 11 * (1 + 1e-12) and errors(gf, Jg, Jc, gv, cv, mu) <= opt.kappa_eps * mu:
            mu = max(opt.tol / 11.0, min(opt.kappa_mu * mu, mu ** opt.theta_mu))
            changed = True
        if changed:
            filt = []
        tau = max(opt.tau_min, 1.0 - mu)
        # condensed KKT system
        sig = z / s
        rc = cv - s
        Hc = W + Jc.T @ (sig[:, None] * Jc)
        rhs_x = -(sf * gf + Jg.T @ lam - Jc.T @ (mu / s - sig * rc))
        first_ever = delta_last == 0.0
        theta = np.abs(gv).sum() + np.abs(rc).sum()
        phi = sf * fv - mu * np.log(s).sum()
        if theta0 is None:
            theta0 = theta
        theta_min = 1e-4 * max(1.0, theta0)
        theta_max = 1e4 * max(1.0, theta0)
        trial = 0.0
        accepted = False
        while not accepted:
            # ---- inertia-correcting factorisation (delta_w ladder)
            while True:
                K = np.zeros((n + me, n + me))
                K[:n, :n] = Hc + trial * np.eye(n)
                K[:n, n:] = Jg.T
                K[n:, :n] = Jg
                try:
                    L, D, perm = sla.ldl(K, lower=True)
                    pos, neg, zero = _inertia(D)
                except Exception:
                    pos = neg = zero = -1
                if pos == n and neg == me:
                    sol = np.linalg.solve(K, np.concatenate([rhs_x, -gv]))
                    dw, dlam = sol[:n], sol[n:]
                    break
                if trial == 0.0:
                    trial = 1e-4 if delta_last == 0.0 else max(1e-20, delta_last / 3.0)
                else:
                    trial *= 100.0 if first_ever else 8.0
                if trial > 1e40:
                    return IpmResult(w, fv, "linesearch_fail", it, err0, mu, lam, z, sf)
            dsl = Jc @ dw + rc
            dz = mu / s - z - sig * dsl
            # fraction to the boundary
            neg_s = dsl < 0
            a_p = min(1.0, float((-tau * s[neg_s] / dsl[neg_s]).min())) if neg_s.any() else 1.0
            neg_z = dz < 0
            a_d = min(1.0, float((-tau * z[neg_z] / dz[neg_z]).min())) if neg_z.any() else 1.0
            # ---- filter line search on (theta, phi)
            dphi = sf * gf @ dw - mu * (dsl / s).sum()
            alpha = a_p
            for _ in range(opt.max_backtrack):
                wt, st = w + alpha * dw, s + alpha * dsl
                ft, gt, ct = nlp.values(wt)
                ok = np.isfinite(ft) and np.isfinite(gt).all() and np.isfinite(ct).all()
                if ok:
                    th_t = np.abs(gt).sum() + np.abs(ct - st).sum()
                    ph_t = sf * ft - mu * np.log(st).sum()
                    in_filter = any(th_t >= ft_ and ph_t >= fp_ for ft_, fp_ in filt) or th_t > theta_max
                    if not in_filter:
                        switching = (theta <= theta_min and dphi < 0 and
                                     alpha * (-dphi) ** opt.s_phi > opt.delta * theta ** opt.s_theta)
                        if switching:
                            if ph_t <= phi + opt.eta_phi * alpha * dphi:
                                accepted = True
                        elif th_t <= (1 - opt.gamma_theta) * theta or ph_t <= phi - opt.gamma_phi * theta:
                            accepted = True
                            filt.append(((1 - opt.gamma_theta) * theta, phi - opt.gamma_phi * theta))
                        # tiny steps at the noise floor: accept if both measures are unchanged to rounding
                        if not accepted and abs(ph_t - phi) <= 1e-13 * max(1.0, abs(phi)) and th_t <= theta + 1e-13:
                            accepted = True
                    if accepted:
                        break
                alpha *= 0.5
            if accepted:
                if trial > 0:
                    delta_last = trial
                break
            # No acceptable step along this direction.  IPOPT would enter its restoration phase here; this
            # restatement instead damps the step (Levenberg-Marquardt style: a larger delta_w turns the
            # direction towards the least-norm feasibility step, which the filter accepts) and retries.
            trial = max(trial, 1e-4) * 100.0
            if trial > 1e12:
                break
        if not accepted:
            status = "linesearch_fail"
            break
        if opt.verbose:
            print(f"      alpha {alpha:.3e} a_p {a_p:.3e} a_d {a_d:.3e} delta_w {trial:.2e} |dw| {np.abs(dw).max():.3e} dphi {dphi:.3e}")
        w, s = wt, st
        lam = lam + alpha * dlam
        z = z + a_d * dz
        z = np.minimum(np.maximum(z, mu / (opt.kappa_sigma * s)), opt.kappa_sigma * mu / s)
    fv, _, _ = nlp.values(w)
    return IpmResult(w, fv, status, it, float(err0), mu, lam, z, sf)
