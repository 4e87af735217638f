// ORACLE (test infrastructure only -- never linked into or loaded by the product path).
//
// Second-order forward-mode automatic differentiation on the host: a Jet<N, S> carries a value, its gradient with
// respect to N stage variables and the packed lower triangle of its Hessian.  This plays the role CasADi's AD plays
// in the reference (cascaded_mpc.py:65 "expand": True -> exact constraint Jacobian and Lagrangian Hessian for IPOPT).
// It is deliberately a different mechanism from the product's hyper-dual (pair-per-thread) evaluation.
#pragma once
#include <cmath>

namespace vco {

template <int N, bool SECOND = true> struct Jet {
    static constexpr int NH = SECOND ? N * (N + 1) / 2 : 0;   // SECOND = false: gradient only (Jacobian sweeps)
    double v;
    double g[N];
    double h[NH > 0 ? NH : 1];   // h[i*(i+1)/2 + j], i >= j
    Jet() {}
    Jet(double c) : v(c) { for (int i = 0; i < N; ++i) g[i] = 0.0; for (int i = 0; i < NH; ++i) h[i] = 0.0; }
    static Jet var(double x, int idx) { Jet r(x); r.g[idx] = 1.0; return r; }
    double hess(int i, int j) const { return i >= j ? h[i * (i + 1) / 2 + j] : h[j * (j + 1) / 2 + i]; }
};

inline double val(double x) { return x; }
template <int N, bool S> inline double val(const Jet<N, S>& x) { return x.v; }

// chain rule for a univariate function with f, f', f'' known at x.v
template <int N, bool S> inline Jet<N, S> lift(const Jet<N, S>& x, double f0, double f1, double f2) {
    Jet<N, S> r;
    r.v = f0;
    for (int i = 0; i < N; ++i) r.g[i] = f1 * x.g[i];
    if (S) {
        int q = 0;
        for (int i = 0; i < N; ++i) {
            const double gi = f2 * x.g[i];
            for (int j = 0; j <= i; ++j, ++q) r.h[q] = f1 * x.h[q] + gi * x.g[j];
        }
    }
    return r;
}

template <int N, bool S> inline Jet<N, S> operator+(const Jet<N, S>& a, const Jet<N, S>& b) {
    Jet<N, S> r; r.v = a.v + b.v;
    for (int i = 0; i < N; ++i) r.g[i] = a.g[i] + b.g[i];
    for (int i = 0; i < Jet<N, S>::NH; ++i) r.h[i] = a.h[i] + b.h[i];
    return r;
}
template <int N, bool S> inline Jet<N, S> operator-(const Jet<N, S>& a, const Jet<N, S>& b) {
    Jet<N, S> r; r.v = a.v - b.v;
    for (int i = 0; i < N; ++i) r.g[i] = a.g[i] - b.g[i];
    for (int i = 0; i < Jet<N, S>::NH; ++i) r.h[i] = a.h[i] - b.h[i];
    return r;
}
template <int N, bool S> inline Jet<N, S> operator-(const Jet<N, S>& a) {
    Jet<N, S> r; r.v = -a.v;
    for (int i = 0; i < N; ++i) r.g[i] = -a.g[i];
    for (int i = 0; i < Jet<N, S>::NH; ++i) r.h[i] = -a.h[i];
    return r;
}
template <int N, bool S> inline Jet<N, S> operator*(const Jet<N, S>& a, const Jet<N, S>& b) {
    Jet<N, S> r; r.v = a.v * b.v;
    for (int i = 0; i < N; ++i) r.g[i] = a.g[i] * b.v + a.v * b.g[i];
    if (S) {
        int q = 0;
        for (int i = 0; i < N; ++i)
            for (int j = 0; j <= i; ++j, ++q)
                r.h[q] = a.h[q] * b.v + a.v * b.h[q] + a.g[i] * b.g[j] + a.g[j] * b.g[i];
    }
    return r;
}
template <int N, bool S> inline Jet<N, S> operator+(const Jet<N, S>& a, double c) { Jet<N, S> r = a; r.v += c; return r; }
template <int N, bool S> inline Jet<N, S> operator+(double c, const Jet<N, S>& a) { Jet<N, S> r = a; r.v += c; return r; }
template <int N, bool S> inline Jet<N, S> operator-(const Jet<N, S>& a, double c) { Jet<N, S> r = a; r.v -= c; return r; }
template <int N, bool S> inline Jet<N, S> operator-(double c, const Jet<N, S>& a) { Jet<N, S> r = -a; r.v += c; return r; }
template <int N, bool S> inline Jet<N, S> operator*(const Jet<N, S>& a, double c) {
    Jet<N, S> r; r.v = a.v * c;
    for (int i = 0; i < N; ++i) r.g[i] = a.g[i] * c;
    for (int i = 0; i < Jet<N, S>::NH; ++i) r.h[i] = a.h[i] * c;
    return r;
}
template <int N, bool S> inline Jet<N, S> operator*(double c, const Jet<N, S>& a) { return a * c; }
template <int N, bool S> inline Jet<N, S> inv(const Jet<N, S>& a) {
    const double r = 1.0 / a.v;
    return lift(a, r, -r * r, 2.0 * r * r * r);
}
inline double inv(double a) { return 1.0 / a; }
template <int N, bool S> inline Jet<N, S> operator/(const Jet<N, S>& a, const Jet<N, S>& b) { return a * inv(b); }
template <int N, bool S> inline Jet<N, S> operator/(const Jet<N, S>& a, double c) { return a * (1.0 / c); }
template <int N, bool S> inline Jet<N, S> operator/(double c, const Jet<N, S>& b) { return inv(b) * c; }

template <int N, bool S> inline Jet<N, S> sqrt(const Jet<N, S>& a) {
    const double s = std::sqrt(a.v);
    return lift(a, s, 0.5 / s, -0.25 / (s * a.v));
}
template <int N, bool S> inline Jet<N, S> sin(const Jet<N, S>& a) { const double s = std::sin(a.v), c = std::cos(a.v); return lift(a, s, c, -s); }
template <int N, bool S> inline Jet<N, S> cos(const Jet<N, S>& a) { const double s = std::sin(a.v), c = std::cos(a.v); return lift(a, c, -s, -c); }
template <int N, bool S> inline Jet<N, S> tan(const Jet<N, S>& a) {
    const double t = std::tan(a.v), d = 1.0 + t * t;
    return lift(a, t, d, 2.0 * t * d);
}
template <int N, bool S> inline Jet<N, S> atan(const Jet<N, S>& a) {
    const double d = 1.0 / (1.0 + a.v * a.v);
    return lift(a, std::atan(a.v), d, -2.0 * a.v * d * d);
}
template <int N, bool S> inline Jet<N, S> tanh(const Jet<N, S>& a) {
    const double t = std::tanh(a.v), d = 1.0 - t * t;
    return lift(a, t, d, -2.0 * t * d);
}
template <int N, bool S> inline Jet<N, S> fabs(const Jet<N, S>& a) { return a.v < 0.0 ? -a : a; }
using std::atan; using std::cos; using std::fabs; using std::sin; using std::sqrt; using std::tan; using std::tanh;

}  // namespace vco
