// Forward-mode automatic-differentiation scalars for the vehicle models.
//
// The reference never forms derivatives itself: CasADi's AD produces the constraint
// Jacobian and the exact Lagrangian Hessian inside nlpsol (cascaded_mpc.py:65 "expand": True).
// Here the same exact chain rule is applied by template instantiation of the model code:
//   double : values
//   Dual1  : value + one directional derivative          (Jacobian columns, one per lane)
//   HDual  : value + d/da + d/db + d2/(da db)            (Hessian entries, one (a,b) pair per thread)
#pragma once
#include <cmath>

#if defined(__CUDACC__)
#define VC_HD __host__ __device__ __forceinline__
#else
#define VC_HD inline
#endif

namespace vc {

struct Dual1 { double v, d; };
struct HDual { double v, a, b, ab; };
template <int K> struct DualK { double v, d[K]; };

// ---- value access / lifting of a univariate function with known f, f', f'' at g.v
VC_HD double val(double x) { return x; }
VC_HD double val(const Dual1& x) { return x.v; }
VC_HD double val(const HDual& x) { return x.v; }

VC_HD double lift(double, double f0, double, double) { return f0; }
VC_HD Dual1 lift(const Dual1& g, double f0, double f1, double) { return Dual1{f0, f1 * g.d}; }
VC_HD HDual lift(const HDual& g, double f0, double f1, double f2) {
    return HDual{f0, f1 * g.a, f1 * g.b, f1 * g.ab + f2 * g.a * g.b};
}

template <class T> VC_HD T cst(double c);
template <> VC_HD double cst<double>(double c) { return c; }
template <> VC_HD Dual1 cst<Dual1>(double c) { return Dual1{c, 0.0}; }
template <> VC_HD HDual cst<HDual>(double c) { return HDual{c, 0.0, 0.0, 0.0}; }
template <> VC_HD DualK<2> cst<DualK<2>>(double c) { return DualK<2>{c, {0.0, 0.0}}; }
template <> VC_HD DualK<4> cst<DualK<4>>(double c) { return DualK<4>{c, {0.0, 0.0, 0.0, 0.0}}; }

// ---- Dual1 arithmetic
VC_HD Dual1 operator+(const Dual1& x, const Dual1& y) { return {x.v + y.v, x.d + y.d}; }
VC_HD Dual1 operator-(const Dual1& x, const Dual1& y) { return {x.v - y.v, x.d - y.d}; }
VC_HD Dual1 operator*(const Dual1& x, const Dual1& y) { return {x.v * y.v, x.d * y.v + x.v * y.d}; }
VC_HD Dual1 operator-(const Dual1& x) { return {-x.v, -x.d}; }
VC_HD Dual1 operator+(const Dual1& x, double c) { return {x.v + c, x.d}; }
VC_HD Dual1 operator+(double c, const Dual1& x) { return {x.v + c, x.d}; }
VC_HD Dual1 operator-(const Dual1& x, double c) { return {x.v - c, x.d}; }
VC_HD Dual1 operator-(double c, const Dual1& x) { return {c - x.v, -x.d}; }
VC_HD Dual1 operator*(const Dual1& x, double c) { return {x.v * c, x.d * c}; }
VC_HD Dual1 operator*(double c, const Dual1& x) { return {x.v * c, x.d * c}; }

// ---- HDual arithmetic
VC_HD HDual operator+(const HDual& x, const HDual& y) { return {x.v + y.v, x.a + y.a, x.b + y.b, x.ab + y.ab}; }
VC_HD HDual operator-(const HDual& x, const HDual& y) { return {x.v - y.v, x.a - y.a, x.b - y.b, x.ab - y.ab}; }
VC_HD HDual operator*(const HDual& x, const HDual& y) {
    return {x.v * y.v, x.a * y.v + x.v * y.a, x.b * y.v + x.v * y.b,
            x.ab * y.v + x.a * y.b + x.b * y.a + x.v * y.ab};
}
VC_HD HDual operator-(const HDual& x) { return {-x.v, -x.a, -x.b, -x.ab}; }
VC_HD HDual operator+(const HDual& x, double c) { return {x.v + c, x.a, x.b, x.ab}; }
VC_HD HDual operator+(double c, const HDual& x) { return {x.v + c, x.a, x.b, x.ab}; }
VC_HD HDual operator-(const HDual& x, double c) { return {x.v - c, x.a, x.b, x.ab}; }
VC_HD HDual operator-(double c, const HDual& x) { return {c - x.v, -x.a, -x.b, -x.ab}; }
VC_HD HDual operator*(const HDual& x, double c) { return {x.v * c, x.a * c, x.b * c, x.ab * c}; }
VC_HD HDual operator*(double c, const HDual& x) { return {x.v * c, x.a * c, x.b * c, x.ab * c}; }


// ---- DualK<K>: value + K directional derivatives carried together (the value part and every transcendental are
// evaluated once for K Jacobian columns; K = 2 / 4 trade registers for fewer redundant evaluations than Dual1)
template <int K> VC_HD double val(const DualK<K>& x) { return x.v; }
template <int K> VC_HD DualK<K> lift(const DualK<K>& g, double f0, double f1, double) {
    DualK<K> r; r.v = f0;
#pragma unroll
    for (int i = 0; i < K; ++i) r.d[i] = f1 * g.d[i];
    return r;
}
template <int K> VC_HD DualK<K> operator+(const DualK<K>& x, const DualK<K>& y) {
    DualK<K> r; r.v = x.v + y.v;
#pragma unroll
    for (int i = 0; i < K; ++i) r.d[i] = x.d[i] + y.d[i];
    return r;
}
template <int K> VC_HD DualK<K> operator-(const DualK<K>& x, const DualK<K>& y) {
    DualK<K> r; r.v = x.v - y.v;
#pragma unroll
    for (int i = 0; i < K; ++i) r.d[i] = x.d[i] - y.d[i];
    return r;
}
template <int K> VC_HD DualK<K> operator*(const DualK<K>& x, const DualK<K>& y) {
    DualK<K> r; r.v = x.v * y.v;
#pragma unroll
    for (int i = 0; i < K; ++i) r.d[i] = x.d[i] * y.v + x.v * y.d[i];
    return r;
}
template <int K> VC_HD DualK<K> operator-(const DualK<K>& x) {
    DualK<K> r; r.v = -x.v;
#pragma unroll
    for (int i = 0; i < K; ++i) r.d[i] = -x.d[i];
    return r;
}
template <int K> VC_HD DualK<K> operator+(const DualK<K>& x, double c) { DualK<K> r = x; r.v += c; return r; }
template <int K> VC_HD DualK<K> operator+(double c, const DualK<K>& x) { DualK<K> r = x; r.v += c; return r; }
template <int K> VC_HD DualK<K> operator-(const DualK<K>& x, double c) { DualK<K> r = x; r.v -= c; return r; }
template <int K> VC_HD DualK<K> operator-(double c, const DualK<K>& x) { DualK<K> r = -x; r.v += c; return r; }
template <int K> VC_HD DualK<K> operator*(const DualK<K>& x, double c) {
    DualK<K> r; r.v = x.v * c;
#pragma unroll
    for (int i = 0; i < K; ++i) r.d[i] = x.d[i] * c;
    return r;
}
template <int K> VC_HD DualK<K> operator*(double c, const DualK<K>& x) { return x * c; }

// ---- elementary functions, written once for all three scalar types
template <class T> VC_HD T recip(const T& x) {
    const double r = 1.0 / val(x);
    return lift(x, r, -r * r, 2.0 * r * r * r);
}
VC_HD Dual1 operator/(const Dual1& x, const Dual1& y) { return x * recip(y); }
VC_HD Dual1 operator/(const Dual1& x, double c) { return x * (1.0 / c); }
VC_HD Dual1 operator/(double c, const Dual1& y) { return c * recip(y); }
VC_HD HDual operator/(const HDual& x, const HDual& y) { return x * recip(y); }
VC_HD HDual operator/(const HDual& x, double c) { return x * (1.0 / c); }
VC_HD HDual operator/(double c, const HDual& y) { return c * recip(y); }

template <int K> VC_HD DualK<K> operator/(const DualK<K>& x, const DualK<K>& y) { return x * recip(y); }
template <int K> VC_HD DualK<K> operator/(const DualK<K>& x, double c) { return x * (1.0 / c); }
template <int K> VC_HD DualK<K> operator/(double c, const DualK<K>& y) { return c * recip(y); }

template <class T> VC_HD T ad_sqrt(const T& x) {
    const double s = sqrt(val(x));
    const double i = 0.5 / s;
    return lift(x, s, i, -0.5 * i / val(x));
}
template <class T> VC_HD void ad_sincos(const T& x, T& sn, T& cs) {
    double s, c;
#if defined(__CUDA_ARCH__)
    sincos(val(x), &s, &c);
#else
    s = sin(val(x)); c = cos(val(x));
#endif
    sn = lift(x, s, c, -s);
    cs = lift(x, c, -s, -c);
}
template <class T> VC_HD T ad_tan(const T& x) {
    const double t = tan(val(x));
    const double d = 1.0 + t * t;
    return lift(x, t, d, 2.0 * t * d);
}
template <class T> VC_HD T ad_atan(const T& x) {
    const double v = val(x);
    const double d = 1.0 / (1.0 + v * v);
    return lift(x, atan(v), d, -2.0 * v * d * d);
}
template <class T> VC_HD T ad_tanh(const T& x) {
    const double t = tanh(val(x));
    const double d = 1.0 - t * t;
    return lift(x, t, d, -2.0 * t * d);
}
template <class T> VC_HD T ad_abs(const T& x) { return val(x) < 0.0 ? -x : x; }
template <class T> VC_HD T ad_log(const T& x) {
    const double r = 1.0 / val(x);
    return lift(x, log(val(x)), r, -r * r);
}

}  // namespace vc
