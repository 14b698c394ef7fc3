/* sanm_b200/csrc/poly.cpp -- see poly.h. */
#include "poly.h"
#include "rt.h"

#include <cfloat>
#include <cmath>

namespace sanm_b200 {
namespace poly {

double eval(const double* f, size_t n, double x) {
    double r = 0;
    for (size_t i = n; i-- > 0;) r = r * x + f[i];
    return r;
}

template <class F>
double brent_zero(double a, double b, double t, F&& f) {
    const double macheps = DBL_EPSILON;
    double sa = a, sb = b, fa = f(sa), fb = f(sb);
    double c = sa, fc = fa, e = sb - sa, d = e;
    for (;;) {
        if (std::fabs(fc) < std::fabs(fb)) {
            sa = sb;
            sb = c;
            c = sa;
            fa = fb;
            fb = fc;
            fc = fa;
        }
        const double tol = 2.0 * macheps * std::fabs(sb) + t;
        const double m = 0.5 * (c - sb);
        if (std::fabs(m) <= tol || fb == 0.0) break;
        if (std::fabs(e) < tol || std::fabs(fa) <= std::fabs(fb)) {
            e = m;
            d = e;
        } else {
            double s = fb / fa, p, q;
            if (sa == c) {
                // secant
                p = 2.0 * m * s;
                q = 1.0 - s;
            } else {
                // inverse quadratic interpolation
                q = fa / fc;
                const double r = fb / fc;
                p = s * (2.0 * m * q * (q - r) - (sb - sa) * (r - 1.0));
                q = (q - 1.0) * (r - 1.0) * (s - 1.0);
            }
            if (0.0 < p) {
                q = -q;
            } else {
                p = -p;
            }
            s = e;
            e = d;
            if (2.0 * p < 3.0 * m * q - std::fabs(tol * q) &&
                p < std::fabs(0.5 * s * q)) {
                d = p / q;
            } else {
                e = m;
                d = e;
            }
        }
        sa = sb;
        fa = fb;
        if (tol < std::fabs(d)) {
            sb += d;
        } else if (0.0 < m) {
            sb += tol;
        } else {
            sb -= tol;
        }
        fb = f(sb);
        if ((0.0 < fb && 0.0 < fc) || (fb <= 0.0 && fc <= 0.0)) {
            c = sa;
            fc = fa;
            e = sb - sa;
            d = e;
        }
    }
    return sb;
}

double solve_eqn(const std::vector<double>& f, double xmin, double xmax,
                 double b, double eps) {
    sanm_check(!f.empty() && xmin < xmax);
    auto fn = [&](double x) { return eval(f, x) - b; };
    double f0 = fn(xmin), f1 = fn(xmax);
    sanm_check(f0 * f1 <= 0, "no zero point: f0=%g f1=%g", f0, f1);
    return brent_zero(xmin, xmax, eps, fn);
}

double stable_x_range(int order) {
    return std::pow(1e15, 1.0 / static_cast<double>(order));
}

namespace {

/*!
 * Ellenberger's simultaneous Bairstow (quadratic) / Newton (linear) synthetic
 * division, written as an explicit state machine.  The arithmetic and the
 * order of operations follow the reference routine statement by statement
 * (unary_polynomial.cpp:154-334): in ANM use the denominator polynomial always
 * has a vanishing leading coefficient (pade.cpp:68-79 with a(i,1) = 0), which
 * sends one root to infinity, and what the routine returns after deflating
 * that root decides the Pade pole -- so the degenerate behaviour has to be the
 * same, not just the well-conditioned roots.
 */
class BairstowNewton {
    enum State { INIT, REVERSE, SETUP, ITERATE, LINEAR, QUADRATIC };

    int n;
    std::vector<double> hs, bs, cs, ds, es;  // logical index range [-2, n]
    double* h;
    double* b;
    double* c;
    double* d;
    double* e;
    const bool only_real;
    const int max_iter;
    std::vector<std::complex<double>>& out;

public:
    BairstowNewton(const std::vector<double>& a, bool only_real_, int max_iter_,
                   std::vector<std::complex<double>>& out_)
            : n{(int)a.size() - 1},
              hs(n + 3, 0.0),
              bs(n + 3, 0.0),
              cs(n + 3, 0.0),
              ds(n + 3, 0.0),
              es(n + 3, 0.0),
              h{hs.data() + 2},
              b{bs.data() + 2},
              c{cs.data() + 2},
              d{ds.data() + 2},
              e{es.data() + 2},
              only_real{only_real_},
              max_iter{max_iter_},
              out{out_} {
        // highest power first
        for (int j = n, m = 0; j >= 0; --j) h[j] = a[m++];
    }

    bool run(double tol) {
        double K = 1 / tol;  // reciprocal of the demanded relative accuracy
        int t = 1;  // -1 while working on the reciprocal-root polynomial
        double ps = 0, qs = 0, pt = 0, qt = 0, s = 0, rev = 1.0, r = 0;
        double p = 0, q = 0;
        for (; h[n] == 0.0; --n) {
            sanm_check(n > -2);
        }
        State st = INIT;
        for (;;) {
            switch (st) {
                case INIT: {
                    if (n == 0) return true;
                    ps = qs = pt = qt = s = 0.0;
                    rev = 1.0;
                    if (n == 1) {
                        r = -h[1] / h[0];
                        st = LINEAR;
                        break;
                    }
                    for (int j = n; j >= 0; --j) {
                        if (h[j] != 0.0) s += std::log(std::fabs(h[j]));
                    }
                    s = std::exp(s / (n + 1));
                    for (int j = n; j >= 0; --j) h[j] /= s;
                    st = std::fabs(h[1] / h[0]) < std::fabs(h[n - 1] / h[n])
                                 ? REVERSE
                                 : SETUP;
                    break;
                }
                case REVERSE: {
                    t = -t;
                    for (int j = (n - 1) / 2; j >= 0; --j) {
                        s = h[j];
                        h[j] = h[n - j];
                        h[n - j] = s;
                    }
                    st = SETUP;
                    break;
                }
                case SETUP: {
                    st = ITERATE;
                    if (qs != 0.0) {
                        p = ps;
                        q = qs;
                    } else {
                        if (h[n - 2] == 0.0) {
                            q = 1.0;
                            p = -2.0;
                        } else {
                            q = h[n] / h[n - 2];
                            p = (h[n - 1] - q * h[n - 3]) / h[n - 2];
                        }
                        if (n == 2) {
                            st = QUADRATIC;
                        } else {
                            r = 0.0;
                        }
                    }
                    break;
                }
                case ITERATE: {
                    State next = REVERSE;  // taken when max_iter is exhausted
                    bool done = false;
                    for (int i = max_iter; i > 0 && !done; --i) {
                        for (int j = 0; j <= n; ++j) {
                            b[j] = h[j] - p * b[j - 1] - q * b[j - 2];
                            c[j] = b[j] - p * c[j - 1] - q * c[j - 2];
                        }
                        if (h[n - 1] != 0.0 && b[n - 1] != 0.0) {
                            if (std::fabs(h[n - 1] / b[n - 1]) >= K) {
                                b[n] = h[n] - q * b[n - 2];
                            }
                            if (b[n] == 0.0 || K < std::fabs(h[n] / b[n])) {
                                next = QUADRATIC;
                                done = true;
                                break;
                            }
                        }
                        for (int j = 0; j <= n; ++j) {
                            d[j] = h[j] + r * d[j - 1];
                            e[j] = d[j] + r * e[j - 1];
                        }
                        if (d[n] == 0.0 || K < std::fabs(h[n] / d[n])) {
                            next = LINEAR;
                            done = true;
                            break;
                        }
                        c[n - 1] = -p * c[n - 2] - q * c[n - 3];
                        s = c[n - 2] * c[n - 2] - c[n - 1] * c[n - 3];
                        if (s == 0.0) {
                            p -= 2.0;
                            q *= (q + 1.0);
                        } else {
                            p += (b[n - 1] * c[n - 2] - b[n] * c[n - 3]) / s;
                            q += (-b[n - 1] * c[n - 1] + b[n] * c[n - 2]) / s;
                        }
                        if (e[n - 1] == 0.0) {
                            r -= 1.0;
                        } else {
                            r -= d[n] / e[n - 1];
                        }
                    }
                    if (!done) {
                        // no convergence: keep the last two (p, q), relax the
                        // accuracy every other attempt, flip the polynomial
                        ps = pt;
                        qs = qt;
                        pt = p;
                        qt = q;
                        if (rev < 0.0) K /= 10.0;
                        if (K < 1e-8) return false;
                        rev = -rev;
                    }
                    st = next;
                    break;
                }
                case LINEAR: {
                    if (t < 0) r = 1.0 / r;
                    --n;
                    out.emplace_back(r, 0.0);
                    for (int j = n; j >= 0; --j) {
                        if (d[j] != 0.0 && std::fabs(h[j] / d[j]) < K) {
                            h[j] = d[j];
                        } else {
                            h[j] = 0.0;
                        }
                    }
                    if (n == 0) return true;
                    st = ITERATE;
                    break;
                }
                case QUADRATIC: {
                    if (t < 0) {
                        p /= q;
                        q = 1.0 / q;
                    }
                    n -= 2;
                    if (0.0 < (q - (p * p / 4.0))) {
                        s = std::sqrt(q - (p * p / 4.0));
                        if (!only_real) {
                            out.emplace_back(-p / 2.0, s);
                            out.emplace_back(-p / 2.0, -s);
                        }
                    } else {
                        s = std::sqrt(((p * p / 4.0)) - q);
                        double qq = p < 0.0 ? -p / 2.0 + s : -p / 2.0 - s;
                        out.emplace_back(qq, 0.0);
                        out.emplace_back(q / qq, 0.0);
                    }
                    for (int j = n; j >= 0; --j) {
                        if (b[j] != 0.0 && std::fabs(h[j] / b[j]) < K) {
                            h[j] = b[j];
                        } else {
                            h[j] = 0.0;
                        }
                    }
                    st = INIT;
                    break;
                }
            }
        }
    }
};

}  // namespace

bool roots(const std::vector<double>& f, bool only_real,
           std::vector<std::complex<double>>& out, int max_iter, double tol) {
    sanm_check(f.size() >= 2);
    out.clear();
    BairstowNewton bn{f, only_real, max_iter, out};
    return bn.run(tol);
}

}  // namespace poly
}  // namespace sanm_b200
