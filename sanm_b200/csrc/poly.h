/* sanm_b200/csrc/poly.h -- scalar polynomial tools of the continuation driver
 * (host side; restates libsanm/unary_polynomial.{h,cpp}). */
#pragma once

#include <complex>
#include <cstddef>
#include <vector>

namespace sanm_b200 {
namespace poly {

//! Horner evaluation of sum f[i] x^i  (unary_polynomial.cpp:84-90)
double eval(const double* f, size_t n, double x);
inline double eval(const std::vector<double>& f, double x) {
    return eval(f.data(), f.size(), x);
}

//! x in [xmin, xmax] with f(x) = b by Brent's zero finder, absolute tolerance
//! eps (unary_polynomial.cpp:101-108; third_party/BRENT `zero`)
double solve_eqn(const std::vector<double>& f, double xmin, double xmax,
                 double b = 0, double eps = 1e-6);

//! |x| below which a degree-`order` polynomial evaluates stably in double
//! (unary_polynomial.cpp:110-116)
double stable_x_range(int order);

/*!
 * all roots by simultaneous Bairstow / Newton synthetic division
 * (K. W. Ellenberger, CACM Algorithm 30, 1960; the method the reference uses,
 * unary_polynomial.cpp:154-334).  f[0] is the constant term.
 * \return false if the iteration did not converge
 */
bool roots(const std::vector<double>& f, bool only_real,
           std::vector<std::complex<double>>& out, int max_iter = 300,
           double tol = 1e-8);

//! Brent's bracketing zero finder (R. P. Brent, Algorithms for Minimization
//! without Derivatives, 1973, ch. 4)
template <class F>
double brent_zero(double a, double b, double t, F&& f);

}  // namespace poly
}  // namespace sanm_b200
