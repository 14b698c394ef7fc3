/* sanm_b200/csrc/pade.h -- vector Pade approximant of the series
 * [x(a); t(a)] (host side; restates libsanm/pade.{h,cpp}).  The K+1
 * coefficient vectors are host copies of the device series; every decision
 * (acceptance, pole, bisection) is taken in the reference's own arithmetic so
 * that ANM step counts match.
 */
#pragma once

#include <cstddef>
#include <vector>

namespace sanm_b200 {

class PadeApproximation {
public:
    /*!
     * \param xs pointers to the K+1 coefficient vectors, each of length len
     *      (the last element is t)
     * \param anm_cond the series satisfies x_1 . x_i = 0 for i > 1
     */
    PadeApproximation(std::vector<const double*> xs, size_t len, bool anm_cond,
                      bool sanity_check);

    //! pade.cpp:107-179
    bool estimate_valid_range(double start, double eps, double limit = 0);
    double get_t_max() const { return m_t_max; }
    double get_t_max_a() const { return m_t_max_a; }
    //! pade.cpp:196-205
    double solve_a(double t) const;
    //! [x(a); t(a)], pade.cpp:219-224
    std::vector<double> eval_xt(double a) const;
    //! denominator coefficients (m_d / m_d_lo of libsanm/pade.h:24-25); empty
    //! when the constructor rejected the series
    const std::vector<double>& denominator(bool lower_order) const {
        return lower_order ? m_d_lo : m_d;
    }

private:
    const bool m_sanity_check;
    std::vector<const double*> m_xs;
    size_t m_len;
    std::vector<double> m_d, m_d_lo, m_t_nume_coeffs;
    double m_t0 = 0, m_t_max = 0, m_t_max_a = 0;

    std::vector<double> eval_nume(double a, const double* d, int n) const;
    double eval_t(double a) const;
};

}  // namespace sanm_b200
