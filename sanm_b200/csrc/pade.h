/* sanm_b200/csrc/pade.h -- vector Pade approximant of the series
 * [x(a); t(a)] (restates libsanm/pade.{h,cpp}).
 *
 * The scalar side (denominator recurrences, roots, bisection, every accept /
 * reject decision) runs on the host in the reference's own arithmetic so that
 * ANM step counts match.  The vector side -- the Gram-Schmidt projections of
 * pade.cpp:30-55, the two numerators and their norms of the range check
 * (:137-147), the evaluation (:181-224) -- goes through PadeVectorOps, which
 * has two implementations: HostPadeOps over host copies of the K+1
 * coefficient vectors (the sanm_pade_* entry points), and DevicePadeOps
 * (pade_dev.cu) over the device-resident series of the ANM drivers, which
 * then never leaves HBM.
 */
#pragma once

#include <cstddef>
#include <memory>
#include <vector>

namespace sanm_b200 {

//! the K+1 coefficient vectors of length len and the passes over them
class PadeVectorOps {
public:
    virtual ~PadeVectorOps() = default;
    virtual size_t nx() const = 0;   //!< number of coefficient vectors
    virtual size_t len() const = 0;  //!< their length (the last entry is t)
    /*! classical Gram-Schmidt of x_1 .. x_n (pade.cpp:30-55): fills
     * a[i * nx + j] = <x_i, u_j> (j < i) and a[i * nx + i] = |u_i| before
     * normalisation; a(i, 1) is asserted small and zeroed under anm_cond.
     * \return false on an exactly zero norm (the reference gives up) */
    virtual bool gram_schmidt(bool anm_cond, double* a) = 0;
    //! last element of x_i
    virtual double t_coeff(size_t i) const = 0;
    /*! pn = sum_i a^(..) x_i sc_n[i] (Horner, i = n .. 1), pl likewise with
     * sc_l over i = n-1 .. 1; returns |pl f - pn|^2 and |pn|^2 */
    virtual void check_norms(double a, const double* sc_n, const double* sc_l,
                             int n, double f, double* diff2, double* pn2) const = 0;
    //! out = (Horner numerator with sc over i = n .. 1) * f + x_0, on the host
    virtual void eval_xt(double a, const double* sc, int n, double f,
                         double* out) const = 0;
};

//! coefficient vectors in host memory
std::unique_ptr<PadeVectorOps> make_host_pade_ops(std::vector<const double*> xs,
                                                  size_t len);

class PadeApproximation {
public:
    /*!
     * \param xs pointers to the K+1 coefficient vectors, each of length len
     *      (the last element is t)
     * \param anm_cond the series satisfies x_1 . x_i = 0 for i > 1
     */
    PadeApproximation(std::vector<const double*> xs, size_t len, bool anm_cond,
                      bool sanity_check)
            : PadeApproximation(make_host_pade_ops(std::move(xs), len), anm_cond,
                                sanity_check) {}
    PadeApproximation(std::unique_ptr<PadeVectorOps> ops, bool anm_cond,
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
    std::unique_ptr<PadeVectorOps> m_ops;
    std::vector<double> m_d, m_d_lo, m_t_nume_coeffs;
    double m_t0 = 0, m_t_max = 0, m_t_max_a = 0;

    double eval_t(double a) const;
};

}  // namespace sanm_b200
