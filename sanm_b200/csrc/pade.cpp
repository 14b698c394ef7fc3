/* sanm_b200/csrc/pade.cpp -- see pade.h. */
#include "pade.h"
#include "poly.h"
#include "rt.h"

#include <algorithm>
#include <cmath>
#include <vector>
#include <limits>

namespace sanm_b200 {

namespace {
constexpr size_t CHUNK = 4096;

//! fixed chunking => the result does not depend on the number of threads
double dot(const double* a, const double* b, size_t n) {
    const size_t nchunk = (n + CHUNK - 1) / CHUNK;
    if (nchunk <= 1) {
        double s = 0;
        for (size_t i = 0; i < n; ++i) s += a[i] * b[i];
        return s;
    }
    std::vector<double> part(nchunk);
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) schedule(static)
    for (int64_t c = 0; c < (int64_t)nchunk; ++c) {
        const size_t i0 = c * CHUNK, i1 = std::min(n, i0 + CHUNK);
        double s = 0;
        for (size_t i = i0; i < i1; ++i) s += a[i] * b[i];
        part[c] = s;
    }
    double s = 0;
    for (double v : part) s += v;
    return s;
}

class HostPadeOps final : public PadeVectorOps {
    std::vector<const double*> m_xs;
    size_t m_len;

public:
    HostPadeOps(std::vector<const double*> xs, size_t len)
            : m_xs{std::move(xs)}, m_len{len} {}
    size_t nx() const override { return m_xs.size(); }
    size_t len() const override { return m_len; }
    double t_coeff(size_t i) const override { return m_xs[i][m_len - 1]; }

    bool gram_schmidt(bool anm_cond, double* astore) override {
        const size_t nx = m_xs.size(), len = m_len;
        auto a = [&](int i, int j) -> double& { return astore[i * nx + j]; };
        const int n = (int)nx - 1;
        constexpr double eps = std::numeric_limits<double>::epsilon();
        std::vector<std::vector<double>> orth(nx);
        // Per i: all projections a(i, j) in ONE pass over x_i (they use the
        // original x_i, so they do not depend on each other), then one pass
        // that subtracts them in the order j = 1, 2, ...  Chunking and
        // summation orders are those of dot() above, so the numbers are the
        // ones of the plain j-by-j loop; only the number of parallel regions
        // drops from O(n^2) to O(n).
        const size_t nchunk = (len + CHUNK - 1) / CHUNK;
        const bool par = nchunk > 1;
        std::vector<double> part((size_t)nx * std::max<size_t>(nchunk, 1));
        for (int i = 1; i <= n; ++i) {
            auto& u = orth[i];
            u.assign(m_xs[i], m_xs[i] + len);
            if (i > 1) {
                const double* xi = m_xs[i];
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) schedule(static) if (par)
                for (int64_t c = 0; c < (int64_t)nchunk; ++c) {
                    const size_t i0 = c * CHUNK, i1 = std::min(len, i0 + CHUNK);
                    for (int j = 1; j < i; ++j) {
                        const double* o = orth[j].data();
                        double sacc = 0;
                        for (size_t q = i0; q < i1; ++q) sacc += xi[q] * o[q];
                        part[(size_t)j * nchunk + c] = sacc;
                    }
                }
                std::vector<double> fac(i, 0.0);
                for (int j = 1; j < i; ++j) {
                    double sacc = 0;
                    for (size_t c = 0; c < nchunk; ++c) sacc += part[(size_t)j * nchunk + c];
                    a(i, j) = sacc;
                    if (anm_cond && j == 1) {
                        sanm_check(std::fabs(a(i, j)) < 1e-4);
                        a(i, j) = 0;
                    } else {
                        fac[j] = -a(i, j);
                    }
                }
                const int j_first = anm_cond ? 2 : 1;
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) schedule(static) if (par)
                for (int64_t c = 0; c < (int64_t)nchunk; ++c) {
                    const size_t i0 = c * CHUNK, i1 = std::min(len, i0 + CHUNK);
                    for (int j = j_first; j < i; ++j) {
                        const double* o = orth[j].data();
                        const double f = fac[j];
                        for (size_t q = i0; q < i1; ++q) u[q] += o[q] * f;
                    }
                }
            }
            double aii = std::sqrt(dot(u.data(), u.data(), len));
            if (aii == 0) return false;
            a(i, i) = aii;
            double inv = 1 / std::max(aii, eps);
            for (auto& v : u) v *= inv;
            if (aii < eps) {
                double nn = 1 / std::sqrt(dot(u.data(), u.data(), len));
                for (auto& v : u) v *= nn;
            }
        }
        return true;
    }

    void check_norms(double a, const double* sc_n, const double* sc_l, int n,
                     double f, double* diff2_out, double* pn2_out) const override {
        // One fused pass per candidate: both numerators by Horner (x_i read
        // once for the two), their weighted difference and the two squared
        // norms.  The element operations and the chunked summation are those
        // of the unfused form (eval_nume + dot).
        const int64_t nchunk = ((int64_t)m_len + CHUNK - 1) / CHUNK;
        std::vector<double> part_d(nchunk), part_n(nchunk);
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) schedule(static) if (nchunk > 1)
        for (int64_t c = 0; c < nchunk; ++c) {
            const int64_t q0 = c * CHUNK, q1 = std::min<int64_t>(m_len, q0 + CHUNK);
            double pn[64], pl[64];
            double sd = 0, sn = 0;
            for (int64_t qb = q0; qb < q1; qb += 64) {
                const int w = (int)std::min<int64_t>(64, q1 - qb);
                for (int t = 0; t < w; ++t) pn[t] = pl[t] = 0.0;
                for (int i = n; i >= 1; --i) {
                    const double* x = m_xs[i] + qb;
                    const double s1 = sc_n[i];
                    for (int t = 0; t < w; ++t) pn[t] = pn[t] * a + x[t] * s1;
                    if (i <= n - 1) {
                        const double s2 = sc_l[i];
                        for (int t = 0; t < w; ++t) pl[t] = pl[t] * a + x[t] * s2;
                    }
                }
                for (int t = 0; t < w; ++t) {
                    const double dv = pl[t] * f - pn[t];
                    sd += dv * dv;
                    sn += pn[t] * pn[t];
                }
            }
            part_d[c] = sd;
            part_n[c] = sn;
        }
        double diff2 = 0, pn2 = 0;
        for (int64_t c = 0; c < nchunk; ++c) {
            diff2 += part_d[c];
            pn2 += part_n[c];
        }
        *diff2_out = diff2;
        *pn2_out = pn2;
    }

    void eval_xt(double a, const double* sc, int n, double f,
                 double* out) const override {
        // pade.cpp:181-194, 219-224
        const int64_t nchunk = ((int64_t)m_len + CHUNK - 1) / CHUNK;
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) schedule(static) if (nchunk > 1)
        for (int64_t c = 0; c < nchunk; ++c) {
            const int64_t q0 = c * CHUNK, q1 = std::min<int64_t>(m_len, q0 + CHUNK);
            for (int64_t q = q0; q < q1; ++q) out[q] = 0.0;
            // Horner per element, i = n .. 1 (same operation order as a pass per i)
            for (int i = n; i >= 1; --i) {
                const double* x = m_xs[i];
                const double s = sc[i];
                for (int64_t q = q0; q < q1; ++q) out[q] = out[q] * a + x[q] * s;
            }
            for (int64_t q = q0; q < q1; ++q) out[q] = out[q] * f + m_xs[0][q];
        }
    }
};
}  // namespace

std::unique_ptr<PadeVectorOps> make_host_pade_ops(std::vector<const double*> xs,
                                                  size_t len) {
    return std::make_unique<HostPadeOps>(std::move(xs), len);
}

PadeApproximation::PadeApproximation(std::unique_ptr<PadeVectorOps> ops,
                                     bool anm_cond, bool sanity_check)
        : m_sanity_check{sanity_check}, m_ops{std::move(ops)} {
    const size_t nx = m_ops->nx(), len = m_ops->len();
    sanm_check(nx >= 3);
    if (len < nx * 2 || nx <= 4) return;  // pade.cpp:19-21
    std::vector<double> astore(nx * nx, 0.0);
    auto a = [&](int i, int j) -> double& { return astore[i * nx + j]; };
    const int n = (int)nx - 1;
    if (!m_ops->gram_schmidt(anm_cond, astore.data())) {
        m_d.clear();
        return;
    }

    auto solve_d = [&](std::vector<double>& d, int nn) {
        // pade.cpp:68-92
        d.assign(nn, 0.0);
        d[0] = 1;
        for (int i = 1; i < nn; ++i) {
            double s = 0;
            for (int j = 0; j < i; ++j) s += a(nn - j, nn - i) * d[j];
            double y = a(nn - i, nn - i);
            d[i] = -s * y / (y * y + 1e-20);
        }
        if (m_sanity_check) {
            for (int i = 1; i < nn; ++i) {
                double s = 0;
                for (int j = 0; j <= nn - i; ++j) s += a(nn - j, i) * d[j];
                sanm_check(std::fabs(s) < 1e-5, "%g", s);
            }
        }
    };
    solve_d(m_d, n);
    solve_d(m_d_lo, n - 1);

    m_t_nume_coeffs.assign(n, 0.0);
    for (int i = 0; i < n; ++i) {
        double ti = m_ops->t_coeff(i);
        if (!i) {
            m_t0 = ti;
        } else {
            for (int j = 0; j < n - i; ++j) m_t_nume_coeffs[i + j] += m_d[j] * ti;
        }
    }
}

double PadeApproximation::eval_t(double a) const {
    return poly::eval(m_t_nume_coeffs, a) / poly::eval(m_d, a) + m_t0;
}

bool PadeApproximation::estimate_valid_range(double start, double eps,
                                             double limit) {
    sanm_check(start > 0 && eps > 0);
    if (m_d.empty()) return false;
    std::vector<std::complex<double>> rts;
    if (!poly::roots(m_d, true, rts)) return false;
    double pole = 0;
    for (auto& r : rts) {
        sanm_check(r.imag() == 0);
        if (r.real() > 0 && (pole == 0 || r.real() < pole)) pole = r.real();
    }
    if (pole == 0) pole = start * 4;
    if (pole <= start) return false;

    const double eps2 = eps * eps;
    const int n = (int)m_ops->nx() - 2;
    std::vector<double> sc_n(n + 1), sc_l(n + 1);
    auto check = [&](double a) {
        // pade.cpp:137-147
        double denom_n = poly::eval(m_d, a), denom_lo = poly::eval(m_d_lo, a);
        const double f = denom_n / denom_lo;
        for (int i = n; i >= 1; --i) sc_n[i] = poly::eval(m_d.data(), n - i + 1, a);
        for (int i = n - 1; i >= 1; --i) sc_l[i] = poly::eval(m_d_lo.data(), n - 1 - i + 1, a);
        double diff2 = 0, pn2 = 0;
        m_ops->check_norms(a, sc_n.data(), sc_l.data(), n, f, &diff2, &pn2);
        return diff2 <= pn2 * eps2;
    };
    double left = start * 1.001, right = start + (pole - start) * 0.99;
    if (!check(left)) return false;
    if (limit && right > limit) right = limit;
    if (right > start * 2) {
        if (check(start * 2)) {
            left = start * 2;
        } else {
            right = start * 2;
        }
    }
    int iter = 0;
    while (iter < 8 && right - left > 1e-3) {
        double mid = (left + right) / 2;
        if (check(mid)) {
            left = mid;
        } else {
            right = mid;
        }
        ++iter;
    }
    m_t_max_a = left;
    m_t_max = eval_t(left);
    return true;
}

double PadeApproximation::solve_a(double t) const {
    sanm_check(t >= m_t0 && t <= m_t_max);
    if (t == m_t_max) return m_t_max_a;
    std::vector<double> c = m_t_nume_coeffs;
    for (size_t i = 0; i < c.size(); ++i) c[i] -= (t - m_t0) * m_d[i];
    return poly::solve_eqn(c, 0, m_t_max_a, 0);
}

std::vector<double> PadeApproximation::eval_xt(double a) const {
    const int n = (int)m_ops->nx() - 2;
    std::vector<double> sc(n + 1, 0.0), ret(m_ops->len());
    for (int i = n; i >= 1; --i) sc[i] = poly::eval(m_d.data(), n - i + 1, a);
    m_ops->eval_xt(a, sc.data(), n, a / poly::eval(m_d, a), ret.data());
    return ret;
}

}  // namespace sanm_b200
