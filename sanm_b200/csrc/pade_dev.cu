/* sanm_b200/csrc/pade_dev.cu -- see pade_dev.h. */
#include "pade_dev.h"

#include "prof.h"

#include <cmath>
#include <cstring>
#include <limits>

namespace sanm_b200 {

namespace {

constexpr int PT = 256;          //!< threads per block
constexpr uint64_t PCH = 2048;   //!< elements per block (fixed: sets the reduction tree)
constexpr int MAXC = 64;         //!< coefficient vectors supported on the device

inline uint64_t nblocks(uint64_t len) { return (len + PCH - 1) / PCH; }

//! sum of v over the block, result valid on every thread after the call
SANM_D inline double block_sum(const rt::BlockCtx& c, double* red, double v) {
    red[c.tid] = v;
    c.sync();
    for (int s = c.nthr / 2; s > 0; s >>= 1) {
        if (c.tid < s) red[c.tid] += red[c.tid + s];
        c.sync();
    }
    const double r = red[0];
    c.sync();
    return r;
}

//! partial[j * nblk + b] = <x_i, u_j> over the block's elements, j = 1 .. nj
struct MultiDotPartial {
    const double* xi;
    const double* u;  //!< u_j = u + j * ld
    uint64_t ld, len, nblk;
    int nj;
    double* partial;
    SANM_D void operator()(const rt::BlockCtx& c) const {
        double* red = reinterpret_cast<double*>(c.smem);
        const uint64_t q0 = c.bid * PCH, q1 = q0 + PCH < len ? q0 + PCH : len;
        for (int j = 1; j <= nj; ++j) {
            const double* uj = u + (uint64_t)j * ld;
            double s = 0;
            for (uint64_t q = q0 + c.tid; q < q1; q += c.nthr) s += xi[q] * uj[q];
            s = block_sum(c, red, s);
            if (c.tid == 0) partial[(uint64_t)j * nblk + c.bid] = s;
        }
    }
};

//! out[j] = sum_b partial[j * nblk + b] in block order, j = j0 .. j0 + nj - 1
struct RowsFinal {
    const double* partial;
    uint64_t nblk;
    int j0, nj;
    double* out;
    SANM_D void operator()(const rt::BlockCtx& c) const {
        for (int j = j0 + c.tid; j < j0 + nj; j += c.nthr) {
            double s = 0;
            for (uint64_t b = 0; b < nblk; ++b) s += partial[(uint64_t)j * nblk + b];
            out[j] = s;
        }
    }
};

//! u_i = x_i - sum_{j = j_first}^{i-1} a[j] u_j ; partial[b] = |u_i|^2 over the block
struct GsUpdate {
    const double* xi;
    double* u;
    uint64_t ld, len;
    int i, j_first;
    const double* a;  //!< device row a(i, .)
    double* partial;
    SANM_D void operator()(const rt::BlockCtx& c) const {
        double* red = reinterpret_cast<double*>(c.smem);
        const uint64_t q0 = c.bid * PCH, q1 = q0 + PCH < len ? q0 + PCH : len;
        double* ui = u + (uint64_t)i * ld;
        double s = 0;
        for (uint64_t q = q0 + c.tid; q < q1; q += c.nthr) {
            double v = xi[q];
            // u += u_j * (-a(i, j)), j ascending (pade.cpp:40)
            for (int j = j_first; j < i; ++j) v += u[(uint64_t)j * ld + q] * (-a[j]);
            ui[q] = v;
            s += v * v;
        }
        s = block_sum(c, red, s);
        if (c.tid == 0) partial[c.bid] = s;
    }
};

//! v *= f ; partial[b] = |v|^2 afterwards
struct ScaleNorm {
    double* v;
    uint64_t len;
    double f;
    double* partial;
    SANM_D void operator()(const rt::BlockCtx& c) const {
        double* red = reinterpret_cast<double*>(c.smem);
        const uint64_t q0 = c.bid * PCH, q1 = q0 + PCH < len ? q0 + PCH : len;
        double s = 0;
        for (uint64_t q = q0 + c.tid; q < q1; q += c.nthr) {
            const double x = v[q] * f;
            v[q] = x;
            s += x * x;
        }
        s = block_sum(c, red, s);
        if (c.tid == 0) partial[c.bid] = s;
    }
};

struct Coefs {
    double v[MAXC];
};

//! the two numerators of the range check and their norms (pade.cpp:137-147)
struct CheckPartial {
    const double* xs;
    uint64_t ld, len, nblk;
    int n;
    double a, f;
    Coefs sc_n, sc_l;
    double* partial;  //!< [2][nblk]: |pl f - pn|^2, |pn|^2
    SANM_D void operator()(const rt::BlockCtx& c) const {
        double* red = reinterpret_cast<double*>(c.smem);
        const uint64_t q0 = c.bid * PCH, q1 = q0 + PCH < len ? q0 + PCH : len;
        double sd = 0, sn = 0;
        for (uint64_t q = q0 + c.tid; q < q1; q += c.nthr) {
            double pn = 0, pl = 0;
            for (int i = n; i >= 1; --i) {
                const double x = xs[(uint64_t)i * ld + q];
                pn = pn * a + x * sc_n.v[i];
                if (i <= n - 1) pl = pl * a + x * sc_l.v[i];
            }
            const double dv = pl * f - pn;
            sd += dv * dv;
            sn += pn * pn;
        }
        sd = block_sum(c, red, sd);
        sn = block_sum(c, red, sn);
        if (c.tid == 0) {
            partial[c.bid] = sd;
            partial[nblk + c.bid] = sn;
        }
    }
};

//! out = (Horner of x_n .. x_1 with weights sc) * f + x_0 * w0
struct EvalKernel {
    const double* xs;
    uint64_t ld;
    int n;
    double a, f, w0;
    Coefs sc;
    double* out;
    SANM_D void operator()(size_t q) const {
        double s = 0;
        for (int i = n; i >= 1; --i) s = s * a + xs[(uint64_t)i * ld + q] * sc.v[i];
        out[q] = s * f + xs[q] * w0;
    }
};

//! plain Horner of the series: out = (((x_m a + x_{m-1}) a + ...) a + x_0
struct HornerKernel {
    const double* xs;
    uint64_t ld;
    int m;  //!< highest coefficient
    double a;
    double* out;
    SANM_D void operator()(size_t q) const {
        double s = xs[(uint64_t)m * ld + q];
        for (int i = m - 1; i >= 0; --i) s = s * a + xs[(uint64_t)i * ld + q];
        out[q] = s;
    }
};

//! per coefficient vector: partial[i * nblk + b] = |x_i|^2 over the block,
//! flag set on a non-finite entry
struct SeriesNormsPartial {
    const double* xs;
    uint64_t ld, len, nblk;
    int nx;
    double* partial;
    int* bad;
    SANM_D void operator()(const rt::BlockCtx& c) const {
        double* red = reinterpret_cast<double*>(c.smem);
        const uint64_t q0 = c.bid * PCH, q1 = q0 + PCH < len ? q0 + PCH : len;
        for (int i = 0; i < nx; ++i) {
            double s = 0;
            bool nf = false;
            for (uint64_t q = q0 + c.tid; q < q1; q += c.nthr) {
                const double x = xs[(uint64_t)i * ld + q];
                nf |= !(x - x == 0.0);  // inf or nan
                s += x * x;
            }
            if (nf) *bad = i + 1;
            s = block_sum(c, red, s);
            if (c.tid == 0) partial[(uint64_t)i * nblk + c.bid] = s;
        }
    }
};

//! out[i] = xs[i * ld + pos]
struct GatherColumn {
    const double* xs;
    uint64_t ld, pos;
    double* out;
    SANM_D void operator()(size_t i) const { out[i] = xs[(uint64_t)i * ld + pos]; }
};

class DevicePadeOps final : public PadeVectorOps {
    const double* m_xs;
    size_t m_nx, m_ld, m_len;
    uint64_t m_nblk;
    mutable rt::DBuf<double> m_u, m_partial, m_row, m_out;

public:
    DevicePadeOps(const double* d_xs, size_t nx, size_t ld, size_t len)
            : m_xs{d_xs}, m_nx{nx}, m_ld{ld}, m_len{len}, m_nblk{nblocks(len)} {
        sanm_check(nx <= (size_t)MAXC);
        m_partial.alloc((nx + 2) * m_nblk);
        m_row.alloc(nx + 2);
    }
    size_t nx() const override { return m_nx; }
    size_t len() const override { return m_len; }
    double t_coeff(size_t i) const override {
        double v;
        rt::d2h(&v, m_xs + i * m_ld + m_len - 1, sizeof(double));
        return v;
    }

    bool gram_schmidt(bool anm_cond, double* astore) override {
        const size_t nx = m_nx;
        const int n = (int)nx - 1;
        auto a = [&](int i, int j) -> double& { return astore[i * nx + j]; };
        constexpr double eps = std::numeric_limits<double>::epsilon();
        m_u.alloc(nx * m_ld);
        const size_t smem = PT * sizeof(double);
        std::vector<double> row(nx + 2);
        for (int i = 1; i <= n; ++i) {
            const double* xi = m_xs + (size_t)i * m_ld;
            if (i > 1) {
                rt::for_each_block<PT>(m_nblk, PT, smem,
                                       MultiDotPartial{xi, m_u.ptr(), m_ld, m_len, m_nblk,
                                                       i - 1, m_partial.ptr()});
                rt::for_each_block<PT>(1, PT, 0,
                                       RowsFinal{m_partial.ptr(), m_nblk, 1, i - 1,
                                                 m_row.ptr()});
            }
            // (under anm_cond the projection on u_1 is only checked, not
            // subtracted: pade.cpp:36-39)
            rt::for_each_block<PT>(m_nblk, PT, smem,
                                   GsUpdate{xi, m_u.ptr(), m_ld, m_len, i,
                                            anm_cond ? 2 : 1, m_row.ptr(),
                                            m_partial.ptr()});
            rt::for_each_block<PT>(1, PT, 0,
                                   RowsFinal{m_partial.ptr(), m_nblk, 0, 1,
                                             m_row.ptr() + nx});
            rt::d2h(row.data(), m_row.ptr(), (nx + 1) * sizeof(double));
            for (int j = 1; j < i; ++j) {
                a(i, j) = row[j];
                if (anm_cond && j == 1) {
                    sanm_check(std::fabs(a(i, j)) < 1e-4);
                    a(i, j) = 0;
                }
            }
            const double aii = std::sqrt(row[nx]);
            if (aii == 0) return false;
            a(i, i) = aii;
            double* ui = m_u.ptr() + (size_t)i * m_ld;
            rt::for_each_block<PT>(m_nblk, PT, smem,
                                   ScaleNorm{ui, m_len, 1 / std::max(aii, eps),
                                             m_partial.ptr()});
            if (aii < eps) {
                rt::for_each_block<PT>(1, PT, 0,
                                       RowsFinal{m_partial.ptr(), m_nblk, 0, 1,
                                                 m_row.ptr() + nx});
                double n2;
                rt::d2h(&n2, m_row.ptr() + nx, sizeof(double));
                rt::for_each_block<PT>(m_nblk, PT, smem,
                                       ScaleNorm{ui, m_len, 1 / std::sqrt(n2),
                                                 m_partial.ptr()});
            }
        }
        return true;
    }

    void check_norms(double a, const double* sc_n, const double* sc_l, int n,
                     double f, double* diff2, double* pn2) const override {
        CheckPartial k{m_xs, m_ld, m_len, m_nblk, n, a, f, {}, {}, m_partial.ptr()};
        for (int i = 1; i <= n; ++i) {
            k.sc_n.v[i] = sc_n[i];
            k.sc_l.v[i] = i <= n - 1 ? sc_l[i] : 0.0;
        }
        rt::for_each_block<PT>(m_nblk, PT, PT * sizeof(double), k);
        rt::for_each_block<PT>(1, PT, 0,
                               RowsFinal{m_partial.ptr(), m_nblk, 0, 2, m_row.ptr()});
        double out[2];
        rt::d2h(out, m_row.ptr(), sizeof(out));
        *diff2 = out[0];
        *pn2 = out[1];
    }

    void eval_xt(double a, const double* sc, int n, double f,
                 double* out) const override {
        m_out.alloc(m_len);
        EvalKernel k{m_xs, m_ld, n, a, f, 1.0, {}, m_out.ptr()};
        for (int i = 1; i <= n; ++i) k.sc.v[i] = sc[i];
        rt::for_each(m_len, k);
        rt::d2h(out, m_out.ptr(), m_len * sizeof(double));
    }
};

}  // namespace

std::unique_ptr<PadeVectorOps> make_device_pade_ops(const double* d_xs, size_t nx,
                                                    size_t ld, size_t len) {
    return std::make_unique<DevicePadeOps>(d_xs, nx, ld, len);
}

bool device_pade_supported(size_t nx) { return nx <= (size_t)MAXC; }

namespace series_dev {

void summary(const double* d_xs, size_t nx, size_t ld, size_t len,
             std::vector<double>& norm2, std::vector<double>& t_coeffs,
             int* first_bad) {
    sanm_check(nx <= (size_t)MAXC);
    const uint64_t nblk = nblocks(len);
    rt::DBuf<double> partial(nx * nblk), out(2 * nx);
    rt::DBuf<int> bad(1);
    bad.zero();
    rt::for_each_block<PT>(nblk, PT, PT * sizeof(double),
                           SeriesNormsPartial{d_xs, ld, len, nblk, (int)nx,
                                              partial.ptr(), bad.ptr()});
    rt::for_each_block<PT>(1, PT, 0,
                           RowsFinal{partial.ptr(), nblk, 0, (int)nx, out.ptr()});
    rt::for_each(nx, GatherColumn{d_xs, ld, len - 1, out.ptr() + nx});
    std::vector<double> h(2 * nx);
    rt::d2h(h.data(), out.ptr(), h.size() * sizeof(double));
    int hb = 0;
    rt::d2h(&hb, bad.ptr(), sizeof(int));
    norm2.assign(h.begin(), h.begin() + nx);
    t_coeffs.assign(h.begin() + nx, h.end());
    *first_bad = hb - 1;
}

void horner(const double* d_xs, size_t ncoef, size_t ld, size_t len, double a,
            double* out_host) {
    sanm_check(ncoef >= 1);
    rt::DBuf<double> out(len);
    rt::for_each(len, HornerKernel{d_xs, ld, (int)ncoef - 1, a, out.ptr()});
    rt::d2h(out_host, out.ptr(), len * sizeof(double));
}

}  // namespace series_dev

}  // namespace sanm_b200
