/* oracle/shim/shim_impl.cpp -- TEST INFRASTRUCTURE (oracle build only).
 *
 * Stand-ins for the two MKL pieces the reference links against, so that the
 * UNMODIFIED reference sources under /root/reference compile into
 * oracle/_ref/:
 *   1. row-major naive cblas_dgemm / cblas_dgemm_batch_strided;
 *   2. sanm::SparseSolver (header libsanm/sparse_solver.h used as is) on the
 *      vendored Eigen::SparseLU instead of MKL PARDISO
 *      (replaces libsanm/sparse_solver.cpp, the only TU that includes
 *      mkl_pardiso.h / mkl_spblas.h).
 * Semantics kept from libsanm/sparse_solver.cpp: rows added in increasing
 * order per builder (:297-303), |coeff|<1e-9 dropped (:291-293), duplicates
 * summed (:250-278), l2 path solves (A'A+pI)x = A'b (:366-395,163-175).
 */
#include "libsanm/sparse_solver.h"
#include "libsanm/tensor_impl_helper.h"

#include <Eigen/Sparse>
#include <Eigen/SparseLU>
#include <cmath>
#include <cstring>
#include <mutex>

extern "C" void cblas_dgemm(CBLAS_LAYOUT layout, CBLAS_TRANSPOSE ta,
                            CBLAS_TRANSPOSE tb, int m, int n, int k,
                            double alpha, const double* a, int lda,
                            const double* b, int ldb, double beta, double* c,
                            int ldc) {
    if (layout != CblasRowMajor) abort();
    for (int i = 0; i < m; ++i) {
        for (int j = 0; j < n; ++j) {
            double s = 0;
            for (int p = 0; p < k; ++p) {
                double av = ta == CblasNoTrans ? a[i * lda + p] : a[p * lda + i];
                double bv = tb == CblasNoTrans ? b[p * ldb + j] : b[j * ldb + p];
                s += av * bv;
            }
            double& dst = c[i * ldc + j];
            dst = beta == 0 ? alpha * s : alpha * s + beta * dst;
        }
    }
}

extern "C" void cblas_dgemm_batch_strided(
        CBLAS_LAYOUT layout, CBLAS_TRANSPOSE ta, CBLAS_TRANSPOSE tb, int m,
        int n, int k, double alpha, const double* a, int lda, int stridea,
        const double* b, int ldb, int strideb, double beta, double* c, int ldc,
        int stridec, int batch) {
    for (int i = 0; i < batch; ++i) {
        cblas_dgemm(layout, ta, tb, m, n, k, alpha, a + (size_t)i * stridea, lda,
                    b + (size_t)i * strideb, ldb, beta, c + (size_t)i * stridec,
                    ldc);
    }
}

using namespace sanm;

namespace {
int g_nr_threads = 0;
int g_verbosity = 0;
using SpMat = Eigen::SparseMatrix<double, Eigen::ColMajor, int>;
using Trip = Eigen::Triplet<double, int>;
SpMat g_last_A;
}  // namespace

class SparseSolver::Impl final : public NonCopyable {
public:
    class Builder final : public SparseSolver::SparseMatBuilder {
    public:
        Impl* owner;
        size_t offset, prev;
        std::vector<Trip> trips;
        Builder(Impl* o, size_t off) : owner{o}, offset{off}, prev{off} {}
        void add_constraint(size_t cidx, size_t xidx, fp_t coeff) override {
            sanm_assert(!owner->prepared);
            sanm_assert(std::isfinite(coeff), "coeff[%zu,%zu]=%g", cidx, xidx,
                        coeff);
            if (std::fabs(coeff) < 1e-9) return;
            sanm_assert(cidx < owner->n && xidx < owner->n);
            if (cidx != prev) {
                sanm_assert(cidx == prev + 1,
                            "constraints not added in order: prev=%zu cur=%zu",
                            prev, cidx);
                prev = cidx;
            }
            trips.emplace_back((int)cidx, (int)xidx, coeff);
        }
    };
    size_t n;
    bool prepared = false;
    fp_t l2 = 0;
    std::vector<std::unique_ptr<Builder>> builders;
    SpMat A, At;
    Eigen::SparseLU<SpMat, Eigen::COLAMDOrdering<int>> lu;
    std::mutex mtx;

    explicit Impl(size_t n_) : n{n_} {}
};

SparseSolver::SparseSolver(size_t nr_xs) : m_pimpl{new Impl{nr_xs}} {}
SparseSolver::~SparseSolver() = default;
void SparseSolver::set_num_threads(int nr) { g_nr_threads = nr; }
int SparseSolver::get_num_threads() {
    return g_nr_threads ? g_nr_threads : sanm::get_num_threads();
}
void SparseSolver::set_verbosity(int v) { g_verbosity = v; }

SparseSolver::SparseMatBuilder* SparseSolver::make_builder(size_t off) {
    std::lock_guard<std::mutex> lg{m_pimpl->mtx};
    m_pimpl->builders.emplace_back(new Impl::Builder{m_pimpl.get(), off});
    return m_pimpl->builders.back().get();
}

SparseSolver& SparseSolver::prepare(fp_t l2_penalty) {
    SANM_SCOPED_PROFILER("sparse_prep");
    ScopedAllowMalloc allow;
    auto& p = *m_pimpl;
    sanm_assert(!p.prepared);
    p.prepared = true;
    p.l2 = l2_penalty;
    std::vector<Trip> all;
    size_t tot = 0;
    for (auto& b : p.builders) tot += b->trips.size();
    all.reserve(tot);
    for (auto& b : p.builders) {
        all.insert(all.end(), b->trips.begin(), b->trips.end());
    }
    p.builders.clear();
    p.A.resize(p.n, p.n);
    p.A.setFromTriplets(all.begin(), all.end());  // sums duplicates
    p.A.makeCompressed();
    g_last_A = p.A;
    if (l2_penalty) {
        p.At = p.A.transpose();
        SpMat M = p.At * p.A;
        for (size_t i = 0; i < p.n; ++i) M.coeffRef(i, i) += l2_penalty;
        M.makeCompressed();
        p.lu.compute(M);
    } else {
        p.lu.compute(p.A);
    }
    sanm_assert(p.lu.info() == Eigen::Success, "SparseLU failed: %s",
                p.lu.lastErrorMessage().c_str());
    return *this;
}

TensorND SparseSolver::solve(const TensorND& b) const {
    SANM_SCOPED_PROFILER("sparse_solve");
    ScopedAllowMalloc allow;
    auto& p = *m_pimpl;
    sanm_assert(p.prepared && b.shape().total_nr_elems() == p.n);
    const fp_t* bp = b.ptr();
    for (size_t i = 0; i < p.n; ++i) {
        sanm_assert(std::isfinite(bp[i]), "b[%zu]=%g", i, bp[i]);
    }
    Eigen::Map<const Eigen::VectorXd> bv{bp, (Eigen::Index)p.n};
    Eigen::VectorXd x;
    if (p.l2) {
        Eigen::VectorXd atb = p.At * bv;
        x = p.lu.solve(atb);
    } else {
        x = p.lu.solve(bv);
    }
    TensorND ret{b.shape()};
    memcpy(ret.woptr(), x.data(), sizeof(double) * p.n);
    return ret;
}

TensorND SparseSolver::apply(const TensorND& x) const {
    ScopedAllowMalloc allow;
    auto& p = *m_pimpl;
    sanm_assert(p.prepared && x.shape().total_nr_elems() == p.n);
    Eigen::Map<const Eigen::VectorXd> xv{x.ptr(), (Eigen::Index)p.n};
    Eigen::VectorXd y = p.A * xv;
    TensorND ret{x.shape()};
    memcpy(ret.woptr(), y.data(), sizeof(double) * p.n);
    return ret;
}

fp_t SparseSolver::coeff_l2() const {
    ScopedAllowMalloc allow;
    return m_pimpl->A.norm();
}

void SparseSolver::dump(const TensorND&, FILE*) const {}

namespace sanm_oracle {
//! CSR (row-major, sorted columns) of the matrix given to the most recent
//! SparseSolver::prepare(); used by the golden-vector dumper / ref_capi only
void last_prepared_csr(std::vector<int>& ia, std::vector<int>& ja,
                       std::vector<double>& a) {
    Eigen::SparseMatrix<double, Eigen::RowMajor, int> R = g_last_A;
    R.makeCompressed();
    ia.assign(R.outerIndexPtr(), R.outerIndexPtr() + R.rows() + 1);
    ja.assign(R.innerIndexPtr(), R.innerIndexPtr() + R.nonZeros());
    a.assign(R.valuePtr(), R.valuePtr() + R.nonZeros());
}
}  // namespace sanm_oracle
