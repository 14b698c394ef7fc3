/* sanm_b200/csrc/remap.h -- device form of sanm::SparseLinearDesc
 * (libsanm/anm.h:24-85): a sparse linear map between flattened tensors, kept as
 * CSR.  apply() is the gather of libsanm/anm.cpp:55-75; both the node->element
 * map (remap_in) and the element->node assembly (remap_out, a scatter written
 * as a per-output gather, so it is deterministic and atomic-free) use it.
 *
 * Either side may be a batched per-item tensor in structure-of-arrays layout
 * ([e][npad], see taylor_dev.h); `e == 0` means a plain flat vector.
 */
#pragma once

#include "graph.h"

namespace sanm_b200 {

struct VecLayout {
    int e = 0;         //!< elements per item; 0 => flat vector
    int64_t npad = 0;  //!< SoA row pitch when e > 0
    SANM_HD inline size_t addr(size_t flat) const {
        return e ? (flat % e) * (size_t)npad + flat / e : flat;
    }
};

class Remap {
public:
    Remap(const Shape& out_shape, const Shape& in_shape, const uint64_t* rowptr,
          const uint64_t* idx, const double* coeff);
    static std::shared_ptr<Remap> make_identity(const Shape& in_shape,
                                                const Shape& out_shape);

    const Shape& out_shape() const { return m_out; }
    const Shape& in_shape() const { return m_in; }
    uint64_t nr_out() const { return m_out.total(); }
    uint64_t nr_in() const { return m_in.total(); }

    //! y = A x on the device; x, y are device pointers with the given layouts
    void apply(double* y, VecLayout ylay, const double* x, VecLayout xlay) const;

    //! y = A x for host vectors, staged in the library's static device scratch
    //! (no allocation; remap.cu, sanm_remap_apply_static)
    void apply_static(double* y_host, const double* x_host) const;

    //! host views (kept for K's symbolic analysis)
    const std::vector<uint32_t>& h_rowptr() const { return m_h_rowptr; }
    const std::vector<uint32_t>& h_idx() const { return m_h_idx; }
    const std::vector<double>& h_coeff() const { return m_h_coeff; }
    //! content hashes: sparsity structure only / structure + coefficients
    uint64_t hash_struct() const { return m_hash_struct; }
    uint64_t hash_full() const { return m_hash_full; }
    const uint32_t* d_rowptr() const { return m_rowptr.ptr(); }
    const uint32_t* d_idx() const { return m_idx.ptr(); }
    const double* d_coeff() const { return m_coeff.ptr(); }

private:
    Shape m_out, m_in;
    uint64_t m_hash_struct = 0, m_hash_full = 0;
    std::vector<uint32_t> m_h_rowptr, m_h_idx;
    std::vector<double> m_h_coeff;
    rt::DBuf<uint32_t> m_rowptr, m_idx;
    rt::DBuf<double> m_coeff;
};

}  // namespace sanm_b200
