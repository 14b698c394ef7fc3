/* sanm_b200/csrc/fused_io.h -- the two mesh remaps of `fea` folded into the
 * fused element kernels (fused.cu):
 *
 *   remap_in  (MeshShapeMatTrans, fea/mesh_template.h:19-111): every entry of
 *     the per-tet shape matrix is  -x[a] + x[b] (+ delta * t); kept per tet as
 *     a fixed-width list of signed unit references (4 B each) plus at most one
 *     general (index, coefficient) reference, read coalesced by the kernel,
 *     which then gathers from the L2-resident nodal vector;
 *   remap_out (MeshForceOutputTrans, fea/mesh_template.h:131-161): the nodal
 *     force is  sum_tets P n_{tet,vertex}; kept per tet as <= 4 vertex slots
 *     (area-weighted normal, 3 doubles) with <= 3 output rows each.  Assembly
 *     is a deterministic two-level segmented reduction: a block of TILE tets
 *     reduces its 12*TILE contributions in shared memory over a precomputed
 *     contributor list into one partial sum per distinct row of the block,
 *     and a second kernel adds the partials of each row in fixed order.  No
 *     atomics: results are bitwise reproducible.
 *
 * Both are derived from the generic CSR of a SparseLinearDesc
 * (libsanm/anm.h:24-85) by analyse(); descriptors without this structure
 * (e.g. the random maps of tests/symbolic.cpp:428-523) keep the staged path:
 * Remap::apply before / after the element kernels.
 */
#pragma once

#include "fused_dev.h"
#include "remap.h"

#include <memory>

namespace sanm_b200 {

struct FusedIO {
    uint64_t n_tets = 0, nin = 0, nout = 0;
    uint64_t hash = 0;  //!< content hash of the two maps it was derived from
    int64_t npad = 0;
    // ---- gather
    int wu = 0;   //!< unit references per shape-matrix entry (<= 2)
    bool has_general = false;
    rt::DBuf<uint32_t> uidx;  //!< [wu][9][npad]: (index << 1 | negative), ~0u = none
    rt::DBuf<uint32_t> gidx;  //!< [9][npad], ~0u = none
    rt::DBuf<double> gcoef;   //!< [9][npad]
    // ---- assembly
    rt::DBuf<double> nrm;      //!< [4][3][npad]: weights of vertex slot d
    rt::DBuf<uint16_t> lrow;   //!< [12][npad]: block-local row of (d, c), 0xFFFF = none
    rt::DBuf<uint32_t> blk_ptr;  //!< [ntiles + 1] first partial of each block
    rt::DBuf<uint32_t> con_ptr;  //!< [nr_partial + 1]
    rt::DBuf<uint16_t> con_idx;  //!< contributor = slot * TILE + lane
    rt::DBuf<uint32_t> fin_ptr;  //!< [nout + 1]
    rt::DBuf<uint32_t> fin_idx;  //!< partial ids of each output row
    rt::DBuf<double> partial;    //!< [nr_partial]
    uint64_t nr_partial = 0;
    //! output rows that receive any contribution from these tets lie in
    //! [row_lo, row_hi): the whole vector for a full mesh, this rank's slab
    //! (plus its interface planes) for a pre-sharded one
    uint64_t row_lo = 0, row_hi = 0;
    // ---- host copies for the element-wise K assembly (k_assemble.cu)
    std::vector<uint32_t> h_uidx, h_gidx;  //!< as uidx / gidx above
    std::vector<uint32_t> h_grow;          //!< [12][npad] global row of (d, c)

    //! nullptr when either map lacks the structure described above
    static std::unique_ptr<FusedIO> analyse(const Remap& rin, const Remap& rout,
                                            uint64_t n_tets);
    //! b[row] = sum of its partials (second level of the assembly).
    //! local_only: touch rows [row_lo, row_hi) only -- the work of a rank is
    //! then proportional to ITS tets, not to the global mesh (rows outside
    //! receive nothing from this rank; with distributed nodal vectors nobody
    //! reads them here)
    void finalize(double* b, bool local_only = false) const;
    //! bytes of index/weight tables one order step reads
    uint64_t table_bytes() const;
};

}  // namespace sanm_b200
