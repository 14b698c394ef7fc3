/* sanm_b200/csrc/sparse_symbolic.h -- host-side symbolic analysis for the
 * multifrontal solver (see sparse_lu.h). */
#pragma once

#include <cstdint>
#include <vector>

namespace sanm_b200 {

struct Symbolic {
    uint32_t n = 0, nsn = 0;
    std::vector<uint32_t> perm;   //!< new index -> original index
    std::vector<uint32_t> iperm;  //!< original index -> new index

    //! supernode s owns new columns [c0, c0+nc); its front is nr x nr
    std::vector<uint32_t> sn_c0, sn_nc, sn_nr;
    std::vector<int32_t> sn_parent;  //!< -1 for roots
    //! row structure of front s (new indices, sorted; the first nc are its own
    //! columns): structs[soff[s] .. soff[s]+nr)
    std::vector<uint64_t> soff;
    std::vector<uint32_t> structs;
    //! column-major nr x nr front of s starts at foff[s] doubles
    std::vector<uint64_t> foff;
    uint64_t front_total = 0;

    //! children of s: child_list[child_ptr[s] .. child_ptr[s+1])
    std::vector<uint32_t> child_ptr, child_list;
    //! for a non-root s: inv[inv_off[s] + i] = index among s's update rows
    //! (0-based after its nc pivot rows) that maps to row i of the parent
    //! front, or -1
    std::vector<uint64_t> inv_off;
    std::vector<int32_t> inv;

    //! level schedule: fronts of level l are level_list[level_ptr[l] ..
    //! level_ptr[l+1]); children are always in a lower level
    std::vector<uint32_t> level_ptr, level_list;

    //! destination (offset into the front array) of every CSR entry
    std::vector<uint64_t> a_dest;

    double flops = 0;
    uint64_t factor_entries = 0;
};

/*!
 * \param rowptr,colidx CSR pattern (n rows), columns sorted and unique; it is
 *      symmetrised internally
 * \param node_keys optional ordering hint, one value per unknown (or null):
 *      where three unknowns share one adjacency (the three coordinates of a
 *      mesh vertex) their values are read as that node's position, and
 *      axis-aligned planes join the candidate bisections of the nested
 *      dissection.  A hint never makes the result invalid; candidates are
 *      compared by separator weight.
 */
Symbolic symbolic_analyze(uint32_t n, const uint32_t* rowptr,
                          const uint32_t* colidx,
                          const double* node_keys = nullptr);

}  // namespace sanm_b200
