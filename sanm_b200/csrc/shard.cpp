/* sanm_b200/csrc/shard.cpp -- restriction of a graph and of the two remaps to
 * a contiguous range of batch items (tets), for the multi-GPU path (comm.h).
 * Mirrors what each worker of the reference's ParallelTaylorCoeffProp sees:
 * its slice of the input (libsanm/symbolic.cpp:525-533) and shard-aware
 * constants (libsanm/oprs/misc.cpp:48-73). */
#include "shard.h"

namespace sanm_b200 {

std::shared_ptr<Graph> shard_graph(const Graph& g, uint64_t N, uint64_t lo,
                                   uint64_t hi) {
    sanm_check(lo < hi && hi <= N);
    auto ret = std::make_shared<Graph>(g);
    for (auto& op : ret->ops) {
        if (op.type != taylor::OP_CONSTANT) continue;
        sanm_check(op.cshape.dim[0] == N,
                   "constant batch %s does not match the sharded batch %zu",
                   op.cshape.str().c_str(), (size_t)N);
        const uint64_t per = op.cshape.per_item();
        std::vector<double> sl(op.cdata.begin() + lo * per,
                               op.cdata.begin() + hi * per);
        op.cdata.swap(sl);
        op.cshape.dim[0] = hi - lo;
    }
    return ret;
}

std::shared_ptr<Remap> shard_remap_rows(const Remap& r, uint64_t N, uint64_t lo,
                                        uint64_t hi) {
    sanm_check(lo < hi && hi <= N && r.out_shape().dim[0] == N);
    const uint64_t per = r.nr_out() / N;
    const auto& rp = r.h_rowptr();
    const auto& ix = r.h_idx();
    const auto& cf = r.h_coeff();
    const uint64_t r0 = lo * per, r1 = hi * per;
    std::vector<uint64_t> rowptr(r1 - r0 + 1), idx(rp[r1] - rp[r0]);
    std::vector<double> coeff(cf.begin() + rp[r0], cf.begin() + rp[r1]);
    for (uint64_t i = r0; i <= r1; ++i) rowptr[i - r0] = rp[i] - rp[r0];
    for (uint64_t p = rp[r0]; p < rp[r1]; ++p) idx[p - rp[r0]] = ix[p];
    Shape os = r.out_shape();
    os.dim[0] = hi - lo;
    return std::make_shared<Remap>(os, r.in_shape(), rowptr.data(), idx.data(),
                                   coeff.data());
}

std::shared_ptr<Remap> shard_remap_cols(const Remap& r, uint64_t N, uint64_t lo,
                                        uint64_t hi) {
    sanm_check(lo < hi && hi <= N && r.in_shape().dim[0] == N);
    const uint64_t per = r.nr_in() / N;
    const auto& rp = r.h_rowptr();
    const auto& ix = r.h_idx();
    const auto& cf = r.h_coeff();
    const uint64_t c0 = lo * per, c1 = hi * per, nrow = r.nr_out();
    std::vector<uint64_t> rowptr(nrow + 1, 0), idx;
    std::vector<double> coeff;
    for (uint64_t i = 0; i < nrow; ++i) {
        for (uint32_t p = rp[i]; p < rp[i + 1]; ++p) {
            if (ix[p] >= c0 && ix[p] < c1) {
                idx.push_back(ix[p] - c0);
                coeff.push_back(cf[p]);
            }
        }
        rowptr[i + 1] = idx.size();
    }
    Shape is = r.in_shape();
    is.dim[0] = hi - lo;
    // Remap needs non-null arrays even when the shard touches nothing
    idx.push_back(0);
    coeff.push_back(0);
    return std::make_shared<Remap>(r.out_shape(), is, rowptr.data(), idx.data(),
                                   coeff.data());
}

}  // namespace sanm_b200
