/* sanm_b200/csrc/comm.cu -- see comm.h. */
#include "comm.h"
#include "capi_util.h"
#include "prof.h"

#include <algorithm>

#ifdef SANM_WITH_NCCL
#include <nccl.h>
#endif

namespace sanm_b200 {
namespace comm {

namespace {
int g_world = 1, g_rank = 0;
sanm_allreduce_fn g_cb = nullptr;
std::vector<double> g_stage;
#ifdef SANM_WITH_NCCL
ncclComm_t g_nccl = nullptr;
#define SANM_NCCL_CHECK(x)                                                   \
    do {                                                                     \
        ncclResult_t r__ = (x);                                              \
        if (r__ != ncclSuccess) {                                            \
            sanm_throw(SANM_ERR_CUDA, "NCCL error at %s:%d: %s", __FILE__,   \
                       __LINE__, ncclGetErrorString(r__));                   \
        }                                                                    \
    } while (0)
#endif
}  // namespace

int world() { return g_world; }
int rank() { return g_rank; }

void allreduce_sum(double* dev, size_t n) {
    if (g_world == 1 || !n) return;
    SANM_PROF("comm_allreduce");
#ifdef SANM_WITH_NCCL
    if (g_nccl) {
        rt::count_launch();
        SANM_NCCL_CHECK(ncclAllReduce(dev, dev, n, ncclDouble, ncclSum, g_nccl,
                                      rt::stream()));
        return;
    }
#endif
    sanm_check(g_cb, "multi-rank run without a communicator");
#ifdef SANM_HOSTEMU
    g_cb(dev, n);
#else
    g_stage.resize(n);
    rt::d2h(g_stage.data(), dev, n * sizeof(double));
    g_cb(g_stage.data(), n);
    rt::h2d(dev, g_stage.data(), n * sizeof(double));
#endif
}

namespace {
struct HaloAdd {
    double* v;
    const double* recv;
    bool recv_first;
    SANM_D void operator()(size_t i) const {
        v[i] = recv_first ? recv[i] + v[i] : v[i] + recv[i];
    }
};
}  // namespace

void halo_setup(Halo& h, uint64_t lo, uint64_t hi, uint64_t n) {
    h.lo = lo;
    h.hi = hi;
    h.n = n;
    h.peers.clear();
    h.ready = true;
    if (g_world == 1) return;
#ifdef SANM_WITH_NCCL
    if (g_nccl) {
        std::vector<uint64_t> mine{lo, hi}, all((size_t)2 * g_world);
        rt::DBuf<uint64_t> dm, da;
        dm.upload(mine);
        da.alloc(all.size());
        SANM_NCCL_CHECK(ncclAllGather(dm.ptr(), da.ptr(), 2, ncclUint64, g_nccl,
                                      rt::stream()));
        da.download(all.data(), all.size());
        size_t off = 0;
        for (int q = 0; q < g_world; ++q) {
            if (q == g_rank) continue;
            const uint64_t olo = std::max(lo, all[2 * q]);
            const uint64_t ohi = std::min(hi, all[2 * q + 1]);
            if (olo >= ohi) continue;
            h.peers.push_back({q, olo, ohi - olo, off});
            off += ohi - olo;
        }
        // The sum of a shared row is formed pairwise with the neighbour (lower
        // rank first), which is the same on both owners only if a row has at
        // most two of them: true for slabs at least two vertex planes thick.
        // Refuse anything else rather than break the bitwise-identical-owners
        // guarantee.
        for (size_t a = 0; a < h.peers.size(); ++a) {
            for (size_t c = a + 1; c < h.peers.size(); ++c) {
                const auto &pa = h.peers[a], &pc = h.peers[c];
                sanm_check(pa.lo + pa.cnt <= pc.lo || pc.lo + pc.cnt <= pa.lo,
                           "halo exchange: rows [%zu, %zu) are shared by more than "
                           "two ranks (%d, %d and this one); use the all-reduce "
                           "exchange for this partition",
                           (size_t)std::max(pa.lo, pc.lo),
                           (size_t)std::min(pa.lo + pa.cnt, pc.lo + pc.cnt), pa.rank,
                           pc.rank);
            }
        }
        h.scratch.alloc(off ? off : 1);
    }
#endif
}

void halo_sum(double* dev, Halo& h) {
    sanm_check(h.ready, "halo_sum() before halo_setup()");
    if (g_world == 1) return;
    SANM_PROF("comm_halo");
#ifdef SANM_WITH_NCCL
    if (g_nccl) {
        if (h.peers.empty()) return;
        rt::count_launch();
        SANM_NCCL_CHECK(ncclGroupStart());
        for (auto& p : h.peers) {
            SANM_NCCL_CHECK(ncclSend(dev + p.lo, p.cnt, ncclDouble, p.rank, g_nccl,
                                     rt::stream()));
            SANM_NCCL_CHECK(ncclRecv(h.scratch.ptr() + p.scratch_off, p.cnt,
                                     ncclDouble, p.rank, g_nccl, rt::stream()));
        }
        SANM_NCCL_CHECK(ncclGroupEnd());
        // peers are in ascending rank order: lower ranks' shares come first
        for (auto& p : h.peers) {
            rt::for_each(p.cnt, HaloAdd{dev + p.lo, h.scratch.ptr() + p.scratch_off,
                                        p.rank < g_rank});
        }
        return;
    }
#endif
    // callback transport (CPU tests): the sum over all ranks is a superset of
    // the halo sum once the rows this rank does not contribute to are zero
    // (in halo mode they are not written by the assembly)
    if (h.lo) rt::dzero(dev, h.lo * sizeof(double));
    if (h.hi < h.n) rt::dzero(dev + h.hi, (h.n - h.hi) * sizeof(double));
    allreduce_sum(dev, h.n);
}

}  // namespace comm
}  // namespace sanm_b200

using namespace sanm_b200;

extern "C" {

int sanm_comm_unique_id(char* out128) {
    return capi::guard([&] {
#ifdef SANM_WITH_NCCL
        static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
        ncclUniqueId id;
        SANM_NCCL_CHECK(ncclGetUniqueId(&id));
        memcpy(out128, &id, 128);
#else
        (void)out128;
        sanm_throw(SANM_ERR_GENERIC, "built without NCCL");
#endif
    });
}

int sanm_comm_init_nccl(const char* id128, int rank, int world) {
    return capi::guard([&] {
#ifdef SANM_WITH_NCCL
        sanm_check(world >= 1 && rank >= 0 && rank < world);
        rt::check_device();
        if (comm::g_nccl) {
            ncclCommDestroy(comm::g_nccl);
            comm::g_nccl = nullptr;
        }
        ncclUniqueId id;
        memcpy(&id, id128, 128);
        SANM_NCCL_CHECK(ncclCommInitRank(&comm::g_nccl, world, id, rank));
        comm::g_world = world;
        comm::g_rank = rank;
#else
        (void)id128;
        (void)rank;
        (void)world;
        sanm_throw(SANM_ERR_GENERIC, "built without NCCL");
#endif
    });
}

int sanm_comm_init_callback(sanm_allreduce_fn fn, int rank, int world) {
    return capi::guard([&] {
        sanm_check(fn && world >= 1 && rank >= 0 && rank < world);
        comm::g_cb = fn;
        comm::g_world = world;
        comm::g_rank = rank;
    });
}

int sanm_comm_finalize(void) {
    return capi::guard([&] {
#ifdef SANM_WITH_NCCL
        if (comm::g_nccl) {
            rt::sync();
            ncclCommDestroy(comm::g_nccl);
            comm::g_nccl = nullptr;
        }
#endif
        comm::g_cb = nullptr;
        comm::g_world = 1;
        comm::g_rank = 0;
    });
}

int sanm_comm_rank(int* rank, int* world) {
    *rank = comm::g_rank;
    *world = comm::g_world;
    return 0;
}

/* test / diagnostics entry: in-place sum over ranks of a host vector */
int sanm_comm_allreduce_host(double* buf, uint64_t n) {
    return capi::guard([&] {
        rt::DBuf<double> d;
        d.upload(buf, n);
        comm::allreduce_sum(d.ptr(), n);
        d.download(buf, n);
    });
}
}
