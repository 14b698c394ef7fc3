/* oracle/shim/presize_pool.cpp -- TEST INFRASTRUCTURE (linked into ref_fea only).
 *
 * The reference's shared zero / one constant pools (TensorStorage::constant<>,
 * libsanm/tensor.cpp:251-293) grow on demand without synchronisation, which
 * races when its worker threads run.  SURVEY 8d's
 * remedy for multi-thread TIMING runs, without touching a reference source:
 * create and fill one tensor of the largest size first; afterwards the pools
 * never reallocate.  This translation unit does exactly that before main()
 * when SANM_REF_PRESIZE_ELEMS is set (bench.py sets it for the sys-mtN legs).
 * It is linked last, so its initialiser runs after the reference's statics. */
#include <cstdlib>

#include "libsanm/tensor.h"

namespace {
struct PresizePool {
    PresizePool() {
        const char* e = getenv("SANM_REF_PRESIZE_ELEMS");
        if (!e) return;
        const size_t n = strtoull(e, nullptr, 10);
        if (!n) return;
        sanm::TensorND pre{sanm::TensorShape{n}};
        pre.fill_with_inplace(0);
        pre.fill_with_inplace(1);
    }
} g_presize_pool;
}  // namespace
