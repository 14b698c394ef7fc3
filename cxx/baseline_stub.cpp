// cxx/baseline_stub.cpp -- fea/main.cpp references the entry points of the
// Newton / Levenberg-Marquardt competitor (fea/baseline/main.h:35-48), which
// needs <mkl.h> and is not the ANM path (DESIGN.md "out of scope").  gpu_fea
// links these stand-ins; configs with a "baseline" key fail loudly.
#include "fea/baseline/main.h"

#include <stdexcept>

namespace baseline {
bool g_hessian_proj = true;
double g_hessian_diag_reg = 0;

namespace {
[[noreturn]] void unavailable() {
    throw std::runtime_error{
            "fea/baseline (Newton / LM on MKL) is not part of the B200 build"};
}
}  // namespace

Stat solve_energy_min(const IndexMat&, const CoordMat&, const CoordMat&,
                      const CoordMat*, const MaskMat&, const MaterialDesc&,
                      double, const IterCallback&) {
    unavailable();
}

Stat solve_force_equ_levmar(const IndexMat&, const CoordMat&, const CoordMat&,
                            const MaskMat&, const MaterialDesc&, double,
                            const IterCallback&) {
    unavailable();
}
}  // namespace baseline
