/* oracle/shim/baseline_stub.cpp -- TEST INFRASTRUCTURE (oracle build only).
 * fea/baseline/* (the Newton / Levenberg-Marquardt competitor) needs <mkl.h>
 * and is out of scope; fea/main.cpp references its entry points
 * (fea/baseline/main.h:35-48), so provide throwing stubs. */
#include "fea/baseline/main.h"
#include <stdexcept>

namespace baseline {
bool g_hessian_proj = true;
double g_hessian_diag_reg = 0;

Stat solve_energy_min(const IndexMat&, const CoordMat&, const CoordMat&,
                      const CoordMat*, const MaskMat&, const MaterialDesc&,
                      double, const IterCallback&) {
    throw std::runtime_error{"baseline solver not built in the oracle (needs MKL)"};
}

Stat solve_force_equ_levmar(const IndexMat&, const CoordMat&, const CoordMat&,
                            const MaskMat&, const MaterialDesc&, double,
                            const IterCallback&) {
    throw std::runtime_error{"baseline solver not built in the oracle (needs MKL)"};
}
}  // namespace baseline
