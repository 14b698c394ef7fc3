/* sanm_b200/csrc/pade_dev.h -- the vector passes of the Pade approximant and
 * of the range estimation on the DEVICE-resident series of the ANM drivers
 * (libsanm/pade.cpp:30-55,137-147,181-224, anm.cpp:117-191).
 *
 * The K+1 series vectors stay in HBM: Gram-Schmidt projections, the two
 * numerators of the range check with their norms and the evaluation are
 * kernels with fixed-tree reductions (blocks of 2048 elements, partials summed
 * in block order: bitwise reproducible); only scalars travel to the host,
 * where pade.cpp takes every decision as before.
 */
#pragma once

#include "pade.h"
#include "rt.h"

namespace sanm_b200 {

//! d_xs: device array, coefficient vector i at d_xs + i * ld, length len
std::unique_ptr<PadeVectorOps> make_device_pade_ops(const double* d_xs, size_t nx,
                                                    size_t ld, size_t len);
bool device_pade_supported(size_t nx);

namespace series_dev {
//! squared norms and last elements (the t coefficients) of the nx vectors;
//! first_bad = index of a vector holding a non-finite value, or -1
void summary(const double* d_xs, size_t nx, size_t ld, size_t len,
             std::vector<double>& norm2, std::vector<double>& t_coeffs,
             int* first_bad);
//! out_host = sum_i x_i a^i (Horner from the highest coefficient)
void horner(const double* d_xs, size_t ncoef, size_t ld, size_t len, double a,
            double* out_host);
}  // namespace series_dev

}  // namespace sanm_b200
