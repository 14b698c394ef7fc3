// cxx/libsanm/unary_polynomial.h -- forwarding header: lets sources written against the
// reference's `#include "libsanm/unary_polynomial.h"` compile against the B200 host layer.
#pragma once
#include "sanm_host/solvers.hpp"
