// cxx/libsanm/sparse_solver.h -- forwarding header: lets sources written against the
// reference's `#include "libsanm/sparse_solver.h"` compile against the B200 host layer.
#pragma once
#include "sanm_host/solvers.hpp"
