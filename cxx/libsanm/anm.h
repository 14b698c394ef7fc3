// cxx/libsanm/anm.h -- forwarding header: lets sources written against the
// reference's `#include "libsanm/anm.h"` compile against the B200 host layer.
#pragma once
#include "sanm_host/solvers.hpp"
