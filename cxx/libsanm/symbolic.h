// cxx/libsanm/symbolic.h -- forwarding header: lets sources written against the
// reference's `#include "libsanm/symbolic.h"` compile against the B200 host layer.
#pragma once
#include "sanm_host/graph.hpp"
