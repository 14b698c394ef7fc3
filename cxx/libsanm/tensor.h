// cxx/libsanm/tensor.h -- forwarding header: lets sources written against the
// reference's `#include "libsanm/tensor.h"` compile against the B200 host layer.
#pragma once
#include "sanm_host/tensor.hpp"
