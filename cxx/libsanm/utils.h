// cxx/libsanm/utils.h -- forwarding header: lets sources written against the
// reference's `#include "libsanm/utils.h"` compile against the B200 host layer.
#pragma once
#include "sanm_host/base.hpp"
