#include "capi_util.h"
namespace sanm_b200 {
namespace capi {
std::string& last_error() {
    static thread_local std::string err;
    return err;
}
}  // namespace capi
}  // namespace sanm_b200
