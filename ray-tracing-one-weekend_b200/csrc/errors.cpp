// errors.cpp -- storage behind rtow_last_error().
#include "errors.h"

namespace rtow {
std::string &last_error_ref() {
    static thread_local std::string msg;
    return msg;
}
}  // namespace rtow

extern "C" const char *rtow_last_error(void) { return rtow::last_error_ref().c_str(); }
