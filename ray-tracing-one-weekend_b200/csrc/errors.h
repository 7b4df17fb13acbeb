// errors.h -- status codes + last-error text shared by the C entry points.
// The reference reports problems by printing and carrying on (bvhNode.h:54-56,103-105;
// scenes.h:202-207 silently swaps a glass sphere in).  Here every entry point returns a
// status and rtow_last_error() holds the message of the last failure on this thread.
#pragma once
#include <string>
#include "../../include/rtow_b200.h"

namespace rtow {
std::string &last_error_ref();
inline int set_error(int code, const std::string &msg) {
    last_error_ref() = msg;
    return code;
}
}  // namespace rtow
