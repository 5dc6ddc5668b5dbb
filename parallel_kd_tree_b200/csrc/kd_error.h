// Per-thread error message behind kd_last_error().
#pragma once
#include <string>

namespace kd {
inline std::string& error_slot() {
    static thread_local std::string s;
    return s;
}
inline int fail(const std::string& msg, int code = 1) {
    error_slot() = msg;
    return code;
}
}  // namespace kd
