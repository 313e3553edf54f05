// common.cpp -- thread-local error state (include/b200ps.h "Conventions").
#include "common.h"

namespace b200ps {

std::string& last_error() {
    static thread_local std::string msg;
    return msg;
}

int fail(int status, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    last_error() = buf;
    return status;
}

}  // namespace b200ps

extern "C" const char* b200ps_last_error(void) { return b200ps::last_error().c_str(); }
extern "C" int b200ps_version(void) { return 1; }
