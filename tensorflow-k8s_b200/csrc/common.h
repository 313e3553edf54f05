// common.h -- error plumbing shared by every translation unit of libb200ps.so.
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>

#include "b200ps.h"

namespace b200ps {

// thread-local message behind b200ps_last_error()
std::string& last_error();
int fail(int status, const char* fmt, ...) __attribute__((format(printf, 2, 3)));

}  // namespace b200ps

#define B200PS_REQUIRE(cond, status, ...)                        \
    do {                                                         \
        if (!(cond)) return ::b200ps::fail((status), __VA_ARGS__); \
    } while (0)
