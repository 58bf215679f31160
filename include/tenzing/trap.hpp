// SIGINT/SIGABRT/SIGTERM hook: run a user function (dump partial results) and
// exit(1) (reference: src/trap.cpp:11-35, include/tenzing/trap.hpp:12-14).
#pragma once

#include <functional>

namespace tenzing {
using TrapHandler = std::function<void(int)>;
void register_handler(TrapHandler fn);
void unregister_handler();
} // namespace tenzing
