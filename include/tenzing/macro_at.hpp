// Logging / error macros (reference: include/tenzing/macro_at.hpp:13-40).
// The rank prefix comes from a process-wide value set by the communicator
// (no MPI query in a macro), and log verbosity is a run-time switch so hot
// search loops are silent by default.
#pragma once

#include <iostream>
#include <sstream>
#include <stdexcept>
#include <string>

namespace tenzing {
namespace log {
int &rank_ref();      // rank shown in the prefix (default 0)
int &verbosity_ref(); // 0 = errors only, 1 = info, 2 = debug
inline void set_rank(int r) { rank_ref() = r; }
inline void set_verbosity(int v) { verbosity_ref() = v; }
} // namespace log
} // namespace tenzing

#define TENZING_STRINGIFY_(x) #x
#define TENZING_STRINGIFY(x) TENZING_STRINGIFY_(x)
#define AT __FILE__ ":" TENZING_STRINGIFY(__LINE__)

#define TENZING_LOG_AT(level, msg)                                                                 \
  do {                                                                                             \
    if (::tenzing::log::verbosity_ref() >= (level)) {                                              \
      std::ostringstream tz_log_ss_;                                                               \
      tz_log_ss_ << "[" << ::tenzing::log::rank_ref() << "] " << AT << ": " << msg << "\n";        \
      std::cerr << tz_log_ss_.str();                                                               \
    }                                                                                              \
  } while (0)

// always printed
#define STDERR(msg) TENZING_LOG_AT(0, msg)
#define TENZING_INFO(msg) TENZING_LOG_AT(1, msg)
#define TENZING_DEBUG(msg) TENZING_LOG_AT(2, msg)

#define THROW_RUNTIME(msg)                                                                         \
  do {                                                                                             \
    std::ostringstream tz_throw_ss_;                                                               \
    tz_throw_ss_ << AT << ": " << msg;                                                             \
    throw std::runtime_error(tz_throw_ss_.str());                                                  \
  } while (0)

#if defined(__GNUC__)
#define TENZING_LIKELY(x) __builtin_expect(!!(x), 1)
#define TENZING_UNLIKELY(x) __builtin_expect(!!(x), 0)
#else
#define TENZING_LIKELY(x) (x)
#define TENZING_UNLIKELY(x) (x)
#endif
