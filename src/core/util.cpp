// L0 odds and ends: logging state, runs test, counters, signal trap,
// reproducibility header, init notice.
#include <csignal>
#include <cstdlib>
#include <limits>

#include "tenzing/counters.hpp"
#include "tenzing/init.hpp"
#include "tenzing/json.hpp"
#include "tenzing/macro_at.hpp"
#include "tenzing/numeric.hpp"
#include "tenzing/randomness.hpp"
#include "tenzing/reproduce.hpp"
#include "tenzing/trap.hpp"
#include "tenzing/version.hpp"

namespace tenzing {

namespace log {
int &rank_ref() {
  static thread_local int r = 0; // per thread: ranks may be threads (LoopbackWorld)
  return r;
}
int &verbosity_ref() {
  static int v = [] {
    const char *e = std::getenv("TENZING_VERBOSE");
    return e ? std::atoi(e) : 0;
  }();
  return v;
}
} // namespace log

// ---- randomness ----
namespace randomness {

double runs_test_z(const std::vector<double> &v) {
  const double nan = std::numeric_limits<double>::quiet_NaN();
  if (v.size() < 2) return nan;
  const double median = med(v);
  size_t nHi = 0, nLo = 0, runs = 1;
  bool prev = false;
  for (size_t i = 0; i < v.size(); ++i) {
    const bool hi = v[i] >= median;
    (hi ? nHi : nLo)++;
    if (i && hi != prev) ++runs;
    prev = hi;
  }
  if (nHi < 10 || nLo < 10) return nan;
  const double a = double(nHi), b = double(nLo);
  const double expected = 2 * a * b / (a + b) + 1;
  const double sd = std::sqrt(2 * a * b * (2 * a * b - a - b) / ((a + b) * (a + b) * (a + b - 1)));
  return std::abs((double(runs) - expected) / sd);
}

bool runs_test(const std::vector<double> &v) {
  const double z = runs_test_z(v);
  if (std::isnan(z)) return true; // test does not apply -> reject, as the reference does
  return z > 1.96;                // alpha = 0.05
}

bool compound_test(const std::vector<double> &v) { return runs_test(v); }

} // namespace randomness

// ---- counters ----
namespace counters {
// per thread: ranks may be threads of one process (LoopbackWorld), and each rank counts its own
Mcts &mcts() {
  static thread_local Mcts m;
  return m;
}
Dfs &dfs() {
  static thread_local Dfs d;
  return d;
}
void reset() {
  mcts() = Mcts();
  dfs() = Dfs();
}
} // namespace counters

// ---- trap ----
namespace {
TrapHandler &trap_fn() {
  static TrapHandler fn;
  return fn;
}
void on_signal(int sig) {
  TrapHandler fn = trap_fn();
  if (fn) fn(sig);
  std::_Exit(1);
}
} // namespace

void register_handler(TrapHandler fn) {
  trap_fn() = std::move(fn);
  std::signal(SIGINT, on_signal);
  std::signal(SIGABRT, on_signal);
  std::signal(SIGTERM, on_signal);
}
void unregister_handler() {
  trap_fn() = nullptr;
  std::signal(SIGINT, SIG_DFL);
  std::signal(SIGABRT, SIG_DFL);
  std::signal(SIGTERM, SIG_DFL);
}

// ---- reproduce ----
namespace reproduce {
std::string version_string() {
  return std::to_string(TENZING_VERSION_MAJOR) + "." + std::to_string(TENZING_VERSION_MINOR) +
         "." + std::to_string(TENZING_VERSION_PATCH) + "-" + TENZING_VERSION_HASH;
}
std::string header_json(const std::vector<std::string> &args) {
  Json j = Json::object();
  Json a = Json::array();
  for (const auto &s : args) a.push_back(s);
  j["args"] = a;
  j["hash"] = TENZING_VERSION_HASH;
  j["major"] = TENZING_VERSION_MAJOR;
  j["minor"] = TENZING_VERSION_MINOR;
  j["patch"] = TENZING_VERSION_PATCH;
  return j.dump();
}
bool Header::same_build() const {
  return valid && major == TENZING_VERSION_MAJOR && minor == TENZING_VERSION_MINOR &&
         patch == TENZING_VERSION_PATCH && hash == TENZING_VERSION_HASH;
}
std::string Header::version() const {
  return std::to_string(major) + "." + std::to_string(minor) + "." + std::to_string(patch) + "-" + hash;
}
Header make_header(const std::vector<std::string> &args) {
  Header h;
  h.args = args;
  h.hash = TENZING_VERSION_HASH;
  h.major = TENZING_VERSION_MAJOR;
  h.minor = TENZING_VERSION_MINOR;
  h.patch = TENZING_VERSION_PATCH;
  h.valid = true;
  return h;
}
Header parse_header(const std::string &line) {
  Header h;
  try {
    const Json j = Json::parse(line);
    if (!j.is_object() || !j.contains("args") || !j.contains("hash") || !j.contains("major") ||
        !j.contains("minor") || !j.contains("patch"))
      return h;
    for (const Json &a : j.at("args").as_array()) h.args.push_back(a.as_string());
    h.hash = j.at("hash").as_string();
    h.major = int(j.at("major").as_int());
    h.minor = int(j.at("minor").as_int());
    h.patch = int(j.at("patch").as_int());
    h.valid = true;
  } catch (const std::exception &) {
    h = Header();
  }
  return h;
}
void dump_with_cli(int argc, char **argv) {
  std::vector<std::string> args;
  for (int i = 0; i < argc; ++i) args.emplace_back(argv[i]);
  std::cout << header_json(args) << std::endl;
}
} // namespace reproduce

// ---- init ----
bool notice_acknowledged() { return std::getenv("TENZING_ACK_NOTICE") != nullptr; }

void init() {
  static bool done = false;
  if (done) return;
  done = true;
  if (!notice_acknowledged()) {
    std::cerr << "tenzing-b200 " << reproduce::version_string()
              << ": research software, provided as is; schedules are benchmarked on YOUR "
                 "hardware and results depend on it. Set TENZING_ACK_NOTICE=1 to silence "
                 "this notice.\n";
  }
}
void init(int &, char **&) { init(); }

} // namespace tenzing
