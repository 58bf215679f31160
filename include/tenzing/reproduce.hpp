// Reproducibility header: one JSON line {"args","hash","major","minor","patch"}
// (reference: src/reproduce.cpp:22-46, include/tenzing/version.hpp.in:8-14).
// Written as the first line of every result CSV; `parse_header` reads it back so
// that a CsvBenchmarker (or the post-processing scripts) can tell which build and
// which command line produced a file.
#pragma once

#include <string>
#include <vector>

namespace tenzing {
namespace reproduce {

struct Header {
  std::vector<std::string> args; // argv of the producing run
  std::string hash;              // git hash the library was built from ("unknown" outside git)
  int major = 0, minor = 0, patch = 0;
  bool valid = false;            // false: the line was not a reproduce header
  // same major.minor.patch and hash as this build
  bool same_build() const;
  std::string version() const;   // "major.minor.patch-hash"
};

std::string version_string();
std::string header_json(const std::vector<std::string> &args);
// header of this build for a given command line
Header make_header(const std::vector<std::string> &args);
// inverse of header_json; any other line gives valid == false (never throws)
Header parse_header(const std::string &line);
// print header_json(argv) on stdout (first line of every driver's output)
void dump_with_cli(int argc, char **argv);

} // namespace reproduce
} // namespace tenzing
