// Reproducibility header: one JSON line {"args","hash","major","minor","patch"}
// (reference: src/reproduce.cpp:22-46, include/tenzing/version.hpp.in:8-14).
#pragma once

#include <string>
#include <vector>

namespace tenzing {
namespace reproduce {
std::string version_string();
std::string header_json(const std::vector<std::string> &args);
void dump_with_cli(int argc, char **argv);
} // namespace reproduce
} // namespace tenzing
