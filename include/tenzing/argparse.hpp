// Tiny command-line parser with the interface the reference's drivers use
// (thirdparty/cwpearson/argparse/argparse.hpp:179-356): add_option, add_flag,
// add_positional, no_unrecognized, parse, help.
#pragma once

#include <cstdint>
#include <sstream>
#include <string>
#include <vector>

namespace tenzing {
namespace argparse {

class Parser {
  enum class Kind { Int, Size, Double, String, Flag };
  struct Opt {
    Kind kind;
    void *dst;
    std::string longName, shortName, help;
  };
  struct Pos {
    Kind kind;
    void *dst;
    bool required;
    bool found = false;
  };
  std::string description_;
  std::vector<Opt> opts_;
  std::vector<Pos> pos_;
  bool noUnrecognized_ = false;

  static bool assign(Kind k, void *dst, const std::string &v) {
    try {
      switch (k) {
      case Kind::Int: *static_cast<int *>(dst) = std::stoi(v); break;
      case Kind::Size: *static_cast<size_t *>(dst) = size_t(std::stoull(v)); break;
      case Kind::Double: *static_cast<double *>(dst) = std::stod(v); break;
      case Kind::String: *static_cast<std::string *>(dst) = v; break;
      case Kind::Flag: *static_cast<bool *>(dst) = true; break;
      }
    } catch (...) {
      return false;
    }
    return true;
  }
  void add(Kind k, void *dst, const std::string &l, const std::string &s, const std::string &h) {
    opts_.push_back(Opt{k, dst, l, s, h});
  }

public:
  explicit Parser(const std::string &description = "") : description_(description) {}
  void add_option(int &v, const std::string &l, const std::string &s = "", const std::string &h = "") { add(Kind::Int, &v, l, s, h); }
  void add_option(size_t &v, const std::string &l, const std::string &s = "", const std::string &h = "") { add(Kind::Size, &v, l, s, h); }
  void add_option(double &v, const std::string &l, const std::string &s = "", const std::string &h = "") { add(Kind::Double, &v, l, s, h); }
  void add_option(std::string &v, const std::string &l, const std::string &s = "", const std::string &h = "") { add(Kind::String, &v, l, s, h); }
  void add_flag(bool &v, const std::string &l, const std::string &s = "", const std::string &h = "") { add(Kind::Flag, &v, l, s, h); }
  void add_positional(int &v, bool required = false) { pos_.push_back(Pos{Kind::Int, &v, required}); }
  void add_positional(std::string &v, bool required = false) { pos_.push_back(Pos{Kind::String, &v, required}); }
  void no_unrecognized() { noUnrecognized_ = true; }

  // consumes recognised arguments from argv (argc is updated); false on error
  bool parse(int &argc, char **argv) {
    std::vector<char *> rest{argv[0]};
    size_t nextPos = 0;
    bool ok = true;
    for (int i = 1; i < argc; ++i) {
      const std::string a = argv[i];
      const Opt *hit = nullptr;
      for (const Opt &o : opts_)
        if (a == o.longName || (!o.shortName.empty() && a == o.shortName)) hit = &o;
      if (hit) {
        if (hit->kind == Kind::Flag) {
          assign(hit->kind, hit->dst, "");
        } else if (i + 1 < argc) {
          ok = assign(hit->kind, hit->dst, argv[++i]) && ok;
        } else {
          ok = false;
        }
      } else if (!a.empty() && a[0] == '-' && a.size() > 1 && !(a[1] >= '0' && a[1] <= '9')) {
        if (noUnrecognized_) ok = false;
        rest.push_back(argv[i]);
      } else if (nextPos < pos_.size()) {
        ok = assign(pos_[nextPos].kind, pos_[nextPos].dst, a) && ok;
        pos_[nextPos++].found = true;
      } else {
        rest.push_back(argv[i]);
      }
    }
    for (const Pos &p : pos_)
      if (p.required && !p.found) ok = false;
    argc = int(rest.size());
    for (int i = 0; i < argc; ++i) argv[i] = rest[size_t(i)];
    return ok;
  }

  std::string help() const {
    std::ostringstream ss;
    if (!description_.empty()) ss << description_ << "\n";
    ss << "options:\n";
    for (const Opt &o : opts_) {
      ss << "  " << o.longName;
      if (!o.shortName.empty()) ss << ", " << o.shortName;
      if (o.kind != Kind::Flag) ss << " <value>";
      if (!o.help.empty()) ss << "    " << o.help;
      ss << "\n";
    }
    return ss.str();
  }
};

} // namespace argparse
} // namespace tenzing
