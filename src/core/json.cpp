#include "tenzing/json.hpp"

#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace tenzing {

namespace {

void dump_string(const std::string &s, std::string &out) {
  out.push_back('"');
  for (unsigned char c : s) {
    switch (c) {
    case '"': out += "\\\""; break;
    case '\\': out += "\\\\"; break;
    case '\n': out += "\\n"; break;
    case '\r': out += "\\r"; break;
    case '\t': out += "\\t"; break;
    case '\b': out += "\\b"; break;
    case '\f': out += "\\f"; break;
    default:
      if (c < 0x20) {
        char buf[8];
        std::snprintf(buf, sizeof(buf), "\\u%04x", c);
        out += buf;
      } else {
        out.push_back(static_cast<char>(c));
      }
    }
  }
  out.push_back('"');
}

// shortest decimal form that round-trips
void dump_double(double d, std::string &out) {
  if (!std::isfinite(d)) {
    out += "null";
    return;
  }
  char buf[40];
  for (int prec = 1; prec <= 17; ++prec) {
    std::snprintf(buf, sizeof(buf), "%.*g", prec, d);
    if (std::strtod(buf, nullptr) == d) break;
  }
  out += buf;
  // keep it recognisable as a float
  if (!std::strpbrk(buf, ".eEn")) out += ".0";
}

struct Parser {
  const std::string &t;
  size_t p = 0;
  explicit Parser(const std::string &text) : t(text) {}

  [[noreturn]] void fail(const char *what) const {
    throw std::runtime_error(std::string("json parse error at offset ") + std::to_string(p) +
                             ": " + what);
  }
  void ws() {
    while (p < t.size() && std::isspace(static_cast<unsigned char>(t[p]))) ++p;
  }
  bool eat(char c) {
    ws();
    if (p < t.size() && t[p] == c) {
      ++p;
      return true;
    }
    return false;
  }
  void lit(const char *s) {
    size_t n = std::strlen(s);
    if (t.compare(p, n, s) != 0) fail("bad literal");
    p += n;
  }
  std::string str() {
    if (!eat('"')) fail("expected string");
    std::string out;
    while (true) {
      if (p >= t.size()) fail("unterminated string");
      char c = t[p++];
      if (c == '"') break;
      if (c != '\\') {
        out.push_back(c);
        continue;
      }
      if (p >= t.size()) fail("bad escape");
      char e = t[p++];
      switch (e) {
      case '"': out.push_back('"'); break;
      case '\\': out.push_back('\\'); break;
      case '/': out.push_back('/'); break;
      case 'n': out.push_back('\n'); break;
      case 'r': out.push_back('\r'); break;
      case 't': out.push_back('\t'); break;
      case 'b': out.push_back('\b'); break;
      case 'f': out.push_back('\f'); break;
      case 'u': {
        if (p + 4 > t.size()) fail("bad \\u escape");
        unsigned cp = static_cast<unsigned>(std::strtoul(t.substr(p, 4).c_str(), nullptr, 16));
        p += 4;
        if (cp < 0x80) {
          out.push_back(static_cast<char>(cp));
        } else if (cp < 0x800) {
          out.push_back(static_cast<char>(0xC0 | (cp >> 6)));
          out.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
        } else {
          out.push_back(static_cast<char>(0xE0 | (cp >> 12)));
          out.push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
          out.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
        }
        break;
      }
      default: fail("unknown escape");
      }
    }
    return out;
  }
  Json value() {
    ws();
    if (p >= t.size()) fail("unexpected end");
    char c = t[p];
    if (c == '{') {
      ++p;
      Json j = Json::object();
      if (eat('}')) return j;
      do {
        ws();
        std::string k = str();
        if (!eat(':')) fail("expected ':'");
        j[k] = value();
      } while (eat(','));
      if (!eat('}')) fail("expected '}'");
      return j;
    }
    if (c == '[') {
      ++p;
      Json j = Json::array();
      if (eat(']')) return j;
      do {
        j.push_back(value());
      } while (eat(','));
      if (!eat(']')) fail("expected ']'");
      return j;
    }
    if (c == '"') return Json(str());
    if (c == 't') { lit("true"); return Json(true); }
    if (c == 'f') { lit("false"); return Json(false); }
    if (c == 'n') { lit("null"); return Json(); }
    // number
    size_t b = p;
    bool isFloat = false;
    if (t[p] == '-' || t[p] == '+') ++p;
    while (p < t.size()) {
      char d = t[p];
      if (std::isdigit(static_cast<unsigned char>(d))) {
        ++p;
      } else if (d == '.' || d == 'e' || d == 'E' || d == '-' || d == '+') {
        isFloat = true;
        ++p;
      } else {
        break;
      }
    }
    if (b == p) fail("unexpected character");
    std::string num = t.substr(b, p - b);
    if (isFloat) return Json(std::strtod(num.c_str(), nullptr));
    return Json(static_cast<long long>(std::strtoll(num.c_str(), nullptr, 10)));
  }
};

void dump_into(const Json &j, std::string &out) {
  switch (j.kind()) {
  case Json::Kind::Null: out += "null"; break;
  case Json::Kind::Bool: out += j.as_bool() ? "true" : "false"; break;
  case Json::Kind::Int: out += std::to_string(j.as_int()); break;
  case Json::Kind::Float: dump_double(j.as_double(), out); break;
  case Json::Kind::String: dump_string(j.as_string(), out); break;
  case Json::Kind::Array: {
    out.push_back('[');
    bool first = true;
    for (const Json &e : j.as_array()) {
      if (!first) out.push_back(',');
      first = false;
      dump_into(e, out);
    }
    out.push_back(']');
    break;
  }
  case Json::Kind::Object: {
    out.push_back('{');
    bool first = true;
    for (const auto &kv : j.as_object()) {
      if (!first) out.push_back(',');
      first = false;
      dump_string(kv.first, out);
      out.push_back(':');
      dump_into(kv.second, out);
    }
    out.push_back('}');
    break;
  }
  }
}

} // namespace

bool Json::operator==(const Json &r) const {
  if (is_number() && r.is_number()) {
    if (kind_ == Kind::Int && r.kind_ == Kind::Int) return i_ == r.i_;
    return as_double() == r.as_double();
  }
  if (kind_ != r.kind_) return false;
  switch (kind_) {
  case Kind::Null: return true;
  case Kind::Bool: return b_ == r.b_;
  case Kind::String: return s_ == r.s_;
  case Kind::Array: return a_ == r.a_;
  case Kind::Object: return o_ == r.o_;
  default: return false;
  }
}

std::string Json::dump() const {
  std::string out;
  dump_into(*this, out);
  return out;
}

Json Json::parse(const std::string &text) {
  Parser ps(text);
  Json j = ps.value();
  ps.ws();
  if (ps.p != text.size()) ps.fail("trailing characters");
  return j;
}

} // namespace tenzing
