// Minimal JSON value used for op / schedule serialisation.
//
// The serialized schedule format is the reference's
// nlohmann::json `dump()` output: compact, object keys in alphabetical order.
// This class reproduces exactly that textual form (std::map keeps keys sorted)
// and parses it back; nothing else from a general JSON library is needed.
#pragma once

#include <cstdint>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace tenzing {

class Json {
public:
  enum class Kind { Null, Bool, Int, Float, String, Array, Object };
  using array_t = std::vector<Json>;
  using object_t = std::map<std::string, Json>;

  Json() : kind_(Kind::Null) {}
  Json(std::nullptr_t) : kind_(Kind::Null) {}
  Json(bool b) : kind_(Kind::Bool), b_(b) {}
  Json(int v) : kind_(Kind::Int), i_(v) {}
  Json(long v) : kind_(Kind::Int), i_(v) {}
  Json(long long v) : kind_(Kind::Int), i_(v) {}
  Json(unsigned v) : kind_(Kind::Int), i_(v) {}
  Json(unsigned long v) : kind_(Kind::Int), i_(static_cast<int64_t>(v)) {}
  Json(unsigned long long v) : kind_(Kind::Int), i_(static_cast<int64_t>(v)) {}
  Json(double v) : kind_(Kind::Float), d_(v) {}
  Json(const char *s) : kind_(Kind::String), s_(s) {}
  Json(const std::string &s) : kind_(Kind::String), s_(s) {}
  Json(const array_t &a) : kind_(Kind::Array), a_(a) {}
  Json(const object_t &o) : kind_(Kind::Object), o_(o) {}

  static Json array() { Json j; j.kind_ = Kind::Array; return j; }
  static Json object() { Json j; j.kind_ = Kind::Object; return j; }

  Kind kind() const { return kind_; }
  bool is_null() const { return kind_ == Kind::Null; }
  bool is_bool() const { return kind_ == Kind::Bool; }
  bool is_int() const { return kind_ == Kind::Int; }
  bool is_number() const { return kind_ == Kind::Int || kind_ == Kind::Float; }
  bool is_string() const { return kind_ == Kind::String; }
  bool is_array() const { return kind_ == Kind::Array; }
  bool is_object() const { return kind_ == Kind::Object; }

  bool as_bool() const { expect(Kind::Bool); return b_; }
  int64_t as_int() const {
    if (kind_ == Kind::Float) return static_cast<int64_t>(d_);
    expect(Kind::Int);
    return i_;
  }
  double as_double() const {
    if (kind_ == Kind::Int) return static_cast<double>(i_);
    expect(Kind::Float);
    return d_;
  }
  const std::string &as_string() const { expect(Kind::String); return s_; }
  const array_t &as_array() const { expect(Kind::Array); return a_; }
  array_t &as_array() { expect(Kind::Array); return a_; }
  const object_t &as_object() const { expect(Kind::Object); return o_; }

  // object access; a Null value silently becomes an object on write access
  Json &operator[](const std::string &key) {
    if (kind_ == Kind::Null) kind_ = Kind::Object;
    expect(Kind::Object);
    return o_[key];
  }
  const Json &at(const std::string &key) const {
    expect(Kind::Object);
    auto it = o_.find(key);
    if (it == o_.end()) throw std::out_of_range("json: no key '" + key + "'");
    return it->second;
  }
  bool contains(const std::string &key) const {
    return kind_ == Kind::Object && o_.count(key) != 0;
  }
  // array access
  void push_back(const Json &v) {
    if (kind_ == Kind::Null) kind_ = Kind::Array;
    expect(Kind::Array);
    a_.push_back(v);
  }
  const Json &at(size_t i) const { expect(Kind::Array); return a_.at(i); }
  size_t size() const {
    if (kind_ == Kind::Array) return a_.size();
    if (kind_ == Kind::Object) return o_.size();
    return 0;
  }

  bool operator==(const Json &r) const;
  bool operator!=(const Json &r) const { return !(*this == r); }

  // compact dump, alphabetical keys: {"a":1,"b":[true,null]}
  std::string dump() const;
  static Json parse(const std::string &text);

private:
  void expect(Kind k) const {
    if (kind_ != k) throw std::runtime_error("json: wrong value kind");
  }
  Kind kind_;
  bool b_ = false;
  int64_t i_ = 0;
  double d_ = 0;
  std::string s_;
  array_t a_;
  object_t o_;
};

} // namespace tenzing
