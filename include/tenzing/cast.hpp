// Checked down-casts (reference: include/tenzing/cast.hpp:14-31).
#pragma once

#include <memory>

#include "tenzing/macro_at.hpp"

namespace tenzing {

template <typename T, typename U> bool isa(const std::shared_ptr<U> &p) {
  return bool(std::dynamic_pointer_cast<T>(p));
}

template <typename T, typename U> std::shared_ptr<T> cast_or_throw(const std::shared_ptr<U> &p) {
  auto r = std::dynamic_pointer_cast<T>(p);
  if (!r) THROW_RUNTIME("bad cast to " << typeid(T).name());
  return r;
}

} // namespace tenzing

#define TENZING_MUST_CAST(T, p) (::tenzing::cast_or_throw<T>(p))
