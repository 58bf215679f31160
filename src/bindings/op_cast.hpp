// Ops defined inside the C++ workload libraries (BoxCopyOp, SpMV, ...) are not
// registered with pybind11 one by one.  When such an op reaches Python through a
// shared_ptr<OpBase>, present it as the most specific REGISTERED base class, so
// that e.g. a compound op has .graph() and a GPU op can be bound to a stream.
#pragma once

#include <pybind11/pybind11.h>

#include "tenzing/tenzing.hpp"

namespace pybind11 {
template <> struct polymorphic_type_hook<tenzing::OpBase> {
  static const void *get(const tenzing::OpBase *src, const std::type_info *&type) {
    using namespace tenzing;
    if (!src) return src;
    const std::type_info &dyn = typeid(*src);
    if (detail::get_type_info(dyn)) { // registered exactly
      type = &dyn;
      return dynamic_cast<const void *>(src);
    }
#define TZ_TRY(T)                                                                                  \
  if (auto p = dynamic_cast<const T *>(src)) {                                                     \
    type = &typeid(T);                                                                             \
    return static_cast<const void *>(p);                                                           \
  }
    TZ_TRY(GpuOp)
    TZ_TRY(CompoundOp)
    TZ_TRY(ChoiceOp)
    TZ_TRY(CpuOp)
    TZ_TRY(BoundOp)
#undef TZ_TRY
    type = &typeid(OpBase);
    return src;
  }
};
} // namespace pybind11
