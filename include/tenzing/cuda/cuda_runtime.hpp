// CUDA error-check macro (reference: include/tenzing/cuda/cuda_runtime.hpp:16-34
// prints and exit(-1)s; here it throws so Python callers and tests see an
// exception instead of losing the process).
#pragma once

#include <cuda_runtime_api.h>

#include "tenzing/macro_at.hpp"

namespace tenzing {
inline void check_cuda(cudaError_t e, const char *file, int line, const char *expr) {
  if (TENZING_UNLIKELY(e != cudaSuccess)) {
    std::ostringstream ss;
    ss << file << ":" << line << ": CUDA error " << int(e) << " (" << cudaGetErrorName(e)
       << "): " << cudaGetErrorString(e) << " in " << expr;
    throw std::runtime_error(ss.str());
  }
}
} // namespace tenzing

#define CUDA_RUNTIME(stmt) ::tenzing::check_cuda((stmt), __FILE__, __LINE__, #stmt)
