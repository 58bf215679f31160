// Wald-Wolfowitz runs test used to reject drifting measurement series
// (reference: src/randomness.cpp:12-64).  Return value true = REJECT the
// hypothesis that the series is random (also true when the test does not
// apply: fewer than 10 samples on either side of the median).
#pragma once

#include <vector>

namespace tenzing {
namespace randomness {
bool runs_test(const std::vector<double> &v);
bool compound_test(const std::vector<double> &v);
// the Z statistic itself (NaN if the test does not apply); for diagnostics/tests
double runs_test_z(const std::vector<double> &v);
} // namespace randomness
} // namespace tenzing
