// Force-included (-include) in front of every reference test file that tests/cpp/build_ref_tests.sh
// compiles where it lies under the reference checkout. The reference's test_utils.h names
// FlexFlow types without qualification BEFORE the test files say `using namespace ::FlexFlow`,
// and leans on a few `utils/` helpers; this header supplies exactly that and nothing else.
#pragma once
#define FFK_NO_TEST_SHAPE_HELPERS
#include <algorithm>
#include <cstring>
#include <numeric>
#include <vector>

#include "kernels/accessor.h"
#include "kernels/device.h"
#include "kernels/local_cuda_allocator.h"
#include "kernels/managed_ff_stream.h"
#include "kernels/managed_per_device_ff_handle.h"

namespace FlexFlow {
// utils/stack_vector / ff_ordered and utils/containers: what the test helpers use of them
template <typename T> using FFOrdered = std::vector<T>;
template <typename C, typename F> bool all_of(C const &c, F const &f) {
  return std::all_of(c.begin(), c.end(), f);
}
template <typename C, typename F> int count(C const &c, F const &f) {
  return (int)std::count_if(c.begin(), c.end(), f);
}
template <typename F, typename Out = decltype(std::declval<F>()())> std::vector<Out> repeat(int n, F const &f) {
  std::vector<Out> out;
  for (int i = 0; i < n; ++i) out.push_back(f());
  return out;
}
}  // namespace FlexFlow

using namespace ::FlexFlow;
