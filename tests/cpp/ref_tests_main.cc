// Entry point of tests/cpp/reference_kernel_tests: the reference's own kernel test files
// (lib/kernels/test/src/test_*.cc + test_utils.cc), compiled UNMODIFIED where they lie under the
// reference checkout against this repository's veneer headers, with doctest from the reference's
// deps/. Built by __graft_entry__.build() when the checkout is present; the binary travels to the
// GPU box. Nothing of the reference is copied into this repository.
#define DOCTEST_CONFIG_IMPLEMENT_WITH_MAIN
#include "doctest/doctest.h"
