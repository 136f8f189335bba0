// Same include path as the header the reference GENERATES from lib/kernels/include/kernels/profiling_settings.struct.toml
// (proj dtgen, not vendored); the type lives in ffk_cpp/types.h.
#pragma once
#include "../ffk_cpp/ops.h"
