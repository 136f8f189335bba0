// Same include path as the reference (lib/kernels/include/kernels/datatype_dispatch.h); the declarations live in
// ffk_cpp/{types,runtime,ops}.h, organised by layer rather than by operator.
#pragma once
#include "../ffk_cpp/ops.h"
