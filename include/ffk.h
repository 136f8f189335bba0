/* ffk.h — C ABI of libffk.so, the B200 (sm_100a) implementation of FlexFlow-train's
 * per-operator training kernels.
 *
 * Every entry point below replaces one function of the reference's lib/kernels API
 * (`FlexFlow::Kernels::<Op>::{init,forward,backward}_kernel`, the loss / optimizer /
 * metrics functions and the parallel-op kernels); the comment above each one cites the
 * reference declaration (header) and definition (.cu) it stands in for. Paths are
 * relative to the reference checkout, i.e. lib/kernels/... .
 *
 * Conventions
 *  - plain pointers and sizes only; all tensor pointers are DEVICE pointers unless the
 *    name ends in `_host`;
 *  - `stream` is a cudaStream_t passed as void* (reference: ffStream_t first argument);
 *  - tensors are dense row-major in FF order (outermost dimension first), which is what
 *    the reference's GenericTensorAccessor{R,W}::ptr points at;
 *  - dtype arguments take ffk_dtype_t, whose first six values are the reference's
 *    `DataType` enum in its own order (op-attrs/datatype.enum.toml) plus BFLOAT16;
 *  - backward kernels ACCUMULATE into input/weight gradients unless stated otherwise
 *    (the reference's beta = 1 convention);
 *  - return value: 0 on success, a negative ffk_status_t on failure, with a message
 *    available from ffk_last_error() (per host thread). Nothing is thrown across the
 *    boundary. The C++ veneer (flexflow-train_b200/include/kernels) turns a failure
 *    into the reference's behaviour: print, abort.
 *  - there is no CPU fallback: if no CUDA device is usable every compute entry point
 *    returns FFK_ERR_CUDA.
 */
#ifndef FFK_H_
#define FFK_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  FFK_BOOL = 0,
  FFK_INT32 = 1,
  FFK_INT64 = 2,
  FFK_HALF = 3,
  FFK_FLOAT = 4,
  FFK_DOUBLE = 5,
  FFK_BFLOAT16 = 6 /* extension: the reference has no bf16 */
} ffk_dtype_t;

typedef enum {
  FFK_OK = 0,
  FFK_ERR_INVALID = -1,     /* bad argument (null pointer, negative size, ...) */
  FFK_ERR_DTYPE = -2,       /* dtype not supported by this entry point */
  FFK_ERR_UNSUPPORTED = -3, /* valid request this build cannot serve */
  FFK_ERR_CUDA = -4,        /* CUDA runtime / driver failure */
  FFK_ERR_LIMIT = -5        /* a documented size limit was exceeded */
} ffk_status_t;

/* Activation codes: the reference's `Activation` enum order (op-attrs/activation.enum.toml);
 * FFK_ACT_NONE stands for std::nullopt. */
typedef enum { FFK_ACT_NONE = -1, FFK_ACT_RELU = 0, FFK_ACT_SIGMOID = 1, FFK_ACT_TANH = 2, FFK_ACT_GELU = 3 } ffk_activation_t;

/* ------------------------------------------------------------------ core */
/* Message of the last failed call on this host thread. The reference prints and aborts instead
 * (checkCUDA / checkCUDNN / checkCUBLAS, include/kernels/device.h:88-108, src/device.h:31-56); the
 * C++ veneer restores that behaviour on top of this (FFK_CHECK). */
char const *ffk_last_error(void);
/* Version of this C ABI (no reference counterpart: libkernels.so is linked by name,
 * lib/kernels/CMakeLists.txt:15-19). */
int ffk_abi_version(void);
/* Number of SMs of the current device, or a negative status (the reference sizes its grids with
 * GET_BLOCKS instead, src/device.h:59-71). */
int ffk_device_sm_count(void);

#ifdef __cplusplus
}
#endif

#include "ffk_dense.h"
#include "ffk_elementwise.h"
#include "ffk_norm.h"
#include "ffk_index.h"
#include "ffk_shape.h"
#include "ffk_train.h"
#include "ffk_parallel.h"
#include "ffk_runtime.h"

#endif /* FFK_H_ */
