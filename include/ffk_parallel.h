#ifndef FFK_PARALLEL_H_
#define FFK_PARALLEL_H_
#include "ffk.h"
#ifdef __cplusplus
extern "C" {
#endif

#ifdef __cplusplus
}
#endif
#endif
