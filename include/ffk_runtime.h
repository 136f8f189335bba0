/* ffk_runtime.h — the host-side pieces of the kernel API: device memory arena, per-device
 * handle, stream and timing helpers. Included by ffk.h. These stand in for
 * LocalCudaAllocator, ManagedPerDeviceFFHandle, ManagedFFStream and the event plumbing of
 * profiling_wrapper; the C++ veneer wraps them in the reference's class names. */
#ifndef FFK_RUNTIME_H_
#define FFK_RUNTIME_H_
#include "ffk.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Device-memory arena — replaces LocalCudaAllocator (include/kernels/local_cuda_allocator.h:6-20,
 * src/local_cuda_allocator.cc: one cudaMalloc per tensor, everything freed in the destructor).
 * Here blocks are carved from large slabs and recycled through power-of-two size classes, so a
 * steady-state training step performs no cudaMalloc at all. All blocks are 256-byte aligned.
 * Not thread-safe across host threads sharing one arena (one process per GPU). */
typedef struct ffk_arena ffk_arena_t;
int ffk_arena_create(ffk_arena_t **out, size_t slab_bytes);
int ffk_arena_allocate(ffk_arena_t *arena, size_t bytes, void **out);
int ffk_arena_deallocate(ffk_arena_t *arena, void *ptr);
int ffk_arena_destroy(ffk_arena_t *arena);
/* Bytes handed out and not yet returned / bytes obtained from the driver / driver allocations made so
 * far (LocalCudaAllocator, src/local_cuda_allocator.cc, makes one per tensor; these counters are how the
 * tests show the arena does not). */
size_t ffk_arena_bytes_in_use(ffk_arena_t const *arena);
size_t ffk_arena_bytes_reserved(ffk_arena_t const *arena);
/* Number of driver allocations (cudaMalloc calls) made so far — what the cost estimator and
 * the tests use to show that steady state is allocation-free (src/local_cuda_allocator.cc allocates
 * per tensor). */
size_t ffk_arena_driver_allocations(ffk_arena_t const *arena);

/* The data members of PerDeviceFFHandle (include/kernels/ff_handle.h:13-24) that the kernels
 * read. `dnn` / `blas` are kept for layout but unused: nothing in libffk.so calls cuDNN/cuBLAS. */
typedef struct {
  void *dnn;
  void *blas;
  void *workSpace;
  size_t workSpaceSize;
  int allowTensorOpMathConversion;
  void *ncclComm;
} ffk_handle_t;
/* ManagedPerDeviceFFHandle — include/kernels/managed_per_device_ff_handle.h:8-27,
 * src/managed_per_device_ff_handle.cc:6-38 (1 MiB workspace by default there; pass 0 for it). */
int ffk_handle_create(ffk_handle_t *out, size_t workspace_bytes, int allow_tensor_op_math_conversion);
int ffk_handle_destroy(ffk_handle_t *handle);

/* ManagedFFStream — include/kernels/managed_ff_stream.h:8-25. */
int ffk_stream_create(void **out_stream);
int ffk_stream_destroy(void *stream);
int ffk_stream_synchronize(void *stream);

/* Event pair used by profiling_wrapper (include/kernels/profiling.h:24-49): begin records the
 * start event on `stream`; end records the stop event, waits for it and returns the elapsed
 * milliseconds of everything enqueued in between — the TOTAL over the measured iterations,
 * which is what the reference returns. Events are created once per timer and reused. */
typedef struct ffk_timer ffk_timer_t;
int ffk_timer_create(ffk_timer_t **out);
int ffk_timer_begin(ffk_timer_t *timer, void *stream);
int ffk_timer_end(ffk_timer_t *timer, void *stream, float *elapsed_ms);
int ffk_timer_destroy(ffk_timer_t *timer);
/* Profiling hook in the spirit of profiling_wrapper (include/kernels/profiling.h:24-49), at the
 * granularity of ONE kernel: the next
 * launch of the CTA-pair tcgen05 GEMM issued by the calling host thread (inside whatever entry
 * point issues it — Linear, MultiHeadAttention, ...) records the timer's start event in front of
 * the kernel and its stop event behind it, on the kernel's own stream. ffk_timer_read waits for
 * the stop event and returns the milliseconds between the two. bench.py uses it to time the
 * dominant kernel live inside its timed steps. */
int ffk_gemm_time_next(ffk_timer_t *timer);
int ffk_timer_read(ffk_timer_t *timer, float *elapsed_ms);

/* zero_init_kernel / constant fill on the device — include/kernels/initializer_kernels.h:10-28
 * (declared there; the reference defines only the CPU zero/constant versions,
 * src/cpu/initializer_kernels.cc:1-52, and leaves the GPU one NOT_IMPLEMENTED). */
int ffk_fill(void *stream, void *ptr, size_t n, double value, int dtype);

#ifdef __cplusplus
}
#endif
#endif
