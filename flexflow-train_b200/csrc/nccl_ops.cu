// NCCL entry points. NCCL is loaded with dlopen at first use so that libffk.so has no link-time
// dependency on it; only the types of <nccl.h> are used at compile time.
#include "common.cuh"

#include <dlfcn.h>
#include <nccl.h>

#include <mutex>

namespace ffk {
namespace {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(void const *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  char const *(*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

NcclApi const &api() {
  static NcclApi a;
  static std::once_flag once;
  std::call_once(once, [] {
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return;
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
    a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(dlsym(h, "ncclAllReduce"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
    a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllReduce && a.GetErrorString;
  });
  return a;
}

int nccl_fail(ncclResult_t r, char const *what) {
  set_error("NCCL failure in %s: %s", what, api().GetErrorString ? api().GetErrorString(r) : "?");
  return FFK_ERR_CUDA;
}

#define FFK_NCCL_READY()                                                                 \
  do {                                                                                   \
    if (!api().ok) {                                                                     \
      set_error("libnccl.so.2 could not be loaded (dlopen): NCCL entry points unavailable"); \
      return FFK_ERR_UNSUPPORTED;                                                        \
    }                                                                                    \
  } while (0)

int nccl_dtype(int dt, ncclDataType_t &out) {
  switch (dt) {
    case FFK_FLOAT: out = ncclFloat; return FFK_OK;
    case FFK_HALF: out = ncclHalf; return FFK_OK;
    case FFK_BFLOAT16: out = ncclBfloat16; return FFK_OK;
    case FFK_DOUBLE: out = ncclDouble; return FFK_OK;
    case FFK_INT32: out = ncclInt32; return FFK_OK;
    case FFK_INT64: out = ncclInt64; return FFK_OK;
    default: set_error("all_reduce: unsupported dtype %d", dt); return FFK_ERR_DTYPE;
  }
}

}  // namespace
}  // namespace ffk

using namespace ffk;

static_assert(sizeof(ncclUniqueId) == FFK_NCCL_UNIQUE_ID_BYTES, "ncclUniqueId size changed");

extern "C" {

int ffk_nccl_get_unique_id(void *id_out_host) {
  FFK_NCCL_READY();
  FFK_REQUIRE(id_out_host, "nccl_get_unique_id: null output");
  ncclUniqueId id;
  ncclResult_t r = api().GetUniqueId(&id);
  if (r != ncclSuccess) return nccl_fail(r, "ncclGetUniqueId");
  std::memcpy(id_out_host, &id, sizeof(id));
  return FFK_OK;
}

int ffk_nccl_comm_create(void **comm_out, void const *id_host, int num_ranks, int my_rank) {
  FFK_NCCL_READY();
  FFK_REQUIRE(comm_out && id_host && num_ranks > 0 && my_rank >= 0 && my_rank < num_ranks,
              "nccl_comm_create: bad arguments");
  ncclUniqueId id;
  std::memcpy(&id, id_host, sizeof(id));
  ncclComm_t comm;
  ncclResult_t r = api().CommInitRank(&comm, num_ranks, id, my_rank);
  if (r != ncclSuccess) return nccl_fail(r, "ncclCommInitRank");
  *comm_out = comm;
  return FFK_OK;
}

int ffk_nccl_comm_destroy(void *comm) {
  if (!comm) return FFK_OK;
  FFK_NCCL_READY();
  ncclResult_t r = api().CommDestroy(static_cast<ncclComm_t>(comm));
  if (r != ncclSuccess) return nccl_fail(r, "ncclCommDestroy");
  return FFK_OK;
}

int ffk_all_reduce_sum(void *stream, void *comm, void *buffer, size_t count, int dtype) {
  FFK_NCCL_READY();
  if (count == 0) return FFK_OK;
  FFK_REQUIRE(comm && buffer, "all_reduce: null communicator or buffer");
  ncclDataType_t dt;
  int rc = nccl_dtype(dtype, dt);
  if (rc) return rc;
  ncclResult_t r = api().AllReduce(buffer, buffer, count, dt, ncclSum, static_cast<ncclComm_t>(comm),
                                   as_stream(stream));
  if (r != ncclSuccess) return nccl_fail(r, "ncclAllReduce");
  return FFK_OK;
}

int ffk_sgd_nccl_update(void *stream, void *comm, float lr, float momentum, int nesterov,
                        float weight_decay, void *weight_grad, size_t size, void *weight,
                        float *sgd_v, int dtype, float *master_weight) {
  int rc = ffk_all_reduce_sum(stream, comm, weight_grad, size, dtype);
  if (rc) return rc;
  return ffk_sgd_update(stream, lr, momentum, nesterov, weight_decay, weight_grad, size, 1, weight,
                        sgd_v, dtype, master_weight);
}

int ffk_adam_nccl_update(void *stream, void *comm, float alpha_t, float beta1, float beta2,
                         float weight_decay, float epsilon, void *weight_grad, size_t size,
                         float *adam_m, float *adam_v, void *weight, int dtype,
                         float *master_weight) {
  int rc = ffk_all_reduce_sum(stream, comm, weight_grad, size, dtype);
  if (rc) return rc;
  return ffk_adam_update(stream, alpha_t, beta1, beta2, weight_decay, epsilon, weight_grad, size, 1,
                         adam_m, adam_v, weight, dtype, master_weight);
}

}  // extern "C"
