// Sum over ranks through NVLink peer memory: every rank maps the other ranks' gradient regions
// (CUDA IPC) and the reduction is done by plain loads from peer memory inside one kernel — no
// NCCL call, no staging copies. Up to 64 KB every rank pushes its input into receive slots of all
// peers and sums what has arrived (one switch traversal); up to 512 KB every rank reads all
// peers and sums in rank order (one shot); larger buffers take two shots (rank r reduces chunk r
// from all peers, then the reduced chunks are gathered), so that each byte crosses the switch
// twice per GPU instead of world times.
//
// Cross-rank ordering uses flags that live in each rank's region and only ever grow: a launch
// takes its epoch from a counter in device memory that its last block increments, so the same
// launch can be replayed (CUDA graph) and no host state has to agree between ranks beyond
// "everybody calls the collectives in the same order". Waits are bounded (FFK_P2P_TIMEOUT_MS,
// 30 s by default). A wait that expires never publishes a sum: the rank raises a sticky failure
// flag in EVERY rank's control block, overwrites its own output range with NaN bit patterns and
// leaves without the done / epoch hand-off; every rank that sees the flag (inside a wait or at
// the entry of a later launch) does the same, so a late or missing rank turns into NaN
// gradients and FFK_ERR_CUDA from ffk_p2p_status on all ranks instead of silently wrong sums.
#include "common.cuh"

#include <cstdlib>
#include <cstring>

namespace ffk {
namespace {

constexpr int kMaxRanks = 16;
constexpr size_t kControlBytes = 4096;       // Control block at the head of every region
constexpr int kPhases = 5;                  // input ready / chunk reduced / done reading / pushed (x2 parities)
constexpr size_t kPushLimit = 64 << 10;     // bytes: below, every rank PUSHES its input to all peers
constexpr size_t kOneShotLimit = 512 << 10;  // bytes: below, every rank reads all peers
// receive slots of the push protocol: [parity][source rank][kPushLimit], after the Control block
constexpr size_t kSlotBytes = 2 * size_t(kMaxRanks) * kPushLimit;
constexpr size_t kFlagBytes = kControlBytes + kSlotBytes;  // data part starts here
constexpr unsigned long long kDefaultTimeoutNs = 30ull * 1000000000ull;  // FFK_P2P_TIMEOUT_MS overrides

struct Control {                 // lives at the head of each rank's region (device memory)
  unsigned flags[kPhases][kMaxRanks];
  unsigned epoch;                // completed launches
  unsigned blocks_done;          // grid-wide arrival counters of the current launch
  unsigned blocks_mid;
  unsigned timed_out;            // sticky: a bounded wait expired on this or on another rank
};
static_assert(sizeof(Control) <= kControlBytes, "control block must fit the reserved head");

struct PeerTable {
  unsigned char *base[kMaxRanks];  // region base of every rank as mapped in THIS process
};

struct Group {
  int rank = 0, world = 1, device = 0;
  size_t data_bytes = 0;
  unsigned char *local = nullptr;     // cudaMalloc'd region: [Control | data]
  void *peer_mapped[kMaxRanks] = {};  // cudaIpcOpenMemHandle results (own slot unused)
  PeerTable table{};
  bool connected = false;
  unsigned long long timeout_ns = kDefaultTimeoutNs;
};

__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(unsigned const *p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer16(void const *p) {
  // peer memory must not be served from a stale L1 line of an earlier launch: .cg skips L1
  // (peer lines are never held in the local L2). Not .volatile — strong system-scope loads ran
  // at 140 GB/s whatever the grid size.
  uint4 v;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Tell every peer that this rank reached `phase` of launch `epoch`.
__device__ void signal_peers(PeerTable const &t, int rank, int world, int phase, unsigned epoch) {
  if (threadIdx.x < world) {
    Control *c = reinterpret_cast<Control *>(t.base[threadIdx.x]);
    st_release_sys(&c->flags[phase][rank], epoch);
  }
}
// A bounded wait expired: raise the sticky failure flag on every rank (own block included).
__device__ void poison_all(PeerTable const &t, int world) {
  for (int p = 0; p < world; ++p)
    st_release_sys(&reinterpret_cast<Control *>(t.base[p])->timed_out, 1u);
  __threadfence_system();
}
// One thread spins until *flag >= epoch. false = the wait expired here (all ranks are then
// poisoned) or another rank has poisoned this one.
__device__ bool spin_until(PeerTable const &t, Control *mine, int world, unsigned const *flag,
                           unsigned epoch, unsigned long long timeout_ns) {
  unsigned long long const t0 = global_ns();
  unsigned polls = 0;
  while (ld_acquire_sys(flag) < epoch) {
    if ((++polls & 255u) == 0) {
      if (ld_acquire_sys(&mine->timed_out)) return false;
      if (global_ns() - t0 > timeout_ns) {
        poison_all(t, world);
        return false;
      }
    }
  }
  return true;
}
// Wait (bounded) until every rank has signalled `phase` of launch `epoch` in OUR control block.
// Returns the same verdict to every thread of the block.
__device__ bool wait_peers(PeerTable const &t, Control *mine, int world, int phase, unsigned epoch,
                           unsigned long long timeout_ns) {
  int ok = 1;
  if (threadIdx.x < world)
    ok = spin_until(t, mine, world, &mine->flags[phase][threadIdx.x], epoch, timeout_ns) ? 1 : 0;
  return __syncthreads_and(ok) != 0;
}
// The failed launch's result: all-ones packets (NaN in fp32, half and bf16) over the output.
__device__ void nan_fill(unsigned char *out_local, size_t npk) {
  uint4 const q = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < npk; i += size_t(gridDim.x) * blockDim.x)
    *reinterpret_cast<uint4 *>(out_local + i * 16) = q;
}

template <typename T> struct Acc16 {
  static constexpr int N = 16 / sizeof(T);
  float v[N];
  __device__ void zero() {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = 0.f;
  }
  __device__ void add(uint4 raw) {
    Vec16<T> p;
    *reinterpret_cast<uint4 *>(&p) = raw;
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] += to_f32(p.v[i]);
  }
  __device__ uint4 pack() const {
    Vec16<T> p;
#pragma unroll
    for (int i = 0; i < N; ++i) p.v[i] = from_f32<T>(v[i]);
    return *reinterpret_cast<uint4 const *>(&p);
  }
};

// in_off / out_off: byte offsets of the summed range and of the result inside every region's data
// part; npk = number of 16-byte packets. two_shot: packets are split into `world` chunks.
// THREADS / MINB: 512 x 1 is the stand-alone kernel (two packets from every peer in flight per
// thread). 128 x 6 is the background form: at most 80 registers per thread and no shared memory,
// so that one CTA fits beside a resident tcgen05 GEMM CTA (166 registers x 320 threads of the
// 64 K register file) and the sum proceeds UNDER the backward GEMMs instead of between them.
template <typename T, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
    p2p_all_reduce_kernel(PeerTable t, int rank, int world, size_t in_off, size_t out_off,
                          size_t npk, int two_shot, unsigned long long timeout_ns) {
  constexpr int PEERS = THREADS >= 512 ? 8 : 4;  // peers whose loads are in flight together
  Control *mine = reinterpret_cast<Control *>(t.base[rank]);
  unsigned const epoch = mine->epoch + 1;  // every block reads it before the last block bumps it
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  size_t const nthreads = size_t(gridDim.x) * blockDim.x;
  unsigned char *out_local = t.base[rank] + kFlagBytes + out_off;

  // a group that has failed once fails fast from then on
  if (__syncthreads_or(threadIdx.x == 0 ? int(ld_acquire_sys(&mine->timed_out)) : 0)) {
    nan_fill(out_local, npk);
    return;
  }
  // ---- entry: every rank's input is complete (stream order on its side) once its kernel runs
  if (blockIdx.x == 0) signal_peers(t, rank, world, 0, epoch);
  if (!wait_peers(t, mine, world, 0, epoch, timeout_ns)) {
    nan_fill(out_local, npk);
    return;
  }

  if (!two_shot) {
    for (size_t i = tid; i < npk; i += nthreads) {
      Acc16<T> acc;
      acc.zero();
      for (int p0 = 0; p0 < world; p0 += PEERS) {  // the peers' loads go out before the first add
        uint4 raw[PEERS];
#pragma unroll
        for (int j = 0; j < PEERS; ++j)
          if (p0 + j < world) raw[j] = ld_peer16(t.base[p0 + j] + kFlagBytes + in_off + i * 16);
#pragma unroll
        for (int j = 0; j < PEERS; ++j)
          if (p0 + j < world) acc.add(raw[j]);
      }
      *reinterpret_cast<uint4 *>(out_local + i * 16) = acc.pack();
    }
  } else {
    size_t const per = (npk + world - 1) / world;
    size_t const lo = size_t(rank) * per, hi = min(npk, lo + per);
    // Two packets per thread and trip from EVERY peer are requested before the first is added
    // (16 loads in flight per thread at 8 ranks); the additions then run in rank order.
    // (A warp's load covers 512 contiguous bytes; a thread's packets are a block apart.)
    size_t const bstride = size_t(blockDim.x);
    for (size_t i0 = lo + size_t(blockIdx.x) * bstride * 2 + threadIdx.x; i0 < hi; i0 += nthreads * 2) {
      Acc16<T> acc[2];
      acc[0].zero();
      acc[1].zero();
      bool const ok1 = i0 + bstride < hi;
      for (int p0 = 0; p0 < world; p0 += PEERS) {
        uint4 raw[PEERS][2];
#pragma unroll
        for (int j = 0; j < PEERS; ++j) {
          if (p0 + j < world) {
            unsigned char const *src = t.base[p0 + j] + kFlagBytes + in_off;
            raw[j][0] = ld_peer16(src + i0 * 16);
            raw[j][1] = ok1 ? ld_peer16(src + (i0 + bstride) * 16) : make_uint4(0, 0, 0, 0);
          }
        }
#pragma unroll
        for (int j = 0; j < PEERS; ++j) {
          if (p0 + j < world) {
            acc[0].add(raw[j][0]);
            acc[1].add(raw[j][1]);
          }
        }
      }
      *reinterpret_cast<uint4 *>(out_local + i0 * 16) = acc[0].pack();
      if (ok1) *reinterpret_cast<uint4 *>(out_local + (i0 + bstride) * 16) = acc[1].pack();
    }
    // ---- middle: my chunk is reduced and visible before any peer gathers it
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      unsigned arrived = atomicAdd(&mine->blocks_mid, 1u) + 1;
      if (arrived == gridDim.x) {
        mine->blocks_mid = 0;
        __threadfence_system();
        st_release_sys(&mine->flags[1][rank], epoch);  // own slot, for the local waiters below
      }
    }
    // the block that completed the count tells the peers; everybody waits on local flags
    int mid_ok = 1;
    if (threadIdx.x == 0)
      mid_ok = spin_until(t, mine, world, &mine->flags[1][rank], epoch, timeout_ns) ? 1 : 0;
    if (!__syncthreads_and(mid_ok)) {
      nan_fill(out_local, npk);
      return;
    }
    if (blockIdx.x == 0) signal_peers(t, rank, world, 1, epoch);
    if (!wait_peers(t, mine, world, 1, epoch, timeout_ns)) {
      nan_fill(out_local, npk);
      return;
    }
    // ---- gather the other ranks' reduced chunks
    for (int q = 1; q < world; ++q) {
      int const p = (rank + q) % world;  // spread the ranks over different peers at any moment
      size_t const plo = size_t(p) * per, phi = min(npk, plo + per);
      unsigned char const *src = t.base[p] + kFlagBytes + out_off;
      for (size_t i0 = plo + size_t(blockIdx.x) * bstride * PEERS + threadIdx.x; i0 < phi;
           i0 += nthreads * PEERS) {
        uint4 raw[PEERS];
#pragma unroll
        for (int u = 0; u < PEERS; ++u)
          raw[u] = (i0 + u * bstride < phi) ? ld_peer16(src + (i0 + u * bstride) * 16)
                                            : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int u = 0; u < PEERS; ++u)
          if (i0 + u * bstride < phi)
            *reinterpret_cast<uint4 *>(out_local + (i0 + u * bstride) * 16) = raw[u];
      }
    }
  }

  // ---- exit: nobody may overwrite what a peer is still reading; the last block of this rank
  // signals "done reading", bumps the epoch, and every block leaves only when all peers are done
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    unsigned arrived = atomicAdd(&mine->blocks_done, 1u) + 1;
    if (arrived == gridDim.x) {
      mine->blocks_done = 0;
      for (int p = 0; p < world; ++p)
        st_release_sys(&reinterpret_cast<Control *>(t.base[p])->flags[2][rank], epoch);
      bool ok = true;
      for (int p = 0; p < world && ok; ++p)
        ok = spin_until(t, mine, world, &mine->flags[2][p], epoch, timeout_ns);
      __threadfence_system();
      // no hand-off after a failure: the epoch stays, later launches leave at their entry check
      if (ok) mine->epoch = epoch;  // the next launch on this stream starts after this kernel has ended
    }
  }
}


// Small buffers: one-way protocol. Block b stores this rank's input into rank b's receive slot
// (remote stores are fire-and-forget) and raises a flag behind it; then every block sums a
// slice of the `world` slots that have arrived HERE, in rank order. One switch traversal instead
// of the two round trips of the read-based kernel. Slots alternate with the epoch's parity: a
// peer can only overwrite a slot two launches later, which it cannot start before this rank has
// begun the launch in between, i.e. finished this one.
template <typename T>
__global__ void __launch_bounds__(512)
    p2p_push_all_reduce_kernel(PeerTable t, int rank, int world, size_t in_off, size_t out_off,
                               size_t npk, unsigned long long timeout_ns) {
  Control *mine = reinterpret_cast<Control *>(t.base[rank]);
  unsigned const epoch = mine->epoch + 1;
  int const parity = int(epoch & 1u);
  unsigned char const *in_local = t.base[rank] + kFlagBytes + in_off;
  unsigned char *out_local = t.base[rank] + kFlagBytes + out_off;
  if (__syncthreads_or(threadIdx.x == 0 ? int(ld_acquire_sys(&mine->timed_out)) : 0)) {
    nan_fill(out_local, npk);
    return;
  }
  for (int p = blockIdx.x; p < world; p += gridDim.x) {
    unsigned char *slot = t.base[p] + kControlBytes + (size_t(parity) * kMaxRanks + rank) * kPushLimit;
    for (size_t i = threadIdx.x; i < npk; i += blockDim.x)
      *reinterpret_cast<uint4 *>(slot + i * 16) = *reinterpret_cast<uint4 const *>(in_local + i * 16);
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      st_release_sys(&reinterpret_cast<Control *>(t.base[p])->flags[3 + parity][rank], epoch);
    }
  }
  if (!wait_peers(t, mine, world, 3 + parity, epoch, timeout_ns)) {
    nan_fill(out_local, npk);
    return;
  }
  unsigned char const *slots = t.base[rank] + kControlBytes + size_t(parity) * kMaxRanks * kPushLimit;
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < npk; i += size_t(gridDim.x) * blockDim.x) {
    Acc16<T> acc;
    acc.zero();
    for (int p = 0; p < world; ++p) acc.add(ld_peer16(slots + size_t(p) * kPushLimit + i * 16));
    *reinterpret_cast<uint4 *>(out_local + i * 16) = acc.pack();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned arrived = atomicAdd(&mine->blocks_done, 1u) + 1;
    if (arrived == gridDim.x) {
      mine->blocks_done = 0;
      __threadfence();
      mine->epoch = epoch;
    }
  }
}

}  // namespace
}  // namespace ffk

using namespace ffk;

extern "C" {

int ffk_p2p_create(void **group_out, int rank, int world, size_t data_bytes,
                   unsigned char *handle_out) {
  FFK_REQUIRE(group_out && handle_out, "p2p: null argument");
  FFK_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world,
              "p2p: rank %d of %d (at most %d ranks)", rank, world, kMaxRanks);
  FFK_REQUIRE(data_bytes % 16 == 0, "p2p: region size must be a multiple of 16 bytes");
  Group *g = new Group();
  g->rank = rank;
  g->world = world;
  g->data_bytes = data_bytes;
  if (char const *e = std::getenv("FFK_P2P_TIMEOUT_MS")) {
    long long ms = atoll(e);
    if (ms > 0) g->timeout_ns = (unsigned long long)ms * 1000000ull;
  }
  cudaIpcMemHandle_t h;
  cudaError_t err = cudaGetDevice(&g->device);
  if (err == cudaSuccess) err = cudaMalloc(reinterpret_cast<void **>(&g->local), kFlagBytes + data_bytes);
  if (err == cudaSuccess) err = cudaMemset(g->local, 0, kControlBytes);
  if (err == cudaSuccess) err = cudaIpcGetMemHandle(&h, g->local);
  if (err != cudaSuccess) {  // nothing of a half-built group survives
    if (g->local) cudaFree(g->local);
    delete g;
    FFK_CUDA(err);
  }
  static_assert(sizeof(h) == FFK_P2P_HANDLE_BYTES, "IPC handle size");
  std::memcpy(handle_out, &h, sizeof(h));
  g->table.base[rank] = g->local;
  *group_out = g;
  return FFK_OK;
}

int ffk_p2p_connect(void *group, unsigned char const *handles_all) {
  Group *g = reinterpret_cast<Group *>(group);
  FFK_REQUIRE(g && handles_all, "p2p: null argument");
  FFK_REQUIRE(!g->connected, "p2p: group already connected");
  for (int p = 0; p < g->world; ++p) {
    if (p == g->rank) continue;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handles_all + size_t(p) * FFK_P2P_HANDLE_BYTES, sizeof(h));
    FFK_CUDA(cudaIpcOpenMemHandle(&g->peer_mapped[p], h, cudaIpcMemLazyEnablePeerAccess));
    g->table.base[p] = reinterpret_cast<unsigned char *>(g->peer_mapped[p]);
  }
  g->connected = true;
  return FFK_OK;
}

int ffk_p2p_local_buffer(void *group, void **data_out, size_t *bytes_out) {
  Group *g = reinterpret_cast<Group *>(group);
  FFK_REQUIRE(g && data_out, "p2p: null argument");
  *data_out = g->local + kFlagBytes;
  if (bytes_out) *bytes_out = g->data_bytes;
  return FFK_OK;
}

}  // extern "C"

namespace {
template <int THREADS, int MINB>
void launch_all_reduce(int dtype, unsigned blocks, cudaStream_t s, Group *g, size_t in_offset,
                       size_t out_offset, size_t npk, int two_shot) {
  switch (dtype) {
    case FFK_FLOAT:
      p2p_all_reduce_kernel<float, THREADS, MINB><<<blocks, THREADS, 0, s>>>(
          g->table, g->rank, g->world, in_offset, out_offset, npk, two_shot, g->timeout_ns);
      break;
    case FFK_HALF:
      p2p_all_reduce_kernel<__half, THREADS, MINB><<<blocks, THREADS, 0, s>>>(
          g->table, g->rank, g->world, in_offset, out_offset, npk, two_shot, g->timeout_ns);
      break;
    default:
      p2p_all_reduce_kernel<__nv_bfloat16, THREADS, MINB><<<blocks, THREADS, 0, s>>>(
          g->table, g->rank, g->world, in_offset, out_offset, npk, two_shot, g->timeout_ns);
      break;
  }
}

int all_reduce_impl(void *stream, void *group, size_t in_offset, size_t out_offset, size_t count,
                    int dtype, int background_blocks) {
  Group *g = reinterpret_cast<Group *>(group);
  FFK_REQUIRE(g, "p2p: null group");
  FFK_REQUIRE(g->connected || g->world == 1, "p2p: ffk_p2p_connect has not been called");
  size_t es = dtype_size(dtype);
  FFK_REQUIRE(dtype == FFK_FLOAT || dtype == FFK_HALF || dtype == FFK_BFLOAT16,
              "p2p all-reduce: dtype %d not supported (FLOAT, HALF, BFLOAT16)", dtype);
  size_t bytes = count * es;
  if (count == 0) return FFK_OK;
  FFK_REQUIRE(bytes % 16 == 0 && in_offset % 16 == 0 && out_offset % 16 == 0,
              "p2p all-reduce: offsets and size must be multiples of 16 bytes");
  FFK_REQUIRE(in_offset + bytes <= g->data_bytes && out_offset + bytes <= g->data_bytes,
              "p2p all-reduce: range exceeds the region");
  FFK_REQUIRE(in_offset + bytes <= out_offset || out_offset + bytes <= in_offset,
              "p2p all-reduce: input and output ranges overlap");
  size_t npk = bytes / 16;
  cudaStream_t s = as_stream(stream);
  if (bytes <= kPushLimit && background_blocks <= 0) {
    unsigned blocks = unsigned(g->world);
    switch (dtype) {
      case FFK_FLOAT:
        p2p_push_all_reduce_kernel<float><<<blocks, 512, 0, s>>>(g->table, g->rank, g->world, in_offset, out_offset, npk, g->timeout_ns);
        break;
      case FFK_HALF:
        p2p_push_all_reduce_kernel<__half><<<blocks, 512, 0, s>>>(g->table, g->rank, g->world, in_offset, out_offset, npk, g->timeout_ns);
        break;
      default:
        p2p_push_all_reduce_kernel<__nv_bfloat16><<<blocks, 512, 0, s>>>(g->table, g->rank, g->world, in_offset, out_offset, npk, g->timeout_ns);
        break;
    }
    FFK_LAUNCH_CHECK();
    return FFK_OK;
  }
  int two_shot = bytes > kOneShotLimit && g->world > 1;
  if (background_blocks > 0) {
    unsigned blocks = unsigned(std::min<size_t>(size_t(background_blocks), (npk + 127) / 128));
    launch_all_reduce<128, 6>(dtype, blocks, s, g, in_offset, out_offset, npk, two_shot);
    FFK_LAUNCH_CHECK();
    return FFK_OK;
  }
  // measured inside bench.py at 2 GPUs (32 MiB bucket under the memory-bound tail of the step):
  // 24 blocks 1.221 ms/step, 48 1.182, 96 1.156, 148 1.165, NCCL 1.168. FFK_P2P_BLOCKS overrides.
  unsigned blocks = two_shot ? 96u : unsigned(std::min<size_t>(16, (npk + 511) / 512));
  if (two_shot) {
    if (char const *e = std::getenv("FFK_P2P_BLOCKS")) blocks = unsigned(std::max(1, atoi(e)));
  }
  launch_all_reduce<512, 1>(dtype, blocks, s, g, in_offset, out_offset, npk, two_shot);
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}
}  // namespace

extern "C" {

int ffk_p2p_all_reduce_sum(void *stream, void *group, size_t in_offset, size_t out_offset,
                           size_t count, int dtype) {
  return all_reduce_impl(stream, group, in_offset, out_offset, count, dtype, 0);
}

int ffk_p2p_all_reduce_sum_background(void *stream, void *group, size_t in_offset,
                                      size_t out_offset, size_t count, int dtype, int blocks) {
  FFK_REQUIRE(blocks > 0 && blocks <= 1024, "p2p background all-reduce: blocks must be in 1..1024");
  return all_reduce_impl(stream, group, in_offset, out_offset, count, dtype, blocks);
}

int ffk_p2p_status(void *group) {
  Group *g = reinterpret_cast<Group *>(group);
  FFK_REQUIRE(g, "p2p: null group");
  Control c;
  FFK_CUDA(cudaMemcpy(&c, g->local, sizeof(c), cudaMemcpyDeviceToHost));
  if (c.timed_out) {
    set_error("p2p: a wait for a peer rank expired on some rank of the group (a rank did not reach a "
              "collective in time); results since then are NaN and the group must be recreated");
    return FFK_ERR_CUDA;
  }
  return FFK_OK;
}

int ffk_p2p_destroy(void *group) {
  Group *g = reinterpret_cast<Group *>(group);
  if (!g) return FFK_OK;
  for (int p = 0; p < g->world; ++p)
    if (g->peer_mapped[p]) cudaIpcCloseMemHandle(g->peer_mapped[p]);
  if (g->local) cudaFree(g->local);
  delete g;
  return FFK_OK;
}

}  // extern "C"
