// tcgen05 / TMEM / TMA GEMM for sm_100a plus the CUDA-core fp32-exact GEMM.
// See gemm_sm100.cuh for the contract.
#include "gemm_sm100.cuh"
#include "philox.cuh"

#include <cudaTypedefs.h>
#include <cstdlib>
#include <mutex>

namespace ffk {
namespace gemm {

namespace {

constexpr int BLOCK_M = 128;       // rows per CTA tile = TMEM lanes
constexpr int SWIZZLE_BYTES = 128; // TMA / UMMA swizzle span
constexpr int MMA_K_BYTES = 32;    // one tcgen05.mma consumes 32 bytes of K per row
constexpr int K_STEPS = SWIZZLE_BYTES / MMA_K_BYTES;
constexpr int GROUP_M = 8;         // tile rasterisation: 8 row-blocks share B columns in L2
constexpr int NUM_THREADS = 320;   // warp0 TMA, warp1 MMA, warps 2..9 epilogue
constexpr int EPI_WARPS = 8;       // two warps per TMEM lane quadrant, each drains half of the tile's columns
constexpr int EPI_STAGE_PER_WARP = 32 * 64;  // 32 rows x 32 sixteen-bit outputs, transposed through smem
constexpr int EPI_STAGING_BYTES = EPI_WARPS * EPI_STAGE_PER_WARP;

template <int BLOCK_N, int STAGES> struct SmemPlan {
  static constexpr int A_BYTES = BLOCK_M * SWIZZLE_BYTES;
  static constexpr int B_BYTES = BLOCK_N * SWIZZLE_BYTES;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
  static constexpr int STAGING_OFFSET = BAR_OFFSET + 256;  // after the barriers
  static constexpr int TOTAL = STAGING_OFFSET + EPI_STAGING_BYTES + 1024;  // + alignment slack
};

struct TileCoord {
  int b, m_idx, n_idx;
};

__device__ __forceinline__ TileCoord decode_tile(int tile, int num_m, int num_n, int block_n) {
  int per_batch = num_m * num_n;
  int b = tile / per_batch;
  int t = tile - b * per_batch;
  int group_span = GROUP_M * num_n;
  int g = t / group_span;
  int first_m = g * GROUP_M;
  int gsize = min(num_m - first_m, GROUP_M);
  int in_g = t - g * group_span;
  TileCoord c;
  c.b = b;
  c.m_idx = (first_m + in_g % gsize) * BLOCK_M;
  c.n_idx = (in_g / gsize) * block_n;
  return c;
}

template <typename T>
__device__ __forceinline__ void epilogue_row_chunk(uint32_t const (&acc)[32], Epilogue const &ep,
                                                   long long row_off, int col0, int ncols_valid,
                                                   bool vec_ok) {
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(acc[i]) * ep.alpha;
  T *drow = reinterpret_cast<T *>(ep.d) + row_off + col0;
  constexpr int VN = 16 / sizeof(T);
  bool const full = vec_ok && ncols_valid == 32;
  if (ep.add_d_before_act) {
    if (full) {
#pragma unroll
      for (int j = 0; j < 32 / VN; ++j) {
        Vec16<T> t = ld_vec(drow + j * VN);
#pragma unroll
        for (int i = 0; i < VN; ++i) v[j * VN + i] += to_f32(t.v[i]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < ncols_valid) v[i] += to_f32(drow[i]);
    }
  }
  if (ep.bias != nullptr) {
    T const *bp = reinterpret_cast<T const *>(ep.bias) + col0;
    if (full && ep.bias_vec_ok) {
#pragma unroll
      for (int j = 0; j < 32 / VN; ++j) {
        Vec16<T> t = ld_vec(bp + j * VN);
#pragma unroll
        for (int i = 0; i < VN; ++i) v[j * VN + i] += to_f32(t.v[i]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < ncols_valid) v[i] += to_f32(bp[i]);
    }
  }
  // one branch per chunk, not per element
  if (ep.act == ACT_RELU) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
  } else if (ep.act != ACT_NONE) {
    if constexpr (sizeof(T) == 2) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = apply_act_fast(v[i], ep.act);
    } else {
#pragma unroll 4
      for (int i = 0; i < 32; ++i) v[i] = apply_act(v[i], ep.act);
    }
  }
  if (full) {
    if (ep.r != nullptr) {
      T const *rp = reinterpret_cast<T const *>(ep.r) + row_off + col0;
#pragma unroll
      for (int j = 0; j < 32 / VN; ++j) {
        Vec16<T> t = ld_vec(rp + j * VN);
#pragma unroll
        for (int i = 0; i < VN; ++i) v[j * VN + i] += ep.lambda * to_f32(t.v[i]);
      }
    }
    if (ep.accumulate) {
#pragma unroll
      for (int j = 0; j < 32 / VN; ++j) {
        Vec16<T> t = ld_vec(drow + j * VN);
#pragma unroll
        for (int i = 0; i < VN; ++i) v[j * VN + i] += to_f32(t.v[i]);
      }
    }
#pragma unroll
    for (int j = 0; j < 32 / VN; ++j) {
      Vec16<T> t;
#pragma unroll
      for (int i = 0; i < VN; ++i) t.v[i] = from_f32<T>(v[j * VN + i]);
      st_vec(drow + j * VN, t);
    }
  } else {
    T const *rp = ep.r ? reinterpret_cast<T const *>(ep.r) + row_off + col0 : nullptr;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      if (i < ncols_valid) {
        float x = v[i];
        if (rp) x += ep.lambda * to_f32(rp[i]);
        if (ep.accumulate) x += to_f32(drow[i]);
        drow[i] = from_f32<T>(x);
      }
    }
  }
}

// 16-bit outputs, full 32-column chunk, 16-byte-aligned rows: the warp turns its 32 x 32 block
// through 2 KB of shared memory so that global memory sees 64-byte row segments written (and,
// for beta = 1 / the regulariser, read) by 4 adjacent lanes — 8 rows per instruction — instead
// of 32 lanes touching 32 different rows with 16 bytes each. Row r keeps its four 16-byte
// pieces at piece ^ ((r >> 1) & 3): conflict-free for both the row-per-thread and the
// 8-rows-by-4-pieces access. Warp-collective: every lane calls it, `row_ok` masks dead rows.
// EXTRAS compiles in the fused-operator steps (pre-activation copy, dropout, gate); the plain
// instantiation is the epilogue every other GEMM runs and carries none of their code.
template <typename T, bool EXTRAS>
__device__ __forceinline__ void epilogue_chunk_staged(uint32_t const (&acc)[32], Epilogue const &ep,
                                                      uint8_t *stage, long long block_off,
                                                      int rows_valid, int col0, bool row_ok) {
  static_assert(sizeof(T) == 2, "staged epilogue is for 16-bit outputs");
  int const lane = threadIdx.x & 31;
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(acc[i]) * ep.alpha;
  if (ep.bias != nullptr) {
    T const *bp = reinterpret_cast<T const *>(ep.bias) + col0;
    if (ep.bias_vec_ok) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        Vec16<T> t = ld_vec(bp + j * 8);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[j * 8 + i] += to_f32(t.v[i]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += to_f32(bp[i]);
    }
  }
  // coalesced side: lane -> (row = it*8 + lane/4, piece = lane%4)
  int const c_row = lane >> 2, c_piece = lane & 3;
  int const my_sw = (lane >> 1) & 3;  // row-per-thread side: this thread's row is `lane`
  // v (row per thread) -> 16 bits -> staging -> 64-byte row segments of `dst`
  auto store_rows = [&](T *dst) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      Vec16<T> t;
#pragma unroll
      for (int i = 0; i < 8; ++i) t.v[i] = from_f32<T>(v[j * 8 + i]);
      *reinterpret_cast<uint4 *>(stage + lane * 64 + ((j ^ my_sw) << 4)) = *reinterpret_cast<uint4 *>(&t);
    }
    __syncwarp();
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      int const r = it * 8 + c_row;
      uint4 t = *reinterpret_cast<uint4 const *>(stage + r * 64 + ((c_piece ^ ((r >> 1) & 3)) << 4));
      if (r < rows_valid) *reinterpret_cast<uint4 *>(dst + block_off + r * ep.ldd + col0 + c_piece * 8) = t;
    }
    __syncwarp();
  };
  if constexpr (EXTRAS) {
    if (ep.aux_out != nullptr) store_rows(reinterpret_cast<T *>(ep.aux_out));  // the pre-activation
  }
  if (ep.act == ACT_RELU) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
  } else if (ep.act != ACT_NONE) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = apply_act_fast(v[i], ep.act);
  }
  if (EXTRAS && ep.drop_on) {
    // this thread's 32 elements are one mask word: element e keeps iff field e % 8 of
    // Philox(e / 8, offset, seed) >= threshold — the rule of the stand-alone dropout kernel
    unsigned long long const e0 = (unsigned long long)(block_off + (long long)lane * ep.ldd + col0);
    uint32_t bits = 0;
    if (ep.drop_thresh < 0x10000u) {
#pragma unroll
      for (int q = 0; q < 4; ++q)
        bits |= philox_keep8(e0 / 8 + q, ep.drop_offset, ep.drop_seed, ep.drop_thresh) << (q * 8);
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = ((bits >> i) & 1u) ? v[i] * ep.drop_scale : 0.f;
    if (ep.drop_mask != nullptr && row_ok) ep.drop_mask[e0 / 32] = bits;
  }
  auto add_from_global = [&](T const *src, float scale, bool gate) {
    uint4 t[4];
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      int const r = it * 8 + c_row;
      t[it] = r < rows_valid
                  ? *reinterpret_cast<uint4 const *>(src + block_off + r * ep.ldd + col0 + c_piece * 8)
                  : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      int const r = it * 8 + c_row;
      *reinterpret_cast<uint4 *>(stage + r * 64 + ((c_piece ^ ((r >> 1) & 3)) << 4)) = t[it];
    }
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      Vec16<T> t2;
      *reinterpret_cast<uint4 *>(&t2) =
          *reinterpret_cast<uint4 const *>(stage + lane * 64 + ((j ^ my_sw) << 4));
      if (EXTRAS && gate) {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[j * 8 + i] = to_f32(t2.v[i]) > 0.f ? v[j * 8 + i] * scale : 0.f;
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[j * 8 + i] += scale * to_f32(t2.v[i]);
      }
    }
    __syncwarp();
  };
  if (ep.r != nullptr) add_from_global(reinterpret_cast<T const *>(ep.r), ep.lambda, ep.r_mode == 1);
  if (ep.accumulate) add_from_global(reinterpret_cast<T const *>(ep.d), 1.f, false);
  store_rows(reinterpret_cast<T *>(ep.d));
}

// 16-bit outputs whose rows are NOT 16-byte aligned (a 30522-column vocabulary): same transposition,
// but global memory is written element by element with the 32 lanes on 32 consecutive columns of
// one row (64 contiguous bytes per instruction) instead of every lane walking its own row.
// Plain epilogues only (alpha, bias, activation); ragged right edge and bottom handled.
template <typename T>
__device__ __forceinline__ void epilogue_chunk_staged_unaligned(uint32_t const (&acc)[32],
                                                                Epilogue const &ep, uint8_t *stage,
                                                                long long block_off, int rows_valid,
                                                                int col0, int valid) {
  int const lane = threadIdx.x & 31;
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(acc[i]) * ep.alpha;
  if (ep.bias != nullptr) {
    T const *bp = reinterpret_cast<T const *>(ep.bias) + col0;
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < valid) v[i] += to_f32(bp[i]);
  }
  if (ep.act == ACT_RELU) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
  } else if (ep.act != ACT_NONE) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = apply_act_fast(v[i], ep.act);
  }
  int const my_sw = (lane >> 1) & 3;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    Vec16<T> t;
#pragma unroll
    for (int i = 0; i < 8; ++i) t.v[i] = from_f32<T>(v[j * 8 + i]);
    *reinterpret_cast<uint4 *>(stage + lane * 64 + ((j ^ my_sw) << 4)) = *reinterpret_cast<uint4 *>(&t);
  }
  __syncwarp();
  T *d = reinterpret_cast<T *>(ep.d) + block_off + col0 + lane;
  int const piece = lane >> 3, within = lane & 7;
  if (lane < valid) {
#pragma unroll 8
    for (int r = 0; r < 32; ++r) {
      if (r < rows_valid) {
        T const x = *reinterpret_cast<T const *>(stage + r * 64 + ((piece ^ ((r >> 1) & 3)) << 4) + within * 2);
        d[static_cast<long long>(r) * ep.ldd] = x;
      }
    }
  }
  __syncwarp();
}

// ---- K ranges of one output tile shared between CTA pairs (see gemm_tc_pair_kernel) ----
// A pair that computed a K range NOT starting at 0 parks its raw fp32 accumulators in its own
// slot of the scratch; the pair that computed the range starting at 0 adds them (in slot order,
// so the sum is reproducible) before the ordinary epilogue. One slot = 16 warp positions
// (2 CTAs x 8 epilogue warps) x 16 KB; float4 index (chunk * 8 + i) * 32 + lane inside a
// position, so that every instruction of a warp touches 512 contiguous bytes.
constexpr int SK_POSITIONS = 2 * EPI_WARPS;
constexpr int SK_POS_FLOAT4 = 4 * 8 * 32;                      // 4 chunks of 32 columns x 32 rows
constexpr size_t SK_FLAG_BYTES = 16384;                        // flags of up to 256 slots
constexpr size_t SK_SLOT_BYTES = size_t(SK_POSITIONS) * SK_POS_FLOAT4 * sizeof(float4);

struct SkEpilogue {
  int mode = 0;                    // 0 whole tile, 1 park the partial, 2 add `count` parked partials
  float4 *park = nullptr;          // mode 1: this warp position inside this pair's slot
  uint32_t *park_flag = nullptr;
  float4 const *take = nullptr;    // mode 2: this warp position inside the first contributor's slot
  uint32_t const *take_flag = nullptr;
  int count = 0;                   // mode 2: contributors occupy consecutive slots
  uint32_t wait_ms = 0;            // mode 2: bound of the wait for each contributor
};

// K-cut schedule. A launch whose tile count does not fill its last wave of CTA pairs keeps the
// leading `whole_tiles` tiles (a multiple of the pair count) whole, dealt round the pairs, and
// cuts the `cut_tiles` tiles of the last wave along K so that every pair works to the end:
//  * few tiles (2 x cut_tiles <= pairs): every cut tile is cut at the same k-blocks into
//    `pieces` runs of `piece_k`, one pair per piece (pair = tile * pieces + piece);
//  * many tiles: pair t < cut_tiles computes k-blocks [0, piece_k) of tile t, and each of the
//    remaining pairs computes the tails [piece_k, num_k) of `tails` consecutive tiles.
// Either way all pairs that share an operand panel sweep it at the same k at the same time —
// the property that lets one read from HBM serve them all, and that a split into equal runs of
// the linearised (tile, k) space does not have (measured: 64 tiles of 512 k-blocks on 74 pairs
// ran 32 % SLOWER that way, HBM-bound on re-reads). A run that does not start at k = 0 is
// parked (role 1) in slot `slot`; the run that does (role 2) adds `count` parked partials
// from consecutive slots. A pair never parks after it has waited, so waits cannot chain.
struct SkArgs {
  int whole_tiles = 0, cut_tiles = 0;
  int pieces = 0;   // >= 2: uniform pieces; 0: heads and tails
  int piece_k = 0;  // k-blocks per piece, or the length of the head
  int tails = 0;    // heads and tails: tails per helper pair
};
struct SkSegment {
  int tile, kb0, kb1;
  int role;         // 0 whole tile, 1 park, 2 own the tile and take `count` partials
  int slot, count;
};
// Host-callable: the CPU tests walk the same code (ffk_gemm_kcut_segments). One `next` per
// segment, so that the three role loops of the kernel keep a single body each.
struct SkWalk {
  SkArgs a;
  int pair, pairs, num_k, next_tile, whole_end, cut_i, cut_end;
  __host__ __device__ SkWalk(int pair_, int pairs_, int num_tiles, int num_k_, SkArgs const &args)
      : a(args), pair(pair_), pairs(pairs_), num_k(num_k_), next_tile(pair_) {
    whole_end = a.cut_tiles > 0 ? a.whole_tiles : num_tiles;
    cut_i = cut_end = 0;
    if (a.cut_tiles > 0) {
      if (a.pieces >= 2) {
        cut_end = pair < a.cut_tiles * a.pieces ? 1 : 0;
      } else if (pair < a.cut_tiles) {
        cut_end = 1;
      } else {
        cut_i = (pair - a.cut_tiles) * a.tails;
        cut_end = cut_i + a.tails < a.cut_tiles ? cut_i + a.tails : a.cut_tiles;
      }
    }
  }
  __host__ __device__ __forceinline__ bool next(SkSegment &g) {
    if (next_tile < whole_end) {
      g.tile = next_tile;
      g.kb0 = 0;
      g.kb1 = num_k;
      g.role = 0;
      g.slot = g.count = 0;
      next_tile += pairs;
      return true;
    }
    if (cut_i >= cut_end) return false;
    if (a.pieces >= 2) {
      int const t = pair / a.pieces, j = pair - t * a.pieces;
      g.tile = a.whole_tiles + t;
      g.kb0 = j * a.piece_k;
      g.kb1 = g.kb0 + a.piece_k < num_k ? g.kb0 + a.piece_k : num_k;
      g.role = j == 0 ? 2 : 1;
      g.slot = t * (a.pieces - 1) + (j == 0 ? 0 : j - 1);
      g.count = j == 0 ? a.pieces - 1 : 0;
    } else if (pair < a.cut_tiles) {
      g.tile = a.whole_tiles + pair;
      g.kb0 = 0;
      g.kb1 = a.piece_k;
      g.role = 2;
      g.slot = pair;
      g.count = 1;
    } else {
      g.tile = a.whole_tiles + cut_i;
      g.kb0 = a.piece_k;
      g.kb1 = num_k;
      g.role = 1;
      g.slot = cut_i;
      g.count = 0;
    }
    ++cut_i;
    return true;
  }
};

__device__ __forceinline__ void sk_publish(uint32_t *flag) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flag), "r"(1u) : "memory");
}
// Bounded wait: a contributor that does not arrive within `wait_ms` turns the tile into NaN
// instead of hanging the device. The bound is long (FFK_GEMM_STREAMK_WAIT_MS, default 120 s):
// a contributor pair may not even be resident yet while another kernel holds its SMs — e.g. a
// peer-memory collective waiting up to its own 30 s limit for a late rank.
__device__ __forceinline__ bool sk_arrived(uint32_t const *flag, uint32_t wait_ms) {
  uint32_t v;
  unsigned long long t0 = 0;
  for (uint32_t spin = 0;; ++spin) {
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
    if (v != 0u) return true;
    if ((spin & 1023u) == 1023u) {
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > 1000000ull * wait_ms) return false;
    }
    __nanosleep(64);
  }
}

// One warp's share of a tile epilogue: NCHUNK chunks of 32 accumulator columns for the 32 rows
// (TMEM lanes) of its quadrant. The TMEM load of chunk c+1 is in flight while chunk c is
// converted and stored; `release` runs once the last load has landed (the accumulator buffer
// may then be overwritten by the next tile's MMAs).
template <int NCHUNK, bool EXTRAS, typename Release>
__device__ __forceinline__ void epilogue_columns(uint32_t taddr, Epilogue const &ep, long long block_off,
                                                 int rows_valid, int N, int col_first, int out_dtype,
                                                 int vec_ok, uint8_t *stage, Release release,
                                                 SkEpilogue const *sk = nullptr) {
  int const lane = threadIdx.x & 31;
  bool const row_ok = lane < rows_valid;
  long long const row_off = block_off + static_cast<long long>(lane) * ep.ldd;
  bool const staged_ok = vec_ok && out_dtype != FFK_FLOAT && !ep.add_d_before_act;
  bool const unaligned_ok = !vec_ok && out_dtype != FFK_FLOAT && !ep.add_d_before_act &&
                            !ep.accumulate && ep.r == nullptr;
  uint32_t r[2][32];
  ptx::tmem_ld_32x32(taddr, r[0]);
  bool sk_lost = false;
  if (sk != nullptr && sk->mode == 2) {
    int ok = 1;
    if (lane == 0)
      for (int j = 0; j < sk->count; ++j) ok &= sk_arrived(sk->take_flag + j * SK_POSITIONS, sk->wait_ms) ? 1 : 0;
    sk_lost = __shfl_sync(0xffffffffu, ok, 0) == 0;
  }
#pragma unroll
  for (int c = 0; c < NCHUNK; ++c) {
    ptx::tmem_ld_wait_for(r[c & 1]);
    if (c + 1 < NCHUNK) {
      ptx::tmem_ld_32x32(taddr + (c + 1) * 32, r[(c + 1) & 1]);
    } else {
      release();
    }
    if (sk != nullptr && sk->mode == 1) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        __stcg(reinterpret_cast<uint4 *>(sk->park) + (c * 8 + i) * 32 + lane,
               make_uint4(r[c & 1][4 * i], r[c & 1][4 * i + 1], r[c & 1][4 * i + 2], r[c & 1][4 * i + 3]));
      continue;
    }
    if (sk != nullptr && sk->mode == 2) {
      for (int j = 0; j < sk->count; ++j) {
        float4 const *src = sk->take + size_t(j) * (SK_POSITIONS * SK_POS_FLOAT4) + c * 8 * 32 + lane;
        float4 t[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) t[i] = __ldcg(src + i * 32);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          r[c & 1][4 * i] = __float_as_uint(__uint_as_float(r[c & 1][4 * i]) + t[i].x);
          r[c & 1][4 * i + 1] = __float_as_uint(__uint_as_float(r[c & 1][4 * i + 1]) + t[i].y);
          r[c & 1][4 * i + 2] = __float_as_uint(__uint_as_float(r[c & 1][4 * i + 2]) + t[i].z);
          r[c & 1][4 * i + 3] = __float_as_uint(__uint_as_float(r[c & 1][4 * i + 3]) + t[i].w);
        }
      }
      if (sk_lost) {
#pragma unroll
        for (int i = 0; i < 32; ++i) r[c & 1][i] = 0x7fc00000u;
      }
    }
    int const col0 = col_first + c * 32;
    int const valid = min(32, N - col0);
    if (valid <= 0 || rows_valid <= 0) continue;
    if (unaligned_ok) {  // warp-uniform
      if (out_dtype == FFK_BFLOAT16)
        epilogue_chunk_staged_unaligned<__nv_bfloat16>(r[c & 1], ep, stage, block_off, rows_valid, col0, valid);
      else
        epilogue_chunk_staged_unaligned<__half>(r[c & 1], ep, stage, block_off, rows_valid, col0, valid);
    } else if (staged_ok && valid == 32) {  // warp-uniform
      if (out_dtype == FFK_BFLOAT16)
        epilogue_chunk_staged<__nv_bfloat16, EXTRAS>(r[c & 1], ep, stage, block_off, rows_valid, col0, row_ok);
      else
        epilogue_chunk_staged<__half, EXTRAS>(r[c & 1], ep, stage, block_off, rows_valid, col0, row_ok);
    } else if (row_ok) {
      switch (out_dtype) {
        case FFK_FLOAT: epilogue_row_chunk<float>(r[c & 1], ep, row_off, col0, valid, vec_ok); break;
        case FFK_BFLOAT16:
          epilogue_row_chunk<__nv_bfloat16>(r[c & 1], ep, row_off, col0, valid, vec_ok);
          break;
        default: epilogue_row_chunk<__half>(r[c & 1], ep, row_off, col0, valid, vec_ok); break;
      }
    }
  }
  if (sk != nullptr && sk->mode == 1) {
    __threadfence();
    __syncwarp();
    if (lane == 0) sk_publish(sk->park_flag);
  }
}

// ESIZE: bytes per input element (2: f16/bf16, 4: fp32 consumed as tf32).
template <int ESIZE, bool A_MN, bool B_MN, int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(NUM_THREADS, 1)
    gemm_tc_kernel(const __grid_constant__ CUtensorMap map_a,
                   const __grid_constant__ CUtensorMap map_b, Epilogue ep, int M, int N, int K,
                   int batch, int batch2, uint32_t idesc, int out_dtype, int vec_ok) {
  using Plan = SmemPlan<BLOCK_N, STAGES>;
  constexpr int BLOCK_K = SWIZZLE_BYTES / ESIZE;    // elements of K per stage
  constexpr int ATOM_MN = SWIZZLE_BYTES / ESIZE;    // MN elements per swizzle row (MN-major)
  constexpr uint32_t TMEM_COLS = 2 * BLOCK_N;       // double-buffered fp32 accumulator
  static_assert(TMEM_COLS == 64 || TMEM_COLS == 128 || TMEM_COLS == 256 || TMEM_COLS == 512, "");

  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + Plan::BAR_OFFSET);
  uint64_t *empty_bar = full_bar + STAGES;
  uint64_t *tmem_full = empty_bar + STAGES;
  uint64_t *tmem_empty = tmem_full + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_empty + 2);

  int const warp = threadIdx.x >> 5;
  int const lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    ptx::prefetch_tensormap(&map_a);
    ptx::prefetch_tensormap(&map_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tmem_full[s], 1);
      ptx::mbar_init(&tmem_empty[s], EPI_WARPS);  // one arrive per epilogue warp
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) {
    ptx::tmem_alloc(tmem_slot, TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t const tmem_base = *tmem_slot;

  int const num_m = (M + BLOCK_M - 1) / BLOCK_M;
  int const num_n = (N + BLOCK_N - 1) / BLOCK_N;
  int const num_tiles = num_m * num_n * batch;
  int const num_k = (K + BLOCK_K - 1) / BLOCK_K;

  if (warp == 0 && lane == 0) {
    // ------------------------------------------------------------ TMA producer
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      TileCoord tc = decode_tile(tile, num_m, num_n, BLOCK_N);
      for (int kb = 0; kb < num_k; ++kb) {
        ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t *sa = smem + stage * Plan::STAGE_BYTES;
        uint8_t *sb = sa + Plan::A_BYTES;
        ptx::mbar_arrive_expect_tx(&full_bar[stage], Plan::STAGE_BYTES);
        int const k_idx = kb * BLOCK_K;
        if constexpr (!A_MN) {
          ptx::tma_load_4d(sa, &map_a, &full_bar[stage], k_idx, tc.m_idx, tc.b % batch2, tc.b / batch2);
        } else {
#pragma unroll
          for (int j = 0; j < BLOCK_M / ATOM_MN; ++j)
            ptx::tma_load_4d(sa + j * BLOCK_K * SWIZZLE_BYTES, &map_a, &full_bar[stage],
                             tc.m_idx + j * ATOM_MN, k_idx, tc.b % batch2, tc.b / batch2);
        }
        if constexpr (!B_MN) {
          ptx::tma_load_4d(sb, &map_b, &full_bar[stage], k_idx, tc.n_idx, tc.b % batch2, tc.b / batch2);
        } else {
#pragma unroll
          for (int j = 0; j < BLOCK_N / ATOM_MN; ++j)
            ptx::tma_load_4d(sb + j * BLOCK_K * SWIZZLE_BYTES, &map_b, &full_bar[stage],
                             tc.n_idx + j * ATOM_MN, k_idx, tc.b % batch2, tc.b / batch2);
        }
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1 && lane == 0) {
    // ------------------------------------------------------------ MMA issuer
    // K-major operand: rows of 128 B, 8-row groups 1024 B apart; a k-step moves 32 B.
    // MN-major operand: 128-B rows hold 128/ESIZE consecutive MN indices of one k;
    //   8-k groups 1024 B apart (SBO), next MN atom BLOCK_K*128 B further (LBO);
    //   a k-step moves (32/ESIZE) rows = (32/ESIZE)*128 B.
    constexpr uint32_t A_LBO = A_MN ? BLOCK_K * SWIZZLE_BYTES : 0;
    constexpr uint32_t B_LBO = B_MN ? BLOCK_K * SWIZZLE_BYTES : 0;
    constexpr uint32_t A_KSTEP = A_MN ? (MMA_K_BYTES / ESIZE) * SWIZZLE_BYTES : MMA_K_BYTES;
    constexpr uint32_t B_KSTEP = B_MN ? (MMA_K_BYTES / ESIZE) * SWIZZLE_BYTES : MMA_K_BYTES;
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      ptx::mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
      ptx::tc_fence_after();
      uint32_t const d_tmem = tmem_base + acc * BLOCK_N;
      for (int kb = 0; kb < num_k; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        uint32_t const sa = ptx::smem_u32(smem + stage * Plan::STAGE_BYTES);
        uint32_t const sb = sa + Plan::A_BYTES;
#pragma unroll
        for (int k = 0; k < K_STEPS; ++k) {
          uint64_t da = ptx::make_smem_desc_sw128(sa + k * A_KSTEP, A_LBO, 1024);
          uint64_t db = ptx::make_smem_desc_sw128(sb + k * B_KSTEP, B_LBO, 1024);
          if constexpr (ESIZE == 2)
            ptx::mma_f16_ss(d_tmem, da, db, idesc, (kb | k) != 0);
          else
            ptx::mma_tf32_ss(d_tmem, da, db, idesc, (kb | k) != 0);
        }
        ptx::mma_commit(&empty_bar[stage]);  // frees the smem slot when these MMAs retire
        if (kb == num_k - 1) ptx::mma_commit(&tmem_full[acc]);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  } else if (warp >= 2) {
    // ------------------------------------------------------------ epilogue
    int const q = warp & 3;             // TMEM lane quadrant this warp may read
    int const half = (warp - 2) >> 2;   // which half of the tile's columns it drains
    constexpr int NCHUNK = BLOCK_N / 64;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      TileCoord tc = decode_tile(tile, num_m, num_n, BLOCK_N);
      ptx::mbar_wait(&tmem_full[acc], acc_phase);
      ptx::tc_fence_after();
      int const row0 = tc.m_idx + q * 32;  // first row of this warp's 32-row block
      long long const block_off = static_cast<long long>(tc.b / batch2) * ep.d_batch +
                                  static_cast<long long>(tc.b % batch2) * ep.d_batch2 +
                                  static_cast<long long>(row0) * ep.ldd;
      int const col_off = half * (BLOCK_N / 2);
      epilogue_columns<NCHUNK, false>(
          tmem_base + (uint32_t(q * 32) << 16) + acc * BLOCK_N + col_off, ep, block_off,
          min(32, M - row0), N, tc.n_idx + col_off, out_dtype, vec_ok,
          smem + Plan::STAGING_OFFSET + (warp - 2) * EPI_STAGE_PER_WARP, [&] {
            // accumulator fully in registers: hand the TMEM buffer back to the MMA warp
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&tmem_empty[acc]);
          });
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------ CTA-pair kernel
// Two CTAs of one cluster (one TPC) share a 256 x 256 output tile: CTA r owns rows
// [r*128, r*128+128) and loads those rows of A plus columns [r*128, r*128+128) of B, so each
// SM pulls 32 KB per k-block from L2 instead of 48 KB. The leader CTA (rank 0) issues
// tcgen05.mma.cta_group::2 for both; TMA completions of both CTAs are counted on the
// leader's full barrier; commits are multicast to both CTAs' empty / tmem-full barriers;
// both epilogues release the accumulator through the leader's tmem-empty barrier.
template <bool A_MN, bool B_MN, int STAGES, bool EXTRAS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
    gemm_tc_pair_kernel(const __grid_constant__ CUtensorMap map_a,
                        const __grid_constant__ CUtensorMap map_b, Epilogue ep, int M, int N, int K,
                        int batch, int batch2, uint32_t idesc, int out_dtype, int vec_ok,
                        SkArgs sk_args, uint8_t *sk_ws, uint32_t sk_wait_ms) {
  constexpr int BLOCK_N = 256, HALF_N = 128, PAIR_M = 256, ESIZE = 2;
  constexpr int BLOCK_K = SWIZZLE_BYTES / ESIZE, ATOM_MN = SWIZZLE_BYTES / ESIZE;
  constexpr int A_BYTES = BLOCK_M * SWIZZLE_BYTES, B_BYTES = HALF_N * SWIZZLE_BYTES;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t TMEM_COLS = 2 * BLOCK_N;

  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
  uint64_t *empty_bar = full_bar + STAGES;
  uint64_t *tmem_full = empty_bar + STAGES;
  uint64_t *tmem_empty = tmem_full + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_empty + 2);

  int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t const rank = ptx::cluster_ctarank();
  bool const leader = rank == 0;

  if (threadIdx.x == 0) {
    ptx::prefetch_tensormap(&map_a);
    ptx::prefetch_tensormap(&map_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full_bar[s], 1);   // leader's producer; both CTAs' TMA bytes are counted on it
      ptx::mbar_init(&empty_bar[s], 1);  // one multicast commit
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tmem_full[s], 1);
      ptx::mbar_init(&tmem_empty[s], 2 * EPI_WARPS);  // the epilogue warps of both CTAs
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) {
    ptx::tmem_alloc_2cta(tmem_slot, TMEM_COLS);
    ptx::tmem_relinquish_2cta();
  }
  ptx::tc_fence_before();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  uint32_t const tmem_base = *tmem_slot;

  int const num_m = (M + PAIR_M - 1) / PAIR_M;
  int const num_n = (N + BLOCK_N - 1) / BLOCK_N;
  int const num_tiles = num_m * num_n * batch;
  int const num_k = (K + BLOCK_K - 1) / BLOCK_K;
  int const tile0 = ptx::cluster_id_x(), tile_step = ptx::num_clusters_x();

  // tile -> (batch, pair row block, column block); 4 pair-row-blocks (1024 rows) share B in L2
  auto decode = [&](int tile, int &b, int &m_idx, int &n_idx) {
    int per_batch = num_m * num_n;
    b = tile / per_batch;
    int t = tile - b * per_batch;
    constexpr int G = 4;
    int span = G * num_n, g = t / span, first = g * G;
    int gsize = min(num_m - first, G), in_g = t - g * span;
    m_idx = (first + in_g % gsize) * PAIR_M;
    n_idx = (in_g / gsize) * BLOCK_N;
  };
  // all three roles walk the same list of (tile, k range) segments
  SkWalk walk(tile0, tile_step, num_tiles, num_k, sk_args);
  SkSegment seg;

  if (warp == 0 && lane == 0) {
    // ------------------------------------------------------------ TMA producer (both CTAs)
    int stage = 0;
    uint32_t phase = 0;
    while (walk.next(seg)) {
      int const kb0 = seg.kb0, kb1 = seg.kb1;
      int b, m_idx, n_idx;
      decode(seg.tile, b, m_idx, n_idx);
      int const my_m = m_idx + rank * BLOCK_M, my_n = n_idx + rank * HALF_N;
      int const b_in = b % batch2, b_out = b / batch2;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t *sa = smem + stage * STAGE_BYTES;
        uint8_t *sb = sa + A_BYTES;
        uint32_t const lead_full = ptx::map_to_cta(ptx::smem_u32(&full_bar[stage]), 0);
        if (leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
        int const k_idx = kb * BLOCK_K;
        if constexpr (!A_MN) {
          ptx::tma_load_4d_2cta(sa, &map_a, lead_full, k_idx, my_m, b_in, b_out);
        } else {
#pragma unroll
          for (int j = 0; j < BLOCK_M / ATOM_MN; ++j)
            ptx::tma_load_4d_2cta(sa + j * BLOCK_K * SWIZZLE_BYTES, &map_a, lead_full,
                                  my_m + j * ATOM_MN, k_idx, b_in, b_out);
        }
        if constexpr (!B_MN) {
          ptx::tma_load_4d_2cta(sb, &map_b, lead_full, k_idx, my_n, b_in, b_out);
        } else {
#pragma unroll
          for (int j = 0; j < HALF_N / ATOM_MN; ++j)
            ptx::tma_load_4d_2cta(sb + j * BLOCK_K * SWIZZLE_BYTES, &map_b, lead_full,
                                  my_n + j * ATOM_MN, k_idx, b_in, b_out);
        }
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1 && lane == 0 && leader) {
    // ------------------------------------------------------------ MMA issuer (leader only)
    constexpr uint32_t A_LBO = A_MN ? BLOCK_K * SWIZZLE_BYTES : 0;
    constexpr uint32_t B_LBO = B_MN ? BLOCK_K * SWIZZLE_BYTES : 0;
    constexpr uint32_t A_KSTEP = A_MN ? (MMA_K_BYTES / ESIZE) * SWIZZLE_BYTES : MMA_K_BYTES;
    constexpr uint32_t B_KSTEP = B_MN ? (MMA_K_BYTES / ESIZE) * SWIZZLE_BYTES : MMA_K_BYTES;
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    while (walk.next(seg)) {
      int const kb0 = seg.kb0, kb1 = seg.kb1;
      ptx::mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
      ptx::tc_fence_after();
      uint32_t const d_tmem = tmem_base + acc * BLOCK_N;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        uint32_t const sa = ptx::smem_u32(smem + stage * STAGE_BYTES);
        uint32_t const sb = sa + A_BYTES;
#pragma unroll
        for (int k = 0; k < K_STEPS; ++k) {
          uint64_t da = ptx::make_smem_desc_sw128(sa + k * A_KSTEP, A_LBO, 1024);
          uint64_t db = ptx::make_smem_desc_sw128(sb + k * B_KSTEP, B_LBO, 1024);
          ptx::mma_f16_ss_2cta(d_tmem, da, db, idesc, ((kb - kb0) | k) != 0);
        }
        ptx::mma_commit_2cta(&empty_bar[stage], 0b11);
        if (kb == kb1 - 1) ptx::mma_commit_2cta(&tmem_full[acc], 0b11);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  } else if (warp >= 2) {
    // ------------------------------------------------------------ epilogue (both CTAs)
    int const q = warp & 3;
    int const half = (warp - 2) >> 2;
    int acc = 0;
    uint32_t acc_phase = 0;
    while (walk.next(seg)) {
      int b, m_idx, n_idx;
      decode(seg.tile, b, m_idx, n_idx);
      // a K range that does not start at 0 is parked in its slot of the scratch; the range that
      // does collects the slots that hold the rest of the tile
      SkEpilogue sk;
      if (seg.role != 0) {
        int const pos = int(rank) * EPI_WARPS + (warp - 2);
        uint32_t *const flag = reinterpret_cast<uint32_t *>(sk_ws) + seg.slot * SK_POSITIONS + pos;
        float4 *const slot = reinterpret_cast<float4 *>(sk_ws + SK_FLAG_BYTES) +
                             (size_t(seg.slot) * SK_POSITIONS + pos) * SK_POS_FLOAT4;
        sk.mode = seg.role;
        if (seg.role == 1) {
          sk.park = slot;
          sk.park_flag = flag;
        } else {
          sk.take = slot;
          sk.take_flag = flag;
          sk.count = seg.count;
          sk.wait_ms = sk_wait_ms;
        }
      }
      ptx::mbar_wait(&tmem_full[acc], acc_phase);
      ptx::tc_fence_after();
      int const row0 = m_idx + rank * BLOCK_M + q * 32;
      long long const block_off = static_cast<long long>(b / batch2) * ep.d_batch +
                                  static_cast<long long>(b % batch2) * ep.d_batch2 +
                                  static_cast<long long>(row0) * ep.ldd;
      int const col_off = half * (BLOCK_N / 2);
      epilogue_columns<BLOCK_N / 64, EXTRAS>(
          tmem_base + (uint32_t(q * 32) << 16) + acc * BLOCK_N + col_off, ep, block_off,
          min(32, M - row0), N, n_idx + col_off, out_dtype, vec_ok,
          smem + STAGES * STAGE_BYTES + 256 + (warp - 2) * EPI_STAGE_PER_WARP, [&] {
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0)
              ptx::mbar_arrive_cluster(ptx::map_to_cta(ptx::smem_u32(&tmem_empty[acc]), 0));
          }, EXTRAS ? nullptr : &sk);
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  ptx::tc_fence_before();
  ptx::cluster_sync();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_2cta(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------ scores + row softmax
// P = softmax_rows(A B^T) for rows of exactly 512 columns (the attention scores of a 512-token
// sequence), both operands K-major. A CTA pair owns 256 rows x all 512 columns: the fp32 score
// tile fills the whole of TMEM (no double buffer), and the eight epilogue warps of each CTA
// normalise it there — row maximum, exp written back to TMEM with tcgen05.st while the row
// sum accumulates, then scale, convert and store — so the scores never travel to HBM in
// fp32 or unnormalised form. Same barrier protocol as gemm_tc_pair_kernel.
template <int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
    gemm_softmax_pair_kernel(const __grid_constant__ CUtensorMap map_a,
                             const __grid_constant__ CUtensorMap map_b, Epilogue ep, int M, int K,
                             int batch, int batch2, uint32_t idesc, int out_dtype) {
  constexpr int N = 512, HALF_N = 128, PAIR_M = 256, ESIZE = 2;
  constexpr int BLOCK_K = SWIZZLE_BYTES / ESIZE;
  constexpr int A_BYTES = BLOCK_M * SWIZZLE_BYTES, B_BYTES = HALF_N * SWIZZLE_BYTES;
  constexpr int STAGE_BYTES = A_BYTES + 2 * B_BYTES;  // this CTA's quarter of each 256-column half
  constexpr uint32_t TMEM_COLS = N;

  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
  uint64_t *empty_bar = full_bar + STAGES;
  uint64_t *tmem_full = empty_bar + STAGES;
  uint64_t *tmem_empty = tmem_full + 1;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_empty + 1);
  float *exch = reinterpret_cast<float *>(smem + STAGES * STAGE_BYTES + 128);  // [2 halves][128 rows]
  uint8_t *staging = smem + STAGES * STAGE_BYTES + 128 + 1024;

  int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t const rank = ptx::cluster_ctarank();
  bool const leader = rank == 0;

  if (threadIdx.x == 0) {
    ptx::prefetch_tensormap(&map_a);
    ptx::prefetch_tensormap(&map_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    ptx::mbar_init(tmem_full, 1);
    ptx::mbar_init(tmem_empty, 2 * EPI_WARPS);
    ptx::fence_mbar_init();
  }
  if (warp == 2) {
    ptx::tmem_alloc_2cta(tmem_slot, TMEM_COLS);
    ptx::tmem_relinquish_2cta();
  }
  ptx::tc_fence_before();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  uint32_t const tmem_base = *tmem_slot;

  int const num_m = (M + PAIR_M - 1) / PAIR_M;
  int const num_tiles = num_m * batch;
  int const num_k = (K + BLOCK_K - 1) / BLOCK_K;
  int const tile0 = ptx::cluster_id_x(), tile_step = ptx::num_clusters_x();

  if (warp == 0 && lane == 0) {
    // ------------------------------------------------------------ TMA producer (both CTAs)
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = tile0; tile < num_tiles; tile += tile_step) {
      int const b = tile / num_m, m_idx = (tile - b * num_m) * PAIR_M;
      int const my_m = m_idx + rank * BLOCK_M;
      int const b_in = b % batch2, b_out = b / batch2;
      for (int kb = 0; kb < num_k; ++kb) {
        ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t *sa = smem + stage * STAGE_BYTES;
        uint8_t *sb = sa + A_BYTES;
        uint32_t const lead_full = ptx::map_to_cta(ptx::smem_u32(&full_bar[stage]), 0);
        if (leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
        int const k_idx = kb * BLOCK_K;
        ptx::tma_load_4d_2cta(sa, &map_a, lead_full, k_idx, my_m, b_in, b_out);
        ptx::tma_load_4d_2cta(sb, &map_b, lead_full, k_idx, int(rank) * HALF_N, b_in, b_out);
        ptx::tma_load_4d_2cta(sb + B_BYTES, &map_b, lead_full, k_idx, 256 + int(rank) * HALF_N, b_in,
                              b_out);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1 && lane == 0 && leader) {
    // ------------------------------------------------------------ MMA issuer (leader only)
    int stage = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int tile = tile0; tile < num_tiles; tile += tile_step) {
      ptx::mbar_wait(tmem_empty, acc_phase ^ 1);
      ptx::tc_fence_after();
      for (int kb = 0; kb < num_k; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        uint32_t const sa = ptx::smem_u32(smem + stage * STAGE_BYTES);
        uint32_t const sb = sa + A_BYTES;
#pragma unroll
        for (int nh = 0; nh < 2; ++nh) {
#pragma unroll
          for (int k = 0; k < K_STEPS; ++k) {
            uint64_t da = ptx::make_smem_desc_sw128(sa + k * MMA_K_BYTES, 0, 1024);
            uint64_t db = ptx::make_smem_desc_sw128(sb + nh * B_BYTES + k * MMA_K_BYTES, 0, 1024);
            ptx::mma_f16_ss_2cta(tmem_base + nh * 256, da, db, idesc, (kb | k) != 0);
          }
        }
        ptx::mma_commit_2cta(&empty_bar[stage], 0b11);
        if (kb == num_k - 1) ptx::mma_commit_2cta(tmem_full, 0b11);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      acc_phase ^= 1;
    }
  } else if (warp >= 2) {
    // ------------------------------------------------------------ softmax epilogue (both CTAs)
    int const q = warp & 3, half = (warp - 2) >> 2;
    int const row_in_cta = q * 32 + lane;
    uint32_t const taddr = tmem_base + (uint32_t(q * 32) << 16) + half * 256;
    uint8_t *stage_buf = staging + (warp - 2) * EPI_STAGE_PER_WARP;
    constexpr float kLog2e = 1.4426950408889634f;
    uint32_t acc_phase = 0;
    for (int tile = tile0; tile < num_tiles; tile += tile_step) {
      int const b = tile / num_m, m_idx = (tile - b * num_m) * PAIR_M;
      ptx::mbar_wait(tmem_full, acc_phase);
      ptx::tc_fence_after();
      // pass 1: row maximum over this warp's 256 columns, then over both halves. The loops over
      // chunks are NOT unrolled: the three passes must stay inside the instruction cache (a fully
      // unrolled version spent most of its time on instruction fetch and ran 1.5x slower).
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 8; c += 2) {
        uint32_t r0[32], r1[32];
        ptx::tmem_ld_32x32(taddr + c * 32, r0);
        ptx::tmem_ld_32x32(taddr + c * 32 + 32, r1);
        ptx::tmem_ld_wait_for(r0);
        float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          m0 = fmaxf(m0, __uint_as_float(r0[i]));
          m1 = fmaxf(m1, __uint_as_float(r0[i + 1]));
        }
        ptx::tmem_ld_wait_for(r1);
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          m0 = fmaxf(m0, __uint_as_float(r1[i]));
          m1 = fmaxf(m1, __uint_as_float(r1[i + 1]));
        }
        mx = fmaxf(mx, fmaxf(m0, m1));
      }
      exch[half * 128 + row_in_cta] = mx;
      ptx::named_barrier(1, EPI_WARPS * 32);
      mx = fmaxf(mx, exch[(half ^ 1) * 128 + row_in_cta]);
      ptx::named_barrier(1, EPI_WARPS * 32);  // both values read before the slots are reused
      // pass 2: p = exp(s - max) back into TMEM, row sum on the side (ex2.approx; moving half of
      // the exps to an FMA-pipe polynomial measured 5 % slower: the pass is issue-bound, not
      // bound by the special-function unit)
      float const shift = -mx * kLog2e;
      float sum = 0.f;
#pragma unroll 1
      for (int c = 0; c < 8; ++c) {
        uint32_t r[32];
        ptx::tmem_ld_32x32(taddr + c * 32, r);
        ptx::tmem_ld_wait_for(r);
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float const p0 = ptx::ex2_approx(fmaf(__uint_as_float(r[i]), kLog2e, shift));
          float const p1 = ptx::ex2_approx(fmaf(__uint_as_float(r[i + 1]), kLog2e, shift));
          s0 += p0;
          s1 += p1;
          r[i] = __float_as_uint(p0);
          r[i + 1] = __float_as_uint(p1);
        }
        sum += s0 + s1;
        ptx::tmem_st_32x32(taddr + c * 32, r);
      }
      ptx::tmem_st_wait();
      exch[half * 128 + row_in_cta] = sum;
      ptx::named_barrier(1, EPI_WARPS * 32);
      sum += exch[(half ^ 1) * 128 + row_in_cta];
      ptx::named_barrier(1, EPI_WARPS * 32);
      // pass 3: scale by 1 / sum, convert, coalesced store
      Epilogue row_ep = ep;
      row_ep.alpha = 1.f / sum;
      int const row0 = m_idx + int(rank) * BLOCK_M + q * 32;
      long long const block_off = static_cast<long long>(b / batch2) * ep.d_batch +
                                  static_cast<long long>(b % batch2) * ep.d_batch2 +
                                  static_cast<long long>(row0) * ep.ldd;
#pragma unroll 1
      for (int cc = 0; cc < 4; ++cc) {
        epilogue_columns<2, false>(taddr + cc * 64, row_ep, block_off, min(32, M - row0), N,
                            half * 256 + cc * 64, out_dtype, 1, stage_buf, [&] {
                              if (cc == 3) {
                                ptx::tc_fence_before();
                                __syncwarp();
                                if (lane == 0)
                                  ptx::mbar_arrive_cluster(
                                      ptx::map_to_cta(ptx::smem_u32(tmem_empty), 0));
                              }
                            });
      }
      acc_phase ^= 1;
    }
  }

  ptx::tc_fence_before();
  ptx::cluster_sync();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_2cta(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------ host side
PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  });
  return fn;
}

// Stored matrix [batch_outer][batch2][rows][inner] with arbitrary element strides between
// rows (`ld`), inner batches (`batch2_stride`) and outer batches (`batch_stride`).
int make_map(CUtensorMap *out, void const *base, int dtype, uint64_t inner, uint64_t rows,
             uint64_t batch_outer, uint64_t batch2, uint64_t ld, uint64_t batch_stride,
             uint64_t batch2_stride, uint32_t box_inner, uint32_t box_rows) {
  auto fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return FFK_ERR_CUDA;
  }
  size_t es = dtype_size(dtype);
  CUtensorMapDataType dt = dtype == FFK_BFLOAT16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                           : dtype == FFK_HALF   ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                                 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  if (batch2 <= 1) { batch2 = 1; batch2_stride = rows * ld; }
  if (batch_outer <= 1) { batch_outer = 1; batch_stride = batch2_stride * batch2; }
  if (batch2_stride == 0) batch2_stride = rows * ld;
  if (batch_stride == 0) batch_stride = batch2_stride * batch2;
  cuuint64_t dims[4] = {inner, rows, batch2, batch_outer};
  cuuint64_t strides[3] = {ld * es, batch2_stride * es, batch_stride * es};
  cuuint32_t box[4] = {box_inner, box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(out, dt, 4, const_cast<void *>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): inner=%llu rows=%llu batch=%llux%llu ld=%llu",
              (int)r, (unsigned long long)inner, (unsigned long long)rows,
              (unsigned long long)batch_outer, (unsigned long long)batch2, (unsigned long long)ld);
    return FFK_ERR_CUDA;
  }
  return FFK_OK;
}

template <int ESIZE, bool A_MN, bool B_MN, int BLOCK_N, int STAGES>
int launch_tc_inst(Problem const &p, Epilogue const &e, cudaStream_t stream) {
  using Plan = SmemPlan<BLOCK_N, STAGES>;
  constexpr int BLOCK_K = SWIZZLE_BYTES / ESIZE;
  constexpr int ATOM = SWIZZLE_BYTES / ESIZE;
  CUtensorMap ma, mb;
  int rc;
  int const b2 = p.batch2 > 1 ? p.batch2 : 1, bo = p.batch / b2;
  if (!A_MN)
    rc = make_map(&ma, p.a, p.in_dtype, p.K, p.M, bo, b2, p.lda, p.a_batch, p.a_batch2, BLOCK_K, BLOCK_M);
  else
    rc = make_map(&ma, p.a, p.in_dtype, p.M, p.K, bo, b2, p.lda, p.a_batch, p.a_batch2, ATOM, BLOCK_K);
  if (rc) return rc;
  if (!B_MN)
    rc = make_map(&mb, p.b, p.in_dtype, p.K, p.N, bo, b2, p.ldb, p.b_batch, p.b_batch2, BLOCK_K, BLOCK_N);
  else
    rc = make_map(&mb, p.b, p.in_dtype, p.N, p.K, bo, b2, p.ldb, p.b_batch, p.b_batch2, ATOM, BLOCK_K);
  if (rc) return rc;

  auto kern = gemm_tc_kernel<ESIZE, A_MN, B_MN, BLOCK_N, STAGES>;
  FFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Plan::TOTAL));
  uint32_t fmt = p.in_dtype == FFK_BFLOAT16 ? ptx::kFmtBF16
                 : p.in_dtype == FFK_HALF   ? ptx::kFmtF16
                                            : ptx::kFmtTF32;
  uint32_t idesc = ptx::make_idesc(fmt, A_MN, B_MN, BLOCK_M, BLOCK_N);
  int num_m = (p.M + BLOCK_M - 1) / BLOCK_M, num_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  long long tiles = 1LL * num_m * num_n * p.batch;
  int grid = (int)std::min<long long>(tiles, device_info().sm_count);
  size_t os = dtype_size(p.out_dtype);
  int vec_ok = aligned16(e.d) && ((e.ldd * os) % 16 == 0) && ((e.d_batch * os) % 16 == 0) &&
               ((e.d_batch2 * os) % 16 == 0) &&
               (e.r == nullptr || aligned16(e.r));
  Epilogue e2 = e;
  e2.bias_vec_ok = e.bias != nullptr && aligned16(e.bias) && vec_ok;
  kern<<<grid, NUM_THREADS, Plan::TOTAL, stream>>>(ma, mb, e2, p.M, p.N, p.K, p.batch, b2, idesc,
                                                   p.out_dtype, vec_ok);
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

template <int ESIZE, int BLOCK_N, int STAGES>
int launch_tc_layout(Problem const &p, Epilogue const &e, cudaStream_t s) {
  if (!p.a_mn_major && !p.b_mn_major) return launch_tc_inst<ESIZE, false, false, BLOCK_N, STAGES>(p, e, s);
  if constexpr (ESIZE == 4) return FFK_ERR_UNSUPPORTED;
  if (!p.a_mn_major && p.b_mn_major) return launch_tc_inst<ESIZE, false, true, BLOCK_N, STAGES>(p, e, s);
  if (p.a_mn_major && !p.b_mn_major) return launch_tc_inst<ESIZE, true, false, BLOCK_N, STAGES>(p, e, s);
  return launch_tc_inst<ESIZE, true, true, BLOCK_N, STAGES>(p, e, s);
}

struct SkPlan {
  SkArgs args;
  uint8_t *ws = nullptr;
  uint32_t wait_ms = 0;
};
// defined with the registered scratch below
bool sk_begin(Problem const &p, bool extras, long long tiles, int pairs, cudaStream_t stream, SkPlan &plan);
void sk_end(cudaStream_t stream);

template <bool A_MN, bool B_MN>
int launch_tc_pair_inst(Problem const &p, Epilogue const &e, cudaStream_t stream) {
  constexpr int STAGES = 6, BLOCK_K = 64, ATOM = 64, HALF_N = 128;
  constexpr int SMEM = STAGES * (BLOCK_M * SWIZZLE_BYTES + HALF_N * SWIZZLE_BYTES) + 256 +
                       EPI_STAGING_BYTES + 1024;
  CUtensorMap ma, mb;
  int rc;
  int const b2 = p.batch2 > 1 ? p.batch2 : 1, bo = p.batch / b2;
  if (!A_MN)
    rc = make_map(&ma, p.a, p.in_dtype, p.K, p.M, bo, b2, p.lda, p.a_batch, p.a_batch2, BLOCK_K, BLOCK_M);
  else
    rc = make_map(&ma, p.a, p.in_dtype, p.M, p.K, bo, b2, p.lda, p.a_batch, p.a_batch2, ATOM, BLOCK_K);
  if (rc) return rc;
  if (!B_MN)
    rc = make_map(&mb, p.b, p.in_dtype, p.K, p.N, bo, b2, p.ldb, p.b_batch, p.b_batch2, BLOCK_K, HALF_N);
  else
    rc = make_map(&mb, p.b, p.in_dtype, p.N, p.K, bo, b2, p.ldb, p.b_batch, p.b_batch2, ATOM, BLOCK_K);
  if (rc) return rc;
  // the fused-operator epilogue is its own instantiation: plain GEMMs do not carry its code
  bool const extras = e.aux_out != nullptr || e.drop_on || e.r_mode == 1;
  auto kern = extras ? gemm_tc_pair_kernel<A_MN, B_MN, STAGES, true> : gemm_tc_pair_kernel<A_MN, B_MN, STAGES, false>;
  FFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  uint32_t fmt = p.in_dtype == FFK_BFLOAT16 ? ptx::kFmtBF16 : ptx::kFmtF16;
  uint32_t idesc = ptx::make_idesc(fmt, A_MN, B_MN, 256, 256);
  long long tiles = 1LL * ((p.M + 255) / 256) * ((p.N + 255) / 256) * p.batch;
  int const pairs = device_info().sm_count / 2;
  int clusters = (int)std::min<long long>(tiles, pairs);
  size_t os = dtype_size(p.out_dtype);
  int vec_ok = aligned16(e.d) && ((e.ldd * os) % 16 == 0) && ((e.d_batch * os) % 16 == 0) &&
               ((e.d_batch2 * os) % 16 == 0) && (e.r == nullptr || aligned16(e.r));
  Epilogue e2 = e;
  e2.bias_vec_ok = e.bias != nullptr && aligned16(e.bias) && vec_ok;
  // a tile count that leaves the last wave of pairs mostly idle: cut the tiles of that wave along K
  SkPlan sk;
  bool const cut = sk_begin(p, extras, tiles, pairs, stream, sk);
  if (cut) clusters = pairs;
  ::ffk_timer *const timer = ::ffk::gemm_timer_slot();   // ffk_gemm_time_next: bracket THIS launch with the timer's events
  if (timer) cudaEventRecord(timer->start, stream);
  kern<<<clusters * 2, NUM_THREADS, SMEM, stream>>>(ma, mb, e2, p.M, p.N, p.K, p.batch, b2, idesc,
                                                    p.out_dtype, vec_ok, sk.args, sk.ws, sk.wait_ms);
  if (timer) {
    cudaEventRecord(timer->stop, stream);
    ::ffk::gemm_timer_slot() = nullptr;
  }
  if (cut) sk_end(stream);
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int launch_tc_pair(Problem const &p, Epilogue const &e, cudaStream_t s) {
  if (!p.a_mn_major && !p.b_mn_major) return launch_tc_pair_inst<false, false>(p, e, s);
  if (!p.a_mn_major && p.b_mn_major) return launch_tc_pair_inst<false, true>(p, e, s);
  if (p.a_mn_major && !p.b_mn_major) return launch_tc_pair_inst<true, false>(p, e, s);
  return launch_tc_pair_inst<true, true>(p, e, s);
}

int launch_softmax_pair(Problem const &p, Epilogue const &e, cudaStream_t stream) {
  constexpr int STAGES = 3, BLOCK_K = 64, HALF_N = 128;
  constexpr int SMEM = STAGES * (BLOCK_M * SWIZZLE_BYTES + 2 * HALF_N * SWIZZLE_BYTES) + 128 + 1024 +
                       EPI_STAGING_BYTES + 1024;
  CUtensorMap ma, mb;
  int const b2 = p.batch2 > 1 ? p.batch2 : 1, bo = p.batch / b2;
  int rc = make_map(&ma, p.a, p.in_dtype, p.K, p.M, bo, b2, p.lda, p.a_batch, p.a_batch2, BLOCK_K, BLOCK_M);
  if (rc) return rc;
  rc = make_map(&mb, p.b, p.in_dtype, p.K, p.N, bo, b2, p.ldb, p.b_batch, p.b_batch2, BLOCK_K, HALF_N);
  if (rc) return rc;
  auto kern = gemm_softmax_pair_kernel<STAGES>;
  FFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  uint32_t fmt = p.in_dtype == FFK_BFLOAT16 ? ptx::kFmtBF16 : ptx::kFmtF16;
  uint32_t idesc = ptx::make_idesc(fmt, false, false, 256, 256);
  long long tiles = 1LL * ((p.M + 255) / 256) * p.batch;
  int clusters = (int)std::min<long long>(tiles, device_info().sm_count / 2);
  kern<<<clusters * 2, NUM_THREADS, SMEM, stream>>>(ma, mb, e, p.M, p.K, p.batch, b2, idesc,
                                                    p.out_dtype);
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

}  // namespace

bool tc_eligible(Problem const &p, Epilogue const &e) {
  size_t es = dtype_size(p.in_dtype);
  if (!(p.in_dtype == FFK_BFLOAT16 || p.in_dtype == FFK_HALF || p.in_dtype == FFK_FLOAT))
    return false;
  if (!(p.out_dtype == FFK_BFLOAT16 || p.out_dtype == FFK_HALF || p.out_dtype == FFK_FLOAT))
    return false;
  if (p.M <= 0 || p.N <= 0 || p.K <= 0 || p.batch <= 0) return false;
  if (!aligned16(p.a) || !aligned16(p.b)) return false;
  if ((p.lda * es) % 16 || (p.ldb * es) % 16) return false;
  if (p.batch > 1 && ((p.a_batch * es) % 16 || (p.b_batch * es) % 16)) return false;
  if (p.batch2 > 1 && ((p.a_batch2 * es) % 16 || (p.b_batch2 * es) % 16)) return false;
  if (p.batch2 > 1 && p.batch % p.batch2 != 0) return false;
  (void)e;
  return true;
}

int launch_tc(Problem const &p, Epilogue const &e, cudaStream_t stream) {
  if (!tc_eligible(p, e)) {
    set_error("gemm: operands do not satisfy TMA alignment (16-byte base and pitch)");
    return FFK_ERR_UNSUPPORTED;
  }
  bool narrow = p.N <= 128;
  if (p.in_dtype == FFK_FLOAT) {
    if (p.a_mn_major || p.b_mn_major) {
      // TF32 tiles are only wired for K-major operands; the fp32 products of the
      // gradient GEMMs go through the exact CUDA-core path instead.
      set_error("gemm: TF32 tensor path needs K-major operands");
      return FFK_ERR_UNSUPPORTED;
    }
    return narrow ? launch_tc_layout<4, 128, 6>(p, e, stream)
                  : launch_tc_layout<4, 256, 4>(p, e, stream);
  }
  // CTA pairs (256 x 256 tiles) whenever a problem has at least one full pair tile; the
  // single-CTA kernel serves small or narrow outputs. FFK_GEMM_NO_PAIR=1 forces the latter.
  static bool const no_pair = std::getenv("FFK_GEMM_NO_PAIR") != nullptr;
  if (!no_pair && !narrow && p.M > 128) return launch_tc_pair(p, e, stream);
  return narrow ? launch_tc_layout<2, 128, 6>(p, e, stream)
                : launch_tc_layout<2, 256, 4>(p, e, stream);
}

bool fused_eligible(Problem const &p, Epilogue const &e) {
  size_t os = dtype_size(p.out_dtype);
  bool sixteen = (p.in_dtype == FFK_BFLOAT16 || p.in_dtype == FFK_HALF) && p.out_dtype == p.in_dtype;
  static bool const no_pair = std::getenv("FFK_GEMM_NO_PAIR") != nullptr;
  // only the CTA-pair kernel (outputs of more than 128 x 128) instantiates the fused epilogue
  bool const pair = !no_pair && p.N > 128 && p.M > 128;
  return pair && sixteen && p.batch == 1 && p.N % 32 == 0 && e.ldd == p.N && !e.add_d_before_act &&
         aligned16(e.d) && (e.ldd * os) % 16 == 0 && (!e.r || aligned16(e.r)) &&
         (!e.aux_out || aligned16(e.aux_out)) && tc_eligible(p, e);
}

bool softmax_gemm_eligible(Problem const &p, Epilogue const &e) {
  static bool const off = std::getenv("FFK_NO_FUSED_SOFTMAX") != nullptr;
  size_t os = dtype_size(p.out_dtype);
  return !off && tc_eligible(p, e) && p.N == 512 && !p.a_mn_major && !p.b_mn_major &&
         (p.in_dtype == FFK_BFLOAT16 || p.in_dtype == FFK_HALF) && p.out_dtype == p.in_dtype &&
         e.bias == nullptr && e.act == ACT_NONE && !e.accumulate && e.r == nullptr &&
         e.alpha == 1.f && aligned16(e.d) && (e.ldd * os) % 16 == 0 && (e.d_batch * os) % 16 == 0 &&
         (e.d_batch2 * os) % 16 == 0;
}

int launch_softmax_gemm(Problem const &p, Epilogue const &e, cudaStream_t stream) {
  if (!softmax_gemm_eligible(p, e)) {
    set_error("gemm+softmax: needs 512-column K-major 16-bit operands with aligned rows");
    return FFK_ERR_UNSUPPORTED;
  }
  return launch_softmax_pair(p, e, stream);
}

// ------------------------------------------------------------------ SIMT GEMM
namespace {

// 64 x 64 output tile per CTA, 4 x 4 outputs per thread, fp32 FMA. Short problems (a handful of
// tiles, long K — the small-batch Linear shapes of BASELINE configs[0]) would leave most SMs idle
// and serialise the whole K loop in a few CTAs, so K is split across the CTAs of a thread-block
// cluster (cluster dim z = `splits`, 1..8): every CTA accumulates its K range, parks the partial
// tile in its own shared memory, and after a cluster barrier CTA r reduces rows [r*64/splits, ...)
// of all partial tiles through distributed shared memory in rank order (deterministic — no
// atomics, no global scratch) and runs the epilogue for them.
template <typename TIn, typename TOut>
__device__ __forceinline__ void simt_epilogue_store(Epilogue const &ep, TOut *d, TOut const *r,
                                                    TOut const *bias, int gm, int gn, float acc) {
  float v = acc * ep.alpha;
  if (bias) v += to_f32(bias[gn]);
  v = apply_act(v, ep.act);
  long long off = gm * ep.ldd + gn;
  if (r) v += ep.lambda * to_f32(r[off]);
  if (ep.accumulate) v += to_f32(d[off]);
  d[off] = from_f32<TOut>(v);
}

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(256)
    gemm_simt_kernel(TIn const *__restrict__ a, TIn const *__restrict__ b, Epilogue ep, int M,
                     int N, int K, long long sa_m, long long sa_k, long long sb_n, long long sb_k,
                     long long a_batch, long long b_batch, int batch2, long long a_batch2,
                     long long b_batch2, int splits) {
  constexpr int TM = 64, TN = 64, TK = 16;
  __shared__ __align__(16) float As[TK][TM + 4];
  __shared__ __align__(16) float Bs[TK][TN + 4];
  __shared__ __align__(16) float Cs[TM][TN];  // partial tile, read by the cluster peers
  int const rank = blockIdx.z % splits, bz = blockIdx.z / splits;
  int const bo = bz / batch2, bi = bz % batch2;
  a += bo * a_batch + bi * a_batch2;
  b += bo * b_batch + bi * b_batch2;
  int const m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  int const tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  int const k_per = ((K + splits - 1) / splits + TK - 1) / TK * TK;
  int const k_begin = rank * k_per, k_end = min(K, k_begin + k_per);
  float acc[4][4] = {};
  bool const a_kmajor = (sa_k == 1), b_kmajor = (sb_k == 1);
  float ra[4], rb[4];
  auto fetch = [&](int k0) {  // global -> registers for the k-block at k0
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int idx = threadIdx.x + i * 256;
      int mm, kk;
      if (a_kmajor) { mm = idx >> 4; kk = idx & 15; } else { kk = idx >> 6; mm = idx & 63; }
      int gm = m0 + mm, gk = k0 + kk;
      ra[i] = (gm < M && gk < k_end) ? to_f32(a[gm * sa_m + gk * sa_k]) : 0.f;
      int nn;
      if (b_kmajor) { nn = idx >> 4; kk = idx & 15; } else { kk = idx >> 6; nn = idx & 63; }
      int gn = n0 + nn;
      gk = k0 + kk;
      rb[i] = (gn < N && gk < k_end) ? to_f32(b[gn * sb_n + gk * sb_k]) : 0.f;
    }
  };
  if (k_begin < k_end) fetch(k_begin);
  for (int k0 = k_begin; k0 < k_end; k0 += TK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int idx = threadIdx.x + i * 256;
      if (a_kmajor) As[idx & 15][idx >> 4] = ra[i]; else As[idx >> 6][idx & 63] = ra[i];
      if (b_kmajor) Bs[idx & 15][idx >> 4] = rb[i]; else Bs[idx >> 6][idx & 63] = rb[i];
    }
    __syncthreads();
    if (k0 + TK < k_end) fetch(k0 + TK);  // next block's loads fly under this block's FMAs
#pragma unroll
    for (int kk = 0; kk < TK; ++kk) {
      float4 const av = *reinterpret_cast<float4 const *>(&As[kk][ty * 4]);
      float4 const bv = *reinterpret_cast<float4 const *>(&Bs[kk][tx * 4]);
      float const avv[4] = {av.x, av.y, av.z, av.w}, bvv[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(avv[i], bvv[j], acc[i][j]);
    }
    __syncthreads();
  }
  long long const d_off = bo * ep.d_batch + bi * ep.d_batch2;
  TOut *d = reinterpret_cast<TOut *>(ep.d) + d_off;
  TOut const *r = ep.r ? reinterpret_cast<TOut const *>(ep.r) + d_off : nullptr;
  TOut const *bias = reinterpret_cast<TOut const *>(ep.bias);
  if (splits == 1) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int gm = m0 + ty * 4 + i;
      if (gm >= M) continue;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        int gn = n0 + tx * 4 + j;
        if (gn < N) simt_epilogue_store<TIn, TOut>(ep, d, r, bias, gm, gn, acc[i][j]);
      }
    }
    return;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
    *reinterpret_cast<float4 *>(&Cs[ty * 4 + i][tx * 4]) =
        make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  ptx::cluster_sync();  // every partial tile is parked
  int const rows_per = TM / splits;
  uint32_t const cs_base = ptx::smem_u32(&Cs[0][0]);
  for (int e = threadIdx.x; e < rows_per * TN; e += 256) {
    int row = rank * rows_per + e / TN, col = e % TN;
    int gm = m0 + row, gn = n0 + col;
    if (gm >= M || gn >= N) continue;
    float sum = 0.f;
    for (int s2 = 0; s2 < splits; ++s2)
      sum += ptx::ld_dsmem_f32(ptx::map_to_cta(cs_base + uint32_t(row * TN + col) * 4u, uint32_t(s2)));
    simt_epilogue_store<TIn, TOut>(ep, d, r, bias, gm, gn, sum);
  }
  ptx::cluster_sync();  // nobody leaves while a peer still reads its tile
}

// K splits for the cluster: fill the machine (two 256-thread CTAs per SM is plenty for a
// latency-bound problem) while every split keeps at least 64 of K.
int simt_splits(Problem const &p, long long tiles) {
  int s = 1;
  while (s < 8 && tiles * s * 2 <= 2LL * device_info().sm_count && p.K / (s * 2) >= 64) s *= 2;
  return s;
}

template <typename TIn, typename TOut>
int launch_simt_t(Problem const &p, Epilogue const &e, cudaStream_t stream) {
  long long sa_m = p.a_mn_major ? 1 : p.lda, sa_k = p.a_mn_major ? p.lda : 1;
  long long sb_n = p.b_mn_major ? 1 : p.ldb, sb_k = p.b_mn_major ? p.ldb : 1;
  unsigned gx = (p.N + 63) / 64, gy = (p.M + 63) / 64;
  int splits = simt_splits(p, 1LL * gx * gy * p.batch);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(gx, gy, unsigned(p.batch) * splits);
  cfg.blockDim = dim3(256);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = unsigned(splits);
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  FFK_CUDA(cudaLaunchKernelEx(
      &cfg, gemm_simt_kernel<TIn, TOut>, reinterpret_cast<TIn const *>(p.a),
      reinterpret_cast<TIn const *>(p.b), e, p.M, p.N, p.K, sa_m, sa_k, sb_n, sb_k, p.a_batch,
      p.b_batch, p.batch2 > 1 ? p.batch2 : 1, p.a_batch2, p.b_batch2, splits));
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

}  // namespace

int launch_simt(Problem const &p, Epilogue const &e, cudaStream_t stream) {
  if (p.M <= 0 || p.N <= 0 || p.batch <= 0) return FFK_OK;
  if (p.in_dtype == FFK_FLOAT && p.out_dtype == FFK_FLOAT) return launch_simt_t<float, float>(p, e, stream);
  if (p.in_dtype == FFK_BFLOAT16 && p.out_dtype == FFK_BFLOAT16)
    return launch_simt_t<__nv_bfloat16, __nv_bfloat16>(p, e, stream);
  if (p.in_dtype == FFK_BFLOAT16 && p.out_dtype == FFK_FLOAT)
    return launch_simt_t<__nv_bfloat16, float>(p, e, stream);
  if (p.in_dtype == FFK_HALF && p.out_dtype == FFK_HALF) return launch_simt_t<__half, __half>(p, e, stream);
  if (p.in_dtype == FFK_HALF && p.out_dtype == FFK_FLOAT) return launch_simt_t<__half, float>(p, e, stream);
  set_error("gemm: unsupported dtype pair in=%d out=%d", p.in_dtype, p.out_dtype);
  return FFK_ERR_DTYPE;
}

// ------------------------------------------------------------------ split-bf16 fp32 path
namespace {

__global__ void __launch_bounds__(256)
    split_bf16_kernel(float const *__restrict__ x, __nv_bfloat16 *__restrict__ hi,
                      __nv_bfloat16 *__restrict__ lo, size_t n) {
  size_t const nvec = n / 4, stride = size_t(gridDim.x) * blockDim.x;
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    Vec16<float> v = ld_stream(x + i * 4);
    __nv_bfloat16 h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      h[j] = __float2bfloat16_rn(v.v[j]);
      l[j] = __float2bfloat16_rn(v.v[j] - __bfloat162float(h[j]));
    }
    *reinterpret_cast<uint2 *>(hi + i * 4) = *reinterpret_cast<uint2 *>(h);
    *reinterpret_cast<uint2 *>(lo + i * 4) = *reinterpret_cast<uint2 *>(l);
  }
  for (size_t i = nvec * 4 + tid; i < n; i += stride) {
    __nv_bfloat16 h = __float2bfloat16_rn(x[i]);
    hi[i] = h;
    lo[i] = __float2bfloat16_rn(x[i] - __bfloat162float(h));
  }
}

// Scratch registered with ffk_set_gemm_workspace: one slot per DEVICE (the registration and
// every use look the slot up by the current device, so a process that drives several GPUs keeps
// them apart). Uses on different streams of one device are serialised with an event recorded
// behind the last consumer; a use inside a stream capture cannot be ordered against another
// stream and is refused there.
struct Scratch {
  void *ptr = nullptr;
  size_t bytes = 0;
  int fp32_tensor_math = 0;      // ffk_set_gemm_fp32_tensor_math: AUTO may split fp32 into bf16 terms
  int kcut_off = 0;              // ffk_set_gemm_kcut(0): whole tiles only on this device
  cudaStream_t last_stream = nullptr;
  bool used = false;             // last_stream is meaningful
  bool last_captured = false;    // the last use was recorded into a CUDA graph
  cudaEvent_t last_use = nullptr;
};
constexpr int kMaxDevices = 64;
Scratch g_scratch_table[kMaxDevices];
std::mutex g_scratch_mu;

Scratch *scratch_slot() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
  return &g_scratch_table[dev];
}

bool stream_capturing(cudaStream_t s) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(s, &st) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return st != cudaStreamCaptureStatusNone;
}

// Order `stream` behind the scratch's previous consumer. Returns FFK_ERR_UNSUPPORTED when that
// is impossible (the caller then takes the path that needs no scratch).
int scratch_acquire(Scratch &sc, cudaStream_t stream) {
  if (!sc.used || sc.last_stream == stream) return FFK_OK;
  if (sc.last_captured || stream_capturing(stream)) {
    set_error("gemm scratch: last used on another stream inside / outside a stream capture");
    return FFK_ERR_UNSUPPORTED;
  }
  if (sc.last_use && cudaStreamWaitEvent(stream, sc.last_use, 0) != cudaSuccess) {
    cudaGetLastError();
    return FFK_ERR_CUDA;
  }
  return FFK_OK;
}
void scratch_release(Scratch &sc, cudaStream_t stream) {
  sc.used = true;
  sc.last_stream = stream;
  sc.last_captured = stream_capturing(stream);
  if (sc.last_captured) return;
  if (!sc.last_use && cudaEventCreateWithFlags(&sc.last_use, cudaEventDisableTiming) != cudaSuccess) {
    cudaGetLastError();
    sc.last_use = nullptr;
    return;
  }
  cudaEventRecord(sc.last_use, stream);
}

// Element count of a stored operand (rows x ld for every batch).
size_t operand_elems(long long rows, long long ld, long long batch_stride, long long batch2_stride,
                     int batch, int batch2) {
  size_t one = size_t(rows) * size_t(ld);
  int b2 = batch2 > 1 ? batch2 : 1, bo = batch / b2;
  size_t span = one;
  if (b2 > 1) span = std::max(span, size_t(b2 - 1) * size_t(batch2_stride) + one);
  if (bo > 1) span = std::max(span, size_t(bo - 1) * size_t(batch_stride) + (b2 > 1 ? size_t(b2 - 1) * size_t(batch2_stride) : 0) + one);
  return span;
}

}  // namespace

int set_workspace(void *ptr, size_t bytes) {
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  Scratch *sc = scratch_slot();
  if (!sc) {
    set_error("gemm workspace: no current CUDA device");
    return FFK_ERR_CUDA;
  }
  if (ptr) {
    cudaPointerAttributes attr{};
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaPointerGetAttributes(&attr, ptr) != cudaSuccess || attr.type != cudaMemoryTypeDevice ||
        attr.device != dev) {
      cudaGetLastError();
      set_error("gemm workspace: the pointer is not device memory of the current device %d", dev);
      return FFK_ERR_INVALID;
    }
  }
  sc->ptr = ptr;
  sc->bytes = ptr ? bytes : 0;
  sc->used = false;
  sc->last_captured = false;
  return FFK_OK;
}

int set_kcut(int enable) {
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  Scratch *sc = scratch_slot();
  if (!sc) {
    set_error("gemm K-cut switch: no current CUDA device");
    return FFK_ERR_CUDA;
  }
  sc->kcut_off = enable ? 0 : 1;
  return FFK_OK;
}

int set_fp32_tensor_math(int allow) {
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  Scratch *sc = scratch_slot();
  if (!sc) {
    set_error("gemm fp32 policy: no current CUDA device");
    return FFK_ERR_CUDA;
  }
  sc->fp32_tensor_math = allow ? 1 : 0;
  return FFK_OK;
}

namespace {

// K-cut schedule of the CTA-pair kernel (gemm_tc_pair_kernel, `SkWalk`). Taken when
//  * a scratch is registered for this device and can be ordered behind its last user,
//  * the plain epilogue runs, one batch, and the scratch is not holding this GEMM's operands,
//  * the last wave of 256 x 256 tiles would leave >= 5 % of the whole launch idle
//    (256 tiles on 74 pairs: 13.5 %; 512 tiles: 1.2 % -> left alone), and, when that wave is
//    more than half full, >= 25 % of the pairs idle in it (see below).
// FFK_GEMM_NO_STREAMK=1 turns it off. sk_begin .. sk_end hold g_scratch_mu.
bool sk_arithmetic(long long tiles, int num_k, int pairs, SkArgs &a) {
  a = SkArgs{};
  if (pairs <= 1 || pairs > 256 || tiles <= 0 || num_k < 8) return false;
  long long const waves = (tiles + pairs - 1) / pairs;
  int const rem = int(tiles % pairs);
  if (rem == 0 || tiles - rem > (1 << 30)) return false;
  if ((pairs - rem) * 100LL < 5LL * pairs * waves) return false;
  SkArgs r;
  r.whole_tiles = int(tiles - rem);
  r.cut_tiles = rem;
  if (2 * rem <= pairs) {
    int const want = pairs / rem;                    // pairs available per tile
    int piece_k = (num_k + want - 1) / want;
    if (piece_k < 4) piece_k = 4;                    // a piece shorter than 4 k-blocks is all epilogue
    r.piece_k = piece_k;
    r.pieces = (num_k + piece_k - 1) / piece_k;      // every piece non-empty
    if (r.pieces < 2) return false;
  } else {
    // Heads and tails pay only when a good part of the wave is idle: with 64 tiles on 74 pairs
    // (13.5 % idle) the cut launch measured 4 % SLOWER than the whole tiles (207.6 against
    // 200.1 us) — the device runs under its power cap, so 64 busy pairs clock higher than 74,
    // and the tails' operands are extra HBM traffic. Taken from 25 % idle pairs upwards.
    int const helpers = pairs - rem;
    if (helpers * 4 < pairs) return false;
    r.tails = (rem + helpers - 1) / helpers;
    int head = int((1LL * num_k * r.tails + r.tails) / (r.tails + 1));   // ceil(num_k q / (q + 1))
    if (head >= num_k) head = num_k - 1;
    if (head < 1) return false;
    r.pieces = 0;
    r.piece_k = head;
  }
  a = r;
  return true;
}

bool sk_begin(Problem const &p, bool extras, long long tiles, int pairs, cudaStream_t stream, SkPlan &plan) {
  static bool const off = std::getenv("FFK_GEMM_NO_STREAMK") != nullptr;
  if (off || extras || p.batch != 1 || p.scratch_busy) return false;
  SkArgs args;
  if (!sk_arithmetic(tiles, (p.K + 63) / 64, pairs, args)) return false;
  size_t const need = SK_FLAG_BYTES + size_t(pairs) * SK_SLOT_BYTES;   // never more slots than pairs
  g_scratch_mu.lock();
  Scratch *sc = scratch_slot();
  if (!sc || sc->kcut_off || !sc->ptr || !aligned16(sc->ptr) || sc->bytes < need ||
      scratch_acquire(*sc, stream) != FFK_OK ||
      cudaMemsetAsync(sc->ptr, 0, size_t(pairs) * SK_POSITIONS * sizeof(uint32_t), stream) != cudaSuccess) {
    cudaGetLastError();
    g_scratch_mu.unlock();
    return false;
  }
  static uint32_t const wait_ms = [] {
    char const *v = std::getenv("FFK_GEMM_STREAMK_WAIT_MS");
    long ms = v ? std::atol(v) : 0;
    return uint32_t(ms > 0 ? ms : 120000);
  }();
  plan.wait_ms = wait_ms;
  plan.args = args;
  plan.ws = static_cast<uint8_t *>(sc->ptr);
  return true;
}
void sk_end(cudaStream_t stream) {
  Scratch *sc = scratch_slot();
  if (sc) scratch_release(*sc, stream);
  g_scratch_mu.unlock();
}

}  // namespace

int kcut_plan(int M, int N, int K, int pairs, int *plan5) {
  long long const tiles = 1LL * ((M + 255) / 256) * ((N + 255) / 256);
  SkArgs a;
  bool const taken = M > 128 && N > 128 && K > 0 && sk_arithmetic(tiles, (K + 63) / 64, pairs, a);
  if (plan5) {
    plan5[0] = a.whole_tiles;
    plan5[1] = a.cut_tiles;
    plan5[2] = a.pieces;
    plan5[3] = a.piece_k;
    plan5[4] = a.tails;
  }
  return taken ? 1 : 0;
}

int kcut_segments(int M, int N, int K, int pairs, int pair, int *six, int max_segments) {
  long long const tiles = 1LL * ((M + 255) / 256) * ((N + 255) / 256);
  SkArgs a;
  if (!(M > 128 && N > 128 && K > 0 && sk_arithmetic(tiles, (K + 63) / 64, pairs, a))) a = SkArgs{};
  int n = 0;
  SkWalk walk(pair, pairs, int(tiles), (K + 63) / 64, a);
  SkSegment g;
  while (walk.next(g)) {
    if (n < max_segments && six) {
      int *o = six + 6 * n;
      o[0] = g.tile; o[1] = g.kb0; o[2] = g.kb1; o[3] = g.role; o[4] = g.slot; o[5] = g.count;
    }
    ++n;
  }
  return n;
}

size_t split_workspace_bytes(Problem const &p) {
  size_t ea = operand_elems(p.a_mn_major ? p.K : p.M, p.lda, p.a_batch, p.a_batch2, p.batch, p.batch2);
  size_t eb = operand_elems(p.b_mn_major ? p.K : p.N, p.ldb, p.b_batch, p.b_batch2, p.batch, p.batch2);
  auto pad = [](size_t n) { return (n + 127) / 128 * 128; };
  return (pad(ea) + pad(eb)) * 2 * sizeof(__nv_bfloat16);
}

int launch_split_bf16(Problem const &p, Epilogue const &e, cudaStream_t stream, void *workspace,
                      size_t workspace_bytes) {
  if (p.M <= 0 || p.N <= 0 || p.batch <= 0) return FFK_OK;
  FFK_REQUIRE(p.in_dtype == FFK_FLOAT && p.out_dtype == FFK_FLOAT, "split-bf16 gemm: fp32 operands only");
  FFK_REQUIRE(!(e.accumulate && (e.act != ACT_NONE || e.bias)), "split-bf16 gemm: accumulate with bias/activation");
  size_t need = split_workspace_bytes(p);
  if (!workspace || workspace_bytes < need || !aligned16(workspace)) {
    set_error("split-bf16 gemm: needs %zu bytes of 16-byte aligned scratch, got %zu", need, workspace_bytes);
    return FFK_ERR_UNSUPPORTED;
  }
  size_t ea = operand_elems(p.a_mn_major ? p.K : p.M, p.lda, p.a_batch, p.a_batch2, p.batch, p.batch2);
  size_t eb = operand_elems(p.b_mn_major ? p.K : p.N, p.ldb, p.b_batch, p.b_batch2, p.batch, p.batch2);
  auto pad = [](size_t n) { return (n + 127) / 128 * 128; };
  __nv_bfloat16 *a_hi = static_cast<__nv_bfloat16 *>(workspace), *a_lo = a_hi + pad(ea);
  __nv_bfloat16 *b_hi = a_lo + pad(ea), *b_lo = b_hi + pad(eb);
  unsigned ga = stream_grid(ea / 4 + 1, 256, 8), gb = stream_grid(eb / 4 + 1, 256, 8);
  split_bf16_kernel<<<ga, 256, 0, stream>>>(static_cast<float const *>(p.a), a_hi, a_lo, ea);
  split_bf16_kernel<<<gb, 256, 0, stream>>>(static_cast<float const *>(p.b), b_hi, b_lo, eb);
  FFK_LAUNCH_CHECK();
  Problem q = p;
  q.in_dtype = FFK_BFLOAT16;
  q.scratch_busy = true;   // the registered scratch may be this workspace (and its lock is held)
  if (!tc_eligible(q, e)) {
    set_error("split-bf16 gemm: operands do not satisfy TMA alignment");
    return FFK_ERR_UNSUPPORTED;
  }
  // pass 1: D (=|+=) A_hi B_hi ; pass 2: D += A_hi B_lo ; pass 3: D = epilogue(D + A_lo B_hi)
  Epilogue plain = e;
  plain.bias = nullptr;
  plain.act = ACT_NONE;
  plain.lambda = 0.f;
  plain.r = nullptr;
  int rc;
  q.a = a_hi; q.b = b_hi;
  if ((rc = launch_tc(q, plain, stream))) return rc;
  plain.accumulate = 1;
  q.b = b_lo;
  if ((rc = launch_tc(q, plain, stream))) return rc;
  Epilogue last = e;
  last.accumulate = 0;
  last.add_d_before_act = 1;
  q.a = a_lo; q.b = b_hi;
  return launch_tc(q, last, stream);
}

int launch_split_bf16_registered(Problem const &p, Epilogue const &e, cudaStream_t stream) {
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  Scratch *sc = scratch_slot();
  if (!sc || !sc->ptr) {
    set_error("split-bf16 gemm: no scratch registered for the current device (ffk_set_gemm_workspace)");
    return FFK_ERR_UNSUPPORTED;
  }
  int rc = scratch_acquire(*sc, stream);
  if (rc) return rc;
  rc = launch_split_bf16(p, e, stream, sc->ptr, sc->bytes);
  scratch_release(*sc, stream);
  return rc;
}

namespace {

// 16-bit operand whose base, row pitch or batch strides break TMA's 16-byte rule (a vocabulary of
// 30522 columns, say): copy it once into registered scratch with the pitch rounded up to 8
// elements so that the product still runs on the tensor cores; the copy costs two passes over the
// operand, the CUDA-core fallback it replaces runs at 2 % of the tensor-core rate.
__global__ void __launch_bounds__(256)
    repack_rows_kernel(uint16_t const *__restrict__ src, uint16_t *__restrict__ dst, long long rows,
                       long long cols, long long ld_src, long long ld_dst, long long batch_outer,
                       long long batch2, long long stride_outer, long long stride_inner) {
  // one CTA per row at a time; a thread moves 8 columns: element loads (the source row is only
  // 2-byte aligned), one 16-byte store (the padded pitch is a multiple of 8 elements)
  long long const total_rows = rows * batch_outer * batch2;
  for (long long row = blockIdx.x; row < total_rows; row += gridDim.x) {
    long long const b = row / rows, r = row - b * rows;
    long long const bo = b / batch2, bi = b - bo * batch2;
    uint16_t const *s = src + bo * stride_outer + bi * stride_inner + r * ld_src;
    uint16_t *d = dst + row * ld_dst;
    for (long long c0 = (long long)threadIdx.x * 8; c0 < ld_dst; c0 += 256 * 8) {
      alignas(16) uint16_t v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = (c0 + j < cols) ? s[c0 + j] : uint16_t(0);
      *reinterpret_cast<uint4 *>(d + c0) = *reinterpret_cast<uint4 *>(v);
    }
  }
}

bool operand_ok(void const *ptr, long long ld, long long bs, long long bs2, int batch, int batch2) {
  if (!aligned16(ptr) || (ld * 2) % 16) return false;
  if (batch > 1 && (bs * 2) % 16) return false;
  if (batch2 > 1 && (bs2 * 2) % 16) return false;
  return true;
}

// Returns FFK_OK and fills `q` when the operands that needed it were repacked into the scratch.
int repack_for_tensor_cores(Problem const &p, Problem &q, cudaStream_t stream, Scratch const &sc) {
  int const b2 = p.batch2 > 1 ? p.batch2 : 1, bo = p.batch / b2;
  if (p.batch % b2 != 0) return FFK_ERR_UNSUPPORTED;
  char *cursor = static_cast<char *>(sc.ptr);
  size_t left = sc.bytes;
  q = p;
  q.scratch_busy = true;
  auto fix = [&](void const *&ptr, long long &ld, long long &bs, long long &bs2, long long rows,
                 long long cols) -> int {
    if (operand_ok(ptr, ld, bs, bs2, p.batch, b2)) return FFK_OK;
    long long const ld_new = (cols + 7) / 8 * 8;
    size_t const bytes = (size_t(rows) * ld_new * p.batch * 2 + 255) / 256 * 256;
    if (bytes > left) return FFK_ERR_UNSUPPORTED;
    unsigned grid = (unsigned)std::min<long long>(rows * p.batch, (long long)device_info().sm_count * 16);
    repack_rows_kernel<<<grid, 256, 0, stream>>>(
        static_cast<uint16_t const *>(ptr), reinterpret_cast<uint16_t *>(cursor), rows, cols, ld,
        ld_new, bo, b2, bs, bs2);
    ptr = cursor;
    ld = ld_new;
    bs2 = rows * ld_new;
    bs = bs2 * b2;
    cursor += bytes;
    left -= bytes;
    return FFK_OK;
  };
  int rc = fix(q.a, q.lda, q.a_batch, q.a_batch2, p.a_mn_major ? p.K : p.M, p.a_mn_major ? p.M : p.K);
  if (rc) return rc;
  rc = fix(q.b, q.ldb, q.b_batch, q.b_batch2, p.b_mn_major ? p.K : p.N, p.b_mn_major ? p.N : p.K);
  if (rc) return rc;
  if (cudaGetLastError() != cudaSuccess) return FFK_ERR_CUDA;
  return FFK_OK;
}

}  // namespace

int launch(Problem const &p, Epilogue const &e, cudaStream_t stream) {
  if (p.M <= 0 || p.N <= 0 || p.batch <= 0) return FFK_OK;
  bool sixteen = p.in_dtype == FFK_BFLOAT16 || p.in_dtype == FFK_HALF;
  if (sixteen && p.K > 0 && tc_eligible(p, e)) return launch_tc(p, e, stream);
  bool const big = p.K >= 64 && double(p.M) * p.N * p.K * p.batch >= double(1 << 27);
  if (big && (sixteen || p.in_dtype == FFK_FLOAT)) {
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    Scratch *sc = scratch_slot();
    if (sc && sc->ptr && aligned16(sc->ptr)) {
      // 16-bit operands that break TMA's pitch rule: repacked into the scratch
      if (sixteen && !p.scratch_busy && scratch_acquire(*sc, stream) == FFK_OK) {
        Problem q;
        if (repack_for_tensor_cores(p, q, stream, *sc) == FFK_OK && tc_eligible(q, e)) {
          int rc = launch_tc(q, e, stream);
          scratch_release(*sc, stream);
          return rc;
        }
        scratch_release(*sc, stream);  // a repack kernel may have been queued
      }
      // large fp32 products: tensor cores through the bf16 split, when the owner of the device
      // handle allows it (PerDeviceFFHandle::allowTensorOpMathConversion)
      if (p.in_dtype == FFK_FLOAT && p.out_dtype == FFK_FLOAT && sc->fp32_tensor_math &&
          e.alpha == 1.f && !(e.accumulate && (e.act != ACT_NONE || e.bias))) {
        Problem q = p;
        q.in_dtype = FFK_BFLOAT16;
        bool d_ok = aligned16(e.d) && (e.ldd * sizeof(float)) % 16 == 0;
        if (tc_eligible(q, e) && d_ok && split_workspace_bytes(p) <= sc->bytes &&
            scratch_acquire(*sc, stream) == FFK_OK) {
          int rc = launch_split_bf16(p, e, stream, sc->ptr, sc->bytes);
          scratch_release(*sc, stream);
          return rc;
        }
      }
    }
  }
  return launch_simt(p, e, stream);
}

}  // namespace gemm
}  // namespace ffk
