// Dropout randomness, shared by the stand-alone dropout kernels (elementwise.cu) and the fused
// GEMM epilogue (gemm_sm100.cu): both must draw the SAME keep bit for the same element.
#pragma once
#include <cstdint>

namespace ffk {

// Philox4x32-10 (Salmon et al.), counter = (element_block, offset), key = seed.
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t (&k)[2]) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
  uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
  uint32_t n0 = hi1 ^ c[1] ^ k[0], n1 = lo1, n2 = hi0 ^ c[3] ^ k[1], n3 = lo0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
  k[0] += 0x9E3779B9u;
  k[1] += 0xBB67AE85u;
}
__device__ __forceinline__ void philox4x32_10(uint64_t ctr, uint64_t offset, uint64_t seed,
                                              uint32_t (&out)[4]) {
  uint32_t c[4] = {uint32_t(ctr), uint32_t(ctr >> 32), uint32_t(offset), uint32_t(offset >> 32)};
  uint32_t k[2] = {uint32_t(seed), uint32_t(seed >> 32)};
#pragma unroll
  for (int i = 0; i < 10; ++i) philox_round(c, k);
#pragma unroll
  for (int i = 0; i < 4; ++i) out[i] = c[i];
}

// Randomness is defined per element: element e takes 16-bit field (e % 8) of
// Philox(counter = e / 8, offset, key = seed) — field f is bits [16*(f&1), +16) of word f/2 —
// and is kept iff that field >= rate * 2^16. (One Philox call serves 8 elements: at 6.5 TB/s
// the generator, not HBM, bounds this kernel, and 16 bits resolve the rate to 1.5e-5.)
__device__ __forceinline__ uint32_t philox_keep8(uint64_t block, uint64_t offset, uint64_t seed,
                                                 uint32_t thresh16) {
  uint32_t rnd[4];
  philox4x32_10(block, offset, seed, rnd);
  uint32_t bits = 0;
#pragma unroll
  for (int f = 0; f < 8; ++f) {
    uint32_t r16 = (rnd[f >> 1] >> ((f & 1) * 16)) & 0xFFFFu;
    if (r16 >= thresh16) bits |= 1u << f;
  }
  return bits;
}


// threshold on a 16-bit field for a drop probability `rate` (keep iff field >= threshold)
inline uint32_t dropout_thresh16(float rate) {
  return rate >= 1.f ? 0x10000u : uint32_t(double(rate) * 65536.0 + 0.5);
}

}  // namespace ffk
