// Counter-based Philox4x32-10: the dropout mask is a pure function of (seed, step, element index), so the
// backward pass regenerates it instead of storing it. The Python oracle (ops/philox.py) mirrors this file.
#pragma once
#include <stdint.h>

namespace dtm {

struct Philox4 {
  uint32_t x, y, z, w;
};

__host__ __device__ inline Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                 uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = static_cast<uint64_t>(M0) * c0;
    const uint64_t p1 = static_cast<uint64_t>(M1) * c2;
    const uint32_t n0 = static_cast<uint32_t>(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n1 = static_cast<uint32_t>(p1);
    const uint32_t n2 = static_cast<uint32_t>(p0 >> 32) ^ c3 ^ k1;
    const uint32_t n3 = static_cast<uint32_t>(p0);
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  return Philox4{c0, c1, c2, c3};
}

// Dropout keep-decision for 4 consecutive elements starting at element index `e4*4`.
// keep iff (r >> 8) < keep_thresh where keep_thresh = floor(keep_prob * 2^24).
__host__ __device__ inline Philox4 dropout_bits(uint64_t e4, uint64_t step, uint64_t seed) {
  return philox4x32_10(static_cast<uint32_t>(e4), static_cast<uint32_t>(e4 >> 32), static_cast<uint32_t>(step),
                       static_cast<uint32_t>(step >> 32), static_cast<uint32_t>(seed),
                       static_cast<uint32_t>(seed >> 32));
}

}  // namespace dtm
