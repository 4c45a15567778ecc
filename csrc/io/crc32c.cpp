// Host-side CRC-32C (Castagnoli) for the TFRecord / tensor-bundle / event-file codecs.
// Hardware path: SSE4.2 crc32 instruction, three interleaved 8-byte chains over 768-byte blocks; portable path: byte table.
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#if defined(__x86_64__)
#include <nmmintrin.h>
#include <cpuid.h>
#endif

namespace {

uint32_t g_table[256];
bool g_table_ready = false;

void init_table() {
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
    g_table[i] = c;
  }
  g_table_ready = true;
}

uint32_t crc_sw(const uint8_t* p, size_t n, uint32_t c) {
  if (!g_table_ready) init_table();
  for (size_t i = 0; i < n; ++i) c = g_table[(c ^ p[i]) & 0xFF] ^ (c >> 8);
  return c;
}

#if defined(__x86_64__)
bool has_sse42() {
  unsigned a, b, c, d;
  if (!__get_cpuid(1, &a, &b, &c, &d)) return false;
  return (c & bit_SSE4_2) != 0;
}

// Three independent crc32 chains hide the instruction's 3-cycle latency (one chain: 8 bytes per 3 cycles).  The lanes are
// LANE bytes each; lane results are joined with the linear map "advance the register through LANE zero bytes", tabulated per
// register byte:  f(s, A|B|C) = shift(shift(f(s, A)) ^ f(0, B)) ^ f(0, C).
constexpr size_t LANE = 256;
uint32_t g_shift[4][256];
bool g_shift_ready = false;

void init_shift() {
  if (!g_table_ready) init_table();
  for (int b = 0; b < 4; ++b) {
    for (uint32_t v = 0; v < 256; ++v) {
      uint32_t c = v << (8 * b);
      for (size_t i = 0; i < LANE; ++i) c = g_table[c & 0xFF] ^ (c >> 8);
      g_shift[b][v] = c;
    }
  }
  g_shift_ready = true;
}

inline uint32_t shift_lane(uint32_t c) {
  return g_shift[0][c & 0xFF] ^ g_shift[1][(c >> 8) & 0xFF] ^ g_shift[2][(c >> 16) & 0xFF] ^ g_shift[3][c >> 24];
}

__attribute__((target("sse4.2"))) uint32_t crc_hw(const uint8_t* p, size_t n, uint32_t c) {
  uint64_t c64 = c;
  while (n >= 3 * LANE) {
    uint64_t a = c64, b = 0, d = 0;
    for (size_t i = 0; i < LANE; i += 8) {
      uint64_t va, vb, vd;
      memcpy(&va, p + i, 8);
      memcpy(&vb, p + LANE + i, 8);
      memcpy(&vd, p + 2 * LANE + i, 8);
      a = _mm_crc32_u64(a, va);
      b = _mm_crc32_u64(b, vb);
      d = _mm_crc32_u64(d, vd);
    }
    uint32_t r = shift_lane(static_cast<uint32_t>(a)) ^ static_cast<uint32_t>(b);
    r = shift_lane(r) ^ static_cast<uint32_t>(d);
    c64 = r;
    p += 3 * LANE;
    n -= 3 * LANE;
  }
  while (n >= 8) {
    uint64_t v;
    memcpy(&v, p, 8);
    c64 = _mm_crc32_u64(c64, v);
    p += 8;
    n -= 8;
  }
  uint32_t c32 = static_cast<uint32_t>(c64);
  while (n--) c32 = _mm_crc32_u8(c32, *p++);
  return c32;
}
#endif

}  // namespace

extern "C" uint32_t dtm_crc32c(const char* data, size_t n, uint32_t crc) {
  const uint8_t* p = reinterpret_cast<const uint8_t*>(data);
  uint32_t c = crc ^ 0xFFFFFFFFu;
#if defined(__x86_64__)
  static const bool hw = [] { init_shift(); return has_sse42(); }();
  c = hw ? crc_hw(p, n, c) : crc_sw(p, n, c);
#else
  c = crc_sw(p, n, c);
#endif
  return c ^ 0xFFFFFFFFu;
}
