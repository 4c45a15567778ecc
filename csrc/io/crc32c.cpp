// Host-side CRC-32C (Castagnoli) for the TFRecord / tensor-bundle / event-file codecs.
// Hardware path: SSE4.2 crc32 instruction, 8 bytes per step; portable path: byte table.
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

__attribute__((target("sse4.2"))) uint32_t crc_hw(const uint8_t* p, size_t n, uint32_t c) {
  uint64_t c64 = c;
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
  static const bool hw = has_sse42();
  c = hw ? crc_hw(p, n, c) : crc_sw(p, n, c);
#else
  c = crc_sw(p, n, c);
#endif
  return c ^ 0xFFFFFFFFu;
}
