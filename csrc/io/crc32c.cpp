// b200ps N8 — CRC-32C (Castagnoli), the checksum of TFRecord/tfevents framing and of the
// tensor-bundle checkpoint format (model_export.py:53 Saver; example.py:154 SummaryWriter).
// Slice-by-8 table implementation; uses the SSE4.2 instruction when the CPU has it.
#include <cstddef>
#include <cstdint>
#if defined(__x86_64__)
#include <nmmintrin.h>
#include <cpuid.h>
#endif

namespace {
uint32_t g_table[8][256];
bool g_init = false;
bool g_hw = false;

void init_tables() {
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
    g_table[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; ++i)
    for (int t = 1; t < 8; ++t) g_table[t][i] = (g_table[t - 1][i] >> 8) ^ g_table[0][g_table[t - 1][i] & 0xff];
#if defined(__x86_64__)
  unsigned a, b, c, d;
  if (__get_cpuid(1, &a, &b, &c, &d)) g_hw = (c & (1u << 20)) != 0;
#endif
  g_init = true;
}

#if defined(__x86_64__)
__attribute__((target("sse4.2"))) uint32_t crc_hw(uint32_t crc, const uint8_t* p, size_t n) {
  uint64_t c = crc;
  while (n >= 8) {
    uint64_t v;
    __builtin_memcpy(&v, p, 8);
    c = _mm_crc32_u64(c, v);
    p += 8; n -= 8;
  }
  uint32_t c32 = (uint32_t)c;
  while (n--) c32 = _mm_crc32_u8(c32, *p++);
  return c32;
}
#endif

uint32_t crc_sw(uint32_t crc, const uint8_t* p, size_t n) {
  while (n >= 8) {
    uint32_t lo, hi;
    __builtin_memcpy(&lo, p, 4);
    __builtin_memcpy(&hi, p + 4, 4);
    lo ^= crc;
    crc = g_table[7][lo & 0xff] ^ g_table[6][(lo >> 8) & 0xff] ^ g_table[5][(lo >> 16) & 0xff] ^
          g_table[4][lo >> 24] ^ g_table[3][hi & 0xff] ^ g_table[2][(hi >> 8) & 0xff] ^
          g_table[1][(hi >> 16) & 0xff] ^ g_table[0][hi >> 24];
    p += 8; n -= 8;
  }
  while (n--) crc = (crc >> 8) ^ g_table[0][(crc ^ *p++) & 0xff];
  return crc;
}
}  // namespace

// crc = b200ps_crc32c_extend(0, data, n) is the plain CRC-32C of data.
extern "C" uint32_t b200ps_crc32c_extend(uint32_t init_crc, const void* data, uint64_t n) {
  if (!g_init) init_tables();
  uint32_t c = init_crc ^ 0xffffffffu;
  const uint8_t* p = static_cast<const uint8_t*>(data);
#if defined(__x86_64__)
  if (g_hw) return crc_hw(c, p, n) ^ 0xffffffffu;
#endif
  return crc_sw(c, p, n) ^ 0xffffffffu;
}
