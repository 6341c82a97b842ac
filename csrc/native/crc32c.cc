// CRC-32C (Castagnoli): SSE4.2 hardware path with a table fallback, plus the masking used by
// TFRecord framing and the tensor bundle (rotate right 15, add 0xa282ead8).
#include <nmmintrin.h>

#include "../../include/dtf_native.h"
#include "util.h"

namespace dtfn {

static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static uint32_t g_table[256];
static bool g_table_ready = false;
static void init_table() {
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82f63b78u : (c >> 1);
    g_table[i] = c;
  }
  g_table_ready = true;
}

static bool has_sse42() {
#if defined(__x86_64__)
  return __builtin_cpu_supports("sse4.2");
#else
  return false;
#endif
}

uint32_t crc32c_extend(uint32_t crc, const uint8_t* p, size_t n) {
  uint32_t c = crc ^ 0xffffffffu;
  static const bool hw = has_sse42();
  if (hw) {
    uint64_t c64 = c;
    while (n >= 8) {
      uint64_t v;
      memcpy(&v, p, 8);
      c64 = _mm_crc32_u64(c64, v);
      p += 8;
      n -= 8;
    }
    c = static_cast<uint32_t>(c64);
    while (n--) c = _mm_crc32_u8(c, *p++);
  } else {
    if (!g_table_ready) init_table();
    while (n--) c = g_table[(c ^ *p++) & 0xff] ^ (c >> 8);
  }
  return c ^ 0xffffffffu;
}

}  // namespace dtfn

extern "C" {
const char* dtfn_last_error(void) { return dtfn::g_err; }
uint32_t dtfn_crc32c(const void* data, size_t n) {
  return dtfn::crc32c_extend(0, static_cast<const uint8_t*>(data), n);
}
uint32_t dtfn_crc32c_extend(uint32_t crc, const void* data, size_t n) {
  return dtfn::crc32c_extend(crc, static_cast<const uint8_t*>(data), n);
}
uint32_t dtfn_crc32c_mask(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }
uint32_t dtfn_crc32c_unmask(uint32_t masked) {
  uint32_t rot = masked - 0xa282ead8u;
  return (rot >> 17) | (rot << 15);
}
}
