// Shared helpers for libdtf_native: error string, varint / fixed coding, tiny protobuf writer+reader.
#pragma once

#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

namespace dtfn {

void set_error(const char* fmt, ...);

inline void put_fixed32(std::string* s, uint32_t v) {
  char b[4];
  memcpy(b, &v, 4);  // little-endian hosts only (x86-64 / aarch64)
  s->append(b, 4);
}
inline void put_fixed64(std::string* s, uint64_t v) {
  char b[8];
  memcpy(b, &v, 8);
  s->append(b, 8);
}
inline uint32_t get_fixed32(const char* p) {
  uint32_t v;
  memcpy(&v, p, 4);
  return v;
}
inline uint64_t get_fixed64(const char* p) {
  uint64_t v;
  memcpy(&v, p, 8);
  return v;
}
inline void put_varint64(std::string* s, uint64_t v) {
  while (v >= 0x80) {
    s->push_back(static_cast<char>(v | 0x80));
    v >>= 7;
  }
  s->push_back(static_cast<char>(v));
}
inline void put_varint32(std::string* s, uint32_t v) { put_varint64(s, v); }
// returns pointer past the varint or nullptr on malformed / truncated input
inline const char* get_varint64(const char* p, const char* limit, uint64_t* v) {
  uint64_t r = 0;
  for (int shift = 0; shift <= 63 && p < limit; shift += 7) {
    uint64_t b = static_cast<unsigned char>(*p++);
    r |= (b & 0x7f) << shift;
    if (!(b & 0x80)) {
      *v = r;
      return p;
    }
  }
  return nullptr;
}
inline const char* get_varint32(const char* p, const char* limit, uint32_t* v) {
  uint64_t t;
  p = get_varint64(p, limit, &t);
  if (p) *v = static_cast<uint32_t>(t);
  return p;
}

// ---- minimal protobuf wire helpers (proto3: zero-valued scalars are omitted by callers) ----
inline void pb_varint_field(std::string* s, int field, uint64_t v) {
  put_varint32(s, static_cast<uint32_t>(field << 3 | 0));
  put_varint64(s, v);
}
inline void pb_fixed32_field(std::string* s, int field, uint32_t v) {
  put_varint32(s, static_cast<uint32_t>(field << 3 | 5));
  put_fixed32(s, v);
}
inline void pb_bytes_field(std::string* s, int field, const std::string& v) {
  put_varint32(s, static_cast<uint32_t>(field << 3 | 2));
  put_varint64(s, v.size());
  s->append(v);
}

struct PbField {
  int number;
  int wire;           // 0 varint, 1 fixed64, 2 bytes, 5 fixed32
  uint64_t value;     // varint / fixed
  const char* data;   // bytes
  size_t size;
};
// Iterates top-level fields of a serialized message. Returns false on malformed input.
inline bool pb_next(const char** p, const char* limit, PbField* f) {
  uint64_t key;
  const char* q = get_varint64(*p, limit, &key);
  if (!q) return false;
  f->number = static_cast<int>(key >> 3);
  f->wire = static_cast<int>(key & 7);
  f->data = nullptr;
  f->size = 0;
  f->value = 0;
  switch (f->wire) {
    case 0:
      q = get_varint64(q, limit, &f->value);
      if (!q) return false;
      break;
    case 1:
      if (limit - q < 8) return false;
      f->value = get_fixed64(q);
      q += 8;
      break;
    case 2: {
      uint64_t n;
      q = get_varint64(q, limit, &n);
      if (!q || static_cast<uint64_t>(limit - q) < n) return false;
      f->data = q;
      f->size = n;
      q += n;
      break;
    }
    case 5:
      if (limit - q < 4) return false;
      f->value = get_fixed32(q);
      q += 4;
      break;
    default:
      return false;
  }
  *p = q;
  return true;
}

}  // namespace dtfn
