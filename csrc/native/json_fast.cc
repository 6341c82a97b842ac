// Fast path for the serving endpoint's JSON: the predict request is an object whose values are
// rectangular nested arrays of numbers (reference README.md:40:
// {"keys": [[11.0], [2.0]], "features": [[1], [2]]}), and the response has the same form.  Python's json
// module spends ~15 ms on a 4096 x 39 request; these two routines parse / format such arrays directly
// into / from flat buffers.  Anything that is not a plain rectangular numeric array makes the parser
// return a negative code and the caller falls back to the generic JSON path.
#include <math.h>
#include <stdlib.h>

#include "../../include/dtf_native.h"
#include "util.h"

namespace {

struct Cur {
  const char* p;
  const char* end;
  void ws() {
    while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
  }
};

// skips a JSON string starting at the opening quote; false on malformed input
bool skip_string(Cur& c) {
  if (c.p >= c.end || *c.p != '"') return false;
  ++c.p;
  while (c.p < c.end) {
    if (*c.p == '\\') {
      c.p += 2;
      continue;
    }
    if (*c.p == '"') {
      ++c.p;
      return true;
    }
    ++c.p;
  }
  return false;
}

// skips any JSON value (used for keys we are not interested in)
bool skip_value(Cur& c) {
  c.ws();
  if (c.p >= c.end) return false;
  if (*c.p == '"') return skip_string(c);
  if (*c.p == '{' || *c.p == '[') {
    int depth = 0;
    while (c.p < c.end) {
      if (*c.p == '"') {
        if (!skip_string(c)) return false;
        continue;
      }
      if (*c.p == '{' || *c.p == '[') ++depth;
      if (*c.p == '}' || *c.p == ']') {
        --depth;
        if (depth == 0) {
          ++c.p;
          return true;
        }
      }
      ++c.p;
    }
    return false;
  }
  while (c.p < c.end && *c.p != ',' && *c.p != '}' && *c.p != ']') ++c.p;   // number / true / false / null
  return true;
}

// number -> double; JSON grammar only (no hex, no inf/nan)
bool parse_number(Cur& c, double* out) {
  const char* s = c.p;
  bool neg = false;
  if (s < c.end && *s == '-') {
    neg = true;
    ++s;
  }
  if (s >= c.end || *s < '0' || *s > '9') return false;
  // fast path: up to 15 significant digits and a small decimal exponent are exact in double arithmetic
  unsigned long long mant = 0;
  int digits = 0, frac = 0;
  const char* q = s;
  while (q < c.end && *q >= '0' && *q <= '9') {
    if (digits < 18) {
      mant = mant * 10 + static_cast<unsigned>(*q - '0');
      ++digits;
    } else {
      digits = 99;
    }
    ++q;
  }
  if (q < c.end && *q == '.') {
    ++q;
    if (q >= c.end || *q < '0' || *q > '9') return false;
    while (q < c.end && *q >= '0' && *q <= '9') {
      if (digits < 18) {
        mant = mant * 10 + static_cast<unsigned>(*q - '0');
        ++digits;
        ++frac;
      } else {
        digits = 99;
      }
      ++q;
    }
  }
  const bool has_exp = q < c.end && (*q == 'e' || *q == 'E');
  if (!has_exp && digits <= 15 && frac <= 22) {
    static const double p10[] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    const double v = static_cast<double>(mant) / p10[frac];
    *out = neg ? -v : v;
    c.p = q;
    return true;
  }
  // general case: strtod on a bounded copy (the buffer is not NUL-terminated)
  if (has_exp) {
    ++q;
    if (q < c.end && (*q == '+' || *q == '-')) ++q;
    if (q >= c.end || *q < '0' || *q > '9') return false;
    while (q < c.end && *q >= '0' && *q <= '9') ++q;
  }
  char tmp[64];
  const size_t n = static_cast<size_t>(q - c.p);
  if (n >= sizeof(tmp)) return false;
  memcpy(tmp, c.p, n);
  tmp[n] = 0;
  char* endp = nullptr;
  const double v = strtod(tmp, &endp);
  if (endp != tmp + n) return false;
  *out = v;
  c.p = q;
  return true;
}

// nested array of numbers at c.p ('['): values go to out in row-major order; every level must be
// rectangular and all numbers must sit at the same depth
struct ArrayState {
  double* out;
  size_t cap;
  size_t count = 0;
  long long dims[4] = {0, 0, 0, 0};
  bool known[4] = {false, false, false, false};
  int leaf_depth = -1;   // depth of the arrays that hold the numbers
  bool overflow = false;
};

bool parse_array(Cur& c, int depth, ArrayState& st) {
  if (depth >= 4) return false;
  c.ws();
  if (c.p >= c.end || *c.p != '[') return false;
  ++c.p;
  long long n = 0;
  c.ws();
  if (c.p < c.end && *c.p == ']') {
    ++c.p;
    if (st.leaf_depth < 0) st.leaf_depth = depth;   // "[]" or "[[], []]": empty leaves
  } else {
    for (;;) {
      c.ws();
      if (c.p >= c.end) return false;
      if (*c.p == '[') {
        if (st.leaf_depth >= 0 && st.leaf_depth <= depth) return false;   // array next to numbers
        if (!parse_array(c, depth + 1, st)) return false;
      } else {
        if (st.leaf_depth >= 0 && st.leaf_depth != depth) return false;   // number at another depth
        st.leaf_depth = depth;
        double v;
        if (!parse_number(c, &v)) return false;
        if (st.count >= st.cap) {
          st.overflow = true;
          return false;
        }
        st.out[st.count++] = v;
      }
      ++n;
      c.ws();
      if (c.p < c.end && *c.p == ',') {
        ++c.p;
        continue;
      }
      if (c.p < c.end && *c.p == ']') {
        ++c.p;
        break;
      }
      return false;
    }
  }
  if (!st.known[depth]) {
    st.known[depth] = true;
    st.dims[depth] = n;
  } else if (st.dims[depth] != n) {
    return false;   // ragged
  }
  return true;
}

}  // namespace

extern "C" {

// Finds "key" at the top level of the JSON object in body (or inside a top-level "data" object, the
// alternative request envelope) and parses its value as a rectangular nested array of numbers.
// Returns the number of values written to out (row-major) and fills dims[0..*ndims); returns -1 when the
// key is absent, -2 when the value is not such an array (caller falls back), -3 when out is too small.
long long dtfn_json_numeric_array(const char* body, size_t n, const char* key, double* out, size_t cap,
                                  long long* dims, int* ndims) {
  const size_t klen = strlen(key);
  Cur c{body, body + n};
  c.ws();
  if (c.p >= c.end || *c.p != '{') return -2;
  ++c.p;
  for (int level = 0; level < 2; ++level) {
    bool descended = false;
    for (;;) {
      c.ws();
      if (c.p < c.end && *c.p == '}') return -1;
      const char* ks = c.p;
      if (!skip_string(c)) return -2;
      const char* ke = c.p;
      c.ws();
      if (c.p >= c.end || *c.p != ':') return -2;
      ++c.p;
      const size_t len = static_cast<size_t>(ke - ks) - 2;
      if (len == klen && memcmp(ks + 1, key, klen) == 0) {
        ArrayState st;
        st.out = out;
        st.cap = cap;
        c.ws();
        if (c.p >= c.end || *c.p != '[') return -2;
        if (!parse_array(c, 0, st)) return st.overflow ? -3 : -2;
        const int nd = st.leaf_depth + 1;
        for (int i = 0; i < nd; ++i) dims[i] = st.dims[i];
        *ndims = nd;
        return static_cast<long long>(st.count);
      }
      if (level == 0 && len == 4 && memcmp(ks + 1, "data", 4) == 0) {
        c.ws();
        if (c.p < c.end && *c.p == '{') {
          ++c.p;
          descended = true;
          break;   // look for the key inside the envelope
        }
      }
      if (!skip_value(c)) return -2;
      c.ws();
      if (c.p < c.end && *c.p == ',') {
        ++c.p;
        continue;
      }
      if (c.p < c.end && *c.p == '}') return -1;
      return -2;
    }
    if (!descended) break;
  }
  return -1;
}

// Formats a row-major array as nested JSON arrays into out (no NUL terminator); returns the number of
// bytes written, or the negative required capacity estimate when cap is too small.  kind: 0 = float32
// values (shortest form that round-trips float32, always with a '.' or exponent), 1 = int64.
long long dtfn_json_format_array(const void* data, int kind, const long long* dims, int ndims, char* out, size_t cap) {
  if (ndims < 1 || ndims > 4) return 0;
  long long total = 1;
  for (int i = 0; i < ndims; ++i) total *= dims[i];
  const size_t need = static_cast<size_t>(total) * 26 + static_cast<size_t>(total) * 2 * ndims + 16;
  if (cap < need) return -static_cast<long long>(need);
  char* w = out;
  long long idx[4] = {0, 0, 0, 0};
  long long flat = 0;
  for (int i = 0; i < ndims; ++i) *w++ = '[';
  if (total == 0) {
    for (int i = 0; i < ndims; ++i) *w++ = ']';
    return w - out;
  }
  for (;;) {
    if (kind == 1) {
      w += snprintf(w, 24, "%lld", static_cast<const long long*>(data)[flat]);
    } else {
      const float f = static_cast<const float*>(data)[flat];
      if (isnan(f)) {
        memcpy(w, "NaN", 3);
        w += 3;
      } else if (isinf(f)) {
        const char* s = f > 0 ? "Infinity" : "-Infinity";
        const size_t l = strlen(s);
        memcpy(w, s, l);
        w += l;
      } else {
        int len = 0;
        for (int prec = 6; prec <= 9; ++prec) {   // shortest %g that reads back as the same float32
          len = snprintf(w, 24, "%.*g", prec, static_cast<double>(f));
          if (strtof(w, nullptr) == f) break;
        }
        bool plain = true;
        for (int i = 0; i < len; ++i)
          if (w[i] == '.' || w[i] == 'e' || w[i] == 'E') plain = false;
        w += len;
        if (plain) {
          *w++ = '.';
          *w++ = '0';
        }
      }
    }
    ++flat;
    // advance the index vector, closing / opening brackets as levels roll over
    int level = ndims - 1;
    while (level >= 0 && ++idx[level] == dims[level]) {
      idx[level] = 0;
      *w++ = ']';
      --level;
    }
    if (level < 0) break;
    *w++ = ',';
    *w++ = ' ';
    for (int i = level + 1; i < ndims; ++i) *w++ = '[';
  }
  return w - out;
}

}  // extern "C"
