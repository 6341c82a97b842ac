// TensorFlow checkpoint-V2 "tensor bundle":  <prefix>.index (a sorted table whose "" key holds a
// BundleHeaderProto and every other key a BundleEntryProto) plus <prefix>.data-SSSSS-of-NNNNN files
// holding the raw little-endian tensor bytes back to back.  Native stand-in for SaveV2 /
// RestoreV2 / MergeV2Checkpoints behind tf.train.Saver (reference trainer/task.py:143,
// trainer/task_supervisor.py:145).  Also hosts the TFRecord writer used for event files.
//
//   BundleHeaderProto { int32 num_shards = 1; Endianness endianness = 2; VersionDef version = 3; }
//   BundleEntryProto  { DataType dtype = 1; TensorShapeProto shape = 2; int32 shard_id = 3;
//                       int64 offset = 4; int64 size = 5; fixed32 crc32c = 6; slices = 7; }
//   TensorShapeProto  { repeated Dim dim = 2 { int64 size = 1; } }
#include <stdlib.h>
#include <unistd.h>

#include <algorithm>
#include <map>
#include <memory>

#include "../../include/dtf_native.h"
#include "util.h"

namespace dtfn {

uint32_t crc32c_extend(uint32_t crc, const uint8_t* p, size_t n);

struct Entry {
  std::string name;
  int dtype = 0;
  std::vector<long long> dims;
  int shard_id = 0;
  long long offset = 0, size = 0;
  uint32_t crc = 0;  // masked
  // Partitioned variable (BundleEntryProto.slices): the entry of the FULL tensor lists the slices that
  // exist; each slice's bytes live under its own key (encode_slice_key).  Per slice, per dimension:
  // (start, length), length -1 = the whole dimension.
  std::vector<std::vector<std::pair<long long, long long>>> slices;
};

// ---- OrderedCode (tensorflow/core/lib/strings/ordered_code.cc), the pieces slice keys need ----------
static void oc_write_num_increasing(std::string* s, uint64_t v) {
  // one length byte, then the value big-endian without leading zero bytes
  unsigned char buf[9];
  int n = 0;
  while (v > 0) {
    buf[8 - n] = static_cast<unsigned char>(v & 0xff);
    v >>= 8;
    ++n;
  }
  s->push_back(static_cast<char>(n));
  s->append(reinterpret_cast<const char*>(buf + 9 - n), n);
}
static void oc_write_string(std::string* s, const std::string& v) {
  for (unsigned char c : v) {
    if (c == 0x00) s->append("\x00\xff", 2);
    else if (c == 0xff) s->append("\xff\x00", 2);
    else s->push_back(static_cast<char>(c));
  }
  s->append("\x00\x01", 2);
}
static void oc_write_signed_num_increasing(std::string* s, long long val) {
  // length-1 encodings hold 6 value bits, every further byte adds 7; the header bits (a run of ones, the
  // run's length = encoded length) are XORed onto the sign-extended big-endian value
  const uint64_t x = val < 0 ? ~static_cast<uint64_t>(val) : static_cast<uint64_t>(val);
  if (x < 64) {
    s->push_back(static_cast<char>(0x80 ^ static_cast<unsigned char>(val)));
    return;
  }
  int bits = 0;
  for (uint64_t t = x; t; t >>= 1) ++bits;             // significant bits of x
  const int len = (bits + 1 + 6) / 7;                  // smallest len with 7 * len - 1 >= bits
  unsigned char buf[10];
  const unsigned char sign = val < 0 ? 0xff : 0x00;
  buf[0] = buf[1] = sign;
  for (int i = 0; i < 8; ++i) buf[2 + i] = static_cast<unsigned char>(static_cast<uint64_t>(val) >> (56 - 8 * i));
  static const unsigned char hdr[11][2] = {{0, 0},    {0x80, 0}, {0xc0, 0}, {0xe0, 0},    {0xf0, 0},   {0xf8, 0},
                                            {0xfc, 0}, {0xfe, 0}, {0xff, 0}, {0xff, 0x80}, {0xff, 0xc0}};
  unsigned char* begin = buf + 10 - len;
  begin[0] ^= hdr[len][0];
  begin[1] ^= hdr[len][1];
  s->append(reinterpret_cast<const char*>(begin), len);
}
// EncodeTensorNameSlice (tensorflow/core/util/saved_tensor_slice_util.cc): the key under which one slice of a
// partitioned tensor is stored.  Starts with a zero byte, so slice keys sort before every tensor name.
static std::string encode_slice_key(const std::string& name, const std::vector<std::pair<long long, long long>>& ext) {
  std::string k;
  oc_write_num_increasing(&k, 0);
  oc_write_string(&k, name);
  oc_write_num_increasing(&k, ext.size());
  for (const auto& e : ext) {
    oc_write_signed_num_increasing(&k, e.first);
    oc_write_signed_num_increasing(&k, e.second);
  }
  return k;
}
static int dtype_size(int dt) {
  switch (dt) {
    case 1: case 3: return 4;                 // float, int32
    case 2: case 9: return 8;                 // double, int64
    case 4: case 6: case 10: return 1;        // uint8, int8, bool
    case 5: case 14: case 19: return 2;       // int16, bfloat16, half
    default: return 0;
  }
}

static std::string data_path(const std::string& prefix, int shard, int num) {
  char buf[64];
  snprintf(buf, sizeof(buf), ".data-%05d-of-%05d", shard, num);
  return prefix + buf;
}

static std::string encode_header(int num_shards) {
  std::string s;
  pb_varint_field(&s, 1, static_cast<uint64_t>(num_shards));
  // endianness = LITTLE (0) is the proto3 default and is omitted
  std::string version;
  pb_varint_field(&version, 1, 1);  // VersionDef.producer = 1 (kTensorBundleVersion)
  pb_bytes_field(&s, 3, version);
  return s;
}

static std::string encode_entry(const Entry& e) {
  std::string s;
  if (e.dtype != 0) pb_varint_field(&s, 1, static_cast<uint64_t>(e.dtype));
  std::string shape;
  for (long long d : e.dims) {
    std::string dim;
    if (d != 0) pb_varint_field(&dim, 1, static_cast<uint64_t>(d));
    pb_bytes_field(&shape, 2, dim);
  }
  pb_bytes_field(&s, 2, shape);
  if (e.shard_id != 0) pb_varint_field(&s, 3, static_cast<uint64_t>(e.shard_id));
  if (e.offset != 0) pb_varint_field(&s, 4, static_cast<uint64_t>(e.offset));
  if (e.size != 0) pb_varint_field(&s, 5, static_cast<uint64_t>(e.size));
  if (e.slices.empty() || e.size != 0) pb_fixed32_field(&s, 6, e.crc);
  for (const auto& sl : e.slices) {
    // TensorSliceProto { repeated Extent extent = 1; }  Extent { int64 start = 1; oneof { int64 length = 2; } }
    std::string slice;
    for (const auto& ex : sl) {
      std::string extent;
      if (ex.first != 0) pb_varint_field(&extent, 1, static_cast<uint64_t>(ex.first));
      if (ex.second >= 0) pb_varint_field(&extent, 2, static_cast<uint64_t>(ex.second));
      pb_bytes_field(&slice, 1, extent);
    }
    pb_bytes_field(&s, 7, slice);
  }
  return s;
}

static bool decode_entry(const char* data, size_t n, Entry* e) {
  const char* p = data;
  const char* limit = data + n;
  PbField f;
  while (p < limit) {
    if (!pb_next(&p, limit, &f)) return false;
    switch (f.number) {
      case 1: e->dtype = static_cast<int>(f.value); break;
      case 2: {
        const char* q = f.data;
        const char* ql = f.data + f.size;
        PbField g;
        while (q < ql) {
          if (!pb_next(&q, ql, &g)) return false;
          if (g.number == 2 && g.wire == 2) {
            long long dim = 0;
            const char* r = g.data;
            const char* rl = g.data + g.size;
            PbField h;
            while (r < rl) {
              if (!pb_next(&r, rl, &h)) return false;
              if (h.number == 1) dim = static_cast<long long>(h.value);
            }
            e->dims.push_back(dim);
          }
        }
        break;
      }
      case 3: e->shard_id = static_cast<int>(f.value); break;
      case 4: e->offset = static_cast<long long>(f.value); break;
      case 5: e->size = static_cast<long long>(f.value); break;
      case 6: e->crc = static_cast<uint32_t>(f.value); break;
      case 7: {
        std::vector<std::pair<long long, long long>> sl;
        const char* q = f.data;
        const char* ql = f.data + f.size;
        PbField g;
        while (q < ql) {
          if (!pb_next(&q, ql, &g)) return false;
          if (g.number != 1 || g.wire != 2) continue;
          long long start = 0, length = -1;
          const char* r = g.data;
          const char* rl = g.data + g.size;
          PbField h;
          while (r < rl) {
            if (!pb_next(&r, rl, &h)) return false;
            if (h.number == 1) start = static_cast<long long>(h.value);
            if (h.number == 2) length = static_cast<long long>(h.value);
          }
          sl.emplace_back(start, length);
        }
        e->slices.push_back(std::move(sl));
        break;
      }
      default: break;
    }
  }
  return true;
}

static bool decode_header(const char* data, size_t n, int* num_shards, int* endianness) {
  const char* p = data;
  const char* limit = data + n;
  PbField f;
  *num_shards = 0;
  *endianness = 0;
  while (p < limit) {
    if (!pb_next(&p, limit, &f)) return false;
    if (f.number == 1) *num_shards = static_cast<int>(f.value);
    if (f.number == 2) *endianness = static_cast<int>(f.value);
  }
  return true;
}

static bool write_index(const std::string& prefix, int num_shards, const std::map<std::string, Entry>& entries) {
  const std::string path = prefix + ".index";
  const std::string tmp = path + ".tempstate";
  dtfn_table_builder* tb = dtfn_table_builder_new(tmp.c_str(), 0);
  if (!tb) return false;
  const std::string header = encode_header(num_shards);
  if (dtfn_table_builder_add(tb, "", 0, header.data(), header.size()) != 0) {
    dtfn_table_builder_abandon(tb);
    return false;
  }
  for (const auto& kv : entries) {
    const std::string enc = encode_entry(kv.second);
    if (dtfn_table_builder_add(tb, kv.first.data(), kv.first.size(), enc.data(), enc.size()) != 0) {
      dtfn_table_builder_abandon(tb);
      return false;
    }
  }
  if (dtfn_table_builder_finish(tb) != 0) return false;
  if (rename(tmp.c_str(), path.c_str()) != 0) {
    set_error("bundle: rename %s -> %s failed", tmp.c_str(), path.c_str());
    return false;
  }
  return true;
}

// "<var>/part_<r>" (and "<var>/part_<r>/<slot>") for r = 0..k-1 are the partitions of one logical variable
// "<var>" ("<var>/<slot>"), cut along dimension 0 -- the naming TF's PartitionedVariable uses for its
// parts.  Besides the per-part entries, the index gets what TF itself writes for a partitioned variable: an
// entry for the full tensor that lists the slices, and one slice-keyed entry per part that points at the SAME
// bytes as "<var>/part_<r>" (no data is duplicated).  A reader that asks for "<var>" gets the whole tensor.
static void add_partitioned_views(std::map<std::string, Entry>* entries) {
  struct Part { int r; const Entry* e; };
  std::map<std::pair<std::string, std::string>, std::vector<Part>> families;   // (base, "/slot" or "") -> parts
  for (const auto& kv : *entries) {
    const std::string& name = kv.first;
    const size_t at = name.find("/part_");
    if (at == std::string::npos || at == 0) continue;
    size_t q = at + 6;
    if (q >= name.size() || !isdigit(static_cast<unsigned char>(name[q]))) continue;
    int r = 0;
    while (q < name.size() && isdigit(static_cast<unsigned char>(name[q]))) r = r * 10 + (name[q++] - '0');
    if (q < name.size() && name[q] != '/') continue;
    families[{name.substr(0, at), name.substr(q)}].push_back(Part{r, &kv.second});
  }
  std::vector<Entry> add;
  for (auto& fam : families) {
    std::vector<Part>& parts = fam.second;
    std::sort(parts.begin(), parts.end(), [](const Part& a, const Part& b) { return a.r < b.r; });
    const Entry& first = *parts[0].e;
    bool ok = !first.dims.empty() && dtype_size(first.dtype) > 0;
    for (size_t i = 0; ok && i < parts.size(); ++i) {
      const Entry& e = *parts[i].e;
      ok = parts[i].r == static_cast<int>(i) && e.dtype == first.dtype && e.dims.size() == first.dims.size() &&
           std::equal(e.dims.begin() + 1, e.dims.end(), first.dims.begin() + 1) && e.slices.empty();
    }
    const std::string logical = fam.first.first + fam.first.second;
    if (!ok || entries->count(logical)) continue;
    Entry full;
    full.name = logical;
    full.dtype = first.dtype;
    full.dims = first.dims;
    if (parts.size() == 1) {
      // one partition: the logical tensor IS that part -- a plain second name for the same bytes
      full.shard_id = first.shard_id; full.offset = first.offset; full.size = first.size; full.crc = first.crc;
      add.push_back(std::move(full));
      continue;
    }
    long long row = 0;
    for (const Part& pt : parts) {
      std::vector<std::pair<long long, long long>> ext(first.dims.size(), {0, -1});
      ext[0] = {row, pt.e->dims[0]};
      Entry sl = *pt.e;                                   // same dtype / slice shape / shard / offset / size / crc
      sl.name = encode_slice_key(logical, ext);
      add.push_back(std::move(sl));
      full.slices.push_back(std::move(ext));
      row += pt.e->dims[0];
    }
    full.dims[0] = row;
    add.push_back(std::move(full));
  }
  for (Entry& e : add) (*entries)[e.name] = std::move(e);
}

}  // namespace dtfn

using namespace dtfn;

struct dtfn_bundle_writer {
  std::string prefix, tmp_data;
  FILE* f = nullptr;
  long long offset = 0;
  std::map<std::string, Entry> entries;
};

extern "C" {

dtfn_bundle_writer* dtfn_bundle_writer_new(const char* prefix) {
  auto w = std::make_unique<dtfn_bundle_writer>();
  w->prefix = prefix;
  w->tmp_data = data_path(w->prefix, 0, 1) + ".tempstate";
  w->f = fopen(w->tmp_data.c_str(), "wb");
  if (!w->f) {
    set_error("bundle: cannot open %s for writing", w->tmp_data.c_str());
    return nullptr;
  }
  return w.release();
}

int dtfn_bundle_writer_add(dtfn_bundle_writer* w, const char* name, int dtype, int ndims, const long long* dims,
                           const void* data, size_t nbytes) {
  if (name == nullptr || name[0] == '\0') {
    set_error("bundle: tensor name must not be empty");
    return -2;
  }
  if (w->entries.count(name)) {
    set_error("bundle: duplicate tensor name '%s'", name);
    return -2;
  }
  Entry e;
  e.name = name;
  e.dtype = dtype;
  e.dims.assign(dims, dims + ndims);
  e.shard_id = 0;
  e.offset = w->offset;
  e.size = static_cast<long long>(nbytes);
  const uint32_t c = crc32c_extend(0, static_cast<const uint8_t*>(data), nbytes);
  e.crc = ((c >> 15) | (c << 17)) + 0xa282ead8u;
  if (nbytes && fwrite(data, 1, nbytes, w->f) != nbytes) {
    set_error("bundle: short write to %s", w->tmp_data.c_str());
    return -1;
  }
  w->offset += static_cast<long long>(nbytes);
  w->entries[name] = e;
  return 0;
}

int dtfn_bundle_writer_finish(dtfn_bundle_writer* w) {
  int rc = 0;
  if (fclose(w->f) != 0) {
    set_error("bundle: close failed for %s", w->tmp_data.c_str());
    rc = -1;
  }
  const std::string final_data = data_path(w->prefix, 0, 1);
  if (rc == 0 && rename(w->tmp_data.c_str(), final_data.c_str()) != 0) {
    set_error("bundle: rename %s -> %s failed", w->tmp_data.c_str(), final_data.c_str());
    rc = -1;
  }
  if (rc == 0 && !write_index(w->prefix, 1, w->entries)) rc = -1;
  delete w;
  return rc;
}

void dtfn_bundle_writer_abandon(dtfn_bundle_writer* w) {
  if (w->f) fclose(w->f);
  remove(w->tmp_data.c_str());
  delete w;
}

}  // extern "C"

// ------------------------------------------------------------------------------- reader
struct dtfn_bundle_reader {
  std::string prefix;
  int num_shards = 0;
  std::vector<Entry> entries;  // sorted by name
};

extern "C" {

dtfn_bundle_reader* dtfn_bundle_reader_open(const char* prefix) {
  const std::string index = std::string(prefix) + ".index";
  dtfn_table* t = dtfn_table_open(index.c_str());
  if (!t) return nullptr;
  auto r = std::make_unique<dtfn_bundle_reader>();
  r->prefix = prefix;
  const long long n = dtfn_table_count(t);
  bool ok = n > 0;
  for (long long i = 0; ok && i < n; ++i) {
    const void *k, *v;
    size_t kl, vl;
    dtfn_table_entry(t, i, &k, &kl, &v, &vl);
    if (i == 0) {
      int endian = 0;
      if (kl != 0 || !decode_header(static_cast<const char*>(v), vl, &r->num_shards, &endian)) {
        set_error("bundle: %s has no header entry", index.c_str());
        ok = false;
      } else if (endian != 0) {
        set_error("bundle: big-endian bundles are not supported");
        ok = false;
      }
      continue;
    }
    Entry e;
    e.name.assign(static_cast<const char*>(k), kl);
    if (!decode_entry(static_cast<const char*>(v), vl, &e)) {
      set_error("bundle: malformed entry for '%s'", e.name.c_str());
      ok = false;
    }
    if (!e.slices.empty() && e.size == 0) {   // full tensor of a partitioned variable: size from shape and dtype
      long long count = 1;
      for (long long d : e.dims) count *= d;
      e.size = count * dtype_size(e.dtype);
    }
    r->entries.push_back(std::move(e));
  }
  if (n == 0) set_error("bundle: %s is empty", index.c_str());
  dtfn_table_close(t);
  return ok ? r.release() : nullptr;
}

int dtfn_bundle_reader_num_shards(const dtfn_bundle_reader* r) { return r->num_shards; }
long long dtfn_bundle_reader_count(const dtfn_bundle_reader* r) { return static_cast<long long>(r->entries.size()); }

int dtfn_bundle_reader_entry(const dtfn_bundle_reader* r, long long i, const char** name, size_t* name_len,
                             int* dtype, int* ndims, long long* dims, int* shard_id, long long* offset,
                             long long* size, uint32_t* masked_crc) {
  if (i < 0 || i >= static_cast<long long>(r->entries.size())) {
    set_error("bundle: entry %lld out of range", i);
    return -1;
  }
  const Entry& e = r->entries[i];
  if (e.dims.size() > 8) {
    set_error("bundle: '%s' has rank %zu > 8", e.name.c_str(), e.dims.size());
    return -2;
  }
  *name = e.name.data();
  *name_len = e.name.size();
  *dtype = e.dtype;
  *ndims = static_cast<int>(e.dims.size());
  for (size_t d = 0; d < e.dims.size(); ++d) dims[d] = e.dims[d];
  *shard_id = e.shard_id;
  *offset = e.offset;
  *size = e.size;
  *masked_crc = e.crc;
  return 0;
}

int dtfn_bundle_reader_lookup(const dtfn_bundle_reader* r, const char* name) {
  const std::string k(name);
  auto it = std::lower_bound(r->entries.begin(), r->entries.end(), k,
                             [](const Entry& e, const std::string& key) { return e.name < key; });
  if (it == r->entries.end() || it->name != k) return -1;
  return static_cast<int>(it - r->entries.begin());
}

int dtfn_bundle_reader_num_slices(const dtfn_bundle_reader* r, long long i) {
  if (i < 0 || i >= static_cast<long long>(r->entries.size())) return -1;
  return static_cast<int>(r->entries[i].slices.size());
}

static int lookup_key(const dtfn_bundle_reader* r, const std::string& k) {
  auto it = std::lower_bound(r->entries.begin(), r->entries.end(), k,
                             [](const Entry& e, const std::string& key) { return e.name < key; });
  if (it == r->entries.end() || it->name != k) return -1;
  return static_cast<int>(it - r->entries.begin());
}

int dtfn_bundle_reader_read(const dtfn_bundle_reader* r, long long i, void* dst, size_t nbytes) {
  if (i < 0 || i >= static_cast<long long>(r->entries.size())) {
    set_error("bundle: entry %lld out of range", i);
    return -1;
  }
  const Entry& e = r->entries[i];
  if (static_cast<long long>(nbytes) != e.size) {
    set_error("bundle: '%s' holds %lld bytes, caller asked for %zu", e.name.c_str(), e.size, nbytes);
    return -2;
  }
  if (!e.slices.empty()) {
    // partitioned variable: every slice is read from its own key into its place.  Supported: slices along
    // dimension 0 that take every other dimension whole (what PartitionedVariable with a row partitioner writes).
    long long row_bytes = dtype_size(e.dtype);
    for (size_t d = 1; d < e.dims.size(); ++d) row_bytes *= e.dims[d];
    for (const auto& sl : e.slices) {
      bool rows_only = sl.size() == e.dims.size() && !sl.empty();
      for (size_t d = 1; rows_only && d < sl.size(); ++d)
        rows_only = sl[d].second < 0 || (sl[d].first == 0 && sl[d].second == e.dims[d]);
      if (!rows_only) {
        set_error("bundle: '%s' is partitioned along a dimension other than 0 (not supported)", e.name.c_str());
        return -2;
      }
      const int si = lookup_key(r, encode_slice_key(e.name, sl));
      if (si < 0) {
        set_error("bundle: slice of '%s' starting at row %lld is missing from the index", e.name.c_str(), sl[0].first);
        return -3;
      }
      const long long start = sl[0].first, len = sl[0].second < 0 ? e.dims[0] : sl[0].second;
      if (start < 0 || start + len > e.dims[0] || r->entries[si].size != len * row_bytes) {
        set_error("bundle: slice of '%s' (rows %lld..%lld) does not fit the tensor", e.name.c_str(), start, start + len);
        return -3;
      }
      const int rc = dtfn_bundle_reader_read(r, si, static_cast<char*>(dst) + start * row_bytes,
                                             static_cast<size_t>(len * row_bytes));
      if (rc != 0) return rc;
    }
    return 0;
  }
  const std::string path = data_path(r->prefix, e.shard_id, r->num_shards);
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) {
    set_error("bundle: cannot open %s", path.c_str());
    return -1;
  }
  int rc = 0;
  if (fseeko(f, static_cast<off_t>(e.offset), SEEK_SET) != 0 || (nbytes && fread(dst, 1, nbytes, f) != nbytes)) {
    set_error("bundle: short read of '%s' from %s", e.name.c_str(), path.c_str());
    rc = -1;
  }
  fclose(f);
  if (rc == 0) {
    const uint32_t c = crc32c_extend(0, static_cast<const uint8_t*>(dst), nbytes);
    const uint32_t m = ((c >> 15) | (c << 17)) + 0xa282ead8u;
    if (m != e.crc) {
      set_error("bundle: checksum mismatch for '%s' (stored %08x, computed %08x)", e.name.c_str(), e.crc, m);
      rc = -3;
    }
  }
  return rc;
}

void dtfn_bundle_reader_close(dtfn_bundle_reader* r) { delete r; }

int dtfn_bundle_merge(const char* const* prefixes, int n, const char* merged_prefix) {
  if (n <= 0) {
    set_error("bundle merge: nothing to merge");
    return -2;
  }
  std::map<std::string, Entry> merged;
  for (int i = 0; i < n; ++i) {
    dtfn_bundle_reader* r = dtfn_bundle_reader_open(prefixes[i]);
    if (!r) return -1;
    if (r->num_shards != 1) {
      set_error("bundle merge: source %s has %d shards (expected 1)", prefixes[i], r->num_shards);
      dtfn_bundle_reader_close(r);
      return -2;
    }
    for (Entry e : r->entries) {
      if (merged.count(e.name)) {
        set_error("bundle merge: tensor '%s' appears in more than one shard", e.name.c_str());
        dtfn_bundle_reader_close(r);
        return -2;
      }
      e.shard_id = i;
      merged[e.name] = e;
    }
    dtfn_bundle_reader_close(r);
  }
  add_partitioned_views(&merged);
  for (int i = 0; i < n; ++i) {
    const std::string src = data_path(prefixes[i], 0, 1);
    const std::string dst = data_path(merged_prefix, i, n);
    if (rename(src.c_str(), dst.c_str()) != 0) {
      set_error("bundle merge: rename %s -> %s failed", src.c_str(), dst.c_str());
      return -1;
    }
  }
  if (!write_index(merged_prefix, n, merged)) return -1;
  for (int i = 0; i < n; ++i) {
    const std::string idx = std::string(prefixes[i]) + ".index";
    if (idx != std::string(merged_prefix) + ".index") remove(idx.c_str());
  }
  return 0;
}

// ------------------------------------------------------------------------------- TFRecord
}  // extern "C"

struct dtfn_record_writer {
  FILE* f = nullptr;
  std::string path;
};

extern "C" {

dtfn_record_writer* dtfn_record_writer_open(const char* path) {
  FILE* f = fopen(path, "ab");
  if (!f) {
    set_error("record writer: cannot open %s", path);
    return nullptr;
  }
  auto* w = new dtfn_record_writer();
  w->f = f;
  w->path = path;
  return w;
}

// u64 length | u32 masked_crc(length) | bytes | u32 masked_crc(bytes)
int dtfn_record_writer_write(dtfn_record_writer* w, const void* data, size_t n) {
  char hdr[12];
  const uint64_t len = n;
  memcpy(hdr, &len, 8);
  uint32_t c = crc32c_extend(0, reinterpret_cast<const uint8_t*>(hdr), 8);
  c = ((c >> 15) | (c << 17)) + 0xa282ead8u;
  memcpy(hdr + 8, &c, 4);
  uint32_t d = crc32c_extend(0, static_cast<const uint8_t*>(data), n);
  d = ((d >> 15) | (d << 17)) + 0xa282ead8u;
  if (fwrite(hdr, 1, 12, w->f) != 12 || (n && fwrite(data, 1, n, w->f) != n) || fwrite(&d, 1, 4, w->f) != 4) {
    set_error("record writer: short write to %s", w->path.c_str());
    return -1;
  }
  return 0;
}

int dtfn_record_writer_flush(dtfn_record_writer* w) { return fflush(w->f) == 0 ? 0 : -1; }

void dtfn_record_writer_close(dtfn_record_writer* w) {
  if (w->f) fclose(w->f);
  delete w;
}

}  // extern "C"
