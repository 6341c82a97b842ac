// Immutable sorted string table in the LevelDB / TensorFlow "table" format (no compression):
//   [data block]* [metaindex block] [index block] [48-byte footer]
// Every block is followed by a 5-byte trailer: compression type (0) + masked crc32c(block||type).
// Block: prefix-compressed entries (varint shared, varint non_shared, varint value_len, key suffix,
// value) with a restart point every `interval` entries, then u32 restart offsets and u32 count.
// This is the container of the checkpoint-V2 ".index" file written by tf.train.Saver
// (reference trainer/task.py:143).
#include <algorithm>
#include <memory>

#include "../../include/dtf_native.h"
#include "util.h"

namespace dtfn {

uint32_t crc32c_extend(uint32_t crc, const uint8_t* p, size_t n);
static const uint64_t kTableMagic = 0xdb4775248b80fb57ull;

static uint32_t masked_crc(const std::string& block, char type) {
  uint32_t c = crc32c_extend(0, reinterpret_cast<const uint8_t*>(block.data()), block.size());
  c = crc32c_extend(c, reinterpret_cast<const uint8_t*>(&type), 1);
  return ((c >> 15) | (c << 17)) + 0xa282ead8u;
}

class BlockBuilder {
 public:
  explicit BlockBuilder(int interval) : interval_(interval) { restarts_.push_back(0); }
  void add(const std::string& key, const std::string& val) {
    size_t shared = 0;
    if (counter_ < interval_) {
      const size_t m = std::min(last_key_.size(), key.size());
      while (shared < m && last_key_[shared] == key[shared]) ++shared;
    } else {
      restarts_.push_back(static_cast<uint32_t>(buf_.size()));
      counter_ = 0;
    }
    put_varint32(&buf_, static_cast<uint32_t>(shared));
    put_varint32(&buf_, static_cast<uint32_t>(key.size() - shared));
    put_varint32(&buf_, static_cast<uint32_t>(val.size()));
    buf_.append(key.data() + shared, key.size() - shared);
    buf_.append(val);
    last_key_ = key;
    ++counter_;
    ++entries_;
  }
  std::string finish() {
    std::string out = buf_;
    for (uint32_t r : restarts_) put_fixed32(&out, r);
    put_fixed32(&out, static_cast<uint32_t>(restarts_.size()));
    return out;
  }
  size_t size_estimate() const { return buf_.size() + restarts_.size() * 4 + 4; }
  bool empty() const { return entries_ == 0; }
  void reset() {
    buf_.clear();
    restarts_.assign(1, 0);
    counter_ = 0;
    entries_ = 0;
    last_key_.clear();
  }

 private:
  int interval_;
  std::string buf_;
  std::vector<uint32_t> restarts_;
  int counter_ = 0;
  int entries_ = 0;
  std::string last_key_;
};

struct Handle {
  uint64_t offset = 0, size = 0;
  void encode(std::string* s) const {
    put_varint64(s, offset);
    put_varint64(s, size);
  }
};

}  // namespace dtfn

using namespace dtfn;

struct dtfn_table_builder {
  FILE* f = nullptr;
  std::string path;
  int block_size = 262144;
  uint64_t offset = 0;
  BlockBuilder data{16};
  BlockBuilder index{1};
  std::string last_key;
  bool have_last = false;
  bool pending = false;
  Handle pending_handle;
  bool failed = false;

  bool write_block(const std::string& contents, Handle* h) {
    h->offset = offset;
    h->size = contents.size();
    std::string trailer(1, '\0');
    put_fixed32(&trailer, masked_crc(contents, '\0'));
    if (fwrite(contents.data(), 1, contents.size(), f) != contents.size() ||
        fwrite(trailer.data(), 1, trailer.size(), f) != trailer.size()) {
      set_error("table: write to %s failed", path.c_str());
      failed = true;
      return false;
    }
    offset += contents.size() + trailer.size();
    return true;
  }
  bool flush_data() {
    if (data.empty()) return true;
    if (!write_block(data.finish(), &pending_handle)) return false;
    data.reset();
    pending = true;
    return true;
  }
};

extern "C" {

dtfn_table_builder* dtfn_table_builder_new(const char* path, int block_size) {
  FILE* f = fopen(path, "wb");
  if (!f) {
    set_error("table: cannot open %s for writing", path);
    return nullptr;
  }
  auto* b = new dtfn_table_builder();
  b->f = f;
  b->path = path;
  if (block_size > 0) b->block_size = block_size;
  return b;
}

int dtfn_table_builder_add(dtfn_table_builder* b, const void* key, size_t klen, const void* val, size_t vlen) {
  if (b->failed) return -1;
  std::string k(static_cast<const char*>(key), klen), v(static_cast<const char*>(val), vlen);
  if (b->have_last && !(b->last_key < k)) {
    set_error("table: keys must be added in strictly increasing order ('%s' after '%s')", k.c_str(),
              b->last_key.c_str());
    return -2;
  }
  if (b->pending) {
    // index key: anything >= last key of the finished block and < k; the last key itself qualifies
    std::string enc;
    b->pending_handle.encode(&enc);
    b->index.add(b->last_key, enc);
    b->pending = false;
  }
  b->data.add(k, v);
  b->last_key = k;
  b->have_last = true;
  if (b->data.size_estimate() >= static_cast<size_t>(b->block_size)) {
    if (!b->flush_data()) return -1;
  }
  return 0;
}

int dtfn_table_builder_finish(dtfn_table_builder* b) {
  int rc = 0;
  if (!b->failed && b->flush_data()) {
    if (b->pending) {
      std::string enc;
      b->pending_handle.encode(&enc);
      b->index.add(b->last_key, enc);
      b->pending = false;
    }
    Handle meta_h, index_h;
    BlockBuilder meta(1);
    if (b->write_block(meta.finish(), &meta_h) && b->write_block(b->index.finish(), &index_h)) {
      std::string footer;
      meta_h.encode(&footer);
      index_h.encode(&footer);
      footer.resize(40, '\0');
      put_fixed64(&footer, kTableMagic);
      if (fwrite(footer.data(), 1, footer.size(), b->f) != footer.size()) {
        set_error("table: footer write failed for %s", b->path.c_str());
        rc = -1;
      }
    } else {
      rc = -1;
    }
  } else {
    rc = -1;
  }
  if (fclose(b->f) != 0 && rc == 0) {
    set_error("table: close failed for %s", b->path.c_str());
    rc = -1;
  }
  delete b;
  return rc;
}

void dtfn_table_builder_abandon(dtfn_table_builder* b) {
  if (b->f) fclose(b->f);
  remove(b->path.c_str());
  delete b;
}

}  // extern "C"

// ------------------------------------------------------------------------------- reader
struct dtfn_table {
  std::vector<std::string> keys, vals;
};

namespace {

bool read_block(const std::string& file, const Handle& h, std::string* out) {
  if (h.offset + h.size + 5 > file.size()) {
    set_error("table: block handle out of range");
    return false;
  }
  out->assign(file.data() + h.offset, h.size);
  const char type = file[h.offset + h.size];
  const uint32_t stored = get_fixed32(file.data() + h.offset + h.size + 1);
  if (type != 0) {
    set_error("table: compressed blocks are not supported (type %d)", (int)type);
    return false;
  }
  if (stored != masked_crc(*out, type)) {
    set_error("table: block checksum mismatch at offset %llu", (unsigned long long)h.offset);
    return false;
  }
  return true;
}

bool parse_block(const std::string& blk, std::vector<std::string>* keys, std::vector<std::string>* vals) {
  if (blk.size() < 4) return false;
  const uint32_t nrestart = get_fixed32(blk.data() + blk.size() - 4);
  if (static_cast<uint64_t>(nrestart) * 4 + 4 > blk.size()) return false;
  const char* p = blk.data();
  const char* limit = blk.data() + blk.size() - 4 - 4ull * nrestart;
  std::string key;
  while (p < limit) {
    uint32_t shared, non_shared, vlen;
    p = get_varint32(p, limit, &shared);
    if (!p) return false;
    p = get_varint32(p, limit, &non_shared);
    if (!p) return false;
    p = get_varint32(p, limit, &vlen);
    if (!p) return false;
    if (shared > key.size() || static_cast<uint64_t>(limit - p) < static_cast<uint64_t>(non_shared) + vlen) return false;
    key.resize(shared);
    key.append(p, non_shared);
    p += non_shared;
    keys->push_back(key);
    vals->emplace_back(p, vlen);
    p += vlen;
  }
  return true;
}

bool decode_handle(const char** p, const char* limit, Handle* h) {
  const char* q = get_varint64(*p, limit, &h->offset);
  if (!q) return false;
  q = get_varint64(q, limit, &h->size);
  if (!q) return false;
  *p = q;
  return true;
}

}  // namespace

extern "C" {

dtfn_table* dtfn_table_open(const char* path) {
  FILE* f = fopen(path, "rb");
  if (!f) {
    set_error("table: cannot open %s", path);
    return nullptr;
  }
  std::string file;
  char buf[1 << 16];
  size_t n;
  while ((n = fread(buf, 1, sizeof(buf), f)) > 0) file.append(buf, n);
  fclose(f);
  if (file.size() < 48) {
    set_error("table: %s is too short to be a table", path);
    return nullptr;
  }
  const char* footer = file.data() + file.size() - 48;
  if (get_fixed64(footer + 40) != kTableMagic) {
    set_error("table: bad magic number in %s", path);
    return nullptr;
  }
  Handle meta_h, index_h;
  const char* p = footer;
  if (!decode_handle(&p, footer + 40, &meta_h) || !decode_handle(&p, footer + 40, &index_h)) {
    set_error("table: malformed footer in %s", path);
    return nullptr;
  }
  std::string index_blk;
  if (!read_block(file, index_h, &index_blk)) return nullptr;
  std::vector<std::string> ikeys, ivals;
  if (!parse_block(index_blk, &ikeys, &ivals)) {
    set_error("table: malformed index block in %s", path);
    return nullptr;
  }
  auto t = std::make_unique<dtfn_table>();
  for (const std::string& hv : ivals) {
    Handle h;
    const char* q = hv.data();
    if (!decode_handle(&q, hv.data() + hv.size(), &h)) {
      set_error("table: malformed block handle in %s", path);
      return nullptr;
    }
    std::string blk;
    if (!read_block(file, h, &blk)) return nullptr;
    if (!parse_block(blk, &t->keys, &t->vals)) {
      set_error("table: malformed data block in %s", path);
      return nullptr;
    }
  }
  return t.release();
}

long long dtfn_table_count(const dtfn_table* t) { return static_cast<long long>(t->keys.size()); }

int dtfn_table_entry(const dtfn_table* t, long long i, const void** key, size_t* klen, const void** val,
                     size_t* vlen) {
  if (i < 0 || i >= static_cast<long long>(t->keys.size())) {
    set_error("table: entry index %lld out of range", i);
    return -1;
  }
  *key = t->keys[i].data();
  *klen = t->keys[i].size();
  *val = t->vals[i].data();
  *vlen = t->vals[i].size();
  return 0;
}

int dtfn_table_find(const dtfn_table* t, const void* key, size_t klen, const void** val, size_t* vlen) {
  std::string k(static_cast<const char*>(key), klen);
  auto it = std::lower_bound(t->keys.begin(), t->keys.end(), k);
  if (it == t->keys.end() || *it != k) return -1;
  const size_t i = it - t->keys.begin();
  *val = t->vals[i].data();
  *vlen = t->vals[i].size();
  return 0;
}

void dtfn_table_close(dtfn_table* t) { delete t; }

}  // extern "C"
