// b200ps N8 — native TensorFlow tensor-bundle (Saver V2) writer and reader.
//
// tf.train.Saver (model_export.py:53; implicit in Supervisor) is C++ inside TensorFlow (BundleWriter /
// BundleReader over a leveldb-style table).  This is the same format written natively, so a
// checkpoint of a multi-gigabyte row-sharded weight streams from the caller's buffers to disk with
// one CRC pass and no intermediate copy:
//
//   <prefix>.data-00000-of-00001   raw little-endian tensor bytes, back to back, in the order added
//   <prefix>.index                 sorted table:  ""     -> BundleHeaderProto{num_shards=1, version{producer=1}}
//                                                 <name> -> BundleEntryProto{dtype, shape, offset, size, crc32c}
//
// Table = data blocks of prefix-compressed entries (restart every 16 keys) + restart array, each block
// followed by a compression byte (0) and a masked CRC-32C; an empty meta-index block; an index block; a
// 48-byte footer ending in 0xdb4775248b80fb57.  Both files are written under a temporary name and renamed
// on finish, so a reader never sees a torn checkpoint.  b200ps/io/tf_bundle.py is the byte-for-byte
// Python twin (the tests compare the two).
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <algorithm>
#include <string>
#include <vector>

#include "../common.h"

using namespace b200ps;

extern "C" uint32_t b200ps_crc32c_extend(uint32_t init_crc, const void* data, uint64_t n);

namespace {

constexpr uint64_t kTableMagic = 0xDB4775248B80FB57ull;
constexpr size_t kBlockSize = 262144;
constexpr int kRestartInterval = 16;

inline uint32_t mask_crc(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }

void put_varint(std::string& s, uint64_t v) {
  while (v >= 0x80) { s.push_back((char)((v & 0x7F) | 0x80)); v >>= 7; }
  s.push_back((char)v);
}
void put_fixed32(std::string& s, uint32_t v) {
  char b[4] = {(char)(v & 0xFF), (char)((v >> 8) & 0xFF), (char)((v >> 16) & 0xFF), (char)((v >> 24) & 0xFF)};
  s.append(b, 4);
}
void put_fixed64(std::string& s, uint64_t v) {
  put_fixed32(s, (uint32_t)(v & 0xFFFFFFFFu));
  put_fixed32(s, (uint32_t)(v >> 32));
}
bool get_varint(const std::string& s, size_t& pos, uint64_t* out) {
  uint64_t v = 0;
  for (int shift = 0; shift <= 63 && pos < s.size(); shift += 7) {
    const uint8_t b = (uint8_t)s[pos++];
    v |= (uint64_t)(b & 0x7F) << shift;
    if (!(b & 0x80)) { *out = v; return true; }
  }
  return false;
}

// protobuf fields as b200ps/io/proto.py writes them (zero varints omitted, sub-messages always present)
void pb_varint(std::string& s, int field, uint64_t v) {
  if (v == 0) return;
  put_varint(s, ((uint64_t)field << 3) | 0);
  put_varint(s, v);
}
void pb_msg(std::string& s, int field, const std::string& payload) {
  put_varint(s, ((uint64_t)field << 3) | 2);
  put_varint(s, payload.size());
  s += payload;
}
void pb_fixed32(std::string& s, int field, uint32_t v) {
  put_varint(s, ((uint64_t)field << 3) | 5);
  put_fixed32(s, v);
}

struct BlockBuilder {
  std::string buf;
  std::vector<uint32_t> restarts{0};
  int counter = 0;
  std::string last_key;
  void add(const std::string& key, const std::string& value) {
    size_t shared = 0;
    if (counter < kRestartInterval) {
      const size_t m = std::min(last_key.size(), key.size());
      while (shared < m && last_key[shared] == key[shared]) ++shared;
    } else {
      restarts.push_back((uint32_t)buf.size());
      counter = 0;
    }
    put_varint(buf, shared);
    put_varint(buf, key.size() - shared);
    put_varint(buf, value.size());
    buf.append(key, shared, std::string::npos);
    buf += value;
    last_key = key;
    ++counter;
  }
  size_t size_estimate() const { return buf.size() + 4 * restarts.size() + 4; }
  bool empty() const { return buf.empty(); }
  std::string finish() const {
    std::string out = buf;
    for (uint32_t r : restarts) put_fixed32(out, r);
    put_fixed32(out, (uint32_t)restarts.size());
    return out;
  }
};

std::string shortest_separator(const std::string& start, const std::string& limit) {
  const size_t m = std::min(start.size(), limit.size());
  size_t d = 0;
  while (d < m && start[d] == limit[d]) ++d;
  if (d < m) {
    const uint8_t b = (uint8_t)start[d];
    if (b < 0xFF && (unsigned)b + 1 < (uint8_t)limit[d]) {
      std::string r = start.substr(0, d);
      r.push_back((char)(b + 1));
      return r;
    }
  }
  return start;
}
std::string short_successor(const std::string& key) {
  for (size_t i = 0; i < key.size(); ++i) {
    if ((uint8_t)key[i] != 0xFF) {
      std::string r = key.substr(0, i);
      r.push_back((char)((uint8_t)key[i] + 1));
      return r;
    }
  }
  return key;
}

// Serialise sorted (key, value) pairs as an uncompressed leveldb/TF table.
std::string build_table(std::vector<std::pair<std::string, std::string>>& items) {
  std::sort(items.begin(), items.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
  std::string out;
  BlockBuilder index, block;
  bool pending = false;
  std::string pending_key, pending_handle;
  auto write_block = [&](const BlockBuilder& b) {
    const std::string raw = b.finish();
    std::string h;
    put_varint(h, out.size());
    put_varint(h, raw.size());
    out += raw;
    const char type = 0;
    uint32_t crc = b200ps_crc32c_extend(0, raw.data(), raw.size());
    crc = b200ps_crc32c_extend(crc, &type, 1);
    out.push_back(type);
    put_fixed32(out, mask_crc(crc));
    return h;
  };
  std::string last_key;
  for (const auto& kv : items) {
    if (pending) {
      index.add(shortest_separator(pending_key, kv.first), pending_handle);
      pending = false;
    }
    block.add(kv.first, kv.second);
    last_key = kv.first;
    if (block.size_estimate() >= kBlockSize) {
      pending_handle = write_block(block);
      pending_key = last_key;
      pending = true;
      block = BlockBuilder();
    }
  }
  if (!block.empty()) {
    pending_handle = write_block(block);
    pending_key = last_key;
    pending = true;
  }
  const std::string meta_h = write_block(BlockBuilder());
  if (pending) index.add(short_successor(pending_key), pending_handle);
  const std::string index_h = write_block(index);
  std::string footer = meta_h + index_h;
  footer.resize(40, '\0');
  put_fixed64(footer, kTableMagic);
  out += footer;
  return out;
}

}  // namespace

// ------------------------------------------------------------------------------------ writer
namespace {

struct Writer {
  std::string prefix;
  FILE* data = nullptr;
  uint64_t offset = 0;
  std::vector<std::pair<std::string, std::string>> items;
};

std::string entry_proto(int dtype, const int64_t* shape, int ndim, uint64_t offset, uint64_t size, uint32_t crc) {
  std::string shape_pb;
  for (int i = 0; i < ndim; ++i) {
    std::string dim;
    pb_varint(dim, 1, (uint64_t)shape[i]);      // TensorShapeProto.Dim.size (zero omitted)
    pb_msg(shape_pb, 2, dim);
  }
  std::string e;
  pb_varint(e, 1, (uint64_t)dtype);
  pb_msg(e, 2, shape_pb);
  pb_varint(e, 4, offset);
  pb_varint(e, 5, size);
  pb_fixed32(e, 6, crc);
  return e;
}

}  // namespace

extern "C" int b200ps_bundle_writer_create(void** out, const char* prefix) {
  B200PS_REQUIRE(out && prefix && *prefix, "bundle_writer_create: null argument");
  Writer* w = new Writer();
  w->prefix = prefix;
  const std::string tmp = w->prefix + ".data-00000-of-00001.tmp";
  w->data = fopen(tmp.c_str(), "wb");
  if (!w->data) {
    delete w;
    return set_error(B200PS_ERR_INVALID, "bundle_writer_create: cannot open %s", tmp.c_str());
  }
  std::string version, header;
  pb_varint(version, 1, 1);                      // VersionDef.producer = kTensorBundleVersion
  pb_varint(header, 1, 1);                       // num_shards; endianness LITTLE (0) is the default
  pb_msg(header, 3, version);
  w->items.emplace_back(std::string(), header);
  *out = w;
  return 0;
}

// Append one tensor: `nbytes` of little-endian data in C order.  dtype is the TensorFlow DataType enum
// (1 float32, 2 float64, 3 int32, 9 int64, ...).  Tensors may be added in any order: the index is sorted on
// finish; the data file keeps the order of the calls.
extern "C" int b200ps_bundle_writer_add(void* writer, const char* name, int32_t dtype, const int64_t* shape,
                                        int32_t ndim, const void* data, uint64_t nbytes) {
  Writer* w = static_cast<Writer*>(writer);
  B200PS_REQUIRE(w && w->data && name && *name && ndim >= 0 && ndim <= 8 && (data || nbytes == 0),
                 "bundle_writer_add: bad arguments");
  uint32_t crc = 0;
  const char* p = static_cast<const char*>(data);
  uint64_t left = nbytes;
  while (left) {                                 // 16 MiB at a time: CRC while the lines are hot, then write
    const uint64_t n = std::min<uint64_t>(left, 16u << 20);
    crc = b200ps_crc32c_extend(crc, p, n);
    if (fwrite(p, 1, n, w->data) != n) return set_error(B200PS_ERR_INVALID, "bundle_writer_add: short write (%s)", name);
    p += n;
    left -= n;
  }
  w->items.emplace_back(std::string(name), entry_proto(dtype, shape, ndim, w->offset, nbytes, mask_crc(crc)));
  w->offset += nbytes;
  return 0;
}

// Write the index, close and rename both files into place.  The writer is destroyed either way.
extern "C" int b200ps_bundle_writer_finish(void* writer) {
  Writer* w = static_cast<Writer*>(writer);
  B200PS_REQUIRE(w, "bundle_writer_finish: null writer");
  int rc = 0;
  const std::string data_path = w->prefix + ".data-00000-of-00001", index_path = w->prefix + ".index";
  if (w->data && fclose(w->data) != 0) rc = set_error(B200PS_ERR_INVALID, "bundle_writer_finish: closing the data file failed");
  w->data = nullptr;
  if (!rc) {
    const std::string table = build_table(w->items);
    FILE* f = fopen((index_path + ".tmp").c_str(), "wb");
    if (!f || fwrite(table.data(), 1, table.size(), f) != table.size() || fclose(f) != 0)
      rc = set_error(B200PS_ERR_INVALID, "bundle_writer_finish: cannot write %s.tmp", index_path.c_str());
  }
  if (!rc && (rename((data_path + ".tmp").c_str(), data_path.c_str()) != 0 ||
              rename((index_path + ".tmp").c_str(), index_path.c_str()) != 0))
    rc = set_error(B200PS_ERR_INVALID, "bundle_writer_finish: rename failed for %s", w->prefix.c_str());
  if (rc) {
    remove((data_path + ".tmp").c_str());
    remove((index_path + ".tmp").c_str());
  }
  delete w;
  return rc;
}

// Abandon a writer without producing a checkpoint.
extern "C" int b200ps_bundle_writer_abort(void* writer) {
  Writer* w = static_cast<Writer*>(writer);
  if (!w) return 0;
  if (w->data) fclose(w->data);
  remove((w->prefix + ".data-00000-of-00001.tmp").c_str());
  delete w;
  return 0;
}

// ------------------------------------------------------------------------------------ reader
namespace {

struct Entry {
  std::string name;
  int dtype = 0, ndim = 0, shard_id = 0;
  int64_t shape[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  uint64_t offset = 0, size = 0;
  uint32_t crc = 0;
};

struct Reader {
  std::string prefix;
  int num_shards = 1;
  std::vector<Entry> entries;
  FILE* data = nullptr;
};

bool read_file(const std::string& path, std::string* out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  char buf[1 << 16];
  size_t n;
  while ((n = fread(buf, 1, sizeof buf, f)) > 0) out->append(buf, n);
  fclose(f);
  return true;
}

// One table block (checksum verified) -> its (key, value) pairs appended to `out`.
bool block_entries(const std::string& file, uint64_t off, uint64_t size,
                   std::vector<std::pair<std::string, std::string>>* out, const char** why) {
  if (off + size + 5 > file.size()) { *why = "truncated table block"; return false; }
  if (file[off + size] != 0) { *why = "compressed table blocks are not supported"; return false; }
  uint32_t stored;
  memcpy(&stored, file.data() + off + size + 1, 4);
  if (stored != mask_crc(b200ps_crc32c_extend(0, file.data() + off, size + 1))) { *why = "table block checksum mismatch"; return false; }
  if (size < 4) { *why = "table block too small"; return false; }
  const std::string raw = file.substr(off, size);
  uint32_t n_restarts;
  memcpy(&n_restarts, raw.data() + raw.size() - 4, 4);
  if ((uint64_t)n_restarts * 4 + 4 > raw.size()) { *why = "bad restart array"; return false; }
  const size_t limit = raw.size() - 4 - 4 * (size_t)n_restarts;
  size_t pos = 0;
  std::string key;
  while (pos < limit) {
    uint64_t shared, non_shared, vlen;
    if (!get_varint(raw, pos, &shared) || !get_varint(raw, pos, &non_shared) || !get_varint(raw, pos, &vlen) ||
        shared > key.size() || pos + non_shared + vlen > limit) { *why = "corrupt table entry"; return false; }
    key.resize(shared);
    key.append(raw, pos, non_shared);
    pos += non_shared;
    out->emplace_back(key, raw.substr(pos, vlen));
    pos += vlen;
  }
  return true;
}

// Minimal protobuf walk: calls f(field, wire_type, varint_value, bytes) for every field of `msg`.
template <typename F>
bool pb_walk(const std::string& msg, F f) {
  size_t pos = 0;
  while (pos < msg.size()) {
    uint64_t tag;
    if (!get_varint(msg, pos, &tag)) return false;
    const int field = (int)(tag >> 3), wt = (int)(tag & 7);
    if (wt == 0) {
      uint64_t v;
      if (!get_varint(msg, pos, &v)) return false;
      f(field, wt, v, std::string());
    } else if (wt == 2) {
      uint64_t n;
      if (!get_varint(msg, pos, &n) || pos + n > msg.size()) return false;
      f(field, wt, 0, msg.substr(pos, n));
      pos += n;
    } else if (wt == 5 || wt == 1) {
      const size_t n = wt == 5 ? 4 : 8;
      if (pos + n > msg.size()) return false;
      uint64_t v = 0;
      memcpy(&v, msg.data() + pos, n);
      f(field, wt, v, std::string());
      pos += n;
    } else {
      return false;
    }
  }
  return true;
}

}  // namespace

extern "C" int b200ps_bundle_reader_open(void** out, const char* prefix) {
  B200PS_REQUIRE(out && prefix && *prefix, "bundle_reader_open: null argument");
  std::string file;
  if (!read_file(std::string(prefix) + ".index", &file))
    return set_error(B200PS_ERR_INVALID, "bundle_reader_open: cannot read %s.index", prefix);
  uint64_t magic = 0;
  if (file.size() >= 48) memcpy(&magic, file.data() + file.size() - 8, 8);
  B200PS_REQUIRE(file.size() >= 48 && magic == kTableMagic, "bundle_reader_open: %s.index is not a TF/leveldb table", prefix);
  const std::string footer = file.substr(file.size() - 48);
  size_t pos = 0;
  uint64_t mo, ms, io, isz;
  B200PS_REQUIRE(get_varint(footer, pos, &mo) && get_varint(footer, pos, &ms) && get_varint(footer, pos, &io) &&
                 get_varint(footer, pos, &isz), "bundle_reader_open: bad footer");
  const char* why = "";
  std::vector<std::pair<std::string, std::string>> index, kv;
  if (!block_entries(file, io, isz, &index, &why)) return set_error(B200PS_ERR_INVALID, "bundle_reader_open: %s", why);
  for (const auto& h : index) {
    size_t p = 0;
    uint64_t off, size;
    if (!get_varint(h.second, p, &off) || !get_varint(h.second, p, &size) || !block_entries(file, off, size, &kv, &why))
      return set_error(B200PS_ERR_INVALID, "bundle_reader_open: %s", *why ? why : "bad block handle");
  }
  Reader* r = new Reader();
  r->prefix = prefix;
  bool ok = true, have_header = false;
  for (const auto& e : kv) {
    if (e.first.empty()) {
      have_header = true;
      ok = ok && pb_walk(e.second, [&](int f, int, uint64_t v, const std::string&) { if (f == 1) r->num_shards = (int)v; });
      continue;
    }
    Entry en;
    en.name = e.first;
    ok = ok && pb_walk(e.second, [&](int f, int, uint64_t v, const std::string& b) {
      if (f == 1) en.dtype = (int)v;
      else if (f == 3) en.shard_id = (int)v;
      else if (f == 4) en.offset = v;
      else if (f == 5) en.size = v;
      else if (f == 6) en.crc = (uint32_t)v;
      else if (f == 2)
        pb_walk(b, [&](int f2, int, uint64_t, const std::string& dim) {
          if (f2 != 2 || en.ndim >= 8) return;
          int64_t sz = 0;
          pb_walk(dim, [&](int f3, int, uint64_t v3, const std::string&) { if (f3 == 1) sz = (int64_t)v3; });
          en.shape[en.ndim++] = sz;
        });
    });
    r->entries.push_back(en);
  }
  if (!ok || !have_header) {
    delete r;
    return set_error(B200PS_ERR_INVALID, "bundle_reader_open: %s", have_header ? "corrupt entry proto" : "bundle has no header entry");
  }
  *out = r;
  return 0;
}

extern "C" int b200ps_bundle_reader_count(void* reader, int32_t* n) {
  Reader* r = static_cast<Reader*>(reader);
  B200PS_REQUIRE(r && n, "bundle_reader_count: null argument");
  *n = (int32_t)r->entries.size();
  return 0;
}

// Entry i (sorted by name): name (NUL-terminated, truncated to name_cap), dtype enum, shape[8], ndim, byte size.
extern "C" int b200ps_bundle_reader_entry(void* reader, int32_t i, char* name, int32_t name_cap, int32_t* dtype,
                                          int64_t* shape8, int32_t* ndim, uint64_t* nbytes) {
  Reader* r = static_cast<Reader*>(reader);
  B200PS_REQUIRE(r && i >= 0 && i < (int)r->entries.size() && name && name_cap > 0 && dtype && shape8 && ndim && nbytes,
                 "bundle_reader_entry: bad arguments");
  const Entry& e = r->entries[i];
  snprintf(name, (size_t)name_cap, "%s", e.name.c_str());
  *dtype = e.dtype;
  *ndim = e.ndim;
  for (int k = 0; k < 8; ++k) shape8[k] = e.shape[k];
  *nbytes = e.size;
  return 0;
}

// Read entry i into dst (nbytes must equal its size); the data checksum is verified unless verify == 0.
extern "C" int b200ps_bundle_reader_read(void* reader, int32_t i, void* dst, uint64_t nbytes, int32_t verify) {
  Reader* r = static_cast<Reader*>(reader);
  B200PS_REQUIRE(r && i >= 0 && i < (int)r->entries.size() && (dst || nbytes == 0), "bundle_reader_read: bad arguments");
  const Entry& e = r->entries[i];
  B200PS_REQUIRE(nbytes == e.size, "bundle_reader_read: %s holds %llu bytes, not %llu", e.name.c_str(),
                 (unsigned long long)e.size, (unsigned long long)nbytes);
  char path[4096];
  snprintf(path, sizeof path, "%s.data-%05d-of-%05d", r->prefix.c_str(), e.shard_id, r->num_shards);
  FILE* f = fopen(path, "rb");
  if (!f) return set_error(B200PS_ERR_INVALID, "bundle_reader_read: cannot open %s", path);
  int rc = 0;
  if (fseeko(f, (off_t)e.offset, SEEK_SET) != 0 || fread(dst, 1, nbytes, f) != nbytes)
    rc = set_error(B200PS_ERR_INVALID, "%s: truncated data file", e.name.c_str());
  fclose(f);
  if (!rc && verify && mask_crc(b200ps_crc32c_extend(0, dst, nbytes)) != e.crc)
    rc = set_error(B200PS_ERR_INVALID, "%s: data checksum mismatch", e.name.c_str());
  return rc;
}

extern "C" int b200ps_bundle_reader_close(void* reader) {
  delete static_cast<Reader*>(reader);
  return 0;
}
