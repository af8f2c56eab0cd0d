// TF1 V2 checkpoint ("tensor bundle") writer/reader.
//
//   <prefix>.data-00000-of-00001 : raw little-endian tensor bytes, concatenated in key order
//   <prefix>.index               : SSTable; key "" -> BundleHeaderProto, key <name> -> BundleEntryProto
//
// Reference: FastSaver(var_list).save(..., write_meta_graph=False)  (utils.py:28-32,
// worker.py:102-103) which writes global/dense/{kernel,bias}, global/dense_1/{kernel,bias} and the
// int32 global/global_step.  Both files are written to a temporary name and renamed.
#include "sstable.h"

#include <string.h>

#include <map>

using namespace dtfe_host;

namespace {

struct TensorRec {
  std::string name;
  int dtype;  // TF DataType enum: DT_FLOAT=1, DT_DOUBLE=2, DT_INT32=3, DT_INT64=9, DT_BFLOAT16=14
  std::vector<int64_t> shape;
  std::string bytes;
  uint64_t offset = 0;
  uint32_t crc = 0;
};

struct Writer {
  std::map<std::string, TensorRec> tensors;  // std::map iterates in bytewise key order
};

struct Reader {
  std::vector<TensorRec> tensors;  // data bytes loaded lazily into .bytes
  std::string data;
  int num_shards = 0;
};

std::string EncodeShape(const std::vector<int64_t>& shape) {
  std::string s;
  for (int64_t d : shape) {
    std::string dim;
    put_varint_field(dim, 1, static_cast<uint64_t>(d));  // TensorShapeProto.Dim.size
    put_bytes_field(s, 2, dim);                          // TensorShapeProto.dim
  }
  return s;
}

std::string EncodeHeader() {
  std::string h;
  put_varint_field(h, 1, 1);  // num_shards = 1; endianness LITTLE(0) is the proto3 default -> omitted
  std::string ver;
  put_varint_field(ver, 1, 1);  // VersionDef.producer = 1 (kTensorBundleVersion)
  put_bytes_field(h, 3, ver);
  return h;
}

std::string EncodeEntry(const TensorRec& t) {
  std::string e;
  put_varint_field(e, 1, static_cast<uint64_t>(t.dtype));
  put_bytes_field(e, 2, EncodeShape(t.shape));  // present even when empty (scalar): "12 00"
  // shard_id (3) = 0 omitted
  if (t.offset != 0) put_varint_field(e, 4, t.offset);
  if (!t.bytes.empty()) put_varint_field(e, 5, t.bytes.size());
  put_tag(e, 6, 5);
  put_fixed32(e, t.crc);
  return e;
}

bool SkipField(const uint8_t*& p, const uint8_t* end, int wt) {
  uint64_t v;
  switch (wt) {
    case 0: return get_varint(p, end, &v);
    case 1: if (end - p < 8) return false; p += 8; return true;
    case 2: if (!get_varint(p, end, &v) || (uint64_t)(end - p) < v) return false; p += v; return true;
    case 5: if (end - p < 4) return false; p += 4; return true;
    default: return false;
  }
}

bool DecodeShape(const uint8_t* p, const uint8_t* end, std::vector<int64_t>* shape) {
  while (p < end) {
    uint64_t tag;
    if (!get_varint(p, end, &tag)) return false;
    int field = tag >> 3, wt = tag & 7;
    if (field == 2 && wt == 2) {
      uint64_t len;
      if (!get_varint(p, end, &len) || (uint64_t)(end - p) < len) return false;
      const uint8_t* q = p;
      const uint8_t* qend = p + len;
      int64_t size = 0;
      while (q < qend) {
        uint64_t t2;
        if (!get_varint(q, qend, &t2)) return false;
        if ((t2 >> 3) == 1 && (t2 & 7) == 0) {
          uint64_t v;
          if (!get_varint(q, qend, &v)) return false;
          size = static_cast<int64_t>(v);
        } else if (!SkipField(q, qend, t2 & 7)) {
          return false;
        }
      }
      shape->push_back(size);
      p = qend;
    } else if (!SkipField(p, end, wt)) {
      return false;
    }
  }
  return true;
}

bool DecodeEntry(const std::string& v, TensorRec* t, uint64_t* size, int* shard) {
  const uint8_t* p = reinterpret_cast<const uint8_t*>(v.data());
  const uint8_t* end = p + v.size();
  *size = 0;
  *shard = 0;
  while (p < end) {
    uint64_t tag, x;
    if (!get_varint(p, end, &tag)) return false;
    int field = tag >> 3, wt = tag & 7;
    if (field == 1 && wt == 0) { if (!get_varint(p, end, &x)) return false; t->dtype = (int)x; }
    else if (field == 2 && wt == 2) {
      if (!get_varint(p, end, &x) || (uint64_t)(end - p) < x) return false;
      if (!DecodeShape(p, p + x, &t->shape)) return false;
      p += x;
    }
    else if (field == 3 && wt == 0) { if (!get_varint(p, end, &x)) return false; *shard = (int)x; }
    else if (field == 4 && wt == 0) { if (!get_varint(p, end, &x)) return false; t->offset = x; }
    else if (field == 5 && wt == 0) { if (!get_varint(p, end, &x)) return false; *size = x; }
    else if (field == 6 && wt == 5) { if (end - p < 4) return false; memcpy(&t->crc, p, 4); p += 4; }
    else if (!SkipField(p, end, wt)) return false;
  }
  return true;
}

bool DecodeHeader(const std::string& v, int* num_shards, int* endianness) {
  const uint8_t* p = reinterpret_cast<const uint8_t*>(v.data());
  const uint8_t* end = p + v.size();
  *num_shards = 0;
  *endianness = 0;
  while (p < end) {
    uint64_t tag, x;
    if (!get_varint(p, end, &tag)) return false;
    int field = tag >> 3, wt = tag & 7;
    if (field == 1 && wt == 0) { if (!get_varint(p, end, &x)) return false; *num_shards = (int)x; }
    else if (field == 2 && wt == 0) { if (!get_varint(p, end, &x)) return false; *endianness = (int)x; }
    else if (!SkipField(p, end, wt)) return false;
  }
  return true;
}

}  // namespace

DTFE_HOST_API void* dtfe_bundle_writer_new() { return new Writer(); }

DTFE_HOST_API int dtfe_bundle_writer_add(void* w, const char* name, int dtype, const int64_t* shape, int rank,
                                         const void* data, size_t nbytes) {
  Writer* wr = static_cast<Writer*>(w);
  std::string key(name);
  if (key.empty()) {
    set_error("bundle: empty tensor name is reserved for the header");
    return -1;
  }
  if (wr->tensors.count(key)) {
    set_error("bundle: duplicate tensor %s", name);
    return -1;
  }
  TensorRec t;
  t.name = key;
  t.dtype = dtype;
  t.shape.assign(shape, shape + rank);
  t.bytes.assign(static_cast<const char*>(data), nbytes);
  wr->tensors.emplace(key, std::move(t));
  return 0;
}

// Writes <prefix>.index and <prefix>.data-00000-of-00001 (tmp + rename) and frees the writer.
DTFE_HOST_API int dtfe_bundle_writer_finish(void* w, const char* prefix) {
  Writer* wr = static_cast<Writer*>(w);
  std::string data;
  std::vector<std::pair<std::string, std::string>> kv;
  kv.emplace_back("", EncodeHeader());
  for (auto& it : wr->tensors) {
    TensorRec& t = it.second;
    t.offset = data.size();
    t.crc = crc32c_mask(crc32c(t.bytes.data(), t.bytes.size()));
    data.append(t.bytes);
    kv.emplace_back(t.name, EncodeEntry(t));
  }
  std::string index;
  bool ok = sstable_build(kv, 256 * 1024, &index);
  std::string p(prefix);
  ok = ok && write_file_atomic(p + ".data-00000-of-00001", data);
  ok = ok && write_file_atomic(p + ".index", index);
  delete wr;
  return ok ? 0 : -1;
}

DTFE_HOST_API void dtfe_bundle_writer_abort(void* w) { delete static_cast<Writer*>(w); }

DTFE_HOST_API void* dtfe_bundle_reader_open(const char* prefix) {
  std::string p(prefix), index;
  if (!read_file(p + ".index", &index)) return nullptr;
  std::vector<std::pair<std::string, std::string>> kv;
  if (!sstable_parse(index, &kv)) return nullptr;
  if (kv.empty() || !kv[0].first.empty()) {
    set_error("bundle: %s.index has no header entry", prefix);
    return nullptr;
  }
  Reader* r = new Reader();
  int endian;
  if (!DecodeHeader(kv[0].second, &r->num_shards, &endian) || endian != 0) {
    set_error("bundle: bad or big-endian header in %s.index", prefix);
    delete r;
    return nullptr;
  }
  if (r->num_shards != 1) {
    set_error("bundle: %d data shards not supported (the reference's Saver is non-sharded)", r->num_shards);
    delete r;
    return nullptr;
  }
  if (!read_file(p + ".data-00000-of-00001", &r->data)) {
    delete r;
    return nullptr;
  }
  for (size_t i = 1; i < kv.size(); ++i) {
    TensorRec t;
    t.name = kv[i].first;
    uint64_t size;
    int shard;
    if (!DecodeEntry(kv[i].second, &t, &size, &shard) || shard != 0 || t.offset + size > r->data.size()) {
      set_error("bundle: corrupt entry for %s", t.name.c_str());
      delete r;
      return nullptr;
    }
    uint32_t actual = crc32c_mask(crc32c(r->data.data() + t.offset, size));
    if (actual != t.crc) {
      set_error("bundle: crc mismatch for tensor %s", t.name.c_str());
      delete r;
      return nullptr;
    }
    t.bytes.assign(r->data.data() + t.offset, size);
    r->tensors.push_back(std::move(t));
  }
  return r;
}

DTFE_HOST_API int dtfe_bundle_reader_count(void* r) { return (int)static_cast<Reader*>(r)->tensors.size(); }

// out_meta: [0]=dtype [1]=rank [2]=nbytes [3]=offset [4]=masked crc; shape gets up to 8 dims.
DTFE_HOST_API const char* dtfe_bundle_reader_entry(void* r, int i, int64_t* out_meta, int64_t* shape8) {
  Reader* rd = static_cast<Reader*>(r);
  if (i < 0 || i >= (int)rd->tensors.size()) return nullptr;
  const TensorRec& t = rd->tensors[i];
  out_meta[0] = t.dtype;
  out_meta[1] = (int64_t)t.shape.size();
  out_meta[2] = (int64_t)t.bytes.size();
  out_meta[3] = (int64_t)t.offset;
  out_meta[4] = (int64_t)t.crc;
  for (size_t d = 0; d < t.shape.size() && d < 8; ++d) shape8[d] = t.shape[d];
  return t.name.c_str();
}

DTFE_HOST_API int dtfe_bundle_reader_copy(void* r, int i, void* dst, size_t cap) {
  Reader* rd = static_cast<Reader*>(r);
  if (i < 0 || i >= (int)rd->tensors.size()) return -1;
  const TensorRec& t = rd->tensors[i];
  if (cap < t.bytes.size()) {
    set_error("bundle: destination too small for %s", t.name.c_str());
    return -1;
  }
  memcpy(dst, t.bytes.data(), t.bytes.size());
  return 0;
}

DTFE_HOST_API void dtfe_bundle_reader_close(void* r) { delete static_cast<Reader*>(r); }

// Raw SSTable access for tests/goldens: build from parallel key/value arrays.
DTFE_HOST_API int dtfe_sstable_write(const char* path, const char* const* keys, const size_t* klens,
                                     const char* const* vals, const size_t* vlens, int n, size_t block_size) {
  std::vector<std::pair<std::string, std::string>> kv;
  for (int i = 0; i < n; ++i) kv.emplace_back(std::string(keys[i], klens[i]), std::string(vals[i], vlens[i]));
  std::string out;
  if (!sstable_build(kv, block_size, &out)) return -1;
  return write_file_atomic(path, out) ? 0 : -1;
}

// Parses a table and returns the entries flattened as: u32 klen | key | u32 vlen | val ... into a malloc'd buffer.
DTFE_HOST_API int dtfe_sstable_read(const char* path, char** out, size_t* out_len, int* n) {
  std::string file;
  if (!read_file(path, &file)) return -1;
  std::vector<std::pair<std::string, std::string>> kv;
  if (!sstable_parse(file, &kv)) return -1;
  std::string flat;
  for (auto& e : kv) {
    put_fixed32(flat, (uint32_t)e.first.size());
    flat.append(e.first);
    put_fixed32(flat, (uint32_t)e.second.size());
    flat.append(e.second);
  }
  *out = static_cast<char*>(malloc(flat.size() ? flat.size() : 1));
  memcpy(*out, flat.data(), flat.size());
  *out_len = flat.size();
  *n = (int)kv.size();
  return 0;
}

DTFE_HOST_API void dtfe_host_free(void* p) { free(p); }
