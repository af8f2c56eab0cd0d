// LevelDB-format sorted string table, the container of a TF1 V2 checkpoint's ".index" file
// (SURVEY Appendix B).  Writer: single pass over keys in ascending order, prefix-compressed
// entries with a restart point every 16, no compression, per-block masked crc32c, empty
// metaindex block, one index entry per data block, 48-byte footer.  Reader: full scan into
// (key, value) pairs with crc verification.
#include "sstable.h"

#include <string.h>

namespace dtfe_host {

static const uint64_t kTableMagic = 0xdb4775248b80fb57ull;
static const int kRestartInterval = 16;
static const size_t kFooterSize = 48;

// ---------------------------------------------------------------- block builder
class BlockBuilder {
 public:
  BlockBuilder() { restarts_.push_back(0); }
  void Add(const std::string& key, const std::string& value) {
    size_t shared = 0;
    if (counter_ < kRestartInterval) {
      size_t lim = std::min(last_key_.size(), key.size());
      while (shared < lim && last_key_[shared] == key[shared]) ++shared;
    } else {
      restarts_.push_back(static_cast<uint32_t>(buf_.size()));
      counter_ = 0;
    }
    put_varint(buf_, shared);
    put_varint(buf_, key.size() - shared);
    put_varint(buf_, value.size());
    buf_.append(key, shared, std::string::npos);
    buf_.append(value);
    last_key_ = key;
    ++counter_;
    ++entries_;
  }
  std::string Finish() {
    for (uint32_t r : restarts_) put_fixed32(buf_, r);
    put_fixed32(buf_, static_cast<uint32_t>(restarts_.size()));
    return buf_;
  }
  size_t CurrentSizeEstimate() const { return buf_.size() + restarts_.size() * 4 + 4; }
  bool empty() const { return entries_ == 0; }
  const std::string& last_key() const { return last_key_; }

 private:
  std::string buf_, last_key_;
  std::vector<uint32_t> restarts_;
  int counter_ = 0;
  size_t entries_ = 0;
};

static void EncodeHandle(std::string& dst, uint64_t offset, uint64_t size) {
  put_varint(dst, offset);
  put_varint(dst, size);
}

// Shortest string >= key: bump the first byte that is not 0xff and cut there (LevelDB's
// BytewiseComparator::FindShortSuccessor, used for the last index key).
static std::string ShortSuccessor(const std::string& key) {
  for (size_t i = 0; i < key.size(); ++i) {
    uint8_t b = static_cast<uint8_t>(key[i]);
    if (b != 0xff) {
      std::string out = key.substr(0, i + 1);
      out[i] = static_cast<char>(b + 1);
      return out;
    }
  }
  return key;
}

// Shortest separator in [start, limit) (FindShortestSeparator), used between two data blocks.
static std::string ShortestSeparator(const std::string& start, const std::string& limit) {
  size_t lim = std::min(start.size(), limit.size()), d = 0;
  while (d < lim && start[d] == limit[d]) ++d;
  if (d >= lim) return start;
  uint8_t b = static_cast<uint8_t>(start[d]);
  if (b < 0xff && b + 1 < static_cast<uint8_t>(limit[d])) {
    std::string out = start.substr(0, d + 1);
    out[d] = static_cast<char>(b + 1);
    return out;
  }
  return start;
}

static void WriteBlock(std::string& file, const std::string& contents, uint64_t* off, uint64_t* size) {
  *off = file.size();
  *size = contents.size();
  file.append(contents);
  char type = 0;  // kNoCompression
  uint32_t crc = crc32c_extend(crc32c(contents.data(), contents.size()), &type, 1);
  file.push_back(type);
  put_fixed32(file, crc32c_mask(crc));
}

bool sstable_build(const std::vector<std::pair<std::string, std::string>>& sorted_kv, size_t block_size,
                   std::string* out) {
  for (size_t i = 1; i < sorted_kv.size(); ++i) {
    if (!(sorted_kv[i - 1].first < sorted_kv[i].first)) {
      set_error("sstable_build: keys not strictly ascending at %zu", i);
      return false;
    }
  }
  std::string file;
  BlockBuilder index;
  BlockBuilder* data = new BlockBuilder();
  bool pending = false;
  std::string pending_last;
  uint64_t p_off = 0, p_size = 0;
  for (const auto& kv : sorted_kv) {
    if (pending) {
      std::string handle;
      EncodeHandle(handle, p_off, p_size);
      index.Add(ShortestSeparator(pending_last, kv.first), handle);
      pending = false;
    }
    data->Add(kv.first, kv.second);
    if (data->CurrentSizeEstimate() >= block_size) {
      pending_last = data->last_key();
      WriteBlock(file, data->Finish(), &p_off, &p_size);
      delete data;
      data = new BlockBuilder();
      pending = true;
    }
  }
  if (!data->empty()) {
    if (pending) {  // cannot happen (pending is resolved on the next Add) but keep the invariant obvious
      std::string handle;
      EncodeHandle(handle, p_off, p_size);
      index.Add(ShortSuccessor(pending_last), handle);
    }
    pending_last = data->last_key();
    WriteBlock(file, data->Finish(), &p_off, &p_size);
    pending = true;
  }
  delete data;
  if (pending) {
    std::string handle;
    EncodeHandle(handle, p_off, p_size);
    index.Add(ShortSuccessor(pending_last), handle);
  }
  uint64_t meta_off, meta_size, idx_off, idx_size;
  BlockBuilder meta;
  WriteBlock(file, meta.Finish(), &meta_off, &meta_size);
  WriteBlock(file, index.Finish(), &idx_off, &idx_size);
  std::string footer;
  EncodeHandle(footer, meta_off, meta_size);
  EncodeHandle(footer, idx_off, idx_size);
  footer.resize(kFooterSize - 8, '\0');
  put_fixed32(footer, static_cast<uint32_t>(kTableMagic & 0xffffffffu));
  put_fixed32(footer, static_cast<uint32_t>(kTableMagic >> 32));
  file.append(footer);
  out->swap(file);
  return true;
}

// ---------------------------------------------------------------- reader
static bool ReadBlock(const std::string& file, uint64_t off, uint64_t size, std::string* contents) {
  if (off + size + 5 > file.size()) {
    set_error("sstable: block handle (%llu,%llu) outside file of %zu bytes", (unsigned long long)off,
              (unsigned long long)size, file.size());
    return false;
  }
  const char* p = file.data() + off;
  if (p[size] != 0) {
    set_error("sstable: compressed blocks (type %d) are not supported", (int)p[size]);
    return false;
  }
  uint32_t stored;
  memcpy(&stored, p + size + 1, 4);
  uint32_t actual = crc32c_mask(crc32c(p, size + 1));
  if (stored != actual) {
    set_error("sstable: block checksum mismatch at offset %llu", (unsigned long long)off);
    return false;
  }
  contents->assign(p, size);
  return true;
}

static bool ParseBlock(const std::string& block, std::vector<std::pair<std::string, std::string>>* out) {
  if (block.size() < 4) {
    set_error("sstable: block too small");
    return false;
  }
  uint32_t num_restarts;
  memcpy(&num_restarts, block.data() + block.size() - 4, 4);
  if ((uint64_t)num_restarts * 4 + 4 > block.size()) {
    set_error("sstable: bad restart count");
    return false;
  }
  const uint8_t* p = reinterpret_cast<const uint8_t*>(block.data());
  const uint8_t* end = p + block.size() - 4 - (size_t)num_restarts * 4;
  std::string key;
  while (p < end) {
    uint64_t shared, non_shared, vlen;
    if (!get_varint(p, end, &shared) || !get_varint(p, end, &non_shared) || !get_varint(p, end, &vlen) ||
        shared > key.size() || p + non_shared + vlen > end) {
      set_error("sstable: corrupt entry");
      return false;
    }
    key.resize(shared);
    key.append(reinterpret_cast<const char*>(p), non_shared);
    p += non_shared;
    out->emplace_back(key, std::string(reinterpret_cast<const char*>(p), vlen));
    p += vlen;
  }
  return true;
}

bool sstable_parse(const std::string& file, std::vector<std::pair<std::string, std::string>>* out) {
  if (file.size() < kFooterSize) {
    set_error("sstable: file shorter than footer");
    return false;
  }
  const uint8_t* f = reinterpret_cast<const uint8_t*>(file.data()) + file.size() - kFooterSize;
  uint32_t lo, hi;
  memcpy(&lo, f + kFooterSize - 8, 4);
  memcpy(&hi, f + kFooterSize - 4, 4);
  if (((uint64_t)hi << 32 | lo) != kTableMagic) {
    set_error("sstable: bad magic");
    return false;
  }
  const uint8_t* p = f;
  const uint8_t* fend = f + kFooterSize - 8;
  uint64_t mo, ms, io, is;
  if (!get_varint(p, fend, &mo) || !get_varint(p, fend, &ms) || !get_varint(p, fend, &io) ||
      !get_varint(p, fend, &is)) {
    set_error("sstable: bad footer handles");
    return false;
  }
  std::string index_block;
  if (!ReadBlock(file, io, is, &index_block)) return false;
  std::vector<std::pair<std::string, std::string>> index;
  if (!ParseBlock(index_block, &index)) return false;
  for (const auto& ie : index) {
    const uint8_t* hp = reinterpret_cast<const uint8_t*>(ie.second.data());
    const uint8_t* hend = hp + ie.second.size();
    uint64_t off, size;
    if (!get_varint(hp, hend, &off) || !get_varint(hp, hend, &size)) {
      set_error("sstable: bad block handle in index");
      return false;
    }
    std::string block;
    if (!ReadBlock(file, off, size, &block)) return false;
    if (!ParseBlock(block, out)) return false;
  }
  return true;
}

}  // namespace dtfe_host
