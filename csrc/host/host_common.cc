// crc32c, protobuf wire primitives, file helpers, error text.
#include "host_common.h"

#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <unistd.h>

namespace dtfe_host {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ---- crc32c: slicing-by-8 tables generated at first use ------------------------------
static uint32_t g_tab[8][256];
static bool g_tab_ready = false;

static void init_tables() {
  const uint32_t poly = 0x82f63b78u;  // reflected Castagnoli polynomial
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ poly : (c >> 1);
    g_tab[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = g_tab[0][i];
    for (int t = 1; t < 8; ++t) {
      c = g_tab[0][c & 0xff] ^ (c >> 8);
      g_tab[t][i] = c;
    }
  }
  g_tab_ready = true;
}

uint32_t crc32c_extend(uint32_t crc, const void* data, size_t n) {
  if (!g_tab_ready) init_tables();
  const uint8_t* p = static_cast<const uint8_t*>(data);
  uint32_t c = crc ^ 0xffffffffu;
  while (n && (reinterpret_cast<uintptr_t>(p) & 7)) {
    c = g_tab[0][(c ^ *p++) & 0xff] ^ (c >> 8);
    --n;
  }
  while (n >= 8) {
    uint64_t w;
    memcpy(&w, p, 8);
    w ^= c;
    c = g_tab[7][w & 0xff] ^ g_tab[6][(w >> 8) & 0xff] ^ g_tab[5][(w >> 16) & 0xff] ^
        g_tab[4][(w >> 24) & 0xff] ^ g_tab[3][(w >> 32) & 0xff] ^ g_tab[2][(w >> 40) & 0xff] ^
        g_tab[1][(w >> 48) & 0xff] ^ g_tab[0][(w >> 56) & 0xff];
    p += 8;
    n -= 8;
  }
  while (n--) c = g_tab[0][(c ^ *p++) & 0xff] ^ (c >> 8);
  return c ^ 0xffffffffu;
}

// ---- protobuf wire ---------------------------------------------------------------------
void put_varint(std::string& s, uint64_t v) {
  while (v >= 0x80) {
    s.push_back(static_cast<char>((v & 0x7f) | 0x80));
    v >>= 7;
  }
  s.push_back(static_cast<char>(v));
}
void put_fixed32(std::string& s, uint32_t v) {
  char b[4];
  for (int i = 0; i < 4; ++i) b[i] = static_cast<char>((v >> (8 * i)) & 0xff);
  s.append(b, 4);
}
void put_fixed64(std::string& s, uint64_t v) {
  char b[8];
  for (int i = 0; i < 8; ++i) b[i] = static_cast<char>((v >> (8 * i)) & 0xff);
  s.append(b, 8);
}
void put_tag(std::string& s, int field, int wire_type) { put_varint(s, (uint64_t(field) << 3) | wire_type); }
void put_bytes_field(std::string& s, int field, const std::string& payload) {
  put_tag(s, field, 2);
  put_varint(s, payload.size());
  s.append(payload);
}
void put_varint_field(std::string& s, int field, uint64_t v) {
  put_tag(s, field, 0);
  put_varint(s, v);
}
bool get_varint(const uint8_t*& p, const uint8_t* end, uint64_t* out) {
  uint64_t v = 0;
  for (int shift = 0; shift < 64 && p < end; shift += 7) {
    uint8_t b = *p++;
    v |= uint64_t(b & 0x7f) << shift;
    if (!(b & 0x80)) {
      *out = v;
      return true;
    }
  }
  return false;
}

// ---- files ------------------------------------------------------------------------------
bool write_file_atomic(const std::string& path, const std::string& data) {
  std::string tmp = path + ".tmp" + std::to_string(getpid());
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) {
    set_error("cannot open %s for writing", tmp.c_str());
    return false;
  }
  bool ok = fwrite(data.data(), 1, data.size(), f) == data.size();
  ok = (fflush(f) == 0) && ok;
  ok = (fclose(f) == 0) && ok;
  if (!ok || rename(tmp.c_str(), path.c_str()) != 0) {
    set_error("write/rename of %s failed", path.c_str());
    remove(tmp.c_str());
    return false;
  }
  return true;
}

bool read_file(const std::string& path, std::string* out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) {
    set_error("cannot open %s", path.c_str());
    return false;
  }
  out->clear();
  char buf[1 << 16];
  size_t n;
  while ((n = fread(buf, 1, sizeof(buf), f)) > 0) out->append(buf, n);
  fclose(f);
  return true;
}

}  // namespace dtfe_host

DTFE_HOST_API const char* dtfe_host_last_error() { return dtfe_host::g_err; }
DTFE_HOST_API uint32_t dtfe_crc32c(const void* data, size_t n) { return dtfe_host::crc32c(data, n); }
DTFE_HOST_API uint32_t dtfe_crc32c_extend(uint32_t crc, const void* data, size_t n) {
  return dtfe_host::crc32c_extend(crc, data, n);
}
DTFE_HOST_API uint32_t dtfe_crc32c_masked(const void* data, size_t n) {
  return dtfe_host::crc32c_mask(dtfe_host::crc32c(data, n));
}
