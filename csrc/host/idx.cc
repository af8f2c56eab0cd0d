// idx.gz reader: big-endian magic 0x00000803 (images, 3 dims) or
// 0x00000801 (labels, 1 dim), then dims, then u8 payload, gzip-wrapped.  Replaces the file
// parsing inside TF1's input_data.read_data_sets (reference main.py:43-44).
#include "host_common.h"

#include <stdlib.h>
#include <string.h>
#include <zlib.h>

using namespace dtfe_host;

static bool read_be32(gzFile f, uint32_t* v) {
  unsigned char b[4];
  if (gzread(f, b, 4) != 4) return false;
  *v = (uint32_t(b[0]) << 24) | (uint32_t(b[1]) << 16) | (uint32_t(b[2]) << 8) | uint32_t(b[3]);
  return true;
}

// dims[0..ndim) receive the shape; *data is malloc'd (free with dtfe_host_free).
DTFE_HOST_API int dtfe_idx_read(const char* path, int* ndim, int64_t* dims4, unsigned char** data, size_t* nbytes) {
  gzFile f = gzopen(path, "rb");
  if (!f) {
    set_error("cannot open %s", path);
    return -1;
  }
  gzbuffer(f, 1 << 20);
  uint32_t magic;
  if (!read_be32(f, &magic)) {
    gzclose(f);
    set_error("%s: truncated header", path);
    return -1;
  }
  int nd;
  if (magic == 0x00000803u) nd = 3;
  else if (magic == 0x00000801u) nd = 1;
  else {
    gzclose(f);
    set_error("%s: bad idx magic 0x%08x", path, magic);
    return -1;
  }
  size_t total = 1;
  for (int i = 0; i < nd; ++i) {
    uint32_t d;
    if (!read_be32(f, &d)) {
      gzclose(f);
      set_error("%s: truncated dims", path);
      return -1;
    }
    dims4[i] = d;
    total *= d;
  }
  unsigned char* buf = static_cast<unsigned char*>(malloc(total ? total : 1));
  size_t got = 0;
  while (got < total) {
    unsigned chunk = static_cast<unsigned>(std::min<size_t>(total - got, 1u << 30));
    int r = gzread(f, buf + got, chunk);
    if (r <= 0) break;
    got += static_cast<size_t>(r);
  }
  unsigned char extra;
  bool trailing = gzread(f, &extra, 1) == 1;
  gzclose(f);
  if (got != total || trailing) {
    free(buf);
    set_error("%s: payload is %s than the %zu bytes the header promises", path, trailing ? "longer" : "shorter", total);
    return -1;
  }
  *ndim = nd;
  *data = buf;
  *nbytes = total;
  return 0;
}

// u8 -> float32 in [0,1] (multiply by 1/255 like TF1's DataSet), optionally gathering rows.
DTFE_HOST_API void dtfe_u8_to_unit_float(const unsigned char* src, float* dst, size_t n) {
  const float s = 1.0f / 255.0f;
  for (size_t i = 0; i < n; ++i) dst[i] = static_cast<float>(src[i]) * s;
}
