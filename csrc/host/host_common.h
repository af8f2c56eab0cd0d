// Shared declarations of the GPU-less native runtime (libdtfe_host.so).
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <string>

#define DTFE_HOST_API extern "C" __attribute__((visibility("default")))

namespace dtfe_host {

void set_error(const char* fmt, ...);

// CRC-32C (Castagnoli), the checksum of TFRecord framing and of LevelDB/tensor-bundle blocks.
uint32_t crc32c_extend(uint32_t crc, const void* data, size_t n);
inline uint32_t crc32c(const void* data, size_t n) { return crc32c_extend(0, data, n); }
// TF/LevelDB "masked" crc: rotate right 15 and add a constant, so that a crc of a string that
// itself embeds crcs does not degenerate.
inline uint32_t crc32c_mask(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xa282ead8u; }
inline uint32_t crc32c_unmask(uint32_t m) { uint32_t r = m - 0xa282ead8u; return (r >> 17) | (r << 15); }

// protobuf wire helpers (just what Event / Summary / BundleHeader / BundleEntry need)
void put_varint(std::string& s, uint64_t v);
void put_fixed32(std::string& s, uint32_t v);
void put_fixed64(std::string& s, uint64_t v);
void put_tag(std::string& s, int field, int wire_type);
void put_bytes_field(std::string& s, int field, const std::string& payload);
void put_varint_field(std::string& s, int field, uint64_t v);
bool get_varint(const uint8_t*& p, const uint8_t* end, uint64_t* out);

bool write_file_atomic(const std::string& path, const std::string& data);
bool read_file(const std::string& path, std::string* out);

}  // namespace dtfe_host
