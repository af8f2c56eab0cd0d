#pragma once
#include "host_common.h"

#include <algorithm>
#include <string>
#include <utility>
#include <vector>

namespace dtfe_host {
bool sstable_build(const std::vector<std::pair<std::string, std::string>>& sorted_kv, size_t block_size,
                   std::string* out);
bool sstable_parse(const std::string& file, std::vector<std::pair<std::string, std::string>>* out);
}  // namespace dtfe_host
