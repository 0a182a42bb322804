// Pattern lint / difficulty heuristic / provider shim for the CLI (see pattern_lint.cpp).
#pragma once
#include <stdint.h>
#include <string>
#include "host_util.h"

namespace vg {
std::string validate_charset(const std::string &pattern, bool case_insensitive, int format);
size_t count_fixed_chars(const std::string &pattern);
uint64_t estimate_difficulty(const std::string &pattern, bool case_insensitive, int format);

struct ProviderResult { std::string address; int format = 0; bool has_range = false; U256 start, end; };
// 1 = resolved, 0 = not a provider reference (treat as regex), -1 = provider error (err set)
int provider_resolve(const std::string &pattern, ProviderResult &out, std::string &err);
std::string provider_build_pattern(const ProviderResult &r, size_t prefix_len);
std::string provider_build_exact_pattern(const ProviderResult &r);
}  // namespace vg
