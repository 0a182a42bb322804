// Result of lowering a (format, pattern) pair; see lowering.cpp.
#pragma once
#include <string>
#include "host_util.h"
#include "regex_dfa.h"
#include "../scan_types.h"

namespace vg {
struct LoweredPattern {
    int format = 0;
    std::string pattern;
    bool case_insensitive = false;
    Dfa dfa;                 // exact matcher on the address string (host confirm + device DFA path)
    PatternInfo info;
    vg_matcher_dev dev;      // device matcher (DFA table pointers are filled by the engine per device)
    double density = 1.0;    // upper estimate of the fraction of keys the DEVICE matcher reports (1 = unknown / may be dense)
    std::string kind_name;   // "intervals" | "exact" | "dfa" | "all" | "none"
    bool host_regex = false; // the DFA is too large for the device: it reports a superset (all keys / prefix intervals), the host decides with the NFA
};
// shared-memory bytes the walk kernel needs to stage this DFA (table + accept + symbol classes), 0 if it is too large
uint32_t dfa_smem_bytes(const vg_dfa_dev &d);
bool lower_pattern(int format, const std::string &pattern, bool case_insensitive, LoweredPattern &out, std::string &err);
}
