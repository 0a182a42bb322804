// Pattern front end of the product: Rust-`regex` syntax subset -> dense DFA with `is_match`
// semantics (unanchored search unless ^/$ are written), mirroring Pattern::new / Pattern::matches
// (src/pattern.rs:21-45). The DFA runs on the host (confirmation, vgen_pattern_is_match) and on the
// device (match.cuh) over address strings. Also extracts the literal-prefix structure that
// lowering.cpp turns into digest interval compares.
#pragma once
#include <stdint.h>
#include <string>
#include <vector>
#include <bitset>
#include <memory>

namespace vg {

struct NfaProgram;
struct Dfa {
    uint32_t n_states = 0, n_classes = 0, start = 0;
    std::vector<uint16_t> table;        // [n_states][n_classes]
    std::vector<uint8_t> accept_end;    // [n_states]: 1 if the input may end here with a match
    uint8_t class_of[256];
    // false: the pattern needs more DFA states than a device table may have. The table is then a one-state "accept everything"
    // stand-in for the device, and match() below decides with the NFA (exact, slower): the host confirmation stays exact.
    bool table_complete = true;
    std::shared_ptr<const NfaProgram> nfa;
    bool match(const char *s, size_t n) const;
};

struct PatternInfo {
    bool anchored_start = false;                 // pattern is  ^ <sets...> <rest>
    std::vector<std::bitset<128>> prefix_sets;   // the leading single-character sets after ^
    bool rest_empty = false;                     // nothing after the sets  (=> prefix match)
    bool rest_is_end = false;                    // exactly `$` after the sets (=> exact match)
};

// Returns false and fills err on a syntax error (the cases Rust rejects). Everything `regex::Regex` accepts that can be decided on
// ASCII haystacks compiles: \b \B, \p{..} (through its ASCII members), nested classes and && -- ~~, flag x, non-ASCII literals
// (which match nothing). Patterns whose DFA would exceed 8,192 states come back with table_complete == false.
// fold_all: case-insensitive everywhere, inline (?-i) included: the language then contains the lower-cased form of every
// string the case-sensitive pattern matches (used for the Ethereum prefilter, lowering.cpp)
bool compile_pattern(const std::string &pattern, bool case_insensitive, Dfa &dfa, PatternInfo &info, std::string &err, bool fold_all = false);

}  // namespace vg
