// Rust-`regex` syntax -> AST -> Thompson NFA -> dense DFA (see regex_dfa.h).
//
// Haystacks are ASCII address strings, which is what makes the whole `regex::Regex` surface reducible to a byte DFA:
//   * Unicode classes (\p{L}, \pN, scripts ...) and case folding only matter through their ASCII members;
//   * a non-ASCII literal can never match (it compiles to the empty set);
//   * \b / \B (Unicode word boundary in Rust) see only ASCII word bytes [0-9A-Za-z_]: the DFA state carries one bit of
//     look-behind ("the previous byte was a word byte") and the assertion is resolved when the next byte is read.
// A pattern whose DFA exceeds the state limit is not an error either: the NFA is kept and the host confirms with it
// (Dfa::table_complete == false; lowering.cpp then lets the device report a superset).
#include "regex_dfa.h"
#include <map>
#include <set>
#include <memory>
#include <algorithm>
#include <tuple>
#include <string.h>

namespace vg {

// ---- Thompson NFA (kept inside Dfa for the host-side fallback matcher) ----
struct NfaProgram {
    struct State { enum Kind { Char, Split, Eps, Bol, Eol, WordB, NotWordB, Match } kind; std::bitset<256> set; int x = -1, y = -1; };
    std::vector<State> st;
    int entry = 0;
    bool has_word_boundary = false;
    int add(State::Kind k);
    struct Ctx { bool at_start, at_end, prev_word, next_word; };
    // follows every transition that consumes no input under `c`; out_chars = Char states reached, returns true if Match is reached
    bool closure(const std::vector<int> &seeds, const Ctx &c, std::vector<int> &out_chars, std::vector<char> &seen) const;
    bool match(const char *s, size_t n) const;
};
struct RegexTooBig { std::string msg; };
int NfaProgram::add(State::Kind k) {
    if (st.size() > 100000) throw RegexTooBig{"compiled regex exceeds size limit"};
    State s; s.kind = k; st.push_back(s);
    return (int)st.size() - 1;
}
static inline bool is_word_byte(int c) { return (c >= '0' && c <= '9') || (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c == '_'; }

bool NfaProgram::closure(const std::vector<int> &seeds, const Ctx &c, std::vector<int> &out_chars, std::vector<char> &seen) const {
    std::fill(seen.begin(), seen.end(), 0);
    std::vector<int> stack(seeds.begin(), seeds.end());
    bool match = false;
    while (!stack.empty()) {
        int s = stack.back(); stack.pop_back();
        if (s < 0 || seen[s]) continue;
        seen[s] = 1;
        const State &n = st[s];
        switch (n.kind) {
            case State::Char: out_chars.push_back(s); break;
            case State::Split: stack.push_back(n.y); stack.push_back(n.x); break;
            case State::Eps: stack.push_back(n.x); break;
            case State::Bol: if (c.at_start) stack.push_back(n.x); break;
            case State::Eol: if (c.at_end) stack.push_back(n.x); break;
            case State::WordB: if (c.prev_word != c.next_word) stack.push_back(n.x); break;
            case State::NotWordB: if (c.prev_word == c.next_word) stack.push_back(n.x); break;
            case State::Match: match = true; break;
        }
    }
    std::sort(out_chars.begin(), out_chars.end());
    return match;
}
// unanchored search (is_match) by NFA simulation: the exact matcher for patterns whose DFA is too large
bool NfaProgram::match(const char *s, size_t n) const {
    std::vector<char> seen(st.size());
    std::vector<int> seeds{entry}, chars;
    bool prev_word = false;
    for (size_t i = 0; i <= n; i++) {
        const bool end = i == n;
        const int c = end ? 0 : (unsigned char)s[i];
        Ctx ctx{i == 0, end, prev_word, !end && is_word_byte(c)};
        chars.clear();
        if (closure(seeds, ctx, chars, seen)) return true;
        if (end) break;
        seeds.clear();
        for (int cs : chars) if (st[cs].set[c]) seeds.push_back(st[cs].x);
        seeds.push_back(entry);                      // a new attempt starts at every position
        prev_word = is_word_byte(c);
    }
    return false;
}

bool Dfa::match(const char *s, size_t n) const {
    if (!table_complete) return nfa && nfa->match(s, n);
    uint32_t st = start;
    for (size_t i = 0; i < n; i++) st = table[(size_t)st * n_classes + class_of[(uint8_t)s[i]]];
    return accept_end[st] != 0;
}

namespace {

typedef std::bitset<256> CharSet;

struct Node {
    enum Kind { Empty, Set, Cat, Alt, Rep, Bol, Eol, WordB, NotWordB } kind = Empty;
    CharSet set;
    std::vector<std::unique_ptr<Node>> kids;
    int min = 0, max = 0;   // Rep; max < 0 = unbounded
};
typedef std::unique_ptr<Node> NodeP;

struct ParseError { std::string msg; };

class Parser {
public:
    Parser(const std::string &s, bool ci, bool fold_all = false) : s_(s), ci_(ci || fold_all), fold_all_(fold_all) {}
    NodeP parse() {
        NodeP n = alt(0);
        if (pos_ != s_.size()) fail("unopened group");
        return n;
    }
private:
    const std::string &s_;
    size_t pos_ = 0;
    bool ci_, dotall_ = false, fold_all_ = false, verbose_ = false;

    [[noreturn]] void fail(const char *m) { throw ParseError{m}; }
    bool more() const { return pos_ < s_.size(); }
    unsigned char peek() const { return (unsigned char)s_[pos_]; }
    // flag x: whitespace and #-comments are insignificant (also inside classes, as regex-syntax does)
    void skip_ws() {
        if (!verbose_) return;
        while (more()) {
            unsigned char c = peek();
            if (c == ' ' || (c >= 9 && c <= 13)) pos_++;
            else if (c == '#') { while (more() && peek() != '\n') pos_++; }
            else break;
        }
    }
    // one whole UTF-8 sequence starting at a byte >= 0x80: it can never occur in an ASCII haystack
    void skip_utf8() { pos_++; while (more() && (peek() & 0xC0) == 0x80) pos_++; }
    static void fold(CharSet &c) {
        for (int ch = 'a'; ch <= 'z'; ch++) {
            if (c[ch]) c.set(ch - 32);
            if (c[ch - 32]) c.set(ch);
        }
    }
    static void range(CharSet &c, int lo, int hi) { for (int ch = lo; ch <= hi; ch++) c.set(ch); }
    static void word(CharSet &c) { range(c, '0', '9'); range(c, 'a', 'z'); range(c, 'A', 'Z'); c.set('_'); }
    static void space(CharSet &c) { c.set(' '); range(c, 9, 13); }
    static int hexv(int ch) {
        if (ch >= '0' && ch <= '9') return ch - '0';
        if (ch >= 'a' && ch <= 'f') return ch - 'a' + 10;
        if (ch >= 'A' && ch <= 'F') return ch - 'A' + 10;
        return -1;
    }
    // \p{Name} / \pX restricted to its ASCII members (general categories, a few binary properties and scripts; names are matched
    // loosely like the regex crate does: case, spaces, '-' and '_' are ignored)
    static bool unicode_class(std::string name, CharSet &c) {
        std::string k;
        for (char ch : name) if (ch != ' ' && ch != '_' && ch != '-') k.push_back((char)tolower((unsigned char)ch));
        for (const char *pre : {"gc=", "generalcategory=", "sc=", "script=", "scx=", "scriptextensions="}) if (k.rfind(pre, 0) == 0) { k = k.substr(strlen(pre)); break; }
        if (k == "l" || k == "letter" || k == "alphabetic" || k == "alpha" || k == "latin" || k == "latn" || k == "lc" || k == "casedletter" || k == "cased") { range(c, 'a', 'z'); range(c, 'A', 'Z'); }
        else if (k == "lu" || k == "uppercaseletter" || k == "uppercase" || k == "upper") range(c, 'A', 'Z');
        else if (k == "ll" || k == "lowercaseletter" || k == "lowercase" || k == "lower") range(c, 'a', 'z');
        else if (k == "n" || k == "number" || k == "nd" || k == "decimalnumber" || k == "digit") range(c, '0', '9');
        else if (k == "alnum") { range(c, '0', '9'); range(c, 'a', 'z'); range(c, 'A', 'Z'); }
        else if (k == "xdigit" || k == "hexdigit" || k == "asciihexdigit" || k == "ahex") { range(c, '0', '9'); range(c, 'a', 'f'); range(c, 'A', 'F'); }
        else if (k == "p" || k == "punctuation" || k == "punct") { for (const char *q = "!\"#%&'()*,-./:;?@[\\]_{}"; *q; q++) c.set((unsigned char)*q); }
        else if (k == "s" || k == "symbol") { for (const char *q = "$+<=>^`|~"; *q; q++) c.set((unsigned char)*q); }
        else if (k == "z" || k == "separator" || k == "zs" || k == "spaceseparator") c.set(' ');
        else if (k == "whitespace" || k == "space" || k == "wspace") space(c);
        else if (k == "cc" || k == "control" || k == "cntrl" || k == "c" || k == "other") { range(c, 0, 31); c.set(127); }
        else if (k == "ascii") range(c, 0, 127);
        else if (k == "any") c.set();
        else if (k == "word" || k == "w") word(c);
        else {
            // scripts and categories without ASCII members: valid, empty over ASCII
            for (const char *e : {"greek", "grek", "cyrillic", "cyrl", "han", "hani", "arabic", "arab", "hebrew", "hebr", "hiragana", "hira", "katakana", "kana", "thai", "hangul",
                                  "hang", "devanagari", "deva", "armenian", "armn", "georgian", "geor", "lm", "lo", "lt", "m", "mark", "mn", "mc", "me", "nl", "no", "zl", "zp",
                                  "cf", "format", "co", "privateuse", "cs", "surrogate", "cn", "unassigned", "emoji"})
                if (k == e) return true;
            return false;
        }
        return true;
    }
    void unicode_escape(bool negate, CharSet &out) {       // pos_ is just past \p or \P
        if (!more()) fail("incomplete unicode class");
        std::string name;
        if (peek() == '{') {
            pos_++;
            bool neg2 = false;
            if (more() && peek() == '^') { neg2 = true; pos_++; }
            size_t e = s_.find('}', pos_);
            if (e == std::string::npos) fail("unclosed unicode class");
            name = s_.substr(pos_, e - pos_); pos_ = e + 1;
            negate = negate != neg2;
        } else { name = std::string(1, (char)peek()); pos_++; }
        CharSet t;
        if (!unicode_class(name, t)) fail("unknown unicode class");
        if (ci_) fold(t);
        out |= negate ? ~t : t;
    }
    // after a backslash: returns a literal byte (>= 0) or -1 after OR-ing a class into `out`
    int escape(CharSet &out, bool in_class) {
        if (!more()) fail("incomplete escape sequence");
        int ch = peek(); pos_++;
        CharSet t;
        switch (ch) {
            case 'd': range(out, '0', '9'); return -1;
            case 'D': range(t, '0', '9'); out |= ~t; return -1;
            case 'w': word(out); return -1;
            case 'W': word(t); out |= ~t; return -1;
            case 's': space(out); return -1;
            case 'S': space(t); out |= ~t; return -1;
            case 'n': return '\n'; case 't': return '\t'; case 'r': return '\r';
            case 'f': return '\f'; case 'v': return '\v'; case 'a': return 7; case '0': return 0;
            case 'x': case 'u': case 'U': {
                int v = 0, nd = 0;
                const int fixed = ch == 'x' ? 2 : ch == 'u' ? 4 : 8;
                if (more() && peek() == '{') {
                    pos_++;
                    while (more() && peek() != '}') { int h = hexv(peek()); pos_++; if (h < 0) fail("invalid hex escape"); v = v * 16 + h; nd++; if (v > 0x10FFFF) fail("invalid hex escape"); }
                    if (!more() || !nd) fail("invalid hex escape");
                    pos_++;
                } else {
                    for (; nd < fixed; nd++) { if (!more()) fail("invalid hex escape"); int h = hexv(peek()); pos_++; if (h < 0) fail("invalid hex escape"); v = v * 16 + h; }
                }
                return v > 0x7f ? -2 : v;           // -2: a non-ASCII code point, matches nothing in an ASCII haystack
            }
            case 'p': unicode_escape(false, out); return -1;
            case 'P': unicode_escape(true, out); return -1;
            case 'b': if (in_class) return 8; fail("word boundary in the wrong place");
            case 'B': fail("word boundary in the wrong place");
            default:
                if ((ch >= '0' && ch <= '9') || (ch >= 'a' && ch <= 'z') || (ch >= 'A' && ch <= 'Z')) fail("unrecognized escape sequence");
                if (ch >= 0x80) { pos_--; skip_utf8(); return -2; }
                return ch;
        }
    }
    static bool posix(const std::string &name, CharSet &c) {
        if (name == "alpha") { range(c, 'a', 'z'); range(c, 'A', 'Z'); }
        else if (name == "digit") range(c, '0', '9');
        else if (name == "alnum") { range(c, '0', '9'); range(c, 'a', 'z'); range(c, 'A', 'Z'); }
        else if (name == "upper") range(c, 'A', 'Z');
        else if (name == "lower") range(c, 'a', 'z');
        else if (name == "xdigit") { range(c, '0', '9'); range(c, 'a', 'f'); range(c, 'A', 'F'); }
        else if (name == "space") space(c);
        else if (name == "punct") { range(c, '!', '/'); range(c, ':', '@'); range(c, '[', '`'); range(c, '{', '~'); }
        else if (name == "word") word(c);
        else if (name == "blank") { c.set(' '); c.set('\t'); }
        else if (name == "cntrl") { range(c, 0, 31); c.set(127); }
        else if (name == "graph") range(c, '!', '~');
        else if (name == "print") range(c, ' ', '~');
        else if (name == "ascii") range(c, 0, 127);
        else return false;
        return true;
    }
    // One bracketed class, pos_ just past '['. Grammar of regex-syntax: items are united by juxtaposition; the binary operators
    // && (intersection), -- (difference) and ~~ (symmetric difference) have equal precedence, associate to the left and bind
    // looser than union; an item may itself be a bracketed class (nesting). Case folding is applied to every operand.
    CharSet klass(int depth) {
        if (depth > 50) fail("class nesting too deep");
        CharSet result, cur;
        bool neg = false, first = true, have_result = false;
        int pending = 0;                               // operator waiting for its right operand: 1 &&, 2 --, 3 ~~
        auto combine = [&]() {
            if (ci_) fold(cur);
            if (!have_result) { result = cur; have_result = true; }
            else if (pending == 1) result &= cur;
            else if (pending == 2) result &= ~cur;
            else if (pending == 3) result ^= cur;
            cur.reset();
        };
        if (more() && peek() == '^') { neg = true; pos_++; }
        for (;;) {
            skip_ws();
            if (!more()) fail("unclosed character class");
            int ch = peek();
            if (ch == ']' && !first) { pos_++; break; }
            if (!first && pos_ + 1 < s_.size() && ((ch == '&' && s_[pos_ + 1] == '&') || (ch == '-' && s_[pos_ + 1] == '-') || (ch == '~' && s_[pos_ + 1] == '~'))) {
                combine();
                pending = ch == '&' ? 1 : ch == '-' ? 2 : 3;
                pos_ += 2;
                continue;
            }
            first = false;
            if (ch == '[') {
                if (pos_ + 1 < s_.size() && s_[pos_ + 1] == ':') {
                    size_t q = pos_ + 2; bool pneg = false;
                    if (q < s_.size() && s_[q] == '^') { pneg = true; q++; }
                    size_t e = s_.find(":]", q);
                    if (e != std::string::npos) {
                        CharSet t;
                        if (!posix(s_.substr(q, e - q), t)) fail("unknown posix class");
                        cur |= pneg ? ~t : t; pos_ = e + 2; continue;
                    }
                }
                pos_++;
                cur |= klass(depth + 1);
                continue;
            }
            pos_++;
            int lo = ch;
            if (ch >= 0x80) { pos_--; skip_utf8(); lo = -2; }
            else if (ch == '\\') { CharSet t; lo = escape(t, true); if (lo == -1) { cur |= t; continue; } }
            // a range lo-hi (hi may be an escape); "-" before "]" or before a "--" operator is a literal
            if (pos_ + 1 < s_.size() && s_[pos_] == '-' && s_[pos_ + 1] != ']' && s_[pos_ + 1] != '-') {
                pos_++;
                int hi = peek(); pos_++;
                if (hi == '[') fail("invalid character class range");
                if (hi >= 0x80) { pos_--; skip_utf8(); hi = -2; }
                else if (hi == '\\') { CharSet t; hi = escape(t, true); if (hi == -1) fail("invalid range in class"); }
                if (lo == -2 && hi == -2) continue;                  // both ends beyond ASCII (order unchecked)
                if (lo == -2) fail("invalid character class range");  // a non-ASCII start cannot reach an ASCII end
                if (hi == -2) hi = 127;                               // a-<non-ASCII>: the ASCII part of the range
                else if (hi < lo) fail("invalid character class range");
                range(cur, lo, hi);
            } else if (lo >= 0) cur.set(lo);
        }
        combine();
        return neg ? ~result : result;
    }
    NodeP mk(Node::Kind k) { NodeP n(new Node); n->kind = k; return n; }
    NodeP atom(int depth) {
        int ch = peek();
        if (ch == '(') {
            pos_++;
            bool sci = ci_, sdot = dotall_, sx = verbose_;
            if (more() && peek() == '?') {
                pos_++;
                if (more() && (peek() == 'P' || peek() == '<')) {
                    if (peek() == 'P') pos_++;
                    if (!more() || peek() != '<') fail("invalid group syntax");
                    pos_++;
                    size_t q = pos_;
                    while (more() && peek() != '>') pos_++;
                    if (!more() || pos_ == q) fail("invalid group name");
                    pos_++;
                } else {
                    bool negf = false; int nf = 0;
                    for (;;) {
                        if (!more()) fail("unclosed group flags");
                        int f = peek(); pos_++;
                        if (f == ')') { if (!nf) fail("empty flags"); return mk(Node::Empty); }   // applies to the rest of the enclosing group
                        if (f == ':') break;
                        nf++;
                        if (f == '-') negf = true;
                        else if (f == 'i') ci_ = fold_all_ || !negf;
                        else if (f == 's') dotall_ = !negf;
                        else if (f == 'x') verbose_ = !negf;
                        else if (f == 'm' || f == 'U' || f == 'u' || f == 'R') {}
                        else fail("unsupported flag");
                    }
                }
            }
            NodeP n = alt(depth + 1);
            if (!more() || peek() != ')') fail("unclosed group");
            pos_++; ci_ = sci; dotall_ = sdot; verbose_ = sx;
            return n;
        }
        pos_++;
        NodeP n;
        switch (ch) {
            case '[': n = mk(Node::Set); n->set = klass(0); return n;
            case '.': n = mk(Node::Set); n->set.set(); if (!dotall_) n->set.reset('\n'); return n;
            case '^': return mk(Node::Bol);
            case '$': return mk(Node::Eol);
            case '\\': {
                if (more() && peek() == 'A') { pos_++; return mk(Node::Bol); }
                if (more() && peek() == 'z') { pos_++; return mk(Node::Eol); }
                if (more() && peek() == 'b') { pos_++; return mk(Node::WordB); }
                if (more() && peek() == 'B') { pos_++; return mk(Node::NotWordB); }
                n = mk(Node::Set);
                int lit = escape(n->set, false);
                if (lit >= 0) n->set.set(lit);
                if (ci_) fold(n->set);
                return n;
            }
            case '*': case '+': case '?': case '{': fail("repetition operator missing expression");
            default:
                n = mk(Node::Set);
                if (ch >= 0x80) { pos_--; skip_utf8(); return n; }       // a non-ASCII literal: the empty set
                n->set.set(ch); if (ci_) fold(n->set);
                return n;
        }
    }
    NodeP repeat(int depth) {
        NodeP n = atom(depth);
        for (;;) {
            skip_ws();
            if (!more()) break;
            int ch = peek(), mn, mx;
            if (ch == '*') { mn = 0; mx = -1; pos_++; }
            else if (ch == '+') { mn = 1; mx = -1; pos_++; }
            else if (ch == '?') { mn = 0; mx = 1; pos_++; }
            else if (ch == '{') {
                size_t q = pos_ + 1; long a = 0, b = 0; int na = 0, nb = 0; bool comma = false;
                auto ws = [&]() { while (verbose_ && q < s_.size() && (s_[q] == ' ' || (s_[q] >= 9 && s_[q] <= 13))) q++; };
                ws();
                while (q < s_.size() && isdigit((unsigned char)s_[q])) { a = a * 10 + (s_[q++] - '0'); na++; if (a > 100000) break; }
                ws();
                if (q < s_.size() && s_[q] == ',') { comma = true; q++; ws(); while (q < s_.size() && isdigit((unsigned char)s_[q])) { b = b * 10 + (s_[q++] - '0'); nb++; if (b > 100000) break; } ws(); }
                if (q >= s_.size() || s_[q] != '}' || !na) fail("invalid counted repetition");
                mn = (int)a; mx = comma ? (nb ? (int)b : -1) : (int)a;
                if (mx >= 0 && mx < mn) fail("invalid repetition range");
                if (mn > 1000 || mx > 1000) fail("repetition too large");
                pos_ = q + 1;
            } else break;
            if (more() && peek() == '?') pos_++;   // lazy form: same language
            if (n->kind == Node::Bol || n->kind == Node::Eol || n->kind == Node::WordB || n->kind == Node::NotWordB) {
                // a repeated assertion: x{0,..} is optional (= empty), x{1,..} is the assertion itself
                if (mn == 0) n = mk(Node::Empty);
                continue;
            }
            NodeP r = mk(Node::Rep); r->min = mn; r->max = mx; r->kids.push_back(std::move(n)); n = std::move(r);
        }
        return n;
    }
    NodeP cat(int depth) {
        NodeP n = mk(Node::Cat);
        for (;;) {
            skip_ws();
            if (!more() || peek() == '|' || peek() == ')') break;
            n->kids.push_back(repeat(depth));
        }
        if (more() && peek() == ')' && depth == 0) fail("unopened group");
        return n;
    }
    NodeP alt(int depth) {
        if (depth > 200) fail("nesting too deep");
        NodeP first = cat(depth);
        if (!more() || peek() != '|') return first;
        NodeP a = mk(Node::Alt);
        a->kids.push_back(std::move(first));
        while (more() && peek() == '|') { pos_++; a->kids.push_back(cat(depth)); }
        return a;
    }
};

// emit into nfa; returns entry state, and appends dangling exits (state idx, which field) to `outs`
struct Builder {
    NfaProgram &nfa;
    typedef NfaProgram::State NState;
    typedef std::pair<int, int> Out;   // (state, 0 = x / 1 = y)
    void patch(const std::vector<Out> &outs, int target) { for (auto &o : outs) (o.second ? nfa.st[o.first].y : nfa.st[o.first].x) = target; }
    int leaf(NState::Kind k, std::vector<Out> &outs) { int s = nfa.add(k); outs.push_back({s, 0}); return s; }
    int build(const Node *n, std::vector<Out> &outs) {
        switch (n->kind) {
            case Node::Empty: return leaf(NState::Eps, outs);
            case Node::Set: { int s = leaf(NState::Char, outs); nfa.st[s].set = n->set; return s; }
            case Node::Bol: return leaf(NState::Bol, outs);
            case Node::Eol: return leaf(NState::Eol, outs);
            case Node::WordB: nfa.has_word_boundary = true; return leaf(NState::WordB, outs);
            case Node::NotWordB: nfa.has_word_boundary = true; return leaf(NState::NotWordB, outs);
            case Node::Cat: {
                if (n->kids.empty()) return leaf(NState::Eps, outs);
                std::vector<Out> cur; int entry = build(n->kids[0].get(), cur);
                for (size_t i = 1; i < n->kids.size(); i++) { std::vector<Out> nx; int e = build(n->kids[i].get(), nx); patch(cur, e); cur.swap(nx); }
                outs.insert(outs.end(), cur.begin(), cur.end());
                return entry;
            }
            case Node::Alt: {
                int entry = -1, prev_split = -1;
                for (size_t i = 0; i < n->kids.size(); i++) {
                    if (i + 1 < n->kids.size()) {
                        int sp = nfa.add(NState::Split);
                        if (prev_split >= 0) nfa.st[prev_split].y = sp; else entry = sp;
                        int e = build(n->kids[i].get(), outs); nfa.st[sp].x = e; prev_split = sp;
                    } else {
                        int e = build(n->kids[i].get(), outs);
                        if (prev_split >= 0) nfa.st[prev_split].y = e; else entry = e;
                    }
                }
                return entry;
            }
            case Node::Rep: {
                const Node *k = n->kids[0].get();
                int entry = -1; std::vector<Out> cur;
                auto chain = [&](int e, std::vector<Out> &nx) { if (entry < 0) entry = e; else patch(cur, e); cur.swap(nx); };
                for (int i = 0; i < n->min; i++) { std::vector<Out> nx; int e = build(k, nx); chain(e, nx); }
                if (n->max < 0) {
                    int sp = nfa.add(NState::Split); std::vector<Out> body; int e = build(k, body);
                    nfa.st[sp].x = e; patch(body, sp);
                    std::vector<Out> nx{{sp, 1}}; chain(sp, nx);
                } else {
                    std::vector<Out> skips;
                    for (int i = n->min; i < n->max; i++) {
                        int sp = nfa.add(NState::Split); std::vector<Out> body; int e = build(k, body);
                        nfa.st[sp].x = e; skips.push_back({sp, 1});
                        chain(sp, body);
                    }
                    cur.insert(cur.end(), skips.begin(), skips.end());
                }
                if (entry < 0) { int s = nfa.add(NState::Eps); cur.push_back({s, 0}); entry = s; }
                outs.insert(outs.end(), cur.begin(), cur.end());
                return entry;
            }
        }
        return -1;
    }
};

void analyse(const Node *root, PatternInfo &info) {
    std::vector<const Node *> items;
    if (root->kind == Node::Cat) { for (auto &k : root->kids) if (k->kind != Node::Empty && !(k->kind == Node::Cat && k->kids.empty())) items.push_back(k.get()); }
    else items.push_back(root);
    if (items.empty() || items[0]->kind != Node::Bol) return;
    info.anchored_start = true;
    size_t i = 1;
    for (; i < items.size() && items[i]->kind == Node::Set; i++) {
        std::bitset<128> b;
        for (int c = 0; c < 128; c++) b[c] = items[i]->set[c];
        info.prefix_sets.push_back(b);
    }
    info.rest_empty = (i == items.size());
    info.rest_is_end = (i + 1 == items.size() && items[i]->kind == Node::Eol);
}

}  // namespace

bool compile_pattern(const std::string &pattern, bool ci, Dfa &dfa, PatternInfo &info, std::string &err, bool fold_all) {
    if (pattern.empty()) { err = "Pattern cannot be empty"; return false; }   // src/pattern.rs:22-24
    dfa = Dfa();
    try {
        Parser p(pattern, ci, fold_all);
        NodeP root = p.parse();
        info = PatternInfo();
        analyse(root.get(), info);
        auto prog = std::make_shared<NfaProgram>();
        NfaProgram &nfa = *prog;
        Builder b{nfa};
        std::vector<Builder::Out> outs;
        nfa.entry = b.build(root.get(), outs);
        int m = nfa.add(NfaProgram::State::Match);
        b.patch(outs, m);
        dfa.nfa = prog;
        // byte classes: bytes that no Char state (and, with \b, the word-byte test) tells apart share a column
        std::map<std::vector<bool>, int> sig2class;
        std::vector<int> char_states;
        for (size_t i = 0; i < nfa.st.size(); i++) if (nfa.st[i].kind == NfaProgram::State::Char) char_states.push_back((int)i);
        for (int c = 0; c < 256; c++) {
            std::vector<bool> sig(char_states.size() + 1);
            for (size_t k = 0; k < char_states.size(); k++) sig[k] = nfa.st[char_states[k]].set[c];
            sig[char_states.size()] = nfa.has_word_boundary && is_word_byte(c);
            auto it = sig2class.find(sig);
            if (it == sig2class.end()) it = sig2class.emplace(sig, (int)sig2class.size()).first;
            dfa.class_of[c] = (uint8_t)it->second;
        }
        if (sig2class.size() > 255) { err = "pattern too complex (byte classes)"; return false; }
        dfa.n_classes = (uint32_t)sig2class.size();
        std::vector<int> rep(dfa.n_classes, -1);
        for (int c = 0; c < 256; c++) if (rep[dfa.class_of[c]] < 0) rep[dfa.class_of[c]] = c;
        // Subset construction. A DFA state is (seed set, at_start, previous byte was a word byte); the zero-width assertions are
        // resolved per transition, when the next byte (or the end of the input) is known. State 0 = absorbing ACCEPT.
        struct Key { std::vector<int> seeds; bool at_start, prev_word; bool operator<(const Key &o) const { return std::tie(seeds, at_start, prev_word) < std::tie(o.seeds, o.at_start, o.prev_word); } };
        std::map<Key, uint32_t> ids;
        std::vector<Key> keys;
        std::vector<char> seen(nfa.st.size());
        dfa.table.assign(dfa.n_classes, 0); dfa.accept_end.assign(1, 1); keys.push_back(Key());   // ACCEPT loops to itself
        const uint32_t kMaxStates = 8192;
        bool overflow = false;
        // canonical form of a seed set: what is reachable without consuming input or passing an assertion (Char, assertion and
        // Match states); ^ can only hold in the start state, so it is dropped elsewhere
        auto frontier = [&](const std::vector<int> &seeds, bool at_start, bool &has_match) {
            std::vector<int> out, stack(seeds.begin(), seeds.end());
            std::fill(seen.begin(), seen.end(), 0);
            has_match = false;
            while (!stack.empty()) {
                int q = stack.back(); stack.pop_back();
                if (q < 0 || seen[q]) continue;
                seen[q] = 1;
                const NfaProgram::State &n = nfa.st[q];
                if (n.kind == NfaProgram::State::Split) { stack.push_back(n.y); stack.push_back(n.x); }
                else if (n.kind == NfaProgram::State::Eps) stack.push_back(n.x);
                else if (n.kind == NfaProgram::State::Match) has_match = true;
                else if (n.kind == NfaProgram::State::Bol && !at_start) {}
                else out.push_back(q);
            }
            std::sort(out.begin(), out.end());
            return out;
        };
        auto intern = [&](const std::vector<int> &raw, bool at_start, bool prev_word) -> uint32_t {
            bool has_match = false;
            std::vector<int> seeds = frontier(raw, at_start, has_match);
            if (has_match) return 0;                                     // matched already, whatever follows
            Key k{seeds, at_start, nfa.has_word_boundary && prev_word};
            auto it = ids.find(k);
            if (it != ids.end()) return it->second;
            if (keys.size() >= kMaxStates) { overflow = true; return 0; }
            uint32_t id = (uint32_t)keys.size();
            ids.emplace(k, id); keys.push_back(k);
            dfa.table.resize((size_t)(id + 1) * dfa.n_classes, 0);
            std::vector<int> chars;
            dfa.accept_end.push_back(nfa.closure(k.seeds, NfaProgram::Ctx{at_start, true, k.prev_word, false}, chars, seen) ? 1 : 0);
            return id;
        };
        dfa.start = intern({nfa.entry}, true, false);
        for (uint32_t s = 1; s < keys.size() && !overflow; s++) {
            for (uint32_t c = 0; c < dfa.n_classes && !overflow; c++) {
                const Key k = keys[s];                                   // copy: intern() may grow `keys`
                const bool next_word = is_word_byte(rep[c]);
                std::vector<int> chars;
                uint32_t t;
                if (nfa.closure(k.seeds, NfaProgram::Ctx{k.at_start, false, k.prev_word, next_word}, chars, seen)) t = 0;   // matched before this byte
                else {
                    std::vector<int> seeds;
                    for (int cs : chars) if (nfa.st[cs].set[rep[c]]) seeds.push_back(nfa.st[cs].x);
                    seeds.push_back(nfa.entry);                          // unanchored search: a new attempt starts at every position
                    t = intern(seeds, false, next_word);
                }
                dfa.table[(size_t)s * dfa.n_classes + c] = (uint16_t)t;
            }
        }
        if (overflow) {
            // too many states for a device table: keep the NFA as the exact (host) matcher; the device reports a superset
            dfa.table.assign(dfa.n_classes, 0); dfa.accept_end.assign(1, 1);
            dfa.n_states = 1; dfa.start = 0; dfa.table_complete = false;
            return true;
        }
        dfa.n_states = (uint32_t)keys.size();
        dfa.table_complete = true;
        return true;
    } catch (const ParseError &e) {
        err = e.msg;
        return false;
    } catch (const RegexTooBig &e) {
        err = e.msg;
        return false;
    }
}

}  // namespace vg
