// Rust-`regex` syntax subset -> AST -> Thompson NFA -> dense DFA (see regex_dfa.h).
#include "regex_dfa.h"
#include <map>
#include <set>
#include <memory>
#include <algorithm>
#include <string.h>

namespace vg {
namespace {

typedef std::bitset<256> CharSet;

struct Node {
    enum Kind { Empty, Set, Cat, Alt, Rep, Bol, Eol } kind = Empty;
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
    bool ci_, dotall_ = false, fold_all_ = false;

    [[noreturn]] void fail(const char *m) { throw ParseError{m}; }
    bool more() const { return pos_ < s_.size(); }
    unsigned char peek() const { return (unsigned char)s_[pos_]; }
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
            case 'x': {
                int v = 0, nd = 0;
                if (more() && peek() == '{') {
                    pos_++;
                    while (more() && peek() != '}') { int h = hexv(peek()); pos_++; if (h < 0) fail("invalid hex escape"); v = v * 16 + h; nd++; if (v > 0x10FFFF) fail("invalid hex escape"); }
                    if (!more() || !nd) fail("invalid hex escape");
                    pos_++;
                } else {
                    for (; nd < 2; nd++) { if (!more()) fail("invalid hex escape"); int h = hexv(peek()); pos_++; if (h < 0) fail("invalid hex escape"); v = v * 16 + h; }
                }
                if (v > 0x7f) fail("non-ASCII escape unsupported");
                return v;
            }
            case 'b': if (in_class) return 8; fail("word boundary unsupported");
            case 'B': fail("word boundary unsupported");
            case 'p': case 'P': fail("unicode class unsupported");
            default:
                if ((ch >= '0' && ch <= '9') || (ch >= 'a' && ch <= 'z') || (ch >= 'A' && ch <= 'Z')) fail("unrecognized escape sequence");
                if (ch >= 0x80) fail("non-ASCII escape unsupported");
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
    CharSet klass() {   // pos_ is just past '['
        CharSet set; bool neg = false, first = true;
        if (more() && peek() == '^') { neg = true; pos_++; }
        for (;;) {
            if (!more()) fail("unclosed character class");
            int ch = peek();
            if (ch == ']' && !first) { pos_++; break; }
            first = false;
            if (ch == '[') {
                if (pos_ + 1 < s_.size() && s_[pos_ + 1] == ':') {
                    size_t q = pos_ + 2; bool pneg = false;
                    if (q < s_.size() && s_[q] == '^') { pneg = true; q++; }
                    size_t e = s_.find(":]", q);
                    if (e == std::string::npos) fail("unclosed posix class");
                    CharSet t;
                    if (!posix(s_.substr(q, e - q), t)) fail("unknown posix class");
                    set |= pneg ? ~t : t; pos_ = e + 2; continue;
                }
                fail("nested character class unsupported");
            }
            if (ch >= 0x80) fail("non-ASCII in class unsupported");
            if (pos_ + 1 < s_.size() && ((ch == '&' && s_[pos_ + 1] == '&') || (ch == '~' && s_[pos_ + 1] == '~') || (ch == '-' && s_[pos_ + 1] == '-')))
                fail("class set operation unsupported");
            pos_++;
            int lo = ch;
            if (ch == '\\') { CharSet t; lo = escape(t, true); if (lo < 0) { set |= t; continue; } }
            if (pos_ + 1 < s_.size() && s_[pos_] == '-' && s_[pos_ + 1] != ']') {
                pos_++;
                int hi = peek(); pos_++;
                if (hi == '[') fail("nested character class unsupported");
                if (hi == '\\') { CharSet t; hi = escape(t, true); if (hi < 0) fail("invalid range in class"); }
                if (hi >= 0x80) fail("non-ASCII in class unsupported");
                if (hi < lo) fail("invalid character class range");
                range(set, lo, hi);
            } else set.set(lo);
        }
        if (ci_) fold(set);
        return neg ? ~set : set;
    }
    NodeP mk(Node::Kind k) { NodeP n(new Node); n->kind = k; return n; }
    NodeP atom(int depth) {
        int ch = peek();
        if (ch == '(') {
            pos_++;
            bool sci = ci_, sdot = dotall_;
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
                        else if (f == 'm' || f == 'U' || f == 'u' || f == 'R') {}
                        else fail("unsupported flag");
                    }
                }
            }
            NodeP n = alt(depth + 1);
            if (!more() || peek() != ')') fail("unclosed group");
            pos_++; ci_ = sci; dotall_ = sdot;
            return n;
        }
        pos_++;
        NodeP n;
        switch (ch) {
            case '[': n = mk(Node::Set); n->set = klass(); return n;
            case '.': n = mk(Node::Set); n->set.set(); if (!dotall_) n->set.reset('\n'); return n;
            case '^': return mk(Node::Bol);
            case '$': return mk(Node::Eol);
            case '\\': {
                if (more() && peek() == 'A') { pos_++; return mk(Node::Bol); }
                if (more() && peek() == 'z') { pos_++; return mk(Node::Eol); }
                n = mk(Node::Set);
                int lit = escape(n->set, false);
                if (lit >= 0) n->set.set(lit);
                if (ci_) fold(n->set);
                return n;
            }
            case '*': case '+': case '?': case '{': fail("repetition operator missing expression");
            default: n = mk(Node::Set); n->set.set(ch); if (ci_) fold(n->set); return n;
        }
    }
    NodeP repeat(int depth) {
        NodeP n = atom(depth);
        while (more()) {
            int ch = peek(), mn, mx;
            if (ch == '*') { mn = 0; mx = -1; pos_++; }
            else if (ch == '+') { mn = 1; mx = -1; pos_++; }
            else if (ch == '?') { mn = 0; mx = 1; pos_++; }
            else if (ch == '{') {
                size_t q = pos_ + 1; long a = 0, b = 0; int na = 0, nb = 0; bool comma = false;
                while (q < s_.size() && isdigit((unsigned char)s_[q])) { a = a * 10 + (s_[q++] - '0'); na++; if (a > 100000) break; }
                if (q < s_.size() && s_[q] == ',') { comma = true; q++; while (q < s_.size() && isdigit((unsigned char)s_[q])) { b = b * 10 + (s_[q++] - '0'); nb++; if (b > 100000) break; } }
                if (q >= s_.size() || s_[q] != '}' || !na) fail("invalid counted repetition");
                mn = (int)a; mx = comma ? (nb ? (int)b : -1) : (int)a;
                if (mx >= 0 && mx < mn) fail("invalid repetition range");
                if (mn > 1000 || mx > 1000) fail("repetition too large");
                pos_ = q + 1;
            } else break;
            if (more() && peek() == '?') pos_++;   // lazy form: same language
            NodeP r = mk(Node::Rep); r->min = mn; r->max = mx; r->kids.push_back(std::move(n)); n = std::move(r);
        }
        return n;
    }
    NodeP cat(int depth) {
        NodeP n = mk(Node::Cat);
        while (more() && peek() != '|' && peek() != ')') n->kids.push_back(repeat(depth));
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

// ---- Thompson NFA ----
struct NState { enum Kind { Char, Split, Eps, Bol, Eol, Match } kind; CharSet set; int x = -1, y = -1; };
struct Nfa {
    std::vector<NState> st;
    int add(NState::Kind k) { if (st.size() > 100000) throw ParseError{"compiled regex exceeds size limit"}; NState s; s.kind = k; st.push_back(s); return (int)st.size() - 1; }
};
struct Frag { int start; std::vector<int *> outs; };   // dangling out-pointers are patched later (indices kept stable by reserving)

// emit into nfa; returns entry state, and appends dangling exits (state idx, which field) to `outs`
struct Builder {
    Nfa &nfa;
    typedef std::pair<int, int> Out;   // (state, 0 = x / 1 = y)
    void patch(const std::vector<Out> &outs, int target) { for (auto &o : outs) (o.second ? nfa.st[o.first].y : nfa.st[o.first].x) = target; }
    int build(const Node *n, std::vector<Out> &outs) {
        switch (n->kind) {
            case Node::Empty: { int s = nfa.add(NState::Eps); outs.push_back({s, 0}); return s; }
            case Node::Set: { int s = nfa.add(NState::Char); nfa.st[s].set = n->set; outs.push_back({s, 0}); return s; }
            case Node::Bol: { int s = nfa.add(NState::Bol); outs.push_back({s, 0}); return s; }
            case Node::Eol: { int s = nfa.add(NState::Eol); outs.push_back({s, 0}); return s; }
            case Node::Cat: {
                if (n->kids.empty()) { int s = nfa.add(NState::Eps); outs.push_back({s, 0}); return s; }
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

struct Closure { std::vector<int> chars; bool match = false; };
void closure(const Nfa &nfa, const std::vector<int> &seeds, bool at_start, bool at_end, Closure &out) {
    std::vector<char> seen(nfa.st.size(), 0);
    std::vector<int> stack(seeds.begin(), seeds.end());
    while (!stack.empty()) {
        int s = stack.back(); stack.pop_back();
        if (s < 0 || seen[s]) continue;
        seen[s] = 1;
        const NState &n = nfa.st[s];
        switch (n.kind) {
            case NState::Char: out.chars.push_back(s); break;
            case NState::Split: stack.push_back(n.y); stack.push_back(n.x); break;
            case NState::Eps: stack.push_back(n.x); break;
            case NState::Bol: if (at_start) stack.push_back(n.x); break;
            case NState::Eol: if (at_end) stack.push_back(n.x); break;
            case NState::Match: out.match = true; break;
        }
    }
    std::sort(out.chars.begin(), out.chars.end());
}

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
    try {
        Parser p(pattern, ci, fold_all);
        NodeP root = p.parse();
        info = PatternInfo();
        analyse(root.get(), info);
        Nfa nfa; Builder b{nfa};
        std::vector<Builder::Out> outs;
        int entry = b.build(root.get(), outs);
        int m = nfa.add(NState::Match);
        b.patch(outs, m);
        // byte classes
        std::map<std::vector<bool>, int> sig2class;
        std::vector<int> char_states;
        for (size_t i = 0; i < nfa.st.size(); i++) if (nfa.st[i].kind == NState::Char) char_states.push_back((int)i);
        for (int c = 0; c < 256; c++) {
            std::vector<bool> sig(char_states.size());
            for (size_t k = 0; k < char_states.size(); k++) sig[k] = nfa.st[char_states[k]].set[c];
            auto it = sig2class.find(sig);
            if (it == sig2class.end()) it = sig2class.emplace(sig, (int)sig2class.size()).first;
            dfa.class_of[c] = (uint8_t)it->second;
        }
        if (sig2class.size() > 255) { err = "pattern too complex (byte classes)"; return false; }
        dfa.n_classes = (uint32_t)sig2class.size();
        std::vector<int> rep(dfa.n_classes, -1);
        for (int c = 0; c < 256; c++) if (rep[dfa.class_of[c]] < 0) rep[dfa.class_of[c]] = c;
        // subset construction; state key = (char-state set, accept_end); state 0 = absorbing ACCEPT
        typedef std::pair<std::vector<int>, bool> Key;
        std::map<Key, uint32_t> ids;
        std::vector<Key> keys;
        dfa.table.assign(dfa.n_classes, 0); dfa.accept_end.assign(1, 1); keys.push_back(Key());   // ACCEPT loops to itself
        const uint32_t kMaxStates = 8192;
        auto intern = [&](const std::vector<int> &seeds, bool at_start) -> uint32_t {
            Closure cn, ce;
            closure(nfa, seeds, at_start, false, cn);
            if (cn.match) return 0;
            closure(nfa, seeds, at_start, true, ce);
            Key k(cn.chars, ce.match);
            auto it = ids.find(k);
            if (it != ids.end()) return it->second;
            if (keys.size() >= kMaxStates) throw ParseError{"pattern too complex (DFA states)"};
            uint32_t id = (uint32_t)keys.size();
            ids.emplace(k, id); keys.push_back(k);
            dfa.table.resize((size_t)(id + 1) * dfa.n_classes, 0);
            dfa.accept_end.push_back(ce.match ? 1 : 0);
            return id;
        };
        dfa.start = intern({entry}, true);
        for (uint32_t s = 1; s < keys.size(); s++) {
            for (uint32_t c = 0; c < dfa.n_classes; c++) {
                std::vector<int> seeds;
                std::vector<int> cur = keys[s].first;   // copy: intern() may grow `keys`
                for (int cs : cur) if (nfa.st[cs].set[rep[c]]) seeds.push_back(nfa.st[cs].x);
                seeds.push_back(entry);                  // unanchored search: a new attempt starts at every position
                uint32_t t = intern(seeds, false);
                dfa.table[(size_t)s * dfa.n_classes + c] = (uint16_t)t;
            }
        }
        dfa.n_states = (uint32_t)keys.size();
        return true;
    } catch (const ParseError &e) {
        err = e.msg;
        return false;
    }
}

}  // namespace vg
