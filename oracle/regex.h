/* ORACLE (test infrastructure, not product): Rust-`regex` syntax subset as a Pike-VM (Thompson NFA)
 * matcher with `is_match` semantics: unanchored search on the whole string unless ^/$ are written
 * (src/pattern.rs:21-45; regex 1.12.2, Cargo.lock:1517). Haystacks are ASCII address strings.
 * Supported: literals, escapes (\d \w \s \D \W \S \n \t \r \f \v \xHH \x{H..} punctuation), `.`,
 * classes with ranges/negation/escapes/[:posix:], nested classes and the set operators && -- ~~, groups ( ) (?: ) (?P<n> )
 * (?<n> ), flags i s m U u x (inline and scoped), alternation, * + ? {m} {m,} {m,n} (+ lazy forms), ^ $ \A \z, \b \B
 * (ASCII word bytes: the haystacks are ASCII), \p{..} / \P{..} for the general categories with ASCII members
 * (L Lu Ll N Nd), non-ASCII literals (they match nothing in an ASCII haystack).
 * The engine is deliberately different from the product's DFA so that tests compare two designs. */
#ifndef VGEN_ORACLE_REGEX_H
#define VGEN_ORACLE_REGEX_H
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

enum { RN_EMPTY, RN_SET, RN_CAT, RN_ALT, RN_REP, RN_BOL, RN_EOL, RN_WORDB, RN_NWORDB };
typedef struct rnode { int type; uint64_t set[4]; struct rnode *a, *b; int min, max; } rnode;
enum { RI_SET, RI_SPLIT, RI_JMP, RI_BOL, RI_EOL, RI_MATCH, RI_WORDB, RI_NWORDB };
typedef struct { int op; uint64_t set[4]; int x, y; } rinst;
typedef struct { rinst *prog; int n, cap; char err[128]; } oregex;

typedef struct { const char *p, *end; int fi, fs, fx; char *err; rnode **pool; int npool, cappool; } rparser;

static inline void set_add(uint64_t s[4], int c) { s[(c >> 6) & 3] |= 1ULL << (c & 63); }
static inline int set_has(const uint64_t s[4], int c) { return (int)((s[(c >> 6) & 3] >> (c & 63)) & 1); }
static inline void set_range(uint64_t s[4], int lo, int hi) { for (int c = lo; c <= hi; c++) set_add(s, c); }
static inline void set_neg(uint64_t s[4]) { for (int i = 0; i < 4; i++) s[i] = ~s[i]; }
static inline void set_or(uint64_t s[4], const uint64_t t[4]) { for (int i = 0; i < 4; i++) s[i] |= t[i]; }
static inline int is_word_c(int c) { return (c >= '0' && c <= '9') || (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c == '_'; }
static inline void set_fold(uint64_t s[4]) {
    for (int c = 'a'; c <= 'z'; c++) {
        if (set_has(s, c)) set_add(s, c - 32);
        if (set_has(s, c - 32)) set_add(s, c);
    }
}
static rnode *rn_new(rparser *ps, int type) {
    rnode *n = (rnode *)calloc(1, sizeof(rnode)); n->type = type;
    if (ps->npool == ps->cappool) {
        ps->cappool = ps->cappool ? ps->cappool * 2 : 64;
        ps->pool = (rnode **)realloc(ps->pool, sizeof(rnode *) * ps->cappool);
    }
    ps->pool[ps->npool++] = n; return n;
}
static int rfail(rparser *ps, const char *msg) { if (!ps->err[0]) snprintf(ps->err, 128, "%s", msg); return -1; }
static int hexval(int c) {
    if (c >= '0' && c <= '9') return c - '0';
    if (c >= 'a' && c <= 'f') return c - 'a' + 10;
    if (c >= 'A' && c <= 'F') return c - 'A' + 10;
    return -1;
}
/* flag x: whitespace and # comments mean nothing */
static void skip_ws(rparser *ps) {
    if (!ps->fx) return;
    while (ps->p < ps->end) {
        int c = (unsigned char)*ps->p;
        if (c == ' ' || (c >= 9 && c <= 13)) ps->p++;
        else if (c == '#') { while (ps->p < ps->end && *ps->p != '\n') ps->p++; }
        else break;
    }
}
/* ps->p at the lead byte of a UTF-8 sequence: skip it (a non-ASCII character matches nothing in an ASCII haystack) */
static void skip_utf8(rparser *ps) { ps->p++; while (ps->p < ps->end && ((unsigned char)*ps->p & 0xC0) == 0x80) ps->p++; }
/* \p{Name} / \pX through its ASCII members; *lit = -2 on return means "a set was added" like the classes above */
static int parse_unicode_class(rparser *ps, uint64_t set[4], int negate) {
    char name[32]; int n = 0;
    if (ps->p >= ps->end) return rfail(ps, "incomplete unicode class");
    if (*ps->p == '{') {
        ps->p++;
        if (ps->p < ps->end && *ps->p == '^') { negate = !negate; ps->p++; }
        while (ps->p < ps->end && *ps->p != '}' && n < 31) name[n++] = *ps->p++;
        if (ps->p >= ps->end || *ps->p != '}') return rfail(ps, "unclosed unicode class");
        ps->p++;
    } else name[n++] = *ps->p++;
    name[n] = 0;
    uint64_t t[4] = {0, 0, 0, 0};
    if (!strcmp(name, "L") || !strcmp(name, "Letter") || !strcmp(name, "Alphabetic")) { set_range(t, 'a', 'z'); set_range(t, 'A', 'Z'); }
    else if (!strcmp(name, "Lu") || !strcmp(name, "Uppercase")) set_range(t, 'A', 'Z');
    else if (!strcmp(name, "Ll") || !strcmp(name, "Lowercase")) set_range(t, 'a', 'z');
    else if (!strcmp(name, "N") || !strcmp(name, "Nd") || !strcmp(name, "Number")) set_range(t, '0', '9');
    else if (!strcmp(name, "Greek") || !strcmp(name, "Cyrillic") || !strcmp(name, "Han")) {}
    else return rfail(ps, "unknown unicode class");
    if (ps->fi) set_fold(t);
    if (negate) set_neg(t);
    set_or(set, t);
    return 0;
}
/* parse an escape after '\\'; either a single byte (*lit>=0) or a class into set */
static int parse_escape(rparser *ps, uint64_t set[4], int *lit, int in_class) {
    if (ps->p >= ps->end) return rfail(ps, "incomplete escape sequence");
    int c = (unsigned char)*ps->p++; *lit = -1; uint64_t t[4] = {0, 0, 0, 0};
    switch (c) {
        case 'd': set_range(set, '0', '9'); return 0;
        case 'D': set_range(t, '0', '9'); set_neg(t); set_or(set, t); return 0;
        case 'w': set_range(set, '0', '9'); set_range(set, 'a', 'z'); set_range(set, 'A', 'Z'); set_add(set, '_'); return 0;
        case 'W': set_range(t, '0', '9'); set_range(t, 'a', 'z'); set_range(t, 'A', 'Z'); set_add(t, '_'); set_neg(t); set_or(set, t); return 0;
        case 's': set_add(set, ' '); set_range(set, 9, 13); return 0;
        case 'S': set_add(t, ' '); set_range(t, 9, 13); set_neg(t); set_or(set, t); return 0;
        case 'n': *lit = '\n'; return 0;
        case 't': *lit = '\t'; return 0;
        case 'r': *lit = '\r'; return 0;
        case 'f': *lit = '\f'; return 0;
        case 'v': *lit = '\v'; return 0;
        case 'a': *lit = 7; return 0;
        case '0': *lit = 0; return 0;
        case 'x': {
            int v = 0, nd = 0;
            if (ps->p < ps->end && *ps->p == '{') {
                ps->p++;
                while (ps->p < ps->end && *ps->p != '}') { int h = hexval(*ps->p++); if (h < 0) return rfail(ps, "invalid hex escape"); v = v * 16 + h; nd++; if (v > 0x10FFFF) return rfail(ps, "invalid hex escape"); }
                if (ps->p >= ps->end || !nd) return rfail(ps, "invalid hex escape");
                ps->p++;
            } else {
                for (; nd < 2; nd++) { if (ps->p >= ps->end) return rfail(ps, "invalid hex escape"); int h = hexval(*ps->p++); if (h < 0) return rfail(ps, "invalid hex escape"); v = v * 16 + h; }
            }
            if (v > 0x7F) return rfail(ps, "non-ASCII escape unsupported");
            *lit = v; return 0;
        }
        case 'b': case 'B': if (in_class && c == 'b') { *lit = 8; return 0; } return rfail(ps, "word boundary in the wrong place");
        case 'p': return parse_unicode_class(ps, set, 0);
        case 'P': return parse_unicode_class(ps, set, 1);
        default:
            if ((c >= '0' && c <= '9') || (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z')) return rfail(ps, "unrecognized escape sequence");
            if (c >= 0x80) return rfail(ps, "non-ASCII escape unsupported");
            *lit = c; return 0;
    }
}
static int posix_class(const char *name, size_t len, uint64_t set[4]) {
#define PC(s) (len == strlen(s) && !memcmp(name, s, len))
    if (PC("alpha")) { set_range(set, 'a', 'z'); set_range(set, 'A', 'Z'); }
    else if (PC("digit")) set_range(set, '0', '9');
    else if (PC("alnum")) { set_range(set, '0', '9'); set_range(set, 'a', 'z'); set_range(set, 'A', 'Z'); }
    else if (PC("upper")) set_range(set, 'A', 'Z');
    else if (PC("lower")) set_range(set, 'a', 'z');
    else if (PC("xdigit")) { set_range(set, '0', '9'); set_range(set, 'a', 'f'); set_range(set, 'A', 'F'); }
    else if (PC("space")) { set_add(set, ' '); set_range(set, 9, 13); }
    else if (PC("punct")) { set_range(set, '!', '/'); set_range(set, ':', '@'); set_range(set, '[', '`'); set_range(set, '{', '~'); }
    else if (PC("word")) { set_range(set, '0', '9'); set_range(set, 'a', 'z'); set_range(set, 'A', 'Z'); set_add(set, '_'); }
    else if (PC("blank")) { set_add(set, ' '); set_add(set, '\t'); }
    else if (PC("cntrl")) { set_range(set, 0, 31); set_add(set, 127); }
    else if (PC("graph")) set_range(set, '!', '~');
    else if (PC("print")) set_range(set, ' ', '~');
    else if (PC("ascii")) set_range(set, 0, 127);
    else return -1;
#undef PC
    return 0;
}
/* one bracketed class, ps->p just after '['. Items unite by juxtaposition; && -- ~~ (same precedence, left to right) bind looser;
 * an item may be a bracketed class itself; every operand is case-folded under flag i. */
static int parse_class(rparser *ps, uint64_t out[4], int depth) {
    uint64_t res[4] = {0, 0, 0, 0}, cur[4] = {0, 0, 0, 0};
    int neg = 0, first = 1, have = 0, pending = 0;
    if (depth > 50) return rfail(ps, "class nesting too deep");
    if (ps->p < ps->end && *ps->p == '^') { neg = 1; ps->p++; }
#define COMBINE() do { if (ps->fi) set_fold(cur); \
        if (!have) { memcpy(res, cur, 32); have = 1; } \
        else for (int q_ = 0; q_ < 4; q_++) res[q_] = pending == 1 ? (res[q_] & cur[q_]) : pending == 2 ? (res[q_] & ~cur[q_]) : (res[q_] ^ cur[q_]); \
        memset(cur, 0, 32); } while (0)
    for (;;) {
        skip_ws(ps);
        if (ps->p >= ps->end) return rfail(ps, "unclosed character class");
        int c = (unsigned char)*ps->p;
        if (c == ']' && !first) { ps->p++; break; }
        if (!first && ps->p + 1 < ps->end && ((c == '&' && ps->p[1] == '&') || (c == '-' && ps->p[1] == '-') || (c == '~' && ps->p[1] == '~'))) {
            COMBINE();
            pending = c == '&' ? 1 : c == '-' ? 2 : 3;
            ps->p += 2;
            continue;
        }
        first = 0;
        int lo = -1;
        if (c == '[') {
            if (ps->p + 1 < ps->end && ps->p[1] == ':') {
                const char *q = ps->p + 2; int pneg = 0;
                if (q < ps->end && *q == '^') { pneg = 1; q++; }
                const char *e = q; while (e + 1 < ps->end && !(e[0] == ':' && e[1] == ']')) e++;
                if (e + 1 < ps->end) {
                    uint64_t t[4] = {0, 0, 0, 0};
                    if (posix_class(q, (size_t)(e - q), t)) return rfail(ps, "unknown posix class");
                    if (pneg) set_neg(t);
                    set_or(cur, t); ps->p = e + 2; continue;
                }
            }
            ps->p++;
            uint64_t t[4];
            if (parse_class(ps, t, depth + 1)) return -1;
            set_or(cur, t);
            continue;
        }
        if (c >= 0x80) { skip_utf8(ps); lo = -2; }
        else {
            ps->p++;
            if (c == '\\') {
                uint64_t t[4] = {0, 0, 0, 0};
                if (parse_escape(ps, t, &lo, 1)) return -1;
                if (lo < 0) { set_or(cur, t); continue; }
            } else lo = c;
        }
        /* range? ("-" before "]" or before a "--" operator is a literal) */
        if (ps->p + 1 < ps->end && ps->p[0] == '-' && ps->p[1] != ']' && ps->p[1] != '-') {
            ps->p++;
            int hc = (unsigned char)*ps->p, hi = hc;
            if (hc == '[') return rfail(ps, "invalid character class range");
            if (hc >= 0x80) { skip_utf8(ps); hi = -2; }
            else {
                ps->p++;
                if (hc == '\\') {
                    uint64_t t[4] = {0, 0, 0, 0};
                    if (parse_escape(ps, t, &hi, 1)) return -1;
                    if (hi < 0) return rfail(ps, "invalid range in class");
                }
            }
            if (lo == -2 && hi == -2) continue;
            if (lo == -2) return rfail(ps, "invalid character class range");
            if (hi == -2) hi = 127;
            else if (hi < lo) return rfail(ps, "invalid character class range");
            set_range(cur, lo, hi);
        } else if (lo >= 0) set_add(cur, lo);
    }
    COMBINE();
#undef COMBINE
    if (neg) set_neg(res);
    memcpy(out, res, sizeof res); return 0;
}
static rnode *parse_alt(rparser *ps, int depth);
static rnode *parse_atom(rparser *ps, int depth) {
    int c = (unsigned char)*ps->p;
    rnode *n;
    if (c == '(') {
        ps->p++;
        int sfi = ps->fi, sfs = ps->fs, sfx = ps->fx;
        if (ps->p < ps->end && *ps->p == '?') {
            ps->p++;
            if (ps->p < ps->end && (*ps->p == 'P' || *ps->p == '<')) {
                if (*ps->p == 'P') ps->p++;
                if (ps->p >= ps->end || *ps->p != '<') { rfail(ps, "invalid group syntax"); return NULL; }
                ps->p++;
                const char *q = ps->p;
                while (ps->p < ps->end && *ps->p != '>') ps->p++;
                if (ps->p >= ps->end || ps->p == q) { rfail(ps, "invalid group name"); return NULL; }
                ps->p++;
            } else {
                int negf = 0, nf = 0;
                for (;;) {
                    if (ps->p >= ps->end) { rfail(ps, "unclosed group flags"); return NULL; }
                    int f = *ps->p++;
                    if (f == ':' || f == ')') {
                        if (f == ')') {           /* (?flags) applies to the rest of the enclosing group */
                            if (!nf) { rfail(ps, "empty flags"); return NULL; }
                            return rn_new(ps, RN_EMPTY);
                        }
                        break;
                    }
                    nf++;
                    if (f == '-') negf = 1;
                    else if (f == 'i') ps->fi = !negf;
                    else if (f == 's') ps->fs = !negf;
                    else if (f == 'x') ps->fx = !negf;
                    else if (f == 'm' || f == 'U' || f == 'u' || f == 'R') {}
                    else { rfail(ps, "unsupported flag"); return NULL; }
                }
            }
        }
        n = parse_alt(ps, depth + 1);
        if (!n) return NULL;
        if (ps->p >= ps->end || *ps->p != ')') { rfail(ps, "unclosed group"); return NULL; }
        ps->p++; ps->fi = sfi; ps->fs = sfs; ps->fx = sfx;
        return n;
    }
    ps->p++;
    if (c == '[') { n = rn_new(ps, RN_SET); if (parse_class(ps, n->set, 0)) return NULL; return n; }
    if (c == '.') {
        n = rn_new(ps, RN_SET); set_range(n->set, 0, 255);
        if (!ps->fs) n->set[0] &= ~(1ULL << '\n');
        return n;
    }
    if (c == '^') return rn_new(ps, RN_BOL);
    if (c == '$') return rn_new(ps, RN_EOL);
    if (c == '\\') {
        if (ps->p < ps->end && *ps->p == 'A') { ps->p++; return rn_new(ps, RN_BOL); }
        if (ps->p < ps->end && *ps->p == 'z') { ps->p++; return rn_new(ps, RN_EOL); }
        if (ps->p < ps->end && *ps->p == 'b') { ps->p++; return rn_new(ps, RN_WORDB); }
        if (ps->p < ps->end && *ps->p == 'B') { ps->p++; return rn_new(ps, RN_NWORDB); }
        n = rn_new(ps, RN_SET); int lit;
        if (parse_escape(ps, n->set, &lit, 0)) return NULL;
        if (lit >= 0) set_add(n->set, lit);
        if (ps->fi) set_fold(n->set);
        return n;
    }
    if (c == '*' || c == '+' || c == '?' || c == '{') { rfail(ps, "repetition operator missing expression"); return NULL; }
    n = rn_new(ps, RN_SET);
    if (c >= 0x80) { ps->p--; skip_utf8(ps); return n; }      /* non-ASCII literal: the empty set */
    set_add(n->set, c); if (ps->fi) set_fold(n->set);
    return n;
}
static rnode *parse_repeat(rparser *ps, int depth) {
    rnode *n = parse_atom(ps, depth);
    if (!n) return NULL;
    for (;;) {
        skip_ws(ps);
        if (ps->p >= ps->end) break;
        int c = *ps->p, mn, mx;
        if (c == '*') { mn = 0; mx = -1; ps->p++; }
        else if (c == '+') { mn = 1; mx = -1; ps->p++; }
        else if (c == '?') { mn = 0; mx = 1; ps->p++; }
        else if (c == '{') {
            const char *q = ps->p + 1; long a = 0, b = 0; int na = 0, nb = 0, comma = 0;
            while (q < ps->end && *q >= '0' && *q <= '9') { a = a * 10 + (*q++ - '0'); na++; if (a > 100000) break; }
            if (q < ps->end && *q == ',') { comma = 1; q++; while (q < ps->end && *q >= '0' && *q <= '9') { b = b * 10 + (*q++ - '0'); nb++; if (b > 100000) break; } }
            if (q >= ps->end || *q != '}' || !na) { rfail(ps, "invalid counted repetition"); return NULL; }
            mn = (int)a; mx = comma ? (nb ? (int)b : -1) : (int)a;
            if (mx >= 0 && mx < mn) { rfail(ps, "invalid repetition range"); return NULL; }
            if (mn > 1000 || mx > 1000) { rfail(ps, "repetition too large"); return NULL; }
            ps->p = q + 1;
        } else break;
        if (ps->p < ps->end && *ps->p == '?') ps->p++;      /* lazy: same language */
        if (n->type == RN_BOL || n->type == RN_EOL || n->type == RN_WORDB || n->type == RN_NWORDB) {
            if (mn == 0) n = rn_new(ps, RN_EMPTY);             /* a repeated assertion: optional = empty, else itself */
            continue;
        }
        rnode *r = rn_new(ps, RN_REP); r->a = n; r->min = mn; r->max = mx; n = r;
    }
    return n;
}
static rnode *parse_cat(rparser *ps, int depth) {
    rnode *n = NULL;
    for (;;) {
        skip_ws(ps);
        if (ps->p >= ps->end || *ps->p == '|' || *ps->p == ')') break;
        rnode *r = parse_repeat(ps, depth);
        if (!r) return NULL;
        if (!n) n = r; else { rnode *c = rn_new(ps, RN_CAT); c->a = n; c->b = r; n = c; }
    }
    if (ps->p < ps->end && *ps->p == ')' && depth == 0) { rfail(ps, "unopened group"); return NULL; }
    return n ? n : rn_new(ps, RN_EMPTY);
}
static rnode *parse_alt(rparser *ps, int depth) {
    if (depth > 200) { rfail(ps, "nesting too deep"); return NULL; }
    int sfi = ps->fi, sfs = ps->fs;
    rnode *n = parse_cat(ps, depth);
    if (!n) return NULL;
    while (ps->p < ps->end && *ps->p == '|') {
        ps->p++;
        rnode *r = parse_cat(ps, depth);
        if (!r) return NULL;
        rnode *a = rn_new(ps, RN_ALT); a->a = n; a->b = r; n = a;
    }
    (void)sfi; (void)sfs;
    return n;
}
static int remit(oregex *re, int op) {
    if (re->n >= 200000) return -1;
    if (re->n == re->cap) { re->cap = re->cap ? re->cap * 2 : 64; re->prog = (rinst *)realloc(re->prog, sizeof(rinst) * re->cap); }
    memset(&re->prog[re->n], 0, sizeof(rinst)); re->prog[re->n].op = op; return re->n++;
}
static int rcompile(oregex *re, const rnode *n) {
    int i, j, k;
    switch (n->type) {
        case RN_EMPTY: return 0;
        case RN_SET: if ((i = remit(re, RI_SET)) < 0) return -1; memcpy(re->prog[i].set, n->set, 32); return 0;
        case RN_BOL: return remit(re, RI_BOL) < 0 ? -1 : 0;
        case RN_EOL: return remit(re, RI_EOL) < 0 ? -1 : 0;
        case RN_WORDB: return remit(re, RI_WORDB) < 0 ? -1 : 0;
        case RN_NWORDB: return remit(re, RI_NWORDB) < 0 ? -1 : 0;
        case RN_CAT: if (rcompile(re, n->a)) return -1; return rcompile(re, n->b);
        case RN_ALT:
            if ((i = remit(re, RI_SPLIT)) < 0) return -1;
            re->prog[i].x = re->n; if (rcompile(re, n->a)) return -1;
            if ((j = remit(re, RI_JMP)) < 0) return -1;
            re->prog[i].y = re->n; if (rcompile(re, n->b)) return -1;
            re->prog[j].x = re->n; return 0;
        case RN_REP:
            for (k = 0; k < n->min; k++) if (rcompile(re, n->a)) return -1;
            if (n->max < 0) {
                if ((i = remit(re, RI_SPLIT)) < 0) return -1;
                re->prog[i].x = re->n; if (rcompile(re, n->a)) return -1;
                if ((j = remit(re, RI_JMP)) < 0) return -1;
                re->prog[j].x = i; re->prog[i].y = re->n; return 0;
            }
            {
                int nopt = n->max - n->min; int *splits = (int *)malloc(sizeof(int) * (nopt + 1));
                for (k = 0; k < nopt; k++) {
                    if ((splits[k] = remit(re, RI_SPLIT)) < 0) { free(splits); return -1; }
                    re->prog[splits[k]].x = re->n;
                    if (rcompile(re, n->a)) { free(splits); return -1; }
                }
                for (k = 0; k < nopt; k++) re->prog[splits[k]].y = re->n;
                free(splits); return 0;
            }
    }
    return -1;
}
static int oregex_compile(oregex *re, const char *pattern, int ci) {
    memset(re, 0, sizeof *re);
    rparser ps; memset(&ps, 0, sizeof ps);
    ps.p = pattern; ps.end = pattern + strlen(pattern); ps.fi = ci; ps.err = re->err;
    rnode *root = parse_alt(&ps, 0);
    int rc = 0;
    if (!root) rc = -1;
    else if (ps.p != ps.end) { rfail(&ps, "unopened group"); rc = -1; }
    else if (rcompile(re, root) || remit(re, RI_MATCH) < 0) { rfail(&ps, "compiled regex exceeds size limit"); rc = -1; }
    for (int i = 0; i < ps.npool; i++) free(ps.pool[i]);
    free(ps.pool);
    if (rc) { free(re->prog); re->prog = NULL; re->n = 0; }
    return rc;
}
static void oregex_free(oregex *re) { free(re->prog); re->prog = NULL; }

typedef struct { int *pc; int n; } rlist;
static int raddthread(const oregex *re, rlist *l, int *mark, int gen, int pc, int at_start, int at_end, int prev_word, int next_word) {
    /* returns 1 if MATCH reachable */
    int stack[64], sp = 0, matched = 0; int *big = NULL, bigcap = 0;
    (void)big; (void)bigcap;
    /* iterative DFS with an explicit heap stack */
    int cap = 256, *st = (int *)malloc(sizeof(int) * cap); (void)stack;
    st[sp++] = pc;
    while (sp) {
        int p = st[--sp];
        if (mark[p] == gen) continue;
        mark[p] = gen;
        const rinst *in = &re->prog[p];
        if (sp + 2 >= cap) { cap *= 2; st = (int *)realloc(st, sizeof(int) * cap); }
        switch (in->op) {
            case RI_JMP: st[sp++] = in->x; break;
            case RI_SPLIT: st[sp++] = in->y; st[sp++] = in->x; break;
            case RI_BOL: if (at_start) st[sp++] = p + 1; break;
            case RI_EOL: if (at_end) st[sp++] = p + 1; break;
            case RI_WORDB: if (prev_word != next_word) st[sp++] = p + 1; break;
            case RI_NWORDB: if (prev_word == next_word) st[sp++] = p + 1; break;
            case RI_MATCH: matched = 1; break;
            default: l->pc[l->n++] = p;
        }
    }
    free(st);
    return matched;
}
static int oregex_match(const oregex *re, const char *s, size_t len) {
    if (!re->prog) return 0;
    int n = re->n, gen = 1, res = 0;
    int *mark = (int *)calloc(n, sizeof(int));
    rlist cur = {(int *)malloc(sizeof(int) * n), 0}, nxt = {(int *)malloc(sizeof(int) * n), 0};
    for (size_t i = 0; i <= len; i++) {
        /* unanchored search: a new thread starts at every position */
        int pw = i > 0 && is_word_c((unsigned char)s[i - 1]), nw = i < len && is_word_c((unsigned char)s[i]);
        if (raddthread(re, &cur, mark, gen, 0, i == 0, i == len, pw, nw)) { res = 1; break; }
        if (i == len) break;
        gen++;
        int c = (unsigned char)s[i]; nxt.n = 0;
        for (int k = 0; k < cur.n; k++) {
            const rinst *in = &re->prog[cur.pc[k]];
            if (set_has(in->set, c) && raddthread(re, &nxt, mark, gen, cur.pc[k] + 1, 0, i + 1 == len, is_word_c(c), i + 1 < len && is_word_c((unsigned char)s[i + 1]))) { res = 1; break; }
        }
        if (res) break;
        rlist t = cur; cur = nxt; nxt = t;
        /* the next iteration adds the fresh start thread into cur using the same gen marks */
    }
    free(mark); free(cur.pc); free(nxt.pc);
    return res;
}
#endif
