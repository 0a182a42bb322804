#!/usr/bin/env python3
"""Generates tests/golden/addresses.json and tests/golden/regex_cases.json.

Independent of oracle/ and of the product: the EC point comes from Python `cryptography`
(OpenSSL) and is cross-checked against pure-Python integer arithmetic; SHA-256 / RIPEMD-160 come
from hashlib; Keccak-256, Base58Check, Bech32(m), EIP-55 and BIP341 are written out here from the
standards. The script first reproduces every known answer the reference ships
(src/address.rs:233-238, src/provider.rs:76-87, README.md:121-127, README.md:107-108), which
validates it, then emits vectors for all six formats of src/address.rs:92-151.

Regex cases are evaluated with Python `re` (same semantics as Rust `regex` for this ASCII subset;
src/pattern.rs:321-350 cases are included verbatim).

Run here (container), commit the JSON; the GPU box only reads the JSON.
"""
import hashlib
import json
import os
import random
import re

from cryptography.hazmat.primitives.asymmetric import ec

P = 2**256 - 2**32 - 977
N = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141
GX = 0x79BE667EF9DCBBAC55A06295CE870B07029BFCDB2DCE28D959F2815B16F81798
GY = 0x483ADA7726A3C4655DA4FBFC0E1108A8FD17B448A68554199C47D08FFB10D4B8


def ec_add(a, b):
    if a is None:
        return b
    if b is None:
        return a
    if a[0] == b[0]:
        if (a[1] + b[1]) % P == 0:
            return None
        lam = 3 * a[0] * a[0] * pow(2 * a[1], -1, P) % P
    else:
        lam = (b[1] - a[1]) * pow(b[0] - a[0], -1, P) % P
    x = (lam * lam - a[0] - b[0]) % P
    return x, (lam * (a[0] - x) - a[1]) % P


def ec_mul_py(k):
    r, q = None, (GX, GY)
    while k:
        if k & 1:
            r = ec_add(r, q)
        q = ec_add(q, q)
        k >>= 1
    return r


def ec_mul(k):
    nums = ec.derive_private_key(k, ec.SECP256K1()).public_key().public_numbers()
    return nums.x, nums.y


def keccak256(data):
    RC = [0x0000000000000001, 0x0000000000008082, 0x800000000000808A, 0x8000000080008000, 0x000000000000808B,
          0x0000000080000001, 0x8000000080008081, 0x8000000000008009, 0x000000000000008A, 0x0000000000000088,
          0x0000000080008009, 0x000000008000000A, 0x000000008000808B, 0x800000000000008B, 0x8000000000008089,
          0x8000000000008003, 0x8000000000008002, 0x8000000000000080, 0x000000000000800A, 0x800000008000000A,
          0x8000000080008081, 0x8000000000008080, 0x0000000080000001, 0x8000000080008008]
    M = (1 << 64) - 1
    rol = lambda v, n: ((v << n) | (v >> (64 - n))) & M if n else v
    rate = 136
    msg = bytearray(data)
    padlen = rate - len(msg) % rate
    msg += b"\x81" if padlen == 1 else b"\x01" + b"\x00" * (padlen - 2) + b"\x80"
    A = [[0] * 5 for _ in range(5)]
    for off in range(0, len(msg), rate):
        for i in range(rate // 8):
            A[i % 5][i // 5] ^= int.from_bytes(msg[off + 8 * i:off + 8 * i + 8], "little")
        for rnd in range(24):
            Cc = [A[x][0] ^ A[x][1] ^ A[x][2] ^ A[x][3] ^ A[x][4] for x in range(5)]
            D = [Cc[(x - 1) % 5] ^ rol(Cc[(x + 1) % 5], 1) for x in range(5)]
            A = [[A[x][y] ^ D[x] for y in range(5)] for x in range(5)]
            B = [[0] * 5 for _ in range(5)]
            x, y = 1, 0
            B[0][0] = A[0][0]
            rot = {}
            cx, cy = 1, 0
            for t in range(24):
                rot[(cx, cy)] = ((t + 1) * (t + 2) // 2) % 64
                cx, cy = cy, (2 * cx + 3 * cy) % 5
            for x in range(5):
                for y in range(5):
                    B[y][(2 * x + 3 * y) % 5] = rol(A[x][y], rot.get((x, y), 0))
            A = [[B[x][y] ^ ((~B[(x + 1) % 5][y]) & B[(x + 2) % 5][y]) for y in range(5)] for x in range(5)]
            A[0][0] ^= RC[rnd]
    out = b""
    for i in range(4):
        out += A[i % 5][i // 5].to_bytes(8, "little")
    return out


B58 = "123456789ABCDEFGHJKLMNPQRSTUVWXYZabcdefghijkmnopqrstuvwxyz"


def b58check(payload):
    data = payload + hashlib.sha256(hashlib.sha256(payload).digest()).digest()[:4]
    n = int.from_bytes(data, "big")
    s = ""
    while n:
        n, r = divmod(n, 58)
        s = B58[r] + s
    return "1" * (len(data) - len(data.lstrip(b"\0"))) + s


CHARSET = "qpzry9x8gf2tvdw0s3jn54khce6mua7l"


def bech32_polymod(values):
    GEN = [0x3B6A57B2, 0x26508E6D, 0x1EA119FA, 0x3D4233DD, 0x2A1462B3]
    chk = 1
    for v in values:
        b = chk >> 25
        chk = (chk & 0x1FFFFFF) << 5 ^ v
        for i in range(5):
            chk ^= GEN[i] if ((b >> i) & 1) else 0
    return chk


def segwit_addr(witver, prog):
    acc = bits = 0
    data = [witver]
    for b in prog:
        acc = (acc << 8) | b
        bits += 8
        while bits >= 5:
            bits -= 5
            data.append((acc >> bits) & 31)
    if bits:
        data.append((acc << (5 - bits)) & 31)
    hrp = "bc"
    vals = [ord(c) >> 5 for c in hrp] + [0] + [ord(c) & 31 for c in hrp] + data
    const = 1 if witver == 0 else 0x2BC830A3
    pm = bech32_polymod(vals + [0] * 6) ^ const
    return hrp + "1" + "".join(CHARSET[d] for d in data + [(pm >> 5 * (5 - i)) & 31 for i in range(6)])


def h160(b):
    return hashlib.new("ripemd160", hashlib.sha256(b).digest()).digest()


def eip55(addr20):
    low = addr20.hex()
    h = keccak256(low.encode()).hex()
    return "0x" + "".join(c.upper() if int(h[i], 16) >= 8 else c for i, c in enumerate(low))


def derive(k):
    x, y = ec_mul(k)
    kb = k.to_bytes(32, "big")
    comp = bytes([2 + (y & 1)]) + x.to_bytes(32, "big")
    unc = b"\x04" + x.to_bytes(32, "big") + y.to_bytes(32, "big")
    out = {"key": kb.hex(), "x": "%064x" % x, "y": "%064x" % y}
    hc = h160(comp)
    out["p2pkh"] = {"digest": hc.hex(), "address": b58check(b"\0" + hc), "wif": b58check(b"\x80" + kb + b"\x01")}
    hu = h160(unc)
    out["p2pkh-uncompressed"] = {"digest": hu.hex(), "address": b58check(b"\0" + hu), "wif": b58check(b"\x80" + kb)}
    out["p2wpkh"] = {"digest": hc.hex(), "address": segwit_addr(0, hc), "wif": out["p2pkh"]["wif"]}
    hs = h160(b"\x00\x14" + hc)
    out["p2sh-p2wpkh"] = {"digest": hs.hex(), "address": b58check(b"\x05" + hs), "wif": out["p2pkh"]["wif"]}
    tag = hashlib.sha256(b"TapTweak").digest()
    t = int.from_bytes(hashlib.sha256(tag + tag + x.to_bytes(32, "big")).digest(), "big")
    q = ec_add((x, y if y % 2 == 0 else P - y), ec_mul_py(t))
    qx = q[0].to_bytes(32, "big")
    out["p2tr"] = {"digest": qx.hex(), "address": segwit_addr(1, qx), "wif": out["p2pkh"]["wif"]}
    ea = keccak256(unc[1:])[12:]
    out["ethereum"] = {"digest": ea.hex(), "address": eip55(ea), "wif": kb.hex()}
    return out


def main():
    here = os.path.dirname(os.path.abspath(__file__))
    assert keccak256(b"").hex() == "c5d2460186f7233c927e7db2dcc703c0e500b653ca82273b7bfad8045d85a470"
    # reference known answers
    k = int.from_bytes(hashlib.sha256(b"correct horse battery staple").digest(), "big")
    assert derive(k)["p2pkh"]["address"] == "1C7zdTfnkzmr13HfA2vNm5SJYRK6nEKyq8"          # address.rs:233-238
    assert derive(1)["p2pkh"]["address"] == "1BgGZ9tcN4rm9KBzDn7KprQz87SZ26SAMH"          # provider.rs:76-87
    rk = 0x0C28FCA386C7A227600B2FE50B7CAE11EC86D3BF1FBE471BE89827E19D72AA1D
    d = derive(rk)["p2pkh-uncompressed"]                                                 # README.md:121-127
    assert d["wif"] == "5HueCGU8rMjxEXxiPuD5BDku4MkFqeZyd4dZ1jvhTVqvbTLvyTJ" and d["address"] == "1GAehh7TsJAHuUAeKZcXf5CnwuGuGgyX2S"
    assert ec_mul(k) == ec_mul_py(k) and ec_mul(N - 1) == ec_mul_py(N - 1)
    rng = random.Random(42)
    scalars = [1, 2, 3, 4, 5, 255, 256, 2**32 - 5, 2**32, 2**64, 2**64 + 1, 2**64 + 12345, 2**65, 2**65 + int(0.83 * 2**40),
               k, rk, N - 1, N - 2, N - 2**20, 2**255, 2**128 + 7]
    scalars += [rng.randrange(1, N) for _ in range(64)]
    scalars += [rng.randrange(1, 2**40) for _ in range(16)]
    vecs = [derive(s) for s in scalars]
    # sequential runs: hash160 (compressed) for consecutive scalars, the reference GPU contract (gpu.rs:660)
    runs = []
    for start, n in ((1, 64), (2**64, 64), (2**32 - 5, 16), (N - 40, 39), (rng.randrange(1, N), 64)):
        hs = []
        for i in range(n):
            x, y = ec_mul(start + i)
            hs.append(h160(bytes([2 + (y & 1)]) + x.to_bytes(32, "big")).hex())
        runs.append({"start": "%064x" % start, "hash160": hs})
    json.dump({"vectors": vecs, "runs": runs,
               "invalid_keys": ["%064x" % 0, "%064x" % N, "%064x" % (N + 1), "f" * 64]},
              open(os.path.join(here, "addresses.json"), "w"), indent=0)

    # regex cases: (pattern, case_insensitive, haystack) -> bool by Python re
    pats = ["^1", "^1Cat", "^1CatDog", "dead$", "^0xdead", "^bc1q", "Cat", "^1[A-Z]", "^1[a-z]{2}", "(?i)^1cat",
            "^1(Cat|Dog)", "^1.*z$", "^1Cat|Dog$", "[0-9]{3}$", "^1[^a-z]", "^3", "^bc1q.*dead$", "^0x[dD][eE]", "a+b*c?",
            "^1[A-Za-z0-9]+$", "^.{34}$", "^.{33}$", "x{2,3}", "(ab){2}", "^1\\d", "\\w+$", "^(?:1|3)", "^bc1q[02-9ac-hj-np-z]{38}$",
            "^1C[a-c]t", "q$", "^1BgGZ9tcN4rm9KBzDn7KprQz87SZ26SAMH$", "^$", "1{2}", "(?i)DEAD", "^1(?i:c)at",
            "^1\\x43", "A|B|C$", "^(1C|1D)(a|o)(t|g)", "z{0}", "^1C?at", "(^1Cat)|(^1Dog)", "."]
    hay = [v[f]["address"] for v in vecs[:24] for f in ("p2pkh", "p2wpkh", "p2sh-p2wpkh", "ethereum")]
    hay += ["1CatDogxyz", "1Catxyz", "1cat", "1CAT", "bc1qdead", "0xdeadbeef", "0xDEADbeef", "", "1", "11", "abc", "aabbc", "1Dog", "1Cot", "xx", "xxx", "abab"]
    results = {}
    for p in pats:
        for ci in (False, True):
            rx = re.compile(p, re.I if ci else 0)
            results["%d|%s" % (ci, p)] = "".join("1" if rx.search(h) else "0" for h in hay)
    # constructs beyond the plain subset (round 2): word boundaries and flag x by Python `re`; nested classes, the set operators
    # && -- ~~ and \\p{..} by the third-party `regex` module in VERSION1 mode. Where Python and Rust `regex` differ the case is
    # left out rather than bent: whitespace inside a class under flag x (Python keeps it, Rust drops it) and mixed set
    # operators without brackets (`regex` gives them different precedences, Rust equal precedence left to right).
    import regex as regex_mod
    hay_x = ["1Cat-Dog", "a cat b", "0xdead beef", "cat", "1Cat Dog", "x_y z", "ABcd", "abCD12", "1-2", "a b"]
    for p in pats:
        for ci in (False, True):
            rx = re.compile(p, re.I if ci else 0)
            results["%d|%s" % (ci, p)] += "".join("1" if rx.search(h) else "0" for h in hay_x)
    hay = hay + hay_x
    ext_re = ["\\b1C", "t\\b", "\\Bat", "\\bcat\\b", "\\b[a-z]+\\b", "^\\b", "\\b$", "\\B$", "d\\b.\\bb", "\\b\\d\\b", "a\\b b", "(?i)\\bCAT\\b", "\\b(Cat|Dog)\\b", "x\\B_",
              "(?x) ^1 C a t", "(?x)^1 [A-Z] # upper\n[a-z]", "(?x: d e a d)$", "(?x)a\\ b", "(?x) c a t | D o g ", "^(?x: 1 C ) at", "(?x)x{ 2 , 3 }"[:0] or "(?x)x{2,3}",
              "\u00e9", "1Cat\u00e9?", "[\u00e9]", "[a-\u00e9]{3}$", "1[^\u00e9]at"]
    ext_v1 = ["[a-z&&[^aeiou]]{3}$", "^1[A-Z--[IO]]", "[0-9a-f~~[a-c]]{4}$", "^1[[A-C][a-c]]", "[^[a-z][0-9]]$", "(?i)[a-c&&[^b]]$", "[a-c&&B]$",
              "^\\p{Lu}", "\\p{Ll}{5}$", "[\\p{L}--\\p{Lu}]{4}$", "\\P{L}$", "\\pN{3}$", "[\\d&&[^5-9]]+$", "[a-z--b--c]$", "[[:alpha:]&&[^a-m]]{2}$", "[\\w--\\d]$"]
    for p in ext_re:
        for ci in (False, True):
            rx = re.compile(p, re.I if ci else 0)
            results["%d|%s" % (ci, p)] = "".join("1" if rx.search(h) else "0" for h in hay)
    # pinned by hand: Python < 3.14 never lets \\B match in an EMPTY string; Rust `regex` does (no word byte on either side)
    for ci in (False, True):
        k = "%d|%s" % (ci, "\\B$")
        i = hay.index("")
        assert results[k][i] == "0"
        results[k] = results[k][:i] + "1" + results[k][i + 1:]
    for p in ext_v1:
        for ci in (False, True):
            if ci and "[^" in p and not p.startswith("(?i)"):
                continue        # negated class under -i: Rust folds, then negates; `regex` V1 does not always agree - left out
            rx = regex_mod.compile(p, regex_mod.V1 | (regex_mod.I if ci else 0))
            results["%d|%s" % (ci, p)] = "".join("1" if rx.search(h) else "0" for h in hay)
    bad = ["", "(", ")", "[", "a{2,1}", "*a", "a**b" if False else "+", "(?P<n", "\\", "[z-a]", "(?z)a", "a{", "\\q", "[[a]", "\\p{Foo}", "(?x", "[a-\\d]", "\\p{L"]
    json.dump({"haystacks": hay, "results": results, "invalid": [b for b in bad if b != ""]},
              open(os.path.join(here, "regex_cases.json"), "w"), indent=0)
    print(len(vecs), "vectors;", len(results) * len(hay), "regex cases")


if __name__ == "__main__":
    main()
