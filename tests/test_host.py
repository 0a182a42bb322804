"""CPU-side tests of the product's host layer (no GPU): the C-ABI library loads and exports every declared symbol,
pattern compile/match semantics (src/pattern.rs:296-350), pattern lowering (intervals / exact / dfa), address and WIF
string encoding against the golden vectors. The oracle appears only as the checker."""
import ctypes as C
import os
import re

import pytest

import oracle
import vgen_b200 as vg
from vgen_b200 import _cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FORMATS = ["p2pkh", "p2pkh-uncompressed", "p2wpkh", "p2sh-p2wpkh", "p2tr", "ethereum"]


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "vgen_cuda.h")).read()
    declared = set(re.findall(r"\b(vgen_[a-z0-9_]+)\s*\(", hdr)) - {"vgen_progress_fn"}
    L = C.CDLL(_cabi.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), name
    assert declared == set(_cabi.SYMBOLS), declared ^ set(_cabi.SYMBOLS)


def test_no_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_cabi.VgenError):
        vg.GpuRunner()
    with pytest.raises(RuntimeError):
        vg.GpuRunner(backend=vg.GpuBackend.Vulkan)


def test_format_strings_golden(golden):
    for v in golden["vectors"]:
        key = bytes.fromhex(v["key"])
        for i, f in enumerate(FORMATS):
            fmt = vg.AddressFormat.parse(f)
            assert int(fmt) == i
            assert vg.address_from_digest(fmt, bytes.fromhex(v[f]["digest"])) == v[f]["address"]
            assert vg.wif_from_key(fmt, key) == v[f]["wif"]


def test_pattern_semantics_reference_cases():
    # src/pattern.rs:301-350
    with pytest.raises(ValueError):
        vg.Pattern("")
    with pytest.raises(ValueError):
        vg.Pattern("[invalid")
    assert vg.Pattern("^1").matches("1BvBMSEYstWetqTFn5Au4m4GFg7xJaNVN2")
    assert not vg.Pattern("^1").matches("3J98t1WpEZ73CNmQviecrnyiWrnqRhWNLy")
    assert vg.Pattern("[0-9]{3}$").matches("1BvBMSEYstWetqTFn5Au4m4GFg7xJaN123")
    assert vg.Pattern("^1cat", True).matches("1CATxyz") and not vg.Pattern("^1cat").matches("1CATxyz")
    p = vg.Pattern("^1[Cc]at")
    assert p.matches("1Catabc") and p.matches("1catabc") and not p.matches("1CATabc")
    assert p.original() == "^1[Cc]at" and not p.is_case_insensitive()


def test_pattern_vs_regex_golden(regex_golden):
    hay = regex_golden["haystacks"]
    for key, bits in regex_golden["results"].items():
        ci, pat = key.split("|", 1)
        p = vg.Pattern(pat, ci == "1")
        got = "".join("1" if p.matches(h) else "0" for h in hay)
        assert got == bits, (pat, ci)
    for bad in regex_golden["invalid"]:
        with pytest.raises(ValueError):
            vg.Pattern(bad)


def test_lowering_kinds():
    F = vg.AddressFormat
    assert vg.Pattern("^1Cat").lowering(F.P2pkh)["kind"] == "intervals"
    assert len(vg.Pattern("^1Cat").lowering(F.P2pkh)["intervals"]) == 2
    assert vg.Pattern("^1").lowering(F.P2pkh)["kind"] == "all"
    assert vg.Pattern(".").lowering(F.P2pkh)["kind"] == "dfa"
    assert vg.Pattern("^3").lowering(F.P2pkh)["kind"] == "none"
    assert vg.Pattern("^3").lowering(F.P2shP2wpkh)["kind"] == "all"
    assert vg.Pattern("dead$").lowering(F.P2wpkh)["kind"] == "dfa"
    assert vg.Pattern("^bc1qdead").lowering(F.P2wpkh)["kind"] == "intervals"
    assert vg.Pattern("^0xdead").lowering(F.Ethereum)["kind"] == "intervals"
    assert len(vg.Pattern("^1cat", True).lowering(F.P2pkh)["intervals"]) > 2
    ex = vg.Pattern(r"^1BgGZ9tcN4rm9KBzDn7KprQz87SZ26SAMH$").lowering(F.P2pkh)
    assert ex["kind"] == "exact" and ex["exact"].hex() == "751e76e8199196d454941c45d1b3a323f1433bd6"
    ex = vg.Pattern(r"^bc1qw508d6qejxtdg4y5r3zarvary0c5xw7kv8f3t4$").lowering(F.P2wpkh)
    assert ex["kind"] == "exact" and ex["exact"].hex() == "751e76e8199196d454941c45d1b3a323f1433bd6"
    assert vg.Pattern(r"^1BgGZ9tcN4rm9KBzDn7KprQz87SZ26SAMh$").lowering(F.P2pkh)["kind"] == "none"   # bad checksum


def test_lowering_alternations_read_off_the_dfa():
    """Anchored patterns that are not a plain run of character sets (alternations, optional tails) still become
    digest-interval prefilters: the prefixes are read off the compiled DFA (lowering.cpp derive_prefixes_from_dfa)."""
    F = vg.AddressFormat
    cases = [
        (F.P2pkh, "^(1Cat|1Dog)", "intervals", 4), (F.P2pkh, "^1(Cat|Dog)s?", "intervals", 4),
        (F.P2wpkh, "^bc1q(aa|zz)", "intervals", 2), (F.Ethereum, "^0x(dead|beef)", "intervals", 2),
        (F.P2tr, "^bc1p(q|p)q", "intervals", 2), (F.P2pkh, "(1Cat|1Dog)", "dfa", None),
        (F.P2pkh, "^(3Cat|3Dog)", "none", None), (F.P2pkh, "^1.*x", "dfa", None),
        (F.P2pkh, "^(1BgGZ9tcN4rm9KBzDn7KprQz87SZ26SAMH|1CUNEBjYrCn2y1SdiUMohaKUi4wpP326Lb)$", "intervals", 2),
    ]
    for fmt, pat, kind, n in cases:
        low = vg.Pattern(pat).lowering(fmt)
        assert low["kind"] == kind, (pat, low)
        if n is not None:
            assert len(low["intervals"]) == n, (pat, low)
    # the union of the branches' intervals equals the alternation's intervals
    a = vg.Pattern("^1Cat").lowering(F.P2pkh)["intervals"]
    b = vg.Pattern("^1Dog").lowering(F.P2pkh)["intervals"]
    assert sorted(a + b) == sorted(vg.Pattern("^(1Cat|1Dog)").lowering(F.P2pkh)["intervals"])


@pytest.mark.parametrize("fmt", ["p2pkh", "p2wpkh", "p2sh-p2wpkh", "ethereum"])
def test_alternation_intervals_are_a_superset(fmt, golden):
    """Completeness of the DFA-derived prefilter: every golden address is accepted by ^(own prefix|other prefix)."""
    F = vg.AddressFormat.parse(fmt)
    vs = golden["vectors"][:30]
    for i, v in enumerate(vs):
        addr, dg = v[fmt]["address"], bytes.fromhex(v[fmt]["digest"])
        other = vs[(i + 7) % len(vs)][fmt]["address"]
        for n in (3, 5, 7, 10):
            pat = "^(" + re.escape(other[:n]) + "|" + re.escape(addr[:n]) + ")"
            p = vg.Pattern(pat)
            assert p.matches(addr)
            low = p.lowering(F)
            assert low["kind"] in ("intervals", "all"), (pat, low)
            assert _covered(low, dg), (pat, low)


def _covered(low, digest):
    if low["kind"] in ("all", "dfa"):
        return True
    if low["kind"] == "exact":
        return digest[:len(low["exact"])] == low["exact"]
    top = int.from_bytes(digest[:8], "big")
    return any(lo <= top <= hi for lo, hi in low["intervals"])


@pytest.mark.parametrize("fmt", ["p2pkh", "p2pkh-uncompressed", "p2wpkh", "p2sh-p2wpkh", "ethereum"])
def test_intervals_are_a_superset_and_tight(fmt, golden):
    """Completeness: for every golden address and every prefix length, the device prefilter derived from ^prefix must
    accept the address's own digest. Tightness: a random digest accepted by the prefilter almost always really matches."""
    import random
    rng = random.Random(7)
    F = vg.AddressFormat.parse(fmt)
    for v in golden["vectors"][:40]:
        addr, dg = v[fmt]["address"], bytes.fromhex(v[fmt]["digest"])
        for n in (1, 2, 3, 4, 5, 6, 8, 12, len(addr)):
            pat = "^" + re.escape(addr[:n]) + ("$" if n == len(addr) else "")
            low = vg.Pattern(pat).lowering(F)
            assert _covered(low, dg), (pat, low)
    # tightness on a 4-char prefix: sample digests inside the intervals, encode, check the prefix
    addr = golden["vectors"][3][fmt]["address"]
    n = {"p2wpkh": 7, "ethereum": 6}.get(fmt, 4)
    low = vg.Pattern("^" + addr[:n]).lowering(F)
    assert low["kind"] == "intervals"
    ok = tot = 0
    for lo, hi in low["intervals"]:
        for _ in range(200):
            top = rng.randint(lo, hi)
            dg = top.to_bytes(8, "big") + bytes(rng.randrange(256) for _ in range(12))
            a = vg.address_from_digest(F, dg)
            tot += 1
            ok += a.lower().startswith(addr[:n].lower()) if fmt == "ethereum" else a.startswith(addr[:n])
    assert ok >= tot * 0.98, (ok, tot)


def test_intervals_exhaustive_small_prefixes():
    """Soundness of base58 interval lowering against brute force: for random digests, membership in the intervals of
    ^1X / ^1XY equals the string test, except within one checksum-width of an interval edge."""
    import random
    rng = random.Random(11)
    F = vg.AddressFormat.P2pkh
    pats = ["^1A", "^1Zz", "^11", "^12", "^1z", "^111", "^1Cat", "^1Q"]
    lows = {p: vg.Pattern(p).lowering(F) for p in pats}
    for _ in range(4000):
        r = rng.random()
        dg = bytes(rng.randrange(256) for _ in range(20))
        if r < 0.2:
            dg = b"\0" + dg[1:]
        if r < 0.05:
            dg = b"\0\0" + dg[2:]
        a = oracle.address_from_digest("p2pkh", dg)
        for p in pats:
            if a.startswith(p[1:]):
                assert _covered(lows[p], dg), (p, a, dg.hex())


# ---- pattern lint / difficulty / provider (CLI support): the reference's own unit tests, src/pattern.rs:355-635 ----
@pytest.mark.parametrize("pat,ci,fmt,want", [
    ("^1Ab", False, "p2pkh", 58 ** 2), ("^1Ab", True, "p2pkh", 34 ** 2), ("^bc1qab", False, "p2wpkh", 32 ** 2), ("^1", False, "p2pkh", 1),
    ("^0xdead", False, "ethereum", 16 ** 4), ("dead$", False, "p2pkh", 58 ** 4), ("Cat", False, "p2pkh", 58 ** 3),
    ("^Cat", False, "p2pkh", 58 ** 3), ("^bc1ab", False, "p2wpkh", 32 ** 2), ("^bc1pab", False, "p2tr", 32 ** 2),
    ("^1[Oo]ri", False, "p2pkh", 58 ** 2), ("^1" + "A" * 30, False, "p2pkh", 2 ** 64 - 1),      # saturating_pow
])
def test_estimate_difficulty_reference_cases(pat, ci, fmt, want):
    assert vg.Pattern(pat, ci).estimate_difficulty(vg.AddressFormat.parse(fmt)) == want


@pytest.mark.parametrize("pat,fmt,want", [
    ("^1OR", "p2pkh", ["O"]), ("^1Cat", "p2pkh", []), ("^10ri", "p2pkh", ["0"]), ("^1OIl0", "p2pkh", ["O", "I", "l", "0"]),
    ("^bc1qAB", "p2wpkh", ["b", "1", "A", "B"]), ("^0xghi", "ethereum", ["g", "h", "i"]), ("^1[0OIl]", "p2pkh", ["0", "O", "I", "l"]),
    ("^1[Aa]", "p2pkh", []), ("^1[a-z]", "p2pkh", []), ("^1[A0]", "p2pkh", []), ("^1[^0]", "p2pkh", []), ("^1[0-9a]", "p2pkh", []),
    ("^1[I-O]", "p2pkh", []), ("^1[0]", "p2pkh", ["0"]), ("^1[-A]", "p2pkh", []), ("^1[A-]", "p2pkh", []), ("^1[-0]", "p2pkh", ["0"]),
    ("^1[.]", "p2pkh", ["."]), ("^1[_]", "p2pkh", ["_"]), ("^1[.A]", "p2pkh", []), ("^1[\\.]", "p2pkh", ["."]), ("^1[\\^^]", "p2pkh", ["^"]),
])
def test_validate_charset_reference_cases(pat, fmt, want):
    assert vg.Pattern(pat).validate_charset(vg.AddressFormat.parse(fmt)) == want


def test_provider_shim():
    """src/provider.rs:12-63,76-87: boha:b1000:1 is the address of key 1; ranges are 2^(N-1)..2^N-1; every embedded
    solved entry is re-derived from its published key by the oracle."""
    import oracle
    from vgen_b200 import provider
    assert provider.resolve("^1Cat") is None and provider.resolve("foo:bar") is None
    r = provider.resolve("boha:b1000:1")
    assert r.address == "1BgGZ9tcN4rm9KBzDn7KprQz87SZ26SAMH" and r.format == vg.AddressFormat.P2pkh and r.key_range == (1, 1)
    r66 = provider.resolve("boha:b1000:66")
    assert r66.address == "13zb1hQbWVsc2S7ZTZnP2G4undNNpdh5so" and r66.key_range == (2 ** 65, 2 ** 66 - 1)      # README.md:107-108
    assert provider.build_pattern(r66, 5) == "^13zb1" and provider.build_exact_pattern(r66) == "^13zb1hQbWVsc2S7ZTZnP2G4undNNpdh5so$"
    assert provider.build_pattern(r66, 999) == "^13zb1hQbWVsc2S7ZTZnP2G4undNNpdh5so"
    solved = {1: 1, 2: 3, 3: 7, 4: 8, 5: 0x15, 6: 0x31, 7: 0x4c, 8: 0xe0, 9: 0x1d3, 10: 0x202, 11: 0x483, 12: 0xa7b, 13: 0x1460, 14: 0x2930,
              15: 0x68f3, 16: 0xc936, 17: 0x1764f, 18: 0x3080d, 19: 0x5749f, 20: 0xd2c55}
    for n, k in solved.items():
        r = provider.resolve("boha:b1000:%d" % n)
        assert r.key_range[0] <= k <= r.key_range[1] and oracle.address("p2pkh", k)[0] == r.address
    for bad in ("boha:b1000:67", "boha:b1000:x", "boha:other:1"):
        with pytest.raises(ValueError):
            provider.resolve(bad)


def test_parse_key():
    import ctypes as C
    from vgen_b200 import _cabi
    L = _cabi.lib()
    key, w, c = C.create_string_buffer(32), C.c_int(0), C.c_int(0)
    # README.md:121-127: WIF 5HueCGU8... <-> hex 0c28fca3...
    assert L.vgen_parse_key(b"5HueCGU8rMjxEXxiPuD5BDku4MkFqeZyd4dZ1jvhTVqvbTLvyTJ", key, C.byref(w), C.byref(c)) == 0
    assert key.raw.hex() == "0c28fca386c7a227600b2fe50b7cae11ec86d3bf1fbe471be89827e19d72aa1d" and w.value == 1 and c.value == 0
    assert L.vgen_parse_key(b"KwDiBf89QgGbjEhKnhXJuH7LrciVrZi3qYjgd9M7rFU73sVHnoWn", key, C.byref(w), C.byref(c)) == 0
    assert int.from_bytes(key.raw, "big") == 1 and c.value == 1
    assert L.vgen_parse_key(b"0x" + b"00" * 31 + b"02", key, C.byref(w), C.byref(c)) == 0 and w.value == 0 and key.raw[-1] == 2
    assert L.vgen_parse_key(b"00" * 32, key, None, None) < 0 and L.vgen_parse_key(b"zz", key, None, None) < 0
    assert L.vgen_parse_key(b"ff" * 32, key, None, None) < 0


def test_match_list_is_a_lazy_sequence():
    from vgen_b200.scanner import MatchList
    text = b"00aa 1Abc Kx1\n00bb 1Def Kx2\n"
    m = MatchList(text, 2, vg.AddressFormat.P2pkh)
    assert len(m) == 2 and m._items is None                    # len() does not parse
    assert m[0] == vg.GeneratedAddress("1Abc", "Kx1", "00aa", vg.AddressFormat.P2pkh) and m[-1].hex == "00bb"
    assert [x.address for x in m] == ["1Abc", "1Def"] and m == list(m) and len(MatchList(b"", 0, vg.AddressFormat.P2pkh)) == 0


def test_literal_suffix_lowering():
    """dead$ on segwit formats / F00$ on Ethereum: the DFA stays the decider, and the host adds a (mask, value) test over the
    six Bech32 checksum symbols / the last 32 digest bits that the device runs first (lowering.cpp, scan_types.h)."""
    F = vg.AddressFormat
    CH = "qpzry9x8gf2tvdw0s3jn54khce6mua7l"
    def bech(lit):
        v = 0
        for ch in lit:
            v = (v << 5) | CH.index(ch)
        return ((1 << (5 * len(lit))) - 1, v)
    for fmt in (F.P2wpkh, F.P2tr):
        for lit in ("dead", "q", "zzzzzz", "7l"):
            low = vg.Pattern(lit + "$").lowering(fmt)
            assert low["kind"] == "dfa" and low["suffix"] == bech(lit), (fmt, lit, low)
        assert vg.Pattern("DEAD$", True).lowering(fmt)["suffix"] == bech("dead")
        # not literal suffixes of Bech32 characters (b, i, o, 1 are not in the alphabet), too long, upper case without -i
        for pat in ("de.d$", "dead", "^dead$", "bead$", "qqqqqqq$", "DEAD$", "d[e]ad$", "$"):
            assert vg.Pattern(pat).lowering(fmt)["suffix"] is None, (fmt, pat)
    for lit in ("F00", "f00", "dEaD", "12345678", "0"):
        low = vg.Pattern(lit + "$").lowering(F.Ethereum)
        assert low["kind"] == "dfa" and low["suffix"] == ((1 << (4 * len(lit))) - 1, int(lit, 16)), (lit, low)
    for pat in ("g00$", "123456789$", "f.0$", "^0xf00$", "f00"):
        assert vg.Pattern(pat).lowering(F.Ethereum)["suffix"] is None, pat
    for fmt in (F.P2pkh, F.P2shP2wpkh, F.P2pkhUncompressed):
        assert vg.Pattern("dead$").lowering(fmt)["suffix"] is None


def test_published_vectors_through_the_product_encoders(published):
    """The product's own host encoders (vgen_format_address / vgen_format_wif / vgen_parse_key, host_util.cpp — what turns a device
    hit into the reported strings) against vectors copied from the specifications (tests/golden/published.json)."""
    F = vg.AddressFormat
    for a in published["eip55"]["addresses"]:                                     # EIP-55 (Keccak-256 inside)
        assert vg.address_from_digest(F.Ethereum, bytes.fromhex(a[2:])) == a
    for v in published["base58check_p2pkh"]["vectors"]:                           # Base58Check (double SHA-256 inside)
        assert vg.address_from_digest(F.P2pkh, bytes.fromhex(v["hash160"])) == v["address"]
        assert vg.address_from_digest(F.P2pkhUncompressed, bytes.fromhex(v["hash160"])) == v["address"]
    for v in published["segwit"]["vectors"]:                                      # BIP173 / BIP350
        prog = bytes.fromhex(v["program"])
        if v["witver"] == 0 and len(prog) == 20:
            assert vg.address_from_digest(F.P2wpkh, prog) == v["address"].lower()
        if v["witver"] == 1 and len(prog) == 32:
            assert vg.address_from_digest(F.P2tr, prog) == v["address"].lower()
    for v in published["taproot"]["vectors"]:                                     # BIP341 / BIP86 output key -> address
        assert vg.address_from_digest(F.P2tr, bytes.fromhex(v["output_key"])) == v["address"]
    for v in published["wif"]["vectors"]:
        k = bytes.fromhex(v["key"])
        assert vg.wif_from_key(F.P2pkhUncompressed, k) == v["wif_uncompressed"]
        assert vg.wif_from_key(F.P2pkh, k) == v["wif_compressed"] == vg.wif_from_key(F.P2wpkh, k)
        from vgen_b200 import _cabi
        for wif, comp in ((v["wif_uncompressed"], 0), (v["wif_compressed"], 1)):
            out, is_wif, compressed = C.create_string_buffer(32), C.c_int(-1), C.c_int(-1)
            assert _cabi.lib().vgen_parse_key(wif.encode(), out, C.byref(is_wif), C.byref(compressed)) == 0
            assert (out.raw, is_wif.value, compressed.value) == (k, 1, comp)
    # exact-address patterns are decoded on the host (Base58Check / Bech32(m) decode + checksum): published address -> digest
    v = published["base58check_p2pkh"]["vectors"][1]
    assert vg.Pattern("^" + v["address"] + "$").lowering(F.P2pkh)["exact"].hex() == v["hash160"]
    t = published["taproot"]["vectors"][0]
    assert vg.Pattern("^" + t["address"] + "$").lowering(F.P2tr)["exact"].hex() == t["output_key"]
    e = published["eip55"]["addresses"][4]
    assert vg.Pattern("^" + e + "$").lowering(F.Ethereum)["exact"].hex() == e[2:].lower()
    assert vg.Pattern("^" + e.lower() + "$").lowering(F.Ethereum)["kind"] == "none"      # wrong case can never match the EIP-55 form


def test_product_encoders_vs_oracle_random_and_leading_zeros():
    """vgen_format_address / vgen_format_wif (host_util.cpp: SHA-256, Keccak, Base58 on 32-bit limbs in groups of five digits,
    Bech32(m)) against the oracle's independent encoders: random digests and keys, runs of leading zero bytes (Base58 '1's, a top
    digit group without leading zeros), all-zero and all-ones inputs."""
    import random
    rnd = random.Random(58)
    F = vg.AddressFormat
    cases = [bytes(32), b"\xff" * 32, b"\0" * 19 + b"\1" + bytes(12), b"\0" * 31 + b"\1"]
    cases += [b"\0" * k + bytes(rnd.randrange(1, 256) for _ in range(32 - k)) for k in range(0, 32)]
    cases += [bytes(rnd.randrange(256) for _ in range(32)) for _ in range(1500)]
    for d in cases:
        for fmt in (F.P2pkh, F.P2pkhUncompressed, F.P2wpkh, F.P2shP2wpkh, F.P2tr, F.Ethereum):
            n = 32 if fmt == F.P2tr else 20
            assert vg.address_from_digest(fmt, d[:n]) == oracle.address_from_digest(int(fmt), d[:n]), (fmt, d.hex())
        # WIF = Base58Check(0x80 || key [|| 0x01]) for any 32 bytes (the encoder does not look at the scalar's range)
        assert vg.wif_from_key(F.P2pkh, d) == oracle.base58check(b"\x80" + d + b"\x01")
        assert vg.wif_from_key(F.P2pkhUncompressed, d) == oracle.base58check(b"\x80" + d)
        assert vg.wif_from_key(F.Ethereum, d) == d.hex()


def test_product_sha256_portable_and_sha_ni_paths_agree():
    """host_util.cpp picks the x86 SHA extensions at run time when the CPU has them; VGEN_NO_SHA_NI=1 keeps the portable
    compression. Both must give the oracle's Base58Check strings (whichever of them this machine runs by default)."""
    import subprocess
    import sys
    code = ("import vgen_b200 as vg\n"
            "F = vg.AddressFormat\n"
            "for i in range(64):\n"
            "    d = bytes((i * 37 + j * 11) & 255 for j in range(32))\n"
            "    print(vg.address_from_digest(F.P2pkh, d[:20]), vg.address_from_digest(F.P2shP2wpkh, d[:20]), vg.wif_from_key(F.P2pkh, d), vg.wif_from_key(F.P2pkhUncompressed, d))\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for extra in ({}, {"VGEN_NO_SHA_NI": "1"}):
        env = dict(os.environ, **extra)
        env.pop("VGEN_NO_SHA_NI", None) if not extra else None
        outs.append(subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, check=True).stdout)
    assert outs[0] == outs[1] and outs[0].count("\n") == 64
    first = outs[0].splitlines()[0].split()
    d = bytes((j * 11) & 255 for j in range(32))
    assert first == [oracle.base58check(b"\x00" + d[:20]), oracle.base58check(b"\x05" + d[:20]), oracle.base58check(b"\x80" + d + b"\x01"), oracle.base58check(b"\x80" + d)]


def test_regex_too_big_for_a_device_table_falls_back_to_the_host_nfa(regex_golden):
    """A pattern vgen accepts must not become an error after the swap (src/pattern.rs:21-40 takes whatever `regex` compiles).
    `a.{13}$` needs 2^14 DFA states, more than a device table holds: Pattern still compiles, matches() decides with the NFA, and
    the lowering says the device will report a superset (every key, or the prefix intervals) for the host to confirm."""
    F = vg.AddressFormat
    for pat in ("a.{13}$", "^1C.*a.{13}$", "(a|b).{12}[0-9]$"):
        p = vg.Pattern(pat)
        for h in regex_golden["haystacks"]:
            assert p.matches(h) == bool(re.search(pat, h)), (pat, h)
    low = vg.Pattern("a.{13}$").lowering(F.P2pkh)
    assert low["kind"] == "all" and low["confirm"] == "host-nfa", low
    low = vg.Pattern("^1Ca.*a.{13}$").lowering(F.P2pkh)
    assert low["kind"] == "intervals" and low["confirm"] == "host-nfa" and len(low["intervals"]) == 2, low
    assert vg.Pattern("a.{5}$").lowering(F.P2pkh)["confirm"] is None          # small enough: a device DFA as before


def test_word_boundary_is_one_bit_of_look_behind_in_the_dfa():
    """\\b / \\B compile into the dense DFA (state = NFA set x "previous byte was a word byte"), so they run on the device too."""
    F = vg.AddressFormat
    d = vg.Pattern(r"\bCat\b").device_dfa(F.P2pkh)
    assert 0 < d["n_states"] < 64
    for pat, text, want in [(r"\bcat\b", "a cat b", True), (r"\bcat\b", "acat", False), (r"\bcat\b", "cat", True), (r"cat\B", "cats", True),
                            (r"cat\B", "cat", False), (r"\B", "", True), (r"\b", "", False), (r"\b", "a", True), (r"^\B", "-", True)]:
        assert vg.Pattern(pat).matches(text) is want, (pat, text)
