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
